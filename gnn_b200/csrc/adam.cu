// K8: the Adam update (torch.optim.Adam(lr) without weight decay or amsgrad: gnn/main.py:318) for up to eight parameter
// tensors in one launch.  torch's fused Adam walks a 5M-element weight in 77 chunks of 65 536 elements, one CTA each (about
// half of the GPU, 41 us per step in the launch list); this kernel spreads every tensor over the whole grid: one pass,
// 16 bytes per thread and tensor stream, 7 streams (read p, g, m, v; write p, m, v) = 28 bytes per element, HBM-bound.
//   m = m + (g - m) (1 - b1);  v = b2 v + (1 - b2) g g;  p -= (lr / (1 - b1^t)) m / (sqrt(v) / sqrt(1 - b2^t) + eps)
// -- the arithmetic of torch's fused kernel, in the same order, so the two agree to rounding.  The step count t lives on
// the device (float32, incremented by the caller before the launch), which keeps the update capturable in a CUDA graph.
#include "common.cuh"

namespace gnnb {

constexpr int ADAM_MAX_TENSORS = 8;
struct AdamArgs {
    float* p[ADAM_MAX_TENSORS];
    const float* g[ADAM_MAX_TENSORS];
    float* m[ADAM_MAX_TENSORS];
    float* v[ADAM_MAX_TENSORS];
    long long n[ADAM_MAX_TENSORS];
    int count;
};

__global__ void __launch_bounds__(256)
adam_kernel(AdamArgs a, const float* __restrict__ step, float lr, float beta1, float beta2, float eps) {
    __shared__ float s_step_size, s_bc2_sqrt;
    if (threadIdx.x == 0) {
        const double t = (double)*step;
        const float bc1 = (float)(1.0 - pow((double)beta1, t));
        const float bc2 = (float)(1.0 - pow((double)beta2, t));
        s_step_size = lr / bc1;
        s_bc2_sqrt = sqrtf(bc2);
    }
    __syncthreads();
    const float step_size = s_step_size, bc2_sqrt = s_bc2_sqrt, w1 = 1.0f - beta1, w2 = 1.0f - beta2;
    auto update = [&](float& p, float g, float& m, float& v) {
        m = m + (g - m) * w1;                               // lerp(m, g, 1 - beta1)
        v = beta2 * v + w2 * g * g;
        const float denom = sqrtf(v) / bc2_sqrt + eps;
        p -= step_size * m / denom;
    };
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x, stride = (long long)gridDim.x * blockDim.x;
    for (int k = 0; k < a.count; ++k) {
        float* p = a.p[k];
        const float* g = a.g[k];
        float* m = a.m[k];
        float* v = a.v[k];
        const long long n = a.n[k];
        const bool vec = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) |
                           reinterpret_cast<uintptr_t>(v)) & 15) == 0;
        const long long n4 = vec ? n / 4 : 0;
        for (long long i = tid; i < n4; i += stride) {
            float4 pp = reinterpret_cast<float4*>(p)[i], mm = reinterpret_cast<float4*>(m)[i], vv = reinterpret_cast<float4*>(v)[i];
            const float4 gg = reinterpret_cast<const float4*>(g)[i];
            update(pp.x, gg.x, mm.x, vv.x);
            update(pp.y, gg.y, mm.y, vv.y);
            update(pp.z, gg.z, mm.z, vv.z);
            update(pp.w, gg.w, mm.w, vv.w);
            reinterpret_cast<float4*>(p)[i] = pp;
            reinterpret_cast<float4*>(m)[i] = mm;
            reinterpret_cast<float4*>(v)[i] = vv;
        }
        for (long long i = 4 * n4 + tid; i < n; i += stride) {
            float pp = p[i], mm = m[i], vv = v[i];
            update(pp, g[i], mm, vv);
            p[i] = pp; m[i] = mm; v[i] = vv;
        }
    }
}

}  // namespace gnnb

using namespace gnnb;

extern "C" {

int gnnb_adam_step(int count, float* const* params, const float* const* grads, float* const* exp_avg, float* const* exp_avg_sq,
                   const int64_t* numel, const float* step, float lr, float beta1, float beta2, float eps, void* stream) {
    GNNB_ENTER();
    GNNB_REQUIRE(count >= 0 && count <= ADAM_MAX_TENSORS && step && (count == 0 || (params && grads && exp_avg && exp_avg_sq && numel)), GNNB_E_BADARG,
                 "adam_step: bad argument (at most %d tensors per call)", ADAM_MAX_TENSORS);
    if (count == 0) return GNNB_OK;
    AdamArgs a = {};
    a.count = count;
    long long total = 0;
    for (int k = 0; k < count; ++k) {
        GNNB_REQUIRE(params[k] && grads[k] && exp_avg[k] && exp_avg_sq[k] && numel[k] >= 0, GNNB_E_BADARG, "adam_step: tensor %d has a null pointer", k);
        a.p[k] = params[k]; a.g[k] = grads[k]; a.m[k] = exp_avg[k]; a.v[k] = exp_avg_sq[k]; a.n[k] = numel[k];
        total += numel[k];
    }
    long long blocks = (total / 4 + 255) / 256;
    const long long cap = (long long)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    adam_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(a, step, lr, beta1, beta2, eps);
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

}  // extern "C"
