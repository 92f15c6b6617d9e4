// K9: gradient averaging across the GPUs of one box through the NVSwitch (NVLS): every rank owns a slice of the
// buffer, pulls the SUM of that slice over all ranks with one multimem.ld_reduce per 16 bytes (the switch reads the
// replicas and adds them in flight), scales it to the mean and pushes it to every rank with one multimem.st.
// Per GPU and direction the links carry the buffer once (20 MB for the C5 gradients: ~25 us at NVLink 5 rates),
// and because each element is reduced by exactly one rank and then broadcast, all replicas receive the same bits.
//
// The reference has no distributed code.  The buffer must be one symmetric allocation mapped
// through a multicast object on all ranks (torch.distributed._symmetric_memory gives both: `multicast_ptr`); the
// caller brackets the launch with cross-GPU barriers (everyone's contribution written before, every slice delivered
// after) -- gnn_b200/parallel.py uses the symmetric-memory handle's barrier for that.
#include "common.cuh"

namespace gnnb {

__device__ __forceinline__ float4 multimem_ld_reduce_add(const float4* mc_addr) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(mc_addr)
                 : "memory");
    return v;
}
__device__ __forceinline__ void multimem_st(float4* mc_addr, float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                 ::"l"(mc_addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}

// n4 = number of float4 elements of the whole buffer; this rank handles [lo, hi)
__global__ void __launch_bounds__(512)
allreduce_mean_multimem_kernel(float4* __restrict__ mc, int64_t lo, int64_t hi, float inv_world) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    int64_t i = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    // four independent 16-byte reductions in flight per thread: the round trip through the switch is long
    for (; i + 3 * stride < hi; i += 4 * stride) {
        float4 a = multimem_ld_reduce_add(mc + i), b = multimem_ld_reduce_add(mc + i + stride);
        float4 c = multimem_ld_reduce_add(mc + i + 2 * stride), d = multimem_ld_reduce_add(mc + i + 3 * stride);
        a.x *= inv_world; a.y *= inv_world; a.z *= inv_world; a.w *= inv_world;
        b.x *= inv_world; b.y *= inv_world; b.z *= inv_world; b.w *= inv_world;
        c.x *= inv_world; c.y *= inv_world; c.z *= inv_world; c.w *= inv_world;
        d.x *= inv_world; d.y *= inv_world; d.z *= inv_world; d.w *= inv_world;
        multimem_st(mc + i, a);
        multimem_st(mc + i + stride, b);
        multimem_st(mc + i + 2 * stride, c);
        multimem_st(mc + i + 3 * stride, d);
    }
    for (; i < hi; i += stride) {
        float4 a = multimem_ld_reduce_add(mc + i);
        a.x *= inv_world; a.y *= inv_world; a.z *= inv_world; a.w *= inv_world;
        multimem_st(mc + i, a);
    }
}

}  // namespace gnnb

using namespace gnnb;

extern "C" {

int gnnb_allreduce_mean_multimem(void* multicast_ptr, int64_t n_floats, int rank, int world, int max_ctas, void* stream) {
    GNNB_ENTER();
    GNNB_REQUIRE(multicast_ptr && n_floats >= 0 && world >= 1 && rank >= 0 && rank < world, GNNB_E_BADARG, "allreduce_mean_multimem: bad argument");
    GNNB_REQUIRE(n_floats % 4 == 0 && (uintptr_t)multicast_ptr % 16 == 0, GNNB_E_BADARG,
                 "allreduce_mean_multimem: the buffer must be 16-byte aligned and a multiple of 4 floats long");
    const int64_t n4 = n_floats / 4;
    const int64_t per = (n4 + world - 1) / world;
    const int64_t lo = per * rank < n4 ? per * rank : n4, hi = lo + per < n4 ? lo + per : n4;
    if (hi <= lo) return GNNB_OK;
    int64_t blocks = (hi - lo + 512 * 4 - 1) / (512 * 4);
    int64_t cap = max_ctas > 0 ? max_ctas : (int64_t)sm_count();
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    allreduce_mean_multimem_kernel<<<(unsigned)blocks, 512, 0, (cudaStream_t)stream>>>(static_cast<float4*>(multicast_ptr), lo, hi,
                                                                                       1.0f / (float)world);
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

}  // extern "C"
