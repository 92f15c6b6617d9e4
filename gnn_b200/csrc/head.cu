// K8: the dense head of the GCN factories, Linear(h -> out) with a handful of outputs (out = 1 in every shipped
// configuration, gnn/networks/networks.py:273-274; gnn/main.py:438).  torch runs it as four library launches per step
// (gemv forward, gemv + gemm-k1 + split-K reduce backward) plus a reduction for the bias; here it is one kernel each way.
// Both are single passes over the [n, h] activations (HBM / L2 bound, ~6 MB at the C5 subgraph): a warp owns a row, a lane
// a float4 of it; the weight rows live in shared memory.
//   forward : out[r, o] = sum_c A[r, c] W[o, c] + b[o]
//   backward: dA[r, c] = sum_o dOut[r, o] W[o, c];  dW[o, c] = sum_r dOut[r, o] A[r, c];  db[o] = sum_r dOut[r, o]
// dW / db are accumulated per lane over the warp's rows, summed over the CTA's warps in warp order, written as one partial
// per CTA, and the CTA that finishes last adds the partials in CTA order: the grid is a function of (n, device) only, so
// the result is reproducible.
#include "common.cuh"

namespace gnnb {

constexpr int HEAD_WARPS = 8, HEAD_THREADS = HEAD_WARPS * 32;

// Dropout in front of the head (torch.nn.Dropout -> Linear, gnn/networks/networks.py:272-274), applied on the fly: no mask and
// no dropped copy of the activations is ever stored.  Counter-based: element (r, c) of call number `counter` keeps its value
// (times 1 / (1 - p)) iff a 32-bit slice of splitmix64(key ^ position) is >= p * 2^32, key = f(seed, counter).  The counter
// lives on the device (state[1]) and is advanced by the forward kernel itself, so a step replayed as a CUDA graph draws a
// fresh mask on every replay without any host work; the forward pass hands (seed, counter) to its backward pass in `used`.
struct DropArgs {
    unsigned long long* state;      // forward: {seed, counter, ticket, -} (device);  backward: unused
    unsigned long long* used;       // forward: writes {seed, counter};  backward: reads them
    uint32_t thresh;                // p * 2^32
    float scale;                    // 1 / (1 - p)
};
__device__ __forceinline__ unsigned long long drop_key(unsigned long long seed, unsigned long long counter) {
    return splitmix64_dev(seed ^ splitmix64_dev(counter));
}
__device__ __forceinline__ float4 drop_mask4(unsigned long long key, int64_t r, int c, uint32_t thresh, float scale) {
    const unsigned long long g = ((unsigned long long)r << 9) | ((unsigned long long)(c >> 2) << 1);     // c < 1024
    const unsigned long long h0 = splitmix64_dev(key ^ g), h1 = splitmix64_dev(key ^ (g | 1ull));
    return make_float4((uint32_t)h0 >= thresh ? scale : 0.f, (uint32_t)(h0 >> 32) >= thresh ? scale : 0.f,
                       (uint32_t)h1 >= thresh ? scale : 0.f, (uint32_t)(h1 >> 32) >= thresh ? scale : 0.f);
}

template <int O, int T, bool DROP>
__global__ void __launch_bounds__(HEAD_THREADS)
head_fwd_kernel(const float* __restrict__ A, int64_t lda, int64_t n, int h, const float* __restrict__ W, int64_t ldw,
                const float* __restrict__ bias, float* __restrict__ out, int64_t ldo, DropArgs drop) {
    __shared__ __align__(16) float w_s[O * T * 128];
    unsigned long long key = 0, counter = 0;
    if (DROP) {
        counter = *reinterpret_cast<volatile unsigned long long*>(drop.state + 1);     // every CTA reads it before it takes its ticket
        key = drop_key(drop.state[0], counter);
    }
    for (int i = threadIdx.x; i < O * T * 128; i += HEAD_THREADS) {
        const int o = i / (T * 128), c = i % (T * 128);
        w_s[i] = c < h ? W[(int64_t)o * ldw + c] : 0.f;
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool vec = (lda % 4 == 0) && ((reinterpret_cast<uintptr_t>(A) & 15) == 0);
    constexpr int R = 4;                                            // rows in flight per warp (one load each: latency-bound otherwise)
    const int64_t warps = (int64_t)gridDim.x * HEAD_WARPS;
    for (int64_t r0 = ((int64_t)blockIdx.x * HEAD_WARPS + warp) * R; r0 < n; r0 += warps * R) {
        float4 a[R][T];
#pragma unroll
        for (int q = 0; q < R; ++q)
#pragma unroll
            for (int t = 0; t < T; ++t) {
                const int c = t * 128 + lane * 4;
                const int64_t r = r0 + q;
                a[q][t] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (r < n) {
                    const float* row = A + r * lda;
                    if (c + 3 < h && vec) a[q][t] = *reinterpret_cast<const float4*>(row + c);
                    else {
                        if (c < h) a[q][t].x = row[c];
                        if (c + 1 < h) a[q][t].y = row[c + 1];
                        if (c + 2 < h) a[q][t].z = row[c + 2];
                        if (c + 3 < h) a[q][t].w = row[c + 3];
                    }
                }
            }
        if (DROP) {
#pragma unroll
            for (int q = 0; q < R; ++q)
#pragma unroll
                for (int t = 0; t < T; ++t) {
                    const float4 m = drop_mask4(key, r0 + q, t * 128 + lane * 4, drop.thresh, drop.scale);
                    a[q][t].x *= m.x; a[q][t].y *= m.y; a[q][t].z *= m.z; a[q][t].w *= m.w;
                }
        }
#pragma unroll
        for (int q = 0; q < R; ++q) {
            float acc[O];
#pragma unroll
            for (int o = 0; o < O; ++o) acc[o] = 0.f;
#pragma unroll
            for (int t = 0; t < T; ++t)
#pragma unroll
                for (int o = 0; o < O; ++o) {
                    const float4 w = *reinterpret_cast<const float4*>(w_s + (o * T + t) * 128 + lane * 4);
                    acc[o] = fmaf(a[q][t].x, w.x, fmaf(a[q][t].y, w.y, fmaf(a[q][t].z, w.z, fmaf(a[q][t].w, w.w, acc[o]))));
                }
#pragma unroll
            for (int o = 0; o < O; ++o)
                for (int sft = 16; sft > 0; sft >>= 1) acc[o] += __shfl_xor_sync(0xffffffffu, acc[o], sft);
            if (lane == 0 && r0 + q < n) {
#pragma unroll
                for (int o = 0; o < O; ++o) out[(r0 + q) * ldo + o] = acc[o] + (bias ? bias[o] : 0.f);
            }
        }
    }
    if (DROP) {
        // the CTA that takes the last ticket knows every CTA has read the counter: it advances it for the next call
        __syncthreads();
        if (threadIdx.x == 0) {
            if (blockIdx.x == 0) { drop.used[0] = drop.state[0]; drop.used[1] = counter; }
            unsigned* ticket = reinterpret_cast<unsigned*>(drop.state + 2);
            if (atomicAdd(ticket, 1u) == gridDim.x - 1) {
                *ticket = 0;
                drop.state[1] = counter + 1;
            }
        }
    }
}

// partial layout per CTA: [O][T * 128] weight-gradient columns, then [O] bias gradients
template <int O, int T, bool DROP>
__global__ void __launch_bounds__(HEAD_THREADS)
head_bwd_kernel(const float* __restrict__ dOut, int64_t lddo, const float* __restrict__ A, int64_t lda, int64_t n, int h,
                const float* __restrict__ W, int64_t ldw, float* __restrict__ dA, int64_t ldda,
                float* __restrict__ dW, int64_t lddw, float* __restrict__ db, float* __restrict__ partial, unsigned* __restrict__ ticket,
                DropArgs drop) {
    const unsigned long long key = DROP ? drop_key(drop.used[0], drop.used[1]) : 0ull;
    constexpr int PER = O * T * 128 + O;
    constexpr int PER_S = (PER + 3) / 4 * 4;                        // rows of `red` stay 16-byte aligned for the float4 stores
    __shared__ __align__(16) float w_s[O * T * 128];
    __shared__ __align__(16) float red[HEAD_WARPS][PER_S];
    __shared__ bool last;
    for (int i = threadIdx.x; i < O * T * 128; i += HEAD_THREADS) {
        const int o = i / (T * 128), c = i % (T * 128);
        w_s[i] = c < h ? W[(int64_t)o * ldw + c] : 0.f;
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool vec_a = (lda % 4 == 0) && ((reinterpret_cast<uintptr_t>(A) & 15) == 0);
    const bool vec_d = dA && (ldda % 4 == 0) && ((reinterpret_cast<uintptr_t>(dA) & 15) == 0);
    float4 gw[O * T];
    float gb[O];
#pragma unroll
    for (int i = 0; i < O * T; ++i) gw[i] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int o = 0; o < O; ++o) gb[o] = 0.f;
    constexpr int R = 2;                                            // rows in flight per warp
    const int64_t warps = (int64_t)gridDim.x * HEAD_WARPS;
    for (int64_t r0 = ((int64_t)blockIdx.x * HEAD_WARPS + warp) * R; r0 < n; r0 += warps * R) {
        float g[R][O];
        float4 a[R][T];
#pragma unroll
        for (int q = 0; q < R; ++q) {
            const int64_t r = r0 + q;
#pragma unroll
            for (int o = 0; o < O; ++o) g[q][o] = r < n ? dOut[r * lddo + o] : 0.f;
#pragma unroll
            for (int t = 0; t < T; ++t) {
                const int c = t * 128 + lane * 4;
                a[q][t] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (r < n) {
                    const float* row = A + r * lda;
                    if (c + 3 < h && vec_a) a[q][t] = *reinterpret_cast<const float4*>(row + c);
                    else {
                        if (c < h) a[q][t].x = row[c];
                        if (c + 1 < h) a[q][t].y = row[c + 1];
                        if (c + 2 < h) a[q][t].z = row[c + 2];
                        if (c + 3 < h) a[q][t].w = row[c + 3];
                    }
                }
            }
        }
#pragma unroll
        for (int q = 0; q < R; ++q) {
            const int64_t r = r0 + q;
#pragma unroll
            for (int o = 0; o < O; ++o) gb[o] += g[q][o];
#pragma unroll
            for (int t = 0; t < T; ++t) {
                const int c = t * 128 + lane * 4;
                float4 d = make_float4(0.f, 0.f, 0.f, 0.f);
                float4 m = make_float4(1.f, 1.f, 1.f, 1.f);
                if (DROP) {                                         // the forward pass saw A * m: so does dW, and dA is masked the same way
                    m = drop_mask4(key, r, c, drop.thresh, drop.scale);
                    a[q][t].x *= m.x; a[q][t].y *= m.y; a[q][t].z *= m.z; a[q][t].w *= m.w;
                }
#pragma unroll
                for (int o = 0; o < O; ++o) {
                    const float4 w = *reinterpret_cast<const float4*>(w_s + (o * T + t) * 128 + lane * 4);
                    const float go = g[q][o];
                    d.x = fmaf(go, w.x, d.x); d.y = fmaf(go, w.y, d.y); d.z = fmaf(go, w.z, d.z); d.w = fmaf(go, w.w, d.w);
                    float4& acc = gw[o * T + t];
                    acc.x = fmaf(go, a[q][t].x, acc.x); acc.y = fmaf(go, a[q][t].y, acc.y);
                    acc.z = fmaf(go, a[q][t].z, acc.z); acc.w = fmaf(go, a[q][t].w, acc.w);
                }
                if (DROP) { d.x *= m.x; d.y *= m.y; d.z *= m.z; d.w *= m.w; }
                if (dA && r < n) {
                    float* drow = dA + r * ldda;
                    if (c + 3 < h && vec_d) *reinterpret_cast<float4*>(drow + c) = d;
                    else {
                        if (c < h) drow[c] = d.x;
                        if (c + 1 < h) drow[c + 1] = d.y;
                        if (c + 2 < h) drow[c + 2] = d.z;
                        if (c + 3 < h) drow[c + 3] = d.w;
                    }
                }
            }
        }
    }
    // warp partials -> shared memory -> one partial per CTA (warps added in warp order)
#pragma unroll
    for (int i = 0; i < O * T; ++i) *reinterpret_cast<float4*>(&red[warp][i * 128 + lane * 4]) = gw[i];
    if (lane == 0) {
#pragma unroll
        for (int o = 0; o < O; ++o) red[warp][O * T * 128 + o] = gb[o];
    }
    __syncthreads();
    float* mine = partial + (int64_t)blockIdx.x * PER;
    for (int i = threadIdx.x; i < PER; i += HEAD_THREADS) {
        float s = red[0][i];
#pragma unroll
        for (int w = 1; w < HEAD_WARPS; ++w) s += red[w][i];
        mine[i] = s;
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!last) return;
    __threadfence();
    // the last CTA: thread = (output element, half of the CTAs); consecutive threads read consecutive floats of a partial
    // (coalesced: one SM must pull every partial, so the number of L2 sectors it requests is what this tail costs); each
    // half is added in CTA order, then the two halves in half order -- a fixed tree
    float* halves = &red[0][0];                                    // [2][128] of the staging area, free again by now
    for (int base = 0; base < PER; base += 128) {
        const int i = base + (threadIdx.x & 127), half = threadIdx.x >> 7;
        float sacc = 0.f;
        if (i < PER) {
#pragma unroll 16
            for (unsigned b = half; b < gridDim.x; b += 2) sacc += __ldcg(partial + (int64_t)b * PER + i);      // 16 loads in flight
        }
        __syncthreads();
        halves[half * 128 + (threadIdx.x & 127)] = sacc;
        __syncthreads();
        if (half == 0 && i < PER) {
            const float total = halves[threadIdx.x & 127] + halves[128 + (threadIdx.x & 127)];
            if (i < O * T * 128) {
                const int o = i / (T * 128), c = i % (T * 128);
                if (c < h) dW[(int64_t)o * lddw + c] = total;
            } else if (db) {
                db[i - O * T * 128] = total;
            }
        }
    }
    if (threadIdx.x == 0) *ticket = 0;                             // ready for the next call
}

static int head_grid(int64_t n) {
    int64_t blocks = (n + HEAD_WARPS * 4 - 1) / (HEAD_WARPS * 4);      // ~4 rows per warp at least
    const int64_t cap = (int64_t)sm_count() * 4;                        // 32 warps per SM: the row loop is latency-bound
    if (blocks > cap) blocks = cap;
    return (int)(blocks < 1 ? 1 : blocks);
}
static int head_t(int64_t h) { return h <= 128 ? 1 : h <= 256 ? 2 : h <= 512 ? 4 : h <= 1024 ? 8 : 0; }
static bool head_ok(int64_t h, int64_t o) { const int t = head_t(h); return t > 0 && o >= 1 && o <= 4 && o * t <= 8; }

#define HEAD_DISPATCH(O_, T_, CALL)                                      \
    if (o == O_ && t == T_) { CALL(O_, T_); launched = true; }
#define HEAD_ALL(CALL)                                                   \
    HEAD_DISPATCH(1, 1, CALL) HEAD_DISPATCH(1, 2, CALL) HEAD_DISPATCH(1, 4, CALL) HEAD_DISPATCH(1, 8, CALL) \
    HEAD_DISPATCH(2, 1, CALL) HEAD_DISPATCH(2, 2, CALL) HEAD_DISPATCH(2, 4, CALL)                             \
    HEAD_DISPATCH(3, 1, CALL) HEAD_DISPATCH(3, 2, CALL) HEAD_DISPATCH(4, 1, CALL) HEAD_DISPATCH(4, 2, CALL)

}  // namespace gnnb

using namespace gnnb;

extern "C" {

int gnnb_head_supported(int64_t h, int64_t out) { return head_ok(h, out) ? 1 : 0; }

int gnnb_head_workspace_bytes(int64_t n, int64_t h, int64_t out, size_t* bytes) {
    GNNB_REQUIRE(bytes && n >= 0 && head_ok(h, out), GNNB_E_BADARG, "head_workspace_bytes: bad argument (out <= 4, h <= 1024, out * ceil(h / 128) <= 8)");
    const int t = head_t(h);
    *bytes = 256 + (size_t)sm_count() * 4 * (size_t)(out * t * 128 + out) * sizeof(float);
    return GNNB_OK;
}

static int drop_args(double p, unsigned long long* state, unsigned long long* used, DropArgs* d, const char* who) {
    GNNB_REQUIRE(p >= 0.0 && p < 1.0 && used && (((uintptr_t)used | (uintptr_t)state) & 7) == 0, GNNB_E_BADARG,
                 "%s: 0 <= p < 1 and 8-byte aligned device pointers for the generator state", who);
    d->state = state;
    d->used = used;
    d->thresh = (uint32_t)(p * 4294967296.0);
    d->scale = (float)(1.0 / (1.0 - p));
    return GNNB_OK;
}

static int head_fwd_impl(const float* A, int64_t lda, int64_t n, int64_t h, const float* W, int64_t ldw, int64_t out_features,
                         const float* bias, float* out, int64_t ldo, const DropArgs* drop, void* stream) {
    GNNB_REQUIRE(A && W && out && n >= 0 && lda >= h && ldw >= h && ldo >= out_features, GNNB_E_BADARG, "head_fwd: bad argument");
    GNNB_REQUIRE(head_ok(h, out_features), GNNB_E_UNSUPPORTED, "head_fwd: out <= 4, h <= 1024 and out * ceil(h / 128) <= 8 (got h = %lld, out = %lld)",
                 (long long)h, (long long)out_features);
    if (n == 0) return GNNB_OK;
    const int o = (int)out_features, t = head_t(h), grid = head_grid(n);
    bool launched = false;
    const DropArgs none = {nullptr, nullptr, 0u, 1.f};
#define CALL_FWD(O_, T_)                                                                                                              \
    if (drop) head_fwd_kernel<O_, T_, true><<<grid, HEAD_THREADS, 0, (cudaStream_t)stream>>>(A, lda, n, (int)h, W, ldw, bias, out, ldo, *drop); \
    else head_fwd_kernel<O_, T_, false><<<grid, HEAD_THREADS, 0, (cudaStream_t)stream>>>(A, lda, n, (int)h, W, ldw, bias, out, ldo, none)
    HEAD_ALL(CALL_FWD)
#undef CALL_FWD
    GNNB_REQUIRE(launched, GNNB_E_UNSUPPORTED, "head_fwd: shape not instantiated");
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

static int head_bwd_impl(const float* dOut, int64_t lddo, const float* A, int64_t lda, int64_t n, int64_t h, const float* W, int64_t ldw,
                         int64_t out_features, float* dA, int64_t ldda, float* dW, int64_t lddw, float* db, const DropArgs* drop,
                         void* workspace, size_t workspace_bytes, void* stream) {
    GNNB_REQUIRE(dOut && A && W && dW && n >= 0 && lda >= h && ldw >= h && lddw >= h && lddo >= out_features && (!dA || ldda >= h), GNNB_E_BADARG,
                 "head_bwd: bad argument");
    GNNB_REQUIRE(head_ok(h, out_features), GNNB_E_UNSUPPORTED, "head_bwd: out <= 4, h <= 1024 and out * ceil(h / 128) <= 8 (got h = %lld, out = %lld)",
                 (long long)h, (long long)out_features);
    size_t need = 0;
    int rc = gnnb_head_workspace_bytes(n, h, out_features, &need);
    if (rc != GNNB_OK) return rc;
    GNNB_REQUIRE(workspace && workspace_bytes >= need, GNNB_E_WORKSPACE, "head_bwd: workspace %zu < %zu bytes", workspace_bytes, need);
    unsigned* ticket = reinterpret_cast<unsigned*>(align_up((size_t)(uintptr_t)workspace, 256));
    float* partial = reinterpret_cast<float*>(reinterpret_cast<char*>(ticket) + 128);
    GNNB_CUDA(cudaMemsetAsync(ticket, 0, sizeof(unsigned), (cudaStream_t)stream));
    const int o = (int)out_features, t = head_t(h), grid = n == 0 ? 1 : head_grid(n);
    bool launched = false;
    const DropArgs none = {nullptr, nullptr, 0u, 1.f};
#define CALL_BWD(O_, T_)                                                                                                     \
    if (drop) head_bwd_kernel<O_, T_, true><<<grid, HEAD_THREADS, 0, (cudaStream_t)stream>>>(dOut, lddo, A, lda, n, (int)h, W, ldw, dA, ldda, dW, lddw, db, partial, ticket, *drop); \
    else head_bwd_kernel<O_, T_, false><<<grid, HEAD_THREADS, 0, (cudaStream_t)stream>>>(dOut, lddo, A, lda, n, (int)h, W, ldw, dA, ldda, dW, lddw, db, partial, ticket, none)
    HEAD_ALL(CALL_BWD)
#undef CALL_BWD
    GNNB_REQUIRE(launched, GNNB_E_UNSUPPORTED, "head_bwd: shape not instantiated");
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

int gnnb_head_fwd(const float* A, int64_t lda, int64_t n, int64_t h, const float* W, int64_t ldw, int64_t out_features,
                  const float* bias, float* out, int64_t ldo, void* stream) {
    GNNB_ENTER();
    return head_fwd_impl(A, lda, n, h, W, ldw, out_features, bias, out, ldo, nullptr, stream);
}

int gnnb_head_bwd(const float* dOut, int64_t lddo, const float* A, int64_t lda, int64_t n, int64_t h, const float* W, int64_t ldw,
                  int64_t out_features, float* dA, int64_t ldda, float* dW, int64_t lddw, float* db,
                  void* workspace, size_t workspace_bytes, void* stream) {
    GNNB_ENTER();
    return head_bwd_impl(dOut, lddo, A, lda, n, h, W, ldw, out_features, dA, ldda, dW, lddw, db, nullptr, workspace, workspace_bytes, stream);
}

int gnnb_head_dropout_fwd(const float* A, int64_t lda, int64_t n, int64_t h, const float* W, int64_t ldw, int64_t out_features,
                          const float* bias, float* out, int64_t ldo, double p, uint64_t* state, uint64_t* used, void* stream) {
    GNNB_ENTER();
    GNNB_REQUIRE(state, GNNB_E_BADARG, "head_dropout_fwd: generator state missing");
    DropArgs d;
    int rc = drop_args(p, (unsigned long long*)state, (unsigned long long*)used, &d, "head_dropout_fwd");
    if (rc != GNNB_OK) return rc;
    return head_fwd_impl(A, lda, n, h, W, ldw, out_features, bias, out, ldo, &d, stream);
}

int gnnb_head_dropout_bwd(const float* dOut, int64_t lddo, const float* A, int64_t lda, int64_t n, int64_t h, const float* W, int64_t ldw,
                          int64_t out_features, float* dA, int64_t ldda, float* dW, int64_t lddw, float* db, double p,
                          const uint64_t* used, void* workspace, size_t workspace_bytes, void* stream) {
    GNNB_ENTER();
    DropArgs d;
    int rc = drop_args(p, nullptr, (unsigned long long*)used, &d, "head_dropout_bwd");
    if (rc != GNNB_OK) return rc;
    return head_bwd_impl(dOut, lddo, A, lda, n, h, W, ldw, out_features, dA, ldda, dW, lddw, db, &d, workspace, workspace_bytes, stream);
}

}  // extern "C"
