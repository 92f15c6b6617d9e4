// K6/K7: neighbourhood aggregation  out[i,:] = sum_e val[e] * H[col[e],:] (+ bias)
//
// Replaces MessagePassing.propagate + torch_scatter.scatter(reduce='add') inside PyG's GCNConv (forward)
// and the autograd of the gather (backward, over the by-source CSR).  Call sites in the reference:
// gnn/networks/networks.py:271-272 and the other GCN factories.
//
// HBM/L2-bound: one warp per output row; a lane owns one float4 of the feature row, so a neighbour row
// is one coalesced 128-bit-per-lane request; neighbour indices and weights are read 32 at a time with
// one coalesced load and handed round with shuffles; GATHER_UNROLL independent row gathers are kept in
// flight per warp.  Summation order is the CSR order (edge order, self loop last) => deterministic.
#include "common.cuh"

namespace gnnb {

#ifndef GATHER_UNROLL_
#define GATHER_UNROLL_ 8
#endif
#ifndef LONG_ROW_
#define LONG_ROW_ 64
#endif
constexpr int GATHER_UNROLL = GATHER_UNROLL_;
constexpr int LONG_ROW = LONG_ROW_;       // rows with more entries than this (hubs of the kNN graph) get a whole CTA

__device__ __forceinline__ float4 ld_row4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }

// warp-per-row part (rows of at most LONG_ROW entries); `block` / `nblocks` index the CTAs that play this role
template <bool HAS_BIAS>
__device__ __forceinline__ void spmm_rows_vec4(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col,
                                               const float* __restrict__ val, int64_t n, const float* __restrict__ H, int64_t ldh, int F,
                                               const float* __restrict__ bias, float* __restrict__ out, int64_t ldo,
                                               int64_t block, int64_t nblocks) {
    const int lane = threadIdx.x & 31;
    const int64_t warp0 = (block * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (nblocks * blockDim.x) >> 5;
    for (int64_t row = warp0; row < n; row += nwarps) {
        const int start = rowptr[row], end = rowptr[row + 1];
        if (end - start > LONG_ROW) continue;                   // spmm_long_rows_kernel owns this row
        for (int fc = 0; fc < F; fc += 128) {
            const int f = fc + lane * 4;
            const bool active = f < F;
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int base = start; base < end; base += 32) {
                const int e = base + lane;
                int c = 0;
                float v = 0.f;
                if (e < end) {
                    c = __ldg(col + e);
                    v = __ldg(val + e);
                }
                const int cnt = min(32, end - base);
                int t = 0;
                for (; t + GATHER_UNROLL <= cnt; t += GATHER_UNROLL) {
                    float4 h[GATHER_UNROLL];
                    float w[GATHER_UNROLL];
#pragma unroll
                    for (int u = 0; u < GATHER_UNROLL; ++u) {
                        int cu = __shfl_sync(0xffffffffu, c, t + u);
                        w[u] = __shfl_sync(0xffffffffu, v, t + u);
                        h[u] = active ? ld_row4(H + (int64_t)cu * ldh + f) : make_float4(0.f, 0.f, 0.f, 0.f);
                    }
#pragma unroll
                    for (int u = 0; u < GATHER_UNROLL; ++u) {
                        acc.x = fmaf(w[u], h[u].x, acc.x);
                        acc.y = fmaf(w[u], h[u].y, acc.y);
                        acc.z = fmaf(w[u], h[u].z, acc.z);
                        acc.w = fmaf(w[u], h[u].w, acc.w);
                    }
                }
                for (; t < cnt; ++t) {
                    int cu = __shfl_sync(0xffffffffu, c, t);
                    float wu = __shfl_sync(0xffffffffu, v, t);
                    if (active) {
                        float4 hu = ld_row4(H + (int64_t)cu * ldh + f);
                        acc.x = fmaf(wu, hu.x, acc.x);
                        acc.y = fmaf(wu, hu.y, acc.y);
                        acc.z = fmaf(wu, hu.z, acc.z);
                        acc.w = fmaf(wu, hu.w, acc.w);
                    }
                }
            }
            if (active) {
                if (HAS_BIAS) {
                    float4 b = ld_row4(bias + f);
                    acc.x += b.x; acc.y += b.y; acc.z += b.z; acc.w += b.w;
                }
                *reinterpret_cast<float4*>(out + row * ldo + f) = acc;
            }
        }
    }
}

// any F / any alignment: a lane owns features lane, lane + 32, ...
template <bool HAS_BIAS>
__global__ void __launch_bounds__(256)
spmm_scalar_kernel(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col, const float* __restrict__ val, int64_t n,
                   const float* __restrict__ H, int64_t ldh, int F, const float* __restrict__ bias,
                   float* __restrict__ out, int64_t ldo) {
    const int lane = threadIdx.x & 31;
    const int64_t warp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t row = warp0; row < n; row += nwarps) {
        const int start = rowptr[row], end = rowptr[row + 1];
        if (end - start > LONG_ROW) continue;
        for (int fc = 0; fc < F; fc += 32) {
            const int f = fc + lane;
            const bool active = f < F;
            float acc = 0.f;
            for (int base = start; base < end; base += 32) {
                const int e = base + lane;
                int c = 0;
                float v = 0.f;
                if (e < end) {
                    c = __ldg(col + e);
                    v = __ldg(val + e);
                }
                const int cnt = min(32, end - base);
                for (int t = 0; t < cnt; ++t) {
                    int cu = __shfl_sync(0xffffffffu, c, t);
                    float wu = __shfl_sync(0xffffffffu, v, t);
                    if (active) acc = fmaf(wu, __ldg(H + (int64_t)cu * ldh + f), acc);
                }
            }
            if (active) out[row * ldo + f] = acc + (HAS_BIAS ? bias[f] : 0.f);
        }
    }
}

// Hub rows: the kNN graphs are heavy-tailed (median degree 35, maximum in the thousands), so one warp per row leaves
// the whole GPU waiting for a handful of warps.  Rows longer than LONG_ROW are handled here by a whole CTA: the 8 warps
// take the row's 32-entry chunks round-robin, and the 8 partial sums are added in warp order (fixed => deterministic).
// CTAs scan blocks of 64 rows for candidates, so no list of long rows has to be kept with the graph (the order in
// which a CTA visits its candidates does not matter: rows are independent).
template <bool HAS_BIAS, bool VEC, int NW>
__device__ __forceinline__ void spmm_long_rows(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col,
                                               const float* __restrict__ val, int64_t n, const float* __restrict__ H, int64_t ldh, int F,
                                               const float* __restrict__ bias, float* __restrict__ out, int64_t ldo,
                                               int64_t block, int64_t nblocks) {
    constexpr int SCAN = 64;
    __shared__ int s_rows[SCAN];
    __shared__ int s_count;
    __shared__ float4 s_part[NW][32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int64_t base_row = block * SCAN; base_row < n; base_row += nblocks * SCAN) {
        if (threadIdx.x == 0) s_count = 0;
        __syncthreads();
        const int64_t r = base_row + threadIdx.x;
        if (threadIdx.x < SCAN && r < n && rowptr[r + 1] - rowptr[r] > LONG_ROW) s_rows[atomicAdd(&s_count, 1)] = (int)threadIdx.x;
        __syncthreads();
        const int count = s_count;
        for (int q = 0; q < count; ++q) {
            const int64_t row = base_row + s_rows[q];
            const int start = rowptr[row], end = rowptr[row + 1];
            for (int fc = 0; fc < F; fc += (VEC ? 128 : 32)) {
                const int f = fc + (VEC ? lane * 4 : lane);
                const bool active = f < F;
                float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
                for (int base = start + warp * 32; base < end; base += NW * 32) {
                    const int e = base + lane;
                    int c = 0;
                    float v = 0.f;
                    if (e < end) {
                        c = __ldg(col + e);
                        v = __ldg(val + e);
                    }
                    const int cnt = min(32, end - base);
                    int t = 0;
                    if (VEC) {
                        for (; t + GATHER_UNROLL <= cnt; t += GATHER_UNROLL) {
                            float4 h[GATHER_UNROLL];
                            float w[GATHER_UNROLL];
#pragma unroll
                            for (int u = 0; u < GATHER_UNROLL; ++u) {
                                int cu = __shfl_sync(0xffffffffu, c, t + u);
                                w[u] = __shfl_sync(0xffffffffu, v, t + u);
                                h[u] = active ? ld_row4(H + (int64_t)cu * ldh + f) : make_float4(0.f, 0.f, 0.f, 0.f);
                            }
#pragma unroll
                            for (int u = 0; u < GATHER_UNROLL; ++u) {
                                acc.x = fmaf(w[u], h[u].x, acc.x);
                                acc.y = fmaf(w[u], h[u].y, acc.y);
                                acc.z = fmaf(w[u], h[u].z, acc.z);
                                acc.w = fmaf(w[u], h[u].w, acc.w);
                            }
                        }
                    }
                    for (; t < cnt; ++t) {
                        int cu = __shfl_sync(0xffffffffu, c, t);
                        float wu = __shfl_sync(0xffffffffu, v, t);
                        if (active) {
                            if (VEC) {
                                float4 hu = ld_row4(H + (int64_t)cu * ldh + f);
                                acc.x = fmaf(wu, hu.x, acc.x);
                                acc.y = fmaf(wu, hu.y, acc.y);
                                acc.z = fmaf(wu, hu.z, acc.z);
                                acc.w = fmaf(wu, hu.w, acc.w);
                            } else {
                                acc.x = fmaf(wu, __ldg(H + (int64_t)cu * ldh + f), acc.x);
                            }
                        }
                    }
                }
                s_part[warp][lane] = acc;
                __syncthreads();
                if (warp == 0 && active) {
                    float4 tot = s_part[0][lane];
#pragma unroll
                    for (int w2 = 1; w2 < NW; ++w2) {
                        tot.x += s_part[w2][lane].x; tot.y += s_part[w2][lane].y;
                        tot.z += s_part[w2][lane].z; tot.w += s_part[w2][lane].w;
                    }
                    if (VEC) {
                        if (HAS_BIAS) {
                            float4 b = ld_row4(bias + f);
                            tot.x += b.x; tot.y += b.y; tot.z += b.z; tot.w += b.w;
                        }
                        *reinterpret_cast<float4*>(out + row * ldo + f) = tot;
                    } else {
                        out[row * ldo + f] = tot.x + (HAS_BIAS ? bias[f] : 0.f);
                    }
                }
                __syncthreads();
            }
        }
        __syncthreads();
    }
}

// Hub rows split into tasks of LONG_ROW entries (the plan is made once per graph by the caller): one warp per task
// writes its partial sum to scratch; the warp that completes a row last adds the partials in task order, so the result
// does not depend on which warp that is.
template <bool HAS_BIAS>
__device__ __forceinline__ void spmm_hub_tasks(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col,
                                               const float* __restrict__ val, const float* __restrict__ H, int64_t ldh, int F,
                                               const float* __restrict__ bias, float* __restrict__ out, int64_t ldo,
                                               const int32_t* __restrict__ task_row, const int32_t* __restrict__ task_chunk,
                                               const int32_t* __restrict__ task_first, const int32_t* __restrict__ task_count,
                                               int64_t n_tasks, float* __restrict__ partial, int* __restrict__ counters,
                                               int64_t block, int64_t nblocks) {
    const int lane = threadIdx.x & 31;
    const int64_t warp0 = (block * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (nblocks * blockDim.x) >> 5;
    for (int64_t t = warp0; t < n_tasks; t += nwarps) {
        const int row = task_row[t];
        const int start = rowptr[row] + task_chunk[t] * LONG_ROW;
        const int end = min(start + LONG_ROW, rowptr[row + 1]);
        for (int fc = 0; fc < F; fc += 128) {
            const int f = fc + lane * 4;
            const bool active = f < F;
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int base = start; base < end; base += 32) {
                const int e = base + lane;
                int c = 0;
                float v = 0.f;
                if (e < end) {
                    c = __ldg(col + e);
                    v = __ldg(val + e);
                }
                const int cnt = min(32, end - base);
                int q = 0;
                for (; q + GATHER_UNROLL <= cnt; q += GATHER_UNROLL) {
                    float4 h[GATHER_UNROLL];
                    float w[GATHER_UNROLL];
#pragma unroll
                    for (int u = 0; u < GATHER_UNROLL; ++u) {
                        int cu = __shfl_sync(0xffffffffu, c, q + u);
                        w[u] = __shfl_sync(0xffffffffu, v, q + u);
                        h[u] = active ? ld_row4(H + (int64_t)cu * ldh + f) : make_float4(0.f, 0.f, 0.f, 0.f);
                    }
#pragma unroll
                    for (int u = 0; u < GATHER_UNROLL; ++u) {
                        acc.x = fmaf(w[u], h[u].x, acc.x);
                        acc.y = fmaf(w[u], h[u].y, acc.y);
                        acc.z = fmaf(w[u], h[u].z, acc.z);
                        acc.w = fmaf(w[u], h[u].w, acc.w);
                    }
                }
                for (; q < cnt; ++q) {
                    int cu = __shfl_sync(0xffffffffu, c, q);
                    float wu = __shfl_sync(0xffffffffu, v, q);
                    if (active) {
                        float4 hu = ld_row4(H + (int64_t)cu * ldh + f);
                        acc.x = fmaf(wu, hu.x, acc.x);
                        acc.y = fmaf(wu, hu.y, acc.y);
                        acc.z = fmaf(wu, hu.z, acc.z);
                        acc.w = fmaf(wu, hu.w, acc.w);
                    }
                }
            }
            if (active) *reinterpret_cast<float4*>(partial + t * F + f) = acc;
        }
        __threadfence();
        __syncwarp();
        const int first = task_first[t], count = task_count[t];
        int prev = 0;
        if (lane == 0) prev = atomicAdd(&counters[first], 1);
        prev = __shfl_sync(0xffffffffu, prev, 0);
        if (prev == count - 1) {                   // this warp completed the row: add the partials in task order
            __threadfence();
            for (int fc = 0; fc < F; fc += 128) {
                const int f = fc + lane * 4;
                if (f < F) {
                    float4 tot = __ldcg(reinterpret_cast<const float4*>(partial + (int64_t)first * F + f));
                    for (int q = 1; q < count; ++q) {
                        float4 pq = __ldcg(reinterpret_cast<const float4*>(partial + (int64_t)(first + q) * F + f));
                        tot.x += pq.x; tot.y += pq.y; tot.z += pq.z; tot.w += pq.w;
                    }
                    if (HAS_BIAS) {
                        float4 b = ld_row4(bias + f);
                        tot.x += b.x; tot.y += b.y; tot.z += b.z; tot.w += b.w;
                    }
                    *reinterpret_cast<float4*>(out + (int64_t)row * ldo + f) = tot;
                }
            }
        }
    }
}

template <bool HAS_BIAS>
__global__ void __launch_bounds__(256, 4)
spmm_planned_vec4_kernel(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col, const float* __restrict__ val, int64_t n,
                         const float* __restrict__ H, int64_t ldh, int F, const float* __restrict__ bias,
                         float* __restrict__ out, int64_t ldo, int hub_blocks,
                         const int32_t* __restrict__ task_row, const int32_t* __restrict__ task_chunk,
                         const int32_t* __restrict__ task_first, const int32_t* __restrict__ task_count, int64_t n_tasks,
                         float* __restrict__ partial, int* __restrict__ counters) {
    if ((int)blockIdx.x < hub_blocks)
        spmm_hub_tasks<HAS_BIAS>(rowptr, col, val, H, ldh, F, bias, out, ldo, task_row, task_chunk, task_first, task_count, n_tasks,
                                 partial, counters, blockIdx.x, hub_blocks);
    else
        spmm_rows_vec4<HAS_BIAS>(rowptr, col, val, n, H, ldh, F, bias, out, ldo, (int64_t)blockIdx.x - hub_blocks,
                                 (int64_t)gridDim.x - hub_blocks);
}

// One launch for both roles: the first `long_blocks` CTAs scan for hub rows (they start first, which is what the tail
// wants), the rest run one warp per ordinary row.
template <bool HAS_BIAS>
__global__ void __launch_bounds__(256, 4)        // <= 64 registers: 32 resident warps per SM for the gather latency
spmm_fused_vec4_kernel(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col, const float* __restrict__ val, int64_t n,
                       const float* __restrict__ H, int64_t ldh, int F, const float* __restrict__ bias,
                       float* __restrict__ out, int64_t ldo, int long_blocks) {
    if ((int)blockIdx.x < long_blocks)
        spmm_long_rows<HAS_BIAS, true, 8>(rowptr, col, val, n, H, ldh, F, bias, out, ldo, blockIdx.x, long_blocks);
    else
        spmm_rows_vec4<HAS_BIAS>(rowptr, col, val, n, H, ldh, F, bias, out, ldo, (int64_t)blockIdx.x - long_blocks,
                                 (int64_t)gridDim.x - long_blocks);
}

template <bool HAS_BIAS>
__global__ void __launch_bounds__(256)
spmm_long_rows_scalar_kernel(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col, const float* __restrict__ val, int64_t n,
                             const float* __restrict__ H, int64_t ldh, int F, const float* __restrict__ bias,
                             float* __restrict__ out, int64_t ldo) {
    spmm_long_rows<HAS_BIAS, false, 8>(rowptr, col, val, n, H, ldh, F, bias, out, ldo, blockIdx.x, gridDim.x);
}

}  // namespace gnnb

using namespace gnnb;

extern "C" int gnnb_spmm_long_row(void) { return LONG_ROW; }

extern "C" int gnnb_spmm_workspace_bytes(int64_t n_hub_tasks, int64_t F, size_t* bytes) {
    GNNB_REQUIRE(bytes && n_hub_tasks >= 0 && F > 0, GNNB_E_BADARG, "spmm_workspace_bytes: bad argument");
    *bytes = align_up((size_t)n_hub_tasks * sizeof(int), 256) + (size_t)n_hub_tasks * (size_t)F * sizeof(float) + 256;
    return GNNB_OK;
}

extern "C" int gnnb_spmm(const int32_t* rowptr, const int32_t* col, const float* val, int64_t n,
                         const float* H, int64_t ldh, int64_t F, const float* bias, float* out, int64_t ldo,
                         const int32_t* hub_task_row, const int32_t* hub_task_chunk, const int32_t* hub_task_first,
                         const int32_t* hub_task_count, int64_t n_hub_tasks, void* workspace, size_t workspace_bytes,
                         void* stream_) {
    GNNB_ENTER();
    cudaStream_t stream = (cudaStream_t)stream_;
    GNNB_REQUIRE(rowptr && col && val && H && out, GNNB_E_BADARG, "spmm: null pointer");
    GNNB_REQUIRE(n >= 0 && F > 0 && F <= 4096 && ldh >= F && ldo >= F, GNNB_E_BADARG, "spmm: bad shape (F must be 1..4096)");
    GNNB_REQUIRE(H != out, GNNB_E_BADARG, "spmm: in-place aggregation is not supported");
    if (n == 0) return GNNB_OK;
    bool vec = (F % 4 == 0) && (ldh % 4 == 0) && (ldo % 4 == 0) && ((uintptr_t)H % 16 == 0) && ((uintptr_t)out % 16 == 0) &&
               (!bias || (uintptr_t)bias % 16 == 0);
    const int64_t sms = sm_count();
    if (vec && n_hub_tasks > 0 && hub_task_row) {
        // planned path: hub rows are split into tasks shared by many warps
        GNNB_REQUIRE(hub_task_chunk && hub_task_first && hub_task_count, GNNB_E_BADARG, "spmm: incomplete hub plan");
        size_t need = 0;
        gnnb_spmm_workspace_bytes(n_hub_tasks, F, &need);
        GNNB_REQUIRE(workspace && workspace_bytes >= need, GNNB_E_WORKSPACE, "spmm: workspace %zu < %zu bytes", workspace_bytes, need);
        int* counters = static_cast<int*>(workspace);
        float* partial = reinterpret_cast<float*>(static_cast<char*>(workspace) + align_up((size_t)n_hub_tasks * sizeof(int), 256));
        GNNB_CUDA(cudaMemsetAsync(counters, 0, (size_t)n_hub_tasks * sizeof(int), stream));
        int64_t hub_blocks = (n_hub_tasks + 7) / 8;
        if (hub_blocks > sms * 8) hub_blocks = sms * 8;
        int64_t main_blocks = (n + 7) / 8;
        if (main_blocks > sms * 4 * 16) main_blocks = sms * 4 * 16;
        unsigned grid = (unsigned)(hub_blocks + main_blocks);
        if (bias)
            spmm_planned_vec4_kernel<true><<<grid, 256, 0, stream>>>(rowptr, col, val, n, H, ldh, (int)F, bias, out, ldo, (int)hub_blocks,
                                                                     hub_task_row, hub_task_chunk, hub_task_first, hub_task_count,
                                                                     n_hub_tasks, partial, counters);
        else
            spmm_planned_vec4_kernel<false><<<grid, 256, 0, stream>>>(rowptr, col, val, n, H, ldh, (int)F, bias, out, ldo, (int)hub_blocks,
                                                                      hub_task_row, hub_task_chunk, hub_task_first, hub_task_count,
                                                                      n_hub_tasks, partial, counters);
        GNNB_LAUNCH_CHECK();
        return GNNB_OK;
    }
    if (vec) {
        int64_t long_blocks = (n + 63) / 64;                // hub-row role: 64 rows scanned per CTA iteration
        if (long_blocks > sms * 8) long_blocks = sms * 8;
        int64_t main_blocks = (n + 7) / 8;                  // 8 warps (rows) per CTA
        if (main_blocks > sms * 4 * 16) main_blocks = sms * 4 * 16;
        unsigned grid = (unsigned)(long_blocks + main_blocks);
        if (bias) spmm_fused_vec4_kernel<true><<<grid, 256, 0, stream>>>(rowptr, col, val, n, H, ldh, (int)F, bias, out, ldo, (int)long_blocks);
        else      spmm_fused_vec4_kernel<false><<<grid, 256, 0, stream>>>(rowptr, col, val, n, H, ldh, (int)F, bias, out, ldo, (int)long_blocks);
        GNNB_LAUNCH_CHECK();
    } else {
        int64_t blocks = (n + 7) / 8;
        int64_t cap = sms * 8 * 4;
        unsigned grid = (unsigned)(blocks < cap ? blocks : cap);
        int64_t lblocks = (n + 63) / 64;
        unsigned lgrid = (unsigned)(lblocks < sms * 8 ? lblocks : sms * 8);
        if (bias) {
            spmm_scalar_kernel<true><<<grid, 256, 0, stream>>>(rowptr, col, val, n, H, ldh, (int)F, bias, out, ldo);
            spmm_long_rows_scalar_kernel<true><<<lgrid, 256, 0, stream>>>(rowptr, col, val, n, H, ldh, (int)F, bias, out, ldo);
        } else {
            spmm_scalar_kernel<false><<<grid, 256, 0, stream>>>(rowptr, col, val, n, H, ldh, (int)F, bias, out, ldo);
            spmm_long_rows_scalar_kernel<false><<<lgrid, 256, 0, stream>>>(rowptr, col, val, n, H, ldh, (int)F, bias, out, ldo);
        }
        GNNB_LAUNCH_CHECK();
    }
    return GNNB_OK;
}
