// K1: k nearest neighbours of int8 genotype rows, canonical (d2, index) order.
//
// Replaces sklearn.neighbors.kneighbors_graph at gnn/loaders/load.py:52-59 and gnn/loaders/util.py:52-57.
// This file holds the CUDA-core (dp4a) tile kernel, the per-row selection shared with the tensor-core
// kernel (knn_tc.cu) and the merge of partial lists.
//
// Layout: a CTA owns a tile of 64 query rows and walks a contiguous range of 64-row key tiles
// ("key split").  Dot products are exact int32; d2 = |q|^2 + |k|^2 - 2 q.k is formed in int64 and, when
// extra float columns are present, extended in float64 with explicitly un-fused multiply/add so that
// it matches the oracle bit for bit.  Each query row keeps its k best (d2, j) pairs sorted in shared
// memory; the partial lists of all key splits are merged by knn_merge_kernel.
#include <math.h>

#include "common.cuh"
#include "knn_select.cuh"

namespace gnnb {

constexpr int TQ = 64, TK = 64, KC_WORDS = 16, STR = KC_WORDS + 1;

__device__ __forceinline__ int load_word(const int8_t* __restrict__ row, int64_t byte0, int64_t p, bool aligned4) {
    if (aligned4 && byte0 + 4 <= p) return __ldg(reinterpret_cast<const int*>(row + byte0));
    int v = 0;
#pragma unroll
    for (int b = 0; b < 4; ++b)
        if (byte0 + b < p) v |= ((int)(unsigned char)row[byte0 + b]) << (8 * b);
    return v;
}

template <bool HAS_EXTRA>
__global__ void __launch_bounds__(256)
knn_dp4a_kernel(const int8_t* __restrict__ Q, int64_t ldq, const int64_t* __restrict__ sq_q, int64_t n_q, int64_t q_off,
                const int8_t* __restrict__ K, int64_t ldk, const int64_t* __restrict__ sq_k, int64_t n_k, int64_t k_off,
                int64_t p, const double* __restrict__ extra_q, const double* __restrict__ extra_k, int n_extra,
                int k, int n_splits, double* __restrict__ part_d2, int32_t* __restrict__ part_idx, int aligned4) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    int* Qs = reinterpret_cast<int*>(smem_raw);                 // [TQ][STR]
    int* Ks = Qs + TQ * STR;                                    // [TK][STR]
    int* D = Ks + TK * STR;                                     // [TQ][TK + 1] dot products
    static_assert(((2 * TQ * STR + TQ * (TK + 1)) * 4) % 8 == 0, "sqk_s must be 8-byte aligned");
    long long* sqk_s = reinterpret_cast<long long*>(D + TQ * (TK + 1));       // [TK]
    double* ek_s = reinterpret_cast<double*>(sqk_s + TK);       // [TK][MAX_EXTRA]
    double* list_d = ek_s + TK * KNN_MAX_EXTRA;                 // [k][TQ]
    int* list_i = reinterpret_cast<int*>(list_d + (size_t)k * TQ);   // [k][TQ]

    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;
    const int64_t q0 = (int64_t)blockIdx.x * TQ;
    const int split = blockIdx.y;
    const int64_t n_ktiles = (n_k + TK - 1) / TK;
    const int64_t kt_begin = n_ktiles * split / n_splits;
    const int64_t kt_end = n_ktiles * (split + 1) / n_splits;

    // selection state: thread r < TQ owns query row q0 + r
    RowSelector sel;
    double eq_r[KNN_MAX_EXTRA];
    long long sqq_r = 0;
    if (tid < TQ) {
        sel.init(list_d, list_i, TQ, tid, k);
        if (q0 + tid < n_q) {
            sqq_r = sq_q[q0 + tid];
            if (HAS_EXTRA)
                for (int c = 0; c < n_extra; ++c) eq_r[c] = extra_q[(q0 + tid) * n_extra + c];
        }
    }

    const int64_t n_chunks = (p + 4 * KC_WORDS - 1) / (4 * KC_WORDS);
    for (int64_t kt = kt_begin; kt < kt_end; ++kt) {
        const int64_t k0 = kt * TK;
        int acc[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = 0;

        for (int64_t ch = 0; ch < n_chunks; ++ch) {
            const int w = tid & 15;
            const int64_t byte0 = ch * (4 * KC_WORDS) + 4 * w;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                int r = (tid >> 4) + 16 * i;
                int64_t gq = q0 + r, gk = k0 + r;
                Qs[r * STR + w] = (gq < n_q) ? load_word(Q + gq * ldq, byte0, p, aligned4) : 0;
                Ks[r * STR + w] = (gk < n_k) ? load_word(K + gk * ldk, byte0, p, aligned4) : 0;
            }
            __syncthreads();
#pragma unroll
            for (int w2 = 0; w2 < KC_WORDS; ++w2) {
                int qv[4], kv[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) qv[i] = Qs[(ty + 16 * i) * STR + w2];
#pragma unroll
                for (int j = 0; j < 4; ++j) kv[j] = Ks[(tx + 16 * j) * STR + w2];
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc[i][j] = __dp4a(qv[i], kv[j], acc[i][j]);
            }
            __syncthreads();
        }
        // publish the tile of dot products and the key-side terms
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) D[(ty + 16 * i) * (TK + 1) + tx + 16 * j] = acc[i][j];
        if (tid < TK) {
            int64_t gk = k0 + tid;
            sqk_s[tid] = (gk < n_k) ? sq_k[gk] : 0;
            if (HAS_EXTRA)
                for (int c = 0; c < n_extra; ++c) ek_s[tid * KNN_MAX_EXTRA + c] = (gk < n_k) ? extra_k[gk * n_extra + c] : 0.0;
        }
        __syncthreads();
        if (tid < TQ && q0 + tid < n_q) {
            const int64_t gi = q_off + q0 + tid;
            const int ncols = (int)((n_k - k0) < TK ? (n_k - k0) : TK);
            for (int c = 0; c < ncols; ++c) {
                const int64_t gj = k_off + k0 + c;
                if (gj == gi) continue;
                long long di = sqq_r + sqk_s[c] - 2LL * (long long)D[tid * (TK + 1) + c];
                double d = (double)di;
                if (HAS_EXTRA) {
                    for (int e = 0; e < n_extra; ++e) {
                        double diff = __dsub_rn(eq_r[e], ek_s[c * KNN_MAX_EXTRA + e]);
                        d = __dadd_rn(d, __dmul_rn(diff, diff));
                    }
                }
                sel.offer(d, (int)gj);
            }
        }
        __syncthreads();
    }
    // write this split's partial lists: part[split][row][t]
    if (tid < TQ && q0 + tid < n_q) {
        double* od = part_d2 + ((int64_t)split * n_q + q0 + tid) * k;
        int32_t* oi = part_idx + ((int64_t)split * n_q + q0 + tid) * k;
        for (int t = 0; t < k; ++t) {
            od[t] = sel.d(t);
            oi[t] = sel.i(t);
        }
    }
}

// Float features (scaled / transformed markers, load.py:45-48): d2 = sum_c fl(fl(q_c - k_c)^2), accumulated left to right
// in float64 with un-fused operations -- the same arithmetic, in the same order, as the oracle, so the lists are
// bit-identical.  CUDA-core tiles (64 x 64 pairs, 16 columns per stage); the tensor-core route for this case is future work.
constexpr int FQ = 64, FK = 64, FC = 16;
__global__ void __launch_bounds__(256)
knn_f64_kernel(const double* __restrict__ Q, int64_t ldq, int64_t n_q, int64_t q_off,
               const double* __restrict__ K, int64_t ldk, int64_t n_k, int64_t k_off, int64_t p,
               int k, int n_splits, double* __restrict__ part_d2, int32_t* __restrict__ part_idx) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* Qs = reinterpret_cast<double*>(smem_raw);            // [FQ][FC + 1]
    double* Ks = Qs + FQ * (FC + 1);                             // [FK][FC + 1]
    double* D = Ks + FK * (FC + 1);                              // [FQ][FK + 1]
    double* list_d = D + FQ * (FK + 1);                          // [k][FQ]
    int* list_i = reinterpret_cast<int*>(list_d + (size_t)k * FQ);
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int64_t q0 = (int64_t)blockIdx.x * FQ;
    const int split = blockIdx.y;
    const int64_t n_ktiles = (n_k + FK - 1) / FK;
    const int64_t kt_begin = n_ktiles * split / n_splits, kt_end = n_ktiles * (split + 1) / n_splits;
    RowSelector sel;
    if (tid < FQ) sel.init(list_d, list_i, FQ, tid, k);
    for (int64_t kt = kt_begin; kt < kt_end; ++kt) {
        const int64_t k0 = kt * FK;
        double acc[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = 0.0;
        for (int64_t c0 = 0; c0 < p; c0 += FC) {
            const int c = tid & 15;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int r = (tid >> 4) + 16 * i;
                const int64_t gq = q0 + r, gk = k0 + r, gc = c0 + c;
                Qs[r * (FC + 1) + c] = (gq < n_q && gc < p) ? Q[gq * ldq + gc] : 0.0;
                Ks[r * (FC + 1) + c] = (gk < n_k && gc < p) ? K[gk * ldk + gc] : 0.0;
            }
            __syncthreads();
            const int cmax = (int)((p - c0) < FC ? (p - c0) : FC);
            for (int cc = 0; cc < cmax; ++cc) {
                double qv[4], kv[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) qv[i] = Qs[(ty + 16 * i) * (FC + 1) + cc];
#pragma unroll
                for (int j = 0; j < 4; ++j) kv[j] = Ks[(tx + 16 * j) * (FC + 1) + cc];
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const double diff = __dsub_rn(qv[i], kv[j]);
                        acc[i][j] = __dadd_rn(acc[i][j], __dmul_rn(diff, diff));
                    }
            }
            __syncthreads();
        }
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) D[(ty + 16 * i) * (FK + 1) + tx + 16 * j] = acc[i][j];
        __syncthreads();
        if (tid < FQ && q0 + tid < n_q) {
            const int64_t gi = q_off + q0 + tid;
            const int ncols = (int)((n_k - k0) < FK ? (n_k - k0) : FK);
            for (int c = 0; c < ncols; ++c) {
                const int64_t gj = k_off + k0 + c;
                if (gj != gi) sel.offer(D[tid * (FK + 1) + c], (int)gj);
            }
        }
        __syncthreads();
    }
    if (tid < FQ && q0 + tid < n_q) {
        double* od = part_d2 + ((int64_t)split * n_q + q0 + tid) * k;
        int32_t* oi = part_idx + ((int64_t)split * n_q + q0 + tid) * k;
        for (int t = 0; t < k; ++t) {
            od[t] = sel.d(t);
            oi[t] = sel.i(t);
        }
    }
}

// include_self=True (load.py:28,57): the k smallest (d2, j) over ALL j are the first k of the self-excluded list with
// (0, i) inserted at its lexicographic place.
__global__ void knn_insert_self_kernel(double* __restrict__ best_d2, int32_t* __restrict__ best_idx, int64_t n, int k, int64_t q_off) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    double* d = best_d2 + r * k;
    int32_t* ix = best_idx + r * k;
    const int me = (int)(q_off + r);
    int pos = 0;
    while (pos < k && d[pos] == 0.0 && ix[pos] >= 0 && ix[pos] < me) ++pos;
    if (pos == k) return;
    for (int t = k - 1; t > pos; --t) {
        d[t] = d[t - 1];
        ix[t] = ix[t - 1];
    }
    d[pos] = 0.0;
    ix[pos] = me;
}

// one thread per query row: merge n_lists sorted partial lists (and the running result when merge != 0)
__global__ void knn_merge_kernel(const double* __restrict__ part_d2, const int32_t* __restrict__ part_idx, int n_lists,
                                 int64_t n_q, int k, double* __restrict__ best_d2, int32_t* __restrict__ best_idx,
                                 int merge, double* __restrict__ tmp_d2, int32_t* __restrict__ tmp_idx) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_q) return;
    // heads of each list; list n_lists (if merge) is the running result, copied to tmp first
    int head[KNN_MAX_SPLITS + 1];
    int total = n_lists + (merge ? 1 : 0);
    for (int l = 0; l < total; ++l) head[l] = 0;
    if (merge) {
        for (int t = 0; t < k; ++t) {
            tmp_d2[r * k + t] = best_d2[r * k + t];
            tmp_idx[r * k + t] = best_idx[r * k + t];
        }
    }
    for (int t = 0; t < k; ++t) {
        double bd = INFINITY;
        unsigned bi = 0xffffffffu;
        int bl = -1;
        for (int l = 0; l < total; ++l) {
            if (head[l] >= k) continue;
            double d;
            unsigned i;
            if (l < n_lists) {
                d = part_d2[((int64_t)l * n_q + r) * k + head[l]];
                i = (unsigned)part_idx[((int64_t)l * n_q + r) * k + head[l]];
            } else {
                d = tmp_d2[r * k + head[l]];
                i = (unsigned)tmp_idx[r * k + head[l]];
            }
            if (bl < 0 || d < bd || (d == bd && i < bi)) { bd = d; bi = i; bl = l; }
        }
        if (bl >= 0) head[bl]++;
        best_d2[r * k + t] = bd;
        best_idx[r * k + t] = (bl >= 0 && bd != INFINITY) ? (int32_t)bi : -1;
    }
}

static size_t dp4a_smem_bytes(int k) {
    size_t b = (size_t)(2 * TQ * STR + TQ * (TK + 1)) * sizeof(int);
    b += TK * sizeof(long long) + (size_t)TK * KNN_MAX_EXTRA * sizeof(double);
    b += (size_t)k * TQ * (sizeof(double) + sizeof(int));
    return b;
}

int knn_pick_splits(int64_t n_q, int64_t n_k, int tile_q, int tile_k) {
    int64_t q_tiles = (n_q + tile_q - 1) / tile_q;
    int64_t k_tiles = (n_k + tile_k - 1) / tile_k;
    int64_t want = (2LL * sm_count() + q_tiles - 1) / q_tiles;
    if (want > KNN_MAX_SPLITS) want = KNN_MAX_SPLITS;
    if (want > k_tiles) want = k_tiles;
    if (want < 1) want = 1;
    return (int)want;
}

// implemented in knn_tc.cu
int knn_tc_launch(const int8_t* Q, int64_t ldq, const int64_t* sq_q, int64_t n_q, int64_t q_off,
                  const int8_t* K, int64_t ldk, const int64_t* sq_k, int64_t n_k, int64_t k_off, int64_t p,
                  const double* extra_q, const double* extra_k, int n_extra, int k, int n_splits,
                  double* part_d2, int32_t* part_idx, cudaStream_t stream, int fp4 = 0);
bool knn_tc_supported(int64_t ldq, int64_t ldk, const void* Q, const void* K, int k);
int knn_tc_splits(int64_t n_q, int64_t n_k);

}  // namespace gnnb

using namespace gnnb;

extern "C" {

int gnnb_knn_workspace_bytes(int64_t n_q, int64_t n_k, int64_t p, int k, int impl, size_t* bytes) {
    GNNB_REQUIRE(bytes && n_q >= 0 && n_k >= 0 && p > 0 && k > 0 && k <= GNNB_KNN_MAX_K, GNNB_E_BADARG,
                 "knn_workspace_bytes: bad argument (k must be 1..%d)", GNNB_KNN_MAX_K);
    // one partial list per key split (the larger plan of the two implementations) + a copy of the running result
    int s1 = knn_pick_splits(n_q > 0 ? n_q : 1, n_k, TQ, TK);
    int s2 = knn_tc_splits(n_q > 0 ? n_q : 1, n_k);
    int splits = impl == 1 ? s1 : (impl == 2 ? s2 : (s1 > s2 ? s1 : s2));
    size_t per_list = (size_t)n_q * k * (sizeof(double) + sizeof(int32_t));
    *bytes = (align_up(per_list, 256) + 512) * ((size_t)splits + 1) + 4096;
    return GNNB_OK;
}

int gnnb_knn_i8(const int8_t* Q, int64_t ldq, const int64_t* sq_q, int64_t n_q, int64_t q_off,
                const int8_t* K, int64_t ldk, const int64_t* sq_k, int64_t n_k, int64_t k_off,
                int64_t p,
                const double* extra_q, const double* extra_k, int n_extra,
                int k, double* best_d2, int32_t* best_idx, int merge, int impl,
                void* workspace, size_t workspace_bytes, void* stream_) {
    GNNB_ENTER();
    cudaStream_t stream = (cudaStream_t)stream_;
    GNNB_REQUIRE(Q && K && sq_q && sq_k && best_d2 && best_idx, GNNB_E_BADARG, "knn_i8: null pointer");
    GNNB_REQUIRE(n_q >= 0 && n_k >= 0 && p > 0 && ldq >= p && ldk >= p, GNNB_E_BADARG, "knn_i8: bad shape");
    GNNB_REQUIRE(k > 0 && k <= GNNB_KNN_MAX_K, GNNB_E_UNSUPPORTED, "knn_i8: k = %d outside 1..%d", k, GNNB_KNN_MAX_K);
    GNNB_REQUIRE(n_extra >= 0 && n_extra <= KNN_MAX_EXTRA, GNNB_E_UNSUPPORTED, "knn_i8: at most %d float columns", KNN_MAX_EXTRA);
    GNNB_REQUIRE(n_extra == 0 || (extra_q && extra_k), GNNB_E_BADARG, "knn_i8: extra columns missing");
    GNNB_REQUIRE(p <= 133000, GNNB_E_UNSUPPORTED, "knn_i8: p = %lld would overflow the int32 dot product", (long long)p);
    GNNB_REQUIRE(impl >= 0 && impl <= 2, GNNB_E_BADARG, "knn_i8: impl must be 0, 1 or 2");
    GNNB_REQUIRE(q_off + n_q < 2147483647LL && k_off + n_k < 2147483647LL, GNNB_E_UNSUPPORTED, "knn_i8: indices exceed int32");
    if (n_q == 0) return GNNB_OK;
    size_t need = 0;
    gnnb_knn_workspace_bytes(n_q, n_k, p, k, impl, &need);
    GNNB_REQUIRE(workspace && workspace_bytes >= need, GNNB_E_WORKSPACE, "knn_i8: workspace %zu < %zu bytes", workspace_bytes, need);

    bool tc_ok = knn_tc_supported(ldq, ldk, Q, K, k);
    if (impl == 2 && !tc_ok)
        return fail(GNNB_E_UNSUPPORTED, "knn_i8: tensor-core path needs 16-byte aligned rows (ld %% 16 == 0) and k <= 64");
    bool use_tc = (impl == 2) || (impl == 0 && tc_ok && n_q >= 128 && n_k >= 128 && p >= 64);

    int n_splits = n_k == 0 ? 0 : (use_tc ? knn_tc_splits(n_q, n_k) : knn_pick_splits(n_q, n_k, TQ, TK));
    Carver cv(workspace, workspace_bytes);
    size_t per = (size_t)n_q * k;
    double* part_d2 = cv.take<double>(per * (n_splits > 0 ? n_splits : 1));
    int32_t* part_idx = cv.take<int32_t>(per * (n_splits > 0 ? n_splits : 1));
    double* tmp_d2 = cv.take<double>(per);
    int32_t* tmp_idx = cv.take<int32_t>(per);
    GNNB_REQUIRE(cv.ok(), GNNB_E_WORKSPACE, "knn_i8: workspace carve overflow");

    if (n_splits > 0) {
        if (use_tc) {
            int rc = knn_tc_launch(Q, ldq, sq_q, n_q, q_off, K, ldk, sq_k, n_k, k_off, p, extra_q, extra_k, n_extra, k,
                                   n_splits, part_d2, part_idx, stream);
            if (rc != GNNB_OK) return rc;
        } else {
            int aligned4 = ((uintptr_t)Q % 4 == 0) && ((uintptr_t)K % 4 == 0) && (ldq % 4 == 0) && (ldk % 4 == 0);
            size_t smem = dp4a_smem_bytes(k);
            dim3 grid((unsigned)((n_q + TQ - 1) / TQ), (unsigned)n_splits);
            if (n_extra > 0) {
                GNNB_SET_DYN_SMEM(knn_dp4a_kernel<true>, smem);
                knn_dp4a_kernel<true><<<grid, 256, smem, stream>>>(Q, ldq, sq_q, n_q, q_off, K, ldk, sq_k, n_k, k_off, p,
                                                                   extra_q, extra_k, n_extra, k, n_splits, part_d2, part_idx, aligned4);
            } else {
                GNNB_SET_DYN_SMEM(knn_dp4a_kernel<false>, smem);
                knn_dp4a_kernel<false><<<grid, 256, smem, stream>>>(Q, ldq, sq_q, n_q, q_off, K, ldk, sq_k, n_k, k_off, p,
                                                                    extra_q, extra_k, n_extra, k, n_splits, part_d2, part_idx, aligned4);
            }
            GNNB_LAUNCH_CHECK();
        }
    }
    knn_merge_kernel<<<(unsigned)((n_q + 127) / 128), 128, 0, stream>>>(part_d2, part_idx, n_splits, n_q, k, best_d2, best_idx,
                                                                        merge, tmp_d2, tmp_idx);
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

// The same search with both operands as e2m1-packed codes (gnnb_pack_fp4) on the FP4 tensor path (tcgen05 kind::mxf4).
int gnnb_knn_fp4(const uint8_t* Q4, int64_t ldq, const int64_t* sq_q, int64_t n_q, int64_t q_off,
                 const uint8_t* K4, int64_t ldk, const int64_t* sq_k, int64_t n_k, int64_t k_off,
                 int64_t p,
                 const double* extra_q, const double* extra_k, int n_extra,
                 int k, double* best_d2, int32_t* best_idx, int merge,
                 void* workspace, size_t workspace_bytes, void* stream_) {
    GNNB_ENTER();
    cudaStream_t stream = (cudaStream_t)stream_;
    GNNB_REQUIRE(Q4 && K4 && sq_q && sq_k && best_d2 && best_idx, GNNB_E_BADARG, "knn_fp4: null pointer");
    GNNB_REQUIRE(n_q >= 0 && n_k >= 0 && p > 0 && ldq >= (p + 1) / 2 && ldk >= (p + 1) / 2, GNNB_E_BADARG, "knn_fp4: bad shape");
    GNNB_REQUIRE(k > 0 && k <= GNNB_KNN_MAX_K, GNNB_E_UNSUPPORTED, "knn_fp4: k = %d outside 1..%d", k, GNNB_KNN_MAX_K);
    GNNB_REQUIRE(n_extra >= 0 && n_extra <= KNN_MAX_EXTRA, GNNB_E_UNSUPPORTED, "knn_fp4: at most %d float columns", KNN_MAX_EXTRA);
    GNNB_REQUIRE(n_extra == 0 || (extra_q && extra_k), GNNB_E_BADARG, "knn_fp4: extra columns missing");
    GNNB_REQUIRE(p <= 4000000, GNNB_E_UNSUPPORTED, "knn_fp4: p = %lld: the f32 dot product of 0/1/2 codes is exact up to 4 p < 2^24", (long long)p);
    GNNB_REQUIRE(q_off + n_q < 2147483647LL && k_off + n_k < 2147483647LL, GNNB_E_UNSUPPORTED, "knn_fp4: indices exceed int32");
    GNNB_REQUIRE(knn_tc_supported(ldq, ldk, Q4, K4, k), GNNB_E_UNSUPPORTED, "knn_fp4: needs 16-byte aligned packed rows (ld %% 16 == 0) and k <= 64");
    if (n_q == 0) return GNNB_OK;
    size_t need = 0;
    gnnb_knn_workspace_bytes(n_q, n_k, p, k, 2, &need);
    GNNB_REQUIRE(workspace && workspace_bytes >= need, GNNB_E_WORKSPACE, "knn_fp4: workspace %zu < %zu bytes", workspace_bytes, need);
    const int n_splits = n_k == 0 ? 0 : knn_tc_splits(n_q, n_k);
    Carver cv(workspace, workspace_bytes);
    size_t per = (size_t)n_q * k;
    double* part_d2 = cv.take<double>(per * (n_splits > 0 ? n_splits : 1));
    int32_t* part_idx = cv.take<int32_t>(per * (n_splits > 0 ? n_splits : 1));
    double* tmp_d2 = cv.take<double>(per);
    int32_t* tmp_idx = cv.take<int32_t>(per);
    GNNB_REQUIRE(cv.ok(), GNNB_E_WORKSPACE, "knn_fp4: workspace carve overflow");
    if (n_splits > 0) {
        int rc = knn_tc_launch(reinterpret_cast<const int8_t*>(Q4), ldq, sq_q, n_q, q_off, reinterpret_cast<const int8_t*>(K4), ldk, sq_k,
                               n_k, k_off, p, extra_q, extra_k, n_extra, k, n_splits, part_d2, part_idx, stream, /*fp4*/ 1);
        if (rc != GNNB_OK) return rc;
    }
    knn_merge_kernel<<<(unsigned)((n_q + 127) / 128), 128, 0, stream>>>(part_d2, part_idx, n_splits, n_q, k, best_d2, best_idx,
                                                                        merge, tmp_d2, tmp_idx);
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

int gnnb_knn_f64(const double* Q, int64_t ldq, int64_t n_q, int64_t q_off,
                 const double* K, int64_t ldk, int64_t n_k, int64_t k_off, int64_t p,
                 int k, double* best_d2, int32_t* best_idx, int merge,
                 void* workspace, size_t workspace_bytes, void* stream_) {
    GNNB_ENTER();
    cudaStream_t stream = (cudaStream_t)stream_;
    GNNB_REQUIRE(Q && K && best_d2 && best_idx, GNNB_E_BADARG, "knn_f64: null pointer");
    GNNB_REQUIRE(n_q >= 0 && n_k >= 0 && p > 0 && ldq >= p && ldk >= p, GNNB_E_BADARG, "knn_f64: bad shape");
    GNNB_REQUIRE(k > 0 && k <= GNNB_KNN_MAX_K, GNNB_E_UNSUPPORTED, "knn_f64: k = %d outside 1..%d", k, GNNB_KNN_MAX_K);
    GNNB_REQUIRE(q_off + n_q < 2147483647LL && k_off + n_k < 2147483647LL, GNNB_E_UNSUPPORTED, "knn_f64: indices exceed int32");
    if (n_q == 0) return GNNB_OK;
    size_t need = 0;
    gnnb_knn_workspace_bytes(n_q, n_k, p, k, 1, &need);
    GNNB_REQUIRE(workspace && workspace_bytes >= need, GNNB_E_WORKSPACE, "knn_f64: workspace %zu < %zu bytes", workspace_bytes, need);
    int n_splits = n_k == 0 ? 0 : knn_pick_splits(n_q, n_k, FQ, FK);
    Carver cv(workspace, workspace_bytes);
    size_t per = (size_t)n_q * k;
    double* part_d2 = cv.take<double>(per * (n_splits > 0 ? n_splits : 1));
    int32_t* part_idx = cv.take<int32_t>(per * (n_splits > 0 ? n_splits : 1));
    double* tmp_d2 = cv.take<double>(per);
    int32_t* tmp_idx = cv.take<int32_t>(per);
    GNNB_REQUIRE(cv.ok(), GNNB_E_WORKSPACE, "knn_f64: workspace carve overflow");
    if (n_splits > 0) {
        size_t smem = (size_t)(2 * FQ * (FC + 1) + FQ * (FK + 1)) * sizeof(double) + (size_t)k * FQ * (sizeof(double) + sizeof(int));
        GNNB_SET_DYN_SMEM(knn_f64_kernel, smem);
        dim3 grid((unsigned)((n_q + FQ - 1) / FQ), (unsigned)n_splits);
        knn_f64_kernel<<<grid, 256, smem, stream>>>(Q, ldq, n_q, q_off, K, ldk, n_k, k_off, p, k, n_splits, part_d2, part_idx);
        GNNB_LAUNCH_CHECK();
    }
    knn_merge_kernel<<<(unsigned)((n_q + 127) / 128), 128, 0, stream>>>(part_d2, part_idx, n_splits, n_q, k, best_d2, best_idx,
                                                                        merge, tmp_d2, tmp_idx);
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

int gnnb_knn_insert_self(double* best_d2, int32_t* best_idx, int64_t n_q, int k, int64_t q_off, void* stream) {
    GNNB_ENTER();
    GNNB_REQUIRE(best_d2 && best_idx && n_q >= 0 && k > 0 && k <= GNNB_KNN_MAX_K && q_off >= 0, GNNB_E_BADARG, "knn_insert_self: bad argument");
    if (n_q == 0) return GNNB_OK;
    knn_insert_self_kernel<<<(unsigned)((n_q + 127) / 128), 128, 0, (cudaStream_t)stream>>>(best_d2, best_idx, n_q, k, q_off);
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

}  // extern "C"
