// K1 on the 5th-generation tensor cores: int8 Gram tiles (tcgen05.mma kind::i8, s32 accumulators in TMEM),
// operands staged by TMA with the 128-byte swizzle, fused distance epilogue + running per-row top-k.
//
// Replaces the O(n^2 p) part of sklearn.neighbors.kneighbors_graph (gnn/loaders/load.py:52-59,
// gnn/loaders/util.py:52-57).  Results are bit-identical to the dp4a kernel in knn.cu and to the oracle:
// the dot products are exact integers either way and the selection code is shared (knn_select.cuh).
//
// CTA = 128 query rows x a range of 256-row key tiles; 8 warps:
//   warp 0    TMA producer      A = Q[128 x 128 B], B = K[256 x 128 B] per k-block into a 2-3 stage ring
//   warp 1    MMA issuer        4 x (M128 N256 K32) per k-block into one of two 256-column TMEM accumulators
//   warp 2    TMEM allocator
//   warps 4-7 epilogue          thread = query row = TMEM lane: tcgen05.ld 32 columns at a time,
//                               d2 = |q|^2 + |k|^2 - 2 q.k, one integer compare rejects almost everything,
//                               survivors are inserted into the row's sorted list in shared memory
// The two accumulators let the epilogue of key tile j overlap the MMAs of key tile j + 1.
// CTAs are rasterised in groups of QG query tiles x n_splits key ranges so that a wave re-uses both operands in L2.
#include <limits.h>
#include <math.h>

#include "knn_select.cuh"
#include "tc_common.cuh"

namespace gnnb {

using namespace tc;

namespace {
constexpr int BM = 128, BN = 256, BKB = 128;              // BKB: bytes (= int8 elements) of K per stage
constexpr int A_BYTES = BM * BKB, B_BYTES = BN * BKB, STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int MAX_STAGES = 4;
constexpr int QG = 18;                                    // query tiles per raster group
constexpr int EPI_THREADS = 128;

struct Smem {
    uint8_t* tiles;
    uint64_t *full, *empty, *tfull, *tempty;
    uint32_t* tmem_ptr;
    long long* sqk;      // [2][BN]
    double* ek;          // [2][BN][KNN_MAX_EXTRA] (only with extra columns)
    double* list_d;      // [k][BM]
    int* list_i;         // [k][BM]
};

__host__ __device__ inline size_t smem_bytes(int stages, int k, bool extra) {
    return 1024 + (size_t)stages * STAGE_BYTES + 128 + 2 * BN * sizeof(long long) +
           (extra ? 2 * BN * KNN_MAX_EXTRA * sizeof(double) : 0) + (size_t)k * BM * (sizeof(double) + sizeof(int));
}

__device__ inline Smem carve(uint8_t* raw, int stages, int k, bool extra) {
    Smem s;
    uint8_t* base = raw + ((1024 - (smem_u32(raw) & 1023)) & 1023);      // 128B-swizzle tiles need 1024-byte alignment
    s.tiles = base;
    uint8_t* p = base + (size_t)stages * STAGE_BYTES;
    s.full = reinterpret_cast<uint64_t*>(p);
    s.empty = s.full + MAX_STAGES;
    s.tfull = s.empty + MAX_STAGES;
    s.tempty = s.tfull + 2;
    s.tmem_ptr = reinterpret_cast<uint32_t*>(s.tempty + 2);
    p += 128;
    s.sqk = reinterpret_cast<long long*>(p);
    p += 2 * BN * sizeof(long long);
    s.ek = reinterpret_cast<double*>(p);
    if (extra) p += 2 * BN * KNN_MAX_EXTRA * sizeof(double);
    s.list_d = reinterpret_cast<double*>(p);
    s.list_i = reinterpret_cast<int*>(s.list_d + (size_t)k * BM);
    return s;
}
}  // namespace

template <bool HAS_EXTRA>
__global__ void __launch_bounds__(256, 1)
knn_tc_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_k,
              const int64_t* __restrict__ sq_q, int64_t n_q, int64_t q_off,
              const int64_t* __restrict__ sq_k, int64_t n_k, int64_t k_off, int64_t p,
              const double* __restrict__ extra_q, const double* __restrict__ extra_k, int n_extra,
              int k, int n_splits, int q_tiles, int stages, double* __restrict__ part_d2, int32_t* __restrict__ part_idx) {
    extern __shared__ uint8_t smem_raw[];
    const int warp = threadIdx.x >> 5;

    // rasterisation: groups of QG query tiles x n_splits key ranges
    const int per_group = QG * n_splits;
    const int group = blockIdx.x / per_group, rem = blockIdx.x % per_group;
    const int q_tile = group * QG + rem % QG;
    const int split = rem / QG;
    if (q_tile >= q_tiles) return;                              // whole CTA leaves before any barrier exists

    Smem sm = carve(smem_raw, stages, k, HAS_EXTRA);
    const int64_t n_ktiles = (n_k + BN - 1) / BN;
    const int64_t kt_begin = n_ktiles * split / n_splits;
    const int64_t kt_end = n_ktiles * (split + 1) / n_splits;
    const int n_kb = (int)((p + BKB - 1) / BKB);

    if (threadIdx.x == 0) {
        for (int s = 0; s < stages; ++s) {
            mbar_init(&sm.full[s], 1);
            mbar_init(&sm.empty[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&sm.tfull[b], 1);
            mbar_init(&sm.tempty[b], EPI_THREADS);
        }
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc<512>(sm.tmem_ptr);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *sm.tmem_ptr;

    if (warp == 0) {
        // ===================================================== TMA producer
        if (elect_one()) {
            prefetch_tensormap(&map_q);
            prefetch_tensormap(&map_k);
            uint32_t it = 0;
            for (int64_t kt = kt_begin; kt < kt_end; ++kt) {
                for (int kb = 0; kb < n_kb; ++kb, ++it) {
                    const int s = it % stages;
                    const uint32_t ph = (it / stages) & 1;
                    mbar_wait(&sm.empty[s], ph ^ 1);
                    mbar_arrive_expect_tx(&sm.full[s], STAGE_BYTES);
                    uint8_t* a = sm.tiles + (size_t)s * STAGE_BYTES;
                    tma_load_2d(a, &map_q, &sm.full[s], kb * BKB, q_tile * BM);
                    tma_load_2d(a + A_BYTES, &map_k, &sm.full[s], kb * BKB, (int32_t)(kt * BN));
                }
            }
        }
    } else if (warp == 1) {
        // ===================================================== MMA issuer (one thread)
        if (elect_one()) {
            constexpr uint32_t idesc = make_idesc(/*S32*/ 2, /*int8*/ 1, 0, 0, BM, BN);
            uint32_t it = 0, tile = 0;
            for (int64_t kt = kt_begin; kt < kt_end; ++kt, ++tile) {
                const uint32_t b = tile & 1, tph = (tile >> 1) & 1;
                mbar_wait(&sm.tempty[b], tph ^ 1);                 // epilogue has drained this accumulator
                tc_fence_after_sync();
                const uint32_t d_tmem = tmem_base + b * BN;
                for (int kb = 0; kb < n_kb; ++kb, ++it) {
                    const int s = it % stages;
                    const uint32_t ph = (it / stages) & 1;
                    mbar_wait(&sm.full[s], ph);                    // TMA bytes have landed
                    tc_fence_after_sync();
                    const uint32_t a_addr = smem_u32(sm.tiles + (size_t)s * STAGE_BYTES);
                    const uint32_t b_addr = a_addr + A_BYTES;
#pragma unroll
                    for (int k4 = 0; k4 < BKB / 32; ++k4) {
                        const uint64_t da = make_smem_desc(a_addr + k4 * 32, 16, 1024);
                        const uint64_t db = make_smem_desc(b_addr + k4 * 32, 16, 1024);
                        mma_i8_ss(d_tmem, da, db, idesc, (kb > 0 || k4 > 0) ? 1u : 0u);
                    }
                    tc_commit(&sm.empty[s]);                       // frees the stage when these MMAs have read it
                }
                tc_commit(&sm.tfull[b]);                           // accumulator complete
            }
        }
    } else if (warp >= 4) {
        // ===================================================== epilogue: distances + selection
        const int e = threadIdx.x - 128;                           // row inside the tile == TMEM lane
        const int wq = warp & 3;
        const int64_t row = (int64_t)q_tile * BM + e;
        const bool valid = row < n_q;
        const int64_t gi = q_off + row;
        RowSelector sel;
        sel.init(sm.list_d, sm.list_i, BM, e, k);
        long long sqq = 0, thr_ll = LLONG_MAX;
        double eq_r[KNN_MAX_EXTRA];
        if (valid) {
            sqq = sq_q[row];
            if (HAS_EXTRA)
                for (int c = 0; c < n_extra; ++c) eq_r[c] = extra_q[row * n_extra + c];
        }
        uint32_t tile = 0;
        for (int64_t kt = kt_begin; kt < kt_end; ++kt, ++tile) {
            const uint32_t b = tile & 1, tph = (tile >> 1) & 1;
            const int64_t k0 = kt * BN;
            long long* sqk_b = sm.sqk + b * BN;
            double* ek_b = sm.ek + (size_t)b * BN * KNN_MAX_EXTRA;
#pragma unroll
            for (int t = e; t < BN; t += EPI_THREADS) {
                const int64_t gk = k0 + t;
                sqk_b[t] = (gk < n_k) ? sq_k[gk] : 0;
                if (HAS_EXTRA)
                    for (int c = 0; c < n_extra; ++c) ek_b[t * KNN_MAX_EXTRA + c] = (gk < n_k) ? extra_k[gk * n_extra + c] : 0.0;
            }
            named_bar_sync(1, EPI_THREADS);
            mbar_wait(&sm.tfull[b], tph);
            tc_fence_after_sync();
            const int ncols = (int)((n_k - k0) < BN ? (n_k - k0) : BN);
#pragma unroll 1
            for (int chunk = 0; chunk < BN / 32; ++chunk) {
                uint32_t v[32];
                tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(32 * wq) << 16) + b * BN + chunk * 32, v);
                tmem_ld_wait();
                if (valid) {
#pragma unroll
                    for (int c = 0; c < 32; ++c) {
                        const int col = chunk * 32 + c;
                        const long long di = sqq + sqk_b[col] - 2LL * (long long)(int)v[c];
                        if (di <= thr_ll && col < ncols) {
                            const int64_t gj = k_off + k0 + col;
                            if (gj != gi) {
                                double d = (double)di;
                                if (HAS_EXTRA) {
                                    for (int x = 0; x < n_extra; ++x) {
                                        double diff = __dsub_rn(eq_r[x], ek_b[col * KNN_MAX_EXTRA + x]);
                                        d = __dadd_rn(d, __dmul_rn(diff, diff));
                                    }
                                }
                                sel.offer(d, (int)gj);
                                thr_ll = (sel.thr_d == INFINITY) ? LLONG_MAX : (long long)sel.thr_d;
                            }
                        }
                    }
                }
            }
            tc_fence_before_sync();
            mbar_arrive(&sm.tempty[b]);
        }
        if (valid) {
            double* od = part_d2 + ((int64_t)split * n_q + row) * k;
            int32_t* oi = part_idx + ((int64_t)split * n_q + row) * k;
            for (int t = 0; t < k; ++t) {
                od[t] = sel.d(t);
                oi[t] = sel.i(t);
            }
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 2) tmem_dealloc<512>(tmem_base);
}

bool knn_tc_supported(int64_t ldq, int64_t ldk, const void* Q, const void* K, int k) {
    return ldq % 16 == 0 && ldk % 16 == 0 && (uintptr_t)Q % 16 == 0 && (uintptr_t)K % 16 == 0 && k >= 1 && k <= 64;
}

int knn_tc_splits(int64_t n_q, int64_t n_k) {
    int64_t q_tiles = (n_q + BM - 1) / BM, k_tiles = (n_k + BN - 1) / BN;
    int64_t s;
    if (q_tiles >= QG) s = 8;                                   // ~QG x 8 = 144 co-resident CTAs share 18 Q tiles and 8 K ranges
    else s = (sm_count() + q_tiles - 1) / q_tiles;
    if (s > KNN_MAX_SPLITS) s = KNN_MAX_SPLITS;
    if (s > k_tiles) s = k_tiles;
    if (s < 1) s = 1;
    return (int)s;
}

int knn_tc_launch(const int8_t* Q, int64_t ldq, const int64_t* sq_q, int64_t n_q, int64_t q_off,
                  const int8_t* K, int64_t ldk, const int64_t* sq_k, int64_t n_k, int64_t k_off, int64_t p,
                  const double* extra_q, const double* extra_k, int n_extra, int k, int n_splits,
                  double* part_d2, int32_t* part_idx, cudaStream_t stream) {
    CUtensorMap map_q, map_k;
    int rc = make_tensor_map_2d(&map_q, Q, 1, (uint64_t)p, (uint64_t)n_q, (uint64_t)ldq, BKB, BM);
    if (rc != GNNB_OK) return rc;
    rc = make_tensor_map_2d(&map_k, K, 1, (uint64_t)p, (uint64_t)n_k, (uint64_t)ldk, BKB, BN);
    if (rc != GNNB_OK) return rc;
    const int stages = k <= 32 ? 3 : 2;
    const bool extra = n_extra > 0;
    const size_t smem = smem_bytes(stages, k, extra);
    const int q_tiles = (int)((n_q + BM - 1) / BM);
    const int groups = (q_tiles + QG - 1) / QG;
    const unsigned grid = (unsigned)(groups * QG * n_splits);
    if (extra) {
        { static size_t hw_ = 0; if ((size_t)(smem) > hw_) { GNNB_CUDA(cudaFuncSetAttribute(knn_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(smem))); hw_ = (size_t)(smem); } }
        knn_tc_kernel<true><<<grid, 256, smem, stream>>>(map_q, map_k, sq_q, n_q, q_off, sq_k, n_k, k_off, p, extra_q, extra_k,
                                                         n_extra, k, n_splits, q_tiles, stages, part_d2, part_idx);
    } else {
        { static size_t hw_ = 0; if ((size_t)(smem) > hw_) { GNNB_CUDA(cudaFuncSetAttribute(knn_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(smem))); hw_ = (size_t)(smem); } }
        knn_tc_kernel<false><<<grid, 256, smem, stream>>>(map_q, map_k, sq_q, n_q, q_off, sq_k, n_k, k_off, p, extra_q, extra_k,
                                                          n_extra, k, n_splits, q_tiles, stages, part_d2, part_idx);
    }
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

}  // namespace gnnb
