// K1 on the 5th-generation tensor cores: int8 Gram tiles (tcgen05.mma kind::i8, s32 accumulators in TMEM),
// operands staged by TMA with the 128-byte swizzle, fused distance epilogue + running per-row top-k.
//
// Replaces the O(n^2 p) part of sklearn.neighbors.kneighbors_graph (gnn/loaders/load.py:52-59,
// gnn/loaders/util.py:52-57).  Results are bit-identical to the dp4a kernel in knn.cu and to the oracle:
// the dot products are exact integers either way and the selection code is shared (knn_select.cuh).
//
// Two kernels: knn_tc2_kernel (CTA pairs, cta_group::2, default) and knn_tc_kernel (one CTA per tile; odd shapes and
// GNNB_KNN_PAIR=0).  Layout of the one-CTA kernel:
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

// Ablation switches (timing experiments that produce INVALID results) exist only in builds made with -DGNNB_ABLATION
// (scripts/build_variant.sh); the product library has none.
#ifdef GNNB_ABLATION
#define KNN_ABLATE(bit) (ablate & (bit))
#else
#define KNN_ABLATE(bit) false
#endif

namespace gnnb {

using namespace tc;

namespace {
constexpr int BM = 128, BN = 256, BKB = 128;              // BKB: bytes (= int8 elements) of K per stage
constexpr int A_BYTES = BM * BKB, B_BYTES = BN * BKB, STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int MAX_STAGES = 5;
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

// One accumulator tile (this thread's row x 256 key columns): distances, one compare per column, rare insertions.
// Exact integer path (no float side columns): RowSelectorU64; otherwise the float64 RowSelector.
template <bool HAS_EXTRA, int TN = BN, bool FP4 = false>
__device__ __forceinline__ void scan_accumulator(uint32_t taddr_row, bool valid, long long sqq, const long long* __restrict__ sqk_b,
                                                 const double* __restrict__ ek_b, const double* eq_r, int n_extra, int ncols,
                                                 int64_t gi, int64_t gk0, RowSelector& sel, long long& thr_ll, RowSelectorU64& selu) {
#pragma unroll 1
    for (int chunk = 0; chunk < TN / 32; ++chunk) {
        uint32_t v[32];
        tmem_ld_32x32b_x32(taddr_row + chunk * 32, v);
        tmem_ld_wait();
        if (valid) {
#pragma unroll
            for (int c = 0; c < 32; ++c) {
                const int col = chunk * 32 + c;
                // int8 tiles accumulate in s32; fp4 tiles in f32, where the dot product of 0/1/2 codes is an exact integer < 2^24
                const long long dot = FP4 ? (long long)__uint_as_float(v[c]) : (long long)(int)v[c];
                const long long di = sqq + sqk_b[col] - 2LL * dot;
                if (HAS_EXTRA) {
                    if (di <= thr_ll && col < ncols) {
                        const int64_t gj = gk0 + col;
                        if (gj != gi) {
                            double d = (double)di;
                            for (int x = 0; x < n_extra; ++x) {
                                double diff = __dsub_rn(eq_r[x], ek_b[col * KNN_MAX_EXTRA + x]);
                                d = __dadd_rn(d, __dmul_rn(diff, diff));
                            }
                            sel.offer(d, (int)gj);
                            thr_ll = (sel.thr_d == INFINITY) ? LLONG_MAX : (long long)sel.thr_d;
                        }
                    }
                } else {
                    if (di <= thr_ll && col < ncols) {
                        const int64_t gj = gk0 + col;
                        if (gj != gi) {
                            selu.offer(RowSelectorU64::make_key(di, (int)gj));
                            thr_ll = (long long)(selu.thr >> 32);          // EMPTY >> 32 = 2^32 - 1 >= any distance
                        }
                    }
                }
            }
        }
    }
}

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
        RowSelectorU64 selu;
        if (HAS_EXTRA) sel.init(sm.list_d, sm.list_i, BM, e, k);
        else           selu.init(sm.list_d, BM, e, k);
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
            scan_accumulator<HAS_EXTRA>(tmem_base + ((uint32_t)(32 * wq) << 16) + b * BN, valid, sqq, sqk_b, ek_b, eq_r, n_extra,
                                        ncols, gi, k_off + k0, sel, thr_ll, selu);
            tc_fence_before_sync();
            mbar_arrive(&sm.tempty[b]);
        }
        if (valid) {
            double* od = part_d2 + ((int64_t)split * n_q + row) * k;
            int32_t* oi = part_idx + ((int64_t)split * n_q + row) * k;
            if (!HAS_EXTRA) selu.sort();
            for (int t = 0; t < k; ++t) {
                od[t] = HAS_EXTRA ? sel.d(t) : selu.d(t);
                oi[t] = HAS_EXTRA ? sel.i(t) : selu.i(t);
            }
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 2) tmem_dealloc<512>(tmem_base);
}

// ------------------------------------------------------------------------------------------------------------------
// CTA-pair version (tcgen05 cta_group::2): two CTAs on the two SMs of a TPC compute a 256-query x 256-key tile together.
// Each CTA stages its own 128 query rows and only HALF of the key tile (128 rows); the pair's MMA reads both halves.
// Per k-block a CTA therefore takes in 32 KB instead of 48 KB for the same 128 x 256 dot products -- the one-CTA kernel
// is bound by exactly that stream (L2 -> SM ~ 8 TB/s, profiles/r01a) -- and the smaller stage buys a fourth ring slot.
// The leader (cluster rank 0) issues every MMA; `full` / `tempty` live in the leader, `empty` / `tfull` in each CTA and
// are signalled by multicast commits.  Each CTA's TMEM holds the accumulator rows of its own queries, so the epilogue
// (distance + selection, one thread per query row) is the one-CTA epilogue unchanged.
namespace {
constexpr int B_HALF_BYTES = (BN / 2) * BKB, STAGE2_BYTES = A_BYTES + B_HALF_BYTES;
#ifndef KNN_FP4_NT
#define KNN_FP4_NT 1       // measured at 15 625 x 50 000, k = 32: 1 -> 6.95 ms, 2 -> 10.4 ms (see knn_tc2_kernel)
#endif
constexpr int FP4_NT = KNN_FP4_NT;                        // key tiles per pass of the FP4 kernel (see knn_tc2_kernel)
constexpr int FP4_TN = 224;                               // key-tile width of the fp4 kernel: 2 x 224 accumulator + 64 scale columns of TMEM
constexpr int QGP = 8;                                    // query-tile PAIRS per raster group (even: a quad = two pairs of one split)

__host__ __device__ inline size_t smem_bytes2(int stages, int k, bool extra, int stage_bytes = STAGE2_BYTES) {
    return 1024 + (size_t)stages * stage_bytes + 128 + 2 * BN * sizeof(long long) +
           (extra ? 2 * BN * KNN_MAX_EXTRA * sizeof(double) : 0) + (size_t)k * BM * (sizeof(double) + sizeof(int));
}
__device__ inline Smem carve2(uint8_t* raw, int stages, int k, bool extra, int stage_bytes = STAGE2_BYTES) {
    Smem s;
    uint8_t* base = raw + ((1024 - (smem_u32(raw) & 1023)) & 1023);
    s.tiles = base;
    uint8_t* p = base + (size_t)stages * stage_bytes;
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

// FP4 = true: the operands are e2m1-packed genotype codes (two per byte, gnnb_pack_fp4) and the tiles run on
// tcgen05.mma kind::mxf4 -- K = 64 codes per instruction instead of 32, so a 128-byte k-block covers 256 markers and both
// the tensor time and every byte stream halve.  Block scale factors are mandatory for that kind: a constant 2^0 (ue8m0
// 0x7F) fills the tensor-memory columns behind the accumulators, which is why the key tile is TN = 224 wide there
// (2 x 224 accumulator columns + 64 scale columns = 512).  Products and sums of 0/1/2 codes are exact in f32.
// NT = key tiles per pass over the query tile's k-blocks.  NT = 1: the two accumulators double-buffer (the epilogue of
// tile j overlaps the MMAs of tile j + 1).  NT = 2: both accumulators are filled in ONE pass -- every query k-block is
// fetched once for two key tiles and 8 MMAs share a commit.  It was built for the FP4 kernel, which at twice the tensor
// rate is bound by the L2 -> SM operand stream (ncu: tensor pipe 49 % active), 70 % of it the query tile; but the epilogue
// of a tile pair then no longer overlaps tensor work, and the selection's insertions (a whole warp pays ~150 cycles
// whenever one of its 32 rows inserts, ~160 times per row and key split) cost more than the saved loads:
// 15 625 x 50 000, k = 32: 6.95 ms with NT = 1, 10.4 ms with NT = 2.  NT = 1 is what ships.
template <bool HAS_EXTRA, int TN, bool FP4, int NT>
__global__ void __launch_bounds__(256, 1)
knn_tc2_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_k,
               const int64_t* __restrict__ sq_q, int64_t n_q, int64_t q_off,
               const int64_t* __restrict__ sq_k, int64_t n_k, int64_t k_off, int64_t p,
               const double* __restrict__ extra_q, const double* __restrict__ extra_k, int n_extra,
               int k, int n_splits, int q_pairs, int stages, double* __restrict__ part_d2, int32_t* __restrict__ part_idx,
               int quad,          // 1: clusters of 4 = two pairs with neighbouring query tiles and the same key range; each key
                                  //    half-tile is fetched once (half by each of the two CTAs that need it) and multicast to both
               int ablate) {      // -DGNNB_ABLATION builds only (results invalid): 1 = no TMA loads, 2 = no MMAs, 4 = no epilogue work
    extern __shared__ uint8_t smem_raw[];
    const int warp = threadIdx.x >> 5;
    const uint32_t crank = cluster_ctarank();                   // 0..1, or 0..3 in quad mode
    const uint32_t rank = crank & 1, lead_rank = crank & ~1u, twin = crank >> 1;
    const bool leader = rank == 0;
    const uint16_t pair_mask = (uint16_t)(3u << lead_rank), all_mask = quad ? 15 : 3;

    const int pair = blockIdx.x >> 1;                           // pair = 2 consecutive CTAs
    const int per_group = QGP * n_splits;
    const int group = pair / per_group, rem = pair % per_group;
    const int q_pair = group * QGP + rem % QGP;
    const int split = rem / QGP;
    // a whole cluster leaves together; in quad mode the second pair of a quad may lie past the last query tile and then
    // runs on zero-filled tiles (its key loads and barrier traffic are part of the first pair's pipeline)
    if ((quad ? (q_pair & ~1) : q_pair) >= q_pairs) return;
    const int q_tile = 2 * q_pair + (int)rank;                  // may be one past the last tile: TMA zero-fills, rows are skipped

    constexpr int TB_HALF = (TN / 2) * BKB, TSTAGE = A_BYTES + NT * TB_HALF;   // this CTA's half of NT key tiles; one ring slot
    Smem sm = carve2(smem_raw, stages, k, HAS_EXTRA, NT == 1 ? STAGE2_BYTES : TSTAGE);
    const int64_t n_ktiles = (n_k + TN - 1) / TN;
    const int64_t kt_begin = n_ktiles * split / n_splits;
    const int64_t kt_end = n_ktiles * (split + 1) / n_splits;
    const int n_kb = (int)((p + BKB - 1) / BKB);                  // p = bytes per operand row (markers, or markers / 2 when FP4)

    if (threadIdx.x == 0) {
        for (int s = 0; s < stages; ++s) {
            mbar_init(&sm.full[s], 2);                          // one arrival per CTA's producer (+ the bytes of both)
            mbar_init(&sm.empty[s], quad ? 2 : 1);              // quad: both pairs must be done before a twin refills the slot
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&sm.tfull[b], 1);
            mbar_init(&sm.tempty[b], 2 * EPI_THREADS);          // the epilogue threads of both CTAs
        }
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc_pair<512>(sm.tmem_ptr);
    tc_fence_before_sync();
    cluster_sync_all();                                         // barriers and TMEM of both CTAs exist from here on
    tc_fence_after_sync();
    const uint32_t tmem_base = *sm.tmem_ptr;
    if (FP4) {
        // kind::mxf4 multiplies every 32-code block by a ue8m0 scale factor read from tensor memory: all of them are 2^0.
        // Warps 4-7 of both CTAs fill columns [2 TN, 2 TN + 64) of their 128 lanes with 0x7F bytes; the MMAs read the A
        // scales from the first 32 of those columns and the B scales from the last 32, whatever the layout inside them.
        if (warp >= 4) {
            uint32_t ones[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) ones[i] = 0x7F7F7F7Fu;
            const uint32_t lanes = tmem_base + ((uint32_t)(32 * (warp & 3)) << 16);
            tmem_st_32x32b_x32(lanes + 2 * TN, ones);
            tmem_st_32x32b_x32(lanes + 2 * TN + 32, ones);
            tmem_st_wait();
        }
        tc_fence_before_sync();
        cluster_sync_all();                                     // the leader's MMAs read the scales of both CTAs
        tc_fence_after_sync();
    }

    if (warp == 0) {
        // ===================================================== TMA producer (both CTAs)
        if (elect_one()) {
            prefetch_tensormap(&map_q);
            prefetch_tensormap(&map_k);
            uint32_t it = 0;
            for (int64_t kt = kt_begin; kt < kt_end; kt += NT) {
                for (int kb = 0; kb < n_kb; ++kb, ++it) {
                    const int s = it % stages;
                    const uint32_t ph = (it / stages) & 1;
                    mbar_wait(&sm.empty[s], ph ^ 1);            // own slot freed by the pair's MMAs (multicast commit)
                    const uint32_t full_leader = mapa_u32(smem_u32(&sm.full[s]), lead_rank);
                    if (KNN_ABLATE(1)) {
                        mbar_arrive_cluster(full_leader);
                        continue;
                    }
                    if (leader) mbar_arrive_expect_tx(&sm.full[s], 2 * TSTAGE);
                    else        mbar_arrive_cluster(full_leader);
                    uint8_t* a = sm.tiles + (size_t)s * TSTAGE;
                    tma_load_2d_pair(a, &map_q, full_leader, kb * BKB, q_tile * BM);
#pragma unroll
                    for (int t = 0; t < NT; ++t) {              // (a key tile past the end reads as zeros and is skipped by the epilogue)
                        uint8_t* bdst = a + A_BYTES + t * TB_HALF;
                        if (quad)
                            tma_load_2d_pair_multicast(bdst + twin * (TB_HALF / 2), &map_k, &sm.full[s], kb * BKB,
                                                       (int32_t)((kt + t) * TN + rank * (TN / 2) + twin * (TN / 4)),
                                                       (uint16_t)((1u << rank) | (1u << (rank + 2))));
                        else
                            tma_load_2d_pair(bdst, &map_k, full_leader, kb * BKB, (int32_t)((kt + t) * TN + rank * (TN / 2)));
                    }
                }
            }
        }
    } else if (warp == 1 && leader) {
        // ===================================================== MMA issuer (one thread of the leader CTA)
        if (elect_one()) {
            constexpr uint32_t idesc = FP4 ? make_idesc_mxf4(2 * BM, TN) : make_idesc(/*S32*/ 2, /*int8*/ 1, 0, 0, 2 * BM, TN);
            const uint32_t tsf = tmem_base + 2 * TN;              // FP4: 64 columns of constant scale factors (see below)
            uint32_t it = 0, tile = 0;
            for (int64_t kt = kt_begin; kt < kt_end; kt += NT, ++tile) {
                // NT = 1: accumulator b = tile parity; NT = 2: both accumulators belong to this pass (barrier pair 0 only)
                const uint32_t b = NT == 1 ? (tile & 1) : 0u, tph = NT == 1 ? ((tile >> 1) & 1) : (tile & 1);
                mbar_wait(&sm.tempty[b], tph ^ 1);              // both epilogues have drained the accumulator(s)
                tc_fence_after_sync();
                const uint32_t d_tmem = tmem_base + b * TN;
                for (int kb = 0; kb < n_kb; ++kb, ++it) {
                    const int s = it % stages;
                    const uint32_t ph = (it / stages) & 1;
                    mbar_wait(&sm.full[s], ph);                 // the bytes of both CTAs have landed
                    tc_fence_after_sync();
                    const uint32_t a_addr = smem_u32(sm.tiles + (size_t)s * TSTAGE);
                    const uint32_t b_addr = a_addr + A_BYTES;
                    if (KNN_ABLATE(2)) {
                        for (uint32_t c = 0; c < (quad ? 4u : 2u); ++c) mbar_arrive_cluster(mapa_u32(smem_u32(&sm.empty[s]), c));
                        continue;
                    }
#pragma unroll
                    for (int t = 0; t < NT; ++t) {
#pragma unroll
                        for (int k4 = 0; k4 < BKB / 32; ++k4) {
                            const uint64_t da = make_smem_desc(a_addr + k4 * 32, 16, 1024);
                            const uint64_t db = make_smem_desc(b_addr + t * TB_HALF + k4 * 32, 16, 1024);
                            if (FP4) mma_mxf4_ss_pair(d_tmem + t * TN, da, db, idesc, tsf, tsf + 32, (kb > 0 || k4 > 0) ? 1u : 0u);
                            else     mma_i8_ss_pair(d_tmem + t * TN, da, db, idesc, (kb > 0 || k4 > 0) ? 1u : 0u);
                        }
                    }
                    // (releasing slots two at a time to halve the commits was measured: 9.3 -> 10.9 ms, the loads need the slot early)
                    tc_commit_pair(&sm.empty[s], all_mask);     // frees the slot in both CTAs (and tells the twin pair in quad mode)
                }
                tc_commit_pair(&sm.tfull[b], pair_mask);        // accumulator(s) complete, both epilogues may read
            }
        }
    } else if (warp >= 4) {
        // ===================================================== epilogue: distances + selection (both CTAs, own rows)
        const int e = threadIdx.x - 128;
        const int wq = warp & 3;
        const int64_t row = (int64_t)q_tile * BM + e;
        const bool valid = row < n_q;
        const int64_t gi = q_off + row;
        RowSelector sel;
        RowSelectorU64 selu;
        if (HAS_EXTRA) sel.init(sm.list_d, sm.list_i, BM, e, k);
        else           selu.init(sm.list_d, BM, e, k);
        long long sqq = 0, thr_ll = LLONG_MAX;
        double eq_r[KNN_MAX_EXTRA];
        if (valid) {
            sqq = sq_q[row];
            if (HAS_EXTRA)
                for (int c = 0; c < n_extra; ++c) eq_r[c] = extra_q[row * n_extra + c];
        }
        uint32_t tile = 0;
        for (int64_t kt = kt_begin; kt < kt_end; kt += NT, ++tile) {
            const uint32_t b = NT == 1 ? (tile & 1) : 0u, tph = NT == 1 ? ((tile >> 1) & 1) : (tile & 1);
            if (NT > 1) named_bar_sync(1, EPI_THREADS);         // both norm buffers are reused every pass: wait for the slowest reader
#pragma unroll
            for (int u = 0; u < NT; ++u) {                      // squared norms (and side columns) of the pass's key rows
                const int64_t k0 = (kt + u) * TN;
                long long* sqk_b = sm.sqk + (b + u) * TN;
                double* ek_b = sm.ek + (size_t)(b + u) * TN * KNN_MAX_EXTRA;
                for (int t = e; t < TN; t += EPI_THREADS) {
                    const int64_t gk = k0 + t;
                    sqk_b[t] = (gk < n_k) ? sq_k[gk] : 0;
                    if (HAS_EXTRA)
                        for (int c = 0; c < n_extra; ++c) ek_b[t * KNN_MAX_EXTRA + c] = (gk < n_k) ? extra_k[gk * n_extra + c] : 0.0;
                }
            }
            named_bar_sync(1, EPI_THREADS);
            mbar_wait(&sm.tfull[b], tph);
            tc_fence_after_sync();
            if (KNN_ABLATE(4)) {                                  // measurement only: no accumulator reads, no selection
                if (leader) mbar_arrive(&sm.tempty[b]);
                else        mbar_arrive_cluster(mapa_u32(smem_u32(&sm.tempty[b]), lead_rank));
                continue;
            }
#pragma unroll
            for (int u = 0; u < NT; ++u) {
                const int64_t k0 = (kt + u) * TN;
                if (k0 >= n_k || kt + u >= kt_end) break;       // the odd tile of the last pass
                const int ncols = (int)((n_k - k0) < TN ? (n_k - k0) : TN);
                scan_accumulator<HAS_EXTRA, TN, FP4>(tmem_base + ((uint32_t)(32 * wq) << 16) + (b + u) * TN, valid, sqq,
                                                     sm.sqk + (b + u) * TN, sm.ek + (size_t)(b + u) * TN * KNN_MAX_EXTRA, eq_r, n_extra,
                                                     ncols, gi, k_off + k0, sel, thr_ll, selu);
            }
            tc_fence_before_sync();
            if (leader) mbar_arrive(&sm.tempty[b]);
            else        mbar_arrive_cluster(mapa_u32(smem_u32(&sm.tempty[b]), lead_rank));
        }
        if (valid) {
            double* od = part_d2 + ((int64_t)split * n_q + row) * k;
            int32_t* oi = part_idx + ((int64_t)split * n_q + row) * k;
            if (!HAS_EXTRA) selu.sort();
            for (int t = 0; t < k; ++t) {
                od[t] = HAS_EXTRA ? sel.d(t) : selu.d(t);
                oi[t] = HAS_EXTRA ? sel.i(t) : selu.i(t);
            }
        }
    }
    tc_fence_before_sync();
    cluster_sync_all();                                         // nobody leaves while the peer can still signal its barriers
    if (warp == 2) tmem_dealloc_pair<512>(tmem_base);
}

bool knn_tc_supported(int64_t ldq, int64_t ldk, const void* Q, const void* K, int k) {
    return ldq % 16 == 0 && ldk % 16 == 0 && (uintptr_t)Q % 16 == 0 && (uintptr_t)K % 16 == 0 && k >= 1 && k <= 64;
}

int knn_tc_splits(int64_t n_q, int64_t n_k) {
    int64_t q_tiles = (n_q + BM - 1) / BM, k_tiles = (n_k + BN - 1) / BN;      // (the fp4 kernel has a few more, narrower key tiles: s <= k_tiles holds for it too)
    int64_t s;
    static int forced = -1;                                     // GNNB_KNN_SPLITS: tuning knob (any value gives the same lists)
    if (forced < 0) { const char* e = getenv("GNNB_KNN_SPLITS"); forced = e ? atoi(e) : 0; }
    // Every key split keeps its own list per query row, and every list pays its own warm-up of insertions (a whole warp
    // stalls ~150 cycles whenever one of its rows inserts), so fewer splits are cheaper -- as long as one wave of CTA pairs
    // fills the GPU (66 pairs = 33 co-resident clusters of four).  Measured, FP4 kernel, ms at splits = 1 / 2 / 4 / 8:
    // 15 625 rows 5.26 / 5.39 / 5.94 / 6.97; 3 534 rows 0.99 / 0.55 / 0.34 / 0.41; 1 814 rows - / 0.18 / 0.14 / 0.12.  From a few
    // waves on, 8 splits win again (90 000 rows: 190 / 181 / - / 178) because a group of 8 query pairs x 8 key ranges keeps
    // its operands in L2.
    const int64_t q_pairs = (q_tiles + 1) / 2, wave = 66;
    if (forced > 0) s = forced;
    else if (q_pairs >= 2 * wave) s = 8;
    else s = (wave + q_pairs / 2) / q_pairs;                    // round(66 / q_pairs)
    if (s > KNN_MAX_SPLITS) s = KNN_MAX_SPLITS;
    if (s > k_tiles) s = k_tiles;
    if (s < 1) s = 1;
    return (int)s;
}

int knn_tc_launch(const int8_t* Q, int64_t ldq, const int64_t* sq_q, int64_t n_q, int64_t q_off,
                  const int8_t* K, int64_t ldk, const int64_t* sq_k, int64_t n_k, int64_t k_off, int64_t p,
                  const double* extra_q, const double* extra_k, int n_extra, int k, int n_splits,
                  double* part_d2, int32_t* part_idx, cudaStream_t stream, int fp4) {
    // fp4 != 0: Q and K are e2m1-packed rows (gnnb_pack_fp4), ldq / ldk in bytes; p stays the marker count
    const int q_tiles = (int)((n_q + BM - 1) / BM);
    static int use_pair = -1, ablate = 0, quad = 1;
    if (use_pair < 0) {
        const char* e = getenv("GNNB_KNN_PAIR");
        use_pair = e ? atoi(e) : 1;
#ifdef GNNB_ABLATION
        const char* a = getenv("GNNB_KNN_ABLATE");             // measurement builds only: 1 = no TMA loads, 2 = no MMAs, 4 = no epilogue
        ablate = a ? atoi(a) : 0;
#endif
        const char* q = getenv("GNNB_KNN_QUAD");
        quad = q ? atoi(q) : -1;                               // -1: pick by shape
    }
    const bool extra = n_extra > 0;
    // quads pay off on long K loops (15625 x 50000: 9.8 -> 9.3 ms); on short ones the lockstep of the two pairs costs more
    // than the shared key tile saves (65536 x 8192: 60 -> 65 ms)
    const int use_quad = quad >= 0 ? quad : (p >= 16384 ? 1 : 0);
    if (fp4 && !(use_pair && q_tiles >= 2))
        return fail(GNNB_E_UNSUPPORTED, "knn_fp4: needs at least 129 query rows (CTA-pair kernel)");
    if (use_pair && q_tiles >= 2) {
        // CTA pairs: 256 queries x 256 (fp4: 224) keys per pair, each CTA stages its 128 queries and one half of the key tile
        const int tn = fp4 ? FP4_TN : BN;
        const int64_t k_bytes = fp4 ? (p + 1) / 2 : p;            // bytes of one operand row = extent of the contraction in memory
        CUtensorMap map_q, map_k;
        int rc = make_tensor_map_2d(&map_q, Q, 1, (uint64_t)k_bytes, (uint64_t)n_q, (uint64_t)ldq, BKB, BM);
        if (rc != GNNB_OK) return rc;
        rc = make_tensor_map_2d(&map_k, K, 1, (uint64_t)k_bytes, (uint64_t)n_k, (uint64_t)ldk, BKB, use_quad ? tn / 4 : tn / 2);
        if (rc != GNNB_OK) return rc;
        const int stage_bytes = (fp4 && FP4_NT > 1) ? A_BYTES + FP4_NT * (FP4_TN / 2) * BKB : STAGE2_BYTES;
        int stages = MAX_STAGES;
        while (stages > 2 && smem_bytes2(stages, k, extra, stage_bytes) > 232448) --stages;
        const size_t smem = smem_bytes2(stages, k, extra, stage_bytes);
        const int q_pairs = (q_tiles + 1) / 2;
        const int groups = (q_pairs + QGP - 1) / QGP;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)(groups * QGP * n_splits * 2));
        cfg.blockDim = dim3(256);
        cfg.dynamicSmemBytes = smem;
        cfg.stream = stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = use_quad ? 4 : 2;             // grid = groups * QGP * n_splits * 2 is a multiple of 4 (QGP even)
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
#define KNN_LAUNCH2(EXTRA_, TN_, FP4_)                                                                                              \
    do {                                                                                                                           \
        GNNB_SET_DYN_SMEM((knn_tc2_kernel<EXTRA_, TN_, FP4_, (FP4_ ? FP4_NT : 1)>), smem);                                         \
        GNNB_CUDA(cudaLaunchKernelEx(&cfg, knn_tc2_kernel<EXTRA_, TN_, FP4_, (FP4_ ? FP4_NT : 1)>, map_q, map_k, sq_q, n_q, q_off, sq_k, n_k, k_off, k_bytes, \
                                     extra_q, extra_k, n_extra, k, n_splits, q_pairs, stages, part_d2, part_idx, use_quad, ablate)); \
    } while (0)
        if (fp4) {
            if (extra) KNN_LAUNCH2(true, FP4_TN, true);
            else       KNN_LAUNCH2(false, FP4_TN, true);
        } else {
            if (extra) KNN_LAUNCH2(true, BN, false);
            else       KNN_LAUNCH2(false, BN, false);
        }
#undef KNN_LAUNCH2
        return GNNB_OK;
    }
    CUtensorMap map_q, map_k;
    int rc = make_tensor_map_2d(&map_q, Q, 1, (uint64_t)p, (uint64_t)n_q, (uint64_t)ldq, BKB, BM);
    if (rc != GNNB_OK) return rc;
    rc = make_tensor_map_2d(&map_k, K, 1, (uint64_t)p, (uint64_t)n_k, (uint64_t)ldk, BKB, BN);
    if (rc != GNNB_OK) return rc;
    const int stages = k <= 32 ? 3 : 2;
    const size_t smem = smem_bytes(stages, k, extra);
    const int groups = (q_tiles + QG - 1) / QG;
    const unsigned grid = (unsigned)(groups * QG * n_splits);
    if (extra) {
        GNNB_SET_DYN_SMEM(knn_tc_kernel<true>, smem);
        knn_tc_kernel<true><<<grid, 256, smem, stream>>>(map_q, map_k, sq_q, n_q, q_off, sq_k, n_k, k_off, p, extra_q, extra_k,
                                                         n_extra, k, n_splits, q_tiles, stages, part_d2, part_idx);
    } else {
        GNNB_SET_DYN_SMEM(knn_tc_kernel<false>, smem);
        knn_tc_kernel<false><<<grid, 256, smem, stream>>>(map_q, map_k, sq_q, n_q, q_off, sq_k, n_k, k_off, p, extra_q, extra_k,
                                                          n_extra, k, n_splits, q_tiles, stages, part_d2, part_idx);
    }
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

}  // namespace gnnb
