// K5 on the 5th-generation tensor cores: the marker projection H = X . W^T and its weight gradient
// dW = dH^T . X for int8 genotype codes X, fp32 W / dH, fp32 accumulation in TMEM.
//
// Replaces GCNConv.lin (PyG Linear(bias=False) -> torch.matmul, and its autograd) for the first layer, where
// p = #markers is large; call sites gnn/networks/networks.py:271-272 and the other GCN factories.
//
// Precision: the genotype codes (0..127) are exact in bf16.  The fp32 operand (W, or dH) is split into T bf16
// terms (hi, mid, lo) whose products with X are accumulated in the same fp32 TMEM accumulator:
// T = 1 is a plain bf16 projection, T = 2 carries ~16 mantissa bits (rel. error ~1e-5), T = 3 is fp32-faithful.
//
// X stays int8 in HBM (1 byte per element, the only large stream).  Per k-block a TMA load brings the raw int8
// tile into a staging ring; converter warps expand it to bf16 (prmt + one packed bf16 subtract per pair) and write
// it with tcgen05.st into an A slot in TENSOR MEMORY; the fp32 operand's bf16 terms arrive by TMA from an L2-resident
// workspace that a small prep kernel fills once per call.  (A first version kept both operands in shared memory and ran
// at 0.46 ms against 0.33 ms; its history is in profiles/README.md.)  These kernels serve precision 1..3; the default
// for int8 genotypes is the int8-digit kernel of linear_i8.cu (precision 4).
//
//   forward  : CTA = 128 rows of X.     A = X tile (TMEM),     B = W terms  [h_pad x 64]  (K-major),  D[128 x h_pad]
//   backward : CTA = 128 columns of X.  A = X^T tile (TMEM),   B = dH terms [64 x 128]    (MN-major), D[128 x 128]
//
// warps: 0 = TMA producer (X), 3 = TMA producer (terms), 1 = MMA issuer, 2 = TMEM allocator,
//        4.. = converters, then epilogue (TMEM -> registers -> global).
#include <cuda_bf16.h>
#include <stdlib.h>

#include "tc_common.cuh"

namespace gnnb {

using namespace tc;

namespace {
constexpr int BM = 128;                 // rows (fwd) / columns (bwd) of X per CTA == UMMA M
constexpr int BK = 64;                  // k elements per block
constexpr int X_STAGE_BYTES = BM * BK;              // int8: fwd 128 rows x 64 B, bwd 64 rows x 128 B
constexpr int MAX_TERMS = 3;
constexpr size_t SMEM_LIMIT = 232448;       // 227 KB of dynamic shared memory per CTA

// fp32 -> up to three bf16 terms: x ~ hi + mid + lo
__device__ __forceinline__ void split3(float x, __nv_bfloat16& hi, __nv_bfloat16& mid, __nv_bfloat16& lo) {
    hi = __float2bfloat16_rn(x);
    float r = x - __bfloat162float(hi);
    mid = __float2bfloat16_rn(r);
    lo = __float2bfloat16_rn(r - __bfloat162float(mid));
}

// out[t][r][c] (t < terms, r < rows_out, c < cols_out, row stride ld_out) = term t of in[r][c], zero outside [rows, cols]
__global__ void split_terms_kernel(const float* __restrict__ in, int64_t ld_in, int64_t rows, int64_t cols,
                                   __nv_bfloat16* __restrict__ out, int64_t rows_out, int64_t cols_out, int64_t ld_out, int terms) {
    int64_t total = rows_out * cols_out;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        int64_t r = i / cols_out, c = i % cols_out;
        float x = (r < rows && c < cols) ? in[r * ld_in + c] : 0.f;
        __nv_bfloat16 t0, t1, t2;
        split3(x, t0, t1, t2);
        out[r * ld_out + c] = t0;
        if (terms > 1) out[rows_out * ld_out + r * ld_out + c] = t1;
        if (terms > 2) out[2 * rows_out * ld_out + r * ld_out + c] = t2;
    }
}

// 16 int8 codes in [0, 127] -> 16 bf16, exactly: bytes are spliced under the exponent of 128.0 (0x4300 | b = 128 + b)
// and 128.0 is subtracted with one packed bf16 op per pair.
__device__ __forceinline__ void cvt16(const uint4& in, uint4& lo, uint4& hi) {
    const __nv_bfloat162 k128 = __halves2bfloat162(__ushort_as_bfloat16((unsigned short)0x4300), __ushort_as_bfloat16((unsigned short)0x4300));
    const uint32_t w[4] = {in.x, in.y, in.z, in.w};
    uint32_t o[8];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        uint32_t p0 = __byte_perm(w[i], 0x43434343u, 0x4140);      // {b0, 0x43, b1, 0x43}
        uint32_t p1 = __byte_perm(w[i], 0x43434343u, 0x4342);      // {b2, 0x43, b3, 0x43}
        __nv_bfloat162 q0 = __hsub2(*reinterpret_cast<__nv_bfloat162*>(&p0), k128);
        __nv_bfloat162 q1 = __hsub2(*reinterpret_cast<__nv_bfloat162*>(&p1), k128);
        o[2 * i] = *reinterpret_cast<uint32_t*>(&q0);
        o[2 * i + 1] = *reinterpret_cast<uint32_t*>(&q1);
    }
    lo = make_uint4(o[0], o[1], o[2], o[3]);
    hi = make_uint4(o[4], o[5], o[6], o[7]);
}
}  // namespace

// ------------------------------------------------------------------------------------------------------------------
// Projection kernels with the A operand (the X tile) in TENSOR MEMORY.
// The converter warps write the bf16 X tile straight into TMEM (tcgen05.st), not into shared memory: the tile never
// costs shared-memory bandwidth again (UMMA reads only the term tiles from shared memory), eight A slots fit beside
// the accumulator (8 x 32 columns), and the shared memory that the bf16 A ring used goes to deeper TMA rings.
//   TMEM columns [0, 256): accumulator (hi | lo halves), [256, 512): 8 A slots of 32 columns (64 bf16 per row)
// Thread = row of the tile = TMEM lane, so all four converter warps work on every k-block (a warp can only touch its
// own 32 lanes).  The int8 staging tile uses the 64-byte TMA swizzle so that 32 rows can be read without bank conflicts.
namespace {
#ifndef TA_NX_
#define TA_NX_ 6
#endif
#ifndef TA_NB_
#define TA_NB_ 8
#endif
constexpr int TA_NX = TA_NX_, TA_NA = 8, TA_NB_MAX = TA_NB_;
#ifndef TA_CW
#define TA_CW 2
#endif
// Converter warps per TMEM lane quarter.  One warp needs ~600 cycles per k-block (two barrier waits, the shared-memory
// reads, the int8 -> bf16 arithmetic, tcgen05.st and its wait, two arrivals -- one dependent chain), more than the
// ~450 cycles of tensor work in a k-block at T = 2; TA_CW warps take k-blocks round-robin.
constexpr int TA_THREADS = 128 + 128 * TA_CW;
// One "full" and one "empty" barrier per k-block slot (ring of TA_NA): `full` collects the four converter warps AND the
// TMA bytes of the term tiles, `empty` is one tcgen05.commit that both the converters and the term producer watch.
// The MMA-issuing thread therefore pays one wait and one commit per k-block (a hop over an mbarrier costs ~230 cycles).
struct TaBars {
    uint64_t x_full[TA_NX], x_empty[TA_NX], full[TA_NA], empty[TA_NA], acc_full;
    uint32_t tmem_ptr;
};
__host__ __device__ inline size_t ta_fixed() { return 1024 + (size_t)TA_NX * X_STAGE_BYTES + sizeof(TaBars) + 64; }
__host__ __device__ inline int ta_nb(int terms, int b_tile_bytes) {
    int nb = (int)((SMEM_LIMIT - ta_fixed()) / ((size_t)terms * b_tile_bytes));
    return nb > TA_NB_MAX ? TA_NB_MAX : nb;
}
}  // namespace

template <bool BWD>
__global__ void __launch_bounds__(TA_THREADS, 1)
linear_ta_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_b,
                 int64_t k_total, int terms, int n_pad, int64_t b_term_rows,
                 float* __restrict__ out, int64_t ld_out, int64_t m_total, int n_valid) {
    extern __shared__ uint8_t smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint8_t* base = smem_raw + ((1024 - (smem_u32(smem_raw) & 1023)) & 1023);
    const int b_tile_bytes = BWD ? (BK * 128 * 2) : (n_pad * BK * 2);
    const int NB = ta_nb(terms, b_tile_bytes);
    uint8_t* x_stage = base;
    uint8_t* b_tiles = x_stage + TA_NX * X_STAGE_BYTES;
    TaBars* bars = reinterpret_cast<TaBars*>(b_tiles + (size_t)NB * terms * b_tile_bytes);
    const int n_kb = (int)((k_total + BK - 1) / BK);
    const int64_t m0 = (int64_t)blockIdx.x * BM;

    if (threadIdx.x == 0) {
        for (int s = 0; s < TA_NX; ++s) { mbar_init(&bars->x_full[s], 1); mbar_init(&bars->x_empty[s], 4); }
        for (int s = 0; s < TA_NA; ++s) { mbar_init(&bars->full[s], 5); mbar_init(&bars->empty[s], 1); }
        mbar_init(&bars->acc_full, 1);
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc<512>(&bars->tmem_ptr);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = bars->tmem_ptr;
    const bool paired = terms >= 2;

    if (warp == 0) {
        if (elect_one()) {                                     // TMA: raw int8 tiles of X, [128 rows x 64 B], 64-byte swizzle
            prefetch_tensormap(&map_x);
            for (int u = 0; u < n_kb; ++u) {
                const int s = u % TA_NX;
                mbar_wait(&bars->x_empty[s], ((u / TA_NX) & 1) ^ 1);
                mbar_arrive_expect_tx(&bars->x_full[s], X_STAGE_BYTES);
                if (BWD) tma_load_2d(x_stage + s * X_STAGE_BYTES, &map_x, &bars->x_full[s], (int32_t)m0, u * BK);   // [64 k-rows x 128 markers]
                else     tma_load_2d(x_stage + s * X_STAGE_BYTES, &map_x, &bars->x_full[s], u * BK, (int32_t)m0);   // [128 rows x 64 k]
            }
        }
    } else if (warp == 3) {
        if (elect_one()) {                                     // TMA: bf16 terms of W, [n_pad rows x 64 k] each
            prefetch_tensormap(&map_b);
            for (int u = 0; u < n_kb; ++u) {
                const int s = u % NB, f = u % TA_NA;
                if (u >= NB) {                                 // term slot s was last read by k-block u - NB
                    const int v = u - NB;
                    mbar_wait(&bars->empty[v % TA_NA], (v / TA_NA) & 1);
                }
                mbar_arrive_expect_tx(&bars->full[f], (uint32_t)(terms * b_tile_bytes));
                uint8_t* dst = b_tiles + (size_t)s * terms * b_tile_bytes;
                for (int t = 0; t < terms; ++t) {
                    if (BWD) {   // dH terms: two 64-feature groups of [64 k-rows x 128 B]
                        const int32_t grow = (int32_t)(t * b_term_rows + (int64_t)u * BK);
                        tma_load_2d(dst + t * b_tile_bytes, &map_b, &bars->full[f], 0, grow);
                        tma_load_2d(dst + t * b_tile_bytes + BK * 128, &map_b, &bars->full[f], 64, grow);
                    } else {
                        tma_load_2d(dst + t * b_tile_bytes, &map_b, &bars->full[f], u * BK, t * n_pad);
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {                                     // MMA issue: A from TMEM, B from shared memory
            // A comes from TMEM (rows = lanes, k along columns); B is K-major (forward) or MN-major (backward) in shared memory
            const uint32_t idesc1 = make_idesc(1, 1, 0, BWD ? 1 : 0, BM, (uint32_t)n_pad);
            const uint32_t idesc2 = make_idesc(1, 1, 0, BWD ? 1 : 0, BM, (uint32_t)(2 * n_pad));
            for (int u = 0; u < n_kb; ++u) {
                const int sa = u % TA_NA, sb = u % NB;
                mbar_wait(&bars->full[sa], (u / TA_NA) & 1);
                tc_fence_after_sync();
                const uint32_t a_tmem = tmem_base + 256 + (uint32_t)sa * 32;
                const uint32_t b_addr = smem_u32(b_tiles + (size_t)sb * terms * b_tile_bytes);
                const uint64_t db0 = BWD ? make_smem_desc(b_addr, BK * 128, 1024) : make_smem_desc(b_addr, 16, 1024);
                const uint64_t term_step = (uint64_t)(b_tile_bytes >> 4);
                const uint64_t k_step = BWD ? (2048 >> 4) : (32 >> 4);      // 16 k: two 8-row atoms (MN-major) / 32 bytes (K-major)
#pragma unroll
                for (int k16 = 0; k16 < BK / 16; ++k16)        // hi (+ lo) terms as one operand of N = n_pad or 2 n_pad
                    mma_bf16_ts(tmem_base, a_tmem + k16 * 8, db0 + k16 * k_step, paired ? idesc2 : idesc1, (u > 0 || k16 > 0) ? 1u : 0u);
                if (terms == 3) {
#pragma unroll
                    for (int k16 = 0; k16 < BK / 16; ++k16)
                        mma_bf16_ts(tmem_base, a_tmem + k16 * 8, db0 + 2 * term_step + k16 * k_step, idesc1, 1u);
                }
                tc_commit(&bars->empty[sa]);
            }
            tc_commit(&bars->acc_full);
        }
    } else if (warp >= 4) {
        // ===================================================== converter: int8 staging (smem) -> bf16 A slot (TMEM)
        const int wq = warp & 3, turn = (warp - 4) >> 2;       // lane quarter; which k-blocks (u % TA_CW == turn) are this warp's
        const int r = 32 * wq + lane;                          // row of the tile == TMEM lane
        const uint32_t sw = (uint32_t)((r >> 1) & 3);          // 64-byte swizzle: 16-byte chunk index ^= (row / 2) % 4
        for (int u = turn; u < n_kb; u += TA_CW) {
            const int sx = u % TA_NX, sa = u % TA_NA;
            mbar_wait(&bars->x_full[sx], (u / TA_NX) & 1);
            mbar_wait(&bars->empty[sa], ((u / TA_NA) & 1) ^ 1);
            tc_fence_after_sync();
            uint32_t v[32];
            if (BWD) {
                // staging [64 k-rows][128 markers]: this thread owns marker column r, i.e. one byte of every row (a warp reads
                // 32 consecutive bytes per row: one wavefront, no conflicts); pairs of k become one bf16x2 column of the A slot
                const uint32_t col_s = smem_u32(x_stage + sx * X_STAGE_BYTES) + (uint32_t)r;
                const __nv_bfloat162 k128 = __halves2bfloat162(__ushort_as_bfloat16((unsigned short)0x4300), __ushort_as_bfloat16((unsigned short)0x4300));
                uint32_t bytes[64];
#pragma unroll
                for (int kk = 0; kk < 64; ++kk)
                    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(bytes[kk]) : "r"(col_s + kk * 128));
#pragma unroll
                for (int q = 0; q < 32; ++q) {
                    uint32_t w = __byte_perm(bytes[2 * q], bytes[2 * q + 1], 0x5410) | 0x43004300u;     // {b_k, 0x43, b_k+1, 0x43}
                    __nv_bfloat162 h2 = __hsub2(*reinterpret_cast<__nv_bfloat162*>(&w), k128);
                    v[q] = *reinterpret_cast<uint32_t*>(&h2);
                }
            } else {
                const uint32_t row_s = smem_u32(x_stage + sx * X_STAGE_BYTES) + (uint32_t)r * 64;
                uint4 in[4];
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                                 : "=r"(in[c].x), "=r"(in[c].y), "=r"(in[c].z), "=r"(in[c].w) : "r"(row_s + (((uint32_t)c ^ sw) << 4)));
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    uint4 lo, hi;
                    cvt16(in[c], lo, hi);
                    v[8 * c + 0] = lo.x; v[8 * c + 1] = lo.y; v[8 * c + 2] = lo.z; v[8 * c + 3] = lo.w;
                    v[8 * c + 4] = hi.x; v[8 * c + 5] = hi.y; v[8 * c + 6] = hi.z; v[8 * c + 7] = hi.w;
                }
            }
            tmem_st_32x32b_x32(tmem_base + ((uint32_t)(32 * wq) << 16) + 256 + (uint32_t)sa * 32, v);
            tmem_st_wait();
            tc_fence_before_sync();
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(&bars->full[sa]);
                mbar_arrive(&bars->x_empty[sx]);
            }
        }
        // ===================================================== epilogue
        mbar_wait(&bars->acc_full, 0);
        tc_fence_after_sync();
        const int64_t m = m0 + r;
#pragma unroll 1
        for (int chunk = turn; chunk * 32 < n_pad; chunk += TA_CW) {      // the warps of a lane quarter share the column chunks
            uint32_t a[32], b[32];
            const uint32_t taddr = tmem_base + ((uint32_t)(32 * wq) << 16) + chunk * 32;
            tmem_ld_32x32b_x32(taddr, a);
            if (paired) tmem_ld_32x32b_x32(taddr + (uint32_t)n_pad, b);
            tmem_ld_wait();
            if (m < m_total) {
#pragma unroll
                for (int c = 0; c < 32; ++c) {
                    const int f = chunk * 32 + c;
                    if (f < n_valid) {
                        const float rr = paired ? __uint_as_float(a[c]) + __uint_as_float(b[c]) : __uint_as_float(a[c]);
                        if (BWD) out[(int64_t)f * ld_out + m] = rr;      // dW[f, marker]: a warp writes 32 consecutive markers
                        else     out[m * ld_out + f] = rr;               // H[row, f]
                    }
                }
            }
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 2) tmem_dealloc<512>(tmem_base);
}

static bool tc_shape_ok(const void* X, int64_t ldx, int64_t h) {
    return (uintptr_t)X % 16 == 0 && ldx % 16 == 0 && h >= 1 && h <= 128;
}

static size_t terms_bytes_fwd(int64_t p, int64_t h, int terms) {
    return align_up((size_t)terms * align_up((size_t)h, 32) * align_up((size_t)p, 8) * 2, 1024);
}
static size_t terms_bytes_bwd(int64_t n, int terms) { return align_up((size_t)terms * (size_t)n * 128 * 2, 1024); }

size_t linear_tc_workspace(int64_t n, int64_t p, int64_t h, int terms) {
    if (h > 128 || terms < 1 || terms > MAX_TERMS) return 0;
    size_t fwd = terms_bytes_fwd(p, h, terms), bwd = terms_bytes_bwd(n, terms);          // bf16 terms of W / of dH
    return (fwd > bwd ? fwd : bwd) + 2048;
}

// simt fallbacks live in linear.cu
int linear_simt_fwd(const void* X, int x_is_i8, int64_t ldx, int64_t n, int64_t p, const float* W, int64_t ldw, int64_t h,
                    float* H, int64_t ldh, void* ws, size_t ws_bytes, cudaStream_t stream);
int linear_simt_bwd_w(const void* X, int x_is_i8, int64_t ldx, int64_t n, int64_t p, const float* dH, int64_t lddh, int64_t h,
                      float* dW, int64_t lddw, void* ws, size_t ws_bytes, cudaStream_t stream);

int linear_tc_fwd(const void* X, int x_is_i8, int64_t ldx, int64_t n, int64_t p, const float* W, int64_t ldw, int64_t h,
                  float* H, int64_t ldh, int terms, void* ws, size_t ws_bytes, cudaStream_t stream) {
    if (!x_is_i8 || !tc_shape_ok(X, ldx, h) || p < 64 || n < 1)
        return linear_simt_fwd(X, x_is_i8, ldx, n, p, W, ldw, h, H, ldh, ws, ws_bytes, stream);    // small / float inputs: exact fp32 tiles
    // h is padded to the UMMA granule of 16 (h = 100 -> hi | lo = one operand of N = 2 x 112)
    const int64_t p_ld = (int64_t)align_up((size_t)p, 8), h_pad = (int64_t)align_up((size_t)h, 16);
    const size_t need = linear_tc_workspace(n, p, h, terms);
    GNNB_REQUIRE(ws && ws_bytes >= need, GNNB_E_WORKSPACE, "linear_fwd: workspace %zu < %zu bytes", ws_bytes, need);
    uint8_t* wsb = reinterpret_cast<uint8_t*>(align_up((size_t)(uintptr_t)ws, 1024));
    __nv_bfloat16* Wt = reinterpret_cast<__nv_bfloat16*>(wsb);
    {
        int64_t total = h_pad * p_ld, blocks = (total + 255) / 256, cap = (int64_t)sm_count() * 16;
        split_terms_kernel<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, stream>>>(W, ldw, h, p, Wt, h_pad, p_ld, p_ld, terms);
        GNNB_LAUNCH_CHECK();
    }
    const int64_t tiles = (n + BM - 1) / BM;
    CUtensorMap mx, mb;
    int rc = make_tensor_map_2d(&mx, X, 1, (uint64_t)p, (uint64_t)n, (uint64_t)ldx, BK, BM, false, /*swizzle64*/ true);
    if (rc != GNNB_OK) return rc;
    rc = make_tensor_map_2d(&mb, Wt, 2, (uint64_t)p_ld, (uint64_t)(terms * h_pad), (uint64_t)p_ld * 2, BK, (uint32_t)h_pad, true);
    if (rc != GNNB_OK) return rc;
    const int btile = (int)(h_pad * BK * 2);
    const size_t smem_ta = ta_fixed() + (size_t)ta_nb(terms, btile) * terms * btile;
    GNNB_SET_DYN_SMEM(linear_ta_kernel<false>, smem_ta);
    linear_ta_kernel<false><<<(unsigned)tiles, TA_THREADS, smem_ta, stream>>>(mx, mb, p, terms, (int)h_pad, h_pad, H, ldh, n, (int)h);
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

int linear_tc_bwd_w(const void* X, int x_is_i8, int64_t ldx, int64_t n, int64_t p, const float* dH, int64_t lddh, int64_t h,
                    float* dW, int64_t lddw, int terms, void* ws, size_t ws_bytes, cudaStream_t stream) {
    if (!x_is_i8 || !tc_shape_ok(X, ldx, h) || p < 128 || n < 64)
        return linear_simt_bwd_w(X, x_is_i8, ldx, n, p, dH, lddh, h, dW, lddw, ws, ws_bytes, stream);
    const size_t need = linear_tc_workspace(n, p, h, terms);
    GNNB_REQUIRE(ws && ws_bytes >= need, GNNB_E_WORKSPACE, "linear_bwd_w: workspace %zu < %zu bytes", ws_bytes, need);
    uint8_t* wsb = reinterpret_cast<uint8_t*>(align_up((size_t)(uintptr_t)ws, 1024));
    __nv_bfloat16* Dt = reinterpret_cast<__nv_bfloat16*>(wsb);
    {
        int64_t total = n * 128, blocks = (total + 255) / 256, cap = (int64_t)sm_count() * 16;
        split_terms_kernel<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, stream>>>(dH, lddh, n, h, Dt, n, 128, 128, terms);
        GNNB_LAUNCH_CHECK();
    }
    const int64_t tiles = (p + BM - 1) / BM;
    CUtensorMap mx, mb;
    int rc = make_tensor_map_2d(&mx, X, 1, (uint64_t)p, (uint64_t)n, (uint64_t)ldx, 128, BK, /*swizzle*/ false);
    if (rc != GNNB_OK) return rc;
    rc = make_tensor_map_2d(&mb, Dt, 2, 128, (uint64_t)(terms * n), 256, 64, BK, true);
    if (rc != GNNB_OK) return rc;
    const int btile = BK * 128 * 2;
    const size_t smem_ta = ta_fixed() + (size_t)ta_nb(terms, btile) * terms * btile;
    GNNB_SET_DYN_SMEM(linear_ta_kernel<true>, smem_ta);
    linear_ta_kernel<true><<<(unsigned)tiles, TA_THREADS, smem_ta, stream>>>(mx, mb, n, terms, 128, n, dW, lddw, p, (int)h);
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

}  // namespace gnnb
