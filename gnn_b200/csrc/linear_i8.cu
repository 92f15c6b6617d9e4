// K5 on the int8 tensor path (precision 4): H = X . W^T and dW = dH^T . X with X as int8 genotype codes.
//
// The float operand (W for the forward pass, dH for the weight gradient) is quantised per output feature to a fixed-point
// integer q = round(w / s_f) and split into three balanced digits ("limbs", each an int8):
//     24-bit form:  |q| <= 8 355 711,  q = d0 + 256 d1 + 65536 d2,  d in [-128, 127],  s_f = max|w_f| / 8 355 711
//     21-bit form:  |q| <= 2^20,       q = d0 + 128 d1 + 16384 d2,  d in [-64, 64],    s_f = max|w_f| / 2^20
// Three int8 MMAs (tcgen05 kind::i8, exact int32 accumulation) give the three digit sums; the epilogue combines them
// as ONE exact 64-bit integer and rounds once: the result is the exact dot product of X with the quantised operand.
// The form is chosen per call from the length K of the contraction and the largest code the X format admits, so that the
// int32 digit sums cannot overflow: 24 bits when K * cmax * 128 < 2^31, else 21 bits when K * cmax * 64 < 2^31, else the
// shape is refused (cmax = 3 for 2-bit packed X, 128 for int8 X).  24 bits relative to the feature's largest element is
// the resolution of the float32 operand itself at that magnitude, i.e. the rounding the reference's own fp32 matmul
// applies to its largest terms.  NaN / Inf in the float operand give NaN in every output of that feature.
// Against the split-bf16 kernels of linear_tc.cu: the 8-bit pipe retires ~2.5x the MACs per clock of the bf16 pipe on
// B200 (scripts/ubench/mma_rate.cu: 7080 vs 2850 MAC/clk/SM), X goes from TMA straight into the MMA (no int8 -> bf16
// converter, no TMEM staging), and three int8 digits beat the 16 bits of two bf16 terms.
//
// Kernel = the CTA-pair pipeline of knn_tc.cu (cta_group::2): a pair owns 256 rows of the M side, each CTA stages its
// own 128-row A tile and HALF of the limb tile; the leader issues, per 128-byte k-block, 4 x [N = 256 + N = 64] (the
// three digit blocks at a stride of 104 rows, see "Digit layout" below).  The kernel is bound by the L2 -> SM operand stream, and the limb tile is the same for every pair:
// clusters of 4 (two pairs, GNNB_LINEAR_QUAD, default on) fetch each limb half-tile once -- each of the two CTAs that
// need it issues half of it as a TMA multicast to both -- and run the two pairs in lockstep per ring slot (the slot's
// `empty` barrier collects the commits of both leaders).  Measured 0.240 -> 0.226 ms forward, 0.245 -> 0.224 backward.
//   forward : M = individuals (A = X tile, K-major),             K = markers,     B = W limbs    [320, p]  K-major
//   backward: M = markers     (A = X tile read MN-major),        K = individuals, B = dH^T limbs [320, n]  K-major
#include <math.h>

#include "tc_common.cuh"

namespace gnnb {

using namespace tc;

// Ablation switches (timing experiments that produce INVALID results) exist only in builds made with -DGNNB_ABLATION
// (scripts/build_variant.sh); the product library has none.
#ifdef GNNB_ABLATION
#define LI_ABLATE(bit) (ablate & (bit))
#else
#define LI_ABLATE(bit) false
#endif

namespace {
constexpr int BM = 128, BKB = 128;                       // A tile: 128 rows (or 128 markers) x 128 bytes of K
// Digit layout.  An MMA costs ~45 + 0.4 N cycles on this pipe (scripts/ubench/mma_rate.cu), so the three digit blocks are
// packed at a stride of 104 rows into 320 rows = one MMA of N = 256 plus one of N = 64 per k-step, whose accumulators are
// adjacent TMEM columns (digit d of feature f = row / column d * 104 + f).  LI_COMPACT = 0 is the first layout (112-row
// blocks for digits 0 and 1 in an N = 224 MMA, digit 2 alone in an N = 128 MMA: 352 rows).
#ifndef LI_COMPACT
#define LI_COMPACT 1
#endif
#if LI_COMPACT
constexpr int HP = 104;                                  // rows per digit (h <= 104)
constexpr int N0 = 256, N1 = 64;                         // the two MMAs of a k-step
constexpr int D2_ROW = 2 * HP;                           // first row / accumulator column of digit 2
#else
constexpr int HP = 112;
constexpr int N0 = 2 * HP, N1 = 128;
constexpr int D2_ROW = N0;
#endif
constexpr int LIMB_ROWS = N0 + N1;                       // rows of the limb matrix
constexpr int F_ROWS = LIMB_ROWS - D2_ROW;               // features slots the quantisers fill (padding rows are zeroed)
constexpr int A_BYTES = BM * BKB, B0_HALF = (N0 / 2) * BKB, B1_HALF = (N1 / 2) * BKB;
constexpr int B_BYTES = B0_HALF + B1_HALF;               // 22 KB: this CTA's half of the digit tile
// Two rings with a producer thread each (X tiles from HBM, digit tiles multicast from L2), so that neither stream waits
// for the other's slot.  Both are latency-bound on the bytes in flight, i.e. on shared memory: measured (XS, BS) ->
// forward / weight gradient ms: (8,3) 0.268/0.263, (10,2) 0.360/0.356, (6,4) 0.221/0.219, (5,5) 0.210/0.202,
// (4,6) 0.228/0.220, (3,7) 0.272/0.260; one ring of 5 whole stages was 0.226/0.224.
#ifndef LI_XS
#define LI_XS 5
#endif
#ifndef LI_BS
#define LI_BS 5
#endif
constexpr int XS = LI_XS, BS = LI_BS;
constexpr int ACC1_COL = 256;                            // TMEM column of the second MMA's accumulator
constexpr int D2_COL = LI_COMPACT ? D2_ROW : ACC1_COL;   // TMEM column of digit 2, feature 0 (compact: N0 == ACC1_COL, so
                                                         // accumulator columns and limb rows are numbered alike)
static_assert(!LI_COMPACT || N0 == ACC1_COL, "the compact layout needs adjacent accumulators");

struct Bars {
    uint64_t xfull[XS], xempty[XS], bfull[BS], bempty[BS], acc_full;
    uint32_t tmem_ptr;
};
constexpr size_t SMEM = 1024 + (size_t)XS * A_BYTES + (size_t)BS * B_BYTES + sizeof(Bars) + 64;
static_assert(SMEM <= 232448, "shared memory of linear_i8_kernel");

// ---- quantisation -----------------------------------------------------------------------------------------------------
// SHIFT = 8: base-256 digits in [-128, 127] (24-bit form); SHIFT = 7: base-128 digits in [-64, 64] (21-bit form)
// Largest |q| whose three balanced base-256 digits fit int8 is 127 * 65536 + 127 * 256 + 127 = 8 355 711.  w * (QMAX / amax)
// is rounded twice in float32 and can land one unit above QMAX for the feature's largest element, so QMAX keeps a margin
// below that bound and q is clamped to the bound: without either, the top digit of the maximum wrapped from 128 to -128.
constexpr int   QLIM24 = 8355711;
constexpr float QMAX24 = 8355000.0f;
constexpr float QMAX21 = 1048576.0f;
template <int SHIFT>
__device__ __forceinline__ void digits3(float w, float inv_scale, int8_t& d0, int8_t& d1, int8_t& d2) {
    constexpr int B = 1 << SHIFT, HALF = B >> 1;
    int q = __float2int_rn(w * inv_scale);               // |q| <= QMAX up to float rounding
    if (SHIFT == 8) q = max(-QLIM24, min(QLIM24, q));
    int t0 = ((q + HALF) & (B - 1)) - HALF;              // balanced digit
    int q1 = (q - t0) >> SHIFT;
    int t1 = ((q1 + HALF) & (B - 1)) - HALF;
    int q2 = (q1 - t1) >> SHIFT;
    d0 = (int8_t)t0; d1 = (int8_t)t1; d2 = (int8_t)q2;
}
// (1 / scale, scale) of a feature whose largest magnitude has the bit pattern `bits`.  A non-finite maximum (NaN or Inf
// somewhere in the feature) quantises everything to zero and makes the scale NaN: every output of the feature is NaN,
// as it is (NaN or Inf) after the reference's fp32 matmul.
__device__ __forceinline__ void feature_scale(unsigned bits, float qmax, float& inv, float& scale) {
    const float a = __uint_as_float(bits);
    if (bits >= 0x7F800000u) { inv = 0.f; scale = __uint_as_float(0x7FC00000u); return; }
    inv = a > 0.f ? qmax / a : 0.f;
    scale = a / qmax;
}

// amax_bits[r] = bits of max_c |in[r, c]|: grid (rows, chunks), one atomicMax per CTA.  The maximum is taken over the BIT
// PATTERNS of |x|: for non-negative floats the integer order is the float order, +Inf sorts above every finite value and
// every NaN above +Inf, so a NaN anywhere in the row survives the reduction (fmaxf would drop it).
__global__ void row_absmax_kernel(const float* __restrict__ in, int64_t ld, int64_t cols, unsigned* __restrict__ amax_bits) {
    const float* row = in + (int64_t)blockIdx.x * ld;
    const int64_t per = ((cols + gridDim.y - 1) / gridDim.y + 3) / 4 * 4;          // chunk starts stay 16-byte aligned
    const int64_t c0 = (int64_t)blockIdx.y * per < cols ? (int64_t)blockIdx.y * per : cols, c1 = c0 + per < cols ? c0 + per : cols;
    unsigned m = 0;
    if (((reinterpret_cast<uintptr_t>(row) | (uintptr_t)(c0 * 4)) & 15) == 0) {        // 128-bit loads over the aligned body
        const int64_t n4 = (c1 - c0) / 4;
        const uint4* v = reinterpret_cast<const uint4*>(row + c0);
        for (int64_t i = threadIdx.x; i < n4; i += blockDim.x) {
            const uint4 q = v[i];
            m = max(max(m, q.x & 0x7FFFFFFFu), max(max(q.y & 0x7FFFFFFFu, q.z & 0x7FFFFFFFu), q.w & 0x7FFFFFFFu));
        }
        for (int64_t c = c0 + 4 * n4 + threadIdx.x; c < c1; c += blockDim.x) m = max(m, __float_as_uint(row[c]) & 0x7FFFFFFFu);
    } else {
        for (int64_t c = c0 + threadIdx.x; c < c1; c += blockDim.x) m = max(m, __float_as_uint(row[c]) & 0x7FFFFFFFu);
    }
    __shared__ unsigned red[32];
    for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x < 32) {
        m = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0u;
        for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
        if (threadIdx.x == 0) atomicMax(&amax_bits[blockIdx.x], m);
    }
}

// amax_bits[c] = bits of max_r |in[r, c]|  (backward: columns of dH); amax must be zeroed
__global__ void col_absmax_kernel(const float* __restrict__ in, int64_t ld, int64_t rows, int cols, unsigned* __restrict__ amax_bits) {
    const int c = threadIdx.x;
    if (c >= cols) return;
    const int64_t r0 = (int64_t)blockIdx.x * 64, r1 = r0 + 64 < rows ? r0 + 64 : rows;
    unsigned m = 0;
    for (int64_t r = r0; r < r1; ++r) m = max(m, __float_as_uint(in[r * ld + c]) & 0x7FFFFFFFu);
    atomicMax(&amax_bits[c], m);
}

// limbs[(digit rows), c] of W[f, c].  One thread = 16 consecutive columns of one limb-matrix row triple (f, HP + f, D2_ROW + f):
// four float4 loads, three 16-byte stores.  Rows f >= h (padding of the MMA N dimension) and columns >= cols are zeroed here,
// so the workspace needs no separate clear.
template <int SHIFT>
__global__ void quantize_rows_kernel(const float* __restrict__ W, int64_t ldw, int h, int64_t cols, const unsigned* __restrict__ amax_bits,
                                     int8_t* __restrict__ limbs, int64_t ldl, float* __restrict__ scale) {
    const int64_t groups = ldl / 16;                     // ldl is a multiple of 16
    const int64_t total = (int64_t)F_ROWS * groups;      // f runs over the feature slots: digits 0/1 for f < HP, digit 2 for all
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int f = (int)(i / groups);
        const int64_t c0 = (i % groups) * 16;
        __align__(16) int8_t d[3][16];
        if (f < h) {
            float inv, sc;
            feature_scale(amax_bits[f], SHIFT == 8 ? QMAX24 : QMAX21, inv, sc);
            const float* src = W + (int64_t)f * ldw + c0;
            const bool vec = (c0 + 16 <= cols) && ((reinterpret_cast<uintptr_t>(src) & 15) == 0);
            float w[16];
            if (vec) {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const float4 v = reinterpret_cast<const float4*>(src)[q];
                    w[4 * q] = v.x; w[4 * q + 1] = v.y; w[4 * q + 2] = v.z; w[4 * q + 3] = v.w;
                }
            } else {
#pragma unroll
                for (int q = 0; q < 16; ++q) w[q] = (c0 + q < cols) ? src[q] : 0.f;
            }
#pragma unroll
            for (int q = 0; q < 16; ++q) digits3<SHIFT>(w[q], inv, d[0][q], d[1][q], d[2][q]);
            if (c0 == 0) scale[f] = sc;
        } else {
#pragma unroll
            for (int q = 0; q < 16; ++q) d[0][q] = d[1][q] = d[2][q] = 0;
        }
        if (f < HP) {
            *reinterpret_cast<uint4*>(limbs + (int64_t)f * ldl + c0) = *reinterpret_cast<const uint4*>(d[0]);
            *reinterpret_cast<uint4*>(limbs + (int64_t)(HP + f) * ldl + c0) = *reinterpret_cast<const uint4*>(d[1]);
        }
        *reinterpret_cast<uint4*>(limbs + (int64_t)(D2_ROW + f) * ldl + c0) = *reinterpret_cast<const uint4*>(d[2]);
    }
}

// The same result as row_absmax_kernel + quantize_rows_kernel in ONE launch, for rows that fit shared memory (4 * cols bytes;
// 200 KB at p = 50 000): a CTA = one feature slot, the row goes through shared memory once (largest magnitude on the way in,
// digits on the way out).  Between the optimiser step and the next projection these launches sit on the critical path of a
// training step; this takes two kernels and a memset (~22 us) to one (~8 us).  Bit-identical digits and scales.
template <int SHIFT>
__global__ void __launch_bounds__(1024, 1)
quantize_rows_fused_kernel(const float* __restrict__ W, int64_t ldw, int h, int64_t cols, int8_t* __restrict__ limbs, int64_t ldl,
                           float* __restrict__ scale) {
    extern __shared__ __align__(16) float q_row[];                 // [ldl] (ldl = cols rounded up to 16), zero beyond cols
    __shared__ unsigned red[32];
    const int f = blockIdx.x;
    const int64_t groups = ldl / 16;
    if (f >= h) {                                                  // padding slot of the MMA N dimension: zero digits
        const uint4 z = make_uint4(0u, 0u, 0u, 0u);
        for (int64_t g = threadIdx.x; g < groups; g += blockDim.x) {
            if (f < HP) {
                *reinterpret_cast<uint4*>(limbs + (int64_t)f * ldl + g * 16) = z;
                *reinterpret_cast<uint4*>(limbs + (int64_t)(HP + f) * ldl + g * 16) = z;
            }
            *reinterpret_cast<uint4*>(limbs + (int64_t)(D2_ROW + f) * ldl + g * 16) = z;
        }
        return;
    }
    const float* row = W + (int64_t)f * ldw;
    unsigned m = 0;
    if ((reinterpret_cast<uintptr_t>(row) & 15) == 0) {
        const int64_t n4 = cols / 4;
        for (int64_t i = threadIdx.x; i < n4; i += blockDim.x) {
            const float4 v = reinterpret_cast<const float4*>(row)[i];
            reinterpret_cast<float4*>(q_row)[i] = v;
            m = max(max(m, __float_as_uint(v.x) & 0x7FFFFFFFu), max(max(__float_as_uint(v.y) & 0x7FFFFFFFu, __float_as_uint(v.z) & 0x7FFFFFFFu),
                                                                    __float_as_uint(v.w) & 0x7FFFFFFFu));
        }
        for (int64_t c = 4 * n4 + threadIdx.x; c < cols; c += blockDim.x) { q_row[c] = row[c]; m = max(m, __float_as_uint(row[c]) & 0x7FFFFFFFu); }
    } else {
        for (int64_t c = threadIdx.x; c < cols; c += blockDim.x) { const float v = row[c]; q_row[c] = v; m = max(m, __float_as_uint(v) & 0x7FFFFFFFu); }
    }
    for (int64_t c = cols + threadIdx.x; c < ldl; c += blockDim.x) q_row[c] = 0.f;
    for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
    __syncthreads();
    m = red[0];
    for (int w = 1; w < (int)(blockDim.x >> 5); ++w) m = max(m, red[w]);
    float inv, sc;
    feature_scale(m, SHIFT == 8 ? QMAX24 : QMAX21, inv, sc);
    if (threadIdx.x == 0) scale[f] = sc;
    // a thread takes four consecutive columns at a time: conflict-free 128-bit reads of the row, and a warp's three 32-bit
    // stores cover 128 consecutive bytes of each digit row
    for (int64_t g = threadIdx.x; g < ldl / 4; g += blockDim.x) {
        const float4 v = reinterpret_cast<const float4*>(q_row)[g];
        __align__(4) int8_t d[3][4];
        digits3<SHIFT>(v.x, inv, d[0][0], d[1][0], d[2][0]);
        digits3<SHIFT>(v.y, inv, d[0][1], d[1][1], d[2][1]);
        digits3<SHIFT>(v.z, inv, d[0][2], d[1][2], d[2][2]);
        digits3<SHIFT>(v.w, inv, d[0][3], d[1][3], d[2][3]);
        if (f < HP) {
            *reinterpret_cast<uint32_t*>(limbs + (int64_t)f * ldl + g * 4) = *reinterpret_cast<const uint32_t*>(d[0]);
            *reinterpret_cast<uint32_t*>(limbs + (int64_t)(HP + f) * ldl + g * 4) = *reinterpret_cast<const uint32_t*>(d[1]);
        }
        *reinterpret_cast<uint32_t*>(limbs + (int64_t)(D2_ROW + f) * ldl + g * 4) = *reinterpret_cast<const uint32_t*>(d[2]);
    }
}

// limbs[(digit rows), r] of dH[r, f]: transposed through shared memory.  CTA = 128 rows x a quarter of the 128 limb-row
// triples; a thread packs 4 consecutive rows of one triple into 32-bit stores (a warp writes 128 consecutive bytes).
template <int SHIFT>
__global__ void quantize_cols_kernel(const float* __restrict__ dH, int64_t ld, int64_t rows, int h, const unsigned* __restrict__ amax_bits,
                                     int8_t* __restrict__ limbs, int64_t ldl, float* __restrict__ scale) {
    __shared__ float tile[128][33];                      // [row][feature of this quarter]
    const int64_t r0 = (int64_t)blockIdx.x * 128;
    const int f0 = blockIdx.y * 32;
    for (int i = threadIdx.x; i < 128 * 32; i += blockDim.x) {
        const int rr = i >> 5, ff = i & 31;
        tile[rr][ff] = (r0 + rr < rows && f0 + ff < h) ? dH[(r0 + rr) * ld + f0 + ff] : 0.f;
    }
    __syncthreads();
    // 256 threads: lane group of 32 covers 128 rows (4 each) of one feature; 8 features per pass
    const int quad = threadIdx.x & 31, fsub = threadIdx.x >> 5;
    for (int ff = fsub; ff < 32; ff += 8) {
        const int f = f0 + ff;
        if (f >= F_ROWS) continue;
        float inv = 0.f, sc = 0.f;
        if (f < h) feature_scale(amax_bits[f], SHIFT == 8 ? QMAX24 : QMAX21, inv, sc);
        __align__(4) int8_t d[3][4];
#pragma unroll
        for (int q = 0; q < 4; ++q) digits3<SHIFT>(tile[4 * quad + q][ff], inv, d[0][q], d[1][q], d[2][q]);
        const int64_t col = r0 + 4 * quad;               // ldl % 16 == 0 and r0 % 128 == 0: the 4-byte store stays inside the row
        if (col < ldl) {
            if (f < HP) {
                *reinterpret_cast<uint32_t*>(limbs + (int64_t)f * ldl + col) = *reinterpret_cast<const uint32_t*>(d[0]);
                *reinterpret_cast<uint32_t*>(limbs + (int64_t)(HP + f) * ldl + col) = *reinterpret_cast<const uint32_t*>(d[1]);
            }
            *reinterpret_cast<uint32_t*>(limbs + (int64_t)(D2_ROW + f) * ldl + col) = *reinterpret_cast<const uint32_t*>(d[2]);
        }
        if (blockIdx.x == 0 && quad == 0 && f < h) scale[f] = sc;
    }
}

// ---- the pair GEMM ------------------------------------------------------------------------------------------------------
template <bool BWD>
__global__ void __launch_bounds__(256, 1)
linear_i8_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_b0,
                 const __grid_constant__ CUtensorMap map_b1, int64_t k_total, const float* __restrict__ scale,
                 float* __restrict__ out, int64_t ld_out, int64_t m_total, int h,
                 int quad,          // 1: clusters of 4 = two pairs; each digit half-tile is fetched once and multicast to both pairs
                                    //    (map_b0 / map_b1 then have boxes of a quarter instead of a half of N0 / N1 rows)
                 int shift,         // log2 of the digit base: 8 (24-bit form) or 7 (21-bit form)
                 int ablate) {      // -DGNNB_ABLATION builds only (results invalid): 1 = no TMA loads, 2 = no MMAs
    extern __shared__ uint8_t smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t crank = cluster_ctarank();                          // 0..1, or 0..3 in quad mode
    const uint32_t rank = crank & 1, lead_rank = crank & ~1u;          // position inside the pair; cluster rank of the pair's leader
    const uint32_t twin = crank >> 1;                                  // which of the two pairs (quad mode)
    const bool leader = rank == 0;
    const uint16_t pair_mask = (uint16_t)(3u << lead_rank), all_mask = quad ? 15 : 3;
    uint8_t* xt = smem_raw + ((1024 - (smem_u32(smem_raw) & 1023)) & 1023);      // X ring
    uint8_t* bt = xt + (size_t)XS * A_BYTES;                                   // digit ring
    Bars* bars = reinterpret_cast<Bars*>(bt + (size_t)BS * B_BYTES);
    const int n_kb = (int)((k_total + BKB - 1) / BKB);
    const int64_t m0 = ((int64_t)(blockIdx.x >> 1) * 2 + rank) * BM;      // this CTA's rows of the M side (may lie past the end)

    if (threadIdx.x == 0) {
        for (int s = 0; s < XS; ++s) { mbar_init(&bars->xfull[s], 2); mbar_init(&bars->xempty[s], 1); }
        for (int s = 0; s < BS; ++s) { mbar_init(&bars->bfull[s], 2); mbar_init(&bars->bempty[s], quad ? 2 : 1); }
        mbar_init(&bars->acc_full, 1);
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc_pair<512>(&bars->tmem_ptr);
    tc_fence_before_sync();
    cluster_sync_all();
    tc_fence_after_sync();
    const uint32_t tmem_base = bars->tmem_ptr;

    if (warp == 0) {
        if (elect_one()) {                                             // TMA producer of the X ring (both CTAs)
            prefetch_tensormap(&map_x);
            for (int u = 0; u < n_kb; ++u) {
                const int s = u % XS;
                mbar_wait(&bars->xempty[s], ((u / XS) & 1) ^ 1);
                const uint32_t full_leader = mapa_u32(smem_u32(&bars->xfull[s]), lead_rank);
                if (LI_ABLATE(1)) { mbar_arrive_cluster(full_leader); continue; }
                if (leader) mbar_arrive_expect_tx(&bars->xfull[s], 2 * A_BYTES);
                else        mbar_arrive_cluster(full_leader);
                uint8_t* a = xt + (size_t)s * A_BYTES;
                if (BWD) tma_load_2d_pair(a, &map_x, full_leader, (int32_t)m0, u * BKB);            // [128 k-rows x 128 markers]
                else     tma_load_2d_pair(a, &map_x, full_leader, u * BKB, (int32_t)m0);            // [128 rows x 128 k]
            }
        }
    } else if (warp == 3) {
        if (elect_one()) {                                             // TMA producer of the digit ring (both CTAs)
            prefetch_tensormap(&map_b0);
            prefetch_tensormap(&map_b1);
            for (int u = 0; u < n_kb; ++u) {
                const int s = u % BS;
                mbar_wait(&bars->bempty[s], ((u / BS) & 1) ^ 1);
                const uint32_t full_leader = mapa_u32(smem_u32(&bars->bfull[s]), lead_rank);
                if (LI_ABLATE(1)) { mbar_arrive_cluster(full_leader); continue; }
                if (leader) mbar_arrive_expect_tx(&bars->bfull[s], 2 * B_BYTES);
                else        mbar_arrive_cluster(full_leader);
                uint8_t* b = bt + (size_t)s * B_BYTES;
                if (quad) {
                    // this CTA fetches half of the digit rows its pair position needs and delivers them to both pairs
                    const uint16_t mc = (uint16_t)((1u << rank) | (1u << (rank + 2)));
                    tma_load_2d_pair_multicast(b + twin * (B0_HALF / 2), &map_b0, &bars->bfull[s], u * BKB,
                                               (int32_t)(rank * (N0 / 2) + twin * (N0 / 4)), mc);
                    tma_load_2d_pair_multicast(b + B0_HALF + twin * (B1_HALF / 2), &map_b1, &bars->bfull[s], u * BKB,
                                               (int32_t)(N0 + rank * (N1 / 2) + twin * (N1 / 4)), mc);
                } else {
                    tma_load_2d_pair(b, &map_b0, full_leader, u * BKB, (int32_t)(rank * (N0 / 2)));
                    tma_load_2d_pair(b + B0_HALF, &map_b1, full_leader, u * BKB, (int32_t)(N0 + rank * (N1 / 2)));
                }
            }
        }
    } else if (warp == 1 && leader) {
        if (elect_one()) {                                             // MMA issuer (leader only)
            constexpr uint32_t idesc0 = make_idesc(/*S32*/ 2, /*int8*/ 1, BWD ? 1 : 0, 0, 2 * BM, N0);
            constexpr uint32_t idesc1 = make_idesc(/*S32*/ 2, /*int8*/ 1, BWD ? 1 : 0, 0, 2 * BM, N1);
            for (int u = 0; u < n_kb; ++u) {
                const int sx = u % XS, sb = u % BS;
                mbar_wait(&bars->xfull[sx], (u / XS) & 1);
                mbar_wait(&bars->bfull[sb], (u / BS) & 1);
                tc_fence_after_sync();
                const uint32_t a_addr = smem_u32(xt + (size_t)sx * A_BYTES);
                const uint32_t b0_addr = smem_u32(bt + (size_t)sb * B_BYTES), b1_addr = b0_addr + B0_HALF;
                if (LI_ABLATE(2)) {
                    for (uint32_t c = 0; c < 2u; ++c) mbar_arrive_cluster(mapa_u32(smem_u32(&bars->xempty[sx]), lead_rank + c));
                    for (uint32_t c = 0; c < (quad ? 4u : 2u); ++c) mbar_arrive_cluster(mapa_u32(smem_u32(&bars->bempty[sb]), c));
                    continue;
                }
#pragma unroll
                for (int k4 = 0; k4 < BKB / 32; ++k4) {
                    // K-major: 32 bytes further along the 128-byte swizzle row; MN-major (A of the backward pass): 32 k-rows =
                    // four 8-row atoms of 1024 bytes further
                    const uint64_t da = BWD ? make_smem_desc(a_addr + k4 * 4096, 16384, 1024) : make_smem_desc(a_addr + k4 * 32, 16, 1024);
                    const uint64_t db0 = make_smem_desc(b0_addr + k4 * 32, 16, 1024);
                    const uint64_t db1 = make_smem_desc(b1_addr + k4 * 32, 16, 1024);
                    const uint32_t acc = (u > 0 || k4 > 0) ? 1u : 0u;
                    mma_i8_ss_pair(tmem_base, da, db0, idesc0, acc);
                    mma_i8_ss_pair(tmem_base + ACC1_COL, da, db1, idesc1, acc);
                }
                tc_commit_pair(&bars->xempty[sx], pair_mask);          // X slots are private to the pair
                tc_commit_pair(&bars->bempty[sb], all_mask);           // digit slots are written by both pairs' producers
            }
            tc_commit_pair(&bars->acc_full, pair_mask);
        }
    } else if (warp >= 4) {
        // ===================================================== epilogue (both CTAs): three digit sums -> one rounding
        const int wq = warp & 3;
        const int64_t m = m0 + 32 * wq + lane;                         // row of the M side == TMEM lane
        mbar_wait(&bars->acc_full, 0);
        tc_fence_after_sync();
        const uint32_t trow = tmem_base + ((uint32_t)(32 * wq) << 16);
#pragma unroll 1
        for (int c0 = 0; c0 < h; c0 += 32) {
            uint32_t a0[32], a1[32], a2[32];
            tmem_ld_32x32b_x32(trow + c0, a0);
            tmem_ld_32x32b_x32(trow + HP + c0, a1);
            tmem_ld_32x32b_x32(trow + D2_COL + c0, a2);
            tmem_ld_wait();
            if (m < m_total) {
#pragma unroll
                for (int c = 0; c < 32; ++c) {
                    const int f = c0 + c;
                    if (f < h) {
                        const long long sum = (long long)(int)a0[c] + ((long long)(int)a1[c] << shift) + ((long long)(int)a2[c] << (2 * shift));
                        const float r = (float)((double)sum * (double)scale[f]);
                        if (BWD) out[(int64_t)f * ld_out + m] = r;     // dW[f, marker]: a warp writes 32 consecutive markers
                        else     out[m * ld_out + f] = r;              // H[row, f]
                    }
                }
            }
        }
    }
    tc_fence_before_sync();
    cluster_sync_all();
    if (warp == 2) tmem_dealloc_pair<512>(tmem_base);
}

// ---- the same GEMM with X as 2-BIT PACKED codes in HBM -----------------------------------------------------------------
// Genotypes are 0/1/2: four codes per byte (gnnb_pack_2bit).  The X stream -- the only large one -- shrinks fourfold in
// HBM, in L2 -> SM traffic and in the shared memory its ring needs.  TMA brings the packed tile [128 rows x 32 B]; four
// converter warps per k-block (thread = row) expand it to the int8 [128 x 128 B] tile in the 128-byte-swizzled UMMA layout
// (~1 ALU operation per code), fence.proxy.async, and arrive on the LEADER's `afull` barrier (8 arrivals: 4 warps x 2 CTAs);
// the MMAs then read it exactly as the TMA-written tile of linear_i8_kernel.  The tile is the same for both directions
// ([128 individuals x 128 markers]; K-major for the forward pass, MN-major for the weight gradient).
#ifndef P2_PXS
#define P2_PXS 8
#endif
#ifndef P2_AS
#define P2_AS 4
#endif
#ifndef P2_BS
#define P2_BS 5
#endif
constexpr int PXS = P2_PXS, AS = P2_AS, BS2 = P2_BS;     // PXS and AS even (two converter groups)
constexpr int PX_BYTES = BM * (BKB / 4);                 // 4 KB packed tile
constexpr int P2_THREADS = 128 + 256;

struct Bars2 {
    uint64_t pxfull[PXS], pxempty[PXS], afull[AS], aempty[AS], bfull[BS2], bempty[BS2], acc_full;
    uint32_t tmem_ptr;
};
constexpr size_t SMEM2 = 1024 + (size_t)PXS * PX_BYTES + (size_t)AS * A_BYTES + (size_t)BS2 * B_BYTES + sizeof(Bars2) + 64;
static_assert(SMEM2 <= 232448, "shared memory of linear_p2_kernel");

// one packed word (16 codes) -> four words of four int8 codes: the four 2-bit planes, then a 4 x 4 byte transpose (prmt)
__device__ __forceinline__ void expand16(uint32_t w, uint32_t (&o)[4]) {
    const uint32_t A = w & 0x03030303u, B = (w >> 2) & 0x03030303u, C = (w >> 4) & 0x03030303u, D = (w >> 6) & 0x03030303u;
    const uint32_t ab_lo = __byte_perm(A, B, 0x5140), cd_lo = __byte_perm(C, D, 0x5140);      // {A0,B0,A1,B1}, {C0,D0,C1,D1}
    const uint32_t ab_hi = __byte_perm(A, B, 0x7362), cd_hi = __byte_perm(C, D, 0x7362);      // {A2,B2,A3,B3}, {C2,D2,C3,D3}
    o[0] = __byte_perm(ab_lo, cd_lo, 0x5410);
    o[1] = __byte_perm(ab_lo, cd_lo, 0x7632);
    o[2] = __byte_perm(ab_hi, cd_hi, 0x5410);
    o[3] = __byte_perm(ab_hi, cd_hi, 0x7632);
}

template <bool BWD>
__global__ void __launch_bounds__(P2_THREADS, 1)
linear_p2_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_b0,
                 const __grid_constant__ CUtensorMap map_b1, int64_t k_total, const float* __restrict__ scale,
                 float* __restrict__ out, int64_t ld_out, int64_t m_total, int h, int quad, int shift) {
    extern __shared__ uint8_t smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t crank = cluster_ctarank();
    const uint32_t rank = crank & 1, lead_rank = crank & ~1u, twin = crank >> 1;
    const bool leader = rank == 0;
    const uint16_t pair_mask = (uint16_t)(3u << lead_rank), all_mask = quad ? 15 : 3;
    uint8_t* at = smem_raw + ((1024 - (smem_u32(smem_raw) & 1023)) & 1023);      // expanded A ring (swizzled: 1024-aligned)
    uint8_t* bt = at + (size_t)AS * A_BYTES;                                   // digit ring
    uint8_t* px = bt + (size_t)BS2 * B_BYTES;                                  // packed X ring
    Bars2* bars = reinterpret_cast<Bars2*>(px + (size_t)PXS * PX_BYTES);
    const int n_kb = (int)((k_total + BKB - 1) / BKB);
    const int64_t m0 = ((int64_t)(blockIdx.x >> 1) * 2 + rank) * BM;

    if (threadIdx.x == 0) {
        for (int s = 0; s < PXS; ++s) { mbar_init(&bars->pxfull[s], 1); mbar_init(&bars->pxempty[s], 4); }
        for (int s = 0; s < AS; ++s) { mbar_init(&bars->afull[s], 8); mbar_init(&bars->aempty[s], 1); }
        for (int s = 0; s < BS2; ++s) { mbar_init(&bars->bfull[s], 2); mbar_init(&bars->bempty[s], quad ? 2 : 1); }
        mbar_init(&bars->acc_full, 1);
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc_pair<512>(&bars->tmem_ptr);
    tc_fence_before_sync();
    cluster_sync_all();
    tc_fence_after_sync();
    const uint32_t tmem_base = bars->tmem_ptr;

    if (warp == 0) {
        if (elect_one()) {                                             // TMA producer of the packed X ring (CTA-local)
            prefetch_tensormap(&map_x);
            for (int u = 0; u < n_kb; ++u) {
                const int s = u % PXS;
                mbar_wait(&bars->pxempty[s], ((u / PXS) & 1) ^ 1);
                mbar_arrive_expect_tx(&bars->pxfull[s], PX_BYTES);
                if (BWD) tma_load_2d(px + (size_t)s * PX_BYTES, &map_x, &bars->pxfull[s], (int32_t)(m0 / 4), u * BKB);   // 128 individuals x 32 B
                else     tma_load_2d(px + (size_t)s * PX_BYTES, &map_x, &bars->pxfull[s], u * (BKB / 4), (int32_t)m0);
            }
        }
    } else if (warp == 3) {
        if (elect_one()) {                                             // TMA producer of the digit ring (as in linear_i8_kernel)
            prefetch_tensormap(&map_b0);
            prefetch_tensormap(&map_b1);
            for (int u = 0; u < n_kb; ++u) {
                const int s = u % BS2;
                mbar_wait(&bars->bempty[s], ((u / BS2) & 1) ^ 1);
                const uint32_t full_leader = mapa_u32(smem_u32(&bars->bfull[s]), lead_rank);
                if (leader) mbar_arrive_expect_tx(&bars->bfull[s], 2 * B_BYTES);
                else        mbar_arrive_cluster(full_leader);
                uint8_t* b = bt + (size_t)s * B_BYTES;
                if (quad) {
                    const uint16_t mc = (uint16_t)((1u << rank) | (1u << (rank + 2)));
                    tma_load_2d_pair_multicast(b + twin * (B0_HALF / 2), &map_b0, &bars->bfull[s], u * BKB,
                                               (int32_t)(rank * (N0 / 2) + twin * (N0 / 4)), mc);
                    tma_load_2d_pair_multicast(b + B0_HALF + twin * (B1_HALF / 2), &map_b1, &bars->bfull[s], u * BKB,
                                               (int32_t)(N0 + rank * (N1 / 2) + twin * (N1 / 4)), mc);
                } else {
                    tma_load_2d_pair(b, &map_b0, full_leader, u * BKB, (int32_t)(rank * (N0 / 2)));
                    tma_load_2d_pair(b + B0_HALF, &map_b1, full_leader, u * BKB, (int32_t)(N0 + rank * (N1 / 2)));
                }
            }
        }
    } else if (warp == 1 && leader) {
        if (elect_one()) {                                             // MMA issuer (leader only)
            constexpr uint32_t idesc0 = make_idesc(/*S32*/ 2, /*int8*/ 1, BWD ? 1 : 0, 0, 2 * BM, N0);
            constexpr uint32_t idesc1 = make_idesc(/*S32*/ 2, /*int8*/ 1, BWD ? 1 : 0, 0, 2 * BM, N1);
            for (int u = 0; u < n_kb; ++u) {
                const int sa = u % AS, sb = u % BS2;
                mbar_wait(&bars->afull[sa], (u / AS) & 1);             // both CTAs' converters have written their tiles
                mbar_wait(&bars->bfull[sb], (u / BS2) & 1);
                tc_fence_after_sync();
                const uint32_t a_addr = smem_u32(at + (size_t)sa * A_BYTES);
                const uint32_t b0_addr = smem_u32(bt + (size_t)sb * B_BYTES), b1_addr = b0_addr + B0_HALF;
#pragma unroll
                for (int k4 = 0; k4 < BKB / 32; ++k4) {
                    const uint64_t da = BWD ? make_smem_desc(a_addr + k4 * 4096, 16384, 1024) : make_smem_desc(a_addr + k4 * 32, 16, 1024);
                    const uint64_t db0 = make_smem_desc(b0_addr + k4 * 32, 16, 1024);
                    const uint64_t db1 = make_smem_desc(b1_addr + k4 * 32, 16, 1024);
                    const uint32_t acc = (u > 0 || k4 > 0) ? 1u : 0u;
                    mma_i8_ss_pair(tmem_base, da, db0, idesc0, acc);
                    mma_i8_ss_pair(tmem_base + ACC1_COL, da, db1, idesc1, acc);
                }
                tc_commit_pair(&bars->aempty[sa], pair_mask);
                tc_commit_pair(&bars->bempty[sb], all_mask);
            }
            tc_commit_pair(&bars->acc_full, pair_mask);
        }
    } else if (warp >= 4) {
        // ===================================================== converters: packed [128 x 32 B] -> int8 [128 x 128 B], swizzled
        // two groups of four warps take the k-blocks alternately (a warp's chain of two barrier waits, the expansion, the proxy
        // fence and the arrivals is longer than the tensor work of one k-block); thread = row of the tile
        const int group = (warp - 4) >> 2, row = (threadIdx.x - 128) & 127;
        const uint32_t afull_leader_base = mapa_u32(smem_u32(&bars->afull[0]), lead_rank);
        const uint32_t row_off = (uint32_t)((row >> 3) * 1024 + (row & 7) * 128), sw = (uint32_t)(row & 7);
        for (int u = group; u < n_kb; u += 2) {
            const int sx = u % PXS, sa = u % AS;                       // PXS and AS are even: a slot always meets the same group
            mbar_wait(&bars->pxfull[sx], (u / PXS) & 1);
            mbar_wait(&bars->aempty[sa], ((u / AS) & 1) ^ 1);
            uint32_t w[8];
            const uint32_t src = smem_u32(px + (size_t)sx * PX_BYTES) + (uint32_t)(row * 32);
            asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "r"(src));
            asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7]) : "r"(src + 16));
            const uint32_t dst = smem_u32(at + (size_t)sa * A_BYTES) + row_off;
#pragma unroll
            for (int c = 0; c < 8; ++c) {                              // 16-byte chunk c of the row: 16 codes = one packed word
                uint32_t o[4];
                expand16(w[c], o);
                asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};"
                             :: "r"(dst + (((uint32_t)c ^ sw) << 4)), "r"(o[0]), "r"(o[1]), "r"(o[2]), "r"(o[3]) : "memory");
            }
            fence_proxy_async();                                       // generic-proxy stores -> visible to the tensor-core proxy
            __syncwarp();
            if (lane == 0) {
                mbar_arrive_cluster(afull_leader_base + (uint32_t)(sa * sizeof(uint64_t)));
                mbar_arrive(&bars->pxempty[sx]);
            }
        }
        if (warp < 8) {
            // ================================================= epilogue (both CTAs): three digit sums -> one rounding
            const int wq = warp & 3;
            const int64_t m = m0 + 32 * wq + lane;
            mbar_wait(&bars->acc_full, 0);
            tc_fence_after_sync();
            const uint32_t trow = tmem_base + ((uint32_t)(32 * wq) << 16);
#pragma unroll 1
            for (int c0 = 0; c0 < h; c0 += 32) {
                uint32_t a0[32], a1[32], a2[32];
                tmem_ld_32x32b_x32(trow + c0, a0);
                tmem_ld_32x32b_x32(trow + HP + c0, a1);
                tmem_ld_32x32b_x32(trow + D2_COL + c0, a2);
                tmem_ld_wait();
                if (m < m_total) {
#pragma unroll
                    for (int c = 0; c < 32; ++c) {
                        const int f = c0 + c;
                        if (f < h) {
                            const long long sum = (long long)(int)a0[c] + ((long long)(int)a1[c] << shift) + ((long long)(int)a2[c] << (2 * shift));
                            const float r = (float)((double)sum * (double)scale[f]);
                            if (BWD) out[(int64_t)f * ld_out + m] = r;
                            else     out[m * ld_out + f] = r;
                        }
                    }
                }
            }
        }
    }
    tc_fence_before_sync();
    cluster_sync_all();
    if (warp == 2) tmem_dealloc_pair<512>(tmem_base);
}

template <bool BWD, bool PACKED>
int launch_i8(const CUtensorMap& mx, const CUtensorMap& mb0, const CUtensorMap& mb1, int64_t k_total, const float* scale, float* out,
              int64_t ld_out, int64_t m_total, int h, int shift, cudaStream_t stream) {
    const int64_t tiles = (m_total + BM - 1) / BM, pairs = (tiles + 1) / 2;
    static int quad = -1, ablate = 0;
    if (quad < 0) {
        const char* e = getenv("GNNB_LINEAR_QUAD");
        quad = e ? atoi(e) : 1;
#ifdef GNNB_ABLATION
        const char* a = getenv("GNNB_LINEAR_ABLATE");          // measurement builds only
        ablate = a ? atoi(a) : 0;
#endif
    }
    const int64_t groups = quad ? (pairs + 1) / 2 : pairs;
    const size_t smem = PACKED ? SMEM2 : SMEM;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)((quad ? 4 : 2) * groups));
    cfg.blockDim = dim3(PACKED ? P2_THREADS : 256);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = quad ? 4 : 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    if (PACKED) {
        GNNB_SET_DYN_SMEM(linear_p2_kernel<BWD>, smem);
        GNNB_CUDA(cudaLaunchKernelEx(&cfg, linear_p2_kernel<BWD>, mx, mb0, mb1, k_total, scale, out, ld_out, m_total, h, quad, shift));
    } else {
        GNNB_SET_DYN_SMEM(linear_i8_kernel<BWD>, smem);
        GNNB_CUDA(cudaLaunchKernelEx(&cfg, linear_i8_kernel<BWD>, mx, mb0, mb1, k_total, scale, out, ld_out, m_total, h, quad, shift, ablate));
    }
    return GNNB_OK;
}
}  // namespace

static uint32_t box_div() {          // quad mode fetches the digit half-tiles in two halves
    static int quad = -1;
    if (quad < 0) { const char* e = getenv("GNNB_LINEAR_QUAD"); quad = e ? atoi(e) : 1; }
    return quad ? 2 : 1;
}
static size_t limb_bytes(int64_t k_len) { return align_up((size_t)LIMB_ROWS * align_up((size_t)k_len, 16), 1024); }

// log2 of the digit base for a contraction of length k_len, or 0 if the int32 digit sums could overflow in either form:
// |sum| <= k_len * cmax * dmax with cmax = 3 (2-bit packed codes) or 128 (any int8) and dmax = 128 / 64 (24- / 21-bit form).
static int digit_shift(int x_format, int64_t k_len) {
    const int64_t cmax = x_format == 2 ? 3 : 128;
    if (k_len * cmax * 128 < (1LL << 31)) return 8;
    if (k_len * cmax * 64 < (1LL << 31)) return 7;
    return 0;
}

// x_format: 1 = int8 codes (ldx in elements = bytes), 2 = 2-bit packed codes (ldx = bytes per packed row).
// The contraction runs over p in the forward pass and over n in the weight gradient: both must be overflow-safe.
bool linear_i8_supported(const void* X, int x_format, int64_t ldx, int64_t n, int64_t p, int64_t h) {
    return (x_format == 1 || x_format == 2) && (uintptr_t)X % 16 == 0 && ldx % 16 == 0 && h >= 1 && h <= HP && n >= 128 && p >= 128 &&
           digit_shift(x_format, p) != 0 && digit_shift(x_format, n) != 0;
}

// Markers one full wave of the weight-gradient kernel covers on the current device: co-resident clusters x the markers a
// cluster owns (4 or 2 M tiles of 128).  Cutting the markers at multiples of this costs no tensor-core time.
int64_t linear_i8_bwd_w_wave_markers(int x_format) {
    const int quad = (int)box_div() == 2 ? 1 : 0;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(quad ? 4 * 64 : 2 * 128);
    cfg.blockDim = dim3(x_format == 2 ? P2_THREADS : 256);
    cfg.dynamicSmemBytes = x_format == 2 ? SMEM2 : SMEM;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = quad ? 4 : 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int clusters = 0;
    cudaError_t e;
    if (x_format == 2) {
        if (cudaFuncSetAttribute(linear_p2_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM2) != cudaSuccess) return 0;
        e = cudaOccupancyMaxActiveClusters(&clusters, linear_p2_kernel<true>, &cfg);
    } else {
        if (cudaFuncSetAttribute(linear_i8_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM) != cudaSuccess) return 0;
        e = cudaOccupancyMaxActiveClusters(&clusters, linear_i8_kernel<true>, &cfg);
    }
    if (e != cudaSuccess || clusters <= 0) { cudaGetLastError(); return 0; }
    return (int64_t)clusters * (quad ? 4 : 2) * BM;
}

size_t linear_i8_workspace(int64_t n, int64_t p, int64_t h) {
    (void)h;
    const size_t k_len = (size_t)(n > p ? n : p);
    return limb_bytes((int64_t)k_len) + 4096 + 2048;      // limbs + (amax, scale) + alignment slack
}

int linear_i8_fwd(const void* X, int x_format, int64_t ldx, int64_t n, int64_t p, const float* W, int64_t ldw, int64_t h,
                  float* H, int64_t ldh, void* ws, size_t ws_bytes, cudaStream_t stream) {
    const size_t need = linear_i8_workspace(n, p, h);
    GNNB_REQUIRE(ws && ws_bytes >= need, GNNB_E_WORKSPACE, "linear_fwd: workspace %zu < %zu bytes", ws_bytes, need);
    const int shift = digit_shift(x_format, p);
    GNNB_REQUIRE(shift != 0, GNNB_E_UNSUPPORTED, "linear_fwd: %lld markers overflow the int32 digit sums", (long long)p);
    uint8_t* wsb = reinterpret_cast<uint8_t*>(align_up((size_t)(uintptr_t)ws, 1024));
    int8_t* limbs = reinterpret_cast<int8_t*>(wsb);
    const int64_t ldl = (int64_t)align_up((size_t)p, 16);
    unsigned* amax_bits = reinterpret_cast<unsigned*>(wsb + limb_bytes(p));
    float* scale = reinterpret_cast<float*>(amax_bits + 512);
    const size_t row_smem = (size_t)ldl * sizeof(float);
    const char* fused_env = getenv("GNNB_QUANT_FUSED");                                     // "0": the two-kernel path (A/B, parity test)
    const bool fused_off = fused_env && fused_env[0] == '0';
    if (row_smem <= 220 * 1024 && !fused_off) {
        // a weight row fits shared memory: maximum, scale and digits in one launch (see quantize_rows_fused_kernel)
        if (shift == 8) {
            GNNB_SET_DYN_SMEM(quantize_rows_fused_kernel<8>, row_smem);
            quantize_rows_fused_kernel<8><<<F_ROWS, 1024, row_smem, stream>>>(W, ldw, (int)h, p, limbs, ldl, scale);
        } else {
            GNNB_SET_DYN_SMEM(quantize_rows_fused_kernel<7>, row_smem);
            quantize_rows_fused_kernel<7><<<F_ROWS, 1024, row_smem, stream>>>(W, ldw, (int)h, p, limbs, ldl, scale);
        }
        GNNB_LAUNCH_CHECK();
    } else {
    GNNB_CUDA(cudaMemsetAsync(amax_bits, 0, 512 * sizeof(unsigned), stream));
    row_absmax_kernel<<<dim3((unsigned)h, 16), 256, 0, stream>>>(W, ldw, p, amax_bits);
    GNNB_LAUNCH_CHECK();
    {
        int64_t total = (int64_t)F_ROWS * (ldl / 16), blocks = (total + 255) / 256, cap = (int64_t)sm_count() * 16;
        const unsigned grid = (unsigned)(blocks < cap ? blocks : cap);
        if (shift == 8) quantize_rows_kernel<8><<<grid, 256, 0, stream>>>(W, ldw, (int)h, p, amax_bits, limbs, ldl, scale);
        else            quantize_rows_kernel<7><<<grid, 256, 0, stream>>>(W, ldw, (int)h, p, amax_bits, limbs, ldl, scale);
        GNNB_LAUNCH_CHECK();
    }
    }
    CUtensorMap mx, mb0, mb1;
    int rc = x_format == 2 ? make_tensor_map_2d(&mx, X, 1, (uint64_t)((p + 3) / 4), (uint64_t)n, (uint64_t)ldx, BKB / 4, BM, /*swizzle*/ false)
                           : make_tensor_map_2d(&mx, X, 1, (uint64_t)p, (uint64_t)n, (uint64_t)ldx, BKB, BM);
    if (rc != GNNB_OK) return rc;
    rc = make_tensor_map_2d(&mb0, limbs, 1, (uint64_t)p, (uint64_t)LIMB_ROWS, (uint64_t)ldl, BKB, N0 / 2 / box_div());
    if (rc != GNNB_OK) return rc;
    rc = make_tensor_map_2d(&mb1, limbs, 1, (uint64_t)p, (uint64_t)LIMB_ROWS, (uint64_t)ldl, BKB, N1 / 2 / box_div());
    if (rc != GNNB_OK) return rc;
    if (x_format == 2) return launch_i8<false, true>(mx, mb0, mb1, p, scale, H, ldh, n, (int)h, shift, stream);
    return launch_i8<false, false>(mx, mb0, mb1, p, scale, H, ldh, n, (int)h, shift, stream);
}

// The weight gradient in two halves, so that a caller can split the markers into ranges (one GEMM launch per range, e.g.
// to overlap the all-reduce of one range with the GEMM of the next) without quantising dH once per range:
// linear_i8_prepare_dh fills the workspace with the digits and scales of dH^T, linear_i8_bwd_w_prepared runs the GEMM for
// the markers X points at (any 16-byte aligned column range of the genotype matrix) against that workspace.
int linear_i8_prepare_dh(int x_format, int64_t n, const float* dH, int64_t lddh, int64_t h, void* ws, size_t ws_bytes,
                         cudaStream_t stream) {
    const size_t need = linear_i8_workspace(n, 128, h);
    GNNB_REQUIRE(ws && ws_bytes >= need, GNNB_E_WORKSPACE, "linear_bwd_w: workspace %zu < %zu bytes", ws_bytes, need);
    const int shift = digit_shift(x_format, n);
    GNNB_REQUIRE(shift != 0, GNNB_E_UNSUPPORTED, "linear_bwd_w: %lld individuals overflow the int32 digit sums", (long long)n);
    uint8_t* wsb = reinterpret_cast<uint8_t*>(align_up((size_t)(uintptr_t)ws, 1024));
    int8_t* limbs = reinterpret_cast<int8_t*>(wsb);
    const int64_t ldl = (int64_t)align_up((size_t)n, 16);
    unsigned* amax_bits = reinterpret_cast<unsigned*>(wsb + limb_bytes(n));
    float* scale = reinterpret_cast<float*>(amax_bits + 512);
    GNNB_CUDA(cudaMemsetAsync(amax_bits, 0, 512 * sizeof(unsigned), stream));
    col_absmax_kernel<<<(unsigned)((n + 63) / 64), 128, 0, stream>>>(dH, lddh, n, (int)h, amax_bits);
    GNNB_LAUNCH_CHECK();
    const dim3 grid((unsigned)((ldl + 127) / 128), 4);
    if (shift == 8) quantize_cols_kernel<8><<<grid, 256, 0, stream>>>(dH, lddh, n, (int)h, amax_bits, limbs, ldl, scale);
    else            quantize_cols_kernel<7><<<grid, 256, 0, stream>>>(dH, lddh, n, (int)h, amax_bits, limbs, ldl, scale);
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

int linear_i8_bwd_w_prepared(const void* X, int x_format, int64_t ldx, int64_t n, int64_t p, int64_t h, float* dW, int64_t lddw,
                             void* ws, size_t ws_bytes, cudaStream_t stream) {
    const size_t need = linear_i8_workspace(n, 128, h);
    GNNB_REQUIRE(ws && ws_bytes >= need, GNNB_E_WORKSPACE, "linear_bwd_w: workspace %zu < %zu bytes", ws_bytes, need);
    const int shift = digit_shift(x_format, n);
    GNNB_REQUIRE(shift != 0, GNNB_E_UNSUPPORTED, "linear_bwd_w: %lld individuals overflow the int32 digit sums", (long long)n);
    uint8_t* wsb = reinterpret_cast<uint8_t*>(align_up((size_t)(uintptr_t)ws, 1024));
    int8_t* limbs = reinterpret_cast<int8_t*>(wsb);
    const int64_t ldl = (int64_t)align_up((size_t)n, 16);
    float* scale = reinterpret_cast<float*>(reinterpret_cast<unsigned*>(wsb + limb_bytes(n)) + 512);
    CUtensorMap mx, mb0, mb1;
    // box: 128 markers (128 B, or 32 packed bytes) x 128 individuals
    int rc = x_format == 2 ? make_tensor_map_2d(&mx, X, 1, (uint64_t)((p + 3) / 4), (uint64_t)n, (uint64_t)ldx, 32, BKB, /*swizzle*/ false)
                           : make_tensor_map_2d(&mx, X, 1, (uint64_t)p, (uint64_t)n, (uint64_t)ldx, 128, BKB);
    if (rc != GNNB_OK) return rc;
    rc = make_tensor_map_2d(&mb0, limbs, 1, (uint64_t)n, (uint64_t)LIMB_ROWS, (uint64_t)ldl, BKB, N0 / 2 / box_div());
    if (rc != GNNB_OK) return rc;
    rc = make_tensor_map_2d(&mb1, limbs, 1, (uint64_t)n, (uint64_t)LIMB_ROWS, (uint64_t)ldl, BKB, N1 / 2 / box_div());
    if (rc != GNNB_OK) return rc;
    if (x_format == 2) return launch_i8<true, true>(mx, mb0, mb1, n, scale, dW, lddw, p, (int)h, shift, stream);
    return launch_i8<true, false>(mx, mb0, mb1, n, scale, dW, lddw, p, (int)h, shift, stream);
}

int linear_i8_bwd_w(const void* X, int x_format, int64_t ldx, int64_t n, int64_t p, const float* dH, int64_t lddh, int64_t h,
                    float* dW, int64_t lddw, void* ws, size_t ws_bytes, cudaStream_t stream) {
    const size_t need = linear_i8_workspace(n, p, h);
    GNNB_REQUIRE(ws && ws_bytes >= need, GNNB_E_WORKSPACE, "linear_bwd_w: workspace %zu < %zu bytes", ws_bytes, need);
    int rc = linear_i8_prepare_dh(x_format, n, dH, lddh, h, ws, ws_bytes, stream);
    if (rc != GNNB_OK) return rc;
    return linear_i8_bwd_w_prepared(X, x_format, ldx, n, p, h, dW, lddw, ws, ws_bytes, stream);
}

}  // namespace gnnb
