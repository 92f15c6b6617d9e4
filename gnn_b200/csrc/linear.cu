// K5: marker projection H = X . W^T and its gradients -- CUDA-core fp32 tiles (precision 0) and the
// dispatcher to the tcgen05 kernels (precision 1..3, linear_tc.cu).
//
// Replaces GCNConv.lin (PyG Linear(bias=False) -> torch.matmul) and its autograd; call sites
// gnn/networks/networks.py:271-272 and the other GCN factories.
//
// The fp32 path is the exact-arithmetic reference point for the tensor-core path and serves the small
// inner layers (h x h).  One generic 64x64x16 tile kernel covers the three products through strides:
//   C(m, n) = sum_k A(m, k) * B(k, n),  A(m, k) = A[m * sam + k * sak],  B(k, n) = B[k * sbk + n * sbn]
// with the unit-stride dimension of each operand chosen at compile time so that global loads coalesce.
#include "common.cuh"

namespace gnnb {

constexpr int BM = 64, BN = 64, BK = 16;

template <typename AT> __device__ __forceinline__ float to_f32(AT v) { return (float)v; }

// A_KC: A is contiguous along k (else along m).  B_KC: B is contiguous along k (else along n).
template <typename AT, bool A_KC, bool B_KC>
__global__ void __launch_bounds__(256)
gemm_simt_kernel(const AT* __restrict__ A, int64_t sam, int64_t sak, const float* __restrict__ B, int64_t sbk, int64_t sbn,
                 float* __restrict__ C, int64_t scm, int64_t scn, int64_t M, int64_t N, int64_t K,
                 int64_t k_per_split, int64_t split_stride) {
    __shared__ float As[BK][BM + 4];
    __shared__ float Bs[BK][BN + 4];
    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;
    const int64_t m0 = (int64_t)blockIdx.y * BM, n0 = (int64_t)blockIdx.x * BN;
    const int64_t kb = (int64_t)blockIdx.z * k_per_split;
    const int64_t ke = (kb + k_per_split < K) ? kb + k_per_split : K;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    for (int64_t k0 = kb; k0 < ke; k0 += BK) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            int m, k;
            if (A_KC) { k = tid & 15; m = (tid >> 4) + 16 * i; }
            else      { m = tid & 63; k = (tid >> 6) + 4 * i; }
            int64_t gm = m0 + m, gk = k0 + k;
            As[k][m] = (gm < M && gk < ke) ? to_f32<AT>(A[gm * sam + gk * sak]) : 0.f;
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            int nn, k;
            if (B_KC) { k = tid & 15; nn = (tid >> 4) + 16 * i; }
            else      { nn = tid & 63; k = (tid >> 6) + 4 * i; }
            int64_t gn = n0 + nn, gk = k0 + k;
            Bs[k][nn] = (gn < N && gk < ke) ? B[gk * sbk + gn * sbn] : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < BK; ++k) {
            float a[4], b[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) a[i] = As[k][ty + 16 * i];
#pragma unroll
            for (int j = 0; j < 4; ++j) b[j] = Bs[k][tx + 16 * j];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        __syncthreads();
    }
    float* Cs = C + (int64_t)blockIdx.z * split_stride;
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            int64_t gm = m0 + ty + 16 * i, gn = n0 + tx + 16 * j;
            if (gm < M && gn < N) Cs[gm * scm + gn * scn] = acc[i][j];
        }
}

// C(m, n) = sum_s part[s][m][n] in a fixed order; part is dense row-major [M, N] per split
__global__ void splitk_reduce_kernel(const float* __restrict__ part, int n_splits, int64_t M, int64_t N,
                                     float* __restrict__ C, int64_t scm, int64_t scn) {
    int64_t total = M * N;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        float s = 0.f;
        for (int sp = 0; sp < n_splits; ++sp) s += part[(int64_t)sp * total + i];
        C[(i / N) * scm + (i % N) * scn] = s;
    }
}

template <typename AT, bool A_KC, bool B_KC>
static int launch_gemm(const AT* A, int64_t sam, int64_t sak, const float* B, int64_t sbk, int64_t sbn,
                       float* C, int64_t scm, int64_t scn, int64_t M, int64_t N, int64_t K,
                       int n_splits, float* partial, cudaStream_t stream) {
    dim3 grid((unsigned)((N + BN - 1) / BN), (unsigned)((M + BM - 1) / BM), (unsigned)(n_splits > 1 ? n_splits : 1));
    if (n_splits <= 1) {
        gemm_simt_kernel<AT, A_KC, B_KC><<<grid, 256, 0, stream>>>(A, sam, sak, B, sbk, sbn, C, scm, scn, M, N, K, K, 0);
        GNNB_LAUNCH_CHECK();
        return GNNB_OK;
    }
    // split-K: dense row-major [M, N] partials per split, reduced afterwards in a fixed order
    int64_t kps = ((K + n_splits - 1) / n_splits + BK - 1) / BK * BK;
    gemm_simt_kernel<AT, A_KC, B_KC><<<grid, 256, 0, stream>>>(A, sam, sak, B, sbk, sbn, partial, N, 1, M, N, K, kps, M * N);
    GNNB_LAUNCH_CHECK();
    int64_t blocks = (M * N + 255) / 256;
    int64_t cap = (int64_t)sm_count() * 16;
    splitk_reduce_kernel<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, stream>>>(partial, n_splits, M, N, C, scm, scn);
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// Inner layers (h x h weights, float activations): both small dimensions fit one CTA tile, so the generic 64x64x16 kernel
// spends its time on tile overhead (29 / 25 / 34+14 us for 15625 x 100 x 100 -- 11 % of a C5 epoch).  Here a CTA keeps BOTH
// operands of a 128-deep product in shared memory and every thread owns an 8 x 8 block of the 128 x 128 output tile
// (32 packed FMAs = 64 fp32 FMAs per 4 128-bit shared loads), fp32 FMA in ascending k like the generic kernel.
// Measured (CUDA graph replays, B200): 15 625 x 100 x 100 forward 14.0 us, dX 12.6, dW 20.1 incl. the chunk sum (cuBLAS through
// torch: 13.5 / 13.2 / 49.0); 180 000 rows 118 / 118 / 181 us (126 / 130 / 300).  Half of a CTA's time is the latency of
// filling its tiles and writing its rows, the other half the shared-memory reads of the product (16 floats per 64 FMAs).
//   MODE 0  C[m, n] = sum_k A[m, k] * W[n, k]      forward      (K, N <= 128; CTA = 128 rows m)
//   MODE 1  C[m, n] = sum_k A[m, k] * W[k, n]      dX           (K, N <= 128; CTA = 128 rows m)
//   MODE 2  P[c][i, j] = sum_{m in chunk c} A[m, i] * B[m, j]   weight gradient partials (I, J <= 128; chunk = 128 rows),
//           summed over chunks in ascending order by splitk_reduce_kernel
constexpr int SK = 128, SK_LD = SK + 4;

// 128-bit tile loaders of the VEC instances (all leading dimensions, widths and base addresses multiples of four floats):
// a thread issues its (up to) 16 loads before the first store, so a tile costs ONE round trip to L2 / HBM instead of the
// eight of the scalar path -- the kernel spent more time waiting for its tiles than multiplying them.
// dst[k][r] = src[(row0 + r) * ld + k] for r < 128 (zero where row0 + r >= rows), k < kdim.  A warp takes 16 rows x 2 quads of
// k: full 32-byte sectors on the global side, and the four transposing stores of a quad hit 32 different banks.
__device__ __forceinline__ void skinny_load_transposed(float* __restrict__ dst, const float* __restrict__ src, int64_t ld,
                                                       int64_t row0, int64_t rows, int kdim, int tid) {
    const int lane = tid & 31, r = (tid >> 5) * 16 + (lane & 15), kq0 = lane >> 4;
    const int64_t m = row0 + r;
    const bool row_ok = m < rows;
    const float* p = src + m * ld;
    float4 v[16];
#pragma unroll
    for (int q = 0; q < 16; ++q) {
        const int k = (q * 2 + kq0) * 4;
        v[q] = (row_ok && k < kdim) ? __ldg(reinterpret_cast<const float4*>(p + k)) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int q = 0; q < 16; ++q) {
        const int k = (q * 2 + kq0) * 4;
        if (k < kdim) {
            dst[(k + 0) * SK_LD + r] = v[q].x;
            dst[(k + 1) * SK_LD + r] = v[q].y;
            dst[(k + 2) * SK_LD + r] = v[q].z;
            dst[(k + 3) * SK_LD + r] = v[q].w;
        }
    }
}
// dst[k][c] = src[(row0 + k) * ld + c] for k < krows (zero where row0 + k >= rows or c >= cols), c < 128.
__device__ __forceinline__ void skinny_load_straight(float* __restrict__ dst, const float* __restrict__ src, int64_t ld,
                                                     int64_t row0, int64_t rows, int cols, int krows, int tid) {
    float4 v[16];
#pragma unroll
    for (int q = 0; q < 16; ++q) {
        const int f = tid + q * 256, k = f >> 5, c = (f & 31) * 4;
        v[q] = (k < krows && row0 + k < rows && c < cols) ? __ldg(reinterpret_cast<const float4*>(src + (row0 + k) * ld + c))
                                                          : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int q = 0; q < 16; ++q) {
        const int f = tid + q * 256, k = f >> 5, c = (f & 31) * 4;
        if (k < krows) *reinterpret_cast<float4*>(dst + k * SK_LD + c) = v[q];
    }
}

template <int MODE, bool VEC>
__global__ void __launch_bounds__(256, 2)
skinny_kernel(const float* __restrict__ A, int64_t lda, const float* __restrict__ B, int64_t ldb,
              float* __restrict__ C, int64_t ldc, int64_t M, int N, int K) {
    extern __shared__ __align__(16) float sk_smem[];
    const int kdim = MODE == 2 ? SK : K;                       // depth of the product held in shared memory
    float* As = sk_smem;                                       // [kdim][SK_LD]: As[k][row of the output tile]
    float* Bs = sk_smem + (size_t)kdim * SK_LD;                // [kdim][SK_LD]: Bs[k][column of the output tile]
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int64_t m0 = (int64_t)blockIdx.x * SK;
    if (VEC) {
        if (MODE == 2) {
            skinny_load_straight(As, A, lda, m0, M, N, SK, tid);
            skinny_load_straight(Bs, B, ldb, m0, M, K, SK, tid);
        } else {
            skinny_load_transposed(As, A, lda, m0, M, kdim, tid);
            if (MODE == 0) skinny_load_transposed(Bs, B, ldb, 0, N, kdim, tid);        // W[n, k] -> Bs[k][n]
            else           skinny_load_straight(Bs, B, ldb, 0, kdim, N, kdim, tid);    // W[k, n] -> Bs[k][n]
        }
    } else if (MODE == 2) {
        // scalar path: loads are issued in batches of 8 per thread before the first store, so that their latencies overlap
        // A[m, i] and B[m, j] are already k-major (k = m): straight copies, zero padding past the matrices
        for (int e0 = tid; e0 < SK * SK; e0 += 256 * 8) {
            float va[8], vb[8];
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                const int e = e0 + q * 256, k = e >> 7, c = e & 127;
                const int64_t m = m0 + k;
                va[q] = (m < M && c < N) ? A[m * lda + c] : 0.f;
                vb[q] = (m < M && c < K) ? B[m * ldb + c] : 0.f;
            }
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                const int e = e0 + q * 256, k = e >> 7, c = e & 127;
                As[k * SK_LD + c] = va[q];
                Bs[k * SK_LD + c] = vb[q];
            }
        }
    } else {
        // A[m, k] is contiguous along k: read along k (rows padded to 128 columns in the index space), store transposed
        for (int e0 = tid; e0 < SK * SK; e0 += 256 * 8) {
            float va[8];
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                const int e = e0 + q * 256, r = e >> 7, k = e & 127;
                const int64_t m = m0 + r;
                va[q] = (m < M && k < kdim) ? A[m * lda + k] : 0.f;
            }
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                const int e = e0 + q * 256, r = e >> 7, k = e & 127;
                if (k < kdim) As[k * SK_LD + r] = va[q];
            }
        }
        for (int e0 = tid; e0 < SK * SK; e0 += 256 * 8) {
            float vb[8];
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                const int e = e0 + q * 256, hi = e >> 7, lo = e & 127;
                if (MODE == 0) vb[q] = (hi < N && lo < kdim) ? B[(int64_t)hi * ldb + lo] : 0.f;      // W[n = hi, k = lo]
                else           vb[q] = (hi < kdim && lo < N) ? B[(int64_t)hi * ldb + lo] : 0.f;      // W[k = hi, n = lo]
            }
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                const int e = e0 + q * 256, hi = e >> 7, lo = e & 127;
                if (MODE == 0) { if (lo < kdim) Bs[lo * SK_LD + hi] = vb[q]; }
                else           { if (hi < kdim) Bs[hi * SK_LD + lo] = vb[q]; }
            }
        }
    }
    __syncthreads();
    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
#pragma unroll 4
    for (int k = 0; k < kdim; ++k) {
        // rows ty*4 + {0..3} and 64 + ty*4 + {0..3}, columns likewise in tx: a quarter warp's 128-bit loads then cover 128
        // contiguous bytes (no bank conflicts); with 8 consecutive columns per thread they were 2-way conflicted
        const float4 a0 = *reinterpret_cast<const float4*>(As + k * SK_LD + ty * 4);
        const float4 a1 = *reinterpret_cast<const float4*>(As + k * SK_LD + 64 + ty * 4);
        const float4 b0 = *reinterpret_cast<const float4*>(Bs + k * SK_LD + tx * 4);
        const float4 b1 = *reinterpret_cast<const float4*>(Bs + k * SK_LD + 64 + tx * 4);
        const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
        const float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
        // packed FMA (fma.rn.f32x2 -> FFMA2 with a broadcast scalar): the three-register scalar FFMA issues every other cycle
        // per scheduler on this part, FFMA2 does two of them in the same slot; each lane is the same IEEE fma as fmaf
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 8; j += 2) ffma2(acc[i][j], acc[i][j + 1], a[i], a[i], b[j], b[j + 1]);
    }
    if (VEC) {
        // rows of the output (or of the chunk's partial) are multiples of four floats wide and 16-byte aligned
        float* out = MODE == 2 ? C + (int64_t)blockIdx.x * N * K : C + m0 * ldc;
        const int64_t ldo = MODE == 2 ? K : ldc;
        const int64_t row_lim = MODE == 2 ? N : M - m0;
        const int col_lim = MODE == 2 ? K : N;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int r = (i < 4 ? 0 : 60) + ty * 4 + i;
            if (r < row_lim) {
                if (tx * 4 < col_lim)
                    *reinterpret_cast<float4*>(out + r * ldo + tx * 4) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
                if (64 + tx * 4 < col_lim)
                    *reinterpret_cast<float4*>(out + r * ldo + 64 + tx * 4) = make_float4(acc[i][4], acc[i][5], acc[i][6], acc[i][7]);
            }
        }
    } else if (MODE == 2) {
        // partial [N (= I) rows][K (= J) columns], dense, one per chunk
        float* P = C + (int64_t)blockIdx.x * N * K;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int r = (i < 4 ? 0 : 60) + ty * 4 + i;
            if (r < N)
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int c = (j < 4 ? 0 : 60) + tx * 4 + j;
                    if (c < K) P[r * K + c] = acc[i][j];
                }
        }
    } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int64_t m = m0 + (i < 4 ? 0 : 60) + ty * 4 + i;
            if (m < M)
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int c = (j < 4 ? 0 : 60) + tx * 4 + j;
                    if (c < N) C[m * ldc + c] = acc[i][j];
                }
        }
    }
}

// C[i] = sum_c part[c][i] for the chunk partials of MODE 2: 64 outputs per CTA, four thread groups each add a quarter of the
// chunks in ascending order, then the four sums are added in group order -- a fixed tree, so the result is reproducible.
__global__ void __launch_bounds__(256)
chunk_reduce_kernel(const float* __restrict__ part, int n_chunks, int64_t total, int64_t N, float* __restrict__ C, int64_t ldc) {
    __shared__ float red[4][64];
    const int o = threadIdx.x & 63, g = threadIdx.x >> 6;
    const int64_t i = (int64_t)blockIdx.x * 64 + o;
    const int per = (n_chunks + 3) / 4, c0 = g * per, c1 = min(n_chunks, c0 + per);
    float s = 0.f;
    if (i < total) {
#pragma unroll 8
        for (int c = c0; c < c1; ++c) s += part[(int64_t)c * total + i];
    }
    red[g][o] = s;
    __syncthreads();
    if (g == 0 && i < total) C[(i / N) * ldc + (i % N)] = ((red[0][o] + red[1][o]) + red[2][o]) + red[3][o];
}

static size_t skinny_smem(int kdim) { return (size_t)2 * kdim * SK_LD * sizeof(float); }

template <int MODE>
static int launch_skinny(const float* A, int64_t lda, const float* B, int64_t ldb, float* C, int64_t ldc, int64_t M, int N, int K,
                         cudaStream_t stream) {
    const size_t smem = skinny_smem(MODE == 2 ? SK : K);
    const unsigned grid = (unsigned)((M + SK - 1) / SK);
    // 128-bit loads and stores when every row of every operand starts on a 16-byte boundary and is a multiple of four floats wide
    // (MODE 2 writes dense [N, K] partials into the workspace: ldc is not used there)
    const bool vec = ((lda | ldb | (MODE == 2 ? 0 : ldc) | N | K) & 3) == 0 &&
                     (((uintptr_t)A | (uintptr_t)B | (uintptr_t)C) & 15) == 0;
    if (vec) {
        GNNB_SET_DYN_SMEM((skinny_kernel<MODE, true>), smem);
        skinny_kernel<MODE, true><<<grid, 256, smem, stream>>>(A, lda, B, ldb, C, ldc, M, N, K);
    } else {
        GNNB_SET_DYN_SMEM((skinny_kernel<MODE, false>), smem);
        skinny_kernel<MODE, false><<<grid, 256, smem, stream>>>(A, lda, B, ldb, C, ldc, M, N, K);
    }
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

static int pick_splits(int64_t M, int64_t N, int64_t K) {
    int64_t tiles = ((M + BM - 1) / BM) * ((N + BN - 1) / BN);
    int64_t want = (2LL * sm_count() + tiles - 1) / tiles;
    int64_t maxs = K / 256;               // at least 256 rows of K per split
    if (want > maxs) want = maxs;
    if (want > 64) want = 64;
    if (want < 1) want = 1;
    return (int)want;
}

// tensor-core implementations (linear_tc.cu)
int linear_tc_fwd(const void* X, int x_is_i8, int64_t ldx, int64_t n, int64_t p, const float* W, int64_t ldw, int64_t h,
                  float* H, int64_t ldh, int terms, void* ws, size_t ws_bytes, cudaStream_t stream);
int linear_tc_bwd_w(const void* X, int x_is_i8, int64_t ldx, int64_t n, int64_t p, const float* dH, int64_t lddh, int64_t h,
                    float* dW, int64_t lddw, int terms, void* ws, size_t ws_bytes, cudaStream_t stream);
size_t linear_tc_workspace(int64_t n, int64_t p, int64_t h, int terms);
// int8-digit implementation (linear_i8.cu), precision 4; shapes it does not take are REFUSED (GNNB_E_UNSUPPORTED), never
// run at another precision behind the caller's back -- gnnb_linear_supported tells a caller beforehand
bool linear_i8_supported(const void* X, int x_is_i8, int64_t ldx, int64_t n, int64_t p, int64_t h);
size_t linear_i8_workspace(int64_t n, int64_t p, int64_t h);
int linear_i8_fwd(const void* X, int x_format, int64_t ldx, int64_t n, int64_t p, const float* W, int64_t ldw, int64_t h,
                  float* H, int64_t ldh, void* ws, size_t ws_bytes, cudaStream_t stream);
int linear_i8_bwd_w(const void* X, int x_format, int64_t ldx, int64_t n, int64_t p, const float* dH, int64_t lddh, int64_t h,
                    float* dW, int64_t lddw, void* ws, size_t ws_bytes, cudaStream_t stream);
int64_t linear_i8_bwd_w_wave_markers(int x_format);
int linear_i8_prepare_dh(int x_format, int64_t n, const float* dH, int64_t lddh, int64_t h, void* ws, size_t ws_bytes, cudaStream_t stream);
int linear_i8_bwd_w_prepared(const void* X, int x_format, int64_t ldx, int64_t n, int64_t p, int64_t h, float* dW, int64_t lddw,
                             void* ws, size_t ws_bytes, cudaStream_t stream);

int linear_simt_fwd(const void* X, int x_is_i8, int64_t ldx, int64_t n, int64_t p, const float* W, int64_t ldw, int64_t h,
                    float* H, int64_t ldh, void* workspace, size_t workspace_bytes, cudaStream_t stream) {
    if (!x_is_i8 && p <= SK && h <= SK && n >= SK)                           // inner layer: both small dimensions in one tile
        return launch_skinny<0>((const float*)X, ldx, W, ldw, H, ldh, n, (int)h, (int)p, stream);
    int splits = pick_splits(n, h, p);
    size_t need = splits > 1 ? (size_t)splits * n * h * sizeof(float) : 0;
    if (need > 0 && (!workspace || workspace_bytes < need)) splits = 1;      // no scratch: single pass over K
    if (x_is_i8)
        return launch_gemm<int8_t, true, true>((const int8_t*)X, ldx, 1, W, 1, ldw, H, ldh, 1, n, h, p, splits,
                                               (float*)workspace, stream);
    return launch_gemm<float, true, true>((const float*)X, ldx, 1, W, 1, ldw, H, ldh, 1, n, h, p, splits,
                                          (float*)workspace, stream);
}

int linear_simt_bwd_w(const void* X, int x_is_i8, int64_t ldx, int64_t n, int64_t p, const float* dH, int64_t lddh, int64_t h,
                      float* dW, int64_t lddw, void* workspace, size_t workspace_bytes, cudaStream_t stream) {
    // dW[f, j] = sum_i dH[i, f] * X[i, j], run as dW^T so that the typed operand X is A:
    //   M = p, N = h, K = n, A(m, k) = X[k, m] (contiguous along m), B(k, n) = dH[k, n] (contiguous along n),
    //   C(m, n) -> dW[n * lddw + m]
    if (!x_is_i8 && p <= SK && h <= SK && n >= SK) {
        // dW[f, j] = sum_i dH[i, f] * X[i, j]: one [h, p] partial per 128-row chunk, then the fixed-order sum over chunks
        const int64_t chunks = (n + SK - 1) / SK;
        const size_t need_sk = (size_t)chunks * h * p * sizeof(float);
        if (workspace && workspace_bytes >= need_sk) {
            int rc = launch_skinny<2>(dH, lddh, (const float*)X, ldx, (float*)workspace, 0, n, (int)h, (int)p, stream);
            if (rc != GNNB_OK) return rc;
            chunk_reduce_kernel<<<(unsigned)((h * p + 63) / 64), 256, 0, stream>>>((const float*)workspace, (int)chunks, h * p, p, dW, lddw);
            GNNB_LAUNCH_CHECK();
            return GNNB_OK;
        }
    }
    int splits = pick_splits(p, h, n);
    size_t need = splits > 1 ? (size_t)splits * h * p * sizeof(float) : 0;
    if (need > 0 && (!workspace || workspace_bytes < need)) splits = 1;
    if (x_is_i8)
        return launch_gemm<int8_t, false, false>((const int8_t*)X, 1, ldx, dH, lddh, 1, dW, 1, lddw, p, h, n, splits,
                                                 (float*)workspace, stream);
    return launch_gemm<float, false, false>((const float*)X, 1, ldx, dH, lddh, 1, dW, 1, lddw, p, h, n, splits,
                                            (float*)workspace, stream);
}

}  // namespace gnnb

using namespace gnnb;

static int refuse_i8(const char* who, int x_format, int64_t n, int64_t p, int64_t h) {
    return fail(GNNB_E_UNSUPPORTED,
                "%s: precision 4 (int8 digits) takes int8 or 2-bit packed X with 16-byte aligned rows, n >= 128, p >= 128, h <= 104 and a "
                "contraction short enough for int32 digit sums (got format %d, n = %lld, p = %lld, h = %lld); ask gnnb_linear_supported "
                "and pick precision 3 (or expand packed X with gnnb_unpack_2bit) for other shapes",
                who, x_format, (long long)n, (long long)p, (long long)h);
}

extern "C" {

int gnnb_linear_workspace_bytes(int64_t n, int64_t p, int64_t h, int precision, size_t* bytes) {
    GNNB_REQUIRE(bytes && n >= 0 && p > 0 && h > 0 && precision >= 0 && precision <= 4, GNNB_E_BADARG,
                 "linear_workspace_bytes: bad argument");
    size_t out_max = (size_t)h * (size_t)(p > n ? p : n);
    // split-K partials (fwd: n*h, bwd_w: h*p): at most 64 splits, and never more than ~2 waves of 64x64 tiles
    size_t tiles_min = (out_max + 4095) / 4096;
    size_t max_splits = (2 * 148 + tiles_min - 1) / tiles_min;
    if (max_splits > 64) max_splits = 64;
    if (max_splits < 1) max_splits = 1;
    size_t simt = max_splits * out_max * sizeof(float) + 256;
    if (p <= SK && h <= SK) {                                  // inner-layer weight gradient: one [h, p] partial per 128 rows
        size_t sk = (size_t)((n + SK - 1) / SK) * (size_t)h * (size_t)p * sizeof(float) + 256;
        if (sk > simt) simt = sk;
    }
    size_t tc = precision > 0 ? linear_tc_workspace(n, p, h, precision == 4 ? 2 : precision) : 0;
    if (precision == 4) {
        size_t i8 = linear_i8_workspace(n, p, h);
        if (i8 > tc) tc = i8;
    }
    *bytes = simt > tc ? simt : tc;
    return GNNB_OK;
}

int gnnb_linear_fwd(const void* X, int x_is_i8, int64_t ldx, int64_t n, int64_t p,
                    const float* W, int64_t ldw, int64_t h, float* H, int64_t ldh, int precision,
                    void* workspace, size_t workspace_bytes, void* stream_) {
    GNNB_ENTER();
    cudaStream_t stream = (cudaStream_t)stream_;
    GNNB_REQUIRE(X && W && H && n >= 0 && p > 0 && h > 0 && ldx >= (x_is_i8 == 2 ? (p + 3) / 4 : p) && ldw >= p && ldh >= h, GNNB_E_BADARG, "linear_fwd: bad argument");
    GNNB_REQUIRE(x_is_i8 >= 0 && x_is_i8 <= 2, GNNB_E_BADARG, "linear_fwd: x format must be 0 (float32), 1 (int8) or 2 (2-bit packed)");
    GNNB_REQUIRE(precision >= 0 && precision <= 4, GNNB_E_BADARG, "linear_fwd: precision must be 0..4");
    if (n == 0) return GNNB_OK;
    if (precision == 4) {
        if (!linear_i8_supported(X, x_is_i8, ldx, n, p, h)) return refuse_i8("linear_fwd", x_is_i8, n, p, h);
        return linear_i8_fwd(X, x_is_i8, ldx, n, p, W, ldw, h, H, ldh, workspace, workspace_bytes, stream);
    }
    if (x_is_i8 == 2)
        return fail(GNNB_E_UNSUPPORTED, "linear_fwd: 2-bit packed X runs at precision 4 only (expand it with gnnb_unpack_2bit otherwise)");
    if (precision > 0) return linear_tc_fwd(X, x_is_i8, ldx, n, p, W, ldw, h, H, ldh, precision, workspace, workspace_bytes, stream);
    return linear_simt_fwd(X, x_is_i8, ldx, n, p, W, ldw, h, H, ldh, workspace, workspace_bytes, stream);
}

int gnnb_linear_bwd_w(const void* X, int x_is_i8, int64_t ldx, int64_t n, int64_t p,
                      const float* dH, int64_t lddh, int64_t h, float* dW, int64_t lddw, int precision,
                      void* workspace, size_t workspace_bytes, void* stream_) {
    GNNB_ENTER();
    cudaStream_t stream = (cudaStream_t)stream_;
    GNNB_REQUIRE(X && dH && dW && n >= 0 && p > 0 && h > 0 && ldx >= (x_is_i8 == 2 ? (p + 3) / 4 : p) && lddh >= h && lddw >= p, GNNB_E_BADARG, "linear_bwd_w: bad argument");
    GNNB_REQUIRE(x_is_i8 >= 0 && x_is_i8 <= 2, GNNB_E_BADARG, "linear_bwd_w: x format must be 0 (float32), 1 (int8) or 2 (2-bit packed)");
    GNNB_REQUIRE(precision >= 0 && precision <= 4, GNNB_E_BADARG, "linear_bwd_w: precision must be 0..4");
    if (n == 0) {
        GNNB_CUDA(cudaMemset2DAsync(dW, lddw * sizeof(float), 0, p * sizeof(float), h, stream));
        return GNNB_OK;
    }
    if (precision == 4) {
        if (!linear_i8_supported(X, x_is_i8, ldx, n, p, h)) return refuse_i8("linear_bwd_w", x_is_i8, n, p, h);
        return linear_i8_bwd_w(X, x_is_i8, ldx, n, p, dH, lddh, h, dW, lddw, workspace, workspace_bytes, stream);
    }
    if (x_is_i8 == 2)
        return fail(GNNB_E_UNSUPPORTED, "linear_bwd_w: 2-bit packed X runs at precision 4 only (expand it with gnnb_unpack_2bit otherwise)");
    if (precision > 0)
        return linear_tc_bwd_w(X, x_is_i8, ldx, n, p, dH, lddh, h, dW, lddw, precision, workspace, workspace_bytes, stream);
    return linear_simt_bwd_w(X, x_is_i8, ldx, n, p, dH, lddh, h, dW, lddw, workspace, workspace_bytes, stream);
}

int gnnb_linear_supported(const void* X, int x_is_i8, int64_t ldx, int64_t n, int64_t p, int64_t h, int precision) {
    if (precision < 0 || precision > 4 || x_is_i8 < 0 || x_is_i8 > 2 || n < 0 || p <= 0 || h <= 0) return 0;
    if (precision == 4) return linear_i8_supported(X, x_is_i8, ldx, n, p, h) ? 1 : 0;
    if (x_is_i8 == 2) return 0;
    if (precision == 0) return 1;
    return x_is_i8 == 1 ? 1 : 0;                               // tcgen05 bf16 tiles read int8 codes; float32 X runs on the fp32 tiles
}

int gnnb_linear_prepare_dh(int x_is_i8, int64_t n, const float* dH, int64_t lddh, int64_t h,
                           void* workspace, size_t workspace_bytes, void* stream_) {
    GNNB_ENTER();
    GNNB_REQUIRE(dH && n >= 128 && h > 0 && h <= 104 && lddh >= h && (x_is_i8 == 1 || x_is_i8 == 2), GNNB_E_BADARG,
                 "linear_prepare_dh: bad argument (int8 or packed X, n >= 128, h <= 104)");
    return linear_i8_prepare_dh(x_is_i8, n, dH, lddh, h, workspace, workspace_bytes, (cudaStream_t)stream_);
}

int gnnb_linear_bwd_w_prepared(const void* X, int x_is_i8, int64_t ldx, int64_t n, int64_t p, int64_t h,
                               float* dW, int64_t lddw, void* workspace, size_t workspace_bytes, void* stream_) {
    GNNB_ENTER();
    GNNB_REQUIRE(X && dW && n >= 128 && p > 0 && h > 0 && lddw >= p && ldx >= (x_is_i8 == 2 ? (p + 3) / 4 : p), GNNB_E_BADARG,
                 "linear_bwd_w_prepared: bad argument");
    if (!linear_i8_supported(X, x_is_i8, ldx, n, p < 128 ? 128 : p, h)) return refuse_i8("linear_bwd_w_prepared", x_is_i8, n, p, h);
    return linear_i8_bwd_w_prepared(X, x_is_i8, ldx, n, p, h, dW, lddw, workspace, workspace_bytes, (cudaStream_t)stream_);
}

int64_t gnnb_linear_bwd_w_wave_markers(int x_is_i8) {
    if (require_sm100() != GNNB_OK || (x_is_i8 != 1 && x_is_i8 != 2)) return 0;
    return linear_i8_bwd_w_wave_markers(x_is_i8);
}

int gnnb_linear_bwd_x(const float* dH, int64_t lddh, int64_t n, int64_t h,
                      const float* W, int64_t ldw, int64_t p, float* dX, int64_t lddx, void* stream_) {
    GNNB_ENTER();
    cudaStream_t stream = (cudaStream_t)stream_;
    GNNB_REQUIRE(dH && W && dX && n >= 0 && p > 0 && h > 0 && lddh >= h && ldw >= p && lddx >= p, GNNB_E_BADARG, "linear_bwd_x: bad argument");
    if (n == 0) return GNNB_OK;
    // dX[i, j] = sum_f dH[i, f] * W[f, j]: A = dH (contiguous along k), B = W (contiguous along n)
    if (p <= SK && h <= SK && n >= SK) return launch_skinny<1>(dH, lddh, W, ldw, dX, lddx, n, (int)p, (int)h, stream);
    return launch_gemm<float, true, false>(dH, lddh, 1, W, ldw, 1, dX, lddx, 1, n, p, h, 1, nullptr, stream);
}

}  // extern "C"
