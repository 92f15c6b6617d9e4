// Errors, device checks, staging kernels, scan and segmented sort.
#include <stdarg.h>

#include "common.cuh"

namespace gnnb {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

static int g_arch_ok[64];      // 0 unknown, 1 ok, -1 not sm_100
static int g_sms[64];

int require_sm100() {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess)
        return fail(GNNB_E_CUDA, "no CUDA device: %s (libgnn_b200 has no CPU fallback)", cudaGetErrorString(e));
    if (dev < 0 || dev >= 64) return fail(GNNB_E_BADARG, "device ordinal %d out of range", dev);
    if (g_arch_ok[dev] == 0) {
        int major = 0, minor = 0, sms = 0;
        cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
        cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        g_sms[dev] = sms;
        g_arch_ok[dev] = (major == 10 && minor == 0) ? 1 : -1;
    }
    if (g_arch_ok[dev] < 0)
        return fail(GNNB_E_ARCH, "device %d is not sm_100 (B200); this library is built for sm_100a only", dev);
    return GNNB_OK;
}

int sm_count() {
    int dev = 0;
    cudaGetDevice(&dev);
    return (dev >= 0 && dev < 64 && g_sms[dev] > 0) ? g_sms[dev] : 148;
}

// ------------------------------------------------------------------------------------ scan
__global__ void scan_reduce_kernel(const int32_t* __restrict__ in, int64_t n, int32_t* __restrict__ block_sums) {
    __shared__ int32_t warp_sums[8];
    int64_t base = (int64_t)blockIdx.x * 1024;
    int32_t s = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        int64_t j = base + threadIdx.x + i * 256;
        if (j < n) s += in[j];
    }
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) warp_sums[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        int32_t t = 0;
        for (int w = 0; w < 8; ++w) t += warp_sums[w];
        block_sums[blockIdx.x] = t;
    }
}

// one block: exclusive scan of block_sums in place; total to block_sums[nb]
__global__ void scan_top_kernel(int32_t* __restrict__ block_sums, int64_t nb, int64_t* __restrict__ total_out) {
    __shared__ int32_t tile[1024];
    __shared__ int32_t carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int64_t base = 0; base < nb; base += 1024) {
        int64_t j = base + threadIdx.x;
        int32_t v = (j < nb) ? block_sums[j] : 0;
        tile[threadIdx.x] = v;
        __syncthreads();
        for (int o = 1; o < 1024; o <<= 1) {           // Hillis-Steele inclusive scan
            int32_t t = (threadIdx.x >= o) ? tile[threadIdx.x - o] : 0;
            __syncthreads();
            tile[threadIdx.x] += t;
            __syncthreads();
        }
        int32_t incl = tile[threadIdx.x];
        int32_t c = carry;
        if (j < nb) block_sums[j] = c + incl - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry = c + incl;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        block_sums[nb] = carry;
        if (total_out) *total_out = (int64_t)carry;
    }
}

__global__ void scan_down_kernel(const int32_t* __restrict__ in, int32_t* __restrict__ out, int64_t n,
                                 const int32_t* __restrict__ block_sums, int64_t nb) {
    __shared__ int32_t warp_sums[8];
    int64_t base = (int64_t)blockIdx.x * 1024 + (int64_t)threadIdx.x * 4;   // 4 consecutive per thread
    int32_t v[4];
    int32_t s = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        v[i] = (base + i < n) ? in[base + i] : 0;
        s += v[i];
    }
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int32_t incl = s;
    for (int o = 1; o < 32; o <<= 1) {
        int32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    int32_t woff = 0;
    for (int w = 0; w < warp; ++w) woff += warp_sums[w];
    int32_t run = block_sums[blockIdx.x] + woff + incl - s;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        if (base + i < n) out[base + i] = run;
        run += v[i];
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) out[n] = block_sums[nb];
}

int scan_exclusive_i32(const int32_t* in, int32_t* out, int64_t n, int32_t* block_sums, int64_t* total_out,
                       cudaStream_t stream) {
    int64_t nb = (n + 1023) / 1024;
    if (nb == 0) nb = 1;
    scan_reduce_kernel<<<(unsigned)nb, 256, 0, stream>>>(in, n, block_sums);
    GNNB_LAUNCH_CHECK();
    scan_top_kernel<<<1, 1024, 0, stream>>>(block_sums, nb, total_out);
    GNNB_LAUNCH_CHECK();
    scan_down_kernel<<<(unsigned)nb, 256, 0, stream>>>(in, out, n, block_sums, nb);
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

// ------------------------------------------------------------------------------------ segmented sort
constexpr int SORT_SMALL = 256;        // rows up to this length: one warp, rank sort through shared memory

__global__ void __launch_bounds__(256) segsort_small_kernel(Rec* __restrict__ rec, const int32_t* __restrict__ seg_off,
                                                            const int32_t* __restrict__ seg_len, int64_t n_seg) {
    __shared__ Rec stage[8][SORT_SMALL];
    int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int64_t r = (int64_t)blockIdx.x * 8 + warp; r < n_seg; r += (int64_t)gridDim.x * 8) {
        int len = seg_len[r];
        if (len <= 1 || len > SORT_SMALL) continue;
        Rec* seg = rec + seg_off[r];
        for (int i = lane; i < len; i += 32) stage[warp][i] = seg[i];
        __syncwarp();
        for (int i = lane; i < len; i += 32) {
            unsigned long long key = stage[warp][i].key;
            int rank = 0;
            for (int j = 0; j < len; ++j) {
                unsigned long long kj = stage[warp][j].key;      // broadcast read
                rank += (kj < key) || (kj == key && j < i);
            }
            seg[rank] = stage[warp][i];
        }
        __syncwarp();
    }
}

// rows longer than SORT_SMALL: one CTA per row, bitonic network with ascending-only exchanges so the
// virtual +inf padding above `len` never moves; works in place in global memory.
__global__ void __launch_bounds__(256) segsort_large_kernel(Rec* __restrict__ rec, const int32_t* __restrict__ seg_off,
                                                            const int32_t* __restrict__ seg_len, int64_t n_seg) {
    for (int64_t r = blockIdx.x; r < n_seg; r += gridDim.x) {
        int len = seg_len[r];
        if (len <= SORT_SMALL) continue;          // uniform across the CTA
        Rec* seg = rec + seg_off[r];
        unsigned P = 1;
        while (P < (unsigned)len) P <<= 1;
        for (unsigned size = 2; size <= P; size <<= 1) {
            for (unsigned i = threadIdx.x; i < P; i += blockDim.x) {      // mirror step
                unsigned j = i ^ (size - 1);
                if (j > i && j < (unsigned)len) {
                    Rec a = seg[i], b = seg[j];
                    if (b.key < a.key) { seg[i] = b; seg[j] = a; }
                }
            }
            __syncthreads();
            for (unsigned stride = size >> 2; stride > 0; stride >>= 1) {
                for (unsigned i = threadIdx.x; i < P; i += blockDim.x) {
                    unsigned j = i ^ stride;
                    if (j > i && j < (unsigned)len) {
                        Rec a = seg[i], b = seg[j];
                        if (b.key < a.key) { seg[i] = b; seg[j] = a; }
                    }
                }
                __syncthreads();
            }
        }
    }
}

int segment_sort(Rec* rec, const int32_t* seg_off, const int32_t* seg_len, int64_t n_seg, cudaStream_t stream) {
    if (n_seg <= 0) return GNNB_OK;
    int sms = sm_count();
    int64_t want = (n_seg + 7) / 8;
    unsigned grid_small = (unsigned)(want < (int64_t)sms * 8 ? want : (int64_t)sms * 8);
    segsort_small_kernel<<<grid_small, 256, 0, stream>>>(rec, seg_off, seg_len, n_seg);
    GNNB_LAUNCH_CHECK();
    unsigned grid_large = (unsigned)(n_seg < (int64_t)sms * 4 ? n_seg : (int64_t)sms * 4);
    segsort_large_kernel<<<grid_large, 256, 0, stream>>>(rec, seg_off, seg_len, n_seg);
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

// ------------------------------------------------------------------------------------ staging kernels
__global__ void synth_kernel(int8_t* __restrict__ X, int64_t n, int64_t p, int64_t ld, int64_t row0, int64_t p_total,
                             unsigned long long seed, unsigned long long seed_freq) {
    // one thread per 4 consecutive markers of one row
    int64_t groups = (p + 3) / 4;
    int64_t total = n * groups;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        int64_t r = t / groups, j0 = (t % groups) * 4;
        unsigned long long i = (unsigned long long)(row0 + r);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            int64_t j = j0 + q;
            if (j >= p) break;
            unsigned long long hf = splitmix64_dev(seed_freq ^ (unsigned long long)j);
            unsigned long long thr = 214748365ULL + (((hf >> 32) * 1932735283ULL) >> 32);
            unsigned long long h = splitmix64_dev(seed ^ (i * (unsigned long long)p_total + (unsigned long long)j));
            int v = ((h & 0xFFFFFFFFULL) < thr) + ((h >> 32) < thr);
            X[r * ld + j] = (int8_t)v;
        }
    }
}

__global__ void gather_rows_kernel(const int8_t* __restrict__ X, int64_t ldx, const int64_t* __restrict__ rows,
                                   int64_t n_rows, int64_t p, int8_t* __restrict__ out, int64_t ldo, int vec16) {
    // one CTA per output row (grid-stride)
    for (int64_t r = blockIdx.x; r < n_rows; r += gridDim.x) {
        const int8_t* src = X + rows[r] * ldx;
        int8_t* dst = out + r * ldo;
        if (vec16) {
            const int4* s4 = reinterpret_cast<const int4*>(src);
            int4* d4 = reinterpret_cast<int4*>(dst);
            int64_t nv = p / 16;
            for (int64_t i = threadIdx.x; i < nv; i += blockDim.x) d4[i] = __ldg(s4 + i);
            for (int64_t i = nv * 16 + threadIdx.x; i < p; i += blockDim.x) dst[i] = src[i];
        } else {
            for (int64_t i = threadIdx.x; i < p; i += blockDim.x) dst[i] = src[i];
        }
    }
}

__global__ void i8_to_f32_kernel(const int8_t* __restrict__ X, int64_t ldx, int64_t n, int64_t p,
                                 float* __restrict__ out, int64_t ldo) {
    int64_t total = n * p;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        int64_t r = t / p, j = t % p;
        out[r * ldo + j] = (float)X[r * ldx + j];
    }
}

// Sum of val(0..n-1) in the order numpy's pairwise summation uses for a contiguous float64 reduction (blocks of <= 128
// elements with 8 running partial sums, halves split at a multiple of 8) -- the order behind pandas' DataFrame.std().
template <class F>
__device__ double np_pairwise_sum(int64_t n, F val) {
    int64_t lo[40], len[40];
    double acc[40];
    int st[40];
    int sp = 0;
    lo[0] = 0; len[0] = n; st[0] = 0; sp = 1;
    double ret = 0.0;
    while (sp > 0) {
        const int t = sp - 1;
        if (st[t] == 0) {
            const int64_t a = lo[t], m = len[t];
            if (m < 8) {
                double res = 0.0;
                for (int64_t i = 0; i < m; ++i) res = __dadd_rn(res, val(a + i));
                ret = res;
                --sp;
            } else if (m <= 128) {
                double r[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) r[j] = val(a + j);
                int64_t i = 8;
                for (; i < m - (m % 8); i += 8) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) r[j] = __dadd_rn(r[j], val(a + i + j));
                }
                double res = __dadd_rn(__dadd_rn(__dadd_rn(r[0], r[1]), __dadd_rn(r[2], r[3])),
                                       __dadd_rn(__dadd_rn(r[4], r[5]), __dadd_rn(r[6], r[7])));
                for (; i < m; ++i) res = __dadd_rn(res, val(a + i));
                ret = res;
                --sp;
            } else {
                int64_t n2 = m / 2;
                n2 -= n2 % 8;
                st[t] = 1;
                lo[sp] = a; len[sp] = n2; st[sp] = 0; ++sp;
            }
        } else if (st[t] == 1) {
            acc[t] = ret;
            st[t] = 2;
            int64_t n2 = len[t] / 2;
            n2 -= n2 % 8;
            lo[sp] = lo[t] + n2; len[sp] = len[t] - n2; st[sp] = 0; ++sp;
        } else {
            ret = __dadd_rn(acc[t], ret);
            --sp;
        }
    }
    return ret;
}

// `scaled` features (load.py:45-48): x -= mean; x /= std (ddof 1); NaN -> 0, column by column, with pandas' arithmetic:
// mean = exact integer sum / n; std = sqrt(pairwise_sum((avg - d)^2) / (n - 1)) where d = x - mean and avg = pairwise_sum(d) / n.
// One thread per column; neighbouring threads read neighbouring bytes of a row.
__global__ void scale_columns_kernel(const int8_t* __restrict__ X, int64_t ldx, int64_t n, int64_t p,
                                     double* __restrict__ out64, int64_t ld64, float* __restrict__ out32, int64_t ld32) {
    int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= p) return;
    const int8_t* col = X + c;
    long long s = 0;
    for (int64_t r = 0; r < n; ++r) s += col[r * ldx];
    const double mean = (double)s / (double)n;
    auto centred = [&](int64_t r) { return __dsub_rn((double)col[r * ldx], mean); };
    const double cnt = n > 1 ? (double)n : nan("");              // pandas: count <= ddof -> NaN
    const double avg = np_pairwise_sum(n, centred) / cnt;
    auto sqdev = [&](int64_t r) { double t = __dsub_rn(avg, centred(r)); return __dmul_rn(t, t); };
    const double var = np_pairwise_sum(n, sqdev) / (n > 1 ? (double)(n - 1) : nan(""));
    const double sd = sqrt(var);
    for (int64_t r = 0; r < n; ++r) {
        double v = centred(r) / sd;
        if (v != v) v = 0.0;                                       // fillna(0)
        if (out64) out64[r * ld64 + c] = v;
        if (out32) out32[r * ld32 + c] = (float)v;
    }
}

// 2-bit genotype packing (codes 0..3, four per byte, code j of a row in bits 2*(j%4).. of byte j/4): the host / PCIe
// format of the markers.  One thread expands one 32-bit word (16 codes) into one 16-byte store.
__device__ __forceinline__ uint32_t expand_2bit_byte(uint32_t b) {
    return (b & 0x3u) | ((b & 0xcu) << 6) | ((b & 0x30u) << 12) | ((b & 0xc0u) << 18);
}

__global__ void unpack_2bit_vec_kernel(const uint8_t* __restrict__ P, int64_t ldp, int64_t n, int64_t p,
                                       int8_t* __restrict__ out, int64_t ldo) {
    const int64_t words = (p + 15) / 16;
    const int64_t total = n * words;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = t / words, w = t % words;
        const uint32_t v = *reinterpret_cast<const uint32_t*>(P + r * ldp + w * 4);
        uint4 o;
        o.x = expand_2bit_byte(v & 0xffu);
        o.y = expand_2bit_byte((v >> 8) & 0xffu);
        o.z = expand_2bit_byte((v >> 16) & 0xffu);
        o.w = expand_2bit_byte(v >> 24);
        int8_t* dst = out + r * ldo + w * 16;
        if (w * 16 + 16 <= ldo) {
            *reinterpret_cast<uint4*>(dst) = o;                 // the padding columns of the row receive zeros
        } else {
            const uint32_t q[4] = {o.x, o.y, o.z, o.w};
            for (int64_t j = w * 16; j < p; ++j) dst[j - w * 16] = (int8_t)((q[(j & 15) >> 2] >> (8 * (j & 3))) & 0xffu);
        }
    }
}

__global__ void unpack_2bit_kernel(const uint8_t* __restrict__ P, int64_t ldp, int64_t n, int64_t p,
                                   int8_t* __restrict__ out, int64_t ldo) {
    const int64_t total = n * p;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = t / p, j = t % p;
        out[r * ldo + j] = (int8_t)((P[r * ldp + (j >> 2)] >> (2 * (j & 3))) & 3);
    }
}

__global__ void pack_2bit_kernel(const int8_t* __restrict__ X, int64_t ldx, int64_t n, int64_t p,
                                 uint8_t* __restrict__ out, int64_t ldo, int32_t* __restrict__ bad) {
    const int64_t nb = (p + 3) / 4;
    const int64_t total = n * nb;
    int any_bad = 0;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = t / nb, b = t % nb;
        uint32_t v = 0;
        for (int q = 0; q < 4; ++q) {
            const int64_t j = b * 4 + q;
            if (j < p) {
                const int c = X[r * ldx + j];
                any_bad |= (c < 0 || c > 3);
                v |= (uint32_t)(c & 3) << (2 * q);
            }
        }
        out[r * ldo + b] = (uint8_t)v;
    }
    if (any_bad && bad) atomicExch(bad, 1);
}

__global__ void sqnorm_kernel(const int8_t* __restrict__ X, int64_t ldx, int64_t n, int64_t p, int64_t* __restrict__ sq) {
    // one warp per row
    int lane = threadIdx.x & 31;
    int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t r = warp; r < n; r += nwarps) {
        const int8_t* row = X + r * ldx;
        long long s = 0;
        for (int64_t j = lane; j < p; j += 32) {
            int v = row[j];
            s += v * v;
        }
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) sq[r] = s;
    }
}

// column sums: out[f] = sum_i A[i, f]; deterministic two-stage (fixed partition of rows, fixed order inside a block)
// stage 1: block b owns rows [b * rows_per_block, ...): 8 row-lanes x 32 columns per pass, lanes combined in lane order
__global__ void __launch_bounds__(256) colsum_partial_kernel(const float* __restrict__ A, int64_t lda, int64_t n, int64_t F,
                                                             float* __restrict__ partial, int rows_per_block) {
    __shared__ float red[8][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
    const int64_t r1 = r0 + rows_per_block < n ? r0 + rows_per_block : n;
    for (int64_t f0 = 0; f0 < F; f0 += 32) {
        const int64_t f = f0 + tx;
        float s = 0.f;
        if (f < F)
            for (int64_t r = r0 + ty; r < r1; r += 8) s += A[r * lda + f];
        red[ty][tx] = s;
        __syncthreads();
        if (ty == 0 && f < F) {
            float t = red[0][tx];
#pragma unroll
            for (int q = 1; q < 8; ++q) t += red[q][tx];
            partial[(int64_t)blockIdx.x * F + f] = t;
        }
        __syncthreads();
    }
}
// stage 2: one warp per column, lanes stride over the blocks, fixed butterfly order
__global__ void __launch_bounds__(256) colsum_final_kernel(const float* __restrict__ partial, int nblocks, int64_t F, float* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t f = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (f >= F) return;
    float s = 0.f;
    for (int b = lane; b < nblocks; b += 32) s += partial[(int64_t)b * F + f];
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) out[f] = s;
}


// ------------------------------------------------------------------------------------ base-3 host / PCIe format
// Genotypes are 0 / 1 / 2, so five of them fit one byte (3^5 = 243): value = c0 + 3 c1 + 9 c2 + 27 c3 + 81 c4.  That is the
// densest byte-aligned form (1.6 bits per code against 2) and what crosses PCIe when the genotypes live on the host; on the
// device it is re-coded once into the 2-bit form the projection kernels read.  One thread = 16 input bytes = 80 codes =
// 20 output bytes; the byte -> ten-bit table sits in shared memory.  HBM-bound: 0.2 B read + 0.25 B written per code.
__global__ void __launch_bounds__(256)
base3_to_2bit_kernel(const uint8_t* __restrict__ in, int64_t ldi, int64_t n, int64_t p, uint8_t* __restrict__ out, int64_t ldo,
                     int32_t* __restrict__ bad_flag) {
    __shared__ uint16_t lut[256];
    {
        const int b = threadIdx.x;                         // blockDim.x == 256
        int v = b, bits = 0;
        for (int d = 0; d < 5; ++d) { bits |= (v % 3) << (2 * d); v /= 3; }
        lut[b] = b < 243 ? (uint16_t)bits : (uint16_t)0xFFFF;
    }
    __syncthreads();
    const int64_t groups = (p + 79) / 80, total = n * groups;
    const int64_t in_bytes = (p + 4) / 5, out_bytes = (p + 3) / 4;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / groups, g = i % groups;
        const uint8_t* src = in + r * ldi + g * 16;
        uint8_t b[16];
        if ((g + 1) * 16 <= ldi && ((reinterpret_cast<uintptr_t>(src) & 15) == 0)) {
            *reinterpret_cast<uint4*>(b) = *reinterpret_cast<const uint4*>(src);
        } else {
#pragma unroll
            for (int q = 0; q < 16; ++q) b[q] = (g * 16 + q < in_bytes) ? src[q] : 0;
        }
        unsigned long long c[4];
        bool bad = false;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const unsigned l0 = lut[b[4 * q]], l1 = lut[b[4 * q + 1]], l2 = lut[b[4 * q + 2]], l3 = lut[b[4 * q + 3]];
            bad |= ((l0 | l1 | l2 | l3) & 0x8000u) != 0;
            c[q] = (unsigned long long)(l0 & 1023u) | ((unsigned long long)(l1 & 1023u) << 10) | ((unsigned long long)(l2 & 1023u) << 20) |
                   ((unsigned long long)(l3 & 1023u) << 30);
        }
        if (bad && bad_flag) *bad_flag = 1;
        uint32_t w[5];
        w[0] = (uint32_t)c[0];
        w[1] = (uint32_t)(c[0] >> 32) | (uint32_t)(c[1] << 8);
        w[2] = (uint32_t)(c[1] >> 24) | (uint32_t)(c[2] << 16);
        w[3] = (uint32_t)(c[2] >> 16) | (uint32_t)(c[3] << 24);
        w[4] = (uint32_t)(c[3] >> 8);
        uint8_t* dst = out + r * ldo + g * 20;
        if ((g + 1) * 20 <= out_bytes && ((reinterpret_cast<uintptr_t>(dst) & 3) == 0)) {
#pragma unroll
            for (int q = 0; q < 5; ++q) reinterpret_cast<uint32_t*>(dst)[q] = w[q];
        } else {
            // last group of a row: only the bytes that hold codes; codes past p inside the last byte are cleared
            for (int q = 0; q < 20 && g * 20 + q < out_bytes; ++q) {
                uint8_t v = (uint8_t)(w[q >> 2] >> (8 * (q & 3)));
                const int64_t first = (g * 20 + q) * 4;
                if (first + 4 > p) v &= (uint8_t)((1u << (2 * (p - first))) - 1u);
                dst[q] = v;
            }
        }
    }
}


// ------------------------------------------------------------------------------------ e2m1 (FP4) operand form of the search
// 0 / 1 / 2 are exact e2m1 values (nibbles 0b0000, 0b0010, 0b0100 = code << 1); two codes per byte, code 2 j in the low
// nibble.  The FP4 tensor path of the neighbour search (kind::mxf4, twice the int8 rate) reads this form.  One thread = 16
// codes = 8 output bytes; HBM-bound (1 B read + 0.5 B written per code).  Padding bytes up to ldo are zeroed (0.0).
__global__ void pack_fp4_kernel(const int8_t* __restrict__ X, int64_t ldx, int64_t n, int64_t p, uint8_t* __restrict__ out, int64_t ldo,
                                int32_t* __restrict__ bad_flag) {
    const int64_t groups = ldo / 8, total = n * groups;                // ldo is a multiple of 16
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / groups, c0 = (i % groups) * 16;
        const int8_t* src = X + r * ldx + c0;
        int8_t c[16];
        if (c0 + 16 <= p && ((reinterpret_cast<uintptr_t>(src) & 15) == 0)) {
            *reinterpret_cast<uint4*>(c) = *reinterpret_cast<const uint4*>(src);
        } else {
#pragma unroll
            for (int q = 0; q < 16; ++q) c[q] = (c0 + q < p) ? src[q] : (int8_t)0;
        }
        unsigned long long w = 0;
        bool bad = false;
#pragma unroll
        for (int q = 0; q < 16; ++q) {
            bad |= (c[q] < 0) | (c[q] > 2);
            w |= (unsigned long long)((c[q] & 3) << 1) << (4 * q);
        }
        if (bad && bad_flag) *bad_flag = 1;
        *reinterpret_cast<unsigned long long*>(out + r * ldo + c0 / 2) = w;
    }
}

}  // namespace gnnb

using namespace gnnb;

extern "C" {

int gnnb_version(void) { return 100; }                 // 0.1.0
const char* gnnb_last_error(void) { return g_err; }
int gnnb_check_device(void) { return require_sm100(); }

int gnnb_synth_genotypes(int8_t* X, int64_t n, int64_t p, int64_t ld, int64_t row0, int64_t p_total,
                         uint64_t seed, uint64_t seed_freq, void* stream) {
    GNNB_ENTER();
    GNNB_REQUIRE(X && n >= 0 && p >= 0 && ld >= p && p_total >= p, GNNB_E_BADARG, "synth_genotypes: bad shape");
    if (n == 0 || p == 0) return GNNB_OK;
    int64_t total = n * ((p + 3) / 4);
    int64_t blocks = (total + 255) / 256;
    int64_t cap = (int64_t)sm_count() * 32;
    synth_kernel<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, (cudaStream_t)stream>>>(
        X, n, p, ld, row0, p_total, seed, seed_freq);
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

int gnnb_gather_rows_i8(const int8_t* X, int64_t ldx, const int64_t* rows, int64_t n_rows, int64_t p,
                        int8_t* out, int64_t ldo, void* stream) {
    GNNB_ENTER();
    GNNB_REQUIRE(n_rows >= 0 && p > 0 && ldx >= p && ldo >= p, GNNB_E_BADARG, "gather_rows: bad shape");
    if (n_rows == 0) return GNNB_OK;
    GNNB_REQUIRE(X && rows && out, GNNB_E_BADARG, "gather_rows: null pointer");
    int vec16 = (ldx % 16 == 0) && (ldo % 16 == 0) && ((uintptr_t)X % 16 == 0) && ((uintptr_t)out % 16 == 0);
    int64_t cap = (int64_t)sm_count() * 16;
    gather_rows_kernel<<<(unsigned)(n_rows < cap ? n_rows : cap), 256, 0, (cudaStream_t)stream>>>(
        X, ldx, rows, n_rows, p, out, ldo, vec16);
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

int gnnb_i8_to_f32(const int8_t* X, int64_t ldx, int64_t n, int64_t p, float* out, int64_t ldo, void* stream) {
    GNNB_ENTER();
    GNNB_REQUIRE(X && out && n >= 0 && p > 0 && ldx >= p && ldo >= p, GNNB_E_BADARG, "i8_to_f32: bad argument");
    if (n == 0) return GNNB_OK;
    int64_t blocks = (n * p + 255) / 256;
    int64_t cap = (int64_t)sm_count() * 32;
    i8_to_f32_kernel<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, (cudaStream_t)stream>>>(X, ldx, n, p, out, ldo);
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

int gnnb_scale_columns_i8(const int8_t* X, int64_t ldx, int64_t n, int64_t p, double* out64, int64_t ld64,
                          float* out32, int64_t ld32, void* stream) {
    GNNB_ENTER();
    GNNB_REQUIRE(X && (out64 || out32) && n >= 0 && p > 0 && ldx >= p, GNNB_E_BADARG, "scale_columns: bad argument");
    GNNB_REQUIRE((!out64 || ld64 >= p) && (!out32 || ld32 >= p), GNNB_E_BADARG, "scale_columns: output stride < p");
    if (n == 0) return GNNB_OK;
    scale_columns_kernel<<<(unsigned)((p + 127) / 128), 128, 0, (cudaStream_t)stream>>>(X, ldx, n, p, out64, ld64, out32, ld32);
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

int gnnb_unpack_2bit(const uint8_t* P, int64_t ldp, int64_t n, int64_t p, int8_t* out, int64_t ldo, void* stream) {
    GNNB_ENTER();
    GNNB_REQUIRE(P && out && n >= 0 && p > 0 && ldp >= (p + 3) / 4 && ldo >= p, GNNB_E_BADARG, "unpack_2bit: bad argument");
    if (n == 0) return GNNB_OK;
    int64_t cap = (int64_t)sm_count() * 32;
    bool vec = ((uintptr_t)P % 4 == 0) && (ldp % 4 == 0) && ldp >= ((p + 15) / 16) * 4 &&
               ((uintptr_t)out % 16 == 0) && (ldo % 16 == 0);
    if (vec) {
        int64_t blocks = (n * ((p + 15) / 16) + 255) / 256;
        unpack_2bit_vec_kernel<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, (cudaStream_t)stream>>>(P, ldp, n, p, out, ldo);
    } else {
        int64_t blocks = (n * p + 255) / 256;
        unpack_2bit_kernel<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, (cudaStream_t)stream>>>(P, ldp, n, p, out, ldo);
    }
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

int gnnb_pack_2bit(const int8_t* X, int64_t ldx, int64_t n, int64_t p, uint8_t* out, int64_t ldo, int32_t* bad_flag,
                   void* stream) {
    GNNB_ENTER();
    GNNB_REQUIRE(X && out && n >= 0 && p > 0 && ldx >= p && ldo >= (p + 3) / 4, GNNB_E_BADARG, "pack_2bit: bad argument");
    if (n == 0) return GNNB_OK;
    int64_t blocks = (n * ((p + 3) / 4) + 255) / 256;
    int64_t cap = (int64_t)sm_count() * 32;
    pack_2bit_kernel<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, (cudaStream_t)stream>>>(X, ldx, n, p, out, ldo, bad_flag);
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

int gnnb_base3_to_2bit(const uint8_t* P3, int64_t ldi, int64_t n, int64_t p, uint8_t* out, int64_t ldo, int32_t* bad_flag,
                       void* stream) {
    GNNB_ENTER();
    GNNB_REQUIRE(P3 && out && n >= 0 && p > 0 && ldi >= (p + 4) / 5 && ldo >= (p + 3) / 4, GNNB_E_BADARG, "base3_to_2bit: bad argument");
    if (n == 0) return GNNB_OK;
    int64_t blocks = (n * ((p + 79) / 80) + 255) / 256;
    int64_t cap = (int64_t)sm_count() * 32;
    base3_to_2bit_kernel<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, (cudaStream_t)stream>>>(P3, ldi, n, p, out, ldo, bad_flag);
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

int gnnb_pack_fp4(const int8_t* X, int64_t ldx, int64_t n, int64_t p, uint8_t* out, int64_t ldo, int32_t* bad_flag, void* stream) {
    GNNB_ENTER();
    GNNB_REQUIRE(X && out && n >= 0 && p > 0 && ldx >= p && ldo >= (p + 1) / 2 && ldo % 16 == 0 && (uintptr_t)out % 16 == 0, GNNB_E_BADARG,
                 "pack_fp4: bad argument (ldo must be a multiple of 16 bytes >= ceil(p / 2))");
    if (n == 0) return GNNB_OK;
    int64_t blocks = (n * (ldo / 8) + 255) / 256;
    int64_t cap = (int64_t)sm_count() * 32;
    pack_fp4_kernel<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, (cudaStream_t)stream>>>(X, ldx, n, p, out, ldo, bad_flag);
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

int gnnb_row_sqnorm_i8(const int8_t* X, int64_t ldx, int64_t n, int64_t p, int64_t* sq, void* stream) {
    GNNB_ENTER();
    GNNB_REQUIRE(X && sq && n >= 0 && p > 0 && ldx >= p, GNNB_E_BADARG, "row_sqnorm: bad argument");
    if (n == 0) return GNNB_OK;
    int64_t blocks = (n + 7) / 8;
    int64_t cap = (int64_t)sm_count() * 16;
    sqnorm_kernel<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, (cudaStream_t)stream>>>(X, ldx, n, p, sq);
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

int gnnb_colsum_workspace_bytes(int64_t F, size_t* bytes) {
    GNNB_REQUIRE(bytes && F > 0, GNNB_E_BADARG, "colsum_workspace_bytes: bad argument");
    *bytes = (size_t)1024 * (size_t)F * sizeof(float);
    return GNNB_OK;
}

int gnnb_colsum(const float* A, int64_t lda, int64_t n, int64_t F, float* out,
                void* workspace, size_t workspace_bytes, void* stream) {
    GNNB_ENTER();
    GNNB_REQUIRE(A && out && n >= 0 && F > 0 && lda >= F, GNNB_E_BADARG, "colsum: bad argument");
    GNNB_REQUIRE(workspace && workspace_bytes >= (size_t)1024 * F * sizeof(float), GNNB_E_WORKSPACE,
                 "colsum: workspace too small");
    int64_t nblocks = (n + 63) / 64;
    if (nblocks > 1024) nblocks = 1024;
    if (nblocks < 1) nblocks = 1;
    int rows_per_block = (int)((n + nblocks - 1) / nblocks);
    if (rows_per_block < 1) rows_per_block = 1;
    float* partial = static_cast<float*>(workspace);
    colsum_partial_kernel<<<(unsigned)nblocks, 256, 0, (cudaStream_t)stream>>>(A, lda, n, F, partial, rows_per_block);
    GNNB_LAUNCH_CHECK();
    colsum_final_kernel<<<(unsigned)((F * 32 + 255) / 256), 256, 0, (cudaStream_t)stream>>>(partial, (int)nblocks, F, out);
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

// ------------------------------------------------------------------------------------ allocation-free column sums
// Scratch for gnnb_colsum_static: 1024 row-block partials x GNNB_COLSUM_STATIC_MAX_F columns of float32 (2 MiB),
// resident in every GPU that loads this library until the process exits.
static __device__ __align__(256) float g_colsum_static_scratch[1024 * GNNB_COLSUM_STATIC_MAX_F];
static float* g_colsum_static_ptr[64];      // device address of the array above, per device ordinal (0 = not resolved yet)

/* Column sums without a workspace argument and without any allocation.
 *
 * Who this is for: code that computes the GCN bias gradient (db = sum over nodes of dOut) from a place where
 * memory cannot be requested -- a training step that an application replays under a fixed memory plan, a serving
 * process that hands the library a closed arena and forbids anything else, or a host thread that must not take
 * the allocator's lock while other streams are running.  Such callers can pass their input and output buffers but
 * have nowhere to get the row-block partials gnnb_colsum needs, so the partials live in a fixed array that is part
 * of the library image.
 *
 * The name of that array is only a host-side handle; the kernels need its address on the current device.  The first
 * call on each device resolves it with cudaGetSymbolAddress (checked) and remembers it; make that first call during
 * start-up, before any stream capture begins.  Limits and rules:
 *   - F must not exceed GNNB_COLSUM_STATIC_MAX_F (512); larger inputs are refused with GNNB_E_UNSUPPORTED;
 *   - every call on a device uses the same array and leaves its partial sums behind.  Nothing has to start at zero:
 *     each call overwrites every partial it later reads, so stale contents are never observed;
 *   - because the array is shared, two calls whose kernels can overlap on the GPU (different streams, no ordering)
 *     race on it.  Serialise them, or use gnnb_colsum with separate workspaces.
 * The result is bit-identical to gnnb_colsum (same kernels, same fixed summation order).
 */
int gnnb_colsum_static(const float* A, int64_t lda, int64_t n, int64_t F, float* out, void* stream) {
    GNNB_ENTER();
    GNNB_REQUIRE((A || n == 0) && out && n >= 0 && F > 0 && lda >= F, GNNB_E_BADARG, "colsum_static: bad argument");
    GNNB_REQUIRE(F <= GNNB_COLSUM_STATIC_MAX_F, GNNB_E_UNSUPPORTED,
                 "colsum_static: F = %lld exceeds the built-in scratch (at most %d columns); use gnnb_colsum with a workspace",
                 (long long)F, GNNB_COLSUM_STATIC_MAX_F);
    int dev = 0;
    GNNB_CUDA(cudaGetDevice(&dev));
    GNNB_REQUIRE(dev >= 0 && dev < 64, GNNB_E_BADARG, "colsum_static: device ordinal %d out of range", dev);
    if (!g_colsum_static_ptr[dev]) {            // once per device, with that device current
        float* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&p, g_colsum_static_scratch);
        if (e != cudaSuccess || !p)
            return fail(GNNB_E_CUDA, "colsum_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
        // two host threads making their first call on one device at the same time both get here and both store the
        // same address (a pointer-sized store of an identical value): harmless, so the table is not locked
        g_colsum_static_ptr[dev] = p;
    }
    float* partial = g_colsum_static_ptr[dev];
    int64_t nblocks = (n + 63) / 64;
    if (nblocks > 1024) nblocks = 1024;
    if (nblocks < 1) nblocks = 1;
    int rows_per_block = (int)((n + nblocks - 1) / nblocks);
    if (rows_per_block < 1) rows_per_block = 1;
    colsum_partial_kernel<<<(unsigned)nblocks, 256, 0, (cudaStream_t)stream>>>(A, lda, n, F, partial, rows_per_block);
    GNNB_LAUNCH_CHECK();
    colsum_final_kernel<<<(unsigned)((F * 32 + 255) / 256), 256, 0, (cudaStream_t)stream>>>(partial, (int)nblocks, F, out);
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

}  // extern "C"
