// Shared helpers for libgnn_b200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>

#include "../../include/gnnb.h"

namespace gnnb {

void set_error(const char* fmt, ...);
int  fail(int code, const char* fmt, ...);
int  require_sm100();                       // GNNB_OK or GNNB_E_ARCH (cached per device)
int  sm_count();

#define GNNB_REQUIRE(cond, code, ...)                       \
    do {                                                    \
        if (!(cond)) return ::gnnb::fail((code), __VA_ARGS__); \
    } while (0)

#define GNNB_CUDA(expr)                                                                   \
    do {                                                                                  \
        cudaError_t _e = (expr);                                                          \
        if (_e != cudaSuccess)                                                            \
            return ::gnnb::fail(GNNB_E_CUDA, "%s failed: %s (%s:%d)", #expr,              \
                                cudaGetErrorString(_e), __FILE__, __LINE__);              \
    } while (0)

#define GNNB_LAUNCH_CHECK()                                                               \
    do {                                                                                  \
        cudaError_t _e = cudaGetLastError();                                              \
        if (_e != cudaSuccess)                                                            \
            return ::gnnb::fail(GNNB_E_CUDA, "kernel launch failed: %s (%s:%d)",          \
                                cudaGetErrorString(_e), __FILE__, __LINE__);              \
    } while (0)

// Opt a kernel in to `bytes` of dynamic shared memory.  The attribute belongs to the (function, DEVICE) pair, so the
// high-water mark is kept per device ordinal: a process that drives several GPUs must set it on each of them.
#define GNNB_MAX_DEVICES 64
#define GNNB_SET_DYN_SMEM(kernel, bytes)                                                                     \
    do {                                                                                                      \
        static size_t hw_[GNNB_MAX_DEVICES] = {0};                                                            \
        int dev_ = 0;                                                                                         \
        GNNB_CUDA(cudaGetDevice(&dev_));                                                                      \
        const bool tracked_ = dev_ >= 0 && dev_ < GNNB_MAX_DEVICES;                                           \
        if (!tracked_ || (size_t)(bytes) > hw_[dev_]) {                                                       \
            GNNB_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(bytes))); \
            if (tracked_) hw_[dev_] = (size_t)(bytes);                                                        \
        }                                                                                                     \
    } while (0)

#define GNNB_ENTER()                                        \
    do {                                                    \
        int _rc = ::gnnb::require_sm100();                  \
        if (_rc != GNNB_OK) return _rc;                     \
    } while (0)

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// Carves a caller-provided workspace into 256-byte aligned pieces.
struct Carver {
    char*  base;
    size_t off;
    size_t cap;
    Carver(void* p, size_t bytes) : base(static_cast<char*>(p)), off(0), cap(bytes) {}
    template <typename T>
    T* take(size_t count) {
        off = align_up(off, 256);
        T* r = reinterpret_cast<T*>(base + off);
        off += count * sizeof(T);
        return r;
    }
    bool ok() const { return off <= cap; }
};

// exclusive scan of int32 counts into int32 offsets (out has n + 1 entries; out[n] = total);
// block_sums needs ceil(n / 1024) + 1 int32.  total is also added to *total_out (int64) when non-null.
int scan_exclusive_i32(const int32_t* in, int32_t* out, int64_t n, int32_t* block_sums, int64_t* total_out,
                       cudaStream_t stream);
static inline size_t scan_blocks(int64_t n) { return (size_t)((n + 1023) / 1024 + 1); }

// 16-byte record sorted per segment by `key`
struct __align__(16) Rec {
    unsigned long long key;
    unsigned long long payload;
};
// sorts rec[seg_off[r] .. seg_off[r] + seg_len[r]) ascending by key for every r < n_seg
int segment_sort(Rec* rec, const int32_t* seg_off, const int32_t* seg_len, int64_t n_seg, cudaStream_t stream);

__device__ __forceinline__ unsigned long long splitmix64_dev(unsigned long long z) {
    z += 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

// (d0, d1) = (a0 * b0 + d0, a1 * b1 + d1), each lane one IEEE fp32 fma (round to nearest even) -- fma.rn.f32x2, SASS FFMA2.
// With a0 == a1 the assembler uses the broadcast-scalar form, so no packing moves are spent on the shared factor.
__device__ __forceinline__ void ffma2(float& d0, float& d1, float a0, float a1, float b0, float b1) {
    unsigned long long d, a, b;
    asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(d0), "f"(d1));
    asm("mov.b64 %0, {%1, %2};" : "=l"(a) : "f"(a0), "f"(a1));
    asm("mov.b64 %0, {%1, %2};" : "=l"(b) : "f"(b0), "f"(b1));
    asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(d) : "l"(a), "l"(b));
    asm("mov.b64 {%0, %1}, %2;" : "=f"(d0), "=f"(d1) : "l"(d));
}

}  // namespace gnnb
