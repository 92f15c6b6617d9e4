// Per-row running selection of the k smallest (d2, index) pairs, shared by the dp4a and tcgen05 kNN kernels.
#pragma once
#include <math.h>

namespace gnnb {

constexpr int KNN_MAX_EXTRA = 4;      // float64 side columns (the phenotype column of util.py:52-53 needs 1)
constexpr int KNN_MAX_SPLITS = 32;    // key splits whose partial lists are merged afterwards

// One thread owns one query row.  The sorted list lives in shared memory, transposed ([slot][row]) so
// that neighbouring threads touch neighbouring banks.  The current k-th best pair is cached in
// registers: almost every candidate is rejected by one compare.
struct RowSelector {
    double* ld;      // [k][rows]
    int*    li;      // [k][rows]
    int     rows, r, k;
    double  thr_d;
    unsigned thr_i;

    __device__ __forceinline__ void init(double* list_d, int* list_i, int rows_, int r_, int k_) {
        ld = list_d; li = list_i; rows = rows_; r = r_; k = k_;
        for (int t = 0; t < k; ++t) {
            ld[t * rows + r] = INFINITY;
            li[t * rows + r] = 0x7fffffff;
        }
        thr_d = INFINITY;
        thr_i = 0x7fffffffu;
    }
    __device__ __forceinline__ double d(int t) const { return ld[t * rows + r]; }
    __device__ __forceinline__ int i(int t) const { return li[t * rows + r]; }

    __device__ __forceinline__ void offer(double dist, int j) {
        if (!(dist < thr_d || (dist == thr_d && (unsigned)j < thr_i))) return;
        int pos = k - 1;
        while (pos > 0) {
            double pd = ld[(pos - 1) * rows + r];
            unsigned pi = (unsigned)li[(pos - 1) * rows + r];
            if (!(dist < pd || (dist == pd && (unsigned)j < pi))) break;
            ld[pos * rows + r] = pd;
            li[pos * rows + r] = (int)pi;
            --pos;
        }
        ld[pos * rows + r] = dist;
        li[pos * rows + r] = j;
        thr_d = ld[(k - 1) * rows + r];
        thr_i = (unsigned)li[(k - 1) * rows + r];
    }
};

// The same selection for exact integer distances, built for the tensor-core epilogue where a whole warp pays whenever
// one of its rows inserts.  The pair (d2, index) is ONE sortable 64-bit key (d2 < 2^32 for int8 codes and p <= 133000);
// the list is kept UNSORTED with its maximum tracked in registers: an insertion overwrites the maximum and rescans the
// k keys with independent loads (~150 cycles) instead of walking a dependent insertion-sort chain through shared
// memory (~1300 cycles).  Unfilled slots hold the all-ones sentinel, so the warm-up needs no special case.  The list is
// sorted once, when the CTA is done.
struct RowSelectorU64 {
    unsigned long long* lk;      // [k][rows]
    int rows, r, k, maxpos;
    unsigned long long thr;      // current maximum key: a candidate enters iff key < thr

    static constexpr unsigned long long EMPTY = ~0ULL;

    __device__ __forceinline__ void init(void* list, int rows_, int r_, int k_) {
        lk = reinterpret_cast<unsigned long long*>(list); rows = rows_; r = r_; k = k_;
        for (int t = 0; t < k; ++t) lk[t * rows + r] = EMPTY;
        thr = EMPTY;
        maxpos = 0;
    }
    __device__ __forceinline__ static unsigned long long make_key(long long d2, int j) {
        return ((unsigned long long)d2 << 32) | (unsigned)j;
    }
    __device__ __forceinline__ void offer(unsigned long long key) {
        if (key >= thr) return;
        lk[maxpos * rows + r] = key;
        unsigned long long m = 0;
        int mp = 0;
#pragma unroll 8
        for (int t = 0; t < k; ++t) {
            const unsigned long long v = lk[t * rows + r];
            if (v >= m) { m = v; mp = t; }
        }
        thr = m;
        maxpos = mp;
    }
    // ascending order in place: repeatedly move the maximum of the unsorted prefix to its end
    __device__ __forceinline__ void sort() {
        for (int end = k - 1; end > 0; --end) {
            unsigned long long m = 0;
            int mp = 0;
            for (int t = 0; t <= end; ++t) {
                const unsigned long long v = lk[t * rows + r];
                if (v >= m) { m = v; mp = t; }
            }
            const unsigned long long last = lk[end * rows + r];
            lk[end * rows + r] = m;
            lk[mp * rows + r] = last;
        }
    }
    __device__ __forceinline__ double d(int t) const {
        const unsigned long long v = lk[t * rows + r];
        return v == EMPTY ? (double)INFINITY : (double)(v >> 32);
    }
    __device__ __forceinline__ int i(int t) const {
        const unsigned long long v = lk[t * rows + r];
        return v == EMPTY ? 0x7fffffff : (int)(unsigned)(v & 0xffffffffULL);
    }
};

}  // namespace gnnb
