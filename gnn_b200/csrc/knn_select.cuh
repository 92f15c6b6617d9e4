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

}  // namespace gnnb
