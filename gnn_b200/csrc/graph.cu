// K2/K3: kNN lists -> similarity -> symmetrise -> Laplacian edge list (gnn/loaders/load.py:61-80)
// K4   : arbitrary edge list -> GCN-normalised CSR in both directions (PyG 2.1.0 gcn_norm, restated in
//        SURVEY.md A.1; call sites gnn/networks/networks.py:271-272 ...)
//
// Everything here is deterministic: atomics are only used for counters and for claiming slots inside a
// row segment, and every segment is sorted afterwards by a unique key, so the result does not depend
// on the order in which threads ran.
#include <math.h>

#include "common.cuh"

namespace gnnb {

__device__ __forceinline__ unsigned long long d2u(double x) { return (unsigned long long)__double_as_longlong(x); }
__device__ __forceinline__ double u2d(unsigned long long x) { return __longlong_as_double((long long)x); }

// ===================================================================================== Laplacian graph
struct GraphWs {
    int32_t *indeg, *outdeg, *cap, *seg_off, *cursor, *ucnt, *rowptr, *block_sums;
    unsigned long long* dmax_bits;   // max squared distance over kept edges (float64 bits; non-negative => monotone)
    unsigned long long* lmax_bits;   // max weighted degree
    int64_t* totals;                 // [0] = sum(cap), [1] = nnz(W)
    double* deg;
    Rec* rec;
};

static size_t graph_ws_layout(void* ws, size_t bytes, int64_t n, int k, GraphWs* out) {
    Carver cv(ws, bytes);
    GraphWs g;
    g.indeg = cv.take<int32_t>(n);
    g.outdeg = cv.take<int32_t>(n);
    g.cap = cv.take<int32_t>(n);
    g.seg_off = cv.take<int32_t>(n + 1);
    g.cursor = cv.take<int32_t>(n);
    g.ucnt = cv.take<int32_t>(n);
    g.rowptr = cv.take<int32_t>(n + 1);
    g.block_sums = cv.take<int32_t>(scan_blocks(n));
    g.dmax_bits = cv.take<unsigned long long>(1);
    g.lmax_bits = cv.take<unsigned long long>(1);
    g.totals = cv.take<int64_t>(2);
    g.deg = cv.take<double>(n);
    g.rec = cv.take<Rec>((size_t)2 * n * k);
    if (out) *out = g;
    return cv.off;
}

__global__ void graph_stats_kernel(const int32_t* __restrict__ idx, const double* __restrict__ d2, int64_t n, int k, int mode,
                                   int32_t* indeg, int32_t* outdeg, unsigned long long* dmax_bits) {
    int64_t total = n * k;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        int j = idx[e];
        if (j < 0) continue;
        double d = d2[e];
        if (mode == 1 && !(d > 0.0)) continue;             // zero distances vanish at .nonzero() (load.py:64)
        atomicAdd(&indeg[j], 1);
        atomicAdd(&outdeg[e / k], 1);
        if (mode == 1) atomicMax(dmax_bits, d2u(d));
    }
}

__global__ void add_i32_kernel(const int32_t* a, const int32_t* b, int32_t* out, int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        out[i] = a[i] + b[i];
}

__global__ void graph_scatter_kernel(const int32_t* __restrict__ idx, const double* __restrict__ d2, int64_t n, int k, int mode,
                                     const int32_t* __restrict__ seg_off, int32_t* cursor,
                                     const unsigned long long* __restrict__ dmax_bits, Rec* __restrict__ rec) {
    int64_t total = n * k;
    double dmax = (mode == 1) ? sqrt(u2d(*dmax_bits)) : 1.0;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        int j = idx[e];
        if (j < 0) continue;
        double dd = d2[e];
        if (mode == 1 && !(dd > 0.0)) continue;
        int i = (int)(e / k);
        double dist = (mode == 1) ? sqrt(dd) : 1.0;
        double nn = dist / dmax;                           // load.py:65
        double s = exp(-(nn * nn) / 2.0);                  // load.py:67 with sigma = 1
        int pos = seg_off[i] + atomicAdd(&cursor[i], 1);
        rec[pos].key = ((unsigned long long)(unsigned)j << 1);            // S[i, j]   (out-edge of row i)
        rec[pos].payload = d2u(s);
        int pos2 = seg_off[j] + atomicAdd(&cursor[j], 1);
        rec[pos2].key = ((unsigned long long)(unsigned)i << 1) | 1ULL;    // S^T[j, i] (in-edge of row j)
        rec[pos2].payload = d2u(s);
    }
}

__global__ void graph_unique_kernel(const Rec* __restrict__ rec, const int32_t* __restrict__ seg_off,
                                    const int32_t* __restrict__ cap, int64_t n, int32_t* __restrict__ ucnt) {
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x) {
        const Rec* seg = rec + seg_off[r];
        int len = cap[r], cnt = 0;
        unsigned long long prev = ~0ULL;
        for (int i = 0; i < len; ++i) {
            unsigned long long c = seg[i].key >> 1;
            cnt += (c != prev) && (c != (unsigned long long)r);      // a self entry (include_self) is not an edge of L
            prev = c;
        }
        ucnt[r] = cnt;
    }
}

__global__ void graph_total_kernel(const int64_t* totals, int64_t n, int64_t* n_edges_out) { *n_edges_out = totals[1] + n; }

// off-diagonal block: W = (S + S^T) / 2 in (row, col) order, L = -W, weighted degree per row
__global__ void graph_emit_offdiag_kernel(const Rec* __restrict__ rec, const int32_t* __restrict__ seg_off,
                                          const int32_t* __restrict__ cap, const int32_t* __restrict__ rowptr, int64_t n,
                                          int64_t n_edges, int64_t* __restrict__ edge_index, double* __restrict__ edge_weight,
                                          double* __restrict__ deg, unsigned long long* lmax_bits) {
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x) {
        const Rec* seg = rec + seg_off[r];
        int len = cap[r];
        int64_t e = rowptr[r];
        double dsum = 0.0, diag = 0.0;
        int i = 0;
        while (i < len) {
            unsigned long long c = seg[i].key >> 1;
            double v = u2d(seg[i].payload);
            if (i + 1 < len && (seg[i + 1].key >> 1) == c) {
                v = __dadd_rn(v, u2d(seg[i + 1].payload));      // S[r,c] + S^T[r,c]
                ++i;
            }
            double w = 0.5 * v;                                  // load.py:68
            if (c == (unsigned long long)r) {                    // include_self: W[r,r] is summed, then removed again
                diag = w;                                        // (csgraph.laplacian: m.sum(axis) - m.diagonal())
                dsum = __dadd_rn(dsum, w);
                ++i;
                continue;
            }
            edge_index[e] = r;
            edge_index[n_edges + e] = (int64_t)c;
            edge_weight[e] = -w;                                 // csgraph.laplacian: L = D - W (load.py:69)
            dsum = __dadd_rn(dsum, w);
            ++e;
            ++i;
        }
        dsum = __dsub_rn(dsum, diag);
        deg[r] = dsum;
        atomicMax(lmax_bits, d2u(dsum));
    }
}

__global__ void graph_emit_diag_kernel(int64_t n, int64_t nnz, int64_t n_edges, int use_weights,
                                       const double* __restrict__ deg, const unsigned long long* __restrict__ lmax_bits,
                                       int64_t* __restrict__ edge_index, double* __restrict__ edge_weight) {
    double mx = (double)(float)u2d(*lmax_bits);                  // max(edge_weight).float() (load.py:75)
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n_edges; e += (int64_t)gridDim.x * blockDim.x) {
        double l;
        if (e >= nnz) {
            int64_t r = e - nnz;
            edge_index[e] = r;
            edge_index[n_edges + e] = r;
            l = deg[r];
        } else {
            l = edge_weight[e];
        }
        edge_weight[e] = use_weights ? __dsub_rn(1.0, __ddiv_rn(fabs(l), mx)) : l;
    }
}

// directed kNN graph as an edge list (smoothing != "laplacian", load.py:76-80)
__global__ void knn_edges_max_kernel(const double* __restrict__ d2, int64_t total, int mode, unsigned long long* mx_bits) {
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x)
        atomicMax(mx_bits, d2u(mode == 1 ? d2[e] : 1.0));
}
__global__ void knn_edges_emit_kernel(const int32_t* __restrict__ idx, const double* __restrict__ d2, int64_t n, int k, int mode,
                                      int use_weights, const unsigned long long* __restrict__ mx_bits,
                                      int64_t* __restrict__ edge_index, double* __restrict__ edge_weight) {
    int64_t total = n * k;
    double mx = (double)(float)sqrt(u2d(*mx_bits));
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        edge_index[e] = e / k;
        edge_index[total + e] = idx[e];
        double v = (mode == 1) ? sqrt(d2[e]) : 1.0;
        edge_weight[e] = use_weights ? __dsub_rn(1.0, __ddiv_rn(fabs(v), mx)) : v;
    }
}

// ===================================================================================== COO -> GCN CSR
struct CsrWs {
    int32_t *cnt_in, *cnt_out, *tot_in, *tot_out, *cur_in, *cur_out, *loop_eid, *block_sums;
    int64_t* totals;
    float* dis;
    Rec *rec_in, *rec_out;
};

static size_t csr_ws_layout(void* ws, size_t bytes, int64_t n, int64_t n_edges, CsrWs* out) {
    Carver cv(ws, bytes);
    CsrWs c;
    c.cnt_in = cv.take<int32_t>(n);
    c.cnt_out = cv.take<int32_t>(n);
    c.tot_in = cv.take<int32_t>(n);
    c.tot_out = cv.take<int32_t>(n);
    c.cur_in = cv.take<int32_t>(n);
    c.cur_out = cv.take<int32_t>(n);
    c.loop_eid = cv.take<int32_t>(n);
    c.block_sums = cv.take<int32_t>(scan_blocks(n));
    c.totals = cv.take<int64_t>(2);
    c.dis = cv.take<float>(n);
    // segments are laid out by the final rowptr, which reserves one slot per row for the self loop
    c.rec_in = cv.take<Rec>((size_t)n_edges + (size_t)n + 1);
    c.rec_out = cv.take<Rec>((size_t)n_edges + (size_t)n + 1);
    if (out) *out = c;
    return cv.off;
}

__global__ void csr_count_nonloop_kernel(const int64_t* __restrict__ ei, int64_t E, unsigned long long* out) {
    unsigned long long local = 0;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < E; e += (int64_t)gridDim.x * blockDim.x)
        local += (ei[e] != ei[E + e]);
    for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
    if ((threadIdx.x & 31) == 0 && local) atomicAdd(out, local);
}
__global__ void csr_nnz_kernel(const unsigned long long* nonloop, int64_t E, int64_t n, int asl, int64_t* nnz_out) {
    *nnz_out = asl ? (int64_t)(*nonloop) + n : E;
}

__global__ void csr_count_kernel(const int64_t* __restrict__ ei, int64_t E, int64_t n, int asl,
                                 int32_t* cnt_in, int32_t* cnt_out, int32_t* loop_eid) {
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < E; e += (int64_t)gridDim.x * blockDim.x) {
        int64_t s = ei[e], d = ei[E + e];
        if (s < 0 || s >= n || d < 0 || d >= n) continue;
        if (asl && s == d) {
            atomicMax(&loop_eid[s], (int)e);      // the last self loop in edge order wins, as on the CPU
        } else {
            atomicAdd(&cnt_in[d], 1);
            atomicAdd(&cnt_out[s], 1);
        }
    }
}

__global__ void csr_tot_kernel(const int32_t* cnt_in, const int32_t* cnt_out, int asl, int32_t* tot_in, int32_t* tot_out, int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        tot_in[i] = cnt_in[i] + (asl ? 1 : 0);
        tot_out[i] = cnt_out[i] + (asl ? 1 : 0);
    }
}

__global__ void csr_fill_kernel(const int64_t* __restrict__ ei, const float* __restrict__ ew, int64_t E, int64_t n, int asl,
                                const int32_t* __restrict__ rowptr, const int32_t* __restrict__ rowptr_t,
                                int32_t* cur_in, int32_t* cur_out, Rec* __restrict__ rec_in, Rec* __restrict__ rec_out) {
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < E; e += (int64_t)gridDim.x * blockDim.x) {
        int64_t s = ei[e], d = ei[E + e];
        if (s < 0 || s >= n || d < 0 || d >= n) continue;
        if (asl && s == d) continue;
        float w = ew ? ew[e] : 1.0f;
        unsigned long long wb = (unsigned long long)__float_as_uint(w) << 32;
        int pi = rowptr[d] + atomicAdd(&cur_in[d], 1);
        rec_in[pi].key = (unsigned long long)e;
        rec_in[pi].payload = wb | (unsigned)s;
        int po = rowptr_t[s] + atomicAdd(&cur_out[s], 1);
        rec_out[po].key = (unsigned long long)e;
        rec_out[po].payload = wb | (unsigned)d;
    }
}

__device__ __forceinline__ float loop_weight(const float* ew, const int32_t* loop_eid, int64_t i) {
    int e = loop_eid[i];
    return (e >= 0 && ew) ? ew[e] : 1.0f;          // add_remaining_self_loops: keep an existing loop's weight, else 1
}

// deg_i = sum of incoming weights in edge order, self loop last; dis = deg^-1/2 with inf -> 0
__global__ void csr_deg_kernel(const Rec* __restrict__ rec_in, const int32_t* __restrict__ rowptr, const int32_t* __restrict__ cnt_in,
                               const float* __restrict__ ew, const int32_t* __restrict__ loop_eid, int asl, int64_t n,
                               float* __restrict__ dis) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const Rec* seg = rec_in + rowptr[i];
        int len = cnt_in[i];
        float deg = 0.f;
        for (int q = 0; q < len; ++q) deg = __fadd_rn(deg, __uint_as_float((unsigned)(seg[q].payload >> 32)));
        if (asl) deg = __fadd_rn(deg, loop_weight(ew, loop_eid, i));
        float d = __frcp_rn(__fsqrt_rn(deg));
        dis[i] = isinf(d) ? 0.f : d;
    }
}

// val = dis[src] * w * dis[dst] (left to right, as PyG writes it); self loop appended last
__global__ void csr_norm_kernel(const Rec* __restrict__ rec, const int32_t* __restrict__ rowptr, const int32_t* __restrict__ cnt,
                                const float* __restrict__ ew, const int32_t* __restrict__ loop_eid, int asl, int64_t n,
                                const float* __restrict__ dis, int transposed, int32_t* __restrict__ col, float* __restrict__ val) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        int base = rowptr[i];
        const Rec* seg = rec + base;
        int len = cnt[i];
        float di = dis[i];
        for (int q = 0; q < len; ++q) {
            unsigned long long pl = seg[q].payload;
            int other = (int)(unsigned)(pl & 0xffffffffULL);
            float w = __uint_as_float((unsigned)(pl >> 32));
            float dsrc = transposed ? di : dis[other];
            float ddst = transposed ? dis[other] : di;
            col[base + q] = other;
            val[base + q] = __fmul_rn(__fmul_rn(dsrc, w), ddst);
        }
        if (asl) {
            col[base + len] = (int)i;
            val[base + len] = __fmul_rn(__fmul_rn(di, loop_weight(ew, loop_eid, i)), di);
        }
    }
}

static unsigned grid_for(int64_t work, int threads = 256, int per_sm = 16) {
    int64_t blocks = (work + threads - 1) / threads;
    int64_t cap = (int64_t)sm_count() * per_sm;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (unsigned)blocks;
}

}  // namespace gnnb

using namespace gnnb;

extern "C" {

int gnnb_graph_workspace_bytes(int64_t n, int k, size_t* bytes) {
    GNNB_REQUIRE(bytes && n > 0 && k > 0, GNNB_E_BADARG, "graph_workspace_bytes: bad argument");
    *bytes = graph_ws_layout(nullptr, 0, n, k, nullptr) + 256;
    return GNNB_OK;
}

int gnnb_graph_count(const int32_t* idx, const double* d2, int64_t n, int k, int mode,
                     int64_t* n_edges_out, void* workspace, size_t workspace_bytes, void* stream_) {
    GNNB_ENTER();
    cudaStream_t stream = (cudaStream_t)stream_;
    GNNB_REQUIRE(idx && d2 && n_edges_out && n > 0 && k > 0 && (mode == 0 || mode == 1), GNNB_E_BADARG, "graph_count: bad argument");
    GNNB_REQUIRE((double)n * k * 2 + n < 2.0e9, GNNB_E_UNSUPPORTED, "graph_count: more than 2^31 edges");
    GraphWs g;
    size_t need = graph_ws_layout(workspace, workspace_bytes, n, k, &g);
    GNNB_REQUIRE(workspace && need <= workspace_bytes, GNNB_E_WORKSPACE, "graph_count: workspace %zu < %zu bytes", workspace_bytes, need);
    GNNB_CUDA(cudaMemsetAsync(g.indeg, 0, sizeof(int32_t) * n, stream));
    GNNB_CUDA(cudaMemsetAsync(g.outdeg, 0, sizeof(int32_t) * n, stream));
    GNNB_CUDA(cudaMemsetAsync(g.cursor, 0, sizeof(int32_t) * n, stream));
    GNNB_CUDA(cudaMemsetAsync(g.dmax_bits, 0, sizeof(unsigned long long), stream));
    GNNB_CUDA(cudaMemsetAsync(g.lmax_bits, 0, sizeof(unsigned long long), stream));
    graph_stats_kernel<<<grid_for(n * k), 256, 0, stream>>>(idx, d2, n, k, mode, g.indeg, g.outdeg, g.dmax_bits);
    GNNB_LAUNCH_CHECK();
    add_i32_kernel<<<grid_for(n), 256, 0, stream>>>(g.indeg, g.outdeg, g.cap, n);
    GNNB_LAUNCH_CHECK();
    int rc = scan_exclusive_i32(g.cap, g.seg_off, n, g.block_sums, &g.totals[0], stream);
    if (rc) return rc;
    graph_scatter_kernel<<<grid_for(n * k), 256, 0, stream>>>(idx, d2, n, k, mode, g.seg_off, g.cursor, g.dmax_bits, g.rec);
    GNNB_LAUNCH_CHECK();
    rc = segment_sort(g.rec, g.seg_off, g.cap, n, stream);
    if (rc) return rc;
    graph_unique_kernel<<<grid_for(n), 256, 0, stream>>>(g.rec, g.seg_off, g.cap, n, g.ucnt);
    GNNB_LAUNCH_CHECK();
    rc = scan_exclusive_i32(g.ucnt, g.rowptr, n, g.block_sums, &g.totals[1], stream);
    if (rc) return rc;
    graph_total_kernel<<<1, 1, 0, stream>>>(g.totals, n, n_edges_out);
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

int gnnb_graph_fill(int64_t n, int k, int use_weights, int64_t n_edges,
                    int64_t* edge_index, double* edge_weight,
                    void* workspace, size_t workspace_bytes, void* stream_) {
    GNNB_ENTER();
    cudaStream_t stream = (cudaStream_t)stream_;
    GNNB_REQUIRE(edge_index && edge_weight && n > 0 && k > 0 && n_edges >= n, GNNB_E_BADARG, "graph_fill: bad argument");
    GraphWs g;
    size_t need = graph_ws_layout(workspace, workspace_bytes, n, k, &g);
    GNNB_REQUIRE(workspace && need <= workspace_bytes, GNNB_E_WORKSPACE, "graph_fill: workspace %zu < %zu bytes", workspace_bytes, need);
    graph_emit_offdiag_kernel<<<grid_for(n, 128), 128, 0, stream>>>(g.rec, g.seg_off, g.cap, g.rowptr, n, n_edges, edge_index,
                                                                    edge_weight, g.deg, g.lmax_bits);
    GNNB_LAUNCH_CHECK();
    graph_emit_diag_kernel<<<grid_for(n_edges), 256, 0, stream>>>(n, n_edges - n, n_edges, use_weights, g.deg, g.lmax_bits,
                                                                  edge_index, edge_weight);
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

int gnnb_graph_knn_edges(const int32_t* idx, const double* d2, int64_t n, int k, int mode, int use_weights,
                         int64_t* edge_index, double* edge_weight, void* workspace, size_t workspace_bytes,
                         void* stream_) {
    GNNB_ENTER();
    cudaStream_t stream = (cudaStream_t)stream_;
    GNNB_REQUIRE(idx && d2 && edge_index && edge_weight && n > 0 && k > 0 && (mode == 0 || mode == 1), GNNB_E_BADARG,
                 "graph_knn_edges: bad argument");
    GNNB_REQUIRE(workspace && workspace_bytes >= 256, GNNB_E_WORKSPACE, "graph_knn_edges: workspace too small");
    unsigned long long* mx = static_cast<unsigned long long*>(workspace);
    GNNB_CUDA(cudaMemsetAsync(mx, 0, sizeof(unsigned long long), stream));
    knn_edges_max_kernel<<<grid_for(n * k), 256, 0, stream>>>(d2, n * k, mode, mx);
    GNNB_LAUNCH_CHECK();
    knn_edges_emit_kernel<<<grid_for(n * k), 256, 0, stream>>>(idx, d2, n, k, mode, use_weights, mx, edge_index, edge_weight);
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

int gnnb_csr_workspace_bytes(int64_t n, int64_t n_edges, size_t* bytes) {
    GNNB_REQUIRE(bytes && n > 0 && n_edges >= 0, GNNB_E_BADARG, "csr_workspace_bytes: bad argument");
    *bytes = csr_ws_layout(nullptr, 0, n, n_edges, nullptr) + 256;
    return GNNB_OK;
}

int gnnb_csr_count(const int64_t* edge_index, int64_t n_edges, int64_t n, int add_self_loops,
                   int64_t* nnz_out, void* workspace, size_t workspace_bytes, void* stream_) {
    GNNB_ENTER();
    cudaStream_t stream = (cudaStream_t)stream_;
    GNNB_REQUIRE(nnz_out && n > 0 && n_edges >= 0 && (edge_index || n_edges == 0), GNNB_E_BADARG, "csr_count: bad argument");
    GNNB_REQUIRE(workspace && workspace_bytes >= 256, GNNB_E_WORKSPACE, "csr_count: workspace too small");
    unsigned long long* cnt = static_cast<unsigned long long*>(workspace);
    GNNB_CUDA(cudaMemsetAsync(cnt, 0, sizeof(unsigned long long), stream));
    if (n_edges > 0) {
        csr_count_nonloop_kernel<<<grid_for(n_edges), 256, 0, stream>>>(edge_index, n_edges, cnt);
        GNNB_LAUNCH_CHECK();
    }
    csr_nnz_kernel<<<1, 1, 0, stream>>>(cnt, n_edges, n, add_self_loops, nnz_out);
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

int gnnb_csr_build(const int64_t* edge_index, const float* edge_weight, int64_t n_edges, int64_t n,
                   int add_self_loops, int64_t nnz,
                   int32_t* rowptr, int32_t* col, float* val,
                   int32_t* rowptr_t, int32_t* col_t, float* val_t,
                   void* workspace, size_t workspace_bytes, void* stream_) {
    GNNB_ENTER();
    cudaStream_t stream = (cudaStream_t)stream_;
    GNNB_REQUIRE(rowptr && col && val && rowptr_t && col_t && val_t && n > 0 && n_edges >= 0 && nnz >= 0, GNNB_E_BADARG,
                 "csr_build: bad argument");
    GNNB_REQUIRE(edge_index || n_edges == 0, GNNB_E_BADARG, "csr_build: edge_index is null");
    GNNB_REQUIRE(nnz < 2147483647LL && n_edges < 2147483647LL, GNNB_E_UNSUPPORTED, "csr_build: more than 2^31 edges");
    int asl = add_self_loops ? 1 : 0;
    CsrWs c;
    size_t need = csr_ws_layout(workspace, workspace_bytes, n, n_edges, &c);
    GNNB_REQUIRE(workspace && need <= workspace_bytes, GNNB_E_WORKSPACE, "csr_build: workspace %zu < %zu bytes", workspace_bytes, need);
    GNNB_CUDA(cudaMemsetAsync(c.cnt_in, 0, sizeof(int32_t) * n, stream));
    GNNB_CUDA(cudaMemsetAsync(c.cnt_out, 0, sizeof(int32_t) * n, stream));
    GNNB_CUDA(cudaMemsetAsync(c.cur_in, 0, sizeof(int32_t) * n, stream));
    GNNB_CUDA(cudaMemsetAsync(c.cur_out, 0, sizeof(int32_t) * n, stream));
    GNNB_CUDA(cudaMemsetAsync(c.loop_eid, 0xff, sizeof(int32_t) * n, stream));
    if (n_edges > 0) {
        csr_count_kernel<<<grid_for(n_edges), 256, 0, stream>>>(edge_index, n_edges, n, asl, c.cnt_in, c.cnt_out, c.loop_eid);
        GNNB_LAUNCH_CHECK();
    }
    csr_tot_kernel<<<grid_for(n), 256, 0, stream>>>(c.cnt_in, c.cnt_out, asl, c.tot_in, c.tot_out, n);
    GNNB_LAUNCH_CHECK();
    int rc = scan_exclusive_i32(c.tot_in, rowptr, n, c.block_sums, &c.totals[0], stream);
    if (rc) return rc;
    rc = scan_exclusive_i32(c.tot_out, rowptr_t, n, c.block_sums, &c.totals[1], stream);
    if (rc) return rc;
    if (n_edges > 0) {
        csr_fill_kernel<<<grid_for(n_edges), 256, 0, stream>>>(edge_index, edge_weight, n_edges, n, asl, rowptr, rowptr_t,
                                                               c.cur_in, c.cur_out, c.rec_in, c.rec_out);
        GNNB_LAUNCH_CHECK();
        rc = segment_sort(c.rec_in, rowptr, c.cnt_in, n, stream);
        if (rc) return rc;
        rc = segment_sort(c.rec_out, rowptr_t, c.cnt_out, n, stream);
        if (rc) return rc;
    }
    csr_deg_kernel<<<grid_for(n, 128), 128, 0, stream>>>(c.rec_in, rowptr, c.cnt_in, edge_weight, c.loop_eid, asl, n, c.dis);
    GNNB_LAUNCH_CHECK();
    csr_norm_kernel<<<grid_for(n, 128), 128, 0, stream>>>(c.rec_in, rowptr, c.cnt_in, edge_weight, c.loop_eid, asl, n, c.dis, 0, col, val);
    GNNB_LAUNCH_CHECK();
    csr_norm_kernel<<<grid_for(n, 128), 128, 0, stream>>>(c.rec_out, rowptr_t, c.cnt_out, edge_weight, c.loop_eid, asl, n, c.dis, 1,
                                                          col_t, val_t);
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

}  // extern "C"
