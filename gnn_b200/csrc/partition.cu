// Subsampling: the parallel half (first occurrence of every node in the neighbour stream) and the
// host cursor walk of naive_partition / split_dataset_graph (gnn/loaders/util.py:15-45, 48-73).
// The O(n + |S|) formulation and why it reproduces the reference's repeated rescans: SURVEY.md A.4.
#include <vector>

#include "common.cuh"

namespace gnnb {

__global__ void first_occurrence_kernel(const int32_t* __restrict__ S, int64_t len, int64_t n, long long* first_pos) {
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < len; e += (int64_t)gridDim.x * blockDim.x) {
        int v = S[e];
        if (v >= 0 && v < n) atomicMin(&first_pos[v], (long long)e);
    }
}

__global__ void fill_i64_kernel(long long* a, int64_t n, long long v) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) a[i] = v;
}

}  // namespace gnnb

using namespace gnnb;

extern "C" {

int gnnb_first_occurrence(const int32_t* S, int64_t len, int64_t n, int64_t* first_pos, void* stream_) {
    GNNB_ENTER();
    cudaStream_t stream = (cudaStream_t)stream_;
    GNNB_REQUIRE(first_pos && n > 0 && len >= 0 && (S || len == 0), GNNB_E_BADARG, "first_occurrence: bad argument");
    int64_t cap = (int64_t)sm_count() * 16;
    int64_t b1 = (n + 255) / 256;
    fill_i64_kernel<<<(unsigned)(b1 < cap ? b1 : cap), 256, 0, stream>>>((long long*)first_pos, n, 0x7fffffffffffffffLL);
    GNNB_LAUNCH_CHECK();
    if (len > 0) {
        int64_t b2 = (len + 255) / 256;
        first_occurrence_kernel<<<(unsigned)(b2 < cap ? b2 : cap), 256, 0, stream>>>(S, len, n, (long long*)first_pos);
        GNNB_LAUNCH_CHECK();
    }
    return GNNB_OK;
}

// Host-only.  `order` lists the nodes that occur in S by ascending first occurrence.
int64_t gnnb_partition_walk_host(const int64_t* order, int64_t n_order, int64_t n_nodes,
                                 const int64_t* sizes, int64_t n_sizes, uint8_t* taken_io,
                                 int64_t* out, int64_t* out_len) {
    if (!order && n_order > 0) return fail(GNNB_E_BADARG, "partition_walk: order is null");
    if (!sizes || !out || !out_len || n_nodes <= 0 || n_sizes < 0 || n_order < 0)
        return fail(GNNB_E_BADARG, "partition_walk: bad argument");
    std::vector<char> taken((size_t)n_nodes, 0);
    int64_t cur = 0, lo = 0, left = n_nodes, written = 0;
    if (taken_io)
        for (int64_t v = 0; v < n_nodes; ++v)
            if (taken_io[v]) { taken[(size_t)v] = 1; --left; }
    auto take_min = [&]() -> int64_t {
        while (taken[(size_t)lo]) ++lo;
        taken[(size_t)lo] = 1;
        --left;
        return lo;
    };
    for (int64_t b = 0; b < n_sizes; ++b) {
        int64_t size = sizes[b], got = 0;
        if (left == 0) { out_len[b] = 0; continue; }
        // util.py:22-27: every call starts from the smallest node not yet taken
        if (written >= n_nodes) return fail(GNNB_E_BADARG, "partition_walk: output overflow");
        out[written + got++] = take_min();
        if (size > 1 && n_order > 0) {
            int64_t took = 0;
            while (got < size && cur < n_order) {           // util.py:29-38: scan S, accept unseen nodes in first-occurrence order
                int64_t v = order[cur++];
                if (v < 0 || v >= n_nodes) return fail(GNNB_E_BADARG, "partition_walk: node id out of range");
                if (taken[(size_t)v]) continue;
                taken[(size_t)v] = 1;
                --left;
                out[written + got++] = v;
                ++took;
            }
            if (got >= size) {
                if (took == 1 && left > 0) out[written + got++] = take_min();   // util.py:36-43 when the scan stops after one pick
            } else {
                while (got < size && left > 0) out[written + got++] = take_min();   // util.py:39-43: nodes nobody points to
            }
        }
        out_len[b] = got;
        written += got;
    }
    if (taken_io)
        for (int64_t v = 0; v < n_nodes; ++v) taken_io[v] = (uint8_t)taken[(size_t)v];
    return written;
}

}  // extern "C"
