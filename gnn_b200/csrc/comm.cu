// K9: gradient averaging across the GPUs of one box through the NVSwitch (NVLS): every rank owns a slice of the
// buffer, pulls the SUM of that slice over all ranks with one multimem.ld_reduce per 16 bytes (the switch reads the
// replicas and adds them in flight), scales it to the mean and pushes it to every rank with one multimem.st.
// Per GPU and direction the links carry the buffer once (20 MB for the C5 gradients: ~25 us at NVLink 5 rates),
// and because each element is reduced by exactly one rank and then broadcast, all replicas receive the same bits.
//
// The reference has no distributed code.  The buffer must be one symmetric allocation mapped
// through a multicast object on all ranks (torch.distributed._symmetric_memory gives both: `multicast_ptr`); the
// caller brackets the launch with cross-GPU barriers (everyone's contribution written before, every slice delivered
// after) -- gnn_b200/parallel.py uses the symmetric-memory handle's barrier for that.
#include "common.cuh"

namespace gnnb {

__device__ __forceinline__ float4 multimem_ld_reduce_add(const float4* mc_addr) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(mc_addr)
                 : "memory");
    return v;
}
__device__ __forceinline__ void multimem_st(float4* mc_addr, float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                 ::"l"(mc_addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}

// n4 = number of float4 elements of the whole buffer; this rank handles [lo, hi)
__global__ void __launch_bounds__(512)
allreduce_mean_multimem_kernel(float4* __restrict__ mc, int64_t lo, int64_t hi, float inv_world) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    int64_t i = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    // four independent 16-byte reductions in flight per thread: the round trip through the switch is long
    for (; i + 3 * stride < hi; i += 4 * stride) {
        float4 a = multimem_ld_reduce_add(mc + i), b = multimem_ld_reduce_add(mc + i + stride);
        float4 c = multimem_ld_reduce_add(mc + i + 2 * stride), d = multimem_ld_reduce_add(mc + i + 3 * stride);
        a.x *= inv_world; a.y *= inv_world; a.z *= inv_world; a.w *= inv_world;
        b.x *= inv_world; b.y *= inv_world; b.z *= inv_world; b.w *= inv_world;
        c.x *= inv_world; c.y *= inv_world; c.z *= inv_world; c.w *= inv_world;
        d.x *= inv_world; d.y *= inv_world; d.z *= inv_world; d.w *= inv_world;
        multimem_st(mc + i, a);
        multimem_st(mc + i + stride, b);
        multimem_st(mc + i + 2 * stride, c);
        multimem_st(mc + i + 3 * stride, d);
    }
    for (; i < hi; i += stride) {
        float4 a = multimem_ld_reduce_add(mc + i);
        a.x *= inv_world; a.y *= inv_world; a.z *= inv_world; a.w *= inv_world;
        multimem_st(mc + i, a);
    }
}

// The same averaging over plain peer memory (NVLink loads and stores, no multicast object): this rank reads element i of
// its slice from every replica, adds the values in rank order, scales, and writes the mean into every replica.  Slices
// are disjoint, so one pass does the reduce-scatter and the all-gather; per GPU and direction the links carry
// (world - 1) / world of the buffer -- a third of what the multimem form moves at 2 GPUs, about the same at 8.
constexpr int P2P_MAX_WORLD = 16;
struct PeerPtrs {
    float4* p[P2P_MAX_WORLD];
};

template <int WORLD>
__global__ void __launch_bounds__(512)
allreduce_mean_p2p_kernel(PeerPtrs peers, int64_t lo, int64_t hi, float inv_world) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += 2 * stride) {
        const int64_t j = i + stride;
        const bool two = j < hi;
        float4 a[WORLD], b[WORLD];
#pragma unroll
        for (int r = 0; r < WORLD; ++r) {                   // all loads of both elements are issued before the first add
            a[r] = __ldcg(peers.p[r] + i);
            if (two) b[r] = __ldcg(peers.p[r] + j);
        }
        float4 s = a[0], t = two ? b[0] : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int r = 1; r < WORLD; ++r) {                   // fixed order: every run and every rank's slice adds alike
            s.x += a[r].x; s.y += a[r].y; s.z += a[r].z; s.w += a[r].w;
            if (two) { t.x += b[r].x; t.y += b[r].y; t.z += b[r].z; t.w += b[r].w; }
        }
        s.x *= inv_world; s.y *= inv_world; s.z *= inv_world; s.w *= inv_world;
        t.x *= inv_world; t.y *= inv_world; t.z *= inv_world; t.w *= inv_world;
#pragma unroll
        for (int r = 0; r < WORLD; ++r) {
            __stcg(peers.p[r] + i, s);
            if (two) __stcg(peers.p[r] + j, t);
        }
    }
}

}  // namespace gnnb

using namespace gnnb;

extern "C" {

int gnnb_allreduce_mean_multimem(void* multicast_ptr, int64_t n_floats, int rank, int world, int max_ctas, void* stream) {
    GNNB_ENTER();
    GNNB_REQUIRE(multicast_ptr && n_floats >= 0 && world >= 1 && rank >= 0 && rank < world, GNNB_E_BADARG, "allreduce_mean_multimem: bad argument");
    GNNB_REQUIRE(n_floats % 4 == 0 && (uintptr_t)multicast_ptr % 16 == 0, GNNB_E_BADARG,
                 "allreduce_mean_multimem: the buffer must be 16-byte aligned and a multiple of 4 floats long");
    const int64_t n4 = n_floats / 4;
    const int64_t per = (n4 + world - 1) / world;
    const int64_t lo = per * rank < n4 ? per * rank : n4, hi = lo + per < n4 ? lo + per : n4;
    if (hi <= lo) return GNNB_OK;
    int64_t blocks = (hi - lo + 512 * 4 - 1) / (512 * 4);
    int64_t cap = max_ctas > 0 ? max_ctas : (int64_t)sm_count();
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    allreduce_mean_multimem_kernel<<<(unsigned)blocks, 512, 0, (cudaStream_t)stream>>>(static_cast<float4*>(multicast_ptr), lo, hi,
                                                                                       1.0f / (float)world);
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

int gnnb_allreduce_mean_p2p(const uint64_t* peer_ptrs_host, int64_t n_floats, int rank, int world, int max_ctas, void* stream) {
    GNNB_ENTER();
    GNNB_REQUIRE(peer_ptrs_host && n_floats >= 0 && world >= 1 && world <= P2P_MAX_WORLD && rank >= 0 && rank < world, GNNB_E_BADARG,
                 "allreduce_mean_p2p: bad argument (at most %d ranks)", P2P_MAX_WORLD);
    GNNB_REQUIRE(n_floats % 4 == 0, GNNB_E_BADARG, "allreduce_mean_p2p: the buffer must be a multiple of 4 floats long");
    PeerPtrs peers = {};
    for (int r = 0; r < world; ++r) {
        GNNB_REQUIRE(peer_ptrs_host[r] != 0 && peer_ptrs_host[r] % 16 == 0, GNNB_E_BADARG, "allreduce_mean_p2p: replica %d is null or unaligned", r);
        peers.p[r] = reinterpret_cast<float4*>(peer_ptrs_host[r]);
    }
    const int64_t n4 = n_floats / 4;
    const int64_t per = (n4 + world - 1) / world;
    const int64_t lo = per * rank < n4 ? per * rank : n4, hi = lo + per < n4 ? lo + per : n4;
    if (hi <= lo) return GNNB_OK;
    int64_t blocks = (hi - lo + 512 * 2 - 1) / (512 * 2);
    int64_t cap = max_ctas > 0 ? max_ctas : (int64_t)sm_count();
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    const float inv = 1.0f / (float)world;
    cudaStream_t st = (cudaStream_t)stream;
    switch (world) {
        case 1: allreduce_mean_p2p_kernel<1><<<(unsigned)blocks, 512, 0, st>>>(peers, lo, hi, inv); break;
        case 2: allreduce_mean_p2p_kernel<2><<<(unsigned)blocks, 512, 0, st>>>(peers, lo, hi, inv); break;
        case 4: allreduce_mean_p2p_kernel<4><<<(unsigned)blocks, 512, 0, st>>>(peers, lo, hi, inv); break;
        case 8: allreduce_mean_p2p_kernel<8><<<(unsigned)blocks, 512, 0, st>>>(peers, lo, hi, inv); break;
        default: return fail(GNNB_E_UNSUPPORTED, "allreduce_mean_p2p: %d ranks (built for 1, 2, 4 and 8)", world);
    }
    GNNB_LAUNCH_CHECK();
    return GNNB_OK;
}

}  // extern "C"
