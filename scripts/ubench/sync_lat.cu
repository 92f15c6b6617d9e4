// Microbenchmark: latencies of the synchronisation hops used by the tensor-core pipelines (B200).
//   ping-pong between two warps over mbarriers (arrive -> try_wait wake-up), with and without tcgen05.commit
#include <cstdio>
#include "../../gnn_b200/csrc/tc_common.cuh"
using namespace gnnb::tc;
namespace gnnb { int fail(int c, const char*, ...) { return c; } void set_error(const char*, ...) {} int sm_count() { return 148; } int require_sm100() { return 0; } }

__device__ __forceinline__ bool mbar_test_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void spin_wait(uint64_t* bar, uint32_t parity) { while (!mbar_test_wait(bar, parity)) {} }
__device__ __forceinline__ bool try_wait_hint(uint64_t* bar, uint32_t parity, uint32_t ns) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity), "r"(ns) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void hint_wait(uint64_t* bar, uint32_t parity, uint32_t ns) { while (!try_wait_hint(bar, parity, ns)) {} }

__global__ void __launch_bounds__(128, 1) pingpong(int iters, int mode, long long* out) {
    __shared__ uint64_t bar_a, bar_b;
    __shared__ uint32_t tmem_ptr;
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { mbar_init(&bar_a, 1); mbar_init(&bar_b, 1); fence_barrier_init(); }
    if (warp == 2) tmem_alloc<32>(&tmem_ptr);
    tc_fence_before_sync(); __syncthreads(); tc_fence_after_sync();
    long long t0 = clock64();
    if (warp == 0) {
        if (elect_one()) {
            for (int i = 0; i < iters; ++i) {
                mbar_arrive(&bar_a);                       // ping
                if (mode == 3) spin_wait(&bar_b, i & 1);
                else if (mode == 4) hint_wait(&bar_b, i & 1, 20);
                else if (mode == 5) { while (!mbar_try_wait(&bar_b, i & 1)) {} }
                else mbar_wait(&bar_b, i & 1);             // wait pong
            }
            out[0] = clock64() - t0;
        }
    } else if (warp == 1) {
        if (elect_one()) {
            for (int i = 0; i < iters; ++i) {
                if (mode == 3) spin_wait(&bar_a, i & 1);
                else if (mode == 4) hint_wait(&bar_a, i & 1, 20);
                else if (mode == 5) { while (!mbar_try_wait(&bar_a, i & 1)) {} }
                else mbar_wait(&bar_a, i & 1);
                if (mode == 0 || mode >= 3) mbar_arrive(&bar_b);                         // plain arrive
                else if (mode == 1) { tc_fence_after_sync(); tc_commit(&bar_b); }   // tcgen05.commit with an empty tensor pipe
                else if (mode == 2) { fence_proxy_async(); mbar_arrive(&bar_b); }   // proxy fence + arrive
            }
        }
    }
    tc_fence_before_sync(); __syncthreads();
    if (warp == 2) tmem_dealloc<32>(tmem_ptr);
}

int main() {
    long long* d; cudaMalloc(&d, 8);
    const char* names[] = {"arrive <-> try_wait round trip", "arrive -> wait -> tcgen05.commit -> wait", "arrive -> wait -> fence.proxy.async + arrive -> wait",
                           "test_wait spin round trip", "try_wait with 20 ns hint round trip", "bare try_wait loop round trip"};
    for (int mode = 0; mode < 6; ++mode) {
        int iters = 20000;
        pingpong<<<1, 128>>>(100, mode, d); cudaDeviceSynchronize();
        pingpong<<<1, 128>>>(iters, mode, d);
        cudaError_t e = cudaDeviceSynchronize();
        long long c; cudaMemcpy(&c, d, 8, cudaMemcpyDeviceToHost);
        printf("%-55s : %.1f cycles per round trip (%s)\n", names[mode], (double)c / iters, cudaGetErrorString(e));
    }
    return 0;
}
