// Microbenchmark: issue rate of tcgen05.mma on B200 with operands resident in shared memory (no loads in the loop).
//   kind: 0 = i8 (S32 accumulate), 1 = f8f6f4 e4m3 (F32), 2 = f16 bf16 (F32)
//   one CTA per SM, M = 128, N given, K = 32 bytes per instruction, K-major SWIZZLE_128B tiles; `per_commit` MMAs
//   between two tcgen05.commit + waits (= the k-block structure of the kNN / projection kernels)
// Prints cycles per MMA and the implied MAC / clk / SM.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 scripts/ubench/mma_rate.cu -o variants/mma_rate -lcuda
#include <cstdio>
#include <cstdlib>
#include "../../gnn_b200/csrc/tc_common.cuh"
using namespace gnnb::tc;
__device__ __forceinline__ void mma_f8_ss(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}\n"
                 ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
namespace gnnb { int fail(int c, const char*, ...) { return c; } void set_error(const char*, ...) {} int sm_count() { return 148; } int require_sm100() { return 0; } }

__global__ void __launch_bounds__(128, 1) mma_loop(int iters, int kind, int n, int per_commit, int depth, int a_bytes_row_step, long long* out) {
    extern __shared__ uint8_t raw[];
    __shared__ uint64_t bars[8];
    __shared__ uint32_t tmem_ptr;
    uint8_t* base = raw + ((1024 - (smem_u32(raw) & 1023)) & 1023);
    const int warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < (128 + 256) * 128 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(base)[i] = 0x01010101u * (i & 1);
    if (threadIdx.x == 0) { for (int s = 0; s < 8; ++s) mbar_init(&bars[s], 1); fence_barrier_init(); }
    fence_proxy_async();
    if (warp == 2) tmem_alloc<512>(&tmem_ptr);
    tc_fence_before_sync(); __syncthreads(); tc_fence_after_sync();
    const uint32_t tmem = tmem_ptr;
    if (warp == 0 && elect_one()) {
        const uint32_t idesc = kind == 0 ? make_idesc(2, 1, 0, 0, 128, n) : kind == 1 ? make_idesc(1, 0, 0, 0, 128, n) : make_idesc(1, 1, 0, 0, 128, n);
        const uint32_t a_addr = smem_u32(base), b_addr = a_addr + 128 * 128;
        long long t0 = clock64();
        for (int i = 0; i < iters; ++i) {
            const int s = i % depth;
            if (i >= depth) mbar_wait(&bars[s], ((i / depth) - 1) & 1);       // the commit of `depth` k-blocks ago
            for (int j = 0; j < per_commit; ++j) {
                const uint64_t da = make_smem_desc(a_addr + (j & 3) * 32, 16, 1024);
                const uint64_t db = make_smem_desc(b_addr + (j & 3) * 32, 16, 1024);
                if (kind == 0) mma_i8_ss(tmem, da, db, idesc, 1u);
                else if (kind == 1) mma_f8_ss(tmem, da, db, idesc, 1u);
                else mma_bf16_ss(tmem, da, db, idesc, 1u);
            }
            tc_commit(&bars[s]);
        }
        for (int s = 0; s < depth && s < iters; ++s) {
            const int last = ((iters - 1 - s) / depth) * depth + s;             // last k-block that used slot s
            mbar_wait(&bars[s], (last / depth) & 1);
        }
        out[blockIdx.x] = clock64() - t0;
    }
    tc_fence_before_sync(); __syncthreads();
    if (warp == 2) tmem_dealloc<512>(tmem);
}

int main(int argc, char** argv) {
    int grid = argc > 1 ? atoi(argv[1]) : 148;
    long long* d; cudaMalloc(&d, 8 * 1024);
    const size_t smem = 1024 + (128 + 256) * 128;
    cudaFuncSetAttribute(mma_loop, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const int iters = 2000;
    const char* names[3] = {"i8", "f8(e4m3)", "bf16"};
    for (int kind = 0; kind < 3; ++kind)
        for (int n : {64, 128, 256})
            for (int per_commit : {4, 16})
                for (int depth : {1, 4}) {
                    mma_loop<<<grid, 128, smem>>>(iters, kind, n, per_commit, depth, 0, d);
                    cudaError_t e = cudaDeviceSynchronize();
                    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
                    long long h[1024]; cudaMemcpy(h, d, 8 * grid, cudaMemcpyDeviceToHost);
                    long long mx = 0; for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
                    const double cyc = (double)mx / ((double)iters * per_commit);
                    const double kelems = kind == 2 ? 16.0 : 32.0;
                    printf("grid %3d %-9s N=%3d per_commit=%2d depth=%d: %7.1f cycles/MMA  %6.0f MAC/clk/SM\n", grid, names[kind], n, per_commit, depth,
                           cyc, 128.0 * n * kelems / cyc);
                }
    return 0;
}
