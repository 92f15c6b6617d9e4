// K5 on tcgen05 tensor cores -- placeholder until the TMA/TMEM kernels land.
#include "common.cuh"
namespace gnnb {
int linear_tc_fwd(const void*, int, int64_t, int64_t, int64_t, const float*, int64_t, int64_t, float*, int64_t, int, void*, size_t,
                  cudaStream_t) {
    return fail(GNNB_E_UNSUPPORTED, "linear_fwd: tensor-core precision modes are not built into this library");
}
int linear_tc_bwd_w(const void*, int, int64_t, int64_t, int64_t, const float*, int64_t, int64_t, float*, int64_t, int, void*, size_t,
                    cudaStream_t) {
    return fail(GNNB_E_UNSUPPORTED, "linear_bwd_w: tensor-core precision modes are not built into this library");
}
size_t linear_tc_workspace(int64_t, int64_t, int64_t, int) { return 0; }
}  // namespace gnnb
