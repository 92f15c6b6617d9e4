// K1 on tcgen05 tensor cores -- placeholder until the TMA/TMEM kernel lands.
#include "common.cuh"
namespace gnnb {
int knn_tc_launch(const int8_t*, int64_t, const int64_t*, int64_t, int64_t, const int8_t*, int64_t, const int64_t*, int64_t, int64_t,
                  int64_t, const double*, const double*, int, int, int, double*, int32_t*, cudaStream_t) {
    return fail(GNNB_E_UNSUPPORTED, "knn_i8: the tensor-core path is not built into this library");
}
bool knn_tc_supported(int64_t, int64_t, const void*, const void*, int) { return false; }
int knn_tc_splits(int64_t, int64_t) { return 1; }
}  // namespace gnnb
