// Host side of tc_common.cuh: tensor-map encoding through the driver entry point (no link against libcuda,
// so the library still loads on a machine without a driver and fails loudly there).
#include "tc_common.cuh"

namespace gnnb {
namespace tc {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

int make_tensor_map_2d(CUtensorMap* map, const void* base, int elem_bytes, uint64_t cols, uint64_t rows, uint64_t ld_bytes,
                       uint32_t box_cols, uint32_t box_rows, bool swizzle128, bool swizzle64) {
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) return fail(GNNB_E_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
    if ((uintptr_t)base % 16 != 0 || ld_bytes % 16 != 0)
        return fail(GNNB_E_UNSUPPORTED, "TMA needs a 16-byte aligned base and row stride (got stride %llu bytes)",
                    (unsigned long long)ld_bytes);
    CUtensorMapDataType dt = elem_bytes == 1 ? CU_TENSOR_MAP_DATA_TYPE_UINT8
                           : elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16
                                             : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {ld_bytes};
    cuuint32_t box[2] = {box_cols, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(map, dt, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : (swizzle64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_NONE), CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(GNNB_E_CUDA, "cuTensorMapEncodeTiled failed with code %d", (int)r);
    return GNNB_OK;
}

}  // namespace tc
}  // namespace gnnb
