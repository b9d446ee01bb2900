// Host-side tensor-map encoder (driver symbol resolved through the runtime, no -lcuda needed).
#include "tma_utils.cuh"

#include <mutex>

namespace pb2 {

typedef CUresult (*encode_tiled_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static encode_tiled_fn g_encode = nullptr;
static std::once_flag g_encode_once;

static void resolve_encode() {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
        g_encode = (encode_tiled_fn)fn;
}

int encode_tmap_2d(CUtensorMap* out, CUtensorMapDataType dtype, unsigned elem_bytes, const void* base,
                   uint64_t dim0, uint64_t dim1, uint64_t ld_elems, uint32_t box0, uint32_t box1,
                   CUtensorMapSwizzle swizzle) {
    std::call_once(g_encode_once, resolve_encode);
    if (!g_encode) return -1;
    cuuint64_t dims[2] = {dim0, dim1};
    cuuint64_t strides[1] = {ld_elems * elem_bytes};
    cuuint32_t box[2] = {box0, box1};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = g_encode(out, dtype, 2, const_cast<void*>(base), dims, strides, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? 0 : -2;
}

}  // namespace pb2
