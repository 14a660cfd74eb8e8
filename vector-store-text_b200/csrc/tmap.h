// tmap.h — host helper: TMA tensor maps over row-major f32 matrices (cuTensorMapEncodeTiled through the
// runtime's driver entry point, so libvsann.so does not link libcuda).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

namespace vsann {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
inline EncodeTiledFn tmap_encode_fn() {
    static EncodeTiledFn fn = [] {
        void* p = nullptr; cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess || qres != cudaDriverEntryPointSuccess) p = nullptr;
        return (EncodeTiledFn)p;
    }();
    return fn;
}

// rows x cols f32 matrix with `row_stride_floats` between rows; box = box_rows x box_cols; zero fill outside
inline bool make_tmap_f32(CUtensorMap* m, const float* base, uint64_t rows, uint32_t cols, size_t row_stride_floats,
                          uint32_t box_cols, uint32_t box_rows, CUtensorMapSwizzle swizzle) {
    EncodeTiledFn fn = tmap_encode_fn();
    if (!fn) return false;
    const cuuint64_t dims[2] = {cols, rows};
    const cuuint64_t strides[1] = {(cuuint64_t)row_stride_floats * 4};
    const cuuint32_t box[2] = {box_cols, box_rows};
    const cuuint32_t estr[2] = {1, 1};
    return fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
              swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

}  // namespace vsann
