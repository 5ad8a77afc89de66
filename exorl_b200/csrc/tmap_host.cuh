// Host side of TMA: a 2-D tiled tensor map with the 128-byte swizzle (cuTensorMapEncodeTiled
// through the runtime's driver entry point — the library does not link libcuda).
#pragma once
#include <cuda.h>
#include "common.cuh"

namespace exb {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static inline EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) ==
            cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// [rows, cols] row-major elements of `dtype` (elem_bytes each, row pitch pitch_bytes, a multiple
// of 16) -> boxes of box_rows x box_cols (box_cols * elem_bytes == 128: one swizzle row), zero
// fill outside the tensor.
static inline int tmap_encode_2d(CUtensorMap* map, CUtensorMapDataType dtype, int elem_bytes,
                                 const void* ptr, int64_t rows, int64_t cols, int64_t pitch_bytes,
                                 int box_cols, int box_rows) {
  EncodeTiledFn enc = encode_fn();
  if (!enc) {
    set_error("cuTensorMapEncodeTiled entry point unavailable");
    return EXB_ECUDA;
  }
  (void)elem_bytes;
  const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)pitch_bytes};
  const cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = enc(map, dtype, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                         CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    return EXB_ECUDA;
  }
  return EXB_OK;
}

}  // namespace exb
