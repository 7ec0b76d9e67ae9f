// Host-side TMA tensor-map construction (cuTensorMapEncodeTiled through the runtime's driver
// entry-point lookup, so the library does not link libcuda directly).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

#include "common.cuh"
#include "mlp_common.cuh"

namespace nb {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static inline EncodeTiledFn encode_tiled_fn() {
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

// A blob of `n_slabs` pre-swizzled 16 KB operand slabs viewed as a [n_slabs*128 rows x 64 bf16] matrix;
// one TMA box = 64 rows x 64 columns = 8 KB, copied verbatim (no swizzle: the image is already swizzled).
static inline int make_slab_tmap(CUtensorMap* tm, const void* blob, int n_slabs) {
  EncodeTiledFn enc = encode_tiled_fn();
  if (!enc) {
    set_error("cuTensorMapEncodeTiled is not available from this driver");
    return NB_E_DEVICE;
  }
  const cuuint64_t gdim[2] = {64, (cuuint64_t)n_slabs * 128};
  const cuuint64_t gstride[1] = {128};
  const cuuint32_t box[2] = {64, 64};
  const cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(blob), gdim, gstride, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (CUresult %d)", (int)r);
    return NB_E_ARG;
  }
  return 0;
}

}  // namespace nb
