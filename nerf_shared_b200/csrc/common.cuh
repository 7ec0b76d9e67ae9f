// Shared helpers for libnerf_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/nerf_b200.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libnerf_b200 is written for sm_100a only"
#endif

namespace nb {

// ---- error plumbing (api.cu owns the storage) ------------------------------------------
void set_error(const char* fmt, ...);
void count_launch(int n = 1);
int check_launch(const char* what);  // cudaGetLastError -> return code, records message
int limit_mlp_clusters();            // nb_set_launch_limits (0 = whole GPU)
int limit_wgrad_ctas();

#define NB_REQUIRE(cond, ...)              \
  do {                                     \
    if (!(cond)) {                         \
      ::nb::set_error(__VA_ARGS__);        \
      return NB_E_ARG;                     \
    }                                      \
  } while (0)

static inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

constexpr int kNumSMs = 148;

static inline int grid_for(int64_t work_items, int per_block, int max_blocks_per_sm = 16) {
  int64_t b = (work_items + per_block - 1) / per_block;
  int64_t cap = (int64_t)kNumSMs * max_blocks_per_sm;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

// ---- device helpers ----------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

}  // namespace nb
