// Frame output tail: utils.to8b (utils.py:30) on the device, so that a rendered frame
// leaves the GPU as H*W*3 bytes instead of H*W*3 floats.
#include "common.cuh"

namespace nb {

// (255 * clip(x, 0, 1)).astype(uint8): fp32 product, truncation toward zero (NumPy's cast); NaN -> 0
__global__ void __launch_bounds__(256) to8b_kernel(const float* __restrict__ x, int64_t n, uint8_t* __restrict__ out) {
  const int64_t i4 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i4 + 3 < n && ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(out)) & 15) == 0) {
    const float4 v = *reinterpret_cast<const float4*>(x + i4);
    const float f[4] = {v.x, v.y, v.z, v.w};
    uint32_t p = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float c = fminf(fmaxf(f[k], 0.0f), 1.0f);   // fmaxf(NaN, 0) = 0
      p |= (uint32_t)(uint8_t)__fmul_rn(255.0f, c) << (8 * k);
    }
    *reinterpret_cast<uint32_t*>(out + i4) = p;
  } else {
    for (int64_t i = i4; i < n && i < i4 + 4; ++i) out[i] = (uint8_t)__fmul_rn(255.0f, fminf(fmaxf(x[i], 0.0f), 1.0f));
  }
}

}  // namespace nb

using namespace nb;

extern "C" {

int nb_to8b(const float* x, int64_t n, uint8_t* out, void* stream) {
  if (n == 0) return 0;
  NB_REQUIRE(x && out, "nb_to8b: null pointer");
  const int64_t blocks = (n + 1023) / 1024;
  NB_REQUIRE(blocks < (1ll << 31), "nb_to8b: too many elements for one launch");
  to8b_kernel<<<(unsigned)blocks, 256, 0, as_stream(stream)>>>(x, n, out);
  return check_launch("nb_to8b");
}

}  // extern "C"
