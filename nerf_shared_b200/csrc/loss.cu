// Photometric loss of the training step (main.py:87-99): img2mse on the fine and the coarse map, their sum, and
// mse2psnr of the fine term (utils.py:24,27), in one launch; the backward writes both map gradients in one launch.
// SURVEY.md section 8 row a11 / "next" f1.
#include "common.cuh"

namespace nb {

// out[0] = mean((rgb - t)^2), out[1] = mean((rgb0 - t)^2) (0 if rgb0 == NULL), out[2] = out[0] + out[1],
// out[3] = -10 log10(out[0]).  One block: the maps of a training batch are a few thousand floats.
__global__ void __launch_bounds__(1024) mse_pair_fwd_kernel(const float* __restrict__ rgb, const float* __restrict__ rgb0,
                                                            const float* __restrict__ target, int64_t n3,
                                                            float* __restrict__ out) {
  __shared__ float red[2][32];
  float a = 0.f, b = 0.f;
  for (int64_t i = threadIdx.x; i < n3; i += blockDim.x) {
    const float t = target[i];
    const float d = rgb[i] - t;
    a = fmaf(d, d, a);
    if (rgb0) {
      const float d0 = rgb0[i] - t;
      b = fmaf(d0, d0, b);
    }
  }
  a = warp_sum(a);
  b = warp_sum(b);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) { red[0][w] = a; red[1][w] = b; }
  __syncthreads();
  if (w == 0) {
    a = lane < (int)(blockDim.x >> 5) ? red[0][lane] : 0.f;
    b = lane < (int)(blockDim.x >> 5) ? red[1][lane] : 0.f;
    a = warp_sum(a);
    b = warp_sum(b);
    if (lane == 0) {
      const float m = a / (float)n3, m0 = b / (float)n3;
      out[0] = m; out[1] = m0; out[2] = m + m0;
      out[3] = -10.0f * logf(m) / logf(10.0f);
    }
  }
}

// d_rgb = (g_total + g_fine) * 2 (rgb - t) / n3, d_rgb0 = (g_total + g_coarse) * 2 (rgb0 - t) / n3; the g_* are the
// upstream gradients of out[0], out[1], out[2] as device scalars (NULL = zero)
__global__ void __launch_bounds__(256) mse_pair_bwd_kernel(const float* __restrict__ rgb, const float* __restrict__ rgb0,
                                                           const float* __restrict__ target, int64_t n3,
                                                           const float* __restrict__ g_fine,
                                                           const float* __restrict__ g_coarse,
                                                           const float* __restrict__ g_total, float* __restrict__ d_rgb,
                                                           float* __restrict__ d_rgb0) {
  const float s = 2.0f / (float)n3;
  const float gt = g_total ? g_total[0] : 0.f;
  const float gf = ((g_fine ? g_fine[0] : 0.f) + gt) * s, gc = ((g_coarse ? g_coarse[0] : 0.f) + gt) * s;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n3; i += (int64_t)gridDim.x * blockDim.x) {
    const float t = target[i];
    d_rgb[i] = gf * (rgb[i] - t);
    if (d_rgb0) d_rgb0[i] = gc * (rgb0[i] - t);
  }
}

}  // namespace nb

using namespace nb;

extern "C" {

int nb_mse_pair_fwd(const float* rgb, const float* rgb0, const float* target, int64_t n3, float* out4, void* stream) {
  NB_REQUIRE(rgb && target && out4 && n3 >= 1, "nb_mse_pair_fwd: bad arguments");
  mse_pair_fwd_kernel<<<1, 1024, 0, as_stream(stream)>>>(rgb, rgb0, target, n3, out4);
  return check_launch("nb_mse_pair_fwd");
}

int nb_mse_pair_bwd(const float* rgb, const float* rgb0, const float* target, int64_t n3, const float* g_fine,
                    const float* g_coarse, const float* g_total, float* d_rgb, float* d_rgb0, void* stream) {
  NB_REQUIRE(rgb && target && d_rgb && n3 >= 1 && (!d_rgb0 || rgb0), "nb_mse_pair_bwd: bad arguments");
  mse_pair_bwd_kernel<<<grid_for(n3, 256), 256, 0, as_stream(stream)>>>(rgb, rgb0, target, n3, g_fine, g_coarse, g_total,
                                                                        d_rgb, d_rgb0);
  return check_launch("nb_mse_pair_bwd");
}

}  // extern "C"
