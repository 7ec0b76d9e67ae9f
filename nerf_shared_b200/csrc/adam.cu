// Fused Adam over flat fp32 buffers (torch.optim.Adam semantics, utils.py:170 / main.py:104).
// Streaming kernel: reads p, g, m, v (16 B/param) and writes p, m, v (12 B/param).
#include "common.cuh"

namespace nb {

__global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                   float* __restrict__ m, float* __restrict__ v, int64_t n,
                                                   float step_size, float beta1, float beta2, float eps,
                                                   float inv_sqrt_bc2, float grad_scale) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float gi = g[i] * grad_scale;
    float mi = beta1 * m[i] + (1.0f - beta1) * gi;
    float vi = beta2 * v[i] + (1.0f - beta2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    float denom = sqrtf(vi) * inv_sqrt_bc2 + eps;
    p[i] = p[i] - step_size * (mi / denom);
  }
}

// Same update with the scalars read from device memory (hyper[0..5] = step_size, beta1, beta2, eps,
// 1/sqrt(bias_correction2), grad_scale): the launch is then identical every step and can sit in a
// CUDA graph while the learning-rate schedule (main.py:108-112) and bias corrections advance.
__global__ void __launch_bounds__(256) adam_dev_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                       float* __restrict__ m, float* __restrict__ v, int64_t n,
                                                       const float* __restrict__ hyper) {
  const float step_size = hyper[0], beta1 = hyper[1], beta2 = hyper[2], eps = hyper[3], inv_sqrt_bc2 = hyper[4],
              grad_scale = hyper[5];
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float gi = g[i] * grad_scale;
    float mi = beta1 * m[i] + (1.0f - beta1) * gi;
    float vi = beta2 * v[i] + (1.0f - beta2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    float denom = sqrtf(vi) * inv_sqrt_bc2 + eps;
    p[i] = p[i] - step_size * (mi / denom);
  }
}

__global__ void adam_hyper_kernel(float* __restrict__ hyper, float step_size, float beta1, float beta2, float eps,
                                  float inv_sqrt_bc2, float grad_scale) {
  hyper[0] = step_size; hyper[1] = beta1; hyper[2] = beta2; hyper[3] = eps; hyper[4] = inv_sqrt_bc2; hyper[5] = grad_scale;
}

}  // namespace nb

using namespace nb;

extern "C" int nb_adam_step(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2,
                            float eps, int step, float grad_scale, void* stream) {
  if (n == 0) return 0;
  NB_REQUIRE(p && g && m && v && step >= 1, "nb_adam_step: bad arguments");
  double bc1 = 1.0 - pow((double)beta1, step), bc2 = 1.0 - pow((double)beta2, step);
  adam_kernel<<<grid_for(n, 256), 256, 0, as_stream(stream)>>>(p, g, m, v, n, (float)(lr / bc1), beta1, beta2, eps,
                                                              (float)(1.0 / sqrt(bc2)), grad_scale);
  return check_launch("nb_adam_step");
}

extern "C" int nb_adam_hyper(float* hyper, float lr, float beta1, float beta2, float eps, int step, float grad_scale,
                             void* stream) {
  NB_REQUIRE(hyper && step >= 1, "nb_adam_hyper: bad arguments");
  double bc1 = 1.0 - pow((double)beta1, step), bc2 = 1.0 - pow((double)beta2, step);
  adam_hyper_kernel<<<1, 1, 0, as_stream(stream)>>>(hyper, (float)(lr / bc1), beta1, beta2, eps, (float)(1.0 / sqrt(bc2)),
                                                    grad_scale);
  return check_launch("nb_adam_hyper");
}

extern "C" int nb_adam_step_dev(float* p, const float* g, float* m, float* v, int64_t n, const float* hyper,
                                void* stream) {
  if (n == 0) return 0;
  NB_REQUIRE(p && g && m && v && hyper, "nb_adam_step_dev: bad arguments");
  adam_dev_kernel<<<grid_for(n, 256), 256, 0, as_stream(stream)>>>(p, g, m, v, n, hyper);
  return check_launch("nb_adam_step_dev");
}
