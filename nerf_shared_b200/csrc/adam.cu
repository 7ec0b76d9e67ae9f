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

// One launch in front of a graph replay: this step's Adam scalars (as nb_adam_hyper) and the copy of the step's two
// inputs (ray batch, targets) into the static buffers the captured graph reads -- three stream operations less
// between two replays of a 1.3 ms step.
namespace nb {
__global__ void __launch_bounds__(256) step_prologue_kernel(float* __restrict__ hyper, float lr_c, float b1, float b2,
                                                            float eps, float inv_sqrt_bc2, float grad_scale,
                                                            const float* __restrict__ src1, float* __restrict__ dst1,
                                                            int64_t n1, const float* __restrict__ src2,
                                                            float* __restrict__ dst2, int64_t n2) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i == 0) {
    hyper[0] = lr_c; hyper[1] = b1; hyper[2] = b2; hyper[3] = eps; hyper[4] = inv_sqrt_bc2; hyper[5] = grad_scale;
  }
  if (i < n1) dst1[i] = src1[i];
  else if (i < n1 + n2) dst2[i - n1] = src2[i - n1];
}
}  // namespace nb

extern "C" int nb_train_step_prologue(float* hyper, float lr, float beta1, float beta2, float eps, int step,
                                      float grad_scale, const float* src1, float* dst1, int64_t n1, const float* src2,
                                      float* dst2, int64_t n2, void* stream) {
  NB_REQUIRE(hyper && step >= 1 && n1 >= 0 && n2 >= 0 && (n1 == 0 || (src1 && dst1)) && (n2 == 0 || (src2 && dst2)),
             "nb_train_step_prologue: bad arguments");
  double bc1 = 1.0 - pow((double)beta1, step), bc2 = 1.0 - pow((double)beta2, step);
  const int64_t n = n1 + n2 > 1 ? n1 + n2 : 1;
  nb::step_prologue_kernel<<<(unsigned)((n + 255) / 256), 256, 0, nb::as_stream(stream)>>>(
      hyper, (float)(lr / bc1), beta1, beta2, eps, (float)(1.0 / sqrt(bc2)), grad_scale, src1, dst1, n1, src2, dst2, n2);
  return nb::check_launch("nb_train_step_prologue");
}

extern "C" int nb_adam_step_dev(float* p, const float* g, float* m, float* v, int64_t n, const float* hyper,
                                void* stream) {
  if (n == 0) return 0;
  NB_REQUIRE(p && g && m && v && hyper, "nb_adam_step_dev: bad arguments");
  adam_dev_kernel<<<grid_for(n, 256), 256, 0, as_stream(stream)>>>(p, g, m, v, n, hyper);
  return check_launch("nb_adam_step_dev");
}
