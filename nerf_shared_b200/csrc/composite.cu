// Volume compositing (Renderer.raw2outputs, render_utils.py:241-290) forward and backward.
//
// One warp per ray.  Lane l owns samples l, l+32, l+64, ... so that every global access of a
// warp is one contiguous run (raw: 32 x float4 = 512 B, z / weights: 128 B).  The exclusive
// transmittance product is a warp shuffle scan per 32-sample chunk with a carried product;
// the backward is the matching reverse (suffix-sum) scan.  HBM-bound: the roofline is
// algorithmic bytes / measured copy bandwidth.
#include "common.cuh"

namespace nb {

struct SampleTerms {
  float c0, c1, c2;  // sigmoid(rgb)
  float alpha;       // 1 - exp(-relu(sigma)*dist)
  float q;           // 1 - alpha + 1e-10
  float rs;          // relu(sigma + noise)
  float dist;        // interval length * |rays_d|
  float dz;          // interval length before the |rays_d| factor
  float z;
};

// ex2.approx + rcp.approx: relative error < 1e-6 (the maps are compared at 2e-5); exp() of the opacity stays exact
__device__ __forceinline__ float sigmoidf_(float x) { return __fdividef(1.0f, 1.0f + __expf(-x)); }

__device__ __forceinline__ SampleTerms load_terms(const float* __restrict__ raw, const float* __restrict__ z,
                                                  const float* __restrict__ noise, int64_t base, int s, int S,
                                                  float dnorm) {
  SampleTerms t;
  if (s < S) {
    float4 r = *reinterpret_cast<const float4*>(raw + (base + s) * 4);
    t.z = z[base + s];
    t.dz = (s + 1 < S) ? (z[base + s + 1] - t.z) : 1e10f;
    t.dist = t.dz * dnorm;
    float sg = r.w;
    if (noise) sg += noise[base + s];
    t.rs = fmaxf(sg, 0.0f);
    t.alpha = 1.0f - expf(-t.rs * t.dist);
    t.q = 1.0f - t.alpha + 1e-10f;
    t.c0 = sigmoidf_(r.x); t.c1 = sigmoidf_(r.y); t.c2 = sigmoidf_(r.z);
  } else {
    t.c0 = t.c1 = t.c2 = 0.f; t.alpha = 0.f; t.q = 1.0f; t.rs = 0.f; t.dist = 0.f; t.dz = 0.f; t.z = 0.f;
  }
  return t;
}

// exclusive prefix product over the warp; returns the exclusive value, `total` = product of all 32
__device__ __forceinline__ float warp_excl_prod(float q, int lane, float& total) {
  float inc = q;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    float up = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc *= up;
  }
  total = __shfl_sync(0xffffffffu, inc, 31);
  float ex = __shfl_up_sync(0xffffffffu, inc, 1);
  return lane == 0 ? 1.0f : ex;
}

// exclusive suffix sum over the warp (sum of lanes > lane); `total` = sum of all 32
__device__ __forceinline__ float warp_excl_suffix_sum(float v, int lane, float& total) {
  float inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    float dn = __shfl_down_sync(0xffffffffu, inc, o);
    if (lane + o < 32) inc += dn;
  }
  total = __shfl_sync(0xffffffffu, inc, 0);
  return inc - v;
}

// terms of one sample from values already in registers (z_next = depth of the following sample)
__device__ __forceinline__ SampleTerms make_terms(const float4& r, float zz, float z_next, bool last, bool valid, float nz,
                                                  float dnorm) {
  SampleTerms t;
  if (valid) {
    t.z = zz;
    t.dz = last ? 1e10f : (z_next - zz);
    t.dist = t.dz * dnorm;
    t.rs = fmaxf(r.w + nz, 0.0f);
    t.alpha = 1.0f - expf(-t.rs * t.dist);
    t.q = 1.0f - t.alpha + 1e-10f;
    t.c0 = sigmoidf_(r.x); t.c1 = sigmoidf_(r.y); t.c2 = sigmoidf_(r.z);
  } else {
    t.c0 = t.c1 = t.c2 = 0.f; t.alpha = 0.f; t.q = 1.0f; t.rs = 0.f; t.dist = 0.f; t.dz = 0.f; t.z = 0.f;
  }
  return t;
}

// All C 32-sample chunks of a ray are requested before any arithmetic (C x 640 B in flight per warp), and the
// depth of the following sample comes from the neighbouring lane instead of a second, misaligned load.
template <int C>
__device__ __forceinline__ void load_ray(const float* __restrict__ raw, const float* __restrict__ z,
                                         const float* __restrict__ noise, int64_t base, int S, int lane, float4 (&r)[C],
                                         float (&zz)[C], float (&zn)[C], float (&nz)[C]) {
#pragma unroll
  for (int j = 0; j < C; ++j) {
    const int s = j * 32 + lane;
    if (s < S) {
      r[j] = __ldcs(reinterpret_cast<const float4*>(raw + (base + s) * 4));
      zz[j] = z[base + s];
      nz[j] = noise ? noise[base + s] : 0.f;
    } else {
      r[j] = make_float4(0.f, 0.f, 0.f, 0.f);
      zz[j] = 0.f;
      nz[j] = 0.f;
    }
  }
#pragma unroll
  for (int j = 0; j < C; ++j) {
    float nx = __shfl_down_sync(0xffffffffu, zz[j], 1);
    const float first_of_next = __shfl_sync(0xffffffffu, j + 1 < C ? zz[j + 1 < C ? j + 1 : j] : 0.f, 0);
    zn[j] = lane == 31 ? first_of_next : nx;
  }
}

template <int C>
__global__ void __launch_bounds__(256) composite_fwd_kernel(const float* __restrict__ raw,
                                                            const float* __restrict__ z,
                                                            const float* __restrict__ rays_d,
                                                            const float* __restrict__ noise, int64_t n, int S,
                                                            int white_bkgd, float* __restrict__ rgb_map,
                                                            float* __restrict__ disp, float* __restrict__ acc_out,
                                                            float* __restrict__ weights, float* __restrict__ depth) {
  int lane = threadIdx.x & 31;
  int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp; r < n; r += nwarps) {
    float d0 = rays_d[r * 3], d1 = rays_d[r * 3 + 1], d2 = rays_d[r * 3 + 2];
    int64_t base = r * S;
    float4 rw[C];
    float zz[C], zn[C], nz[C];
    load_ray<C>(raw, z, noise, base, S, lane, rw, zz, zn, nz);
    float dnorm = sqrtf(d0 * d0 + d1 * d1 + d2 * d2);
    float carry = 1.0f, a0 = 0.f, a1 = 0.f, a2 = 0.f, adepth = 0.f, aacc = 0.f;
#pragma unroll
    for (int j = 0; j < C; ++j) {
      int s = j * 32 + lane;
      SampleTerms t = make_terms(rw[j], zz[j], zn[j], s + 1 >= S, s < S, nz[j], dnorm);
      float total;
      float T = carry * warp_excl_prod(t.q, lane, total);
      carry *= total;
      float w = t.alpha * T;
      if (s < S) {
        if (weights) __stcs(weights + base + s, w);
        a0 += w * t.c0; a1 += w * t.c1; a2 += w * t.c2;
        adepth += w * t.z;
        aacc += w;
      }
    }
    a0 = warp_sum(a0); a1 = warp_sum(a1); a2 = warp_sum(a2);
    adepth = warp_sum(adepth); aacc = warp_sum(aacc);
    if (lane == 0) {
      if (rgb_map) {
        float bg = white_bkgd ? (1.0f - aacc) : 0.0f;
        rgb_map[r * 3] = a0 + bg; rgb_map[r * 3 + 1] = a1 + bg; rgb_map[r * 3 + 2] = a2 + bg;
      }
      if (acc_out) acc_out[r] = aacc;
      if (depth) depth[r] = adepth;
      if (disp) {
        float ratio = adepth / aacc;                              // nan when acc == 0 (reference keeps it)
        float m = (ratio != ratio) ? ratio : fmaxf(1e-10f, ratio);
        disp[r] = 1.0f / m;
      }
    }
  }
}

// C = number of 32-sample chunks held in registers (S <= 32*C).  Only the inputs and the transmittance
// stay live between the forward and the reverse scan; the per-sample terms are recomputed (the kernel is
// bandwidth-bound, registers buy occupancy).
template <int C>
__global__ void __launch_bounds__(256, C <= 6 ? 3 : 1) composite_bwd_kernel(const float* __restrict__ raw,
                                                            const float* __restrict__ z,
                                                            const float* __restrict__ rays_d,
                                                            const float* __restrict__ noise, int64_t n, int S,
                                                            int white_bkgd, const float* __restrict__ g_rgb,
                                                            const float* __restrict__ g_disp,
                                                            const float* __restrict__ g_acc,
                                                            const float* __restrict__ g_depth,
                                                            float* __restrict__ d_raw, float* __restrict__ d_rays_d) {
  int lane = threadIdx.x & 31;
  int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp; r < n; r += nwarps) {
    float d0 = rays_d[r * 3], d1 = rays_d[r * 3 + 1], d2 = rays_d[r * 3 + 2];
    int64_t base = r * S;
    float4 rw[C];
    float zz[C], zn[C], nz[C], T[C];
    load_ray<C>(raw, z, noise, base, S, lane, rw, zz, zn, nz);
    float dnorm = sqrtf(d0 * d0 + d1 * d1 + d2 * d2);
    float carry = 1.0f, adepth = 0.f, aacc = 0.f;
#pragma unroll
    for (int j = 0; j < C; ++j) {
      int s = j * 32 + lane;
      SampleTerms t = make_terms(rw[j], zz[j], zn[j], s + 1 >= S, s < S, nz[j], dnorm);
      float total;
      T[j] = carry * warp_excl_prod(t.q, lane, total);
      carry *= total;
      float w = t.alpha * T[j];
      adepth += w * t.z;
      aacc += w;
    }
    adepth = warp_sum(adepth);
    aacc = warp_sum(aacc);
    // upstream -> per-ray scalars
    float gr0 = g_rgb ? g_rgb[r * 3] : 0.f, gr1 = g_rgb ? g_rgb[r * 3 + 1] : 0.f, gr2 = g_rgb ? g_rgb[r * 3 + 2] : 0.f;
    float gdepth = g_depth ? g_depth[r] : 0.f;
    float gacc = g_acc ? g_acc[r] : 0.f;
    if (white_bkgd) gacc -= (gr0 + gr1 + gr2);
    if (g_disp) {
      float ratio = adepth / aacc;
      if (ratio > 1e-10f) {
        float gratio = -g_disp[r] / (ratio * ratio);
        gdepth += gratio / aacc;
        gacc -= gratio * adepth / (aacc * aacc);
      }
    }
    // reverse scan of g_j * w_j
    float rcarry = 0.f, gnorm = 0.f;
#pragma unroll
    for (int j = C - 1; j >= 0; --j) {
      int s = j * 32 + lane;
      SampleTerms t = make_terms(rw[j], zz[j], zn[j], s + 1 >= S, s < S, nz[j], dnorm);
      float w = t.alpha * T[j];
      float g = gr0 * t.c0 + gr1 * t.c1 + gr2 * t.c2 + gdepth * t.z + gacc;  // dL/dw_s
      float total;
      float suffix = rcarry + warp_excl_suffix_sum(g * w, lane, total);
      rcarry += total;
      float galpha = g * T[j] - suffix / t.q;
      // alpha = 1 - exp(-rs*dist): d/d(rs) = dist*(1-alpha), d/d(dist) = rs*(1-alpha)
      float e = 1.0f - t.alpha;
      float gsigma = (t.rs > 0.f) ? galpha * t.dist * e : 0.f;
      gnorm += galpha * t.rs * e * t.dz;
      if (s < S) {
        float4 o;
        o.x = w * gr0 * t.c0 * (1.0f - t.c0);
        o.y = w * gr1 * t.c1 * (1.0f - t.c1);
        o.z = w * gr2 * t.c2 * (1.0f - t.c2);
        o.w = gsigma;
        __stcs(reinterpret_cast<float4*>(d_raw + (base + s) * 4), o);
      }
    }
    if (d_rays_d) {
      gnorm = warp_sum(gnorm);
      if (lane == 0) {
        float inv = 1.0f / dnorm;
        d_rays_d[r * 3] = gnorm * d0 * inv; d_rays_d[r * 3 + 1] = gnorm * d1 * inv; d_rays_d[r * 3 + 2] = gnorm * d2 * inv;
      }
    }
  }
}

}  // namespace nb

using namespace nb;

extern "C" {

int nb_composite_fwd(const float* raw, const float* z, const float* rays_d, const float* noise, int64_t n, int S,
                     int white_bkgd, float* rgb_map, float* disp, float* acc, float* weights, float* depth,
                     void* stream) {
  if (n == 0) return 0;
  NB_REQUIRE(raw && z && rays_d && S >= 1, "nb_composite_fwd: bad arguments");
  if (S > 512) {
    set_error("nb_composite_fwd: S=%d > 512 samples per ray is not implemented", S);
    return NB_E_UNSUPPORTED;
  }
  const int grid = grid_for(n, 8);
  cudaStream_t st = as_stream(stream);
#define NB_LAUNCH_FWD(C) \
  composite_fwd_kernel<C><<<grid, 256, 0, st>>>(raw, z, rays_d, noise, n, S, white_bkgd, rgb_map, disp, acc, weights, depth)
  if (S <= 64) NB_LAUNCH_FWD(2);
  else if (S <= 128) NB_LAUNCH_FWD(4);
  else if (S <= 192) NB_LAUNCH_FWD(6);
  else if (S <= 256) NB_LAUNCH_FWD(8);
  else NB_LAUNCH_FWD(16);
#undef NB_LAUNCH_FWD
  return check_launch("nb_composite_fwd");
}

int nb_composite_bwd(const float* raw, const float* z, const float* rays_d, const float* noise, int64_t n, int S,
                     int white_bkgd, const float* d_rgb, const float* d_disp, const float* d_acc,
                     const float* d_depth, float* d_raw, float* d_rays_d, void* stream) {
  if (n == 0) return 0;
  NB_REQUIRE(raw && z && rays_d && d_raw && S >= 1, "nb_composite_bwd: bad arguments");
  if (S > 512) {
    set_error("nb_composite_bwd: S=%d > 512 samples per ray is not implemented", S);
    return NB_E_UNSUPPORTED;
  }
  int grid = grid_for(n, 8);
  cudaStream_t st = as_stream(stream);
#define NB_LAUNCH_BWD(C)                                                                                         \
  composite_bwd_kernel<C><<<grid, 256, 0, st>>>(raw, z, rays_d, noise, n, S, white_bkgd, d_rgb, d_disp, d_acc, \
                                                d_depth, d_raw, d_rays_d)
  if (S <= 64) NB_LAUNCH_BWD(2);
  else if (S <= 128) NB_LAUNCH_BWD(4);
  else if (S <= 192) NB_LAUNCH_BWD(6);
  else if (S <= 256) NB_LAUNCH_BWD(8);
  else NB_LAUNCH_BWD(16);
#undef NB_LAUNCH_BWD
  return check_launch("nb_composite_bwd");
}

}  // extern "C"
