// Volume compositing (Renderer.raw2outputs, render_utils.py:241-290) forward and backward.
//
// One warp per ray.  Lane l owns the C = S/32 consecutive samples l*C .. l*C+C-1: the exclusive
// transmittance product is a serial product inside the lane plus one warp shuffle scan of the
// lane totals, the backward is the matching reverse (suffix-sum) scan.  A warp reads one
// contiguous 20*S-byte run per ray.  HBM-bound: the roofline is algorithmic bytes / measured
// copy bandwidth.
#include "common.cuh"

namespace nb {

// ex2.approx + rcp.approx: relative error < 1e-6 (the maps are compared at 2e-5); exp() of the opacity stays exact
__device__ __forceinline__ float sigmoidf_(float x) { return __fdividef(1.0f, 1.0f + __expf(-x)); }

// exclusive prefix product over the warp (branch-free); `total` = product of all 32 lanes
__device__ __forceinline__ float warp_excl_prod(float q, int lane, float& total) {
  float inc = q;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const float up = __shfl_up_sync(0xffffffffu, inc, o);
    inc *= (lane >= o) ? up : 1.0f;
  }
  total = __shfl_sync(0xffffffffu, inc, 31);
  const float ex = __shfl_up_sync(0xffffffffu, inc, 1);
  return lane == 0 ? 1.0f : ex;
}

// exclusive suffix sum over the warp (sum of lanes > lane), branch-free
__device__ __forceinline__ float warp_excl_suffix_sum(float v, int lane) {
  float inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const float dn = __shfl_down_sync(0xffffffffu, inc, o);
    inc += (lane + o < 32) ? dn : 0.0f;
  }
  return inc - v;
}

// One ray in registers.  Lane l owns the C consecutive samples l*C .. l*C+C-1, so the transmittance product is a
// serial product inside the lane plus ONE warp scan of the 32 lane totals per ray (not one scan per 32 samples),
// and every lane reads / writes one contiguous 16*C-byte run of raw (4*C bytes of z).  After load():
//   c[j] = (sigmoid(r), sigmoid(g), sigmoid(b), relu(sigma + noise)), zz[j] = depth, alpha[j], T[j] = transmittance.
// EXACT: S == 32*C (vector loads, no bounds tests); otherwise samples >= S are empty (alpha 0).
template <int C, bool EXACT>
struct RayRegs {
  float4 c[C];
  float zz[C], alpha[C], T[C];
  float dnorm, znext;  // |rays_d|; depth of the first sample of the next lane

  __device__ __forceinline__ bool valid(int j, int lane, int S) const { return EXACT || lane * C + j < S; }
  __device__ __forceinline__ float dz(int j, int lane, int S) const {
    const float nx = (j + 1 < C) ? zz[j + 1 < C ? j + 1 : j] : znext;
    return (lane * C + j + 1 >= S) ? 1e10f : nx - zz[j];
  }

  __device__ __forceinline__ void load(const float* __restrict__ raw, const float* __restrict__ z,
                                       const float* __restrict__ noise, const float* __restrict__ rays_d, int64_t r,
                                       int S, int lane) {
    const int64_t base = r * S + (int64_t)lane * C;
    float nz[C];
    if constexpr (EXACT) {
      const float4* rp = reinterpret_cast<const float4*>(raw) + base;
      const float2* zp = reinterpret_cast<const float2*>(z + base);
#pragma unroll
      for (int j = 0; j < C; ++j) c[j] = __ldcs(rp + j);
#pragma unroll
      for (int j = 0; j < C / 2; ++j) {
        const float2 v = __ldcs(zp + j);
        zz[2 * j] = v.x; zz[2 * j + 1] = v.y;
      }
      if (noise) {
        const float2* np_ = reinterpret_cast<const float2*>(noise + base);
#pragma unroll
        for (int j = 0; j < C / 2; ++j) {
          const float2 v = __ldcs(np_ + j);
          nz[2 * j] = v.x; nz[2 * j + 1] = v.y;
        }
      } else {
#pragma unroll
        for (int j = 0; j < C; ++j) nz[j] = 0.f;
      }
    } else {
#pragma unroll
      for (int j = 0; j < C; ++j) {
        const bool ok = lane * C + j < S;
        c[j] = ok ? __ldcs(reinterpret_cast<const float4*>(raw) + base + j) : make_float4(0.f, 0.f, 0.f, 0.f);
        zz[j] = ok ? z[base + j] : 0.f;
        nz[j] = (ok && noise) ? noise[base + j] : 0.f;
      }
    }
    const float d0 = rays_d[r * 3], d1 = rays_d[r * 3 + 1], d2 = rays_d[r * 3 + 2];
    dnorm = sqrtf(d0 * d0 + d1 * d1 + d2 * d2);
    znext = __shfl_down_sync(0xffffffffu, zz[0], 1);
    float prod = 1.0f;
#pragma unroll
    for (int j = 0; j < C; ++j) {
      const float rs = fmaxf(c[j].w + nz[j], 0.0f);
      const float a = valid(j, lane, S) ? 1.0f - expf(-rs * (dz(j, lane, S) * dnorm)) : 0.0f;
      c[j] = make_float4(sigmoidf_(c[j].x), sigmoidf_(c[j].y), sigmoidf_(c[j].z), rs);
      alpha[j] = a;
      T[j] = prod;
      prod *= 1.0f - a + 1e-10f;
    }
    float total;
    const float ex = warp_excl_prod(prod, lane, total);
#pragma unroll
    for (int j = 0; j < C; ++j) T[j] *= ex;
  }
};

template <int C, bool EXACT>
__global__ void __launch_bounds__(256) composite_fwd_kernel(const float* __restrict__ raw,
                                                            const float* __restrict__ z,
                                                            const float* __restrict__ rays_d,
                                                            const float* __restrict__ noise, int64_t n, int S,
                                                            int white_bkgd, float* __restrict__ rgb_map,
                                                            float* __restrict__ disp, float* __restrict__ acc_out,
                                                            float* __restrict__ weights, float* __restrict__ depth) {
  const int lane = threadIdx.x & 31;
  const int64_t r = (int64_t)blockIdx.x * 8 + __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  if (r >= n) return;
  RayRegs<C, EXACT> ray;
  ray.load(raw, z, noise, rays_d, r, S, lane);
  float a0 = 0.f, a1 = 0.f, a2 = 0.f, adepth = 0.f, aacc = 0.f;
  float w[C];
#pragma unroll
  for (int j = 0; j < C; ++j) {
    w[j] = ray.alpha[j] * ray.T[j];
    a0 += w[j] * ray.c[j].x; a1 += w[j] * ray.c[j].y; a2 += w[j] * ray.c[j].z;
    adepth += w[j] * ray.zz[j];
    aacc += w[j];
  }
  if (weights) {
    float* wp = weights + r * S + (int64_t)lane * C;
    if constexpr (EXACT) {
#pragma unroll
      for (int j = 0; j < C / 2; ++j) __stcs(reinterpret_cast<float2*>(wp) + j, make_float2(w[2 * j], w[2 * j + 1]));
    } else {
#pragma unroll
      for (int j = 0; j < C; ++j)
        if (lane * C + j < S) wp[j] = w[j];
    }
  }
  a0 = warp_sum(a0); a1 = warp_sum(a1); a2 = warp_sum(a2);
  adepth = warp_sum(adepth); aacc = warp_sum(aacc);
  if (lane == 0) {
    if (rgb_map) {
      const float bg = white_bkgd ? (1.0f - aacc) : 0.0f;
      rgb_map[r * 3] = a0 + bg; rgb_map[r * 3 + 1] = a1 + bg; rgb_map[r * 3 + 2] = a2 + bg;
    }
    if (acc_out) acc_out[r] = aacc;
    if (depth) depth[r] = adepth;
    if (disp) {
      const float ratio = adepth / aacc;                              // nan when acc == 0 (reference keeps it)
      const float m = (ratio != ratio) ? ratio : fmaxf(1e-10f, ratio);
      disp[r] = 1.0f / m;
    }
  }
}

// S == 64 (the coarse pass of every shipped config): half a warp per ray, lane owns 4 consecutive samples.  Same
// arithmetic per sample as RayRegs; the scan and the five reductions are 16 lanes wide and every instruction serves
// two rays, which halves the instructions per ray (the 32-lane version of this size is issue-bound, not HBM-bound).
__global__ void __launch_bounds__(256) composite_fwd_s64_kernel(const float* __restrict__ raw,
                                                                const float* __restrict__ z,
                                                                const float* __restrict__ rays_d,
                                                                const float* __restrict__ noise, int64_t n,
                                                                int white_bkgd, float* __restrict__ rgb_map,
                                                                float* __restrict__ disp, float* __restrict__ acc_out,
                                                                float* __restrict__ weights, float* __restrict__ depth) {
  constexpr int S = 64, C = 4;
  const int lane = threadIdx.x & 31, sub = lane & 15;
  const int64_t r_raw = ((int64_t)blockIdx.x * 8 + (threadIdx.x >> 5)) * 2 + (lane >> 4);
  const bool live = r_raw < n;
  const int64_t r = live ? r_raw : n - 1;
  const int64_t base = r * S + sub * C;
  float4 c[C];
  const float4* rp = reinterpret_cast<const float4*>(raw) + base;
#pragma unroll
  for (int j = 0; j < C; ++j) c[j] = __ldcs(rp + j);
  const float4 zv = __ldcs(reinterpret_cast<const float4*>(z + base));
  const float4 nv = noise ? __ldcs(reinterpret_cast<const float4*>(noise + base)) : make_float4(0.f, 0.f, 0.f, 0.f);
  const float zz[C] = {zv.x, zv.y, zv.z, zv.w}, nz[C] = {nv.x, nv.y, nv.z, nv.w};
  const float d0 = rays_d[r * 3], d1 = rays_d[r * 3 + 1], d2 = rays_d[r * 3 + 2];
  const float dnorm = sqrtf(d0 * d0 + d1 * d1 + d2 * d2);
  const float znext = __shfl_down_sync(0xffffffffu, zz[0], 1);
  float alpha[C], T[C], prod = 1.0f;
#pragma unroll
  for (int j = 0; j < C; ++j) {
    const float nx = j + 1 < C ? zz[j + 1 < C ? j + 1 : j] : znext;
    const float dz = (sub * C + j + 1 >= S) ? 1e10f : nx - zz[j];
    const float rs = fmaxf(c[j].w + nz[j], 0.0f);
    const float a = 1.0f - expf(-rs * (dz * dnorm));
    c[j] = make_float4(sigmoidf_(c[j].x), sigmoidf_(c[j].y), sigmoidf_(c[j].z), rs);
    alpha[j] = a;
    T[j] = prod;
    prod *= 1.0f - a + 1e-10f;
  }
  float inc = prod;
#pragma unroll
  for (int o = 1; o < 16; o <<= 1) {
    const float up = __shfl_up_sync(0xffffffffu, inc, o, 16);
    inc *= (sub >= o) ? up : 1.0f;
  }
  const float exu = __shfl_up_sync(0xffffffffu, inc, 1, 16);
  const float ex = sub == 0 ? 1.0f : exu;
  float a0 = 0.f, a1 = 0.f, a2 = 0.f, adepth = 0.f, aacc = 0.f, w[C];
#pragma unroll
  for (int j = 0; j < C; ++j) {
    w[j] = alpha[j] * (T[j] * ex);
    a0 += w[j] * c[j].x; a1 += w[j] * c[j].y; a2 += w[j] * c[j].z;
    adepth += w[j] * zz[j];
    aacc += w[j];
  }
  if (weights && live) __stcs(reinterpret_cast<float4*>(weights + base), make_float4(w[0], w[1], w[2], w[3]));
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) {
    a0 += __shfl_xor_sync(0xffffffffu, a0, o);
    a1 += __shfl_xor_sync(0xffffffffu, a1, o);
    a2 += __shfl_xor_sync(0xffffffffu, a2, o);
    adepth += __shfl_xor_sync(0xffffffffu, adepth, o);
    aacc += __shfl_xor_sync(0xffffffffu, aacc, o);
  }
  if (sub == 0 && live) {
    if (rgb_map) {
      const float bg = white_bkgd ? (1.0f - aacc) : 0.0f;
      rgb_map[r * 3] = a0 + bg; rgb_map[r * 3 + 1] = a1 + bg; rgb_map[r * 3 + 2] = a2 + bg;
    }
    if (acc_out) acc_out[r] = aacc;
    if (depth) depth[r] = adepth;
    if (disp) {
      const float ratio = adepth / aacc;                              // nan when acc == 0 (reference keeps it)
      const float m = (ratio != ratio) ? ratio : fmaxf(1e-10f, ratio);
      disp[r] = 1.0f / m;
    }
  }
}

// Reverse pass: dL/dw_s = g_s, dL/dalpha_s = g_s T_s - (sum_{k>s} g_k w_k) / q_s; the suffix sum is a serial sum
// inside the lane plus one warp scan of the lane totals.  Nothing but the ray's registers is live between the two
// directions (g_s is recomputed: five FMAs).
template <int C, bool EXACT>
__global__ void __launch_bounds__(256) composite_bwd_kernel(const float* __restrict__ raw,
                                                            const float* __restrict__ z,
                                                            const float* __restrict__ rays_d,
                                                            const float* __restrict__ noise, int64_t n, int S,
                                                            int white_bkgd, const float* __restrict__ g_rgb,
                                                            const float* __restrict__ g_disp,
                                                            const float* __restrict__ g_acc,
                                                            const float* __restrict__ g_depth,
                                                            float* __restrict__ d_raw, float* __restrict__ d_rays_d) {
  const int lane = threadIdx.x & 31;
  const int64_t r = (int64_t)blockIdx.x * 8 + __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  if (r >= n) return;
  // upstream -> per-ray scalars (requested before the ray itself)
  const float gr0 = g_rgb ? g_rgb[r * 3] : 0.f, gr1 = g_rgb ? g_rgb[r * 3 + 1] : 0.f, gr2 = g_rgb ? g_rgb[r * 3 + 2] : 0.f;
  float gdepth = g_depth ? g_depth[r] : 0.f;
  float gacc = g_acc ? g_acc[r] : 0.f;
  const float gd = g_disp ? g_disp[r] : 0.f;
  RayRegs<C, EXACT> ray;
  ray.load(raw, z, noise, rays_d, r, S, lane);
  if (white_bkgd) gacc -= (gr0 + gr1 + gr2);
  if (g_disp) {
    float adepth = 0.f, aacc = 0.f;
#pragma unroll
    for (int j = 0; j < C; ++j) {
      const float w = ray.alpha[j] * ray.T[j];
      adepth += w * ray.zz[j];
      aacc += w;
    }
    adepth = warp_sum(adepth);
    aacc = warp_sum(aacc);
    const float ratio = adepth / aacc;
    if (ratio > 1e-10f) {
      const float gratio = -gd / (ratio * ratio);
      gdepth += gratio / aacc;
      gacc -= gratio * adepth / (aacc * aacc);
    }
  }
  float lane_gw = 0.f;
#pragma unroll
  for (int j = 0; j < C; ++j) {
    const float g = gr0 * ray.c[j].x + gr1 * ray.c[j].y + gr2 * ray.c[j].z + gdepth * ray.zz[j] + gacc;
    lane_gw += g * (ray.alpha[j] * ray.T[j]);
  }
  float suffix = warp_excl_suffix_sum(lane_gw, lane);
  float gnorm = 0.f;
  float4* op = reinterpret_cast<float4*>(d_raw) + r * S + (int64_t)lane * C;
#pragma unroll
  for (int j = C - 1; j >= 0; --j) {
    const float4 c = ray.c[j];
    const float a = ray.alpha[j], T = ray.T[j], w = a * T;
    const float g = gr0 * c.x + gr1 * c.y + gr2 * c.z + gdepth * ray.zz[j] + gacc;
    const float e = 1.0f - a;  // exp(-rs*dist)
    const float galpha = g * T - __fdividef(suffix, e + 1e-10f);
    suffix += g * w;
    // alpha = 1 - exp(-rs*dist): d/d(rs) = dist*(1-alpha), d/d(dist) = rs*(1-alpha)
    const float dzj = ray.dz(j, lane, S);
    const float gsigma = (c.w > 0.f) ? galpha * (dzj * ray.dnorm) * e : 0.f;
    gnorm += galpha * c.w * e * dzj;
    if (ray.valid(j, lane, S)) {
      float4 o;
      o.x = w * gr0 * c.x * (1.0f - c.x);
      o.y = w * gr1 * c.y * (1.0f - c.y);
      o.z = w * gr2 * c.z * (1.0f - c.z);
      o.w = gsigma;
      __stcs(op + j, o);
    }
  }
  if (d_rays_d) {
    gnorm = warp_sum(gnorm);
    if (lane == 0) {
      const float d0 = rays_d[r * 3], d1 = rays_d[r * 3 + 1], d2 = rays_d[r * 3 + 2];
      const float inv = 1.0f / ray.dnorm;
      d_rays_d[r * 3] = gnorm * d0 * inv; d_rays_d[r * 3 + 1] = gnorm * d1 * inv; d_rays_d[r * 3 + 2] = gnorm * d2 * inv;
    }
  }
}

// Training fusion (main.py:77-103 between the MLP forward and its backward): compositing forward, the MSE term of
// this map against the target pixels, and the compositing backward of that term, one pass over raw.  The loss is a
// sum of per-map MSE terms (main.py:87-99), so d loss / d rgb_map = 2 (rgb_map - target) * inv_n3 is known as soon as
// the ray's own map value is; nothing else of the batch is needed.  Replaces composite_fwd + mse_pair_fwd +
// mse_pair_bwd + composite_bwd (four launches, raw read twice) on the critical path of the step.
// loss_sum: one float, += sum over the batch of (rgb_map - target)^2 * inv_n3 (the caller zeroes it).
template <int C, bool EXACT>
__global__ void __launch_bounds__(256) composite_train_kernel(const float* __restrict__ raw, const float* __restrict__ z,
                                                              const float* __restrict__ rays_d,
                                                              const float* __restrict__ noise,
                                                              const float* __restrict__ target, int64_t n, int S,
                                                              int white_bkgd, float inv_n3, float* __restrict__ rgb_map,
                                                              float* __restrict__ disp, float* __restrict__ acc_out,
                                                              float* __restrict__ weights, float* __restrict__ d_raw,
                                                              float* __restrict__ loss_sum) {
  __shared__ float blk_loss[8];
  const int lane = threadIdx.x & 31, wid = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int64_t r = (int64_t)blockIdx.x * 8 + wid;
  float sq = 0.f;
  if (r < n) {
    const float t0 = target[r * 3], t1 = target[r * 3 + 1], t2 = target[r * 3 + 2];
    RayRegs<C, EXACT> ray;
    ray.load(raw, z, noise, rays_d, r, S, lane);
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, adepth = 0.f, aacc = 0.f;
    float w[C];
#pragma unroll
    for (int j = 0; j < C; ++j) {
      w[j] = ray.alpha[j] * ray.T[j];
      a0 += w[j] * ray.c[j].x; a1 += w[j] * ray.c[j].y; a2 += w[j] * ray.c[j].z;
      adepth += w[j] * ray.zz[j];
      aacc += w[j];
    }
    if (weights) {
      float* wp = weights + r * S + (int64_t)lane * C;
#pragma unroll
      for (int j = 0; j < C; ++j)
        if (ray.valid(j, lane, S)) wp[j] = w[j];
    }
    a0 = warp_sum(a0); a1 = warp_sum(a1); a2 = warp_sum(a2);
    adepth = warp_sum(adepth); aacc = warp_sum(aacc);
    const float bg = white_bkgd ? (1.0f - aacc) : 0.0f;
    const float m0 = a0 + bg, m1 = a1 + bg, m2 = a2 + bg;
    if (lane == 0) {
      if (rgb_map) { rgb_map[r * 3] = m0; rgb_map[r * 3 + 1] = m1; rgb_map[r * 3 + 2] = m2; }
      if (acc_out) acc_out[r] = aacc;
      if (disp) {
        const float ratio = adepth / aacc;
        const float m = (ratio != ratio) ? ratio : fmaxf(1e-10f, ratio);
        disp[r] = 1.0f / m;
      }
    }
    const float e0 = m0 - t0, e1 = m1 - t1, e2 = m2 - t2;
    sq = (e0 * e0 + e1 * e1 + e2 * e2) * inv_n3;
    const float s2 = 2.0f * inv_n3;
    const float gr0 = s2 * e0, gr1 = s2 * e1, gr2 = s2 * e2;
    const float gacc = white_bkgd ? -(gr0 + gr1 + gr2) : 0.f;
    float lane_gw = 0.f;
#pragma unroll
    for (int j = 0; j < C; ++j) lane_gw += (gr0 * ray.c[j].x + gr1 * ray.c[j].y + gr2 * ray.c[j].z + gacc) * w[j];
    float suffix = warp_excl_suffix_sum(lane_gw, lane);
    float4* op = reinterpret_cast<float4*>(d_raw) + r * S + (int64_t)lane * C;
#pragma unroll
    for (int j = C - 1; j >= 0; --j) {
      const float4 c = ray.c[j];
      const float a = ray.alpha[j], T = ray.T[j];
      const float g = gr0 * c.x + gr1 * c.y + gr2 * c.z + gacc;
      const float e = 1.0f - a;
      const float galpha = g * T - __fdividef(suffix, e + 1e-10f);
      suffix += g * w[j];
      const float gsigma = (c.w > 0.f) ? galpha * (ray.dz(j, lane, S) * ray.dnorm) * e : 0.f;
      if (ray.valid(j, lane, S)) {
        float4 o;
        o.x = w[j] * gr0 * c.x * (1.0f - c.x);
        o.y = w[j] * gr1 * c.y * (1.0f - c.y);
        o.z = w[j] * gr2 * c.z * (1.0f - c.z);
        o.w = gsigma;
        __stcs(op + j, o);
      }
    }
  }
  if (lane == 0) blk_loss[wid] = sq;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += blk_loss[i];
    atomicAdd(loss_sum, t);
  }
}

// out4 = (mse_fine, mse_coarse, their sum, -10 log10(mse_fine)) from the two accumulators of composite_train, which
// are reset for the next step (main.py:87-99)
__global__ void loss_finalize_kernel(float* __restrict__ acc_fine, float* __restrict__ acc_coarse, float* __restrict__ out4) {
  const float m = acc_fine[0], m0 = acc_coarse ? acc_coarse[0] : 0.f;
  out4[0] = m; out4[1] = m0; out4[2] = m + m0;
  out4[3] = -10.0f * logf(m) / logf(10.0f);
  acc_fine[0] = 0.f;
  if (acc_coarse) acc_coarse[0] = 0.f;
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
  NB_REQUIRE((n + 7) / 8 < (1ll << 31), "nb_composite_fwd: too many rays for one launch");
  const unsigned grid = (unsigned)((n + 7) / 8);   // one warp per ray
  cudaStream_t st = as_stream(stream);
#define NB_LAUNCH_FWD(C)                                                                                              \
  do {                                                                                                                \
    if (S == 32 * C)                                                                                                  \
      composite_fwd_kernel<C, true><<<grid, 256, 0, st>>>(raw, z, rays_d, noise, n, S, white_bkgd, rgb_map, disp, acc, \
                                                          weights, depth);                                            \
    else                                                                                                              \
      composite_fwd_kernel<C, false><<<grid, 256, 0, st>>>(raw, z, rays_d, noise, n, S, white_bkgd, rgb_map, disp,    \
                                                           acc, weights, depth);                                      \
  } while (0)
  const bool al16 = (((uintptr_t)raw | (uintptr_t)z | (uintptr_t)noise | (uintptr_t)weights) & 15) == 0;
  if (S == 64 && al16 && (n + 15) / 16 < (1ll << 31))
    composite_fwd_s64_kernel<<<(unsigned)((n + 15) / 16), 256, 0, st>>>(raw, z, rays_d, noise, n, white_bkgd, rgb_map, disp,
                                                                      acc, weights, depth);
  else if (S <= 64) NB_LAUNCH_FWD(2);
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
  NB_REQUIRE((n + 7) / 8 < (1ll << 31), "nb_composite_bwd: too many rays for one launch");
  const unsigned grid = (unsigned)((n + 7) / 8);
  cudaStream_t st = as_stream(stream);
#define NB_LAUNCH_BWD(C)                                                                                              \
  do {                                                                                                                \
    if (S == 32 * C)                                                                                                  \
      composite_bwd_kernel<C, true><<<grid, 256, 0, st>>>(raw, z, rays_d, noise, n, S, white_bkgd, d_rgb, d_disp,     \
                                                          d_acc, d_depth, d_raw, d_rays_d);                           \
    else                                                                                                              \
      composite_bwd_kernel<C, false><<<grid, 256, 0, st>>>(raw, z, rays_d, noise, n, S, white_bkgd, d_rgb, d_disp,    \
                                                           d_acc, d_depth, d_raw, d_rays_d);                          \
  } while (0)
  if (S <= 64) NB_LAUNCH_BWD(2);
  else if (S <= 128) NB_LAUNCH_BWD(4);
  else if (S <= 192) NB_LAUNCH_BWD(6);
  else if (S <= 256) NB_LAUNCH_BWD(8);
  else NB_LAUNCH_BWD(16);
#undef NB_LAUNCH_BWD
  return check_launch("nb_composite_bwd");
}

int nb_composite_train(const float* raw, const float* z, const float* rays_d, const float* noise, const float* target,
                       int64_t n, int S, int white_bkgd, float inv_n3, float* rgb_map, float* disp, float* acc,
                       float* weights, float* d_raw, float* loss_sum, void* stream) {
  if (n == 0) return 0;
  NB_REQUIRE(raw && z && rays_d && target && d_raw && loss_sum && S >= 1, "nb_composite_train: bad arguments");
  if (S > 512) {
    set_error("nb_composite_train: S=%d > 512 samples per ray is not implemented", S);
    return NB_E_UNSUPPORTED;
  }
  NB_REQUIRE((n + 7) / 8 < (1ll << 31), "nb_composite_train: too many rays for one launch");
  const unsigned grid = (unsigned)((n + 7) / 8);
  cudaStream_t st = as_stream(stream);
#define NB_LAUNCH_TR(C)                                                                                               \
  do {                                                                                                                \
    if (S == 32 * C)                                                                                                  \
      composite_train_kernel<C, true><<<grid, 256, 0, st>>>(raw, z, rays_d, noise, target, n, S, white_bkgd, inv_n3,  \
                                                            rgb_map, disp, acc, weights, d_raw, loss_sum);            \
    else                                                                                                              \
      composite_train_kernel<C, false><<<grid, 256, 0, st>>>(raw, z, rays_d, noise, target, n, S, white_bkgd, inv_n3, \
                                                             rgb_map, disp, acc, weights, d_raw, loss_sum);           \
  } while (0)
  if (S <= 64) NB_LAUNCH_TR(2);
  else if (S <= 128) NB_LAUNCH_TR(4);
  else if (S <= 192) NB_LAUNCH_TR(6);
  else if (S <= 256) NB_LAUNCH_TR(8);
  else NB_LAUNCH_TR(16);
#undef NB_LAUNCH_TR
  return check_launch("nb_composite_train");
}

int nb_loss_finalize(float* acc_fine, float* acc_coarse, float* out4, void* stream) {
  NB_REQUIRE(acc_fine && out4, "nb_loss_finalize: bad arguments");
  loss_finalize_kernel<<<1, 1, 0, as_stream(stream)>>>(acc_fine, acc_coarse, out4);
  return check_launch("nb_loss_finalize");
}

}  // extern "C"
