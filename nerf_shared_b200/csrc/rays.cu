// Ray generation, NDC warp, stratified sampling and sample positions.
//
// All of these are streaming (HBM-bound) kernels: one thread per output element group,
// consecutive threads touch consecutive addresses.  Where BASELINE.json asks for bit-exact
// results (get_rays, ndc_rays, z values) every fp32 op of the reference is issued as its own
// IEEE round-to-nearest intrinsic (__fmul_rn / __fadd_rn / __fdiv_rn): the compiler must not
// contract them into FMAs, because ATen rounds after every op.
#include "common.cuh"

namespace nb {

// dirs of pixel (x, y): utils.py:37
__device__ __forceinline__ void pixel_dir(int x, int y, float fx, float fy, float cx, float cy, float& d0,
                                          float& d1, float& d2) {
  d0 = __fdiv_rn(__fsub_rn((float)x, cx), fx);
  d1 = -__fdiv_rn(__fsub_rn((float)y, cy), fy);
  d2 = -1.0f;
}

// rays_d[k] = (d0*R[k][0] + d1*R[k][1]) + d2*R[k][2]: utils.py:39 (ATen sums 3 terms left to right)
__device__ __forceinline__ float rot_row(float d0, float d1, float d2, float r0, float r1, float r2) {
  return __fadd_rn(__fadd_rn(__fmul_rn(d0, r0), __fmul_rn(d1, r1)), __fmul_rn(d2, r2));
}

__global__ void __launch_bounds__(256) get_rays_kernel(int H, int W, float fx, float fy, float cx, float cy,
                                                       const float* __restrict__ c2w, int ld,
                                                       const int64_t* __restrict__ pix, int64_t n,
                                                       float* __restrict__ rays_o, float* __restrict__ rays_d,
                                                       float* __restrict__ viewdirs) {
  __shared__ float R[12];
  if (threadIdx.x < 12) R[threadIdx.x] = c2w[(threadIdx.x / 4) * ld + (threadIdx.x % 4)];
  __syncthreads();
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t p = pix ? pix[i] : i;
    int y = (int)(p / W), x = (int)(p - (int64_t)y * W);
    float d0, d1, d2;
    pixel_dir(x, y, fx, fy, cx, cy, d0, d1, d2);
    float r0 = rot_row(d0, d1, d2, R[0], R[1], R[2]);
    float r1 = rot_row(d0, d1, d2, R[4], R[5], R[6]);
    float r2 = rot_row(d0, d1, d2, R[8], R[9], R[10]);
    float* o = rays_o + i * 3;
    float* d = rays_d + i * 3;
    o[0] = R[3]; o[1] = R[7]; o[2] = R[11];
    d[0] = r0; d[1] = r1; d[2] = r2;
    if (viewdirs) {
      float nrm = sqrtf(r0 * r0 + r1 * r1 + r2 * r2);
      float* v = viewdirs + i * 3;
      v[0] = __fdiv_rn(r0, nrm); v[1] = __fdiv_rn(r1, nrm); v[2] = __fdiv_rn(r2, nrm);
    }
  }
}

// d_c2w[k][m] += sum_pix dirs[m] * d_rays_d[pix][k] ; d_c2w[k][3] += sum_pix d_rays_o[pix][k]
__global__ void __launch_bounds__(256) get_rays_bwd_kernel(int H, int W, float fx, float fy, float cx, float cy,
                                                           const int64_t* __restrict__ pix, int64_t n,
                                                           const float* __restrict__ g_o,
                                                           const float* __restrict__ g_d, float* __restrict__ d_c2w) {
  float acc[12];
#pragma unroll
  for (int k = 0; k < 12; ++k) acc[k] = 0.f;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t p = pix ? pix[i] : i;
    int y = (int)(p / W), x = (int)(p - (int64_t)y * W);
    float d0, d1, d2;
    pixel_dir(x, y, fx, fy, cx, cy, d0, d1, d2);
    if (g_d) {
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        float g = g_d[i * 3 + k];
        acc[k * 4 + 0] += d0 * g;
        acc[k * 4 + 1] += d1 * g;
        acc[k * 4 + 2] += d2 * g;
      }
    }
    if (g_o) {
#pragma unroll
      for (int k = 0; k < 3; ++k) acc[k * 4 + 3] += g_o[i * 3 + k];
    }
  }
  __shared__ float red[8][12];
  int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < 12; ++k) {
    float v = warp_sum(acc[k]);
    if (lane == 0) red[warp][k] = v;
  }
  __syncthreads();
  if (threadIdx.x < 12) {
    float v = 0.f;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) v += red[w][threadIdx.x];
    atomicAdd(d_c2w + threadIdx.x, v);
  }
}

__global__ void __launch_bounds__(256) normalize_fwd_kernel(const float* __restrict__ v, int64_t n,
                                                            float* __restrict__ out) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float a = v[i * 3], b = v[i * 3 + 1], c = v[i * 3 + 2];
    float nrm = sqrtf(a * a + b * b + c * c);
    out[i * 3] = __fdiv_rn(a, nrm); out[i * 3 + 1] = __fdiv_rn(b, nrm); out[i * 3 + 2] = __fdiv_rn(c, nrm);
  }
}

__global__ void __launch_bounds__(256) normalize_bwd_kernel(const float* __restrict__ v,
                                                            const float* __restrict__ g, int64_t n,
                                                            float* __restrict__ dv) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float a = v[i * 3], b = v[i * 3 + 1], c = v[i * 3 + 2];
    float inv = 1.0f / sqrtf(a * a + b * b + c * c);
    float ua = a * inv, ub = b * inv, uc = c * inv;
    float ga = g[i * 3], gb = g[i * 3 + 1], gc = g[i * 3 + 2];
    float dot = ua * ga + ub * gb + uc * gc;
    dv[i * 3] = (ga - ua * dot) * inv; dv[i * 3 + 1] = (gb - ub * dot) * inv; dv[i * 3 + 2] = (gc - uc * dot) * inv;
  }
}

// Ray rows of Renderer.render (render_utils.py:211,223-226): o(3) d(3) near far viewdirs(3), the view direction
// normalised from `view_src` (the rays before the NDC warp) with normalize_fwd_kernel's arithmetic; one launch
// instead of normalise + two fills + cat.  `viewdirs` (contiguous copy for the MLP) may be NULL.
__global__ void __launch_bounds__(256) ray_rows_kernel(const float* __restrict__ ro, const float* __restrict__ rd,
                                                       const float* __restrict__ view_src, float near_, float far_,
                                                       int64_t n, float* __restrict__ rows, float* __restrict__ viewdirs) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float a = view_src[i * 3], b = view_src[i * 3 + 1], c = view_src[i * 3 + 2];
  const float nrm = sqrtf(a * a + b * b + c * c);
  const float va = __fdiv_rn(a, nrm), vb = __fdiv_rn(b, nrm), vc = __fdiv_rn(c, nrm);
  float* r = rows + i * 11;
  r[0] = ro[i * 3]; r[1] = ro[i * 3 + 1]; r[2] = ro[i * 3 + 2];
  r[3] = rd[i * 3]; r[4] = rd[i * 3 + 1]; r[5] = rd[i * 3 + 2];
  r[6] = near_; r[7] = far_;
  r[8] = va; r[9] = vb; r[10] = vc;
  if (viewdirs) { viewdirs[i * 3] = va; viewdirs[i * 3 + 1] = vb; viewdirs[i * 3 + 2] = vc; }
}

// utils.py:54-71, op by op.
__global__ void __launch_bounds__(256) ndc_fwd_kernel(float sx, float sy, float near_, float two_near,
                                                      float neg_two_near, const float* __restrict__ ro,
                                                      const float* __restrict__ rd, int64_t n,
                                                      float* __restrict__ oo, float* __restrict__ od) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float ox = ro[i * 3], oy = ro[i * 3 + 1], oz = ro[i * 3 + 2];
    float dx = rd[i * 3], dy = rd[i * 3 + 1], dz = rd[i * 3 + 2];
    float t = __fdiv_rn(-__fadd_rn(near_, oz), dz);
    float px = __fadd_rn(ox, __fmul_rn(t, dx));
    float py = __fadd_rn(oy, __fmul_rn(t, dy));
    float pz = __fadd_rn(oz, __fmul_rn(t, dz));
    oo[i * 3 + 0] = __fdiv_rn(__fmul_rn(sx, px), pz);
    oo[i * 3 + 1] = __fdiv_rn(__fmul_rn(sy, py), pz);
    oo[i * 3 + 2] = __fadd_rn(1.0f, __fdiv_rn(two_near, pz));
    od[i * 3 + 0] = __fmul_rn(sx, __fsub_rn(__fdiv_rn(dx, dz), __fdiv_rn(px, pz)));
    od[i * 3 + 1] = __fmul_rn(sy, __fsub_rn(__fdiv_rn(dy, dz), __fdiv_rn(py, pz)));
    od[i * 3 + 2] = __fdiv_rn(neg_two_near, pz);
  }
}

__global__ void __launch_bounds__(256) ndc_bwd_kernel(float sx, float sy, float near_, const float* __restrict__ ro,
                                                      const float* __restrict__ rd, const float* __restrict__ gO,
                                                      const float* __restrict__ gD, int64_t n,
                                                      float* __restrict__ g_ro, float* __restrict__ g_rd) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float ox = ro[i * 3], oy = ro[i * 3 + 1], oz = ro[i * 3 + 2];
    float dx = rd[i * 3], dy = rd[i * 3 + 1], dz = rd[i * 3 + 2];
    float t = -(near_ + oz) / dz;
    float px = ox + t * dx, py = oy + t * dy, pz = oz + t * dz;
    float gO0 = gO ? gO[i * 3] : 0.f, gO1 = gO ? gO[i * 3 + 1] : 0.f, gO2 = gO ? gO[i * 3 + 2] : 0.f;
    float gD0 = gD ? gD[i * 3] : 0.f, gD1 = gD ? gD[i * 3 + 1] : 0.f, gD2 = gD ? gD[i * 3 + 2] : 0.f;
    float ipz = 1.0f / pz, idz = 1.0f / dz;
    float a0 = (gO0 - gD0) * sx, a1 = (gO1 - gD1) * sy;  // d/d(px/pz), d/d(py/pz)
    float gpx = a0 * ipz, gpy = a1 * ipz;
    float gpz = -(a0 * px + a1 * py) * ipz * ipz + (gD2 - gO2) * 2.0f * near_ * ipz * ipz;
    float gdx = gD0 * sx * idz, gdy = gD1 * sy * idz;
    float gdz = -(gD0 * sx * dx + gD1 * sy * dy) * idz * idz;
    float gt = gpx * dx + gpy * dy + gpz * dz;
    gdx += t * gpx; gdy += t * gpy; gdz += t * gpz;
    float gox = gpx, goy = gpy, goz = gpz - gt * idz;
    gdz += gt * (near_ + oz) * idz * idz;
    g_ro[i * 3] = gox; g_ro[i * 3 + 1] = goy; g_ro[i * 3 + 2] = goz;
    g_rd[i * 3] = gdx; g_rd[i * 3 + 1] = gdy; g_rd[i * 3 + 2] = gdz;
  }
}

// ---- stratified sampling: render_utils.py:105-131 ------------------------------------------
__device__ __forceinline__ float z_at(float near_, float far_, float t, int lindisp) {
  float omt = __fsub_rn(1.0f, t);
  if (!lindisp) return __fadd_rn(__fmul_rn(near_, omt), __fmul_rn(far_, t));
  float a = __fmul_rn(__fdiv_rn(1.0f, near_), omt);
  float b = __fmul_rn(__fdiv_rn(1.0f, far_), t);
  return __fdiv_rn(1.0f, __fadd_rn(a, b));
}

// One thread per sample (plenty of independent loads in flight); IDX = uint32_t whenever n*S fits, so the
// sample -> (ray, s) split is a 32-bit division.
template <typename IDX>
__global__ void __launch_bounds__(256) sample_coarse_kernel(const float* __restrict__ rays, int64_t n, int ld,
                                                            const float* __restrict__ t_vals,
                                                            const float* __restrict__ t_rand, int S, int lindisp,
                                                            float* __restrict__ z_out, float* __restrict__ pts) {
  const IDX total = (IDX)(n * S);
  const IDX stride = (IDX)gridDim.x * blockDim.x;
  for (IDX i = (IDX)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const IDX r = i / (IDX)S;
    const int s = (int)(i - r * (IDX)S);
    const float* ray = rays + (int64_t)r * ld;
    const float tr = t_rand ? __ldcs(t_rand + i) : 0.f;
    float near_ = ray[6], far_ = ray[7];
    float z = z_at(near_, far_, t_vals[s], lindisp);
    if (t_rand) {
      float lower = z, upper = z;
      if (s > 0) lower = __fmul_rn(0.5f, __fadd_rn(z, z_at(near_, far_, t_vals[s - 1], lindisp)));
      if (s < S - 1) upper = __fmul_rn(0.5f, __fadd_rn(z_at(near_, far_, t_vals[s + 1], lindisp), z));
      z = __fadd_rn(lower, __fmul_rn(__fsub_rn(upper, lower), tr));
    }
    __stcs(z_out + i, z);
    if (pts) {
      float* p = pts + (int64_t)i * 3;
      p[0] = __fadd_rn(ray[0], __fmul_rn(ray[3], z));
      p[1] = __fadd_rn(ray[1], __fmul_rn(ray[4], z));
      p[2] = __fadd_rn(ray[2], __fmul_rn(ray[5], z));
    }
  }
}

// A thread that owns 4 consecutive samples holds 12 consecutive floats of pts; written straight from registers the
// three 16-byte stores of a warp each touch all 48 sectors of the warp's 1,536 bytes (lane stride 48 B).  Staged
// through shared memory every store instruction covers 512 contiguous bytes (a third of the L2 write requests).
__device__ __forceinline__ void store_points_warp(float* __restrict__ pts, int64_t warp_q0, int valid_quads,
                                                  const float (&p)[12], float4* sm) {
  const int lane = threadIdx.x & 31;
#pragma unroll
  for (int k = 0; k < 3; ++k) sm[lane * 3 + k] = make_float4(p[4 * k], p[4 * k + 1], p[4 * k + 2], p[4 * k + 3]);
  __syncwarp();
  float4* out = reinterpret_cast<float4*>(pts) + warp_q0 * 3;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const int idx = k * 32 + lane;
    if (idx < 3 * valid_quads) __stcs(out + idx, sm[idx]);
  }
}

// Four consecutive samples of one ray per thread (S % 4 == 0): one 16-byte load of the jitter, one 16-byte store
// of z (three more for the points), six instead of twelve z_at evaluations.  Same arithmetic, bit for bit.
template <typename IDX>
__global__ void __launch_bounds__(256) sample_coarse_vec4_kernel(const float* __restrict__ rays, int64_t n, int ld,
                                                                 const float* __restrict__ t_vals,
                                                                 const float* __restrict__ t_rand, int S, int lindisp,
                                                                 float* __restrict__ z_out, float* __restrict__ pts) {
  __shared__ float4 stage[8][96];
  const int qpr = S >> 2;
  const IDX total = (IDX)(n * qpr);
  const IDX q_raw = (IDX)blockIdx.x * blockDim.x + threadIdx.x;   // quad index
  const bool live = q_raw < total;
  const IDX q = live ? q_raw : total - 1;
  const IDX r = q / (IDX)qpr;
  const int s0 = (int)(q - r * (IDX)qpr) << 2;
  float4 tr = make_float4(0.f, 0.f, 0.f, 0.f);
  if (t_rand) tr = __ldcs(reinterpret_cast<const float4*>(t_rand) + q);
  const float* ray = rays + (int64_t)r * ld;
  const float near_ = ray[6], far_ = ray[7];
  float zs[6];   // z_at(s0 - 1 .. s0 + 4)
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    const int s = min(max(s0 - 1 + k, 0), S - 1);
    zs[k] = z_at(near_, far_, t_vals[s], lindisp);
  }
  float zo[4];
  const float trs[4] = {tr.x, tr.y, tr.z, tr.w};
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    float zv = zs[k + 1];
    if (t_rand) {
      const int s = s0 + k;
      const float lower = s > 0 ? __fmul_rn(0.5f, __fadd_rn(zv, zs[k])) : zv;
      const float upper = s < S - 1 ? __fmul_rn(0.5f, __fadd_rn(zs[k + 2], zv)) : zv;
      zv = __fadd_rn(lower, __fmul_rn(__fsub_rn(upper, lower), trs[k]));
    }
    zo[k] = zv;
  }
  if (live) __stcs(reinterpret_cast<float4*>(z_out) + q, make_float4(zo[0], zo[1], zo[2], zo[3]));
  if (pts) {
    float p[12];
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
      for (int c = 0; c < 3; ++c) p[k * 3 + c] = __fadd_rn(ray[c], __fmul_rn(ray[3 + c], zo[k]));
    const IDX warp_q0 = q_raw - (threadIdx.x & 31);
    const int valid = warp_q0 < total ? (int)min((IDX)32, total - warp_q0) : 0;
    store_points_warp(pts, (int64_t)warp_q0, valid, p, stage[threadIdx.x >> 5]);
  }
}

// z only (no points wanted): two quads per thread, both 16-byte jitter loads issued before any arithmetic -- the
// one-quad form has a single load in flight per thread and sits at 0.69 of HBM, latency-bound.  Same arithmetic.
template <typename IDX>
__global__ void __launch_bounds__(256) sample_coarse_z_kernel(const float* __restrict__ rays, int64_t n, int ld,
                                                              const float* __restrict__ t_vals,
                                                              const float* __restrict__ t_rand, int S, int lindisp,
                                                              float* __restrict__ z_out) {
  const int qpr = S >> 2;
  const IDX total = (IDX)(n * qpr);
  IDX qs[2];
  bool live[2];
  float4 tr[2];
#pragma unroll
  for (int rep = 0; rep < 2; ++rep) {
    const IDX q_raw = ((IDX)blockIdx.x * 2 + rep) * blockDim.x + threadIdx.x;
    live[rep] = q_raw < total;
    qs[rep] = live[rep] ? q_raw : total - 1;
    tr[rep] = t_rand ? __ldcs(reinterpret_cast<const float4*>(t_rand) + qs[rep]) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
#pragma unroll
  for (int rep = 0; rep < 2; ++rep) {
    const IDX q = qs[rep];
    const IDX r = q / (IDX)qpr;
    const int s0 = (int)(q - r * (IDX)qpr) << 2;
    const float* ray = rays + (int64_t)r * ld;
    const float near_ = ray[6], far_ = ray[7];
    float zs[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      const int s = min(max(s0 - 1 + k, 0), S - 1);
      zs[k] = z_at(near_, far_, t_vals[s], lindisp);
    }
    float zo[4];
    const float trs[4] = {tr[rep].x, tr[rep].y, tr[rep].z, tr[rep].w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float zv = zs[k + 1];
      if (t_rand) {
        const int s = s0 + k;
        const float lower = s > 0 ? __fmul_rn(0.5f, __fadd_rn(zv, zs[k])) : zv;
        const float upper = s < S - 1 ? __fmul_rn(0.5f, __fadd_rn(zs[k + 2], zv)) : zv;
        zv = __fadd_rn(lower, __fmul_rn(__fsub_rn(upper, lower), trs[k]));
      }
      zo[k] = zv;
    }
    if (live[rep]) __stcs(reinterpret_cast<float4*>(z_out) + q, make_float4(zo[0], zo[1], zo[2], zo[3]));
  }
}

template <typename IDX>
__global__ void __launch_bounds__(256) ray_points_fwd_vec4_kernel(const float* __restrict__ ro,
                                                                  const float* __restrict__ rd,
                                                                  const float* __restrict__ z, int64_t n, int S,
                                                                  float* __restrict__ pts) {
  __shared__ float4 stage[8][96];
  const int qpr = S >> 2;
  const IDX total = (IDX)(n * qpr);
  const IDX q_raw = (IDX)blockIdx.x * blockDim.x + threadIdx.x;
  const IDX q = q_raw < total ? q_raw : total - 1;
  const IDX r = q / (IDX)qpr;
  const float4 zz = __ldcs(reinterpret_cast<const float4*>(z) + q);
  const float* op = ro + (int64_t)r * 3;
  const float* dp = rd + (int64_t)r * 3;
  const float o[3] = {op[0], op[1], op[2]}, d[3] = {dp[0], dp[1], dp[2]};
  const float zv[4] = {zz.x, zz.y, zz.z, zz.w};
  float p[12];
#pragma unroll
  for (int k = 0; k < 4; ++k)
#pragma unroll
    for (int c = 0; c < 3; ++c) p[k * 3 + c] = __fadd_rn(o[c], __fmul_rn(d[c], zv[k]));
  const IDX warp_q0 = q_raw - (threadIdx.x & 31);
  const int valid = warp_q0 < total ? (int)min((IDX)32, total - warp_q0) : 0;
  store_points_warp(pts, (int64_t)warp_q0, valid, p, stage[threadIdx.x >> 5]);
}

__global__ void __launch_bounds__(256) ray_points_fwd_kernel(const float* __restrict__ ro,
                                                             const float* __restrict__ rd,
                                                             const float* __restrict__ z, int64_t n, int S,
                                                             float* __restrict__ pts) {
  int64_t total = n * S;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = i / S;
    float zz = z[i];
    pts[i * 3 + 0] = __fadd_rn(ro[r * 3 + 0], __fmul_rn(rd[r * 3 + 0], zz));
    pts[i * 3 + 1] = __fadd_rn(ro[r * 3 + 1], __fmul_rn(rd[r * 3 + 1], zz));
    pts[i * 3 + 2] = __fadd_rn(ro[r * 3 + 2], __fmul_rn(rd[r * 3 + 2], zz));
  }
}

// one warp per ray: d_o = sum_s g, d_d = sum_s z*g
__global__ void __launch_bounds__(256) ray_points_bwd_kernel(const float* __restrict__ z,
                                                             const float* __restrict__ g, int64_t n, int S,
                                                             float* __restrict__ d_o, float* __restrict__ d_d) {
  int lane = threadIdx.x & 31;
  int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp; r < n; r += nwarps) {
    float so[3] = {0.f, 0.f, 0.f}, sd[3] = {0.f, 0.f, 0.f};
    for (int s = lane; s < S; s += 32) {
      int64_t i = r * S + s;
      float zz = z[i];
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        float gv = g[i * 3 + k];
        so[k] += gv;
        sd[k] += zz * gv;
      }
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      so[k] = warp_sum(so[k]);
      sd[k] = warp_sum(sd[k]);
    }
    if (lane == 0) {
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        if (d_o) d_o[r * 3 + k] = so[k];
        if (d_d) d_d[r * 3 + k] = sd[k];
      }
    }
  }
}

}  // namespace nb

using namespace nb;

extern "C" {

int nb_get_rays_fwd(int H, int W, float fx, float fy, float cx, float cy, const float* c2w, int c2w_ld,
                    float* rays_o, float* rays_d, float* viewdirs, void* stream) {
  return nb_get_rays_sel_fwd(H, W, fx, fy, cx, cy, c2w, c2w_ld, nullptr, (int64_t)H * W, rays_o, rays_d, viewdirs,
                             stream);
}

int nb_get_rays_sel_fwd(int H, int W, float fx, float fy, float cx, float cy, const float* c2w, int c2w_ld,
                        const int64_t* pix, int64_t n, float* rays_o, float* rays_d, float* viewdirs, void* stream) {
  if (n == 0) return 0;
  NB_REQUIRE(H > 0 && W > 0 && c2w && rays_o && rays_d && c2w_ld >= 4, "nb_get_rays: bad arguments");
  get_rays_kernel<<<grid_for(n, 256), 256, 0, as_stream(stream)>>>(H, W, fx, fy, cx, cy, c2w, c2w_ld, pix, n, rays_o,
                                                                    rays_d, viewdirs);
  return check_launch("nb_get_rays");
}

int nb_get_rays_bwd(int H, int W, float fx, float fy, float cx, float cy, const int64_t* pix, int64_t n,
                    const float* d_rays_o, const float* d_rays_d, float* d_c2w, void* stream) {
  if (n == 0 || (!d_rays_o && !d_rays_d)) return 0;
  NB_REQUIRE(H > 0 && W > 0 && d_c2w, "nb_get_rays_bwd: bad arguments");
  get_rays_bwd_kernel<<<grid_for(n, 256 * 8, 4), 256, 0, as_stream(stream)>>>(H, W, fx, fy, cx, cy, pix, n, d_rays_o,
                                                                               d_rays_d, d_c2w);
  return check_launch("nb_get_rays_bwd");
}

int nb_normalize_fwd(const float* v, int64_t n, float* out, void* stream) {
  if (n == 0) return 0;
  NB_REQUIRE(v && out, "nb_normalize_fwd: null pointer");
  normalize_fwd_kernel<<<grid_for(n, 256), 256, 0, as_stream(stream)>>>(v, n, out);
  return check_launch("nb_normalize_fwd");
}

int nb_ray_rows(const float* rays_o, const float* rays_d, const float* view_src, float near_, float far_, int64_t n,
                float* rows, float* viewdirs, void* stream) {
  if (n == 0) return 0;
  NB_REQUIRE(rays_o && rays_d && view_src && rows, "nb_ray_rows: null pointer");
  NB_REQUIRE((n + 255) / 256 < (1ll << 31), "nb_ray_rows: too many rays for one launch");
  ray_rows_kernel<<<(unsigned)((n + 255) / 256), 256, 0, as_stream(stream)>>>(rays_o, rays_d, view_src, near_, far_, n, rows,
                                                                              viewdirs);
  return check_launch("nb_ray_rows");
}

int nb_normalize_bwd(const float* v, const float* d_out, int64_t n, float* d_v, void* stream) {
  if (n == 0) return 0;
  NB_REQUIRE(v && d_out && d_v, "nb_normalize_bwd: null pointer");
  normalize_bwd_kernel<<<grid_for(n, 256), 256, 0, as_stream(stream)>>>(v, d_out, n, d_v);
  return check_launch("nb_normalize_bwd");
}

int nb_ndc_rays_fwd(int H, int W, double focal, double near_, const float* rays_o, const float* rays_d, int64_t n,
                    float* out_o, float* out_d, void* stream) {
  if (n == 0) return 0;
  NB_REQUIRE(rays_o && rays_d && out_o && out_d, "nb_ndc_rays_fwd: null pointer");
  // the Python-scalar sub-expressions of utils.py:60-66 are evaluated in double, then enter fp32
  float sx = (float)(-1.0 / (W / (2.0 * focal)));
  float sy = (float)(-1.0 / (H / (2.0 * focal)));
  ndc_fwd_kernel<<<grid_for(n, 256), 256, 0, as_stream(stream)>>>(sx, sy, (float)near_, (float)(2.0 * near_),
                                                                   (float)(-2.0 * near_), rays_o, rays_d, n, out_o,
                                                                   out_d);
  return check_launch("nb_ndc_rays_fwd");
}

int nb_ndc_rays_bwd(int H, int W, double focal, double near_, const float* rays_o, const float* rays_d,
                    const float* d_out_o, const float* d_out_d, int64_t n, float* d_rays_o, float* d_rays_d,
                    void* stream) {
  if (n == 0) return 0;
  NB_REQUIRE(rays_o && rays_d && d_rays_o && d_rays_d, "nb_ndc_rays_bwd: null pointer");
  float sx = (float)(-1.0 / (W / (2.0 * focal)));
  float sy = (float)(-1.0 / (H / (2.0 * focal)));
  ndc_bwd_kernel<<<grid_for(n, 256), 256, 0, as_stream(stream)>>>(sx, sy, (float)near_, rays_o, rays_d, d_out_o,
                                                                   d_out_d, n, d_rays_o, d_rays_d);
  return check_launch("nb_ndc_rays_bwd");
}

int nb_sample_coarse(const float* rays, int64_t n, int ld, const float* t_vals, const float* t_rand, int S,
                     int lindisp, float* z, float* pts, void* stream) {
  if (n == 0) return 0;
  NB_REQUIRE(rays && t_vals && z && ld >= 8 && S >= 1, "nb_sample_coarse: bad arguments");
  const int64_t quads = n * (S >> 2);
  if ((S & 3) == 0 && (quads + 255) / 256 < (1ll << 31) && ((uintptr_t)z & 15) == 0 && ((uintptr_t)t_rand & 15) == 0 &&
      ((uintptr_t)pts & 15) == 0)
  {
    if (!pts && quads < (1ll << 32) - 512)
      sample_coarse_z_kernel<uint32_t><<<(unsigned)((quads + 511) / 512), 256, 0, as_stream(stream)>>>(rays, n, ld, t_vals, t_rand,
                                                                                                      S, lindisp, z);
    else if (quads < (1ll << 32) - 256)
      sample_coarse_vec4_kernel<uint32_t><<<(unsigned)((quads + 255) / 256), 256, 0, as_stream(stream)>>>(
          rays, n, ld, t_vals, t_rand, S, lindisp, z, pts);
    else
      sample_coarse_vec4_kernel<int64_t><<<(unsigned)((quads + 255) / 256), 256, 0, as_stream(stream)>>>(
          rays, n, ld, t_vals, t_rand, S, lindisp, z, pts);
  } else if (n * S < (1ll << 31))
    sample_coarse_kernel<uint32_t><<<grid_for(n * S, 256), 256, 0, as_stream(stream)>>>(rays, n, ld, t_vals, t_rand, S,
                                                                                        lindisp, z, pts);
  else
    sample_coarse_kernel<int64_t><<<grid_for(n * S, 256), 256, 0, as_stream(stream)>>>(rays, n, ld, t_vals, t_rand, S,
                                                                                       lindisp, z, pts);
  return check_launch("nb_sample_coarse");
}

int nb_ray_points_fwd(const float* rays_o, const float* rays_d, const float* z, int64_t n, int S, float* pts,
                      void* stream) {
  if (n == 0) return 0;
  NB_REQUIRE(rays_o && rays_d && z && pts && S >= 1, "nb_ray_points_fwd: bad arguments");
  const int64_t quads = n * (S >> 2);
  if ((S & 3) == 0 && (quads + 255) / 256 < (1ll << 31) && ((uintptr_t)z & 15) == 0 && ((uintptr_t)pts & 15) == 0)
  {
    if (quads < (1ll << 32) - 256)
      ray_points_fwd_vec4_kernel<uint32_t><<<(unsigned)((quads + 255) / 256), 256, 0, as_stream(stream)>>>(rays_o, rays_d, z, n, S, pts);
    else
      ray_points_fwd_vec4_kernel<int64_t><<<(unsigned)((quads + 255) / 256), 256, 0, as_stream(stream)>>>(rays_o, rays_d, z, n, S, pts);
  } else
    ray_points_fwd_kernel<<<grid_for(n * S, 256), 256, 0, as_stream(stream)>>>(rays_o, rays_d, z, n, S, pts);
  return check_launch("nb_ray_points_fwd");
}

int nb_ray_points_bwd(const float* z, const float* d_pts, int64_t n, int S, float* d_rays_o, float* d_rays_d,
                      void* stream) {
  if (n == 0) return 0;
  NB_REQUIRE(z && d_pts && S >= 1, "nb_ray_points_bwd: bad arguments");
  ray_points_bwd_kernel<<<grid_for(n, 8), 256, 0, as_stream(stream)>>>(z, d_pts, n, S, d_rays_o, d_rays_d);
  return check_launch("nb_ray_points_bwd");
}

}  // extern "C"
