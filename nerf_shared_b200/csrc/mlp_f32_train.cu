// fp32 training path of the NeRF MLP: layer-by-layer CUDA-core GEMMs with saved fp32 activations.
//
// replaces NeRF.forward / MLP and its autograd (nerf.py:96-134, main.py:103) for the model variants the
// tensor-core kernels do not tile: encodings wider than 63 / 27 columns (configs/stonehenge.txt: multires 15 / 6
// -> 93 / 39) and use_viewdirs=False (nerf.py:93-94,131-132: trunk -> output_linear).  It is also the exact-operand
// reference of the bf16 kernels' gradients on the device.  Not the fast path: fp32 FMAs on the CUDA cores.
//
// Saved activations per sample (floats): enc(pts) | h0 .. h7 (256 each) | [feature 256 | enc(dirs) | hv 128].
#include "common.cuh"

namespace nb {

// ------------------------------------------------------------------------------------------
// positional encoding (nerf.py:32-41): out[m] = [x, sin(2^0 x), cos(2^0 x), ...], ld floats per row
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) encode_rows_kernel(const float* __restrict__ x, int64_t M, int div, int L,
                                                          float* __restrict__ out, int ld) {
  for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < M * 3; t += (int64_t)gridDim.x * blockDim.x) {
    const int64_t m = t / 3;
    const int c = (int)(t - m * 3);
    const float v = x[(m / div) * 3 + c];
    float* e = out + m * ld;
    e[c] = v;
    float f = 1.0f;
    for (int l = 0; l < L; ++l, f *= 2.0f) {
      const float a = v * f;
      e[3 + 6 * l + c] = sinf(a);
      e[3 + 6 * l + 3 + c] = cosf(a);
    }
  }
}

// ------------------------------------------------------------------------------------------
// One tiled fp32 GEMM for the three shapes of a linear layer.  C[i][j] (+)= sum_r A(i,r) * B(r,j)
//   A(i,r) = a[i * sa_i + r * sa_r],  B(r,j) = b[r * sb_r + j * sb_j],  C row-major with leading dimension ldc.
// Epilogue: + bias[j], += old C (ACC), relu (RELU), * (mask[i * ldm + j] > 0) (the relu' of a saved activation).
// SPLIT > 1: the reduction range is cut over blockIdx.z and the partial results are added atomically (weight
// gradients: the reduction runs over the samples).
// ------------------------------------------------------------------------------------------
constexpr int kBM = 64, kBN = 64, kBK = 16;

struct GemmArgs {
  const float* a; int64_t sa_i, sa_r;
  const float* b; int64_t sb_r, sb_j;
  float* c; int64_t ldc;
  const float* bias;
  const float* mask; int64_t ldm;
  int64_t I, J, R;
  int acc, relu, atomic;
};

__global__ void __launch_bounds__(256) gemm_f32_kernel(GemmArgs g) {
  __shared__ float As[kBK][kBM + 4];
  __shared__ float Bs[kBK][kBN + 4];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;      // 16 x 16 threads, 4 x 4 outputs each
  const int64_t i0 = (int64_t)blockIdx.y * kBM, j0 = (int64_t)blockIdx.x * kBN;
  const int64_t rchunk = (g.R + gridDim.z - 1) / gridDim.z;
  const int64_t r_begin = (int64_t)blockIdx.z * rchunk, r_end = min(g.R, r_begin + rchunk);
  float acc[4][4];
#pragma unroll
  for (int p = 0; p < 4; ++p)
#pragma unroll
    for (int q = 0; q < 4; ++q) acc[p][q] = 0.f;
  for (int64_t r0 = r_begin; r0 < r_end; r0 += kBK) {
    // 256 threads load 64 x 16 of A and 16 x 64 of B (4 elements each); the faster-varying thread index follows the
    // unit-stride dimension of the operand
    for (int e = threadIdx.x; e < kBM * kBK; e += 256) {
      int ii, rr;
      if (g.sa_r == 1) { rr = e % kBK; ii = e / kBK; } else { ii = e % kBM; rr = e / kBM; }
      const int64_t i = i0 + ii, r = r0 + rr;
      As[rr][ii] = (i < g.I && r < r_end) ? g.a[i * g.sa_i + r * g.sa_r] : 0.f;
    }
    for (int e = threadIdx.x; e < kBK * kBN; e += 256) {
      int jj, rr;
      if (g.sb_r == 1) { rr = e % kBK; jj = e / kBK; } else { jj = e % kBN; rr = e / kBN; }
      const int64_t j = j0 + jj, r = r0 + rr;
      Bs[rr][jj] = (j < g.J && r < r_end) ? g.b[r * g.sb_r + j * g.sb_j] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int rr = 0; rr < kBK; ++rr) {
      float av[4], bv[4];
#pragma unroll
      for (int p = 0; p < 4; ++p) av[p] = As[rr][ty * 4 + p];
#pragma unroll
      for (int q = 0; q < 4; ++q) bv[q] = Bs[rr][tx * 4 + q];
#pragma unroll
      for (int p = 0; p < 4; ++p)
#pragma unroll
        for (int q = 0; q < 4; ++q) acc[p][q] = fmaf(av[p], bv[q], acc[p][q]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int p = 0; p < 4; ++p) {
    const int64_t i = i0 + ty * 4 + p;
    if (i >= g.I) continue;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int64_t j = j0 + tx * 4 + q;
      if (j >= g.J) continue;
      float v = acc[p][q];
      float* cp = g.c + i * g.ldc + j;
      if (g.atomic) {
        atomicAdd(cp, v);
        continue;
      }
      if (g.bias) v += g.bias[j];
      if (g.acc) v += *cp;
      if (g.relu) v = fmaxf(v, 0.f);
      if (g.mask) v = g.mask[i * g.ldm + j] > 0.f ? v : 0.f;
      *cp = v;
    }
  }
}

// db[j] += sum_i y[i * ld + j]
__global__ void __launch_bounds__(256) colsum_kernel(const float* __restrict__ y, int64_t I, int J, int64_t ld,
                                                     float* __restrict__ db) {
  const int j = blockIdx.x * 32 + (threadIdx.x & 31);
  const int64_t rows_per = (I + gridDim.y - 1) / gridDim.y;
  const int64_t i_begin = (int64_t)blockIdx.y * rows_per, i_end = min(I, i_begin + rows_per);
  float s = 0.f;
  if (j < J)
    for (int64_t i = i_begin + (threadIdx.x >> 5); i < i_end; i += 8) s += y[i * ld + j];
  __shared__ float red[8][32];
  red[threadIdx.x >> 5][threadIdx.x & 31] = s;
  __syncthreads();
  if (threadIdx.x < 32 && j < J) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += red[w][threadIdx.x];
    atomicAdd(db + j, t);
  }
}

static int launch_gemm(const GemmArgs& g, int split, cudaStream_t st) {
  dim3 grid((unsigned)((g.J + kBN - 1) / kBN), (unsigned)((g.I + kBM - 1) / kBM), (unsigned)split);
  gemm_f32_kernel<<<grid, 256, 0, st>>>(g);
  return check_launch("gemm_f32_kernel");
}

// Y[M, N] (ld ldy) = act( X[M, K] (ld ldx) . W[N, K]^T (ld ldw) + b (+ Y) )
static int linear_fwd(const float* X, int64_t ldx, const float* W, int64_t ldw, const float* b, float* Y, int64_t ldy,
                      int64_t M, int N, int K, bool acc, bool relu, cudaStream_t st) {
  GemmArgs g{X, ldx, 1, W, 1, ldw, Y, ldy, b, nullptr, 0, M, N, K, acc ? 1 : 0, relu ? 1 : 0, 0};
  return launch_gemm(g, 1, st);
}
// dX[M, K] (ld lddx) = ( dY[M, N] (ld ldy) . W[N, K] (ld ldw) (+ dX) ) * (mask > 0)
static int linear_dgrad(const float* dY, int64_t ldy, const float* W, int64_t ldw, float* dX, int64_t lddx, int64_t M, int N,
                        int K, bool acc, const float* mask, int64_t ldm, cudaStream_t st) {
  GemmArgs g{dY, ldy, 1, W, ldw, 1, dX, lddx, nullptr, mask, ldm, M, K, N, acc ? 1 : 0, 0, 0};
  return launch_gemm(g, 1, st);
}
// dW[N, K] (ld ldw) += dY[M, N]^T . X[M, K] ; db[N] += column sums of dY
static int linear_wgrad(const float* dY, int64_t ldy, const float* X, int64_t ldx, float* dW, int64_t ldw, float* db,
                        int64_t M, int N, int K, cudaStream_t st) {
  int split = (int)((M + 4095) / 4096);
  if (split > 256) split = 256;
  if (split < 1) split = 1;
  GemmArgs g{dY, 1, ldy, X, ldx, 1, dW, ldw, nullptr, nullptr, 0, N, K, M, 0, 0, 1};
  int rc = launch_gemm(g, split, st);
  if (rc || !db) return rc;
  dim3 grid((unsigned)((N + 31) / 32), (unsigned)split);
  colsum_kernel<<<grid, 256, 0, st>>>(dY, M, N, ldy, db);
  return check_launch("colsum_kernel");
}

struct F32Layout {
  int in_ch, in_views, per_sample;
  int64_t enc, h[8], feat, encd, hv;   // float offsets of the [M, width] sections (multiply by nothing: already * M)
};
static F32Layout f32_layout(int L_pts, int L_dirs, bool noview, int64_t M) {
  F32Layout o;
  o.in_ch = 3 + 6 * L_pts;
  o.in_views = noview ? 0 : 3 + 6 * L_dirs;
  int64_t off = 0;
  o.enc = off; off += M * o.in_ch;
  for (int l = 0; l < 8; ++l) { o.h[l] = off; off += M * 256; }
  o.feat = off; if (!noview) off += M * 256;
  o.encd = off; if (!noview) off += M * o.in_views;
  o.hv = off; if (!noview) off += M * 128;
  o.per_sample = (int)(off / (M > 0 ? M : 1));
  return o;
}

}  // namespace nb

using namespace nb;

extern "C" {

int64_t nb_mlp_f32_act_floats(int L_pts, int L_dirs, int noview, int64_t M) {
  const int in_ch = 3 + 6 * L_pts, in_views = noview ? 0 : 3 + 6 * L_dirs;
  return M * (int64_t)(in_ch + 8 * 256 + (noview ? 0 : 256 + in_views + 128));
}

#define NB_RC(x) do { int rc_ = (x); if (rc_) return rc_; } while (0)

// Forward with saved activations.  out_w / out_b / out_ch: the output_linear of a use_viewdirs=False model (then
// viewdirs is ignored and raw is [M, out_ch]); NULL / 0 for the view-dependent model (raw [M, 4]).
int nb_mlp_f32_fwd_train(const NbMlpWeights* w, const float* out_w, const float* out_b, int out_ch, const float* pts,
                         const float* viewdirs, int64_t M, int S, float* raw, float* acts, void* stream) {
  if (M == 0) return 0;
  const bool noview = out_w != nullptr;
  NB_REQUIRE(w && pts && raw && acts && S >= 1 && (noview ? (out_b && out_ch >= 1 && out_ch <= 64) : viewdirs != nullptr),
             "nb_mlp_f32_fwd_train: bad arguments");
  NB_REQUIRE(w->L_pts >= 0 && w->L_pts <= 15 && w->L_dirs >= 0 && w->L_dirs <= 6, "nb_mlp_f32_fwd_train: encoding octaves (%d,%d) outside [0,15]x[0,6]",
             w->L_pts, w->L_dirs);
  cudaStream_t st = as_stream(stream);
  const F32Layout L = f32_layout(w->L_pts, w->L_dirs, noview, M);
  float* enc = acts + L.enc;
  encode_rows_kernel<<<grid_for(M * 3, 256), 256, 0, st>>>(pts, M, 1, w->L_pts, enc, L.in_ch);
  NB_RC(check_launch("encode_rows_kernel"));
  const int ic = L.in_ch;
  NB_RC(linear_fwd(enc, ic, w->pts_w[0], ic, w->pts_b[0], acts + L.h[0], 256, M, 256, ic, false, true, st));
  for (int l = 1; l < 8; ++l) {
    if (l == 5) {   // cat[input_pts, h] (nerf.py:117-118): two partial products, the relu after the second
      NB_RC(linear_fwd(enc, ic, w->pts_w[5], ic + 256, w->pts_b[5], acts + L.h[5], 256, M, 256, ic, false, false, st));
      NB_RC(linear_fwd(acts + L.h[4], 256, w->pts_w[5] + ic, ic + 256, nullptr, acts + L.h[5], 256, M, 256, 256, true, true, st));
    } else {
      NB_RC(linear_fwd(acts + L.h[l - 1], 256, w->pts_w[l], 256, w->pts_b[l], acts + L.h[l], 256, M, 256, 256, false, true, st));
    }
  }
  const float* h7 = acts + L.h[7];
  if (noview) return linear_fwd(h7, 256, out_w, 256, out_b, raw, out_ch, M, out_ch, 256, false, false, st);
  float* encd = acts + L.encd;
  encode_rows_kernel<<<grid_for(M * 3, 256), 256, 0, st>>>(viewdirs, M, S, w->L_dirs, encd, L.in_views);
  NB_RC(check_launch("encode_rows_kernel"));
  const int iv = L.in_views;
  NB_RC(linear_fwd(h7, 256, w->alpha_w, 256, w->alpha_b, raw + 3, 4, M, 1, 256, false, false, st));                 // sigma
  NB_RC(linear_fwd(h7, 256, w->feature_w, 256, w->feature_b, acts + L.feat, 256, M, 256, 256, false, false, st));
  NB_RC(linear_fwd(acts + L.feat, 256, w->views_w, 256 + iv, w->views_b, acts + L.hv, 128, M, 128, 256, false, false, st));
  NB_RC(linear_fwd(encd, iv, w->views_w + 256, 256 + iv, nullptr, acts + L.hv, 128, M, 128, iv, true, true, st));
  return linear_fwd(acts + L.hv, 128, w->rgb_w, 128, w->rgb_b, raw, 4, M, 3, 128, false, false, st);                 // rgb
}

// Backward: weight / bias gradients ACCUMULATE into grads + off[i] (i in _ops.mlp_params order: 0..7 pts_w, 8..15 pts_b,
// 16 feature_w, 17 feature_b, 18 alpha_w, 19 alpha_b, 20 views_w, 21 views_b, 22 rgb_w, 23 rgb_b; for a
// use_viewdirs=False model 16 = output_linear.weight, 17 = output_linear.bias and 18..23 are ignored).
// scratch: 2 * M * 256 floats.
int nb_mlp_f32_bwd(const NbMlpWeights* w, const float* out_w, int out_ch, const float* acts, const float* d_raw, int64_t M,
                   const int64_t* off, float* grads, float* scratch, void* stream) {
  if (M == 0) return 0;
  const bool noview = out_w != nullptr;
  NB_REQUIRE(w && acts && d_raw && off && grads && scratch, "nb_mlp_f32_bwd: null pointer");
  cudaStream_t st = as_stream(stream);
  const F32Layout L = f32_layout(w->L_pts, w->L_dirs, noview, M);
  const int ic = L.in_ch, iv = L.in_views;
  const float* enc = acts + L.enc;
  const float* h7 = acts + L.h[7];
  float* ga = scratch;                 // gradient wrt the current layer's pre-activation, [M, 256]
  float* gb = scratch + M * 256;
  if (noview) {
    NB_RC(linear_wgrad(d_raw, out_ch, h7, 256, grads + off[16], 256, grads + off[17], M, out_ch, 256, st));
    NB_RC(linear_dgrad(d_raw, out_ch, out_w, 256, ga, 256, M, out_ch, 256, false, h7, 256, st));
  } else {
    const float* hv = acts + L.hv;
    const float* feat = acts + L.feat;
    const float* encd = acts + L.encd;
    // rgb_linear (nerf.py:129)
    NB_RC(linear_wgrad(d_raw, 4, hv, 128, grads + off[22], 128, grads + off[23], M, 3, 128, st));
    NB_RC(linear_dgrad(d_raw, 4, w->rgb_w, 128, gb, 128, M, 3, 128, false, hv, 128, st));                 // d(view hidden pre-act)
    // views_linears[0] on cat[feature, dirs] (nerf.py:123-127)
    NB_RC(linear_wgrad(gb, 128, feat, 256, grads + off[20], 256 + iv, grads + off[21], M, 128, 256, st));
    NB_RC(linear_wgrad(gb, 128, encd, iv, grads + off[20] + 256, 256 + iv, nullptr, M, 128, iv, st));
    NB_RC(linear_dgrad(gb, 128, w->views_w, 256 + iv, ga, 256, M, 128, 256, false, nullptr, 0, st));      // d feature (linear)
    // feature_linear + alpha_linear both read h7 (nerf.py:120-122)
    NB_RC(linear_wgrad(ga, 256, h7, 256, grads + off[16], 256, grads + off[17], M, 256, 256, st));
    NB_RC(linear_wgrad(d_raw + 3, 4, h7, 256, grads + off[18], 256, grads + off[19], M, 1, 256, st));
    NB_RC(linear_dgrad(ga, 256, w->feature_w, 256, gb, 256, M, 256, 256, false, nullptr, 0, st));
    NB_RC(linear_dgrad(d_raw + 3, 4, w->alpha_w, 256, gb, 256, M, 1, 256, true, h7, 256, st));            // + sigma term, * relu'(h7)
    float* t = ga; ga = gb; gb = t;
  }
  // trunk: ga = d(pre-activation of layer l)
  for (int l = 7; l >= 1; --l) {
    const float* hprev = acts + L.h[l - 1];
    if (l == 5) {
      NB_RC(linear_wgrad(ga, 256, enc, ic, grads + off[5], ic + 256, grads + off[13], M, 256, ic, st));
      NB_RC(linear_wgrad(ga, 256, hprev, 256, grads + off[5] + ic, ic + 256, nullptr, M, 256, 256, st));
      NB_RC(linear_dgrad(ga, 256, w->pts_w[5] + ic, ic + 256, gb, 256, M, 256, 256, false, hprev, 256, st));
    } else {
      NB_RC(linear_wgrad(ga, 256, hprev, 256, grads + off[l], 256, grads + off[8 + l], M, 256, 256, st));
      NB_RC(linear_dgrad(ga, 256, w->pts_w[l], 256, gb, 256, M, 256, 256, false, hprev, 256, st));
    }
    float* t = ga; ga = gb; gb = t;
  }
  return linear_wgrad(ga, 256, enc, ic, grads + off[0], ic, grads + off[8], M, 256, ic, st);
}

}  // extern "C"
