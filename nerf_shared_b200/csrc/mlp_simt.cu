// fp32 CUDA-core evaluation of the NeRF MLP with the positional encoding fused in.
// replaces NeRF.forward / MLP (nerf.py:96-134) with exact fp32 operands.
//
// This is the precise path (Renderer/NeRF `precision="fp32"`): fp32 weights straight from
// the nn.Linear parameters, fp32 activations in shared memory, fmaf accumulation.  It is the
// on-device numerical reference for the tcgen05 kernel (mlp_tc.cu) at sizes the CPU oracle
// cannot reach, and serves the one shipped config whose encoding widths (multires 15/6,
// configs/stonehenge.txt) the tensor-core kernel does not tile.  Not the fast path.
// The same kernel evaluates the use_viewdirs=False model (nerf.py:93-94,131-132: trunk ->
// output_linear), which no shipped config trains; forward only.
#include "common.cuh"

namespace nb {

constexpr int kTS = 64;        // samples per CTA tile
constexpr int kEncLd = 96;     // >= 3+6*15 = 93
constexpr int kDirLd = 40;     // >= 3+6*6 = 39
constexpr int kW = 256;

__device__ __forceinline__ void accum_segment(float (&acc)[kTS], const float* __restrict__ wrow,
                                              const float* __restrict__ in, int ld, int K) {
  int k = 0;
  for (; k + 4 <= K; k += 4) {
    float w0 = __ldg(wrow + k), w1 = __ldg(wrow + k + 1), w2 = __ldg(wrow + k + 2), w3 = __ldg(wrow + k + 3);
#pragma unroll
    for (int s = 0; s < kTS; ++s) {
      float4 a = *reinterpret_cast<const float4*>(in + s * ld + k);
      acc[s] = fmaf(a.x, w0, acc[s]);
      acc[s] = fmaf(a.y, w1, acc[s]);
      acc[s] = fmaf(a.z, w2, acc[s]);
      acc[s] = fmaf(a.w, w3, acc[s]);
    }
  }
  for (; k < K; ++k) {
    float w = __ldg(wrow + k);
#pragma unroll
    for (int s = 0; s < kTS; ++s) acc[s] = fmaf(in[s * ld + k], w, acc[s]);
  }
}

// out[s][n] = act(bias[n] + in1[s][:k1] . W[n][:k1] + in2[s][:k2] . W[n][k1:k1+k2])
__device__ __forceinline__ void dense(const float* __restrict__ W, const float* __restrict__ bias, int n_out,
                                      const float* in1, int ld1, int k1, const float* in2, int ld2, int k2,
                                      float* out, bool relu) {
  int n = threadIdx.x;
  if (n < n_out) {
    float acc[kTS];
    float b = __ldg(bias + n);
#pragma unroll
    for (int s = 0; s < kTS; ++s) acc[s] = b;
    const float* wrow = W + (size_t)n * (k1 + k2);
    accum_segment(acc, wrow, in1, ld1, k1);
    if (k2 > 0) accum_segment(acc, wrow + k1, in2, ld2, k2);
#pragma unroll
    for (int s = 0; s < kTS; ++s) out[s * kW + n] = relu ? fmaxf(acc[s], 0.f) : acc[s];
  }
  __syncthreads();
}

__global__ void __launch_bounds__(256, 1) mlp_fwd_f32_kernel(NbMlpWeights w, const float* __restrict__ pts,
                                                             const float* __restrict__ viewdirs, int64_t M, int S,
                                                             float* __restrict__ raw, const float* __restrict__ out_w,
                                                             const float* __restrict__ out_b, int out_ch) {
  extern __shared__ __align__(16) float sm[];
  float* enc = sm;                       // [kTS][kEncLd]
  float* dirs = enc + kTS * kEncLd;      // [kTS][kDirLd]
  float* hA = dirs + kTS * kDirLd;       // [kTS][256]
  float* hB = hA + kTS * kW;             // [kTS][256]
  float* alpha = hB + kTS * kW;          // [kTS]
  const int in_ch = 3 + 6 * w.L_pts, in_views = 3 + 6 * w.L_dirs;
  const int64_t ntiles = (M + kTS - 1) / kTS;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t m0 = tile * kTS;
    // positional encodings (nerf.py:32-41): thread -> (sample, coordinate)
    for (int t = threadIdx.x; t < kTS * 3; t += blockDim.x) {
      int s = t / 3, c = t - s * 3;
      int64_t m = m0 + s;
      float x = 0.f, v = 0.f;
      if (m < M) {
        x = pts[m * 3 + c];
        if (viewdirs != nullptr) v = viewdirs[(m / S) * 3 + c];
      }
      float* e = enc + s * kEncLd;
      e[c] = x;
      float f = 1.0f;
      for (int l = 0; l < w.L_pts; ++l, f *= 2.0f) {
        float a = x * f;
        e[3 + 6 * l + c] = sinf(a);
        e[3 + 6 * l + 3 + c] = cosf(a);
      }
      float* d = dirs + s * kDirLd;
      d[c] = v;
      f = 1.0f;
      for (int l = 0; l < w.L_dirs; ++l, f *= 2.0f) {
        float a = v * f;
        d[3 + 6 * l + c] = sinf(a);
        d[3 + 6 * l + 3 + c] = cosf(a);
      }
    }
    __syncthreads();
    dense(w.pts_w[0], w.pts_b[0], kW, enc, kEncLd, in_ch, nullptr, 0, 0, hA, true);
    dense(w.pts_w[1], w.pts_b[1], kW, hA, kW, kW, nullptr, 0, 0, hB, true);
    dense(w.pts_w[2], w.pts_b[2], kW, hB, kW, kW, nullptr, 0, 0, hA, true);
    dense(w.pts_w[3], w.pts_b[3], kW, hA, kW, kW, nullptr, 0, 0, hB, true);
    dense(w.pts_w[4], w.pts_b[4], kW, hB, kW, kW, nullptr, 0, 0, hA, true);
    dense(w.pts_w[5], w.pts_b[5], kW, enc, kEncLd, in_ch, hA, kW, kW, hB, true);  // skip: cat[input_pts, h]
    dense(w.pts_w[6], w.pts_b[6], kW, hB, kW, kW, nullptr, 0, 0, hA, true);
    dense(w.pts_w[7], w.pts_b[7], kW, hA, kW, kW, nullptr, 0, 0, hB, true);
    if (out_w != nullptr) {   // use_viewdirs=False: raw = output_linear(h) (nerf.py:131-132); thread -> (sample, channel)
      int s = threadIdx.x >> 2;
      int64_t m = m0 + s;
      for (int o = threadIdx.x & 3; o < out_ch; o += 4) {
        float a = __ldg(out_b + o);
        for (int k = 0; k < kW; ++k) a = fmaf(hB[s * kW + k], __ldg(out_w + o * kW + k), a);
        if (m < M) raw[m * out_ch + o] = a;
      }
      __syncthreads();
      continue;
    }
    // sigma head (nerf.py:121): 4 threads per sample
    {
      int s = threadIdx.x >> 2, part = threadIdx.x & 3;
      float a = 0.f;
      for (int k = part; k < kW; k += 4) a = fmaf(hB[s * kW + k], __ldg(w.alpha_w + k), a);
      a += __shfl_xor_sync(0xffffffffu, a, 1);
      a += __shfl_xor_sync(0xffffffffu, a, 2);
      if (part == 0) alpha[s] = a + __ldg(w.alpha_b);
    }
    dense(w.feature_w, w.feature_b, kW, hB, kW, kW, nullptr, 0, 0, hA, false);
    dense(w.views_w, w.views_b, kW / 2, hA, kW, kW, dirs, kDirLd, in_views, hB, true);  // cat[feature, dirs]
    // rgb head (nerf.py:129): thread -> (sample, channel); channel 3 carries sigma
    {
      int s = threadIdx.x >> 2, o = threadIdx.x & 3;
      float a;
      if (o < 3) {
        a = __ldg(w.rgb_b + o);
        for (int k = 0; k < kW / 2; ++k) a = fmaf(hB[s * kW + k], __ldg(w.rgb_w + o * (kW / 2) + k), a);
      } else {
        a = alpha[s];
      }
      int64_t m = m0 + s;
      if (m < M) raw[m * 4 + o] = a;
    }
    __syncthreads();
  }
}

constexpr size_t kSimtSmem = (size_t)(kTS * kEncLd + kTS * kDirLd + 2 * kTS * kW + kTS) * sizeof(float);

int init_mlp_simt(int) {
  cudaError_t e = cudaFuncSetAttribute(mlp_fwd_f32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSimtSmem);
  if (e != cudaSuccess) {
    set_error("init_mlp_simt: %s", cudaGetErrorString(e));
    return (int)e;
  }
  return 0;
}

}  // namespace nb

using namespace nb;

extern "C" int nb_mlp_fwd_f32(const NbMlpWeights* w, const float* pts, const float* viewdirs, int64_t M, int S,
                              float* raw, void* stream) {
  if (M == 0) return 0;
  NB_REQUIRE(w && pts && viewdirs && raw && S >= 1, "nb_mlp_fwd_f32: bad arguments");
  NB_REQUIRE(M % S == 0, "nb_mlp_fwd_f32: M=%lld is not a multiple of S=%d", (long long)M, S);
  if (w->L_pts < 0 || w->L_pts > 15 || w->L_dirs < 0 || w->L_dirs > 6) {
    set_error("nb_mlp_fwd_f32: encoding octaves (%d,%d) outside [0,15]x[0,6]", w->L_pts, w->L_dirs);
    return NB_E_UNSUPPORTED;
  }
  int64_t ntiles = (M + kTS - 1) / kTS;
  int grid = (int)(ntiles < (int64_t)kNumSMs ? ntiles : (int64_t)kNumSMs);
  mlp_fwd_f32_kernel<<<grid, 256, kSimtSmem, as_stream(stream)>>>(*w, pts, viewdirs, M, S, raw, nullptr, nullptr, 0);
  return check_launch("nb_mlp_fwd_f32");
}

extern "C" int nb_mlp_fwd_f32_noview(const NbMlpWeights* w, const float* out_w, const float* out_b, int out_ch,
                                     const float* pts, int64_t M, float* raw, void* stream) {
  if (M == 0) return 0;
  NB_REQUIRE(w && out_w && out_b && pts && raw, "nb_mlp_fwd_f32_noview: bad arguments");
  if (w->L_pts < 0 || w->L_pts > 15 || out_ch < 1 || out_ch > 64) {
    set_error("nb_mlp_fwd_f32_noview: encoding octaves %d outside [0,15] or output_ch %d outside [1,64]", w->L_pts, out_ch);
    return NB_E_UNSUPPORTED;
  }
  NbMlpWeights t = *w;
  t.L_dirs = 0;
  int64_t ntiles = (M + kTS - 1) / kTS;
  int grid = (int)(ntiles < (int64_t)kNumSMs ? ntiles : (int64_t)kNumSMs);
  mlp_fwd_f32_kernel<<<grid, 256, kSimtSmem, as_stream(stream)>>>(t, pts, nullptr, M, 1, raw, out_w, out_b, out_ch);
  return check_launch("nb_mlp_fwd_f32_noview");
}
