// Hierarchical resampling: per-ray CDF, searchsorted and merge in shared memory.
// replaces utils.sample_pdf (utils.py:74-117) plus sort/cat/std (render_utils.py:141-147,168).
//
// One warp per ray.  The ray's bins, cdf and the (S+I)-entry merge buffer live in shared
// memory; nothing of the reference's [n,I,63] gather expansion is materialised.  Algorithmic
// HBM traffic: read z + weights (2*S*4 B) [+ u: I*4 B], write z_out ((S+I)*4 B) + z_std.
#include <math_constants.h>

#include "common.cuh"

namespace nb {

// searchsorted(cdf, u, right=True): number of entries <= u  (utils.py:101)
__device__ __forceinline__ int upper_bound(const float* cdf, int B, float u) {
  int lo = 0, hi = B;
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if (cdf[mid] <= u) lo = mid + 1; else hi = mid;
  }
  return lo;
}

// utils.py:102-115 with the reference's rounding order
__device__ __forceinline__ float invert_one(const float* bins, const float* cdf, int B, float u, int& inds) {
  inds = upper_bound(cdf, B, u);
  int below = max(inds - 1, 0);
  int above = min(B - 1, inds);
  float cb = cdf[below], ca = cdf[above];
  float bb = bins[below], ba = bins[above];
  float denom = __fsub_rn(ca, cb);
  if (denom < 1e-5f) denom = 1.0f;
  float t = __fdiv_rn(__fsub_rn(u, cb), denom);
  return __fadd_rn(bb, __fmul_rn(t, __fsub_rn(ba, bb)));
}

__global__ void __launch_bounds__(256) invert_cdf_kernel(const float* __restrict__ bins, const float* __restrict__ cdf,
                                                         const float* __restrict__ u, int64_t n, int B, int I,
                                                         float* __restrict__ samples, int64_t* __restrict__ inds) {
  int64_t total = n * I;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = i / I;
    int k;
    float s = invert_one(bins + r * B, cdf + r * B, B, u[i], k);
    if (samples) samples[i] = s;
    if (inds) inds[i] = k;
  }
}

// cdf[0] = 0, cdf[m] = sum_{j<m} (w[j]+1e-5)/sum(w+1e-5), m < B  (utils.py:76-79); w has B-1 entries
__device__ __forceinline__ void build_cdf(const float* __restrict__ w, int B, float* cdf, int lane) {
  float part = 0.f;
  for (int m = lane; m < B - 1; m += 32) {
    float v = __fadd_rn(w[m], 1e-5f);
    cdf[m + 1] = v;  // staged, normalised below
    part += v;
  }
  float wsum = warp_sum(part);
  __syncwarp();
  if (lane == 0) {
    float c = 0.f;
    cdf[0] = 0.f;
    for (int m = 1; m < B; ++m) {  // torch.cumsum is a sequential fp32 sum
      c = __fadd_rn(c, __fdiv_rn(cdf[m], wsum));
      cdf[m] = c;
    }
  }
  __syncwarp();
}

// utils.sample_pdf on explicit bins: one warp per row, cdf in shared memory
__global__ void __launch_bounds__(256) sample_pdf_kernel(const float* __restrict__ bins_g,
                                                         const float* __restrict__ weights,
                                                         const float* __restrict__ u,
                                                         const float* __restrict__ u_det, int64_t n, int B, int I,
                                                         int Bp, float* __restrict__ samples) {
  extern __shared__ float smem[];
  int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  float* bins = smem + (size_t)wib * 2 * Bp;
  float* cdf = bins + Bp;
  int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp; r < n; r += nwarps) {
    for (int m = lane; m < B; m += 32) bins[m] = bins_g[r * B + m];
    build_cdf(weights + r * (B - 1), B, cdf, lane);
    for (int i = lane; i < I; i += 32) {
      int k;
      samples[r * I + i] = invert_one(bins, cdf, B, u ? u[r * I + i] : u_det[i], k);
    }
    __syncwarp();
  }
}

// smem per warp: bins[Bp] cdf[Bp] buf[P]   (Bp = S rounded up to 32, P = pow2 >= S+I)
__global__ void __launch_bounds__(256) sample_pdf_merge_kernel(const float* __restrict__ z,
                                                               const float* __restrict__ weights,
                                                               const float* __restrict__ u,
                                                               const float* __restrict__ u_det, int64_t n, int S, int I,
                                                               int Bp, int P, float* __restrict__ z_out,
                                                               float* __restrict__ z_std, float* __restrict__ z_samples) {
  extern __shared__ float smem[];
  int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  float* bins = smem + (size_t)wib * (2 * Bp + P);
  float* cdf = bins + Bp;
  float* buf = cdf + Bp;
  const int B = S - 1;  // number of bins / cdf entries (63)
  int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp; r < n; r += nwarps) {
    const float* zr = z + r * S;
    const float* wr = weights + r * S;
    // coarse depths go straight into the merge buffer; mid-points into bins
    for (int s = lane; s < S; s += 32) buf[s] = zr[s];
    __syncwarp();
    for (int m = lane; m < B; m += 32) bins[m] = __fmul_rn(0.5f, __fadd_rn(buf[m + 1], buf[m]));
    build_cdf(wr + 1, B, cdf, lane);
    float sum = 0.f;
    for (int i = lane; i < I; i += 32) {
      float uu = u ? u[r * I + i] : u_det[i];
      int k;
      float smp = invert_one(bins, cdf, B, uu, k);
      buf[S + i] = smp;
      sum += smp;
      if (z_samples) z_samples[r * I + i] = smp;
    }
    for (int i = S + I + lane; i < P; i += 32) buf[i] = CUDART_INF_F;
    __syncwarp();
    if (z_std) {
      float mean = warp_sum(sum) / (float)I;
      float var = 0.f;
      for (int i = lane; i < I; i += 32) {
        float d = buf[S + i] - mean;
        var += d * d;
      }
      var = warp_sum(var) / (float)I;
      if (lane == 0) z_std[r] = sqrtf(var);
    }
    // bitonic sort of the P-entry buffer (ascending); value-deterministic like torch.sort
    for (int k = 2; k <= P; k <<= 1) {
      for (int j = k >> 1; j > 0; j >>= 1) {
        for (int t = lane; t < (P >> 1); t += 32) {
          int i = ((t / j) * (j << 1)) + (t % j);
          int l = i + j;
          float a = buf[i], b = buf[l];
          bool up = ((i & k) == 0);
          if ((a > b) == up) { buf[i] = b; buf[l] = a; }
        }
        __syncwarp();
      }
    }
    for (int i = lane; i < S + I; i += 32) z_out[r * (int64_t)(S + I) + i] = buf[i];
    __syncwarp();
  }
}

// ---- fast path: S <= 64 coarse samples, I = 32 * SPT importance samples (every shipped config) --------------
// Same arithmetic, different data movement: the cdf's divisions run on all lanes and only the 62 dependent
// additions of torch.cumsum stay serial (as register shuffles); the importance samples are sorted in registers
// with a shuffle bitonic network (random u only: deterministic u yields ascending samples) and merged with the
// ascending coarse depths by rank (two binary searches) instead of a 256-entry shared-memory bitonic sort.
template <int SPT>
__global__ void __launch_bounds__(256) sample_pdf_merge_fast_kernel(const float* __restrict__ z,
                                                                    const float* __restrict__ weights,
                                                                    const float* __restrict__ u,
                                                                    const float* __restrict__ u_det, int64_t n, int S,
                                                                    float* __restrict__ z_out, float* __restrict__ z_std,
                                                                    float* __restrict__ z_samples) {
  constexpr int I = 32 * SPT;
  __shared__ float smem[8][64 + 64 + 64 + I + 64 + I];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  float* zc = smem[wib];
  float* bins = zc + 64;
  float* cdf = bins + 64;
  float* srt = cdf + 64;
  float* outb = srt + I;
  const int B = S - 1;
  int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp; r < n; r += nwarps) {
    const float* zr = z + r * S;
    const float* wr = weights + r * S + 1;   // weights[..., 1:-1]: B - 1 = S - 2 entries
    float uu[SPT];
#pragma unroll
    for (int k = 0; k < SPT; ++k) uu[k] = u ? __ldcs(u + r * I + k * 32 + lane) : u_det[k * 32 + lane];
    const float z0 = lane < S ? zr[lane] : 0.f, z1 = lane + 32 < S ? zr[lane + 32] : 0.f;
    // v[m] = w[m] + 1e-5, m < B - 1 (lane holds m = lane and lane + 32); sum as build_cdf does
    const float v0 = lane < B - 1 ? __fadd_rn(wr[lane], 1e-5f) : 0.f;
    const float v1 = lane + 32 < B - 1 ? __fadd_rn(wr[lane + 32], 1e-5f) : 0.f;
    float part = 0.f;
    if (lane < B - 1) part += v0;
    if (lane + 32 < B - 1) part += v1;
    const float wsum = warp_sum(part);
    const float q0 = __fdiv_rn(v0, wsum), q1 = __fdiv_rn(v1, wsum);
    if (lane < S) zc[lane] = z0;
    if (lane + 32 < S) zc[lane + 32] = z1;
    // cdf[0] = 0, cdf[m + 1] = cdf[m] + q[m]: the sequential fp32 sum of torch.cumsum
    float c = 0.f, c_lo = 0.f, c_hi = 0.f;   // cdf[lane], cdf[lane + 32]
    for (int m = 0; m < B - 1; ++m) {
      const float qm = __shfl_sync(0xffffffffu, m < 32 ? q0 : q1, m & 31);
      c = __fadd_rn(c, qm);
      if (m + 1 == lane) c_lo = c;
      if (m + 1 == lane + 32) c_hi = c;
    }
    if (lane < B) cdf[lane] = c_lo;
    if (lane + 32 < B) cdf[lane + 32] = c_hi;
    __syncwarp();
    if (lane < B) bins[lane] = __fmul_rn(0.5f, __fadd_rn(zc[lane + 1], zc[lane]));
    if (lane + 32 < B) bins[lane + 32] = __fmul_rn(0.5f, __fadd_rn(zc[lane + 33], zc[lane + 32]));
    __syncwarp();
    float smp[SPT];
    float sum = 0.f;
#pragma unroll
    for (int k = 0; k < SPT; ++k) {
      int kk;
      smp[k] = invert_one(bins, cdf, B, uu[k], kk);
      sum += smp[k];
      if (z_samples) z_samples[r * I + k * 32 + lane] = smp[k];
    }
    if (z_std) {
      const float mean = warp_sum(sum) / (float)I;
      float var = 0.f;
#pragma unroll
      for (int k = 0; k < SPT; ++k) {
        const float d = smp[k] - mean;
        var += d * d;
      }
      var = warp_sum(var) / (float)I;
      if (lane == 0) z_std[r] = sqrtf(var);
    }
    if (u) {  // random draws: sort (element index e = k * 32 + lane)
#pragma unroll
      for (int kk = 2; kk <= I; kk <<= 1) {
#pragma unroll
        for (int j = kk >> 1; j > 0; j >>= 1) {
          if (j >= 32) {
            const int dj = j >> 5;
#pragma unroll
            for (int k = 0; k < SPT; ++k) {
              if ((k & dj) == 0) {
                const bool up = (((k * 32 + lane) & kk) == 0);
                const float a = smp[k], b = smp[k | dj];
                if ((a > b) == up) { smp[k] = b; smp[k | dj] = a; }
              }
            }
          } else {
#pragma unroll
            for (int k = 0; k < SPT; ++k) {
              const float o = __shfl_xor_sync(0xffffffffu, smp[k], j);
              const bool up = (((k * 32 + lane) & kk) == 0);
              const bool low = (lane & j) == 0;
              smp[k] = (low == up) ? fminf(smp[k], o) : fmaxf(smp[k], o);
            }
          }
        }
      }
    }
#pragma unroll
    for (int k = 0; k < SPT; ++k) srt[k * 32 + lane] = smp[k];
    __syncwarp();
    // merge by rank: coarse c -> c + #{samples < z_c}; sample e -> e + #{coarse <= sample}
    {
      const float zz[2] = {z0, z1};
#pragma unroll
      for (int t = 0; t < 2; ++t) {
        const int ci = lane + 32 * t;
        if (ci < S) {
          int lo = 0, hi = I;
          while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (srt[mid] < zz[t]) lo = mid + 1; else hi = mid;
          }
          outb[ci + lo] = zz[t];
        }
      }
#pragma unroll
      for (int k = 0; k < SPT; ++k) {
        int lo = 0, hi = S;
        while (lo < hi) {
          const int mid = (lo + hi) >> 1;
          if (zc[mid] <= smp[k]) lo = mid + 1; else hi = mid;
        }
        outb[k * 32 + lane + lo] = smp[k];
      }
    }
    __syncwarp();
    for (int i = lane; i < S + I; i += 32) __stcs(z_out + r * (int64_t)(S + I) + i, outb[i]);
    __syncwarp();
  }
}

// one step of the branch-free upper-bound search: addr += 4*STEP if cdf[(addr - base)/4 + STEP - 1] <= u
// (LDS [reg + imm], FSETP, predicated IADD: three instructions)
template <int STEP>
__device__ __forceinline__ void search_step(uint32_t& addr, float u) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      ".reg .f32 c;\n"
      "ld.shared.f32 c, [%0 + %2];\n"
      "setp.le.f32 p, c, %1;\n"
      "@p add.u32 %0, %0, %3;\n"
      "}\n"
      : "+r"(addr)
      : "f"(u), "n"((STEP - 1) * 4), "n"(STEP * 4)
      : "memory");
}

// ---- sorted-u path (no per-sample output requested) ---------------------------------------------------------
// The inverse cdf is monotone in u, so sorting u first (registers, blocked bitonic network: lane l owns the
// SPT consecutive slots l*SPT..l*SPT+SPT-1, 13 of the 28 stages for I = 128 never leave the lane) yields the
// importance samples already in ascending order.  The merge with the 64 ascending coarse depths is then pure
// ranking: sample e goes to e + a_e, a_e = #{coarse <= sample} (the sample lies between the mid-points of its
// cdf bin, so a_e is `below + 1` plus two comparisons), coarse c goes to c + #{e : a_e <= c}, a prefix sum over
// a 65-entry histogram of the a_e.  The cdf's 62 dependent additions (torch.cumsum order) run on one lane from
// shared memory as float4 reads / writes; every binary search is six fixed, branch-free steps.
// Same bits as sample_pdf_merge_fast_kernel for z_out (the multiset of samples is identical).
// SC: the number of coarse samples as a compile-time constant (64: every shipped config; the bounds predicates fold
// away) or 0 = the run-time value.  The host launches it with fewer than 2^31 rays: row indices are 32-bit.
template <int SPT, int SC>
__global__ void __launch_bounds__(256) sample_pdf_merge_sorted_kernel(const float* __restrict__ z,
                                                                      const float* __restrict__ weights,
                                                                      const float* __restrict__ u,
                                                                      const float* __restrict__ u_det, int64_t n, int S_rt,
                                                                      float* __restrict__ z_out, float* __restrict__ z_std) {
  constexpr int I = 32 * SPT;
  const int S = SC ? SC : S_rt;
  // + 4: the row stride is 20 (mod 32) banks, so the eight lanes of the block-wide cumsum below read / write their
  // rows' float4 without bank conflicts
  constexpr int kQ = 0, kCdf = 64, kBins = 128, kZc = 192, kHist = 264, kOut = 336, kPerWarp = kOut + 64 + I + 4;
  __shared__ __align__(16) float smem[8][kPerWarp];
  const int lane = threadIdx.x & 31;
  const int wib = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const uint32_t r_raw = blockIdx.x * 8u + (uint32_t)wib, n32 = (uint32_t)n;
  const bool live = r_raw < n32;               // no early exit: the block meets at two barriers
  const uint32_t r = live ? r_raw : n32 - 1u;
  float* qv = smem[wib] + kQ;
  float* cdf = smem[wib] + kCdf;
  float* bins = smem[wib] + kBins;
  float* zc = smem[wib] + kZc;
  int* hist = reinterpret_cast<int*>(smem[wib] + kHist);
  float* outb = smem[wib] + kOut;
  const int B = S - 1;
  const float* zr = z + (size_t)r * (uint32_t)S;
  const float* wr = weights + (size_t)r * (uint32_t)S + 1;   // weights[..., 1:-1]: B - 1 = S - 2 entries
  float uu[SPT];
  {
    const float* up = (u ? u + (size_t)r * I : u_det) + lane * SPT;
    if constexpr (SPT == 4) {
      const float4 v = u ? __ldcs(reinterpret_cast<const float4*>(up)) : *reinterpret_cast<const float4*>(up);
      uu[0] = v.x; uu[1] = v.y; uu[2] = v.z; uu[3] = v.w;
    } else {
      const float2 v = u ? __ldcs(reinterpret_cast<const float2*>(up)) : *reinterpret_cast<const float2*>(up);
      uu[0] = v.x; uu[1] = v.y;
    }
  }
  const float z0 = lane < S ? zr[lane] : CUDART_INF_F, z1 = lane + 32 < S ? zr[lane + 32] : CUDART_INF_F;
  const float v0 = lane < B - 1 ? __fadd_rn(wr[lane], 1e-5f) : 0.f;
  const float v1 = lane + 32 < B - 1 ? __fadd_rn(wr[lane + 32], 1e-5f) : 0.f;
  const float wsum = warp_sum(v0 + v1);
  qv[lane] = __fdiv_rn(v0, wsum);
  qv[lane + 32] = __fdiv_rn(v1, wsum);
  zc[lane] = z0;
  zc[lane + 32] = z1;
  if (lane < 8) zc[64 + lane] = CUDART_INF_F;
  hist[lane] = 0;
  hist[lane + 32] = 0;
  if (lane < 8) hist[64 + lane] = 0;
  __syncthreads();
  // cdf[m] = sum_{i<m} q[i], the sequential fp32 sum of torch.cumsum: 62 dependent additions per ray.  One lane per
  // ray of the block (eight lanes of warp 0) instead of lane 0 of every warp: a divergent lane costs its warp the
  // same issue slots as 32 lanes, so this is 1/8 of the instructions; the other warps sort meanwhile.
  if (wib == 0 && lane < 8) {
    const float* qrow = smem[lane] + kQ;
    float* crow = smem[lane] + kCdf;
    float c = 0.f;
#pragma unroll
    for (int g = 0; g < 16; ++g) {
      const float4 q = reinterpret_cast<const float4*>(qrow)[g];
      float4 o;
      o.x = c; c = __fadd_rn(c, q.x);
      o.y = c; c = __fadd_rn(c, q.y);
      o.z = c; c = __fadd_rn(c, q.z);
      o.w = c; c = __fadd_rn(c, q.w);
      reinterpret_cast<float4*>(crow)[g] = o;
    }
  }
  if (lane < B) bins[lane] = __fmul_rn(0.5f, __fadd_rn(zc[lane + 1], z0));
  if (lane + 32 < B) bins[lane + 32] = __fmul_rn(0.5f, __fadd_rn(zc[lane + 33], z1));
  if (u) {  // random draws: ascending sort, slot e = lane * SPT + k
    // Bitonic network in its "mirror" form: the first step of every merge stage pairs slot i with slot i ^ (kk - 1),
    // all later steps pair i with i ^ j, and every compare-exchange is ascending -- the steps that stay inside a
    // lane are two FMNMX with no direction select; only the cross-lane steps depend on the lane (keep min or max).
#pragma unroll
    for (int kk = 2; kk <= I; kk <<= 1) {
      // mirror step
      if (kk <= SPT) {
#pragma unroll
        for (int k = 0; k < SPT; ++k) {
          const int k2 = k ^ (kk - 1);
          if (k < k2) {
            const float lo = fminf(uu[k], uu[k2]), hi = fmaxf(uu[k], uu[k2]);
            uu[k] = lo;
            uu[k2] = hi;
          }
        }
      } else {
        const int dl = kk / SPT - 1;                 // partner lane = lane ^ dl, partner slot = SPT - 1 - k
        const bool keep_min = (lane & (kk / SPT / 2)) == 0;
        float o[SPT];
#pragma unroll
        for (int k = 0; k < SPT; ++k) o[k] = __shfl_xor_sync(0xffffffffu, uu[SPT - 1 - k], dl);
#pragma unroll
        for (int k = 0; k < SPT; ++k) uu[k] = keep_min ? fminf(uu[k], o[k]) : fmaxf(uu[k], o[k]);
      }
#pragma unroll
      for (int j = kk >> 2; j > 0; j >>= 1) {
        if (j < SPT) {
#pragma unroll
          for (int k = 0; k < SPT; ++k) {
            if ((k & j) == 0) {
              const float lo = fminf(uu[k], uu[k | j]), hi = fmaxf(uu[k], uu[k | j]);
              uu[k] = lo;
              uu[k | j] = hi;
            }
          }
        } else {
          const int dl = j / SPT;
          const bool keep_min = (lane & dl) == 0;
#pragma unroll
          for (int k = 0; k < SPT; ++k) {
            const float o = __shfl_xor_sync(0xffffffffu, uu[k], dl);
            uu[k] = keep_min ? fminf(uu[k], o) : fmaxf(uu[k], o);
          }
        }
      }
    }
  }
  __syncthreads();                                // the cdf rows written by warp 0 are visible
  if (lane >= B) cdf[lane] = CUDART_INF_F;        // entries past the cdf: every search stops before them
  if (lane + 32 >= B) cdf[lane + 32] = CUDART_INF_F;
  __syncwarp();
  int pos[SPT];
  {  // searchsorted(right=True): six fixed steps on a shared-memory byte address
    const uint32_t cdf_s = (uint32_t)__cvta_generic_to_shared(cdf);
#pragma unroll
    for (int k = 0; k < SPT; ++k) {
      uint32_t a = cdf_s;
      search_step<32>(a, uu[k]);
      search_step<16>(a, uu[k]);
      search_step<8>(a, uu[k]);
      search_step<4>(a, uu[k]);
      search_step<2>(a, uu[k]);
      search_step<1>(a, uu[k]);
      pos[k] = (int)((a - cdf_s) >> 2);
    }
  }
  float smp[SPT], sum = 0.f;
  int cnt[SPT];
  bool any2 = false;
#pragma unroll
  for (int k = 0; k < SPT; ++k) {
    const int below = max(pos[k] - 1, 0), above = min(B - 1, pos[k]);
    const float cb = cdf[below], ca = cdf[above], bb = bins[below], ba = bins[above];
    float denom = __fsub_rn(ca, cb);
    if (denom < 1e-5f) denom = 1.0f;
    const float t = __fdiv_rn(__fsub_rn(uu[k], cb), denom);
    smp[k] = __fadd_rn(bb, __fmul_rn(t, __fsub_rn(ba, bb)));
    sum += smp[k];
    // a_e = #{coarse <= sample}: coarse[below] <= bins[below] <= sample <= bins[above] <= coarse[below + 2]
    const bool c1 = zc[below + 1] <= smp[k], c2 = zc[below + 2] <= smp[k];
    cnt[k] = below + 1 + (c1 ? 1 : 0) + (c2 ? 1 : 0);
    any2 |= c2;
  }
  // c2 needs sample == bins[above] == coarse[below + 2], i.e. equal coarse depths (degenerate rays): count on, for the
  // whole warp at once and almost never
  if (__any_sync(0xffffffffu, any2)) {
#pragma unroll
    for (int k = 0; k < SPT; ++k)
      while (cnt[k] < S && zc[cnt[k]] <= smp[k]) ++cnt[k];
  }
#pragma unroll
  for (int k = 0; k < SPT; ++k) {
    atomicAdd(hist + cnt[k], 1);
    outb[lane * SPT + k + cnt[k]] = smp[k];
  }
  if (z_std) {
    const float mean = warp_sum(sum) / (float)I;
    float var = 0.f;
#pragma unroll
    for (int k = 0; k < SPT; ++k) {
      const float d = smp[k] - mean;
      var += d * d;
    }
    var = warp_sum(var) / (float)I;
    if (lane == 0 && live) z_std[r] = sqrtf(var);
  }
  __syncwarp();
  {  // coarse c = 2*lane, 2*lane+1 -> c + #{e : a_e <= c}
    const int2 h = reinterpret_cast<const int2*>(hist)[lane];
    const float2 zz = reinterpret_cast<const float2*>(zc)[lane];
    const int pair = h.x + h.y;
    int inc = pair;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int up_ = __shfl_up_sync(0xffffffffu, inc, o);
      inc += (lane >= o) ? up_ : 0;
    }
    const int b0 = inc - pair + h.x, b1 = inc;
    if (2 * lane < S) outb[2 * lane + b0] = zz.x;
    if (2 * lane + 1 < S) outb[2 * lane + 1 + b1] = zz.y;
  }
  __syncwarp();
  if (!live) return;
  float* zo = z_out + (size_t)r * (uint32_t)(S + I);
  if (S == 64 && ((uintptr_t)z_out & 15) == 0) {  // 16-byte stores ((64 + I) * 4 bytes per ray keeps the rows aligned)
    const float4* ob = reinterpret_cast<const float4*>(outb);
#pragma unroll
    for (int i = 0; i < (64 + I) / 4; i += 32)
      if (i + lane < (64 + I) / 4) __stcs(reinterpret_cast<float4*>(zo) + i + lane, ob[i + lane]);
  } else {
    for (int i = lane; i < S + I; i += 32) __stcs(zo + i, outb[i]);
  }
}

}  // namespace nb

using namespace nb;

extern "C" {

int nb_invert_cdf(const float* bins, const float* cdf, const float* u, int64_t n, int B, int I, float* samples,
                  int64_t* inds, void* stream) {
  if (n == 0) return 0;
  NB_REQUIRE(bins && cdf && u && B >= 1 && I >= 1, "nb_invert_cdf: bad arguments");
  invert_cdf_kernel<<<grid_for(n * I, 256), 256, 0, as_stream(stream)>>>(bins, cdf, u, n, B, I, samples, inds);
  return check_launch("nb_invert_cdf");
}

int nb_sample_pdf(const float* bins, const float* weights, const float* u, const float* u_det, int64_t n, int B, int I,
                  float* samples, void* stream) {
  if (n == 0) return 0;
  NB_REQUIRE(bins && weights && samples && B >= 2 && I >= 1 && (u || u_det), "nb_sample_pdf: bad arguments");
  int Bp = (B + 31) / 32 * 32;
  if (Bp > 1024) {
    set_error("nb_sample_pdf: %d bins > 1024 is not implemented", B);
    return NB_E_UNSUPPORTED;
  }
  sample_pdf_kernel<<<grid_for(n, 4), 128, 4 * 2 * Bp * sizeof(float), as_stream(stream)>>>(bins, weights, u, u_det, n,
                                                                                         B, I, Bp, samples);
  return check_launch("nb_sample_pdf");
}

int nb_sample_pdf_merge(const float* z, const float* weights, const float* u, const float* u_det, int64_t n, int S,
                        int I, float* z_out, float* z_std, float* z_samples, void* stream) {
  if (n == 0) return 0;
  NB_REQUIRE(z && weights && z_out && S >= 3 && I >= 1, "nb_sample_pdf_merge: bad arguments");
  NB_REQUIRE(u || u_det, "nb_sample_pdf_merge: need u or u_det");
  if (S <= 64 && (I == 64 || I == 128) && !z_samples && n < (1ll << 31) - 8) {
    const unsigned grid = (unsigned)((n + 7) / 8);   // one warp per ray
    cudaStream_t st = as_stream(stream);
    if (I == 128 && S == 64)
      sample_pdf_merge_sorted_kernel<4, 64><<<grid, 256, 0, st>>>(z, weights, u, u_det, n, S, z_out, z_std);
    else if (I == 128)
      sample_pdf_merge_sorted_kernel<4, 0><<<grid, 256, 0, st>>>(z, weights, u, u_det, n, S, z_out, z_std);
    else if (S == 64)
      sample_pdf_merge_sorted_kernel<2, 64><<<grid, 256, 0, st>>>(z, weights, u, u_det, n, S, z_out, z_std);
    else
      sample_pdf_merge_sorted_kernel<2, 0><<<grid, 256, 0, st>>>(z, weights, u, u_det, n, S, z_out, z_std);
    return check_launch("nb_sample_pdf_merge");
  }
  if (S <= 64 && (I == 64 || I == 128)) {   // per-sample output wanted (u order): sort after the inversion
    if (I == 128)
      sample_pdf_merge_fast_kernel<4><<<grid_for(n, 8), 256, 0, as_stream(stream)>>>(z, weights, u, u_det, n, S, z_out, z_std,
                                                                                       z_samples);
    else
      sample_pdf_merge_fast_kernel<2><<<grid_for(n, 8), 256, 0, as_stream(stream)>>>(z, weights, u, u_det, n, S, z_out, z_std,
                                                                                       z_samples);
    return check_launch("nb_sample_pdf_merge");
  }
  int P = 2;
  while (P < S + I) P <<= 1;
  int Bp = (S + 31) / 32 * 32;
  if (P > 4096) {
    set_error("nb_sample_pdf_merge: S+I=%d > 4096 is not implemented", S + I);
    return NB_E_UNSUPPORTED;
  }
  // warps per block chosen so that the merge buffers fit the default 48 KB dynamic smem
  int per_warp = (2 * Bp + P) * (int)sizeof(float);
  int wpb = 8;
  while (wpb > 1 && wpb * per_warp > 48 * 1024) wpb >>= 1;
  sample_pdf_merge_kernel<<<grid_for(n, wpb), wpb * 32, wpb * per_warp, as_stream(stream)>>>(
      z, weights, u, u_det, n, S, I, Bp, P, z_out, z_std, z_samples);
  return check_launch("nb_sample_pdf_merge");
}

}  // extern "C"
