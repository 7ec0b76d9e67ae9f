// Backward of the NeRF MLP on the tensor cores (autograd of nerf.py:110-134, main.py:103).
//
// Three kernels, all reading the tiled/swizzled activation slabs the forward kernel saved:
//   1. dgrad chain  (mlp_dgrad_kernel): persistent fused tcgen05 kernel, mirror image of the
//      forward: d_raw -> d(view hidden) -> d(feature) -> d(h7) ... -> d(h0) pre-activation
//      gradients stay on the SM between layers; each layer's dA tile is also written to HBM
//      (same slab image) for the weight-gradient pass.  Transposed bf16 weight slabs are
//      streamed through a 5-stage bulk-copy ring.
//   2. wgrad        (mlp_wgrad_kernel): dW_l = dA_l^T X_l as tcgen05 GEMMs whose reduction
//      dimension is the sample axis.  Both operands are consumed MN-major straight from the
//      saved slabs (a [128 samples x 64 features] swizzled slab is a valid MN-major operand),
//      fp32 accumulation in TMEM (up to all 512 columns), partial results reduced into the
//      fp32 gradient buffer with vector atomics.
//      The layer biases (column sums of dA) are taken from the operand stages while they sit in
//      shared memory.
//   3. small grads  (mlp_small_grads_kernel): the two CUDA-core heads (alpha_linear, rgb_linear)
//      and their biases; one CTA per SM, whole tiles of h7 / hv / d_raw staged by bulk copies.
#include <cuda_bf16.h>
#include <stdio.h>
#include <stdlib.h>

#include "common.cuh"
#include "mlp_common.cuh"
#include "ptx.cuh"
#include "tmap.cuh"

namespace nb {

using namespace ptx;

constexpr int kBwdSlabs = 68;         // views^T (4) + feature^T (8) + pts_linears[7..1]^T (7 x 8)
constexpr int kBwdConstFloats = 640;  // rgb_w [3][128] | alpha_w [256]
constexpr int kBwdStages = 5;

// dgrad shared memory map
constexpr uint32_t kBSmAct0 = 0, kBSmAct1 = 65536, kBSmRing = 131072;
constexpr uint32_t kBSmConst = kBSmRing + kBwdStages * kSlabBytes;  // 212992
constexpr uint32_t kBSmBar = kBSmConst + kBwdConstFloats * 4;       // 215552
constexpr uint32_t kBSmTmemPtr = kBSmBar + 120;
constexpr uint32_t kBSmemBytes = kBSmBar + 128 + 1024;

__host__ __device__ constexpr int bwd_ks_count(int step) { return step == 0 ? 2 : 4; }
__host__ __device__ constexpr int bwd_slab_base(int step) { return step == 0 ? 0 : 4 + 8 * (step - 1); }

// ------------------------------------------------------------------------------------------
// transposed weight packing: slab element (n, k) = W[out0 + k][in0 + n]
// ------------------------------------------------------------------------------------------
struct SlabSrcT {
  const float* w;
  int ld, in0, out0, kvalid;
};
struct PackArgsT {
  SlabSrcT slab[kBwdSlabs];
};

__global__ void __launch_bounds__(256) pack_bwd_kernel(PackArgsT a, NbMlpWeights w, uint8_t* __restrict__ blob) {
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx < kBwdSlabs * 128 * 8) {
    int s = idx / 1024, rem = idx - s * 1024;
    int n = rem >> 3, g = rem & 7;
    const SlabSrcT src = a.slab[s];
    uint32_t p[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int k = g * 8 + j * 2;
      float lo = (k < src.kvalid) ? src.w[(size_t)(src.out0 + k) * src.ld + src.in0 + n] : 0.f;
      float hi = (k + 1 < src.kvalid) ? src.w[(size_t)(src.out0 + k + 1) * src.ld + src.in0 + n] : 0.f;
      p[j] = pack_bf16x2(lo, hi);
    }
    *reinterpret_cast<uint4*>(blob + (size_t)s * kSlabBytes + sw128_chunk_off(n, g)) = make_uint4(p[0], p[1], p[2], p[3]);
  }
  float* c = reinterpret_cast<float*>(blob + (size_t)kBwdSlabs * kSlabBytes);
  if (idx < kBwdConstFloats) c[idx] = idx < 384 ? w.rgb_w[idx] : w.alpha_w[idx - 384];
}

// ------------------------------------------------------------------------------------------
// dgrad chain
// ------------------------------------------------------------------------------------------
// d (bf16x2) * (h > 0): relu'(h) applied to a packed pair
__device__ __forceinline__ uint32_t relu_mask2(uint32_t d, uint32_t h) {
  __nv_bfloat162 d2 = *reinterpret_cast<__nv_bfloat162*>(&d);
  __nv_bfloat162 h2 = *reinterpret_cast<__nv_bfloat162*>(&h);
  __nv_bfloat162 m = __hgt2(h2, __float2bfloat162_rn(0.f));
  __nv_bfloat162 r = __hmul2(d2, m);
  return *reinterpret_cast<uint32_t*>(&r);
}

// MODE 0: plain copy (d_feature); 1: + d_sigma * alpha_w, masked by relu'(h7); 2: masked
// This warp's four chunks [cbeg, cbeg + 4) of its 32 rows (two warps share a lane quarter).  relu' comes from
// the bit masks the forward saved: one coalesced 32-bit word per row and 32-column chunk (`mask` points at
// the [chunk][row] words of this layer), all four requested by the caller before it waits for the accumulator.
template <int MODE>
__device__ __forceinline__ void dgrad_epilogue(uint32_t taddr, uint32_t act, int row, int cbeg, const uint32_t (&mw)[4],
                                               bool live, float d_sigma, const float* __restrict__ alpha_w) {
#pragma unroll
  for (int cc = 0; cc < 4; ++cc) {
    const int c = cbeg + cc;
    uint32_t v[32];
    tmem_ld32(taddr + c * 32, v);
    tmem_ld_wait();
    const uint32_t m = (MODE == 0) ? 0xFFFFFFFFu : (live ? mw[cc] : 0u);
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      float f[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] = __uint_as_float(v[g * 8 + j]);
      if (MODE == 1) {
#pragma unroll
        for (int j = 0; j < 8; j += 4) {
          float4 a = *reinterpret_cast<const float4*>(alpha_w + c * 32 + g * 8 + j);
          f[j] = fmaf(d_sigma, a.x, f[j]); f[j + 1] = fmaf(d_sigma, a.y, f[j + 1]);
          f[j + 2] = fmaf(d_sigma, a.z, f[j + 2]); f[j + 3] = fmaf(d_sigma, a.w, f[j + 3]);
        }
      }
      if (MODE != 0) {
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] = ((m >> (31 - (g * 8 + j))) & 1u) ? f[j] : 0.f;
      }
      const uint32_t p0 = pack_bf16x2(f[0], f[1]), p1 = pack_bf16x2(f[2], f[3]), p2 = pack_bf16x2(f[4], f[5]),
                     p3 = pack_bf16x2(f[6], f[7]);
      const uint32_t off = (c >> 1) * kSlabBytes + sw128_chunk_off(row, (c & 1) * 4 + g);
      st_shared_v4(act + off, p0, p1, p2, p3);
    }
  }
}

__device__ __forceinline__ void pair_sync(int id) { asm volatile("bar.sync %0, 64;" ::"r"(id) : "memory"); }
constexpr int kBwdThreads = 640;  // 4 control warps + 16 epilogue warps (see mlp_tc.cu)

// CTA-pair form (tcgen05 cta_group::2), same protocol as the forward kernel (mlp_tc.cu): M = 256
// samples per MMA (128 per CTA), each CTA streams its half (128 of the 256 output columns) of every
// transposed weight slab, a cluster works on a quad of four tiles per iteration.
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kBwdThreads, 1)
    mlp_dgrad_kernel(const __grid_constant__ CUtensorMap wmap, const uint8_t* __restrict__ blob,
                     const uint8_t* __restrict__ acts, const float* __restrict__ d_raw, int64_t M,
                     uint8_t* __restrict__ dA, int num_quads, int* __restrict__ ready) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gbase = smem_raw + (base - smem_u32(smem_raw));
  const float* cst = reinterpret_cast<const float*>(gbase + kBSmConst);
  const uint32_t bar_full = base + kBSmBar, bar_empty = bar_full + 8 * kBwdStages, bar_act = bar_empty + 8 * kBwdStages,
                 bar_acc = bar_act + 16, bar_const = bar_acc + 16;
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int cluster = (int)cluster_id_x(), num_clusters = (int)cluster_count_x();

  if (threadIdx.x == 0) {
    for (int i = 0; i < kBwdStages; ++i) {
      mbar_init(bar_full + 8 * i, 1);
      mbar_init(bar_empty + 8 * i, 2);   // a weight slab is fetched once per step and released by BOTH tiles' MMAs
    }
    for (int x = 0; x < 2; ++x) {
      mbar_init(bar_act + 8 * x, 16);  // one arrival per epilogue warp of the tile (8), both CTAs
      mbar_init(bar_acc + 8 * x, 1);
    }
    mbar_init(bar_const, 1);
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc_2sm(base + kBSmTmemPtr, 512);
    tmem_relinquish_2sm();
  }
  if (warp == 0 && lane == 0) tma_prefetch_desc(&wmap);
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(gbase + kBSmTmemPtr);

  // producer and MMA warps run converged and issue under elect.sync (see mlp_tc.cu)
  if (warp == 0) {
    if (elect_one()) {
      mbar_expect_tx(bar_const, kBwdConstFloats * 4);
      bulk_g2s(base + kBSmConst, blob + (size_t)kBwdSlabs * kSlabBytes, kBwdConstFloats * 4, bar_const);
    }
    __syncwarp();
    uint32_t stage = 0, phase = 0;
    const uint32_t full_leader0 = mapa(bar_full, 0);
    for (int quad = cluster; quad < num_quads; quad += num_clusters) {
      for (int step = 0; step < 9; ++step) {
        const int nks = bwd_ks_count(step), sb = bwd_slab_base(step);
        // one fetch per step: the step's slabs (<= 4 of the 5 stages) stay in the ring for the MMAs of both tiles
        for (int ks = 0; ks < nks; ++ks) {
          mbar_wait(bar_empty + 8 * stage, phase ^ 1);
          const int row0 = (sb + ks * 2 + (int)rank) * 128;  // this CTA's 128 output columns
          if (elect_one()) {
            if (rank == 0) mbar_expect_tx(bar_full + 8 * stage, 2 * kSlabBytes);
            const uint32_t full_leader = full_leader0 + 8 * stage;
            const uint32_t dst = base + kBSmRing + stage * kSlabBytes;
            tma_load_2d_2sm(dst, &wmap, 0, row0, full_leader);
            tma_load_2d_2sm(dst + 8192, &wmap, 0, row0 + 64, full_leader);
          }
          __syncwarp();
          if (++stage == kBwdStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (rank == 0) {
      const uint32_t idesc = make_idesc_bf16(256, 256);
      const uint64_t desc0 = make_desc_kmajor_sw128(0);
      uint32_t stage = 0, phase = 0, par_act0 = 0, par_act1 = 0;
      for (int quad = cluster; quad < num_quads; quad += num_clusters) {
        for (int step = 0; step < 9; ++step) {
          const int nks = bwd_ks_count(step);
          const uint32_t stage0 = stage, phase0 = phase;   // both tiles walk the same stages of the ring
          for (int x = 0; x < 2; ++x) {
            stage = stage0;
            phase = phase0;
            const uint32_t sAct = base + (x ? kBSmAct1 : kBSmAct0);
            if (x == 0) { mbar_wait_cluster(bar_act, par_act0); par_act0 ^= 1; }
            else { mbar_wait_cluster(bar_act + 8, par_act1); par_act1 ^= 1; }
            tc_fence_after();
            const uint32_t d = tmem + x * 256;
            for (int ks = 0; ks < nks; ++ks) {
              const uint64_t adesc = desc0 | (uint64_t)(((sAct + ks * kSlabBytes) & 0x3FFFFu) >> 4);
              const uint64_t bdesc = desc0 | (uint64_t)(((base + kBSmRing + stage * kSlabBytes) & 0x3FFFFu) >> 4);
              mbar_wait(bar_full + 8 * stage, phase);
              tc_fence_after();
              if (elect_one()) {
                umma_bf16_2sm(d, adesc, bdesc, idesc, ks != 0);
                umma_bf16_2sm(d, adesc + 2, bdesc + 2, idesc, 1);
                umma_bf16_2sm(d, adesc + 4, bdesc + 4, idesc, 1);
                umma_bf16_2sm(d, adesc + 6, bdesc + 6, idesc, 1);
                umma_commit_2sm(bar_empty + 8 * stage);
                if (ks == nks - 1) umma_commit_2sm(bar_acc + 8 * x);
              }
              __syncwarp();
              if (++stage == kBwdStages) { stage = 0; phase ^= 1; }
            }
          }
        }
      }
    }
  } else if (warp >= 4) {
    // eight epilogue warps per tile: lane quarter q, column half h (chunks 4h .. 4h+3 = slabs 2h, 2h+1)
    const int x = (warp - 4) >> 3, h = ((warp - 4) >> 2) & 1, q = warp & 3;
    const int row = (q << 5) | lane;
    const int pair_id = 1 + x * 4 + q;
    const uint32_t sAct = base + (x ? kBSmAct1 : kBSmAct0);
    const uint32_t taddr = tmem + x * 256 + ((uint32_t)(q * 32) << 16);
    const uint32_t act_leader = mapa(bar_act + 8 * x, 0), my_acc = bar_acc + 8 * x;
    const uint32_t wrow = (uint32_t)q * 4096u;
    uint32_t par_acc = 0;
    mbar_wait(bar_const, 0);
    const float* wr = cst;
    const float* wa = cst + 384;
    for (int quad = cluster; quad < num_quads; quad += num_clusters) {
      const int64_t tile = (int64_t)quad * 4 + rank * 2 + x;
      const int64_t m = tile * kTileM + row;
      const bool real = tile * kTileM < M;
      // relu' bit masks of this tile ([group][chunk][row] words, after all activation tiles)
      const uint32_t* mk = real ? reinterpret_cast<const uint32_t*>(acts + (size_t)((M + kTileM - 1) / kTileM) * kActSlabs * kSlabBytes +
                                                                    (size_t)tile * kMaskTileBytes)
                                : nullptr;
      uint8_t* dt = real ? dA + (size_t)tile * kDaSlabs * kSlabBytes : nullptr;
      // every warp streams its own 32 rows (4 KB of each of its slabs) of the freshly written dA tile to HBM
      auto flush = [&](int src_slab0, int dst_slab0, int nslabs) {
        if (dt && lane == 0) {
          for (int j = 0; j < nslabs; ++j)
            bulk_s2g(dt + (size_t)(dst_slab0 + j) * kSlabBytes + wrow, sAct + (src_slab0 + j) * kSlabBytes + wrow, 4096);
          bulk_commit();
        }
      };
      auto reuse = [&]() {
        if (lane == 0) bulk_wait_read_all();
        __syncwarp();
      };
      float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
      if (m < M) g = *reinterpret_cast<const float4*>(d_raw + m * 4);
      reuse();
      pair_sync(pair_id);   // slab 1 below belongs to the h = 1 warp; step 8 of the previous quad flushed it from h = 0
      // prologue: d(view hidden) = (d_rgb . W_rgb) * relu'(hv)   (nerf.py:125-129 backwards); chunks 2h, 2h+1
#pragma unroll
      for (int cc = 0; cc < 2; ++cc) {
        const int c = 2 * h + cc;
        const uint32_t m = mk ? mk[(8 * 8 + c) * 128 + row] : 0u;
#pragma unroll
        for (int qq = 0; qq < 4; ++qq) {
          const int col = c * 32 + qq * 8;
          float f[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float t = g.x * wr[col + j] + g.y * wr[128 + col + j] + g.z * wr[256 + col + j];
            f[j] = ((m >> (31 - (qq * 8 + j))) & 1u) ? t : 0.f;
          }
          const uint32_t p0 = pack_bf16x2(f[0], f[1]), p1 = pack_bf16x2(f[2], f[3]), p2 = pack_bf16x2(f[4], f[5]),
                         p3 = pack_bf16x2(f[6], f[7]);
          const uint32_t off = (c >> 1) * kSlabBytes + sw128_chunk_off(row, (c & 1) * 4 + qq);
          st_shared_v4(sAct + off, p0, p1, p2, p3);
        }
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive_remote(act_leader);
      flush(h, kDaSlabHv + h, 1);
#pragma unroll 1
      for (int step = 0; step < 9; ++step) {
        const int l = 8 - step;  // step >= 1 produces d_a_l, masked by h_l: step 1 -> l = 7 ... step 8 -> l = 0
        uint32_t mw[4] = {0u, 0u, 0u, 0u};
        if (step > 0 && mk) {   // in flight while the MMAs of this step run
#pragma unroll
          for (int cc = 0; cc < 4; ++cc) mw[cc] = mk[(l * 8 + 4 * h + cc) * 128 + row];
        }
        mbar_wait(my_acc, par_acc);
        par_acc ^= 1;
        tc_fence_after();
        reuse();
        if (step == 0) pair_sync(pair_id);   // slab 1 was flushed (d hv) by the h = 1 warp, h = 0 overwrites it now
        int dst_slab;
        if (step == 0) {
          dgrad_epilogue<0>(taddr, sAct, row, 4 * h, mw, true, 0.f, nullptr);
          dst_slab = kDaSlabFeat;
        } else {
          if (step == 1) dgrad_epilogue<1>(taddr, sAct, row, 4 * h, mw, mk != nullptr, g.w, wa);
          else dgrad_epilogue<2>(taddr, sAct, row, 4 * h, mw, mk != nullptr, 0.f, nullptr);
          dst_slab = l * 4;
        }
        fence_proxy_async();
        tc_fence_before();
        __syncwarp();
        if (step < 8 && lane == 0) mbar_arrive_remote(act_leader);
        flush(2 * h, dst_slab + 2 * h, 2);
      }
      // streamed backward: this warp's share of the tile's dA has landed in global memory -> one of the tile's eight
      // arrivals; the weight-gradient kernel running beside this one picks the tile up while it is still in L2
      if (ready != nullptr && dt != nullptr && lane == 0) {
        bulk_wait_all();
        __threadfence();
        atomicAdd(ready + tile, 1);
      }
    }
    if (lane == 0) bulk_wait_all();
  }
  tc_fence_before();
  cluster_sync_all();
  if (warp == 2) tmem_dealloc_2sm(tmem, 512);
}

// ------------------------------------------------------------------------------------------
// wgrad: dW[out, in] += sum_samples dA[s, out] * X[s, in]
// ------------------------------------------------------------------------------------------
// A job streams, per 64-sample stage, an A region (dA blocks: out features) and a B region (activation blocks: in
// features), 8 KB per 64-feature block, and accumulates up to four cells: one 128-row dA block against one activation
// segment each.  Jobs are shaped so that every stage carries 7..9 blocks -- the narrow products share a stage with a
// wide operand instead of running as 3- and 5-block jobs that cannot keep the ring full (they ran 1.15-1.4x slower per
// byte and held the kernel's tail: tools/wgrad_skew.py):
//   enc job   A = dA_0 | dA_5 (8 blocks), B = encoded points (1):   dW_0 and dW_5[:, :in_ch], four 128 x 64 cells
//   layer l   A = dA_l (4), B = h_(l-1) (4):                        two 128 x 256 cells  (l = 1..7, feature_linear)
//   views     A = d hv (2), B = feature | encoded dirs (5):         128 x 256 and 128 x 64
struct WCell {
  int a_blk;     // first block of this cell's 128-row dA block inside the A region
  int b_blk;     // first block of this cell's activation segment inside the B region
  int n_cols;    // in features of the segment rounded up to 64 (64 or 256) = TMEM columns
  int n_valid;   // in features actually present
  int tcol;      // first TMEM column of the accumulator
  int ld;        // row stride of the fp32 weight gradient
  int scratch;   // 1: `off` is an offset into the launch's padded scratch block (rows of the real matrix are not 16-byte
                 // aligned: the cell is reduced there with vector atomics and added to the gradient by wgrad_unpad_kernel)
  int64_t off;   // float offset of dW[first out feature of the block][first in feature of the segment]
};
struct WJob {
  int a_slab[2], a_n[2];   // dA slab ranges of the A region
  int b_slab[2], b_n[2];   // activation slab ranges of the B region
  int n_cells;
  int bias_blocks;         // leading A blocks whose column sums are a bias gradient (<= 4)
  int64_t bias_off;        // float offset of that bias gradient, or -1
  WCell cell[4];
};
constexpr int kNumWJobs = 10;
struct WJobs {
  WJob job[kNumWJobs];
};

constexpr int kWStages = 3;
constexpr uint32_t kWStageBytes = 9 * 8192;  // 64 samples x up to nine 64-feature blocks
constexpr uint32_t kWSmRed = kWStages * kWStageBytes;       // [4 sample groups][256 features] fp32 bias partials
constexpr uint32_t kWSmBar = kWSmRed + 4096;
constexpr uint32_t kWSmTmemPtr = kWSmBar + 120;
constexpr uint32_t kWSmemBytes = kWSmBar + 128 + 1024;

// MN-major operand, 128-byte swizzle: 64-feature blocks `lbo` bytes apart, 8-sample groups 1024 B apart
__device__ __forceinline__ uint64_t make_desc_mnmajor_sw128(uint32_t smem_addr, uint32_t lbo) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)(lbo >> 4) << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
__host__ __device__ constexpr uint32_t make_idesc_bf16_mn(int M, int N) {
  return make_idesc_bf16(M, N) | (1u << 15) | (1u << 16);  // A and B MN-major
}

// Work split.  A job's cost per tile is the number of 8 KB operand blocks it streams (dA blocks + activation
// blocks); the jobs are laid end to end on one axis (job-major, `tiles` tiles each) and the axis is cut into
// gridDim.x equal pieces, so every CTA streams the same number of bytes and owns at most one partial range of two
// neighbouring jobs (<= 2 accumulator flushes).  ceil() on both ends of a piece assigns every (job, tile) once.
struct WSplit {
  int64_t cum[kNumWJobs + 1];  // cost before job j, for one tile of every job (units: blocks)
};
__device__ __forceinline__ void wgrad_unit(const WSplit& sp, const WJob& j, int ji, int num_tiles, int& t0, int& t1) {
  const int64_t total = sp.cum[kNumWJobs] * num_tiles;
  const int64_t lo = total * blockIdx.x / gridDim.x, hi = total * (blockIdx.x + 1) / gridDim.x;
  const int64_t w = sp.cum[ji + 1] - sp.cum[ji], start = sp.cum[ji] * num_tiles;
  const int64_t a = (lo - start + w - 1) / w, b = (hi - start + w - 1) / w;   // ceil; negative -> clamped below
  t0 = (int)min(max(lo <= start ? (int64_t)0 : a, (int64_t)0), (int64_t)num_tiles);
  t1 = (int)min(max(hi <= start ? (int64_t)0 : b, (int64_t)0), (int64_t)num_tiles);
  (void)j;
}

__device__ __forceinline__ void red_add_v4(float* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

// Streamed form (STREAM = true): the kernel runs BESIDE the dgrad chain on its own SMs and consumes dA tiles in the
// order the chain produces them, while they are still in L2.  Every CTA owns ONE job (layer block) and every
// parts[job]-th tile of it (CTAs per job in proportion to the blocks the job streams), so each job's accumulator is
// still flushed once per CTA; the producer warp waits for a tile's eight arrivals in `ready` before it issues the loads.
struct WStream {
  unsigned char job[kNumSMs];    // job of CTA b
  unsigned char part[kNumSMs];   // its first tile
  unsigned char parts[kNumWJobs];  // CTAs working on the job = tile stride
};
template <bool STREAM>
__device__ __forceinline__ void wgrad_range(const WSplit& sp, const WStream& ws, const WJob& j, int ji, int num_tiles, int& tb,
                                            int& te, int& ts) {
  if (STREAM) {
    const bool mine = ws.job[blockIdx.x] == ji;
    tb = mine ? (int)ws.part[blockIdx.x] : 0;
    te = mine ? num_tiles : 0;
    ts = mine ? (int)ws.parts[ji] : 1;
  } else {
    wgrad_unit(sp, j, ji, num_tiles, tb, te);
    ts = 1;
  }
}

// Padded scratch of a launch.  Three weight matrices have rows that are not 16-byte aligned (row strides in_ch,
// 256 + in_ch and 256 + in_views floats); scalar atomics from the ~16 CTAs of such a job, all flushing the same elements
// at the same moment, took 40-70 us (tools/wgrad_skew.py).  Their cells are reduced into padded blocks here with vector
// atomics, and wgrad_unpad_kernel adds the blocks to the real rows.  The scratch is the tail of the caller's dA buffer
// (nb_mlp_da_bytes counts it), zeroed by the call before the weight-gradient kernel starts: private to the call, so
// launches that overlap on the GPU (the coarse and the fine branch of a step) cannot meet in it.
//   [256][256] dW_5[:, in_ch:] | [256][64] dW_0 | [256][64] dW_5[:, :in_ch] | [128][256] dW_v[:, :256] | [128][64] dW_v[:, 256:]
constexpr int kWScr5 = 0, kWScr0 = kWScr5 + 256 * 256, kWScr5e = kWScr0 + 256 * 64, kWScrV = kWScr5e + 256 * 64,
              kWScrVd = kWScrV + 128 * 256, kWScratchFloats = kWScrVd + 128 * 64;
struct WUnpad {
  int start[6];      // first scratch float of region k (start[5] = end)
  int cols_log2[5];  // padded row length
  int valid[5];      // real row length
  int ld[5];         // row stride of the destination
  int64_t off[5];    // destination of element (0, 0)
};

// adds the padded scratch cells to the real gradient rows
__global__ void __launch_bounds__(256) wgrad_unpad_kernel(const float* __restrict__ scratch, float* __restrict__ grads, WUnpad u) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= kWScratchFloats) return;
  const float v = scratch[i];
  int k = 0;
  while (i >= u.start[k + 1]) ++k;
  const int e = i - u.start[k], r = e >> u.cols_log2[k], c = e & ((1 << u.cols_log2[k]) - 1);
  if (c < u.valid[k]) atomicAdd(grads + u.off[k] + (int64_t)r * u.ld[k] + c, v);
}

// start / end (globaltimer, ns) of every CTA of the last weight-gradient launch: nb_debug_wgrad_trace
__device__ unsigned long long g_wgrad_trace[2 * kNumSMs];
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

template <bool STREAM>
__global__ void __launch_bounds__(384, 1) mlp_wgrad_kernel(WJobs jobs, WSplit split, WStream ws, const uint8_t* __restrict__ acts,
                                                           const uint8_t* __restrict__ dA, int num_tiles,
                                                           float* __restrict__ grads, const int* __restrict__ ready,
                                                           float* __restrict__ scratch) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gbase = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t bar_full = base + kWSmBar, bar_empty = bar_full + 8 * kWStages, bar_acc = bar_empty + 8 * kWStages,
                 bar_flush = bar_acc + 8;
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int i = 0; i < kWStages; ++i) {
      mbar_init(bar_full + 8 * i, 1);
      mbar_init(bar_empty + 8 * i, 2);  // tcgen05.commit + the bias warps
    }
    mbar_init(bar_acc, 1);
    mbar_init(bar_flush, 128);
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc(base + kWSmTmemPtr, 512);
    tmem_relinquish();
  }
  if (threadIdx.x == 0 && blockIdx.x < kNumSMs) g_wgrad_trace[2 * blockIdx.x] = globaltimer_ns();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(gbase + kWSmTmemPtr);

  // producer and MMA warps run converged and issue under elect.sync (see mlp_tc.cu)
  if (warp == 0) {
    uint32_t stage = 0, phase = 0;
    for (int ji = 0; ji < kNumWJobs; ++ji) {
      const WJob& j = jobs.job[ji];
      int t0, t1, ts;
      wgrad_range<STREAM>(split, ws, j, ji, num_tiles, t0, t1, ts);
      const int na = j.a_n[0] + j.a_n[1], nb_ = j.b_n[0] + j.b_n[1];
      for (int t = t0; t < t1; t += ts) {
        if (STREAM) {   // the dgrad chain has written this tile (8 epilogue warps): acquire, then order the async-proxy loads
          long long spins = 0;
          int seen;
          do {
            asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(seen) : "l"(ready + t) : "memory");
            if (seen < 8) {
              if (++spins > (1ll << 24)) __trap();   // seconds: the producer kernel is gone
              __nanosleep(200);
            }
          } while (seen < 8);
          asm volatile("fence.proxy.async;" ::: "memory");
        }
        const uint8_t* da_t = dA + (size_t)t * kDaSlabs * kSlabBytes;
        const uint8_t* ac_t = acts + (size_t)t * kActSlabs * kSlabBytes;
        for (int half = 0; half < 2; ++half) {
          mbar_wait(bar_empty + 8 * stage, phase ^ 1);
          if (elect_one()) {
            mbar_expect_tx(bar_full + 8 * stage, (na + nb_) * 8192);
            uint32_t dst = base + stage * kWStageBytes;
            for (int r = 0; r < 2; ++r)
              for (int s = 0; s < j.a_n[r]; ++s, dst += 8192)
                bulk_g2s(dst, da_t + (size_t)(j.a_slab[r] + s) * kSlabBytes + half * 8192, 8192, bar_full + 8 * stage);
            for (int r = 0; r < 2; ++r)
              for (int s = 0; s < j.b_n[r]; ++s, dst += 8192)
                bulk_g2s(dst, ac_t + (size_t)(j.b_slab[r] + s) * kSlabBytes + half * 8192, 8192, bar_full + 8 * stage);
          }
          __syncwarp();
          if (++stage == kWStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    uint32_t stage = 0, phase = 0, par_flush = 0;
    bool first_unit = true;
    const uint64_t desc0 = make_desc_mnmajor_sw128(0, 8192);
    for (int ji = 0; ji < kNumWJobs; ++ji) {
      const WJob& j = jobs.job[ji];
      int t0, t1, ts;
      wgrad_range<STREAM>(split, ws, j, ji, num_tiles, t0, t1, ts);
      if (t0 >= t1) continue;
      const int na = j.a_n[0] + j.a_n[1];
      if (!first_unit) {  // the previous unit's accumulators must be drained before they are overwritten
        mbar_wait(bar_flush, par_flush);
        par_flush ^= 1;
        tc_fence_after();
      }
      first_unit = false;
      uint32_t acc = 0;
      for (int t = t0; t < t1; t += ts) {
        for (int half = 0; half < 2; ++half) {
          mbar_wait(bar_full + 8 * stage, phase);
          tc_fence_after();
          const uint32_t sa = base + stage * kWStageBytes, sb = sa + na * 8192;
          if (elect_one()) {
            for (int ci = 0; ci < j.n_cells; ++ci) {
              const WCell& cl = j.cell[ci];
              const uint32_t d = tmem + cl.tcol;
              const uint32_t idesc = make_idesc_bf16_mn(128, cl.n_cols);
              const uint64_t ad = desc0 | (uint64_t)(((sa + cl.a_blk * 8192) & 0x3FFFFu) >> 4);
              const uint64_t bd = desc0 | (uint64_t)(((sb + cl.b_blk * 8192) & 0x3FFFFu) >> 4);
              umma_bf16(d, ad, bd, idesc, acc);
              umma_bf16(d, ad + (2048 >> 4), bd + (2048 >> 4), idesc, 1);
              umma_bf16(d, ad + 2 * (2048 >> 4), bd + 2 * (2048 >> 4), idesc, 1);
              umma_bf16(d, ad + 3 * (2048 >> 4), bd + 3 * (2048 >> 4), idesc, 1);
            }
            umma_commit(bar_empty + 8 * stage);
            if (t + ts >= t1 && half == 1) umma_commit(bar_acc);
          }
          __syncwarp();
          acc = 1;
          if (++stage == kWStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp >= 4 && warp < 8) {
    // flush warps: TMEM lane = out feature within the 128-block
    const int row = ((warp & 3) << 5) | lane;
    uint32_t par_acc = 0;
    for (int ji = 0; ji < kNumWJobs; ++ji) {
      const WJob j = jobs.job[ji];
      int t0, t1, ts;
      wgrad_range<STREAM>(split, ws, j, ji, num_tiles, t0, t1, ts);
      if (t0 >= t1) continue;
      mbar_wait(bar_acc, par_acc);
      par_acc ^= 1;
      tc_fence_after();
      for (int ci = 0; ci < j.n_cells; ++ci) {
        const WCell cl = j.cell[ci];
        // 16-byte reductions where every row of this cell starts on a 16-byte boundary
        float* cbase = (cl.scratch ? scratch : grads) + cl.off;
        const bool vec = (cl.scratch || cl.n_valid == cl.n_cols) && (cl.ld & 3) == 0 && (((uintptr_t)cbase) & 15) == 0;
        float* grow = cbase + (size_t)row * cl.ld;
        // the CTAs of a job flush the same elements at about the same time: each starts at another column chunk
        const int nch = cl.n_cols >> 5, rot = (int)blockIdx.x;
        for (int cc = 0; cc < nch; ++cc) {
          const int c = ((cc + rot) & (nch - 1)) << 5;
          uint32_t v[32];
          tmem_ld32(tmem + ((uint32_t)((warp & 3) * 32) << 16) + cl.tcol + c, v);
          tmem_ld_wait();
          if (vec) {
#pragma unroll
            for (int q = 0; q < 32; q += 4) red_add_v4(grow + c + q, v[q], v[q + 1], v[q + 2], v[q + 3]);
          } else {
#pragma unroll
            for (int q = 0; q < 32; ++q)
              if (c + q < cl.n_valid) atomicAdd(grow + c + q, __uint_as_float(v[q]));
          }
        }
      }
      tc_fence_before();
      mbar_arrive(bar_flush);
    }
  } else if (warp >= 8) {
    // bias warps: db = column sums of dA over the samples, taken from the operand stages while they
    // sit in shared memory (no second pass over dA in HBM).  Thread (c, q): 8 features 8c..8c+7,
    // samples q, q+4, ... of each 64-sample stage.
    const int t = threadIdx.x - 256, c = t & 31, q = t >> 5;
    float* red = reinterpret_cast<float*>(gbase + kWSmRed);
    uint32_t stage = 0, phase = 0;
    for (int ji = 0; ji < kNumWJobs; ++ji) {
      const WJob j = jobs.job[ji];
      int t0, t1, ts;
      wgrad_range<STREAM>(split, ws, j, ji, num_tiles, t0, t1, ts);
      if (t0 >= t1) continue;
      const bool mine = j.bias_off >= 0 && c < j.bias_blocks * 8;
      float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      for (int st = 2 * ((t1 - t0 + ts - 1) / ts); st > 0; --st) {
        mbar_wait(bar_full + 8 * stage, phase);
        if (mine) {
          const uint32_t blk = base + stage * kWStageBytes + (c >> 3) * 8192;
#pragma unroll 4
          for (int i = 0; i < 16; ++i) {
            const int sm = q + 4 * i;
            uint4 v;
            asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                         : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                         : "r"(blk + (sm >> 3) * 1024 + (sm & 7) * 128 + (((c & 7) ^ (sm & 7)) << 4)));
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              acc[2 * k] += __uint_as_float(w[k] << 16);
              acc[2 * k + 1] += __uint_as_float(w[k] & 0xFFFF0000u);
            }
          }
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");  // all four bias warps are done reading the stage
        if (t == 0) mbar_arrive(bar_empty + 8 * stage);
        if (++stage == kWStages) { stage = 0; phase ^= 1; }
      }
      if (j.bias_off >= 0) {
        if (mine) {
#pragma unroll
          for (int k = 0; k < 8; ++k) red[q * 256 + c * 8 + k] = acc[k];
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
        for (int f = t; f < j.bias_blocks * 64; f += 128)
          atomicAdd(grads + j.bias_off + f, red[f] + red[256 + f] + red[512 + f] + red[768 + f]);
        asm volatile("bar.sync 1, 128;" ::: "memory");
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0 && blockIdx.x < kNumSMs) g_wgrad_trace[2 * blockIdx.x + 1] = globaltimer_ns();
  if (warp == 2) tmem_dealloc(tmem, 512);
}

// ------------------------------------------------------------------------------------------
// biases and CUDA-core heads
// ------------------------------------------------------------------------------------------
struct SmallOffsets {
  int64_t pts_b[8], feature_b, views_b, alpha_w, alpha_b, rgb_w, rgb_b;
};

__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    f[2 * j] = __uint_as_float(w[j] << 16);
    f[2 * j + 1] = __uint_as_float(w[j] & 0xFFFF0000u);
  }
}

// One CTA per SM, 8 consumer warps + 1 producer warp.  The producer stages whole tiles -- h7 (4 slabs, 64 KB
// contiguous), hv (2 slabs, 32 KB contiguous) and the tile's 128 d_raw rows (2 KB) -- with 1-D bulk copies into a
// two-stage ring (2 x 98 KB in flight per SM: the register-staged version of this kernel had ~23 KB in flight and
// ran at 3.4 TB/s, latency-bound); the consumers read the swizzled slab image from shared memory.
// alpha_linear: dW = sum d_sigma * h7, db = sum d_sigma ; rgb_linear: dW[k] = sum d_rgb[k] * hv, db = sum d_rgb.
constexpr int kHgStages = 2;
constexpr uint32_t kHgStageBytes = 6 * kSlabBytes + kTileM * 16;
constexpr uint32_t kHgSmRed = kHgStages * kHgStageBytes;   // [8 warps][80 values][8 chunks] fp32
constexpr uint32_t kHgSmBar = kHgSmRed + 8 * 80 * 8 * 4;
constexpr uint32_t kHgSmemBytes = kHgSmBar + 64 + 1024;

__global__ void __launch_bounds__(288, 1) mlp_small_grads_kernel(const uint8_t* __restrict__ acts,
                                                                 const float* __restrict__ d_raw, int64_t M,
                                                                 int num_tiles, SmallOffsets o,
                                                                 float* __restrict__ grads) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gbase = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t bar_full = base + kHgSmBar, bar_empty = bar_full + 8 * kHgStages;
  const int t = threadIdx.x, lane = t & 31;
  const int warp = __shfl_sync(0xffffffffu, t >> 5, 0);
  const int t0 = (int)((int64_t)num_tiles * blockIdx.x / gridDim.x), t1 = (int)((int64_t)num_tiles * (blockIdx.x + 1) / gridDim.x);
  if (t == 0) {
    for (int i = 0; i < kHgStages; ++i) {
      mbar_init(bar_full + 8 * i, 1);
      mbar_init(bar_empty + 8 * i, 8);   // one arrival per consumer warp
    }
    fence_mbar_init();
  }
  __syncthreads();
  if (warp == 8) {
    uint32_t stage = 0, phase = 0;
    for (int tile = t0; tile < t1; ++tile) {
      mbar_wait(bar_empty + 8 * stage, phase ^ 1);
      if (elect_one()) {
        const int64_t left = M - (int64_t)tile * kTileM;
        const uint32_t rows = (uint32_t)(left < kTileM ? left : kTileM);
        const uint8_t* tb = acts + (size_t)tile * kActSlabs * kSlabBytes;
        const uint32_t dst = base + stage * kHgStageBytes;
        mbar_expect_tx(bar_full + 8 * stage, 6 * kSlabBytes + rows * 16);
        bulk_g2s(dst, tb + (size_t)28 * kSlabBytes, 4 * kSlabBytes, bar_full + 8 * stage);
        bulk_g2s(dst + 4 * kSlabBytes, tb + (size_t)kActSlabHv * kSlabBytes, 2 * kSlabBytes, bar_full + 8 * stage);
        bulk_g2s(dst + 6 * kSlabBytes, d_raw + (size_t)tile * kTileM * 4, rows * 16, bar_full + 8 * stage);
      }
      __syncwarp();
      if (++stage == kHgStages) { stage = 0; phase ^= 1; }
    }
  } else {
    const int g = t & 7, rg = t >> 3;
    float accA[4][8], accR[3][2][8], a4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int s = 0; s < 4; ++s)
#pragma unroll
      for (int j = 0; j < 8; ++j) accA[s][j] = 0.f;
#pragma unroll
    for (int k = 0; k < 3; ++k)
#pragma unroll
      for (int s = 0; s < 2; ++s)
#pragma unroll
        for (int j = 0; j < 8; ++j) accR[k][s][j] = 0.f;
    uint32_t stage = 0, phase = 0;
    for (int tile = t0; tile < t1; ++tile) {
      mbar_wait(bar_full + 8 * stage, phase);
      const uint8_t* sb = gbase + stage * kHgStageBytes;
#pragma unroll 2
      for (int i = 0; i < 4; ++i) {
        const int r = rg + 32 * i;
        const int64_t m = (int64_t)tile * kTileM + r;
        const uint32_t off = sw128_chunk_off(r, g);
        uint4 v[6];
#pragma unroll
        for (int s = 0; s < 6; ++s) v[s] = *reinterpret_cast<const uint4*>(sb + (size_t)s * kSlabBytes + off);
        const float4 d = m < M ? *reinterpret_cast<const float4*>(sb + 6 * kSlabBytes + r * 16) : make_float4(0.f, 0.f, 0.f, 0.f);
        const float dk[3] = {d.x, d.y, d.z};
        if (g == 0) { a4[0] += d.x; a4[1] += d.y; a4[2] += d.z; a4[3] += d.w; }   // head biases: column sums of d_raw
#pragma unroll
        for (int s = 0; s < 4; ++s) {
          float f[8];
          unpack8(v[s], f);
#pragma unroll
          for (int j = 0; j < 8; ++j) accA[s][j] = fmaf(d.w, f[j], accA[s][j]);
        }
#pragma unroll
        for (int s = 0; s < 2; ++s) {
          float f[8];
          unpack8(v[4 + s], f);
#pragma unroll
          for (int k = 0; k < 3; ++k)
#pragma unroll
            for (int j = 0; j < 8; ++j) accR[k][s][j] = fmaf(dk[k], f[j], accR[k][s][j]);
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_empty + 8 * stage);
      if (++stage == kHgStages) { stage = 0; phase ^= 1; }
    }
    // sum over the block's 32 row groups: the four of a warp by shuffles (lane bits 3, 4), the eight warps through
    // shared memory in one pass
    float* red = reinterpret_cast<float*>(gbase + kHgSmRed);
    float* flat = &accA[0][0];
    float* flatR = &accR[0][0][0];
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      flat[i] += __shfl_xor_sync(0xffffffffu, flat[i], 8);
      flat[i] += __shfl_xor_sync(0xffffffffu, flat[i], 16);
    }
#pragma unroll
    for (int i = 0; i < 48; ++i) {
      flatR[i] += __shfl_xor_sync(0xffffffffu, flatR[i], 8);
      flatR[i] += __shfl_xor_sync(0xffffffffu, flatR[i], 16);
    }
    if (lane < 8) {  // lane == g; value index v = s*8+j (alpha, 32 values) or 32 + (k*2+s)*8+j (rgb, 48 values)
#pragma unroll
      for (int i = 0; i < 32; ++i) red[(warp * 80 + i) * 8 + lane] = flat[i];
#pragma unroll
      for (int i = 0; i < 48; ++i) red[(warp * 80 + 32 + i) * 8 + lane] = flatR[i];
    }
    asm volatile("bar.sync 1, 256;" ::: "memory");   // the eight consumer warps
    for (int idx = t; idx < 640; idx += 256) {
      const int v = idx >> 3, gg = idx & 7;
      float sum = 0.f;
#pragma unroll
      for (int w = 0; w < 8; ++w) sum += red[(w * 80 + v) * 8 + gg];
      float* dst;
      if (v < 32) {
        dst = grads + o.alpha_w + (v >> 3) * 64 + gg * 8 + (v & 7);
      } else {
        const int u = v - 32, k = u >> 4, sl = (u >> 3) & 1;
        dst = grads + o.rgb_w + k * 128 + sl * 64 + gg * 8 + (u & 7);
      }
      atomicAdd(dst, sum);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) a4[k] = warp_sum(a4[k]);
    if (lane == 0) {
      atomicAdd(grads + o.rgb_b + 0, a4[0]);
      atomicAdd(grads + o.rgb_b + 1, a4[1]);
      atomicAdd(grads + o.rgb_b + 2, a4[2]);
      atomicAdd(grads + o.alpha_b, a4[3]);
    }
  }
}

// ------------------------------------------------------------------------------------------
// MN-major UMMA self-test: D[128,128] = A[K,128]^T * B[K,128], K in {64, 128}
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128, 1) umma_mn_selftest_kernel(const float* __restrict__ A, const float* __restrict__ B,
                                                                  int K, float* __restrict__ D) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gbase = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t slab = K * 128;  // bytes of one [K samples x 64 features] block
  const uint32_t sA = base, sB = base + 2 * slab, sBar = sB + 2 * slab, sPtr = sBar + 8;
  const int warp = threadIdx.x >> 5;
  // thread t fills sample row (t % K) for feature blocks
  for (int idx = threadIdx.x; idx < K * 2 * 8; idx += 128) {
    const int k = idx % K, fb = (idx / K) % 2, g = idx / (2 * K);
    uint32_t pa[4], pb[4];
    for (int j = 0; j < 4; ++j) {
      const int f = fb * 64 + g * 8 + j * 2;
      pa[j] = pack_bf16x2(A[k * 128 + f], A[k * 128 + f + 1]);
      pb[j] = pack_bf16x2(B[k * 128 + f], B[k * 128 + f + 1]);
    }
    st_shared_v4(sA + fb * slab + sw128_chunk_off(k, g), pa[0], pa[1], pa[2], pa[3]);
    st_shared_v4(sB + fb * slab + sw128_chunk_off(k, g), pb[0], pb[1], pb[2], pb[3]);
  }
  fence_proxy_async();
  if (threadIdx.x == 0) {
    mbar_init(sBar, 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    tmem_alloc(sPtr, 128);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(gbase + (sPtr - base));
  if (threadIdx.x == 0) {
    const uint32_t idesc = make_idesc_bf16_mn(128, 128);
    for (int k16 = 0; k16 < K / 16; ++k16)
      umma_bf16(tmem, make_desc_mnmajor_sw128(sA + k16 * 2048, slab), make_desc_mnmajor_sw128(sB + k16 * 2048, slab),
                idesc, k16 != 0);
    umma_commit(sBar);
  }
  mbar_wait(sBar, 0);
  tc_fence_after();
  const int row = threadIdx.x;
  for (int c = 0; c < 4; ++c) {
    uint32_t v[32];
    tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + c * 32, v);
    tmem_ld_wait();
    for (int j = 0; j < 32; ++j) D[row * 128 + c * 32 + j] = __uint_as_float(v[j]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 128);
}

int init_mlp_bwd(int) {
  cudaError_t e = cudaFuncSetAttribute(mlp_dgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kBSmemBytes);
  if (e == cudaSuccess)
    e = cudaFuncSetAttribute(mlp_wgrad_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kWSmemBytes);
  if (e == cudaSuccess)
    e = cudaFuncSetAttribute(mlp_wgrad_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kWSmemBytes);
  if (e == cudaSuccess)
    e = cudaFuncSetAttribute(mlp_small_grads_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kHgSmemBytes);
  if (e == cudaSuccess)
    e = cudaFuncSetAttribute(umma_mn_selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 * 16384 + 2048);
  if (e != cudaSuccess) {
    set_error("init_mlp_bwd: %s", cudaGetErrorString(e));
    return (int)e;
  }
  return 0;
}

static int build_pack_args_t(const NbMlpWeights* w, PackArgsT& a) {
  const int in_ch = 3 + 6 * w->L_pts, in_views = 3 + 6 * w->L_dirs;
  if (in_ch > 64 || in_views > 32 || w->L_pts < 0 || w->L_dirs < 0) {
    set_error("tensor-core MLP supports multires <= 10 and multires_views <= 4 (got %d, %d)", w->L_pts, w->L_dirs);
    return NB_E_UNSUPPORTED;
  }
  int s = 0;
  for (int step = 0; step < 9; ++step) {
    const float* mat;
    int ld, in_base = 0, kout;
    if (step == 0) { mat = w->views_w; ld = 256 + in_views; kout = 128; }
    else if (step == 1) { mat = w->feature_w; ld = 256; kout = 256; }
    else {
      const int l = 9 - step;  // pts_linears[7] ... pts_linears[1]
      mat = w->pts_w[l];
      ld = l == 5 ? 256 + in_ch : 256;
      in_base = l == 5 ? in_ch : 0;  // the h part of cat[input_pts, h]
      kout = 256;
    }
    for (int ks = 0; ks < bwd_ks_count(step); ++ks)
      for (int nh = 0; nh < 2; ++nh) a.slab[s++] = SlabSrcT{mat, ld, in_base + nh * 128, ks * 64, kout - ks * 64 < 64 ? kout - ks * 64 : 64};
  }
  return s == kBwdSlabs ? 0 : NB_E_ARG;
}

// the ten weight-gradient jobs and the cost axis of the static split
static WJob square_job(int a_slab, int b_slab, int64_t off, int ld, int64_t bias_off) {
  WJob j = {};
  j.a_slab[0] = a_slab; j.a_n[0] = 4;
  j.b_slab[0] = b_slab; j.b_n[0] = 4;
  j.n_cells = 2;
  j.bias_blocks = 4;
  j.bias_off = bias_off;
  for (int mb = 0; mb < 2; ++mb) j.cell[mb] = WCell{2 * mb, 0, 256, 256, 256 * mb, ld, 0, off + (int64_t)mb * 128 * ld};
  return j;
}
static int build_wjobs(const int64_t* off, int in_ch, int in_views, WJobs& jobs, WSplit& split) {
  int n = 0;
  {  // pts_linears[0] (+ its bias) and the skip layer's encoded-points columns W5[:, :in_ch]: both multiply the encoding
    WJob j = {};
    j.a_slab[0] = 0; j.a_n[0] = 4;
    j.a_slab[1] = 20; j.a_n[1] = 4;
    j.b_slab[0] = kActSlabEncP; j.b_n[0] = 1;
    j.n_cells = 4;
    j.bias_blocks = 4;
    j.bias_off = off[8];
    for (int mb = 0; mb < 2; ++mb) {
      j.cell[mb] = WCell{2 * mb, 0, 64, in_ch, 64 * mb, 64, 1, kWScr0 + mb * 128 * 64};
      j.cell[2 + mb] = WCell{4 + 2 * mb, 0, 64, in_ch, 128 + 64 * mb, 64, 1, kWScr5e + mb * 128 * 64};
    }
    jobs.job[n++] = j;
  }
  for (int l = 1; l < 8; ++l) {
    if (l == 5) {
      jobs.job[n++] = square_job(20, 16, kWScr5, 256, off[13]);                              // W5[:, in_ch:]  (h4)
      for (int mb = 0; mb < 2; ++mb) jobs.job[n - 1].cell[mb].scratch = 1;
    } else {
      jobs.job[n++] = square_job(4 * l, 4 * (l - 1), off[l], 256, off[8 + l]);
    }
  }
  jobs.job[n++] = square_job(kDaSlabFeat, 28, off[16], 256, off[17]);                       // feature_linear (h7)
  {  // views_linears[0]: [feature | encoded dirs] against d hv
    WJob j = {};
    j.a_slab[0] = kDaSlabHv; j.a_n[0] = 2;
    j.b_slab[0] = kActSlabFeat; j.b_n[0] = 4;
    j.b_slab[1] = kActSlabEncD; j.b_n[1] = 1;
    j.n_cells = 2;
    j.bias_blocks = 2;
    j.bias_off = off[21];
    j.cell[0] = WCell{0, 0, 256, 256, 0, 256, 1, kWScrV};
    j.cell[1] = WCell{0, 4, 64, in_views, 256, 64, 1, kWScrVd};
    jobs.job[n++] = j;
  }
  if (n != kNumWJobs) return -1;
  // Cost of a job per tile = the 8 KB operand blocks it streams, times a weight in 1/64 for how fast its stages stream
  // (7 / 8 / 9 blocks in flight per stage); NB_WGRAD_W="w7,w9" (floats) overrides the weights of the two odd shapes.
  static int w64[10] = {0};
  if (w64[8] == 0) {
    double w7 = 1.0, w9 = 1.0;
    if (const char* e = getenv("NB_WGRAD_W")) sscanf(e, "%lf,%lf", &w7, &w9);
    for (int b = 0; b < 10; ++b) w64[b] = 64;
    w64[7] = (int)(64 * w7 + 0.5); w64[9] = (int)(64 * w9 + 0.5);
  }
  split.cum[0] = 0;
  for (int i = 0; i < kNumWJobs; ++i) {
    const WJob& j = jobs.job[i];
    const int blocks = j.a_n[0] + j.a_n[1] + j.b_n[0] + j.b_n[1];
    split.cum[i + 1] = split.cum[i] + (int64_t)blocks * w64[blocks < 10 ? blocks : 8];
  }
  return 0;
}

// the padded scratch of a call: the tail of its dA buffer, zeroed on the stream the weight-gradient kernel runs on
static float* wgrad_scratch(void* dA, int64_t tiles, cudaStream_t st) {
  float* scratch = reinterpret_cast<float*>((uint8_t*)dA + (size_t)tiles * kDaSlabs * kSlabBytes);
  cudaMemsetAsync(scratch, 0, sizeof(float) * (size_t)kWScratchFloats, st);
  return scratch;
}
static int launch_unpad(const float* scratch, float* grads, const int64_t* off, int in_ch, int in_views, cudaStream_t st) {
  WUnpad u;
  const int start[6] = {kWScr5, kWScr0, kWScr5e, kWScrV, kWScrVd, kWScratchFloats};
  for (int k = 0; k < 6; ++k) u.start[k] = start[k];
  const int lg[5] = {8, 6, 6, 8, 6}, valid[5] = {256, in_ch, in_ch, 256, in_views};
  const int ld[5] = {256 + in_ch, in_ch, 256 + in_ch, 256 + in_views, 256 + in_views};
  const int64_t dst[5] = {off[5] + in_ch, off[0], off[5], off[20], off[20] + 256};
  for (int k = 0; k < 5; ++k) { u.cols_log2[k] = lg[k]; u.valid[k] = valid[k]; u.ld[k] = ld[k]; u.off[k] = dst[k]; }
  wgrad_unpad_kernel<<<(kWScratchFloats + 255) / 256, 256, 0, st>>>(scratch, grads, u);
  return check_launch("wgrad_unpad_kernel");
}

static int launch_small_grads(const int64_t* off, const void* acts, const float* d_raw, int64_t M, int64_t tiles, float* grads,
                              cudaStream_t st) {
  SmallOffsets o;
  for (int l = 0; l < 8; ++l) o.pts_b[l] = off[8 + l];
  o.feature_b = off[17]; o.views_b = off[21]; o.alpha_w = off[18]; o.alpha_b = off[19];
  o.rgb_w = off[22]; o.rgb_b = off[23];
  NB_REQUIRE(((uintptr_t)d_raw & 15) == 0, "nb_mlp_bwd_tc: d_raw must be 16-byte aligned");
  const int grid = (int)(tiles < kNumSMs ? tiles : kNumSMs);
  mlp_small_grads_kernel<<<grid, 288, kHgSmemBytes, st>>>((const uint8_t*)acts, d_raw, M, (int)tiles, o, grads);
  return check_launch("mlp_small_grads_kernel");
}

// per-device auxiliary stream + events of the streamed backward (library-owned state, created lazily)
struct AuxStream {
  cudaStream_t s = nullptr;
  cudaEvent_t fork = nullptr, join = nullptr;
};
static AuxStream* aux_for_device() {
  static AuxStream aux[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  AuxStream& a = aux[dev];
  if (!a.s) {
    if (cudaStreamCreateWithFlags(&a.s, cudaStreamNonBlocking) != cudaSuccess) return nullptr;
    if (cudaEventCreateWithFlags(&a.fork, cudaEventDisableTiming) != cudaSuccess) return nullptr;
    if (cudaEventCreateWithFlags(&a.join, cudaEventDisableTiming) != cudaSuccess) return nullptr;
  }
  return &a;
}

}  // namespace nb

using namespace nb;

extern "C" {

int64_t nb_mlp_bwd_packed_bytes(int L_pts, int L_dirs) {
  (void)L_pts; (void)L_dirs;
  return (int64_t)kBwdSlabs * kSlabBytes + kBwdConstFloats * 4;
}

// the dA tiles + the padded weight-gradient scratch of the call (see g_wgrad scratch layout above)
int64_t nb_mlp_da_bytes(int64_t M) {
  return ((M + kTileM - 1) / kTileM) * (int64_t)kDaSlabs * kSlabBytes + (int64_t)sizeof(float) * kWScratchFloats;
}

int nb_mlp_pack_bwd(const NbMlpWeights* w, void* packed_bwd, void* stream) {
  NB_REQUIRE(w && packed_bwd, "nb_mlp_pack_bwd: null pointer");
  NB_REQUIRE(((uintptr_t)packed_bwd & 15) == 0, "nb_mlp_pack_bwd: buffer must be 16-byte aligned");
  PackArgsT a;
  int rc = build_pack_args_t(w, a);
  if (rc) return rc;
  const int threads = kBwdSlabs * 128 * 8;
  pack_bwd_kernel<<<(threads + 255) / 256, 256, 0, as_stream(stream)>>>(a, *w, (uint8_t*)packed_bwd);
  return check_launch("nb_mlp_pack_bwd");
}

int nb_mlp_bwd_tc(const void* packed_bwd, int L_pts, int L_dirs, const void* acts, const float* d_raw, int64_t M,
                  void* dA, const int64_t* grad_off_host24, float* grads, void* stream) {
  return nb_mlp_bwd_tc_parts(packed_bwd, L_pts, L_dirs, acts, d_raw, M, dA, grad_off_host24, grads, 3, stream);
}

int nb_mlp_bwd_tc_parts(const void* packed_bwd, int L_pts, int L_dirs, const void* acts, const float* d_raw, int64_t M,
                        void* dA, const int64_t* grad_off_host24, float* grads, int parts, void* stream) {
  if (M == 0) return 0;
  NB_REQUIRE(parts >= 1 && parts <= 3, "nb_mlp_bwd_tc_parts: parts must be 1 (dgrad), 2 (wgrad + heads) or 3 (both)");
  NB_REQUIRE(packed_bwd && acts && d_raw && dA, "nb_mlp_bwd_tc: null pointer");
  if (L_pts != 10 || L_dirs != 4) {
    set_error("nb_mlp_bwd_tc: only multires=10 / multires_views=4 is instantiated (got %d, %d)", L_pts, L_dirs);
    return NB_E_UNSUPPORTED;
  }
  const int in_ch = 3 + 6 * L_pts, in_views = 3 + 6 * L_dirs;
  cudaStream_t st = as_stream(stream);
  const int64_t tiles = (M + kTileM - 1) / kTileM;
  const int64_t quads = (tiles + 3) / 4;
  NB_REQUIRE(quads < (1ll << 30), "nb_mlp_bwd_tc: M too large for one launch");
  if (parts & 1) {
    CUtensorMap wmap;
    int rc = make_slab_tmap(&wmap, packed_bwd, kBwdSlabs);
    if (rc) return rc;
    int clusters = (int)(quads < kNumSMs / 2 ? quads : kNumSMs / 2);
    if (limit_mlp_clusters() > 0 && clusters > limit_mlp_clusters()) clusters = limit_mlp_clusters();
    mlp_dgrad_kernel<<<2 * clusters, kBwdThreads, kBSmemBytes, st>>>(wmap, (const uint8_t*)packed_bwd, (const uint8_t*)acts, d_raw,
                                                             M, (uint8_t*)dA, (int)quads, nullptr);
    rc = check_launch("mlp_dgrad_kernel");
    if (rc) return rc;
  }
  if (!grads || !(parts & 2)) return 0;  // frozen weights / dgrad-only call: only the dA chain was wanted
  NB_REQUIRE(grad_off_host24, "nb_mlp_bwd_tc: gradient offsets missing");
  const int64_t* off = grad_off_host24;  // 0..7 pts_w, 8..15 pts_b, 16 feat_w, 17 feat_b, 18 alpha_w, 19 alpha_b,
                                         // 20 views_w, 21 views_b, 22 rgb_w, 23 rgb_b
  {
    WJobs jobs;
    WSplit split;
    NB_REQUIRE(build_wjobs(off, in_ch, in_views, jobs, split) == 0, "internal: wgrad job table");
    // at least ~4 tiles of the average job per CTA, else fewer CTAs (the flush is a fixed cost per range)
    int64_t grid = split.cum[kNumWJobs] * tiles / (4 * (split.cum[kNumWJobs] / kNumWJobs));
    if (grid > kNumSMs) grid = kNumSMs;
    if (const char* e = getenv("NB_WGRAD_GRID")) {   // experiment knob (tools/wgrad_sm_sweep.sh): fewer CTAs
      const int64_t lim = atoi(e);
      if (lim > 0 && lim < grid) grid = lim;
    }
    if (grid < 1) grid = 1;
    float* scratch = wgrad_scratch(dA, tiles, st);
    if (limit_wgrad_ctas() > 0 && grid > limit_wgrad_ctas()) {
      // SM-limited launch (a concurrent kernel owns the other SMs): CTAs are placed as pairs so that they fill whole
      // TPCs and leave the remaining TPCs free for the 2-CTA clusters of the forward / dgrad kernels
      grid = limit_wgrad_ctas() & ~1;
      if (grid < 2) grid = 2;
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3((unsigned)grid);
      cfg.blockDim = dim3(384);
      cfg.dynamicSmemBytes = kWSmemBytes;
      cfg.stream = st;
      cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeClusterDimension;
      at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
      cfg.attrs = at;
      cfg.numAttrs = 1;
      cudaLaunchKernelEx(&cfg, mlp_wgrad_kernel<false>, jobs, split, WStream{}, (const uint8_t*)acts, (const uint8_t*)dA, (int)tiles, grads,
                         (const int*)nullptr, scratch);
    } else {
      mlp_wgrad_kernel<false><<<(int)grid, 384, kWSmemBytes, st>>>(jobs, split, WStream{}, (const uint8_t*)acts, (const uint8_t*)dA,
                                                                   (int)tiles, grads, nullptr, scratch);
    }
    int rc = check_launch("mlp_wgrad_kernel");
    if (rc) return rc;
    rc = launch_unpad(scratch, grads, off, in_ch, in_views, st);
    if (rc) return rc;
  }
  {
    SmallOffsets o;
    for (int l = 0; l < 8; ++l) o.pts_b[l] = off[8 + l];
    o.feature_b = off[17]; o.views_b = off[21]; o.alpha_w = off[18]; o.alpha_b = off[19];
    o.rgb_w = off[22]; o.rgb_b = off[23];
    NB_REQUIRE(((uintptr_t)d_raw & 15) == 0, "nb_mlp_bwd_tc: d_raw must be 16-byte aligned");
    const int grid = (int)(tiles < kNumSMs ? tiles : kNumSMs);
    mlp_small_grads_kernel<<<grid, 288, kHgSmemBytes, st>>>((const uint8_t*)acts, d_raw, M, (int)tiles, o, grads);
    int rc = check_launch("mlp_small_grads_kernel");
    if (rc) return rc;
  }
  return 0;
}

// Streamed backward: the dgrad chain on `dgrad_clusters` CTA pairs and, BESIDE it on an auxiliary stream, the
// weight-gradient kernel on `wgrad_ctas` CTAs (placed as pairs on the other TPCs) consuming each tile's activation
// gradients as soon as the chain has written them (per-tile arrival counters in `ready`, [tiles] ints, zeroed here),
// i.e. while they are still in L2: the weight-gradient pass reads dA from L2 instead of HBM and its bandwidth-bound time
// overlaps the shared-memory-bound chain.  Same results as nb_mlp_bwd_tc (the accumulation order of the atomics differs).
int nb_mlp_bwd_tc_streamed(const void* packed_bwd, int L_pts, int L_dirs, const void* acts, const float* d_raw, int64_t M,
                           void* dA, const int64_t* grad_off_host24, float* grads, int* ready, int dgrad_clusters,
                           int wgrad_ctas, void* stream) {
  if (M == 0) return 0;
  NB_REQUIRE(packed_bwd && acts && d_raw && dA && grad_off_host24 && grads && ready, "nb_mlp_bwd_tc_streamed: null pointer");
  if (L_pts != 10 || L_dirs != 4) {
    set_error("nb_mlp_bwd_tc_streamed: only multires=10 / multires_views=4 is instantiated (got %d, %d)", L_pts, L_dirs);
    return NB_E_UNSUPPORTED;
  }
  const int in_ch = 3 + 6 * L_pts, in_views = 3 + 6 * L_dirs;
  cudaStream_t st = as_stream(stream);
  const int64_t tiles = (M + kTileM - 1) / kTileM;
  const int64_t quads = (tiles + 3) / 4;
  NB_REQUIRE(quads < (1ll << 30), "nb_mlp_bwd_tc_streamed: M too large for one launch");
  if (dgrad_clusters <= 0) dgrad_clusters = kNumSMs / 4;          // 37 pairs
  if (wgrad_ctas <= 0) wgrad_ctas = kNumSMs / 2;                  // 74 CTAs
  wgrad_ctas &= ~1;
  NB_REQUIRE(dgrad_clusters >= 1 && wgrad_ctas >= 24 && 2 * dgrad_clusters + wgrad_ctas <= kNumSMs,
             "nb_mlp_bwd_tc_streamed: the two kernels must fit the GPU side by side (%d pairs + %d CTAs)", dgrad_clusters, wgrad_ctas);
  AuxStream* aux = aux_for_device();
  NB_REQUIRE(aux != nullptr, "nb_mlp_bwd_tc_streamed: could not create the auxiliary stream");
  WJobs jobs;
  WSplit split;
  NB_REQUIRE(build_wjobs(grad_off_host24, in_ch, in_views, jobs, split) == 0, "internal: wgrad job table");
  // CTAs per job in proportion to the blocks a job streams per tile (largest remainder), at least one each
  WStream ws;
  {
    const int64_t total = split.cum[kNumWJobs];
    int given = 0, cnt[kNumWJobs];
    double rem[kNumWJobs];
    for (int j = 0; j < kNumWJobs; ++j) {
      const double share = (double)(split.cum[j + 1] - split.cum[j]) * wgrad_ctas / (double)total;
      cnt[j] = (int)share < 1 ? 1 : (int)share;
      rem[j] = share - cnt[j];
      given += cnt[j];
    }
    while (given < wgrad_ctas) {
      int best = 0;
      for (int j = 1; j < kNumWJobs; ++j) if (rem[j] > rem[best]) best = j;
      ++cnt[best]; rem[best] -= 1.0; ++given;
    }
    while (given > wgrad_ctas) {
      int best = -1;
      for (int j = 0; j < kNumWJobs; ++j) if (cnt[j] > 1 && (best < 0 || rem[j] < rem[best])) best = j;
      --cnt[best]; rem[best] += 1.0; --given;
    }
    int b = 0;
    for (int j = 0; j < kNumWJobs; ++j) {
      ws.parts[j] = (unsigned char)cnt[j];
      for (int p = 0; p < cnt[j]; ++p, ++b) { ws.job[b] = (unsigned char)j; ws.part[b] = (unsigned char)p; }
    }
  }
  cudaMemsetAsync(ready, 0, sizeof(int) * (size_t)tiles, st);
  cudaEventRecord(aux->fork, st);
  cudaStreamWaitEvent(aux->s, aux->fork, 0);
  {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)wgrad_ctas);
    cfg.blockDim = dim3(384);
    cfg.dynamicSmemBytes = kWSmemBytes;
    cfg.stream = aux->s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    float* scratch = wgrad_scratch(dA, tiles, aux->s);
    cudaLaunchKernelEx(&cfg, mlp_wgrad_kernel<true>, jobs, split, ws, (const uint8_t*)acts, (const uint8_t*)dA, (int)tiles, grads,
                       (const int*)ready, scratch);
    int rc = check_launch("mlp_wgrad_kernel<stream>");
    if (rc) return rc;
    rc = launch_unpad(scratch, grads, grad_off_host24, in_ch, in_views, aux->s);
    if (rc) return rc;
  }
  {
    CUtensorMap wmap;
    int rc = make_slab_tmap(&wmap, packed_bwd, kBwdSlabs);
    if (rc) return rc;
    int clusters = (int)(quads < dgrad_clusters ? quads : dgrad_clusters);
    mlp_dgrad_kernel<<<2 * clusters, kBwdThreads, kBSmemBytes, st>>>(wmap, (const uint8_t*)packed_bwd, (const uint8_t*)acts, d_raw,
                                                             M, (uint8_t*)dA, (int)quads, ready);
    rc = check_launch("mlp_dgrad_kernel");
    if (rc) return rc;
  }
  int rc = launch_small_grads(grad_off_host24, acts, d_raw, M, tiles, grads, aux->s);
  if (rc) return rc;
  cudaEventRecord(aux->join, aux->s);
  cudaStreamWaitEvent(st, aux->join, 0);
  return 0;
}

int nb_debug_wgrad_trace(unsigned long long* host_out, int n) {
  NB_REQUIRE(host_out && n >= 1 && n <= 2 * kNumSMs, "nb_debug_wgrad_trace: bad arguments");
  cudaError_t e = cudaMemcpyFromSymbol(host_out, g_wgrad_trace, sizeof(unsigned long long) * (size_t)n);
  if (e != cudaSuccess) {
    set_error("nb_debug_wgrad_trace: %s", cudaGetErrorString(e));
    return (int)e;
  }
  return 0;
}

int nb_debug_umma_mn(const float* A, const float* B, int K, float* D, void* stream) {
  NB_REQUIRE(A && B && D && (K == 64 || K == 128), "nb_debug_umma_mn: bad arguments");
  umma_mn_selftest_kernel<<<1, 128, 4 * K * 128 + 2048, as_stream(stream)>>>(A, B, K, D);
  return check_launch("nb_debug_umma_mn");
}

}  // extern "C"
