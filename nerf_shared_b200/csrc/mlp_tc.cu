// The NeRF MLP on the 5th-generation tensor cores: weight packing, a UMMA self-test and the
// persistent fused forward kernel.
//
// replaces NeRF.forward / MLP (nerf.py:96-134) including both Embedder calls (nerf.py:98,103)
// and the netchunk loop (nerf.py:106): the encodings are computed in registers and written
// straight into the first layer's shared-memory operand, activations stay on the SM for all
// ten GEMM layers, and only raw [M,4] (and, for training, the bf16 activations the backward
// needs) reach HBM.
//
// Tiling.  A tile is 128 samples (UMMA M = 128, one TMEM lane per sample).  A CTA works on two
// tiles A/B at a time so that the epilogue of one overlaps the MMAs of the other; each tile
// owns 256 TMEM columns (128 x 256 fp32 accumulator), a 64 KB activation operand (4 K-slabs of
// [128 x 64] bf16, K-major, 128-byte swizzle) and one 16 KB auxiliary slab that holds the
// position encoding (layers 0 and 5) and later the view-direction encoding (view branch).
// Weights are pre-packed (nb_mlp_pack) into 16 KB slabs [128 out x 64 in] in exactly the
// shared-memory image the UMMA descriptor expects, in consumption order, and streamed
// L2 -> SMEM with 1-D bulk async copies through a 3-stage mbarrier ring.
//
// Warp roles (640 threads): warp 0 = weight producer, warp 1 = MMA issuer, warp 2 = TMEM
// allocator, warps 4-11 = epilogue of tile A, warps 12-19 = epilogue of tile B (two warps per TMEM lane
// quarter, each taking half of the 256 accumulator columns).
#include "common.cuh"
#include "ptx.cuh"
#include "mlp_common.cuh"
#include "tmap.cuh"
#include "encode.cuh"

namespace nb {

using namespace ptx;

// fp32 side table offsets (floats)
constexpr int kOffBias = 0;        // 8 x 256 trunk biases
constexpr int kOffFeatB = 2048;    // 256
constexpr int kOffViewB = 2304;    // 128
constexpr int kOffAlphaW = 2432;   // 256
constexpr int kOffRgbW = 2688;     // 3 x 128
constexpr int kOffRgbB = 3072;     // 3
constexpr int kOffAlphaB = 3075;   // 1

// shared memory map (bytes from a 1024-aligned base)
constexpr uint32_t kSmAct0 = 0, kSmEnc0 = 65536, kSmAct1 = 81920, kSmEnc1 = 147456;
constexpr uint32_t kSmRing = 163840;
constexpr uint32_t kSmConst = kSmRing + kStages * kSlabBytes;  // 212992
constexpr uint32_t kSmBar = kSmConst + kConstFloats * 4;       // 225792
constexpr uint32_t kSmTmemPtr = kSmBar + 120;
constexpr uint32_t kSmPart = kSmBar + 128;                     // [2 tiles][128 rows] float4 head partial sums
constexpr uint32_t kSmemBytes = kSmPart + 4096 + 1024;         // + alignment slack
constexpr int kFwdThreads = 640;                               // 4 control warps + 16 epilogue warps

__host__ __device__ constexpr int ks_count(int layer) { return layer == 0 ? 1 : ((layer == 5 || layer == 9) ? 5 : 4); }
__host__ __device__ constexpr int nh_count(int layer) { return layer == 9 ? 1 : 2; }
__host__ __device__ constexpr int slab_base(int layer) {
  int b = 0;
  for (int l = 0; l < layer; ++l) b += ks_count(l) * nh_count(l);
  return b;
}
static_assert(slab_base(10) == kNumSlabs, "slab schedule");

// ------------------------------------------------------------------------------------------
// weight packing
// ------------------------------------------------------------------------------------------
struct SlabSrc {
  const float* w;  // fp32 [out,in] matrix
  int ld;          // in-features (row stride)
  int n0;          // first output row of this slab
  int k0;          // first input column
  int kvalid;      // columns actually present (rest is zero padding)
};
struct PackArgs {
  SlabSrc slab[kNumSlabs];
};

__global__ void __launch_bounds__(256) pack_weights_kernel(PackArgs a, NbMlpWeights w, uint8_t* __restrict__ blob) {
  // one thread per 16-byte chunk: slab s, row n (0..127), chunk g (0..7)
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx < kNumSlabs * 128 * 8) {
    int s = idx / 1024, rem = idx - s * 1024;
    int n = rem >> 3, g = rem & 7;
    const SlabSrc src = a.slab[s];
    const float* row = src.w + (size_t)(src.n0 + n) * src.ld + src.k0;
    uint32_t p[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int k = g * 8 + j * 2;
      float lo = (k < src.kvalid) ? row[k] : 0.f;
      float hi = (k + 1 < src.kvalid) ? row[k + 1] : 0.f;
      p[j] = pack_bf16x2(lo, hi);
    }
    *reinterpret_cast<uint4*>(blob + (size_t)s * kSlabBytes + sw128_chunk_off(n, g)) = make_uint4(p[0], p[1], p[2], p[3]);
  }
  // fp32 side table
  float* c = reinterpret_cast<float*>(blob + (size_t)kNumSlabs * kSlabBytes);
  if (idx < kConstFloats) {
    float v = 0.f;
    if (idx < 2048) v = w.pts_b[idx >> 8][idx & 255];
    else if (idx < kOffViewB) v = w.feature_b[idx - kOffFeatB];
    else if (idx < kOffAlphaW) v = w.views_b[idx - kOffViewB];
    else if (idx < kOffRgbW) v = w.alpha_w[idx - kOffAlphaW];
    else if (idx < kOffRgbB) v = w.rgb_w[idx - kOffRgbW];
    else if (idx < kOffAlphaB) v = w.rgb_b[idx - kOffRgbB];
    else if (idx == kOffAlphaB) v = w.alpha_b[0];
    c[idx] = v;
  }
}

// ------------------------------------------------------------------------------------------
// UMMA self-test: D[128,128] = A[128,K] * B[128,K]^T through the same swizzle helper,
// descriptors, tcgen05.mma / commit / ld sequence as the fused kernel (K multiple of 64 <= 256).
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128, 1) umma_selftest_kernel(const float* __restrict__ A, const float* __restrict__ B,
                                                               int K, float* __restrict__ D) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gbase = smem_raw + (base - smem_u32(smem_raw));
  const int nslab = K / 64;
  const uint32_t sA = base, sB = base + nslab * kSlabBytes, sBar = sB + nslab * kSlabBytes, sPtr = sBar + 8;
  const int row = threadIdx.x, warp = threadIdx.x >> 5;
  for (int s = 0; s < nslab; ++s)
    for (int g = 0; g < 8; ++g) {
      uint32_t pa[4], pb[4];
      for (int j = 0; j < 4; ++j) {
        int k = s * 64 + g * 8 + j * 2;
        pa[j] = pack_bf16x2(A[row * K + k], A[row * K + k + 1]);
        pb[j] = pack_bf16x2(B[row * K + k], B[row * K + k + 1]);
      }
      st_shared_v4(sA + s * kSlabBytes + sw128_chunk_off(row, g), pa[0], pa[1], pa[2], pa[3]);
      st_shared_v4(sB + s * kSlabBytes + sw128_chunk_off(row, g), pb[0], pb[1], pb[2], pb[3]);
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
    const uint32_t idesc = make_idesc_bf16(128, 128);
    for (int s = 0; s < nslab; ++s) {
      uint64_t ad = make_desc_kmajor_sw128(sA + s * kSlabBytes), bd = make_desc_kmajor_sw128(sB + s * kSlabBytes);
      for (int k16 = 0; k16 < 4; ++k16) umma_bf16(tmem, ad + 2 * k16, bd + 2 * k16, idesc, (s | k16) != 0);
    }
    umma_commit(sBar);
  }
  mbar_wait(sBar, 0);
  tc_fence_after();
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

template <int N>
__device__ __forceinline__ void store_operand_row(uint32_t slab, int row, const float (&e)[N]) {
#pragma unroll
  for (int g = 0; g < N / 8; ++g) {
    const uint32_t p0 = pack_bf16x2(e[8 * g], e[8 * g + 1]), p1 = pack_bf16x2(e[8 * g + 2], e[8 * g + 3]),
                   p2 = pack_bf16x2(e[8 * g + 4], e[8 * g + 5]), p3 = pack_bf16x2(e[8 * g + 6], e[8 * g + 7]);
    const uint32_t off = sw128_chunk_off(row, g);
    st_shared_v4(slab + off, p0, p1, p2, p3);
  }
}

// ------------------------------------------------------------------------------------------
// epilogues: TMEM accumulator row -> bias (+relu) -> bf16 -> next layer's A operand in SMEM
// ------------------------------------------------------------------------------------------
// MODE 0: relu hidden layer; 1: relu hidden layer + fp32 sigma head; 2: linear (feature layer)
// A warp handles `nchunks` 32-column chunks starting at chunk `cbeg` of its 32 rows (two warps share a
// lane quarter and split the 256 columns).  The bias row of a chunk is fetched between the TMEM load and
// its wait, so the shared-memory latency hides behind the tcgen05.ld.
template <int MODE>
__device__ __forceinline__ void epilogue_hidden(uint32_t taddr, const float* __restrict__ bias,
                                                const float* __restrict__ alpha_w, uint32_t act, int row, int cbeg,
                                                float& sigma, uint32_t* __restrict__ mask_out) {
#pragma unroll
  for (int cc = 0; cc < 4; ++cc) {
    const int c = cbeg + cc;
    uint32_t v[32];
    tmem_ld32(taddr + c * 32, v);
    float4 bq[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) bq[j] = *reinterpret_cast<const float4*>(bias + c * 32 + j * 4);
    tmem_ld_wait();
    uint32_t neg = 0;   // sign bits of the 32 pre-activations, column j at bit 31 - j
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      float f[8];
#pragma unroll
      for (int j = 0; j < 8; j += 4) {
        const float4 b = bq[g * 2 + (j >> 2)];
        f[j] = __uint_as_float(v[g * 8 + j]) + b.x;
        f[j + 1] = __uint_as_float(v[g * 8 + j + 1]) + b.y;
        f[j + 2] = __uint_as_float(v[g * 8 + j + 2]) + b.z;
        f[j + 3] = __uint_as_float(v[g * 8 + j + 3]) + b.w;
      }
      if (MODE != 2 && mask_out) {
#pragma unroll
        for (int j = 0; j < 8; ++j) neg = __funnelshift_l(__float_as_uint(f[j]), neg, 1);   // one SHF per value
      }
      if (MODE == 1) {
#pragma unroll
        for (int j = 0; j < 8; j += 4) {
          float4 a = *reinterpret_cast<const float4*>(alpha_w + c * 32 + g * 8 + j);
          sigma = fmaf(fmaxf(f[j], 0.f), a.x, sigma);
          sigma = fmaf(fmaxf(f[j + 1], 0.f), a.y, sigma);
          sigma = fmaf(fmaxf(f[j + 2], 0.f), a.z, sigma);
          sigma = fmaf(fmaxf(f[j + 3], 0.f), a.w, sigma);
        }
      }
      uint32_t p0, p1, p2, p3;
      if (MODE == 2) {
        p0 = pack_bf16x2(f[0], f[1]); p1 = pack_bf16x2(f[2], f[3]);
        p2 = pack_bf16x2(f[4], f[5]); p3 = pack_bf16x2(f[6], f[7]);
      } else {
        p0 = pack_bf16x2_relu(f[0], f[1]); p1 = pack_bf16x2_relu(f[2], f[3]);
        p2 = pack_bf16x2_relu(f[4], f[5]); p3 = pack_bf16x2_relu(f[6], f[7]);
      }
      const uint32_t off = (c >> 1) * kSlabBytes + sw128_chunk_off(row, (c & 1) * 4 + g);
      st_shared_v4(act + off, p0, p1, p2, p3);
    }
    if (MODE != 2 && mask_out) mask_out[c * 128 + row] = ~neg;
  }
}

// view branch: hv = relu(acc + b) (128 wide), rgb = hv . Wrgb^T in fp32 (nerf.py:125-129); this warp's two
// chunks [cbeg, cbeg + 2), partial sums returned in r / g_ / b_ (the caller adds the bias and the other half)
__device__ __forceinline__ void epilogue_views(uint32_t taddr, const float* __restrict__ cst, float& r, float& g_,
                                               float& b_, int row, int cbeg, uint32_t act, bool save,
                                               uint32_t* __restrict__ mask_out) {
  const float* bias = cst + kOffViewB;
  const float* wr = cst + kOffRgbW;
  r = 0.f; g_ = 0.f; b_ = 0.f;
#pragma unroll
  for (int cc = 0; cc < 2; ++cc) {
    const int c = cbeg + cc;
    uint32_t v[32];
    tmem_ld32(taddr + c * 32, v);
    float4 bq[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) bq[j] = *reinterpret_cast<const float4*>(bias + c * 32 + j * 4);
    tmem_ld_wait();
    uint32_t pos = 0;   // sign bits of the pre-activations, column j of the chunk at bit 31 - j (stored inverted)
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      float f[8];
#pragma unroll
      for (int j = 0; j < 8; j += 4) {
        const int col = c * 32 + g * 8 + j;
        const float4 b = bq[g * 2 + (j >> 2)];
        float4 w0 = *reinterpret_cast<const float4*>(wr + col);
        float4 w1 = *reinterpret_cast<const float4*>(wr + 128 + col);
        float4 w2 = *reinterpret_cast<const float4*>(wr + 256 + col);
        f[j] = __uint_as_float(v[g * 8 + j]) + b.x;
        f[j + 1] = __uint_as_float(v[g * 8 + j + 1]) + b.y;
        f[j + 2] = __uint_as_float(v[g * 8 + j + 2]) + b.z;
        f[j + 3] = __uint_as_float(v[g * 8 + j + 3]) + b.w;
        if (save) {
#pragma unroll
          for (int jj = 0; jj < 4; ++jj) pos = __funnelshift_l(__float_as_uint(f[j + jj]), pos, 1);
        }
        f[j] = fmaxf(f[j], 0.f); f[j + 1] = fmaxf(f[j + 1], 0.f); f[j + 2] = fmaxf(f[j + 2], 0.f); f[j + 3] = fmaxf(f[j + 3], 0.f);
        r = fmaf(f[j], w0.x, r); r = fmaf(f[j + 1], w0.y, r); r = fmaf(f[j + 2], w0.z, r); r = fmaf(f[j + 3], w0.w, r);
        g_ = fmaf(f[j], w1.x, g_); g_ = fmaf(f[j + 1], w1.y, g_); g_ = fmaf(f[j + 2], w1.z, g_); g_ = fmaf(f[j + 3], w1.w, g_);
        b_ = fmaf(f[j], w2.x, b_); b_ = fmaf(f[j + 1], w2.y, b_); b_ = fmaf(f[j + 2], w2.z, b_); b_ = fmaf(f[j + 3], w2.w, b_);
      }
      if (save) {
        st_shared_v4(act + (c >> 1) * kSlabBytes + sw128_chunk_off(row, (c & 1) * 4 + g), pack_bf16x2(f[0], f[1]),
                     pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]), pack_bf16x2(f[6], f[7]));
      }
    }
    if (save) mask_out[c * 128 + row] = ~pos;
  }
}

// chunks [4H, 4H + 4) (32 of the 64 columns) of an operand row from the full encoding
template <int H>
__device__ __forceinline__ void store_operand_half(uint32_t slab, int row, const float (&e)[64]) {
#pragma unroll
  for (int gg = 0; gg < 4; ++gg) {
    const int g = 4 * H + gg;
    const uint32_t p0 = pack_bf16x2(e[8 * g], e[8 * g + 1]), p1 = pack_bf16x2(e[8 * g + 2], e[8 * g + 3]),
                   p2 = pack_bf16x2(e[8 * g + 4], e[8 * g + 5]), p3 = pack_bf16x2(e[8 * g + 6], e[8 * g + 7]);
    st_shared_v4(slab + sw128_chunk_off(row, g), p0, p1, p2, p3);
  }
}

__device__ __forceinline__ void pair_sync(int id) { asm volatile("bar.sync %0, 64;" ::"r"(id) : "memory"); }

// ------------------------------------------------------------------------------------------
// the fused forward kernel (CTA pairs, tcgen05 cta_group::2)
// ------------------------------------------------------------------------------------------
// Two CTAs of a cluster work as one MMA unit: M = 256 samples (128 per CTA), N = 256 output
// features, each CTA holding its own 128 activation rows and HALF of every weight slab (the
// leader the output features 0..127, the peer 128..255), so a 16 KB ring stage feeds twice the
// MMA work of the single-CTA form and the L2 -> SMEM weight traffic per SM halves.  Each CTA
// still ping-pongs two 128-sample tiles (x = 0/1) so that epilogues overlap the other tile's MMAs;
// a cluster therefore works on a "quad" of four tiles per iteration.
//
// Barriers (same shared-memory offsets in both CTAs):
//   full[s]   leader only: weight stage s landed in BOTH CTAs (TMA complete_tx from both + expect_tx)
//   empty[s]  per CTA: the MMAs reading stage s retired (tcgen05.commit multicast to both CTAs)
//   act[x]    leader only: tile x's A operand is written in both CTAs (8 warp arrivals, 4 remote)
//   acc[x]    per CTA: tile x's accumulator is complete (tcgen05.commit multicast)
#ifdef NB_TRACE
// timing experiments (make EXTRA_FLAGS=-DNB_TRACE): SM-clock timestamps of CTA 0's first quads,
// slot = ((quad_iter * 10 + layer) * 2 + x) * 8 + event
__device__ long long g_trace[4 * 10 * 2 * 8];
#define NB_TR(qi, layer, x, ev) \
  do { if (blockIdx.x == 0 && (qi) < 4) g_trace[((((qi) * 10 + (layer)) * 2 + (x)) << 3) + (ev)] = clock64(); } while (0)
#else
#define NB_TR(qi, layer, x, ev)
#endif

template <int L_PTS, int L_DIRS>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kFwdThreads, 1)
    mlp_fwd_tc_kernel(const __grid_constant__ CUtensorMap wmap, const uint8_t* __restrict__ blob,
                      const float* __restrict__ pts, const float* __restrict__ viewdirs, int64_t M, int S,
                      float* __restrict__ raw, uint8_t* __restrict__ acts, int num_quads) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gbase = smem_raw + (base - smem_u32(smem_raw));
  const float* cst = reinterpret_cast<const float*>(gbase + kSmConst);
  const uint32_t bar_full = base + kSmBar, bar_empty = bar_full + 8 * kStages, bar_act = bar_empty + 8 * kStages,
                 bar_acc = bar_act + 16, bar_const = bar_acc + 16;
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;  // warp: provably uniform
  const uint32_t rank = cluster_ctarank();
  const int cluster = (int)cluster_id_x(), num_clusters = (int)cluster_count_x();

  if (threadIdx.x == 0) {
    for (int i = 0; i < kStages; ++i) {
      mbar_init(bar_full + 8 * i, 1);   // leader producer's arrive.expect_tx
      mbar_init(bar_empty + 8 * i, 1);  // tcgen05.commit
    }
    for (int x = 0; x < 2; ++x) {
      mbar_init(bar_act + 8 * x, 16);   // one arrival per epilogue warp of the tile (8), both CTAs
      mbar_init(bar_acc + 8 * x, 1);    // tcgen05.commit
    }
    mbar_init(bar_const, 1);
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc_2sm(base + kSmTmemPtr, 512);
    tmem_relinquish_2sm();
  }
  if (warp == 0 && lane == 0) tma_prefetch_desc(&wmap);
  tc_fence_before();
  cluster_sync_all();  // barriers of both CTAs initialised before any remote arrive / TMA signal
  tc_fence_after();
  const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(gbase + kSmTmemPtr);

  // The producer and MMA warps run CONVERGED (all 32 lanes execute the loops and the barrier waits) and
  // issue the asynchronous instructions under elect.sync: with a single provably-active lane and
  // warp-uniform operands the compiler feeds UTMALDG / UTCHMMA / UTCBAR from uniform registers directly;
  // a `lane == 0` branch instead costs an ELECT + R2UR waterfall loop per instruction (~100 cycles per MMA
  // on the issuing thread, which then cannot keep ahead of the 128-cycle MMAs).
  if (warp == 0) {
    // ---------------- weight producer: this CTA's half of every slab ----------------
    if (elect_one()) {
      mbar_expect_tx(bar_const, kConstFloats * 4);
      bulk_g2s(base + kSmConst, blob + (size_t)kNumSlabs * kSlabBytes, kConstFloats * 4, bar_const);
    }
    __syncwarp();
    uint32_t stage = 0, phase = 0;
    const uint32_t full_leader0 = mapa(bar_full, 0);
    for (int quad = cluster; quad < num_quads; quad += num_clusters) {
      for (int layer = 0; layer < 10; ++layer) {
        const int nks = ks_count(layer), sb = slab_base(layer);
        const uint32_t half_bytes = layer == 9 ? 8192u : 16384u;
        for (int x = 0; x < 2; ++x) {
          for (int ks = 0; ks < nks; ++ks) {
            mbar_wait(bar_empty + 8 * stage, phase ^ 1);
            // layers 0..8: slab (ks, nh = rank), 128 rows; layer 9 (N = 128): rows 64*rank.. of slab ks
            const int row0 = layer == 9 ? (sb + ks) * 128 + 64 * (int)rank : (sb + ks * 2 + (int)rank) * 128;
            if (elect_one()) {
              if (rank == 0) mbar_expect_tx(bar_full + 8 * stage, 2 * half_bytes);
              const uint32_t full_leader = full_leader0 + 8 * stage;
              const uint32_t dst = base + kSmRing + stage * kSlabBytes;
              tma_load_2d_2sm(dst, &wmap, 0, row0, full_leader);
              if (layer != 9) tma_load_2d_2sm(dst + 8192, &wmap, 0, row0 + 64, full_leader);
            }
            __syncwarp();
            if (++stage == kStages) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ---------------- MMA issuer (leader CTA only) ----------------
    if (rank == 0) {
      const uint32_t idesc256 = make_idesc_bf16(256, 256), idesc128 = make_idesc_bf16(256, 128);
      const uint64_t desc0 = make_desc_kmajor_sw128(0);
      uint32_t stage = 0, phase = 0, par_act0 = 0, par_act1 = 0;
      int qi = -1;
      for (int quad = cluster; quad < num_quads; quad += num_clusters) {
        ++qi;
        for (int layer = 0; layer < 10; ++layer) {
          const int nks = ks_count(layer);
          const uint32_t idesc = layer == 9 ? idesc128 : idesc256;
          for (int x = 0; x < 2; ++x) {
            const uint32_t sAct = base + (x ? kSmAct1 : kSmAct0), sEnc = base + (x ? kSmEnc1 : kSmEnc0);
            if (x == 0) { mbar_wait_cluster(bar_act, par_act0); par_act0 ^= 1; }
            else { mbar_wait_cluster(bar_act + 8, par_act1); par_act1 ^= 1; }
            tc_fence_after();
            if (lane == 0) NB_TR(qi, layer, x, 0);
            const uint32_t d = tmem + x * 256;
            for (int ks = 0; ks < nks; ++ks) {
              uint32_t a_addr;
              if (layer == 0) a_addr = sEnc;
              else if (layer == 5) a_addr = ks == 0 ? sEnc : sAct + (ks - 1) * kSlabBytes;  // cat[input_pts, h]
              else if (layer == 9 && ks == 4) a_addr = sEnc;                              // cat[feature, dirs]
              else a_addr = sAct + ks * kSlabBytes;
              const bool half_k = layer == 9 && ks == 4;   // 32-wide view encoding: two K=16 steps
              const uint64_t adesc = desc0 | (uint64_t)((a_addr & 0x3FFFFu) >> 4);
              const uint64_t bdesc = desc0 | (uint64_t)(((base + kSmRing + stage * kSlabBytes) & 0x3FFFFu) >> 4);
              mbar_wait(bar_full + 8 * stage, phase);
              tc_fence_after();
              if (elect_one()) {
                umma_bf16_2sm(d, adesc, bdesc, idesc, ks != 0);
                umma_bf16_2sm(d, adesc + 2, bdesc + 2, idesc, 1);
                if (!half_k) {
                  umma_bf16_2sm(d, adesc + 4, bdesc + 4, idesc, 1);
                  umma_bf16_2sm(d, adesc + 6, bdesc + 6, idesc, 1);
                }
                umma_commit_2sm(bar_empty + 8 * stage);  // frees the stage in both CTAs
                if (ks == nks - 1) umma_commit_2sm(bar_acc + 8 * x);  // accumulator of (tile x, layer) complete
              }
              __syncwarp();
              if (++stage == kStages) { stage = 0; phase ^= 1; }
            }
            if (lane == 0) NB_TR(qi, layer, x, 1);
          }
        }
      }
    }
  } else if (warp >= 4) {
    // ---------------- epilogue warps: tile x, row = TMEM lane, column half h ----------------
    // Eight warps per tile: the two warps of a lane quarter split the 256 accumulator columns (h = 0: chunks
    // 0..3, h = 1: chunks 4..7), which halves the accumulator -> operand latency on the MMA's critical chain.
    const int x = (warp - 4) >> 3, h = ((warp - 4) >> 2) & 1, q = warp & 3;
    const int row = (q << 5) | lane;
    const int pair_id = 1 + x * 4 + q;   // named barrier shared by the two warps of (tile, lane quarter)
    const uint32_t sAct = base + (x ? kSmAct1 : kSmAct0), sEnc = base + (x ? kSmEnc1 : kSmEnc0);
    const uint32_t taddr = tmem + x * 256 + ((uint32_t)(q * 32) << 16);
    const uint32_t act_leader = mapa(bar_act + 8 * x, 0), my_acc = bar_acc + 8 * x;
    float4* part = reinterpret_cast<float4*>(gbase + kSmPart) + x * 128 + row;
    uint32_t par_acc = 0;
    mbar_wait(bar_const, 0);
    int qi = -1;
    for (int quad = cluster; quad < num_quads; quad += num_clusters) {
      ++qi;
      const int64_t tile = (int64_t)quad * 4 + rank * 2 + x;
      const int64_t m = tile * kTileM + row;
      const bool valid = m < M;
      // whole tiles are saved (rows past M hold finite values computed from zero inputs).  The SMEM
      // operand image IS the saved image, so every warp streams its own 32 rows of its own two slabs
      // to HBM with bulk async stores; `flush` follows the fence that publishes the rows to the MMA.
      uint8_t* save = (acts && tile * kTileM < M) ? acts + (size_t)tile * kActSlabs * kSlabBytes : nullptr;
      // relu' bit masks of this tile (after all activation tiles of the launch)
      uint32_t* mask = save ? reinterpret_cast<uint32_t*>(acts + (size_t)((M + kTileM - 1) / kTileM) * kActSlabs * kSlabBytes +
                                                          (size_t)tile * kMaskTileBytes)
                            : nullptr;
#if defined(NB_EXP) && (NB_EXP & 4)   // ablation build: every CTA stores to its own first tile (L2-resident target)
      if (save) save = acts + (size_t)((blockIdx.x * 2 + x) & 31) * kActSlabs * kSlabBytes;
#endif
#if defined(NB_EXP) && (NB_EXP & 2)   // ablation build: no relu' masks
      mask = nullptr;
#endif
      const uint32_t wrow = (uint32_t)q * 4096u;
      auto flush = [&](uint32_t src_slab0, int dst_slab0, int nslabs) {
#if defined(NB_EXP) && (NB_EXP & 1)   // ablation build: no activation stores (timing only, results are wrong)
        return;
#endif
        if (save && lane == 0) {
          for (int j = 0; j < nslabs; ++j)
            bulk_s2g(save + (size_t)(dst_slab0 + j) * kSlabBytes + wrow, src_slab0 + j * kSlabBytes + wrow, 4096);
          bulk_commit();
        }
      };
      auto reuse = [&]() {  // before overwriting SMEM that earlier bulk stores may still be reading
        if (save) {
          if (lane == 0) bulk_wait_read_all();
          __syncwarp();
        }
      };
      // the encoding slab is flushed by the h = 0 warp but written by both: order the other warp's bulk reads /
      // stores against it with the pair barrier
      reuse();
      if (save) pair_sync(pair_id);
      {
        float p0 = 0.f, p1 = 0.f, p2 = 0.f;
        if (valid) { p0 = pts[m * 3]; p1 = pts[m * 3 + 1]; p2 = pts[m * 3 + 2]; }
        float e[64];
        encode3<L_PTS>(p0, p1, p2, e);
        if (h == 0) store_operand_half<0>(sEnc, row, e);
        else store_operand_half<1>(sEnc, row, e);
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive_remote(act_leader);
      if (save) {
        pair_sync(pair_id);
        if (h == 0) flush(sEnc, kActSlabEncP, 1);
      }
      float sigma = 0.f;
#pragma unroll 1
      for (int layer = 0; layer < 9; ++layer) {
        mbar_wait(my_acc, par_acc);
        par_acc ^= 1;
        tc_fence_after();
        if (q == 0 && h == 0 && lane == 0) NB_TR(qi, layer, x, 2);
        reuse();
        if (q == 0 && h == 0 && lane == 0) NB_TR(qi, layer, x, 5);
        uint32_t* mk = mask ? mask + layer * 1024 : nullptr;
        if (layer < 7) epilogue_hidden<0>(taddr, cst + kOffBias + layer * 256, nullptr, sAct, row, 4 * h, sigma, mk);
        else if (layer == 7) epilogue_hidden<1>(taddr, cst + kOffBias + 7 * 256, cst + kOffAlphaW, sAct, row, 4 * h, sigma, mk);
        else {
          epilogue_hidden<2>(taddr, cst + kOffFeatB, nullptr, sAct, row, 4 * h, sigma, nullptr);
          // view-direction encoding replaces the position encoding in the auxiliary slab
          // (its last reader, layer 5's MMA, retired before layer 5's accumulator was signalled)
          if (save) pair_sync(pair_id);   // the h = 0 warp's flush of enc(pts) has been waited for (reuse above)
          if (h == 0) {
            float v0 = 0.f, v1 = 0.f, v2 = 0.f;
            if (valid) {
              const int64_t ray = m / S;
              v0 = viewdirs[ray * 3]; v1 = viewdirs[ray * 3 + 1]; v2 = viewdirs[ray * 3 + 2];
            }
            float e[64];
            encode3<L_DIRS>(v0, v1, v2, e);
            store_operand_half<0>(sEnc, row, e);
          } else {   // columns 32..63 are zero: the wgrad kernel reads the whole 64-wide slab
#pragma unroll
            for (int g = 4; g < 8; ++g) st_shared_v4(sEnc + sw128_chunk_off(row, g), 0u, 0u, 0u, 0u);
          }
        }
        if (q == 0 && h == 0 && lane == 0) NB_TR(qi, layer, x, 3);
        fence_proxy_async();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_remote(act_leader);
        if (q == 0 && h == 0 && lane == 0) NB_TR(qi, layer, x, 4);
        flush(sAct + 2 * h * kSlabBytes, layer * 4 + 2 * h, 2);
        if (layer == 8 && save) {
          pair_sync(pair_id);
          if (h == 0) flush(sEnc, kActSlabEncD, 1);
        }
      }
      mbar_wait(my_acc, par_acc);
      par_acc ^= 1;
      tc_fence_after();
      reuse();
      if (save) pair_sync(pair_id);   // slab 1 (hv of the h = 1 warp) was flushed as layer 8's output by the h = 0 warp
      float r, g, b;
      // the view-branch hidden layer is staged in the (now idle) activation buffer only to be saved
      epilogue_views(taddr, cst, r, g, b, row, 2 * h, sAct, save != nullptr && mask != nullptr, mask ? mask + 8 * 1024 : nullptr);
      tc_fence_before();
      // heads: the h = 1 warp hands its partial sums to the h = 0 warp of the same rows
      if (h == 1) *part = make_float4(r, g, b, sigma);
      pair_sync(pair_id);
      if (h == 0) {
        const float4 o = *part;
        if (valid)
          *reinterpret_cast<float4*>(raw + m * 4) = make_float4(cst[kOffRgbB] + r + o.x, cst[kOffRgbB + 1] + g + o.y,
                                                                cst[kOffRgbB + 2] + b + o.z, cst[kOffAlphaB] + sigma + o.w);
      }
      pair_sync(pair_id);   // the partial slot may be rewritten next iteration
      if (save) {
        fence_proxy_async();
        __syncwarp();
        flush(sAct + h * kSlabBytes, kActSlabHv + h, 1);
      }
    }
    if (acts && lane == 0) bulk_wait_all();
  }
  // no CTA may exit (or free TMEM) while its peer can still signal its barriers or read its SMEM
  tc_fence_before();
  cluster_sync_all();
  if (warp == 2) tmem_dealloc_2sm(tmem, 512);
}

static int build_pack_args(const NbMlpWeights* w, PackArgs& a) {
  const int in_ch = 3 + 6 * w->L_pts, in_views = 3 + 6 * w->L_dirs;
  if (in_ch > 64 || in_views > 32 || w->L_pts < 0 || w->L_dirs < 0) {
    set_error("tensor-core MLP supports multires <= 10 and multires_views <= 4 (got %d, %d); use the fp32 path: nerf.set_default_precision(\"fp32\") or NERF_B200_PRECISION=fp32, inference only",
              w->L_pts, w->L_dirs);
    return NB_E_UNSUPPORTED;
  }
  int s = 0;
  for (int layer = 0; layer < 10; ++layer) {
    const float* mat;
    int ld;
    if (layer < 8) { mat = w->pts_w[layer]; ld = layer == 0 ? in_ch : (layer == 5 ? 256 + in_ch : 256); }
    else if (layer == 8) { mat = w->feature_w; ld = 256; }
    else { mat = w->views_w; ld = 256 + in_views; }
    for (int ks = 0; ks < ks_count(layer); ++ks) {
      int k0, kvalid;
      if (layer == 0) { k0 = 0; kvalid = in_ch; }
      else if (layer == 5) { k0 = ks == 0 ? 0 : in_ch + 64 * (ks - 1); kvalid = ks == 0 ? in_ch : 64; }
      else if (layer == 9 && ks == 4) { k0 = 256; kvalid = in_views; }
      else { k0 = 64 * ks; kvalid = 64; }
      for (int nh = 0; nh < nh_count(layer); ++nh) a.slab[s++] = SlabSrc{mat, ld, nh * 128, k0, kvalid};
    }
  }
  return s == kNumSlabs ? 0 : NB_E_ARG;
}

int init_mlp_tc(int) {
  cudaError_t e = cudaFuncSetAttribute(mlp_fwd_tc_kernel<10, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)kSmemBytes);
  if (e == cudaSuccess)
    e = cudaFuncSetAttribute(umma_selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * kSlabBytes + 2048);
  if (e != cudaSuccess) {
    set_error("init_mlp_tc: %s", cudaGetErrorString(e));
    return (int)e;
  }
  return 0;
}

}  // namespace nb

using namespace nb;

extern "C" {

int64_t nb_mlp_packed_bytes(int L_pts, int L_dirs) {
  (void)L_pts; (void)L_dirs;
  return (int64_t)kNumSlabs * kSlabBytes + kConstFloats * 4;
}

int64_t nb_mlp_act_bytes(int64_t M) {
  return ((M + kTileM - 1) / kTileM) * ((int64_t)kActSlabs * kSlabBytes + kMaskTileBytes);
}

int nb_mlp_pack(const NbMlpWeights* w, void* packed, void* stream) {
  NB_REQUIRE(w && packed, "nb_mlp_pack: null pointer");
  NB_REQUIRE(((uintptr_t)packed & 15) == 0, "nb_mlp_pack: packed buffer must be 16-byte aligned");
  PackArgs a;
  int rc = build_pack_args(w, a);
  if (rc) return rc;
  const int threads = kNumSlabs * 128 * 8;
  pack_weights_kernel<<<(threads + 255) / 256, 256, 0, as_stream(stream)>>>(a, *w, (uint8_t*)packed);
  return check_launch("nb_mlp_pack");
}

int nb_mlp_fwd_tc(const void* packed, int L_pts, int L_dirs, const float* pts, const float* viewdirs, int64_t M,
                  int S, float* raw, void* acts, void* stream) {
  if (M == 0) return 0;
  NB_REQUIRE(packed && pts && viewdirs && raw && S >= 1, "nb_mlp_fwd_tc: bad arguments");
  NB_REQUIRE(M % S == 0, "nb_mlp_fwd_tc: M=%lld is not a multiple of S=%d", (long long)M, S);
  if (L_pts != 10 || L_dirs != 4) {
    set_error("nb_mlp_fwd_tc: only multires=10 / multires_views=4 is instantiated (got %d, %d); other widths run on the fp32 kernel "
              "(nerf.set_default_precision(\"fp32\") or NERF_B200_PRECISION=fp32, inference only)", L_pts, L_dirs);
    return NB_E_UNSUPPORTED;
  }
  const int64_t tiles = (M + kTileM - 1) / kTileM;
  const int64_t quads = (tiles + 3) / 4;  // a CTA pair works on four 128-sample tiles at a time
  NB_REQUIRE(quads < (1ll << 30), "nb_mlp_fwd_tc: M too large for one launch");
  CUtensorMap wmap;
  int rc = make_slab_tmap(&wmap, packed, kNumSlabs);
  if (rc) return rc;
  int clusters = (int)(quads < kNumSMs / 2 ? quads : kNumSMs / 2);
  if (limit_mlp_clusters() > 0 && clusters > limit_mlp_clusters()) clusters = limit_mlp_clusters();
  mlp_fwd_tc_kernel<10, 4><<<2 * clusters, kFwdThreads, kSmemBytes, as_stream(stream)>>>(
      wmap, (const uint8_t*)packed, pts, viewdirs, M, S, raw, (uint8_t*)acts, (int)quads);
  return check_launch("nb_mlp_fwd_tc");
}

#ifdef NB_TRACE
__attribute__((visibility("default"))) int nb_trace_read(long long* out_host) {
  return (int)cudaMemcpyFromSymbol(out_host, g_trace, sizeof(long long) * 4 * 10 * 2 * 8);
}
#endif

// test hook: D[128,128] = A[128,K] B[128,K]^T with bf16-rounded operands (K in {64,128,192,256})
int nb_debug_umma(const float* A, const float* B, int K, float* D, void* stream) {
  NB_REQUIRE(A && B && D && K >= 64 && K <= 256 && K % 64 == 0, "nb_debug_umma: bad arguments");
  umma_selftest_kernel<<<1, 128, 2 * (K / 64) * kSlabBytes + 2048, as_stream(stream)>>>(A, B, K, D);
  return check_launch("nb_debug_umma");
}

}  // extern "C"
