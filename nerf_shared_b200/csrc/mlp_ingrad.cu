// Gradient of the NeRF MLP wrt its INPUTS (sample positions and view directions): the last link of
// the pose-estimation backward (autograd of nerf.py:97-118,124 as driven by
// examples/relative_pose_estimation_demo/demo_est_rel_pose.py:87-97, weights frozen).
//
// The dgrad chain (mlp_bwd.cu) leaves the pre-activation gradients of every layer in HBM in the
// kernels' slab image.  Three of them touch the encoded inputs:
//     d enc(pts)  [M,63] = dA_0 . W_0  +  dA_5 . W_5[:, :63]        (first layer + skip connection)
//     d enc(dirs) [M,27] = dA_hv . W_views[:, 256:283]             (view branch)
// Both are tcgen05 GEMMs with the slabs as K-major A operands (M = 128 samples per tile, K = 512 /
// 128 features) and small transposed weight slabs (N = 64 / 32) resident in shared memory.  The
// epilogue reads the fp32 accumulator row of its sample from TMEM and folds the encoder's
// derivative in registers, with the very sin/cos values the forward used (encode.cuh):
//     d x_c = d e[c] + sum_l 2^l ( cos(2^l x_c) d e[sin_l,c] - sin(2^l x_c) d e[cos_l,c] )
// d_pts [M,3] is written; d_viewdirs [M/S,3] is the sum over the S samples of each ray (the
// reference broadcasts viewdirs over the samples, nerf.py:101) and is accumulated with atomics.
//
// Warp roles (192 threads): warp 0 = dA slab producer (bulk copies, 6-stage ring), warp 1 = TMEM
// allocator + MMA issuer, warps 2-5 = epilogue.  Two 128-column accumulator buffers ping-pong.
#include "common.cuh"
#include "encode.cuh"
#include "mlp_common.cuh"
#include "ptx.cuh"

namespace nb {

using namespace ptx;

constexpr int kInSlabs = 10;             // K-slabs per tile: dA_0 (4), dA_5 (4), dA_hv (2)
constexpr int kInWBytes = 8192;          // one [64 x 64] bf16 weight slab
constexpr int kInStages = 6;
constexpr uint32_t kISmW = 0;                                    // 10 x 8 KB resident weights
constexpr uint32_t kISmRing = kInSlabs * kInWBytes;              // 81920
constexpr uint32_t kISmBar = kISmRing + kInStages * kSlabBytes;  // 180224
constexpr uint32_t kISmTmemPtr = kISmBar + 120;
constexpr uint32_t kISmemBytes = kISmBar + 128 + 1024;

__host__ __device__ constexpr int in_da_slab(int j) { return j < 4 ? j : (j < 8 ? 20 + (j - 4) : kDaSlabHv + (j - 8)); }

// weight slab j, element (n, k): j<4: W0[64j+k][n]; j<8: W5[64(j-4)+k][n]; else Wv[64(j-8)+k][256+n]
__global__ void __launch_bounds__(256) pack_ingrad_kernel(NbMlpWeights w, uint8_t* __restrict__ blob) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= kInSlabs * 64 * 8) return;
  const int j = idx / 512, rem = idx - j * 512, n = rem >> 3, g = rem & 7;
  const int in_ch = 3 + 6 * w.L_pts, in_views = 3 + 6 * w.L_dirs;
  const float* mat;
  int ld, col, nvalid, out0;
  if (j < 4) { mat = w.pts_w[0]; ld = in_ch; col = n; nvalid = in_ch; out0 = 64 * j; }
  else if (j < 8) { mat = w.pts_w[5]; ld = 256 + in_ch; col = n; nvalid = in_ch; out0 = 64 * (j - 4); }
  else { mat = w.views_w; ld = 256 + in_views; col = 256 + n; nvalid = in_views; out0 = 64 * (j - 8); }
  uint32_t p[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const int k = g * 8 + q * 2;
    const float lo = n < nvalid ? mat[(size_t)(out0 + k) * ld + col] : 0.f;
    const float hi = n < nvalid ? mat[(size_t)(out0 + k + 1) * ld + col] : 0.f;
    p[q] = pack_bf16x2(lo, hi);
  }
  *reinterpret_cast<uint4*>(blob + (size_t)j * kInWBytes + sw128_chunk_off(n, g)) = make_uint4(p[0], p[1], p[2], p[3]);
}

// d v_c from the gradient of the encoding row (fp32), given the encoding itself
template <int L, int N>
__device__ __forceinline__ void encode3_bwd(const float (&e)[N], const float (&de)[N], float (&dv)[3]) {
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    float acc = de[c];
#pragma unroll
    for (int l = 0; l < L; ++l) {
      const float s = e[3 + 6 * l + c], co = e[3 + 6 * l + 3 + c];
      acc = fmaf((float)(1 << l), co * de[3 + 6 * l + c] - s * de[3 + 6 * l + 3 + c], acc);
    }
    dv[c] = acc;
  }
}

template <int L_PTS, int L_DIRS>
__global__ void __launch_bounds__(192, 1)
    mlp_ingrad_kernel(const uint8_t* __restrict__ wblob, const uint8_t* __restrict__ dA, const float* __restrict__ pts,
                      const float* __restrict__ viewdirs, int64_t M, int S, int num_tiles, float* __restrict__ d_pts,
                      float* __restrict__ d_viewdirs) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gbase = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t bar_full = base + kISmBar, bar_empty = bar_full + 8 * kInStages, bar_acc = bar_empty + 8 * kInStages,
                 bar_free = bar_acc + 16, bar_w = bar_free + 16;
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int i = 0; i < kInStages; ++i) {
      mbar_init(bar_full + 8 * i, 1);
      mbar_init(bar_empty + 8 * i, 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(bar_acc + 8 * b, 1);    // tcgen05.commit: accumulator b complete
      mbar_init(bar_free + 8 * b, 128); // the four epilogue warps have drained accumulator b
    }
    mbar_init(bar_w, 1);
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(base + kISmTmemPtr, 256);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(gbase + kISmTmemPtr);

  // producer and MMA warps run converged and issue under elect.sync (see mlp_tc.cu)
  if (warp == 0) {
    if (elect_one()) {
      mbar_expect_tx(bar_w, kInSlabs * kInWBytes);
      bulk_g2s(base + kISmW, wblob, kInSlabs * kInWBytes, bar_w);
    }
    __syncwarp();
    uint32_t stage = 0, phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const uint8_t* dt = dA + (size_t)tile * kDaSlabs * kSlabBytes;
      for (int j = 0; j < kInSlabs; ++j) {
        mbar_wait(bar_empty + 8 * stage, phase ^ 1);
        if (elect_one()) {
          mbar_expect_tx(bar_full + 8 * stage, kSlabBytes);
          bulk_g2s(base + kISmRing + stage * kSlabBytes, dt + (size_t)in_da_slab(j) * kSlabBytes, kSlabBytes,
                   bar_full + 8 * stage);
        }
        __syncwarp();
        if (++stage == kInStages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    const uint32_t idesc64 = make_idesc_bf16(128, 64), idesc32 = make_idesc_bf16(128, 32);
    const uint64_t desc0 = make_desc_kmajor_sw128(0);
    uint32_t stage = 0, phase = 0, par_free0 = 0, par_free1 = 0;
    int it = 0;
    mbar_wait(bar_w, 0);
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const int b = it & 1;
      if (it >= 2) {  // accumulator b was used two tiles ago: wait until its epilogue has read it
        if (b == 0) { mbar_wait(bar_free, par_free0); par_free0 ^= 1; }
        else { mbar_wait(bar_free + 8, par_free1); par_free1 ^= 1; }
        tc_fence_after();
      }
      for (int j = 0; j < kInSlabs; ++j) {
        mbar_wait(bar_full + 8 * stage, phase);
        tc_fence_after();
        const uint64_t ad = desc0 | (uint64_t)(((base + kISmRing + stage * kSlabBytes) & 0x3FFFFu) >> 4);
        const uint64_t bd = desc0 | (uint64_t)(((base + kISmW + j * kInWBytes) & 0x3FFFFu) >> 4);
        const uint32_t d = tmem + b * 128 + (j < 8 ? 0 : 64);
        const uint32_t idesc = j < 8 ? idesc64 : idesc32;
        if (elect_one()) {
          umma_bf16(d, ad, bd, idesc, (j & 7) != 0);
          umma_bf16(d, ad + 2, bd + 2, idesc, 1);
          umma_bf16(d, ad + 4, bd + 4, idesc, 1);
          umma_bf16(d, ad + 6, bd + 6, idesc, 1);
          umma_commit(bar_empty + 8 * stage);
          if (j == kInSlabs - 1) umma_commit(bar_acc + 8 * b);
        }
        __syncwarp();
        if (++stage == kInStages) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    const int q = warp & 3;  // TMEM lane quarter this warp may read
    const int row = (q << 5) | lane;
    uint32_t par_acc[2] = {0, 0};
    int it = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const int b = it & 1;
      const int64_t m = (int64_t)tile * kTileM + row;
      const bool valid = m < M;
      float p0 = 0.f, p1 = 0.f, p2 = 0.f, v0 = 0.f, v1 = 0.f, v2 = 0.f;
      int64_t ray = -1;
      if (valid) {
        p0 = pts[m * 3]; p1 = pts[m * 3 + 1]; p2 = pts[m * 3 + 2];
        ray = m / S;
        v0 = viewdirs[ray * 3]; v1 = viewdirs[ray * 3 + 1]; v2 = viewdirs[ray * 3 + 2];
      }
      mbar_wait(bar_acc + 8 * b, par_acc[b]);
      par_acc[b] ^= 1;
      tc_fence_after();
      const uint32_t taddr = tmem + b * 128 + ((uint32_t)(q * 32) << 16);
      uint32_t r0[32], r1[32], r2[32];
      tmem_ld32(taddr, r0);
      tmem_ld32(taddr + 32, r1);
      tmem_ld32(taddr + 64, r2);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(bar_free + 8 * b);
      float dp[3], dv[3];
      {
        float e[64], de[64];
        encode3<L_PTS>(p0, p1, p2, e);
#pragma unroll
        for (int i = 0; i < 32; ++i) { de[i] = __uint_as_float(r0[i]); de[32 + i] = __uint_as_float(r1[i]); }
        encode3_bwd<L_PTS>(e, de, dp);
      }
      {
        float e[32], de[32];
        encode3<L_DIRS>(v0, v1, v2, e);
#pragma unroll
        for (int i = 0; i < 32; ++i) de[i] = __uint_as_float(r2[i]);
        encode3_bwd<L_DIRS>(e, de, dv);
      }
      if (valid) {
        d_pts[m * 3] = dp[0]; d_pts[m * 3 + 1] = dp[1]; d_pts[m * 3 + 2] = dp[2];
      } else {
        dv[0] = dv[1] = dv[2] = 0.f;
      }
      // d_viewdirs: one atomic per warp when its 32 samples belong to one ray, else one per sample
      const int64_t ray0 = __shfl_sync(0xffffffffu, ray, 0);
      const bool uniform = __all_sync(0xffffffffu, ray == ray0 || !valid);
      if (ray0 >= 0 && uniform) {
        const float s0 = warp_sum(dv[0]), s1 = warp_sum(dv[1]), s2 = warp_sum(dv[2]);
        if (lane == 0) {
          atomicAdd(d_viewdirs + ray0 * 3, s0);
          atomicAdd(d_viewdirs + ray0 * 3 + 1, s1);
          atomicAdd(d_viewdirs + ray0 * 3 + 2, s2);
        }
      } else if (valid) {
        atomicAdd(d_viewdirs + ray * 3, dv[0]);
        atomicAdd(d_viewdirs + ray * 3 + 1, dv[1]);
        atomicAdd(d_viewdirs + ray * 3 + 2, dv[2]);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 256);
}

int init_mlp_ingrad(int) {
  cudaError_t e = cudaFuncSetAttribute(mlp_ingrad_kernel<10, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)kISmemBytes);
  if (e != cudaSuccess) {
    set_error("init_mlp_ingrad: %s", cudaGetErrorString(e));
    return (int)e;
  }
  return 0;
}

}  // namespace nb

using namespace nb;

extern "C" {

int64_t nb_mlp_ingrad_packed_bytes(int L_pts, int L_dirs) {
  (void)L_pts; (void)L_dirs;
  return (int64_t)kInSlabs * kInWBytes;
}

int nb_mlp_pack_ingrad(const NbMlpWeights* w, void* packed_in, void* stream) {
  NB_REQUIRE(w && packed_in, "nb_mlp_pack_ingrad: null pointer");
  NB_REQUIRE(((uintptr_t)packed_in & 15) == 0, "nb_mlp_pack_ingrad: buffer must be 16-byte aligned");
  if (3 + 6 * w->L_pts > 64 || 3 + 6 * w->L_dirs > 32 || w->L_pts < 0 || w->L_dirs < 0) {
    set_error("tensor-core MLP supports multires <= 10 and multires_views <= 4 (got %d, %d)", w->L_pts, w->L_dirs);
    return NB_E_UNSUPPORTED;
  }
  const int threads = kInSlabs * 64 * 8;
  pack_ingrad_kernel<<<(threads + 255) / 256, 256, 0, as_stream(stream)>>>(*w, (uint8_t*)packed_in);
  return check_launch("nb_mlp_pack_ingrad");
}

int nb_mlp_bwd_inputs(const void* packed_in, int L_pts, int L_dirs, const void* dA, const float* pts,
                      const float* viewdirs, int64_t M, int S, float* d_pts, float* d_viewdirs, void* stream) {
  if (M == 0) return 0;
  NB_REQUIRE(packed_in && dA && pts && viewdirs && d_pts && d_viewdirs && S >= 1, "nb_mlp_bwd_inputs: bad arguments");
  NB_REQUIRE(M % S == 0, "nb_mlp_bwd_inputs: M=%lld is not a multiple of S=%d", (long long)M, S);
  if (L_pts != 10 || L_dirs != 4) {
    set_error("nb_mlp_bwd_inputs: only multires=10 / multires_views=4 is instantiated (got %d, %d)", L_pts, L_dirs);
    return NB_E_UNSUPPORTED;
  }
  const int64_t tiles = (M + kTileM - 1) / kTileM;
  NB_REQUIRE(tiles < (1ll << 30), "nb_mlp_bwd_inputs: M too large for one launch");
  const int grid = (int)(tiles < kNumSMs ? tiles : kNumSMs);
  mlp_ingrad_kernel<10, 4><<<grid, 192, kISmemBytes, as_stream(stream)>>>(
      (const uint8_t*)packed_in, (const uint8_t*)dA, pts, viewdirs, M, S, (int)tiles, d_pts, d_viewdirs);
  return check_launch("nb_mlp_bwd_inputs");
}

}  // extern "C"
