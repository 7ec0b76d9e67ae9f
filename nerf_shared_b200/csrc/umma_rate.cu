// Micro-benchmark (debug hook, not on the product path): how fast does the tensor pipe retire a pure
// stream of tcgen05.mma instructions with the operand layouts the MLP kernels use (K-major SWIZZLE_128B
// slabs in shared memory, fp32 accumulators in TMEM)?  Every CTA (pair) issues `iters` MMAs back to back
// with no barrier but the final commit and reports its elapsed SM cycles.
//   mode 0: cta_group::2, M=256, N=256      mode 1: cta_group::1, M=128, N=256
//   mode 2: cta_group::2, M=256, N=128      mode 3: cta_group::1, M=128, N=128
//   mode 4: cta_group::2, M=256, N=256, a commit every 4 MMAs (as the fused kernels do)
//   mode 5: as 4, and before a group of 4 MMAs wait for the completion of the group issued 3 groups earlier
//           (a 3-stage ring whose stages are released by tcgen05.commit, no producer in between)
//   mode 6: as 5 with ring depth 1 (commit, wait, issue: exposes the commit -> mbarrier -> wake-up latency)
//   mode 7: the 3-stage ring of the fused kernels: commit -> empty[s] -> a producer thread in another warp
//           (both CTAs) -> arrive full[s] (leader) -> the MMA thread waits full[s]
#include "common.cuh"
#include "mlp_common.cuh"
#include "ptx.cuh"

namespace nb {

using namespace ptx;

constexpr uint32_t kRSmA = 0, kRSmB = 65536, kRSmBar = 65536 + 65536, kRSmPtr = kRSmBar + 64;
constexpr uint32_t kRSmemBytes = kRSmPtr + 64 + 1024;

template <int CG>
__global__ void __launch_bounds__(128, 1) umma_rate_kernel(int mode, int iters, long long* __restrict__ cycles) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gbase = smem_raw + (base - smem_u32(smem_raw));
  const int warp = threadIdx.x >> 5;
  // operands: a bf16 bit pattern of small finite numbers (realistic toggling, no inf/nan)
  for (uint32_t i = threadIdx.x; i < 131072 / 4; i += 128) {
    const uint32_t h = (i * 2654435761u) >> 7;
    reinterpret_cast<uint32_t*>(gbase)[i] = 0x3C003C00u ^ (h & 0x807F807Fu);
  }
  fence_proxy_async();
  const uint32_t bar = base + kRSmBar, bar2 = bar + 8, bar_e = bar + 16, bar_f = bar + 40;
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    mbar_init(bar2, 1);
    for (int i = 0; i < 3; ++i) { mbar_init(bar_e + 8 * i, 1); mbar_init(bar_f + 8 * i, 1); }
    fence_mbar_init();
  }
  if (warp == 0) {
    if (CG == 2) { tmem_alloc_2sm(base + kRSmPtr, 512); tmem_relinquish_2sm(); }
    else { tmem_alloc(base + kRSmPtr, 512); tmem_relinquish(); }
  }
  tc_fence_before();
  if (CG == 2) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(gbase + kRSmPtr);
  const uint32_t rank = CG == 2 ? cluster_ctarank() : 0;
  const int N = (mode == 2 || mode == 3) ? 128 : 256;
  if (mode < 8 && threadIdx.x == 0 && rank == 0) {
    const uint32_t idesc = make_idesc_bf16(CG == 2 ? 256 : 128, N);
    const long long t0 = clock64();
    for (int i = 0; i < (mode >= 5 ? 0 : iters); ++i) {
      const int slab = (i >> 2) & 3, k16 = i & 3;
      const uint64_t ad = make_desc_kmajor_sw128(base + kRSmA + slab * kSlabBytes) + 2 * k16;
      const uint64_t bd = make_desc_kmajor_sw128(base + kRSmB + ((i >> 2) % 3) * kSlabBytes) + 2 * k16;
      const uint32_t d = tmem + ((i >> 4) & 1) * 256;
      if (CG == 2) umma_bf16_2sm(d, ad, bd, idesc, (i & 15) != 0);
      else umma_bf16(d, ad, bd, idesc, (i & 15) != 0);
      if (mode == 4 && (i & 3) == 3) umma_commit_2sm(bar2);
    }
    if (CG == 2 && mode >= 5) {
      const int depth = mode == 6 ? 1 : 3;
      const int groups = iters / 4;
      for (int g = 0; g < groups; ++g) {
        const int st = g % depth;
        if (mode == 7) {
          mbar_wait_cluster(bar_f + 8 * st, (g / depth) & 1);
        } else if (g >= depth) {
          mbar_wait_cluster(bar_e + 8 * st, ((g / depth) - 1) & 1);
        }
        tc_fence_after();
        for (int k16 = 0; k16 < 4; ++k16) {
          const uint64_t ad = make_desc_kmajor_sw128(base + kRSmA + (g & 3) * kSlabBytes) + 2 * k16;
          const uint64_t bd = make_desc_kmajor_sw128(base + kRSmB + st * kSlabBytes) + 2 * k16;
          umma_bf16_2sm(tmem + ((g >> 2) & 1) * 256, ad, bd, idesc, ((g & 3) | k16) != 0);
        }
        umma_commit_2sm(bar_e + 8 * st);
      }
    }
    if (CG == 2) umma_commit_2sm(bar); else umma_commit(bar);
    const long long t1 = clock64();
    mbar_wait(bar, 0);
    const long long t2 = clock64();
    cycles[2 * (blockIdx.x / CG)] = t2 - t0;
    cycles[2 * (blockIdx.x / CG) + 1] = t1 - t0;
  } else if (mode < 8 && threadIdx.x == 0 && CG == 2) {
    mbar_wait(bar, 0);  // the multicast commit also arrives here: keep the peer alive until the MMAs are done
  } else if (CG == 2 && mode == 7 && threadIdx.x == 32) {
    // producer of the fused kernels without the copy: stage free in this CTA -> (leader) signal it full
    const int groups = iters / 4;
    for (int g = 0; g < groups; ++g) {
      const int st = g % 3;
      mbar_wait_cluster(bar_e + 8 * st, ((g / 3) & 1) ^ 1);
      if (rank == 0) mbar_arrive(bar_f + 8 * st);
    }
  }
  if (mode == 8 || mode == 9) {
    // tcgen05.ld throughput: mode 8 = all four warps, mode 9 = warp 0 alone; each reads its 32 lanes x 256
    // columns (32 KB per warp and iteration) `iters` times
    __syncthreads();
    if (mode == 8 || warp == 0) {
      const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16);
      uint32_t sink = 0;
      const long long t0 = clock64();
      for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          uint32_t v[32];
          tmem_ld32(taddr + c * 32, v);
          tmem_ld_wait();
#pragma unroll
          for (int q = 0; q < 32; ++q) sink ^= v[q];
        }
      }
      const long long t1 = clock64();
      if ((threadIdx.x & 31) == 0 && (CG == 1 || rank == 0)) {
        cycles[2 * (blockIdx.x / CG)] = t1 - t0;
        cycles[2 * (blockIdx.x / CG) + 1] = sink;
      }
    }
  }
  tc_fence_before();
  if (CG == 2) cluster_sync_all(); else __syncthreads();
  if (warp == 0) {
    if (CG == 2) tmem_dealloc_2sm(tmem, 512); else tmem_dealloc(tmem, 512);
  }
}

}  // namespace nb

using namespace nb;

extern "C" {

// cycles: 2 int64 per CTA (pair): [total incl. drain, issue only]
// opt-in symbol (not in include/nerf_b200.h): exported explicitly
__attribute__((visibility("default"))) int nb_debug_umma_rate(int mode, int iters, int ctas, long long* cycles, void* stream) {
  NB_REQUIRE(cycles && iters > 0 && mode >= 0 && mode <= 9 && ctas >= 1 && ctas <= kNumSMs, "nb_debug_umma_rate: bad arguments");
  cudaError_t e;
  if (mode == 0 || mode == 2 || mode >= 4) {
    e = cudaFuncSetAttribute(umma_rate_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kRSmemBytes);
    if (e != cudaSuccess) return (int)e;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((ctas / 2) * 2);
    cfg.blockDim = dim3(128);
    cfg.dynamicSmemBytes = kRSmemBytes;
    cfg.stream = as_stream(stream);
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    e = cudaLaunchKernelEx(&cfg, umma_rate_kernel<2>, mode, iters, cycles);
    if (e != cudaSuccess) { set_error("nb_debug_umma_rate: %s", cudaGetErrorString(e)); return (int)e; }
  } else {
    e = cudaFuncSetAttribute(umma_rate_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kRSmemBytes);
    if (e != cudaSuccess) return (int)e;
    umma_rate_kernel<1><<<ctas, 128, kRSmemBytes, as_stream(stream)>>>(mode, iters, cycles);
  }
  return check_launch("nb_debug_umma_rate");
}

}  // extern "C"
