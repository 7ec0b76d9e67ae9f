// Fused Adam over flat fp32 buffers (torch.optim.Adam semantics, utils.py:170 / main.py:104).
// Streaming kernel: reads p, g, m, v (16 B/param) and writes p, m, v (12 B/param).
#include "common.cuh"

namespace nb {

__global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                   float* __restrict__ m, float* __restrict__ v, int64_t n,
                                                   float step_size, float beta1, float beta2, float eps,
                                                   float inv_sqrt_bc2, float grad_scale) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float gi = g[i] * grad_scale;
    float mi = beta1 * m[i] + (1.0f - beta1) * gi;
    float vi = beta2 * v[i] + (1.0f - beta2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    float denom = sqrtf(vi) * inv_sqrt_bc2 + eps;
    p[i] = p[i] - step_size * (mi / denom);
  }
}

// Same update with the scalars read from device memory (hyper[0..5] = step_size, beta1, beta2, eps,
// 1/sqrt(bias_correction2), grad_scale): the launch is then identical every step and can sit in a
// CUDA graph while the learning-rate schedule (main.py:108-112) and bias corrections advance.
__global__ void __launch_bounds__(256) adam_dev_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                       float* __restrict__ m, float* __restrict__ v, int64_t n,
                                                       const float* __restrict__ hyper) {
  const float step_size = hyper[0], beta1 = hyper[1], beta2 = hyper[2], eps = hyper[3], inv_sqrt_bc2 = hyper[4],
              grad_scale = hyper[5];
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float gi = g[i] * grad_scale;
    float mi = beta1 * m[i] + (1.0f - beta1) * gi;
    float vi = beta2 * v[i] + (1.0f - beta2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    float denom = sqrtf(vi) * inv_sqrt_bc2 + eps;
    p[i] = p[i] - step_size * (mi / denom);
  }
}

__global__ void adam_hyper_kernel(float* __restrict__ hyper, float step_size, float beta1, float beta2, float eps,
                                  float inv_sqrt_bc2, float grad_scale, int step) {
  hyper[0] = step_size; hyper[1] = beta1; hyper[2] = beta2; hyper[3] = eps; hyper[4] = inv_sqrt_bc2; hyper[5] = grad_scale;
  hyper[6] = __int_as_float(step);   // epoch of the cross-GPU flags of adam_allreduce_p2p_kernel (never repeats)
}

}  // namespace nb

using namespace nb;

extern "C" int nb_adam_step(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2,
                            float eps, int step, float grad_scale, void* stream) {
  if (n == 0) return 0;
  NB_REQUIRE(p && g && m && v && step >= 1, "nb_adam_step: bad arguments");
  double bc1 = 1.0 - pow((double)beta1, step), bc2 = 1.0 - pow((double)beta2, step);
  adam_kernel<<<grid_for(n, 256), 256, 0, as_stream(stream)>>>(p, g, m, v, n, (float)(lr / bc1), beta1, beta2, eps,
                                                              (float)(1.0 / sqrt(bc2)), grad_scale);
  return check_launch("nb_adam_step");
}

extern "C" int nb_adam_hyper(float* hyper, float lr, float beta1, float beta2, float eps, int step, float grad_scale,
                             int epoch, void* stream) {
  NB_REQUIRE(hyper && step >= 1, "nb_adam_hyper: bad arguments");
  double bc1 = 1.0 - pow((double)beta1, step), bc2 = 1.0 - pow((double)beta2, step);
  adam_hyper_kernel<<<1, 1, 0, as_stream(stream)>>>(hyper, (float)(lr / bc1), beta1, beta2, eps, (float)(1.0 / sqrt(bc2)),
                                                    grad_scale, epoch);
  return check_launch("nb_adam_hyper");
}

// One launch in front of a graph replay: this step's Adam scalars (as nb_adam_hyper) and the copy of the step's two
// inputs (ray batch, targets) into the static buffers the captured graph reads -- three stream operations less
// between two replays of a 1.3 ms step.
namespace nb {
__global__ void __launch_bounds__(256) step_prologue_kernel(float* __restrict__ hyper, float lr_c, float b1, float b2,
                                                            float eps, float inv_sqrt_bc2, float grad_scale, int step,
                                                            const float* __restrict__ src1, float* __restrict__ dst1,
                                                            int64_t n1, const float* __restrict__ src2,
                                                            float* __restrict__ dst2, int64_t n2) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i == 0) {
    hyper[0] = lr_c; hyper[1] = b1; hyper[2] = b2; hyper[3] = eps; hyper[4] = inv_sqrt_bc2; hyper[5] = grad_scale;
    hyper[6] = __int_as_float(step);
  }
  if (i < n1) dst1[i] = src1[i];
  else if (i < n1 + n2) dst2[i - n1] = src2[i - n1];
}
}  // namespace nb

extern "C" int nb_train_step_prologue(float* hyper, float lr, float beta1, float beta2, float eps, int step,
                                      float grad_scale, int epoch, const float* src1, float* dst1, int64_t n1, const float* src2,
                                      float* dst2, int64_t n2, void* stream) {
  NB_REQUIRE(hyper && step >= 1 && n1 >= 0 && n2 >= 0 && (n1 == 0 || (src1 && dst1)) && (n2 == 0 || (src2 && dst2)),
             "nb_train_step_prologue: bad arguments");
  double bc1 = 1.0 - pow((double)beta1, step), bc2 = 1.0 - pow((double)beta2, step);
  const int64_t n = n1 + n2 > 1 ? n1 + n2 : 1;
  nb::step_prologue_kernel<<<(unsigned)((n + 255) / 256), 256, 0, nb::as_stream(stream)>>>(
      hyper, (float)(lr / bc1), beta1, beta2, eps, (float)(1.0 / sqrt(bc2)), grad_scale, epoch, src1, dst1, n1, src2, dst2, n2);
  return nb::check_launch("nb_train_step_prologue");
}

extern "C" int nb_adam_step_dev(float* p, const float* g, float* m, float* v, int64_t n, const float* hyper,
                                void* stream) {
  if (n == 0) return 0;
  NB_REQUIRE(p && g && m && v && hyper, "nb_adam_step_dev: bad arguments");
  adam_dev_kernel<<<grid_for(n, 256), 256, 0, as_stream(stream)>>>(p, g, m, v, n, hyper);
  return check_launch("nb_adam_step_dev");
}

// ------------------------------------------------------------------------------------------
// Gradient all-reduce FUSED with the optimizer step, over NVLink peer memory ("one
// all-reduce of the flat gradient buffer ... fused into the epilogue").  Replaces, for world > 1,
//     ncclAllReduce(flat_g) ; adam(flat_p, flat_g)        (two launches, the first a 50-60 us NCCL ring at 4.8 MB)
// by ONE kernel per rank working on the symmetric gradient buffers of all ranks:
//   barrier A   every rank's weight-gradient kernels have finished (stream order + a flag per peer);
//   phase 1     two-shot all-reduce: rank r sums slice r of ALL ranks' gradients with coalesced 16-byte peer loads
//               and writes the reduced slice back into slice r of EVERY rank's buffer (peer stores).  Slice r of any
//               buffer is read and written by rank r only, so the exchange is race-free in place;
//   barrier B   all slices have landed everywhere (grid-wide locally, flag per peer across ranks);
//   phase 2     Adam on the full buffer from the reduced gradient: replicated, bit-identical on every rank.
// Per rank 2 (W-1)/W x 4.8 MB cross NVLink instead of a ring's latency chain.  Flags carry the step number as epoch
// (monotonic, hyper[6], identical on all ranks), so nothing has to be reset between steps.  Spins are bounded: a
// rank that never arrives traps the kernel instead of hanging the GPU.
// ------------------------------------------------------------------------------------------
namespace nb {

constexpr int kMaxPeers = 16;
struct PeerPtrs {
  float* grad[kMaxPeers];   // every rank's symmetric gradient buffer
  int* flags[kMaxPeers];    // every rank's flag block: [0..W) = barrier A, [16..16+W) = barrier B
  float* mc;                // multicast (NVLS) mapping of the same buffers, or NULL: the NVSwitch reduces / broadcasts
};

// NVLink SHARP: one load returns the sum over all ranks' copies (reduced inside the switch), one store writes all copies
__device__ __forceinline__ float4 multimem_ld_reduce_add(const float* p) {
  float4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ void multimem_st(float* p, const float4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}

__device__ __forceinline__ void st_release_sys(int* p, int v) {
  asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int ld_acquire_sys(const int* p) {
  int v;
  asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// all W flags of `mine` (written by the peers) have reached `epoch`
__device__ __forceinline__ void wait_flags(const int* mine, int world, int epoch) {
  if (threadIdx.x < world) {
    long long spins = 0;
    while (ld_acquire_sys(mine + threadIdx.x) < epoch) {
      if (++spins > (1ll << 25)) __trap();   // ~ 10 s: a peer is gone
      __nanosleep(64);
    }
  }
  __syncthreads();
}

__global__ void __launch_bounds__(512) adam_allreduce_p2p_kernel(PeerPtrs peers, int rank, int world, float* __restrict__ p,
                                                                  float* __restrict__ m, float* __restrict__ v, int64_t n,
                                                                  const float* __restrict__ hyper,
                                                                  unsigned int* __restrict__ done_counter) {
  const int epoch = __float_as_int(hyper[6]);
  int* my_flags = peers.flags[rank];
  // ---- barrier A: tell every peer that this rank's gradients are complete (they are: this kernel is stream-ordered
  //      behind the weight-gradient kernels), wait until every peer said so
  if (blockIdx.x == 0 && threadIdx.x < world) {
    __threadfence_system();
    st_release_sys(peers.flags[threadIdx.x] + rank, epoch);
  }
  wait_flags(my_flags, world, epoch);

  // ---- phase 1: reduce slice `rank` over all ranks, write it back to everyone (float4 granularity)
  const int64_t n4 = n >> 2;                                  // n is a multiple of 4
  const int64_t per = (n4 + world - 1) / world;
  const int64_t lo = per * rank, hi = (lo + per < n4) ? lo + per : n4;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nth = (int64_t)gridDim.x * blockDim.x;
  if (peers.mc != nullptr) {
    // in-switch reduction: 1/W of the buffer per rank crosses this GPU's links once in each direction
    for (int64_t i = lo + tid; i < hi; i += nth) multimem_st(peers.mc + 4 * i, multimem_ld_reduce_add(peers.mc + 4 * i));
  } else
  for (int64_t i = lo + tid; i < hi; i += nth) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int q0 = 0; q0 < world; q0 += 8) {                   // all peer loads of a group in flight before the first add
      float4 g[8];
#pragma unroll
      for (int j = 0; j < 8; ++j)
        g[j] = (q0 + j < world) ? __ldcv(reinterpret_cast<const float4*>(peers.grad[q0 + j]) + i) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int j = 0; j < 8; ++j) {                           // fixed rank order: the same sum bit for bit on every rank
        acc.x += g[j].x; acc.y += g[j].y; acc.z += g[j].z; acc.w += g[j].w;
      }
    }
    for (int q = 0; q < world; ++q) reinterpret_cast<float4*>(peers.grad[q])[i] = acc;
  }
  // ---- barrier B: this rank's slice is visible everywhere once all its blocks are past the stores
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int arrived = atomicAdd(done_counter, 1u);
    if (arrived == gridDim.x - 1) {                           // last block of this rank
      *done_counter = 0u;                                     // (next step)
      __threadfence_system();
      for (int q = 0; q < world; ++q) st_release_sys(peers.flags[q] + 16 + rank, epoch);
    }
  }
  wait_flags(my_flags + 16, world, epoch);

  // ---- phase 2: Adam from the reduced gradient (now in this rank's own buffer)
  const float step_size = hyper[0], beta1 = hyper[1], beta2 = hyper[2], eps = hyper[3], inv_sqrt_bc2 = hyper[4],
              grad_scale = hyper[5];
  const float4* g4 = reinterpret_cast<const float4*>(peers.grad[rank]);
  float4* p4 = reinterpret_cast<float4*>(p);
  float4* m4 = reinterpret_cast<float4*>(m);
  float4* v4 = reinterpret_cast<float4*>(v);
  for (int64_t i = tid; i < n4; i += nth) {
    const float4 gq = __ldcv(g4 + i);                          // written by peers: bypass stale L1 lines
    float4 mq = m4[i], vq = v4[i], pq = p4[i];
    const float gi[4] = {gq.x * grad_scale, gq.y * grad_scale, gq.z * grad_scale, gq.w * grad_scale};
    float* mm = &mq.x; float* vv = &vq.x; float* pp = &pq.x;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float mi = beta1 * mm[k] + (1.0f - beta1) * gi[k];
      const float vi = beta2 * vv[k] + (1.0f - beta2) * gi[k] * gi[k];
      mm[k] = mi;
      vv[k] = vi;
      pp[k] = pp[k] - step_size * (mi / (sqrtf(vi) * inv_sqrt_bc2 + eps));
    }
    m4[i] = mq; v4[i] = vq; p4[i] = pq;
  }
}

}  // namespace nb

extern "C" int nb_adam_allreduce_p2p(float* p, float* m, float* v, int64_t n, const float* hyper, void* const* peer_grads,
                                     void* const* peer_flags, void* multicast_grads, int rank, int world,
                                     unsigned int* done_counter, void* stream) {
  NB_REQUIRE(p && m && v && hyper && peer_grads && peer_flags && done_counter, "nb_adam_allreduce_p2p: null pointer");
  NB_REQUIRE(world >= 1 && world <= nb::kMaxPeers && rank >= 0 && rank < world, "nb_adam_allreduce_p2p: bad rank / world (max %d ranks)",
             nb::kMaxPeers);
  NB_REQUIRE(n > 0 && (n & 3) == 0, "nb_adam_allreduce_p2p: n must be a positive multiple of 4 (pad the flat buffers)");
  NB_REQUIRE((((uintptr_t)p | (uintptr_t)m | (uintptr_t)v) & 15) == 0, "nb_adam_allreduce_p2p: p / m / v must be 16-byte aligned");
  nb::PeerPtrs pp;
  pp.mc = (float*)multicast_grads;
  NB_REQUIRE(((uintptr_t)pp.mc & 15) == 0, "nb_adam_allreduce_p2p: unaligned multicast pointer");
  for (int q = 0; q < nb::kMaxPeers; ++q) {
    pp.grad[q] = q < world ? (float*)peer_grads[q] : nullptr;
    pp.flags[q] = q < world ? (int*)peer_flags[q] : nullptr;
    NB_REQUIRE(q >= world || (pp.grad[q] && pp.flags[q] && ((uintptr_t)pp.grad[q] & 15) == 0),
               "nb_adam_allreduce_p2p: peer %d: null or unaligned pointer", q);
  }
  // one block per SM at most: every block spins on the flags, so all of them have to be resident
  int grid = (int)((n / 4 + 511) / 512);
  if (grid > nb::kNumSMs) grid = nb::kNumSMs;
  nb::adam_allreduce_p2p_kernel<<<grid, 512, 0, nb::as_stream(stream)>>>(pp, rank, world, p, m, v, n, hyper, done_counter);
  return nb::check_launch("nb_adam_allreduce_p2p");
}
