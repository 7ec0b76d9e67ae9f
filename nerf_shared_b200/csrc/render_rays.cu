// nb_render_rays_fwd: Renderer.render_rays (render_utils.py:67-174) for inference as ONE C-ABI call.
//
// SURVEY.md section 8b lists a fused `render_rays_fwd` among the library's entry points: rays in, maps out, nothing of
// the per-stage plumbing visible to the caller.  The call issues the same kernels the per-stage entry points launch, in
// the same order on the caller's stream -- ray rows, stratified depths + points, coarse MLP (tensor cores), compositing,
// inverse-CDF resampling + merge, fine points, fine MLP, compositing -- on a caller-provided workspace, so one host call
// (one ctypes crossing instead of nine, no intermediate torch allocations) renders a chunk of rays; results are
// bit-identical to the per-stage sequence.
#include "common.cuh"

extern "C" {

static inline int64_t al256(int64_t x) { return (x + 255) / 256 * 256; }

// workspace layout (bytes, each section 256-byte aligned):
//   rows [n,11] | viewdirs [n,3] | z_c [n,S] | pts_c [n,S,3] | raw_c [n,S,4] | weights_c [n,S] | z_f [n,F] | pts_f [n,F,3] | raw_f [n,F,4]
int64_t nb_render_rays_workspace_bytes(int64_t n, int S, int I) {
  const int64_t F = S + I;
  return al256(n * 11 * 4) + al256(n * 3 * 4) + al256(n * S * 4) + al256(n * S * 12) + al256(n * S * 16) + al256(n * S * 4) +
         al256(n * F * 4) + al256(n * F * 12) + al256(n * F * 16);
}

// packed_coarse / packed_fine: nb_mlp_pack blobs (multires <= 10 / multires_views <= 4).  t_vals: linspace(0,1,S);
// u_det: linspace(0,1,I); t_rand [n,S], u [n,I], noise_c [n,S], noise_f [n,S+I]: the caller's random draws or NULL
// (perturb = 0 / raw_noise_std = 0).  rays_o / rays_d: [n,3] (NDC-warped already if the renderer does that), view_src
// [n,3]: the directions the view encoding is taken from (normalised here, render_utils.py:211).
// Outputs (fine maps, coarse maps, z_std; any may be NULL): rgb [n,3], disp [n], acc [n], rgb0 [n,3], disp0 [n],
// acc0 [n], z_std [n]; raw_out [n,S+I,4] (NULL: stays in the workspace).  I = 0: the coarse pass only (maps in rgb/disp/acc).
int nb_render_rays_fwd(const void* packed_coarse, const void* packed_fine, int L_pts, int L_dirs, const float* rays_o,
                       const float* rays_d, const float* view_src, float near_, float far_, int64_t n, int S, int I,
                       const float* t_vals, const float* t_rand, const float* u_det, const float* u, const float* noise_c,
                       const float* noise_f, int lindisp, int white_bkgd, void* workspace, int64_t workspace_bytes,
                       float* rgb, float* disp, float* acc, float* rgb0, float* disp0, float* acc0, float* z_std,
                       float* raw_out, void* stream) {
  if (n == 0) return 0;
  NB_REQUIRE(packed_coarse && rays_o && rays_d && view_src && t_vals && workspace && S >= 1 && I >= 0,
             "nb_render_rays_fwd: bad arguments");
  NB_REQUIRE(I == 0 || (u_det || u), "nb_render_rays_fwd: need u or u_det for the importance samples");
  NB_REQUIRE(workspace_bytes >= nb_render_rays_workspace_bytes(n, S, I) && ((uintptr_t)workspace & 255) == 0,
             "nb_render_rays_fwd: workspace too small (%lld < %lld bytes) or not 256-byte aligned", (long long)workspace_bytes,
             (long long)nb_render_rays_workspace_bytes(n, S, I));
  const int64_t F = S + I;
  char* w = (char*)workspace;
  auto take = [&](int64_t bytes) { float* p = (float*)w; w += al256(bytes); return p; };
  float* rows = take(n * 11 * 4);
  float* vd = take(n * 3 * 4);
  float* z_c = take(n * S * 4);
  float* pts_c = take(n * S * 12);
  float* raw_c = take(n * S * 16);
  float* w_c = take(n * S * 4);
  float* z_f = take(n * F * 4);
  float* pts_f = take(n * F * 12);
  float* raw_f = raw_out ? raw_out : take(n * F * 16);
#define NB_RC(x) do { int rc_ = (x); if (rc_) return rc_; } while (0)
  NB_RC(nb_ray_rows(rays_o, rays_d, view_src, near_, far_, n, rows, vd, stream));
  NB_RC(nb_sample_coarse(rows, n, 11, t_vals, t_rand, S, lindisp, z_c, pts_c, stream));
  NB_RC(nb_mlp_fwd_tc(packed_coarse, L_pts, L_dirs, pts_c, vd, n * S, S, raw_c, nullptr, stream));
  if (I == 0)
    return nb_composite_fwd(raw_c, z_c, rays_d, noise_c, n, S, white_bkgd, rgb, disp, acc, nullptr, nullptr, stream);
  NB_RC(nb_composite_fwd(raw_c, z_c, rays_d, noise_c, n, S, white_bkgd, rgb0, disp0, acc0, w_c, nullptr, stream));
  NB_RC(nb_sample_pdf_merge(z_c, w_c, u, u_det, n, S, I, z_f, z_std, nullptr, stream));
  NB_RC(nb_ray_points_fwd(rays_o, rays_d, z_f, n, (int)F, pts_f, stream));
  NB_RC(nb_mlp_fwd_tc(packed_fine ? packed_fine : packed_coarse, L_pts, L_dirs, pts_f, vd, n * F, (int)F, raw_f, nullptr, stream));
  return nb_composite_fwd(raw_f, z_f, rays_d, noise_f, n, (int)F, white_bkgd, rgb, disp, acc, nullptr, nullptr, stream);
#undef NB_RC
}

}  // extern "C"

// ---- nb_render_rays_fwd_noalloc: the same call for callers that cannot allocate --------------------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The workspace of nb_render_rays_fwd
// lives in this file-scope device array: NB_RENDER_RAYS_NOALLOC_MAX_BYTES (16 MiB) on every GPU that loads the library,
// held until the process exits.  In host code `g_render_scratch` is only a stand-in for the symbol, not a device address:
// the address on a device comes from cudaGetSymbolAddress, taken once per device with that device current.
static constexpr int64_t kRenderScratchBytes = 16ll << 20;
static __device__ __align__(256) unsigned char g_render_scratch[kRenderScratchBytes];

namespace {
constexpr int kMaxDevices = 64;
void* g_render_scratch_ptr[kMaxDevices];   // device address of g_render_scratch per device ordinal, NULL = not taken yet

// the current device's address of the scratch array (looked up on the first call on that device, then cached)
int render_scratch_for_current_device(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    nb::set_error("nb_render_rays_fwd_noalloc: cudaGetDevice failed or device ordinal %d out of range: %s", dev,
                  cudaGetErrorString(e));
    return e != cudaSuccess ? (int)e : NB_E_DEVICE;
  }
  void* p = __atomic_load_n(&g_render_scratch_ptr[dev], __ATOMIC_ACQUIRE);
  if (p == nullptr) {
    e = cudaGetSymbolAddress((void**)&p, g_render_scratch);
    if (e != cudaSuccess || p == nullptr) {
      nb::set_error("nb_render_rays_fwd_noalloc: cudaGetSymbolAddress(g_render_scratch) on device %d: %s", dev,
                    cudaGetErrorString(e));
      return e != cudaSuccess ? (int)e : NB_E_DEVICE;
    }
    __atomic_store_n(&g_render_scratch_ptr[dev], p, __ATOMIC_RELEASE);   // racing first calls store the same address
  }
  *out = p;
  return 0;
}
}  // namespace

extern "C" {

// Same arguments and results as nb_render_rays_fwd without `workspace` / `workspace_bytes`.  Calls no cudaMalloc / cudaFree
// and no allocating library routine.  Every call on a device shares the one array, which keeps the previous call's
// contents: every section of the workspace is written in full by the stage that produces it before the next stage reads
// it (nothing has to start at zero, so nothing is cleared), but two calls that overlap on the GPU (different streams or
// threads, same device) race on it -- serialise them.  Sizes whose workspace exceeds the array are rejected.
int nb_render_rays_fwd_noalloc(const void* packed_coarse, const void* packed_fine, int L_pts, int L_dirs, const float* rays_o,
                               const float* rays_d, const float* view_src, float near_, float far_, int64_t n, int S, int I,
                               const float* t_vals, const float* t_rand, const float* u_det, const float* u,
                               const float* noise_c, const float* noise_f, int lindisp, int white_bkgd, float* rgb, float* disp,
                               float* acc, float* rgb0, float* disp0, float* acc0, float* z_std, float* raw_out, void* stream) {
  if (n == 0) return 0;
  NB_REQUIRE(n > 0 && S >= 1 && I >= 0, "nb_render_rays_fwd_noalloc: bad sizes (n = %lld, S = %d, I = %d)", (long long)n, S, I);
  const int64_t need = nb_render_rays_workspace_bytes(n, S, I);
  if (need > kRenderScratchBytes) {
    nb::set_error("nb_render_rays_fwd_noalloc: %lld rays x (%d + %d) samples need %lld bytes of workspace, the built-in scratch "
                  "holds %lld (NB_RENDER_RAYS_NOALLOC_MAX_BYTES): render fewer rays per call or use nb_render_rays_fwd",
                  (long long)n, S, I, (long long)need, (long long)kRenderScratchBytes);
    return NB_E_ARG;
  }
  void* ws = nullptr;
  int rc = render_scratch_for_current_device(&ws);
  if (rc) return rc;
  return nb_render_rays_fwd(packed_coarse, packed_fine, L_pts, L_dirs, rays_o, rays_d, view_src, near_, far_, n, S, I, t_vals,
                            t_rand, u_det, u, noise_c, noise_f, lindisp, white_bkgd, ws, kRenderScratchBytes, rgb, disp, acc,
                            rgb0, disp0, acc0, z_std, raw_out, stream);
}

}  // extern "C"
