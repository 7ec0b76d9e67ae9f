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
