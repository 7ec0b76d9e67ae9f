/* nerf_b200.h -- C ABI of libnerf_b200.so (sm_100a only).
 *
 * The reference (stanford-iprl-lab/nerf_shared) has no FFI of its own: its hot path is a
 * stream of ATen ops launched from Python.  Each entry point below replaces the ATen call
 * sites of one stage of that path; the `replaces:` line cites them (file:line in the
 * reference tree).  The Python side (nerf_shared_b200/_lib.py) binds these with ctypes;
 * INTEGRATION.md shows the stub a reference maintainer would add.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless its name ends in _host; row-major, fp32
 *     unless stated; the library never allocates or frees caller-visible memory;
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*), does no
 *     host synchronisation and is CUDA-graph capturable;
 *   - return value: 0 = ok, <0 = argument error (NB_E_*), >0 = cudaError_t;
 *     nb_last_error() returns a thread-local message for the last non-zero return;
 *   - there is no CPU fallback: nb_init() fails unless the device is compute capability 10.x.
 */
#ifndef NERF_B200_H_
#define NERF_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NB_VERSION 100
#define NB_API __attribute__((visibility("default")))

#define NB_E_ARG (-1)       /* bad pointer / size / flag */
#define NB_E_UNSUPPORTED (-2) /* configuration the sm_100a kernels do not implement */
#define NB_E_DEVICE (-3)    /* not an sm_100 device */

/* ---- library state --------------------------------------------------------------- */
NB_API int nb_version(void);
NB_API const char* nb_last_error(void);
/* Checks that `device` is sm_100 and sets the kernels' dynamic shared-memory attributes. */
NB_API int nb_init(int device);
NB_API int nb_shutdown(void);
/* SM budget of the NEXT launches issued by the calling thread (0 = whole GPU): the persistent tensor-core MLP kernels
 * (forward, dgrad) use at most `mlp_clusters` CTA pairs, the weight-gradient kernel at most `wgrad_ctas` CTAs (placed as
 * pairs).  Used to run a bandwidth-bound and a tensor-bound kernel side by side on disjoint SMs; results do not depend
 * on it.  New in this library (the reference has no launch control). */
NB_API int nb_set_launch_limits(int mlp_clusters, int wgrad_ctas);
/* Number of kernels this library has launched since load (all streams); for bench.py. */
NB_API int64_t nb_launch_count(void);

/* ---- rays ------------------------------------------------------------------------ */
/* Pinhole rays for the full H x W grid.
 * replaces: utils.get_rays utils.py:33-42 (+ viewdir normalisation render_utils.py:211).
 * c2w: 3 rows x 4 cols with row stride `c2w_ld` floats.  rays_o/rays_d: [H*W,3].
 * viewdirs: [H*W,3] or NULL.  Bit-exact with the reference's CPU result for rays_o/rays_d. */
NB_API int nb_get_rays_fwd(int H, int W, float fx, float fy, float cx, float cy, const float* c2w, int c2w_ld,
                    float* rays_o, float* rays_d, float* viewdirs, void* stream);
/* Same for a list of `n` pixels, pix[k] = y*W + x (int64): only the selected rays are made.
 * replaces: get_rays + fancy-index at utils.py:419,437-438 and demo_est_rel_pose.py:82-84. */
NB_API int nb_get_rays_sel_fwd(int H, int W, float fx, float fy, float cx, float cy, const float* c2w, int c2w_ld,
                        const int64_t* pix, int64_t n, float* rays_o, float* rays_d, float* viewdirs, void* stream);
/* Gradient of the above wrt c2w[:3,:4]; pix may be NULL (full grid, n = H*W).
 * d_rays_o / d_rays_d: [n,3] or NULL.  d_c2w: 12 floats, row-major 3x4, ACCUMULATED into
 * (caller zeroes).  replaces: autograd of utils.py:37-41 (demo_est_rel_pose.py:97). */
NB_API int nb_get_rays_bwd(int H, int W, float fx, float fy, float cx, float cy, const int64_t* pix, int64_t n,
                    const float* d_rays_o, const float* d_rays_d, float* d_c2w, void* stream);
/* viewdirs = rays_d / |rays_d| and its backward.  replaces: render_utils.py:211. */
NB_API int nb_normalize_fwd(const float* v, int64_t n, float* out, void* stream);
NB_API int nb_normalize_bwd(const float* v, const float* d_out, int64_t n, float* d_v, void* stream);
/* The [n,11] ray rows Renderer.render builds: o(3) d(3) near far viewdirs(3), viewdirs = view_src / |view_src|
 * (the rays before the NDC warp); `viewdirs` ([n,3], contiguous copy) may be NULL.
 * replaces: render_utils.py:211 + :223-226 (normalise, two fills, cat) for rays that carry no gradient. */
NB_API int nb_ray_rows(const float* rays_o, const float* rays_d, const float* view_src, float near_, float far_, int64_t n,
                float* rows, float* viewdirs, void* stream);
/* NDC reparametrisation.  replaces: utils.ndc_rays utils.py:54-71 (bit-exact forward). */
NB_API int nb_ndc_rays_fwd(int H, int W, double focal, double near_, const float* rays_o, const float* rays_d, int64_t n,
                    float* out_o, float* out_d, void* stream);
NB_API int nb_ndc_rays_bwd(int H, int W, double focal, double near_, const float* rays_o, const float* rays_d,
                    const float* d_out_o, const float* d_out_d, int64_t n, float* d_rays_o, float* d_rays_d,
                    void* stream);

/* ---- sampling along the ray ---------------------------------------------------------- */
/* Stratified depths and sample positions.
 * replaces: render_utils.py:105-131.
 * rays: [n, ld] rows = o(3) d(3) near far [viewdirs(3)] exactly as Renderer.render builds
 * them (render_utils.py:223-226).  t_vals: [S] = torch.linspace(0,1,S) (passed in because
 * linspace is not reproducible by a closed form).  t_rand: [n,S] uniform
 * draws or NULL (perturb == 0).  z: [n,S] out (bit-exact).  pts: [n,S,3] out or NULL. */
NB_API int nb_sample_coarse(const float* rays, int64_t n, int ld, const float* t_vals, const float* t_rand, int S,
                     int lindisp, float* z, float* pts, void* stream);
/* pts[r,s,:] = o[r] + d[r]*z[r,s] and its backward (d_o = sum_s d_pts, d_d = sum_s z d_pts).
 * replaces: render_utils.py:131,148 and their autograd. */
NB_API int nb_ray_points_fwd(const float* rays_o, const float* rays_d, const float* z, int64_t n, int S, float* pts,
                      void* stream);
NB_API int nb_ray_points_bwd(const float* z, const float* d_pts, int64_t n, int S, float* d_rays_o, float* d_rays_d,
                      void* stream);
/* Hierarchical resampling: inverse-CDF samples of the coarse weights, merged into the
 * coarse depths and sorted.  replaces: utils.sample_pdf utils.py:74-117 and the
 * sort/cat/std at render_utils.py:141-147,168.
 * z: [n,S], weights: [n,S] (the interior S-2 are used), u: [n,I] or NULL (det: u=u_det[I],
 * u_det = torch.linspace(0,1,I)).  Outputs: z_out [n,S+I] sorted; z_std [n] or NULL;
 * z_samples [n,I] or NULL (unsorted, as sample_pdf returns them). */
NB_API int nb_sample_pdf_merge(const float* z, const float* weights, const float* u, const float* u_det, int64_t n, int S,
                        int I, float* z_out, float* z_std, float* z_samples, void* stream);
/* utils.sample_pdf on explicit bins (utils.py:74-117): bins [n,B], weights [n,B-1],
 * u [n,I] or NULL (det, u_det [I]) -> samples [n,I]. */
NB_API int nb_sample_pdf(const float* bins, const float* weights, const float* u, const float* u_det, int64_t n, int B,
                         int I, float* samples, void* stream);
/* The inverse-CDF step alone, from a caller-supplied cdf (bit-exact searchsorted indices).
 * bins/cdf: [n,B]; u: [n,I]; samples: [n,I]; inds: [n,I] int64 (searchsorted right=True).
 * replaces: utils.py:100-115. */
NB_API int nb_invert_cdf(const float* bins, const float* cdf, const float* u, int64_t n, int B, int I, float* samples,
                  int64_t* inds, void* stream);

/* ---- volume compositing -------------------------------------------------------------- */
/* replaces: Renderer.raw2outputs render_utils.py:241-290.
 * raw: [n,S,4]; z: [n,S]; rays_d: [n,3]; noise: [n,S] (already scaled by raw_noise_std) or NULL.
 * Outputs (any may be NULL): rgb_map [n,3], disp [n], acc [n], weights [n,S], depth [n]. */
NB_API int nb_composite_fwd(const float* raw, const float* z, const float* rays_d, const float* noise, int64_t n, int S,
                     int white_bkgd, float* rgb_map, float* disp, float* acc, float* weights, float* depth,
                     void* stream);
/* Hand-written reverse-scan backward.  Upstream grads d_rgb [n,3], d_disp, d_acc, d_depth
 * [n] (any may be NULL = zero).  Outputs: d_raw [n,S,4]; d_rays_d [n,3] or NULL
 * (through |rays_d| in the interval lengths, render_utils.py:259). */
NB_API int nb_composite_bwd(const float* raw, const float* z, const float* rays_d, const float* noise, int64_t n, int S,
                     int white_bkgd, const float* d_rgb, const float* d_disp, const float* d_acc,
                     const float* d_depth, float* d_raw, float* d_rays_d, void* stream);

/* ---- photometric loss of the training step --------------------------------------------- */
/* out4 = { mean((rgb-t)^2), mean((rgb0-t)^2) (0 when rgb0 is NULL), their sum, -10 log10(out4[0]) } over n3 floats.
 * replaces: utils.img2mse / mse2psnr utils.py:24,27 as used at main.py:87-99. */
NB_API int nb_mse_pair_fwd(const float* rgb, const float* rgb0, const float* target, int64_t n3, float* out4, void* stream);
/* g_fine / g_coarse / g_total = upstream gradients of out4[0] / [1] / [2], device scalars, NULL = zero.
 * d_rgb = (g_fine+g_total) * 2 (rgb-t)/n3, d_rgb0 = (g_coarse+g_total) * 2 (rgb0-t)/n3 (d_rgb0 may be NULL).
 * replaces: autograd of the above (main.py:103). */
NB_API int nb_mse_pair_bwd(const float* rgb, const float* rgb0, const float* target, int64_t n3, const float* g_fine,
                    const float* g_coarse, const float* g_total, float* d_rgb, float* d_rgb0, void* stream);

/* Training fusion: compositing forward + the MSE term of this map against `target` [n,3] + the compositing backward of
 * that term, one pass over raw.  inv_n3 = 1 / (3 * rays of the GLOBAL batch the mean runs over).  loss_sum (one float,
 * zeroed by the caller) += sum (rgb_map - target)^2 * inv_n3.  d_raw = d(that MSE term)/d raw.  rgb_map / disp / acc /
 * weights may be NULL.
 * replaces: raw2outputs render_utils.py:241-290 + img2mse utils.py:24 (main.py:87,93) + their autograd (main.py:103). */
NB_API int nb_composite_train(const float* raw, const float* z, const float* rays_d, const float* noise, const float* target,
                       int64_t n, int S, int white_bkgd, float inv_n3, float* rgb_map, float* disp, float* acc,
                       float* weights, float* d_raw, float* loss_sum, void* stream);
/* out4 = { acc_fine[0], acc_coarse[0] (0 when NULL), their sum, -10 log10(acc_fine[0]) }; both accumulators are reset
 * to zero for the next step.  replaces: main.py:87-99 (loss sum, mse2psnr utils.py:27). */
NB_API int nb_loss_finalize(float* acc_fine, float* acc_coarse, float* out4, void* stream);

/* ---- the inference path as one call ---------------------------------------------------------------------------- */
/* replaces: Renderer.render_rays render_utils.py:67-174 under torch.no_grad() (ray rows, stratified depths, coarse MLP,
 * compositing, inverse-CDF resampling + merge, fine points, fine MLP, compositing) for one chunk of rays: the kernels of
 * the per-stage entry points, issued back to back on `stream` on a caller-provided workspace
 * (nb_render_rays_workspace_bytes, 256-byte aligned); results are bit-identical to the per-stage sequence.
 * packed_*: nb_mlp_pack blobs (packed_fine NULL: the coarse network is reused, render_utils.py:156); t_vals / u_det:
 * linspace(0,1,S) / linspace(0,1,I); t_rand [n,S], u [n,I], noise_c [n,S], noise_f [n,S+I]: the caller's random draws
 * (noise already scaled by raw_noise_std) or NULL.  Outputs may be NULL; raw_out [n,S+I,4] NULL keeps raw in the workspace.
 * I = 0: coarse pass only, its maps in rgb / disp / acc. */
NB_API int64_t nb_render_rays_workspace_bytes(int64_t n, int S, int I);
NB_API int nb_render_rays_fwd(const void* packed_coarse, const void* packed_fine, int L_pts, int L_dirs, const float* rays_o,
                              const float* rays_d, const float* view_src, float near_, float far_, int64_t n, int S, int I,
                              const float* t_vals, const float* t_rand, const float* u_det, const float* u,
                              const float* noise_c, const float* noise_f, int lindisp, int white_bkgd, void* workspace,
                              int64_t workspace_bytes, float* rgb, float* disp, float* acc, float* rgb0, float* disp0,
                              float* acc0, float* z_std, float* raw_out, void* stream);

/* nb_render_rays_fwd for callers that cannot allocate.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same arguments, same kernels and bit-identical results as nb_render_rays_fwd, minus `workspace`: the
 * workspace is a file-scope device array of NB_RENDER_RAYS_NOALLOC_MAX_BYTES inside the library (every GPU that loads the
 * library holds it until the process exits; its device address is taken with cudaGetSymbolAddress once per device).  The
 * call allocates nothing (no cudaMalloc / cudaFree, no allocating library routine).  Sizes with
 * nb_render_rays_workspace_bytes(n, S, I) > NB_RENDER_RAYS_NOALLOC_MAX_BYTES (about 1,900 rays at S = 64, I = 128) return
 * NB_E_ARG with a message.  All calls on a device share the array and it keeps the last call's contents; every section
 * is rewritten in full before it is read, so nothing needs clearing -- but calls that OVERLAP on the GPU (two streams or
 * threads on one device) race on it and must be serialised by the caller. */
#define NB_RENDER_RAYS_NOALLOC_MAX_BYTES (16ll << 20)
NB_API int nb_render_rays_fwd_noalloc(const void* packed_coarse, const void* packed_fine, int L_pts, int L_dirs,
                                      const float* rays_o, const float* rays_d, const float* view_src, float near_, float far_,
                                      int64_t n, int S, int I, const float* t_vals, const float* t_rand, const float* u_det,
                                      const float* u, const float* noise_c, const float* noise_f, int lindisp, int white_bkgd,
                                      float* rgb, float* disp, float* acc, float* rgb0, float* disp0, float* acc0, float* z_std,
                                      float* raw_out, void* stream);

/* ---- frame output ------------------------------------------------------------------------ */
/* out[i] = (uint8)(255 * clip(x[i], 0, 1)), NumPy's truncating cast.  replaces: utils.to8b utils.py:30 (host NumPy) in
 * render_path / render_from_batch_poses (utils.py:330-357, render_utils.py:293-319). */
NB_API int nb_to8b(const float* x, int64_t n, uint8_t* out, void* stream);

/* ---- the NeRF MLP (positional encoding fused) ---------------------------------------- */
/* Network shape (nerf.py:61-94, use_viewdirs=True): D=8, W=256, skip after layer 4,
 * L_pts / L_dirs encoding octaves (10 / 4 in every shipped config but one). */
typedef struct NbMlpWeights {
  /* fp32 nn.Linear parameters, PyTorch layout [out,in] row-major */
  const float* pts_w[8];
  const float* pts_b[8];
  const float* feature_w; const float* feature_b; /* [256,256] */
  const float* alpha_w;   const float* alpha_b;   /* [1,256]   */
  const float* views_w;   const float* views_b;   /* [128,256+in_views] */
  const float* rgb_w;     const float* rgb_b;     /* [3,128]   */
  int L_pts;   /* multires        -> input_ch = 3+6*L_pts */
  int L_dirs;  /* multires_views  -> input_ch_views = 3+6*L_dirs */
} NbMlpWeights;

/* fp32 CUDA-core evaluation (parity / debugging path, exact fp32 operands).
 * pts: [M,3]; viewdirs: [M/S,3] (one per ray, S consecutive samples share it); raw: [M,4].
 * replaces: NeRF.forward nerf.py:96-134. */
NB_API int nb_mlp_fwd_f32(const NbMlpWeights* w, const float* pts, const float* viewdirs, int64_t M, int S, float* raw,
                   void* stream);
/* The use_viewdirs=False model (nerf.py:93-94, 131-132): the same trunk, then raw = output_linear(h).
 * Only pts_w / pts_b / L_pts of `w` are read; out_w [out_ch,256], out_b [out_ch]; raw: [M,out_ch].
 * fp32, forward only (no shipped config trains this model). */
NB_API int nb_mlp_fwd_f32_noview(const NbMlpWeights* w, const float* out_w, const float* out_b, int out_ch,
                                 const float* pts, int64_t M, float* raw, void* stream);

/* fp32 TRAINING path (CUDA cores, layer-by-layer GEMMs, saved fp32 activations): the model variants the tensor-core
 * kernels do not tile -- encodings up to multires 15 / 6 (configs/stonehenge.txt) and use_viewdirs=False -- and the
 * exact-operand reference of the bf16 kernels' gradients.  acts: nb_mlp_f32_act_floats(...) floats.
 * out_w / out_b / out_ch: output_linear of a use_viewdirs=False model (raw is then [M, out_ch], viewdirs unused), else NULL / 0.
 * replaces: NeRF.forward / MLP nerf.py:96-134 and its autograd (main.py:103). */
NB_API int64_t nb_mlp_f32_act_floats(int L_pts, int L_dirs, int noview, int64_t M);
NB_API int nb_mlp_f32_fwd_train(const NbMlpWeights* w, const float* out_w, const float* out_b, int out_ch, const float* pts,
                                const float* viewdirs, int64_t M, int S, float* raw, float* acts, void* stream);
/* Weight / bias gradients ACCUMULATE into grads + off[i], i in the order of nb_mlp_bwd_tc's offsets (use_viewdirs=False:
 * 16 / 17 = output_linear weight / bias, 18..23 ignored).  scratch: 2 * M * 256 floats. */
NB_API int nb_mlp_f32_bwd(const NbMlpWeights* w, const float* out_w, int out_ch, const float* acts, const float* d_raw, int64_t M,
                          const int64_t* off, float* grads, float* scratch, void* stream);

/* Size in bytes of the packed bf16 operand blob + fp32 side table used by the tensor-core
 * kernels for one model. */
NB_API int64_t nb_mlp_packed_bytes(int L_pts, int L_dirs);
/* Re-pack fp32 parameters into `packed` (call after every optimizer step). */
NB_API int nb_mlp_pack(const NbMlpWeights* w, void* packed, void* stream);
/* Persistent fused tcgen05/TMEM forward: encoding -> 10 GEMM layers -> raw, activations
 * never leave the SM.  bf16 operands, fp32 accumulation, sigma and rgb heads in fp32.
 * If `acts` is non-NULL the bf16 layer activations needed by the backward pass are stored
 * there (nb_mlp_act_bytes(M) bytes: per 128-sample tile 40 slabs of 16 KB in the kernels' own
 * swizzled operand layout; opaque to the caller). */
NB_API int nb_mlp_fwd_tc(const void* packed, int L_pts, int L_dirs, const float* pts, const float* viewdirs, int64_t M,
                  int S, float* raw, void* acts, void* stream);
NB_API int64_t nb_mlp_act_bytes(int64_t M);

/* Backward of the MLP from the activations nb_mlp_fwd_tc saved (autograd of nerf.py:110-134).
 * packed_bwd: transposed bf16 weight slabs (nb_mlp_pack_bwd, nb_mlp_bwd_packed_bytes bytes).
 * d_raw: [M,4] upstream gradient.  dA: scratch of nb_mlp_da_bytes(M) bytes (pre-activation
 * gradients of every layer, tiled like the activations, followed by 544 KB in which the weight gradients whose rows
 * are not 16-byte aligned are reduced before they are added to `grads`; private to the call).  grads: flat fp32 buffer the weight /
 * bias gradients are ACCUMULATED into (caller zeroes), laid out by grad_off_host24 (HOST array of
 * 24 float offsets: pts_linears weights 0..7, their biases 8..15, feature w/b 16,17, alpha w/b
 * 18,19, views w/b 20,21, rgb w/b 22,23); grads == NULL runs only the dgrad chain. */
NB_API int64_t nb_mlp_bwd_packed_bytes(int L_pts, int L_dirs);
NB_API int nb_mlp_pack_bwd(const NbMlpWeights* w, void* packed_bwd, void* stream);
NB_API int64_t nb_mlp_da_bytes(int64_t M);
NB_API int nb_mlp_bwd_tc(const void* packed_bwd, int L_pts, int L_dirs, const void* acts, const float* d_raw, int64_t M,
                         void* dA, const int64_t* grad_off_host24, float* grads, void* stream);
/* The same in two separately launchable parts (bit 0: the dgrad chain, writes dA; bit 1: weight gradients + head
 * gradients from acts / dA / d_raw), so that a caller can put other work between them or on another stream. */
NB_API int nb_mlp_bwd_tc_parts(const void* packed_bwd, int L_pts, int L_dirs, const void* acts, const float* d_raw, int64_t M,
                        void* dA, const int64_t* grad_off_host24, float* grads, int parts, void* stream);
/* Streamed backward: the dgrad chain on `dgrad_clusters` CTA pairs and, beside it on a library-owned auxiliary stream,
 * the weight-gradient kernel on `wgrad_ctas` CTAs consuming every tile's activation gradients as soon as the chain has
 * written them (per-tile arrival counters in `ready`: [ceil(M/128)] ints, zeroed by the call) -- from L2 instead of HBM.
 * 0 for either count = the default split (37 pairs / 74 CTAs).  Same results as nb_mlp_bwd_tc up to the order of the
 * gradient atomics; the call returns with `stream` waiting for both kernels. */
NB_API int nb_mlp_bwd_tc_streamed(const void* packed_bwd, int L_pts, int L_dirs, const void* acts, const float* d_raw, int64_t M,
                           void* dA, const int64_t* grad_off_host24, float* grads, int* ready, int dgrad_clusters,
                           int wgrad_ctas, void* stream);

/* Gradient wrt the MLP's INPUTS from the dA tiles nb_mlp_bwd_tc left behind (pose estimation:
 * demo_est_rel_pose.py:87-97 backpropagates through frozen weights to the sample positions and view
 * directions).  replaces: autograd of the two Embedder calls, the first layer, the skip connection and
 * the view branch input, nerf.py:97-104,110-118,124.
 * packed_in: nb_mlp_pack_ingrad output (nb_mlp_ingrad_packed_bytes bytes).  pts [M,3], viewdirs [M/S,3]
 * as given to nb_mlp_fwd_tc.  d_pts [M,3] is written; d_viewdirs [M/S,3] is ACCUMULATED into (caller
 * zeroes): the reference broadcasts one view direction over the S samples of a ray (nerf.py:101). */
NB_API int64_t nb_mlp_ingrad_packed_bytes(int L_pts, int L_dirs);
NB_API int nb_mlp_pack_ingrad(const NbMlpWeights* w, void* packed_in, void* stream);
NB_API int nb_mlp_bwd_inputs(const void* packed_in, int L_pts, int L_dirs, const void* dA, const float* pts,
                             const float* viewdirs, int64_t M, int S, float* d_pts, float* d_viewdirs, void* stream);

/* ---- optimizer ------------------------------------------------------------------------- */
/* torch.optim.Adam (utils.py:170) on flat fp32 buffers; step is 1-based.
 * replaces: Adam.step main.py:104. */
NB_API int nb_adam_step(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2,
                 float eps, int step, float grad_scale, void* stream);

/* CUDA-graph form of the same update: nb_adam_hyper (outside the graph, once per step) writes the
 * step's scalars -- lr/bias_correction1, beta1, beta2, eps, 1/sqrt(bias_correction2), grad_scale --
 * into `hyper` (8 device floats); nb_adam_step_dev (inside the graph) reads them from there.
 * replaces: Adam.step + the per-iteration lr write main.py:104-112. */
NB_API int nb_adam_hyper(float* hyper, float lr, float beta1, float beta2, float eps, int step, float grad_scale,
                         int epoch, void* stream);
/* nb_adam_hyper plus the copy of a training step's two inputs (n1 floats src1 -> dst1: the ray batch; n2 floats
 * src2 -> dst2: the target colours) into the static buffers a captured step graph reads, in ONE launch
 * (train.GraphedTrainStep issues it in front of every replay). */
NB_API int nb_train_step_prologue(float* hyper, float lr, float beta1, float beta2, float eps, int step, float grad_scale,
                                  int epoch, const float* src1, float* dst1, int64_t n1, const float* src2, float* dst2, int64_t n2,
                                  void* stream);
NB_API int nb_adam_step_dev(float* p, const float* g, float* m, float* v, int64_t n, const float* hyper, void* stream);
/* Multi-GPU: the gradient all-reduce fused with the Adam update, one launch per rank over NVLink peer memory
 * (two-shot all-reduce in place on the ranks' symmetric gradient buffers, then the replicated update; see adam.cu).
 * peer_grads[world] / peer_flags[world]: HOST arrays of device pointers to every rank's gradient buffer (n floats, n a
 * multiple of 4, 16-byte aligned) and flag block (32 zero-initialised ints), own rank included, mapped into this
 * process (torch symmetric memory / CUDA IPC); done_counter: one zero-initialised device uint32 of this rank; hyper as
 * nb_adam_step_dev, with hyper[6] = `epoch` of nb_adam_hyper / nb_train_step_prologue: a counter that grows by one per call of this
 * function and never repeats (unlike `step`, which a caller may roll back), identical on all ranks.
 * multicast_grads: the NVLS multicast mapping of the same gradient buffers (the NVSwitch then reduces and broadcasts:
 * multimem.ld_reduce / multimem.st) or NULL (peer loads / stores).  Every rank must launch it once per step.  replaces: the dist.all_reduce a data-parallel main.py would add before
 * optimizer.step() (main.py:103-104) + Adam.step. */
NB_API int nb_adam_allreduce_p2p(float* p, float* m, float* v, int64_t n, const float* hyper, void* const* peer_grads,
                                 void* const* peer_flags, void* multicast_grads, int rank, int world,
                                 unsigned int* done_counter, void* stream);

/* ---- self test ------------------------------------------------------------------------- */
/* D[128,128] = A[128,K] * B[128,K]^T (bf16-rounded operands, fp32 accumulate) through the same
 * swizzle / descriptor / tcgen05 sequence as nb_mlp_fwd_tc; K in {64,128,192,256}. */
NB_API int nb_debug_umma(const float* A, const float* B, int K, float* D, void* stream);
/* D[128,128] = A[K,128]^T * B[K,128] with both operands MN-major (the wgrad operand mode); K in {64,128}. */
NB_API int nb_debug_umma_mn(const float* A, const float* B, int K, float* D, void* stream);

/* start / end timestamps (globaltimer, ns: [2 * cta], [2 * cta + 1]) of the CTAs of the last weight-gradient launch on the
 * current device, copied to host memory (synchronises the device); n <= 296 values.  tools/wgrad_skew.py */
NB_API int nb_debug_wgrad_trace(unsigned long long* host_out, int n);
/* (The tensor-pipe micro-benchmark nb_debug_umma_rate lives in csrc/umma_rate.cu and is NOT part of the default build:
 * `make EXTRA_SRCS=umma_rate.cu`; tools/umma_rate.py declares its prototype itself.) */

#ifdef __cplusplus
}
#endif
#endif /* NERF_B200_H_ */
