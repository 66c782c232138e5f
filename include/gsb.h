/* gsb.h -- C ABI of libgsb.so, the sm_100a kernels behind gslam's rasterization
 * hot path (reference: /root/reference/gslam/rasterization.py).
 *
 * Conventions:
 *  - every pointer is a DEVICE pointer unless its name ends in _host;
 *  - the caller owns all memory; the library never allocates device memory and
 *    keeps no global mutable state except a per-thread last-error string;
 *  - all work is enqueued on `stream` (a cudaStream_t passed as void*); no
 *    call synchronises the device except gsb_read_i64_host;
 *  - every entry point returns 0 on success, a positive cudaError_t or a
 *    negative GSB_E* code on failure; nothing throws or exits.
 *
 * Splat record ("rec"), one per (camera, Gaussian), 12 floats = 48 bytes:
 *    [0] x  [1] y  [2] opacity  [3] conic a  [4] conic b  [5] conic c
 *    [6] r  [7] g  [8] b        [9] depth   [10] beta    [11] unused (0)
 * It replaces the separate means2d / conics / opacities / colors tensors that
 * the reference builds at gslam/rasterization.py:153-256; the gradient record
 * v_rec has the same layout.
 */
#ifndef GSB_H
#define GSB_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GSB_REC 12            /* floats per splat record              */
#define GSB_CH 5              /* composited channels: r g b depth beta */
#define GSB_TILE 16           /* tile edge in pixels (reference tile_size=16) */

#define GSB_CSTRIDE 32         /* ints between per-tile counters: one 128-byte L2 line each,
                                 because the L2 atomic unit serialises atomics per line */

#define GSB_EINVAL (-1)       /* bad argument                         */
#define GSB_EWORKSPACE (-2)   /* workspace too small                  */

int gsb_version(void);
const char* gsb_last_error(void);

/* Blocking 8-byte device->host read (the one explicit sync of the path: the
 * intersection count that sizes the key buffers; upstream hides the same sync
 * inside isect_tiles, reference gslam/rasterization.py:261). */
int gsb_read_i64_host(const int64_t* src, int64_t* dst_host, void* stream);

/* K1+K8: activations + EWA projection + tile counting, unpacked.
 * Replaces gslam/rasterization.py:145-170, :187, :225-256 and the counting
 * half of isect_tiles (:261).  log_unc may be NULL (beta record slot = 0).
 * Outputs: radii[C,N] i32, means2d[C,N,2], rec[C,N,12], tiles_per_gauss[C,N] i32.
 * tile_counts[C*tile_h*tile_w*GSB_CSTRIDE] i32 (may be NULL; zeroed by the caller) receives
 * at index i*GSB_CSTRIDE the
 * number of intersections per (camera, tile) for the tile-binned sort below.
 * View-dependent colours (extension, SURVEY 8a "SH note"; the reference has none): when sh_coeffs
 * [N, sh_stride, 3] is not NULL the record colour is max(SH_deg(d) + 0.5, 0) with
 * d = normalize(mean - camera centre), real SH basis of degree sh_degree <= 3 over the first
 * (sh_degree+1)^2 coefficients, and logit_colors may be NULL. */
int gsb_project_fwd(const float* means, const float* quats, const float* log_scales,
                    const float* logit_opacities, const float* logit_colors,
                    const float* log_unc, const float* viewmats, const float* Ks,
                    int C, int N, int width, int height, float eps2d, float near_plane,
                    float far_plane, float radius_clip, int tile_w, int tile_h,
                    int32_t* radii, float* means2d, float* rec, int32_t* tiles_per_gauss,
                    int32_t* tile_counts, const float* sh_coeffs, int sh_degree, int sh_stride,
                    void* stream);

/* K2+K8 backward.  v_rec[C,N,12] carries the gradients of every record slot
 * except x,y, which come from v_means2d (element stride `v_means2d_stride`
 * floats between consecutive (c,n) rows, so a [.., 0:2] view of v_rec works).
 * Outputs (overwritten, summed over cameras): v_means[N,3] v_quats[N,4]
 * v_log_scales[N,3] v_logit_opacities[N] v_logit_colors[N,3] v_log_unc[N]
 * (any may be NULL when need_gauss_grads == 0) and v_viewmats[C,4,4] (must be
 * zeroed by the caller; accumulated atomically; may be NULL).  With SH colours pass
 * logit_colors = v_logit_colors = NULL and call gsb_sh_bwd afterwards.
 * Deterministic mode: det_partials (NULL = atomics) is scratch of gsb_project_bwd_det_floats(C, N)
 * floats; the per-block pose-gradient sums go there and are added in a fixed order, so v_viewmats is
 * bit-identical from run to run.
 * Row layout (the mapping step): with row_stride > 0 (a multiple of 4) every v_* pointer of the six
 * Gaussian gradients is the FIRST COLUMN of that parameter inside one row-major [N, row_stride] float
 * record (element k of Gaussian n at ptr[n * row_stride + k]; v_quats 16-byte aligned; v_log_unc is
 * written as 0 when log_unc is NULL), so a Gaussian's whole gradient is one contiguous row; v_vis
 * (may be NULL, addressed like v_logit_opacities) receives the number of views with radii > 0. */
size_t gsb_project_bwd_det_floats(int C, int N);
int gsb_project_bwd(const float* means, const float* quats, const float* log_scales,
                    const float* logit_opacities, const float* logit_colors,
                    const float* log_unc, const float* viewmats, const float* Ks,
                    int C, int N, int width, int height, float eps2d,
                    const int32_t* radii, const float* rec,
                    const float* v_means2d, int64_t v_means2d_stride, const float* v_rec,
                    int need_gauss_grads,
                    float* v_means, float* v_quats, float* v_log_scales,
                    float* v_logit_opacities, float* v_logit_colors, float* v_log_unc,
                    float* v_viewmats, float* det_partials, int64_t row_stride, float* v_vis,
                    void* stream);

/* Backward of the SH colours of gsb_project_fwd: v_sh[N, sh_stride, 3] is overwritten (summed over
 * cameras), the view-direction term is ADDED to v_means[N,3] (may be NULL), so call it after
 * gsb_project_bwd on the same stream.  The camera centre is treated as a constant.  row_stride > 0:
 * v_sh and v_means are columns of the row-major gradient record of gsb_project_bwd (same stride). */
int gsb_sh_bwd(const float* means, const float* sh_coeffs, int sh_degree, int sh_stride,
               const float* viewmats, const int32_t* radii, const float* v_rec, int C, int N,
               float* v_sh, float* v_means, int64_t row_stride, void* stream);

/* K3: packed projection (reference gslam/rasterization.py:390-419).  Two calls:
 * gsb_project_fwd (unpacked) then gsb_pack_visible compacts the rows with
 * radii > 0 in (c, n) order.  `cum` is the inclusive scan of (radii > 0). */
int gsb_pack_visible(const int32_t* radii, const float* means2d, const float* rec,
                     const int64_t* cum, int C, int N, int64_t* camera_ids,
                     int64_t* gaussian_ids, int32_t* radii_out, float* means2d_out,
                     float* depths_out, void* stream);

/* Inclusive int32 -> int64 scan (cumsum of tiles_per_gauss; the last element is
 * the intersection count I).  flag_positive != 0 scans (x > 0) instead of x. */
size_t gsb_scan_workspace_bytes(int64_t n);
int gsb_scan_i32_i64(const int32_t* in, int64_t n, int flag_positive, int64_t* out,
                     void* workspace, size_t workspace_bytes, void* stream);

/* K4: emit one (key, value) per (Gaussian, tile).  key = cam << (32+tile_n_bits)
 * | tile << 32 | bits(depth); value = c*N + n. */
int gsb_isect_emit(const float* means2d, const int32_t* radii, const float* rec,
                   const int64_t* cum, int C, int N, int tile_w, int tile_h,
                   int tile_n_bits, int64_t* isect_ids, int32_t* flatten_ids, void* stream);

/* K4: stable LSD radix sort of (u64 key, i32 value) pairs over bits [0,end_bit).
 * Sorted data ends in keys_out/vals_out; keys_in/vals_in are clobbered. */
size_t gsb_sort_workspace_bytes(int64_t n);
int gsb_sort_pairs(int64_t* keys_in, int32_t* vals_in, int64_t* keys_out, int32_t* vals_out,
                   int64_t n, int end_bit, void* workspace, size_t workspace_bytes,
                   void* stream);

/* K4+K5, tile-binned form (the product path; bit-identical to emit + stable sort +
 * offset encode, see csrc/binning.cu).  No step needs the intersection count I on
 * the host: it is offsets[C*T] (and *n_isects).
 *   gsb_tile_scan    : exclusive scan of tile_counts -> offsets[C*T+1] i32 (dense) and
 *                      cursors[C*T*GSB_CSTRIDE] i32 (strided like tile_counts)
 *   gsb_isect_scatter: (depth bits << 32 | flatten_id) into each tile's segment of
 *                      depth_id[capacity]; sets *overflow when I > capacity
 *   gsb_segment_sort : sorts every segment (bucket sort in shared memory; segments
 *                      longer than 2048 keys go through scratch[capacity]) and writes
 *                      isect_ids[capacity] i64, flatten_ids[capacity] i32; scratch may
 *                      alias isect_ids */
int gsb_tile_scan(const int32_t* tile_counts, int C, int tile_w, int tile_h, int32_t* offsets,
                  int32_t* cursors, int64_t* n_isects, void* stream);
int gsb_isect_scatter(const float* means2d, const int32_t* radii, const float* rec, int C, int N,
                      int tile_w, int tile_h, int32_t* cursors, int64_t capacity,
                      int64_t* depth_id, int32_t* overflow, void* stream);
int gsb_segment_sort(int64_t* depth_id, int64_t* scratch, const int32_t* offsets, int C,
                     int tile_w, int tile_h, int tile_n_bits, int64_t capacity,
                     int64_t* isect_ids, int32_t* flatten_ids, void* stream);

/* K5: isect_offset_encode (reference gslam/rasterization.py:274). */
int gsb_isect_offsets(const int64_t* isect_ids, int64_t n_isects, int C, int tile_w,
                      int tile_h, int tile_n_bits, int32_t* offsets, void* stream);

/* K6: per-tile front-to-back compositing (reference gslam/rasterization.py:325-339).
 * offsets holds C*T entries and n_isects closes the last tile; pass n_isects < 0 when
 * offsets has C*T+1 entries (gsb_tile_scan), so that no host-side count is needed:
 * -n_isects - 1 is then the number of entries flatten_ids can hold (0 = trust offsets), and
 * tile ranges are clipped to it, so a step with more intersections than the buffers were
 * sized for (overflow flag of gsb_isect_scatter) stays in bounds.
 * backgrounds[C,5] may be NULL.  n_touched[C,N] i64 (the reference's dtype,
 * gslam/rasterization.py:352) must be zeroed by the caller
 * (NULL skips the count).  Outputs render[C,H,W,5], alphas[C,H,W], last_ids[C,H,W].
 * pair_count (NULL in the product path) accumulates P, the number of (pixel,
 * Gaussian) pairs evaluated before each pixel stopped: the work unit of K6/K7
 * (SURVEY.md 8d); the caller zeroes it. */
int gsb_composite_fwd(const float* rec, const int32_t* flatten_ids, const int32_t* offsets,
                      const float* backgrounds, int C, int N, int width, int height,
                      int tile_w, int tile_h, int64_t n_isects, float visibility_min_T,
                      float* render, float* alphas, int32_t* last_ids, int64_t* n_touched,
                      int64_t* pair_count, void* stream);

/* K7: compositing backward.  v_rec[C,N,12] must be zeroed by the caller and is
 * accumulated atomically (x,y slots hold v_means2d).  v_backgrounds[C,5] may be NULL.
 * Deterministic mode (det_workspace != NULL, 16-byte aligned, gsb_composite_bwd_det_bytes(capacity)
 * bytes, capacity = n_isects or, for n_isects < 0, -n_isects - 1 > 0): every (tile unit, list entry)
 * gradient row is stored instead of added atomically, and a second kernel sums each Gaussian's rows
 * in a fixed order (tiles row-major, units in order), finding the entry in each tile's sorted list
 * by its depth bits in isect_ids[capacity] (the low word) and its flatten id.  v_rec is then
 * overwritten (no zeroing needed) and bit-identical from run to run; radii[C,N] and isect_ids are
 * required, v_backgrounds must be NULL (it is sum_pixels (1 - alpha) v_render, a plain reduction
 * the caller can do in any deterministic way). */
size_t gsb_composite_bwd_det_bytes(int64_t capacity);
int gsb_composite_bwd(const float* rec, const int32_t* flatten_ids, const int32_t* offsets,
                      const float* backgrounds, int C, int N, int width, int height,
                      int tile_w, int tile_h, int64_t n_isects, const float* render_alphas,
                      const int32_t* last_ids, const float* v_render, const float* v_alphas,
                      float* v_rec, float* v_backgrounds, const int32_t* radii,
                      const int64_t* isect_ids, void* det_workspace, size_t det_workspace_bytes,
                      void* stream);

/* K9: fused_ssim (reference gslam/backend.py:303-307; rahul-goel/fused-ssim@30fb258).
 * img1/img2 are [N,C,H,W] addressed through element strides (host arrays of 4: n, c, h, w),
 * so channels-last views need no copy.  valid != 0 is padding='valid' (the 5-pixel border is
 * left out of the mean).  ssim_sum (zeroed by the caller) receives the SUM of the SSIM map over
 * the counted pixels.  dm_* [N,C,H,W] (all three or none) receive the partial derivatives the
 * backward pass needs.  gsb_ssim_bwd writes (accumulate == 0) or adds (!= 0)
 *   (*grad) * coef * d(ssim_sum)/d(img1)   into v_img1 (element strides strides_out_host);
 * grad is a DEVICE scalar (NULL = 1) so the upstream gradient needs no host sync. */
/* Deterministic mode: det_partials (NULL = one float atomic per CTA) is scratch of
 * gsb_ssim_fwd_det_floats(N, C, H, W) floats; the CTA sums are added in a fixed order and ssim_sum is
 * overwritten. */
size_t gsb_ssim_fwd_det_floats(int N, int C, int H, int W);
int gsb_ssim_fwd(const float* img1, const float* img2, const int64_t* strides1_host,
                 const int64_t* strides2_host, int N, int C, int H, int W, int valid,
                 float* ssim_sum, float* dm_dmu1, float* dm_ds1, float* dm_ds12,
                 float* det_partials, void* stream);
int gsb_ssim_bwd(const float* img1, const float* img2, const int64_t* strides1_host,
                 const int64_t* strides2_host, int N, int C, int H, int W, const float* dm_dmu1,
                 const float* dm_ds1, const float* dm_ds12, const float* grad, float coef,
                 float* v_img1, const int64_t* strides_out_host, int accumulate, void* stream);

/* K10: photometric loss, value AND gradient in one pass (reference gslam/backend.py:273-283,
 * gslam/frontend.py:121-129).  render[C,H,W,5] (r g b depth beta), gt[C,H,W,3],
 * exposure[C,2] = (a, b) or NULL; e = rgb * exp(a) + b - gt.
 *   mode 0: mean(sum_ch e^2 / (2 beta^2)) + mean(0.5 ln^2 beta)      (mapping, active-NeRF)
 *   mode 1: mean(sum_ch e^2 beta^-2)   mode 2: mean |e|   mode 3: mean e^2   (tracking)
 * loss_sum[3] (zeroed by the caller) receives the two un-normalised sums (+ a block ticket);
 * loss_out[2] (may be NULL) receives (coef * loss, loss), written by the last block; v_render[C,H,W,5]
 * is overwritten with coef * dLoss/drender; v_exposure[C,2] (zeroed; may be NULL) accumulates
 * coef * dLoss/dexposure.
 * Deterministic mode: det_partials (NULL = float atomics) is scratch of
 * gsb_photo_loss_det_floats(C, H, W) floats; the CTA sums are added in a fixed order by a second
 * launch that overwrites loss_sum[0..1], loss_out and v_exposure. */
size_t gsb_photo_loss_det_floats(int C, int H, int W);
int gsb_photo_loss(const float* render, const float* gt, const float* exposure, int mode, int C,
                   int H, int W, float coef, float* loss_sum, float* loss_out, float* v_render,
                   float* v_exposure, float* det_partials, void* stream);
int gsb_scale_by_scalar(float* x, int64_t n, const float* grad, void* stream);

/* K12/K13: ONE launch of Adam over up to 8 tensors (reference gslam/backend.py:559-602 steps
 * seven optimizers) with torch.optim.Adam default semantics; `step` is the 1-based count of this
 * update.  Tensor `decay_tensor` (-1: none) additionally gets the opacity decay of
 * gslam/backend.py:356-359: element i is multiplied by `decay` when vis_count[i / vis_div] > 1.
 * step_dev (NULL, or int32[2] = {updates done so far, 0}) keeps the step count on the device instead:
 * the launch uses step_dev[0] + 1 and increments it, so a captured CUDA graph can be replayed. */
typedef struct {
    float* param;
    const float* grad;
    float* exp_avg;
    float* exp_avg_sq;
    int64_t n;
    float lr;
} gsb_adam_tensor;
int gsb_adam_multi(const gsb_adam_tensor* tensors_host, int n_tensors, float beta1, float beta2,
                   float eps, int step, int decay_tensor, const int32_t* vis_count, int vis_div,
                   float decay, int32_t* step_dev, void* stream);

/* ---- per-Gaussian pieces of the mapping step between the big kernels ----------------------
 * gsb_vis_count: vis_count[n] = #views with radii[c,n] > 0 (gslam/backend.py:287, :357).
 * gsb_isotropic_reg: the isotropic regulariser of gslam/backend.py:288-296 on the Gaussians with
 * vis_count > 0: adds weight * d/ds sum|exp(s) - exp(mean s).detach()| to v_log_scales[N,3] and the
 * unweighted sum to *loss_sum (float, may be NULL, accumulated atomically; with det_partials, scratch
 * of ceil(N / 256) floats, the CTA sums are added in a fixed order and *loss_sum is overwritten). */
int gsb_vis_count(const int32_t* radii, int C, int64_t N, int32_t* vis_count, void* stream);
int gsb_isotropic_reg(const float* log_scales, const int32_t* vis_count, int64_t N, float weight,
                      float* v_log_scales, float* loss_sum, float* det_partials, void* stream);

/* ---- the per-Gaussian tail of a mapping iteration, fused and range-limited ----------------------
 * One launch over the Gaussians [n0, n1): visibility = grad_rows[n, vis_col] (the reduced view count
 * written by gsb_project_bwd's v_vis) -> vis_count_out[n] (int32, may be NULL); on tensor
 * `scales_tensor` (width 3, -1: none) the isotropic-regulariser gradient of gslam/backend.py:286-296
 * is added where visibility > 0 and the unweighted sum of |exp(s) - exp(mean s)| is added to
 * *iso_loss_sum (may be NULL; det_partials: scratch of ceil((n1-n0)/256) floats for a fixed-order sum
 * that overwrites it); then torch.optim.Adam (defaults; `step` / `step_dev` as gsb_adam_multi) on
 * every listed tensor, gradient of element k of Gaussian n at grad_rows[n * row_stride + col + k];
 * tensor `decay_tensor` (-1: none) is multiplied by `decay` where visibility > 1
 * (gslam/backend.py:356-359).  A view-sharded step calls it on its own range after the
 * reduce-scatter of the rows; one GPU calls it on [0, N). */
typedef struct {
    float* param;        /* [N, width] */
    float* exp_avg;
    float* exp_avg_sq;
    int width;
    int col;             /* first column of this tensor's gradient in a row */
    float lr;
} gsb_map_tensor;
int gsb_map_update(const gsb_map_tensor* tensors_host, int n_tensors, const float* grad_rows,
                   int64_t row_stride, int vis_col, int64_t n0, int64_t n1, int scales_tensor,
                   float iso_weight, int decay_tensor, float decay, float beta1, float beta2, float eps,
                   int step, int32_t* step_dev, int32_t* vis_count_out, float* iso_loss_sum,
                   float* det_partials, void* stream);

/* gsb_isotropic_reg in deterministic mode WITHOUT caller scratch and without any allocation.  It has no
 * counterpart in the reference and was added for callers that cannot
 * call cudaMalloc (CUDA-graph capture, tight serving loops, harnesses that forbid it).  The per-block partial
 * sums live in a file-scope `static __device__` array of the library (256 KiB, held by every GPU that loads the
 * library until exit), so N <= gsb_isotropic_reg_static_max_n() = 16 777 216; larger N returns GSB_EINVAL
 * with a message.  *loss_sum (required) is OVERWRITTEN with the fixed-order sum, v_log_scales is added to
 * as in gsb_isotropic_reg; results are bit-identical to gsb_isotropic_reg with det_partials.  All calls on
 * a device share the array (it keeps the last call's partials; each call overwrites the range it reads, so
 * nothing needs clearing): calls that overlap on one GPU -- different streams with no ordering -- RACE. */
int64_t gsb_isotropic_reg_static_max_n(void);
int gsb_isotropic_reg_static(const float* log_scales, const int32_t* vis_count, int64_t N, float weight,
                             float* v_log_scales, float* loss_sum, void* stream);

/* The same update as ONE kernel that is also the reduce-scatter before it and the all-gather after it, over
 * NVLink multicast (NVLS; symmetric allocations mapped at a multicast address, e.g. torch symmetric memory):
 * grad_rows_mc is the MULTICAST address of the gradient rows -- every load returns the sum over the GPUs,
 * added inside the NVSwitch (multimem.ld_reduce) -- and the updated parameters of [n0, n1), the visibility
 * counts (vis_count_mc, may be NULL) and this rank's loss scalars (stats_in[2] -> row `rank` of
 * stats_mc[world][4] = total share, photometric share, isotropic sum, 0; may be NULL) are stored through the
 * multicast addresses param_mc_host[k] / vis_count_mc / stats_mc into EVERY GPU's copy (multimem.st).
 * tensors_host[k].param / exp_avg / exp_avg_sq are this GPU's own arrays (read for the update).  The caller
 * puts a cross-GPU barrier before the launch (all rows written) and after it (all stores visible).  scratch:
 * gsb_map_update_mc_scratch_floats(n0, n1) floats of this GPU, zeroed once.  The narrow tensors (width <= 4)
 * must lie in the first 16 columns of a row (ops.RowLayout does). */
size_t gsb_map_update_mc_scratch_floats(int64_t n0, int64_t n1);
int gsb_map_update_mc(const gsb_map_tensor* tensors_host, const uint64_t* param_mc_host, int n_tensors,
                      const float* grad_rows_mc, int64_t row_stride, int vis_col, int64_t n0, int64_t n1,
                      int scales_tensor, float iso_weight, int decay_tensor, float decay, float beta1,
                      float beta2, float eps, int step, int32_t* step_dev, int32_t* vis_count_mc,
                      const float* stats_in, float* stats_mc, int rank, float* scratch, void* stream);

/* ---- densify / prune surgery ------------------------------------------------
 * Replaces the per-tensor `value[keep_mask]` / `torch.cat` loops of gslam/pruning.py:10-53 and
 * gslam/insertion.py:27-62 over the 6 parameters and their 12 Adam states: every listed tensor
 * [n_old, width] is compacted to the rows kept by the mask (keep_cum = inclusive int64 scan of the
 * mask, NULL = keep all; n_keep = its last element) and n_add rows are appended from `add`
 * (NULL = zeros, i.e. fresh optimizer state).  dst has n_keep + n_add rows and must not alias src. */
typedef struct {
    const float* src;
    float* dst;
    const float* add;
    int width;
} gsb_row_tensor;
int gsb_rows_resize(const gsb_row_tensor* tensors_host, int n_tensors, int64_t n_old,
                    const int64_t* keep_cum, int64_t n_keep, int64_t n_add, void* stream);

/* ---- K11: camera pose chain -------------------------------------------------------------
 * Replaces PoseZhou.forward + rotation_6d_to_matrix, reference gslam/primitives.py:15-36, 82-92
 * (batched over the views that gslam/map.py:57,86 stacks with create_batch):
 *   viewmats[c] = Rt[c] . [[GS(dR[c] + (1,0,0,0,1,0)), dt[c]], [0 0 0 1]]
 * Rt [C,4,4], dR [C,6], dt [C,3], viewmats [C,4,4], row-major f32.  gsb_pose_bwd is its gradient:
 * v_viewmats [C,4,4] -> v_dR [C,6], v_dt [C,3] (either output may be NULL). */
int gsb_pose_fwd(const float* Rt, const float* dR, const float* dt, int C, float* viewmats, void* stream);
int gsb_pose_bwd(const float* Rt, const float* dR, const float* v_viewmats, int C, float* v_dR,
                 float* v_dt, void* stream);

/* ---- L-BFGS with a strong-Wolfe search as a device-side state machine ------------
 * The optimizer of gslam/frontend.py:604-662 / gslam/backend.py:436-506 (torch.optim.LBFGS, strong_wolfe)
 * with its whole state in `state` (gsb_lbfgs_state_bytes() bytes of device memory).  gsb_lbfgs_init starts
 * a minimisation from theta[n] (n <= 96, history_size <= 16; max_eval = torch's default max_iter * 5 / 4).
 * Then repeat max_eval times, without any host sync: evaluate the objective at theta -> result[1 + n] =
 * (loss, gradient); gsb_lbfgs_advance consumes it and writes the next point to evaluate into theta.  Once
 * the search has finished theta holds the final iterate and further calls change nothing.  summary[5]
 * (may be NULL) = (finished, evaluations used, iterations, first loss, loss of the last evaluation). */
size_t gsb_lbfgs_state_bytes(void);
int gsb_lbfgs_init(void* state, const float* theta, int n, int history_size, int max_iter, int max_eval,
                   double lr, double tolerance_grad, double tolerance_change, void* stream);
int gsb_lbfgs_advance(void* state, float* theta, const float* result, float* summary, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GSB_H */
