/*
 * bop_b200.h -- C ABI of the B200-native BOP pose-error hot path.
 *
 * Every entry point replaces one Python-level call of the reference
 * (gist-ailab/bop_toolkit); the reference file:line each one stands in for is
 * cited beside it.  All paths are relative to the reference repository root.
 *
 * Conventions
 *   - Plain C types only: pointers, sizes, POD structs.  No torch/pybind types.
 *   - `*_dev` functions take DEVICE pointers and enqueue work on `stream`
 *     (a cudaStream_t passed as void*; NULL = legacy default stream).  They never
 *     synchronise the device.
 *   - `*_host` functions take HOST pointers, stage them through the device on an
 *     internal stream, run the same kernels and return after the results are back
 *     in host memory.
 *   - Return value: 0 on success, non-zero on failure; bop_last_error() then
 *     returns a human-readable message (thread-local).
 *   - Poses are `double[12]`: R row-major (9 values) followed by t (3 values, mm),
 *     exactly the (R, t) pairs the reference passes to pose_error.* .
 *   - Units: millimetres and pixels, OpenCV camera frame (docs/bop_datasets_format.md).
 *   - There is NO CPU fallback: every compute entry point fails if no CUDA device
 *     is usable.
 */
#ifndef BOP_B200_H_
#define BOP_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BOP_B200_ABI_VERSION 6
#define BOP_MAX_TAUS 16      /* taus per launch; the host layer chunks longer lists */
#define BOP_MAX_SYMS_PER_CTA 16

/* Cost types of pose_error.vsd (bop_toolkit_lib/pose_error.py:79-88). */
#define BOP_COST_STEP 0
#define BOP_COST_TLINEAR 1
/* Visibility modes of visibility._estimate_visib_mask (bop_toolkit_lib/visibility.py:29-37). */
#define BOP_VISIB_BOP19 0
#define BOP_VISIB_BOP18 1
/* Depth image element types accepted by the VSD entry points. */
#define BOP_DEPTH_F32 0      /* float32 millimetres (what inout.load_depth returns, inout.py:39-46) */
#define BOP_DEPTH_U16 1      /* raw uint16 PNG samples; multiplied by depth_scale[im] in float32
                                (scripts/eval_calc_errors.py:239-243) */

/*
 * One object model resident in device memory.
 * Mirrors what the reference keeps per obj_id: models[obj_id]['pts'] (inout.load_ply,
 * inout.py:401-583), the renderer's vertex/index buffers (renderer_vispy.py:299-420),
 * models_sym[obj_id] (misc.get_symmetry_transformations, misc.py:43-90) and
 * models_info[obj_id]['diameter'].
 */
typedef struct bop_model_dev {
  const void*    pts;       /* [n_pts][3] float64 if pts_f64 else float32                     */
  const float*   verts4;    /* [n_pts][4] float32 (x,y,z,0): rasteriser vertices, ADI targets */
  const int32_t* faces4;    /* [n_faces][4] int32 (i0,i1,i2,0); may be NULL if n_faces == 0   */
  const double*  syms;      /* [n_syms][12] R row-major then t; may be NULL if n_syms == 0    */
  double         diameter;  /* models_info[obj_id]['diameter'] (mm)                          */
  int32_t        n_pts;
  int32_t        n_faces;
  int32_t        n_syms;
  int32_t        pts_f64;
  int32_t        cull;      /* 0: draw every triangle (the reference's GL state: no face culling,
                               renderer_vispy.py:470).  +1 / -1: skip triangles whose projected
                               signed area is negative / positive.  Only set for meshes verified to
                               be closed and consistently oriented, where the nearest surface along
                               every ray is front-facing and the depth image is unchanged.        */
  float          cull_min_dist; /* culling is applied to a pose only if |t| > cull_min_dist, i.e. the
                               camera centre lies outside the model's bounding sphere around its
                               origin (inside a mesh the visible faces ARE the back faces); set it
                               to max |pts| or more.  0 with cull != 0 culls every pose.  The
                               same radius tells bop_mspd_dev whether a pose keeps every vertex
                               well in front of the camera plane (float32 path) or not (float64). */
} bop_model_dev;

/* Parameters of pose_error.vsd (bop_toolkit_lib/pose_error.py:17-18). */
typedef struct bop_vsd_params {
  double  taus[BOP_MAX_TAUS];
  double  delta;
  int32_t n_taus;
  int32_t normalized_by_diameter;
  int32_t cost_type;        /* BOP_COST_*  */
  int32_t visib_mode;       /* BOP_VISIB_* */
  int32_t height;
  int32_t width;
  int32_t gate;             /* 1: apply the driver's sphere-projection gate
                               (scripts/eval_calc_errors.py:296-313) -> [1.0]*n_taus   */
  int32_t depth_dtype;      /* BOP_DEPTH_* */
} bop_vsd_params;

/* ---- library / device ---------------------------------------------------- */
int         bop_abi_version(void);
const char* bop_last_error(void);
/* Number of usable CUDA devices (0 if none); never fails. */
int         bop_device_count(void);
/* Number of kernels launched by this library in the calling process so far. */
uint64_t    bop_kernel_launch_count(void);

/* Per-kernel device timing (CUDA events on the launching stream), used by bench.py for the
 * roofline figure.  Slots: 0 fused VSD kernel, 1 VSD bbox+plan, 2 MSSD, 3 MSPD, 4 K2 pixel pass,
 * 5 render, 6 ADI, 7 mask overlap, 8 depth-PNG decode (inflate + unfilter).  bop_profile_sum_ms synchronises on the recorded events and returns the summed
 * duration of the last `last_n` timed launches of the slot (0 = all kept, ring of 256).        */
int bop_profile_enable(int on);
int bop_profile_reset(void);
int bop_profile_sum_ms(int slot, int last_n, double* sum_ms, int* count);

/* ---- K5: re / te ----------------------------------------------------------
 * pose_error.re (pose_error.py:187-202), pose_error.te (pose_error.py:205-214).
 * re_out / te_out: [n_pairs] float64; either may be NULL.                      */
int bop_re_te_dev(const double* pose_est, const double* pose_gt, int n_pairs,
                  double* re_out, double* te_out, void* stream);
int bop_re_te_host(const double* pose_est, const double* pose_gt, int n_pairs,
                   double* re_out, double* te_out);

/* ---- K3: per-vertex errors -------------------------------------------------
 * pose_error.mssd (pose_error.py:96-118), mspd (:121-144), add (:147-161),
 * proj (:217-232).  models: device table; obj_idx: [n_pairs] row of `models` per
 * pair.  gate != 0 applies `||t_e - t_g|| < diameter` else +inf
 * (scripts/eval_calc_errors.py:305-309,321-322,334-339); mspd/proj are never gated.
 * K (mspd/proj): [n_pairs][9] row-major 3x3 intrinsics per pair.
 * out: [n_pairs] float64.                                                      */
int bop_mssd_dev(const bop_model_dev* models, const int32_t* obj_idx,
                 const double* pose_est, const double* pose_gt, int n_pairs,
                 int max_syms, int gate, double* out, void* stream);
int bop_mspd_dev(const bop_model_dev* models, const int32_t* obj_idx,
                 const double* pose_est, const double* pose_gt, const double* K,
                 int n_pairs, int max_syms, double* out, void* stream);
int bop_add_dev(const bop_model_dev* models, const int32_t* obj_idx,
                const double* pose_est, const double* pose_gt, int n_pairs,
                int max_pts, int gate, double* out, void* stream);
int bop_proj_dev(const bop_model_dev* models, const int32_t* obj_idx,
                 const double* pose_est, const double* pose_gt, const double* K,
                 int n_pairs, int max_pts, double* out, void* stream);

/* ---- K4: ADI ----------------------------------------------------------------
 * pose_error.adi (pose_error.py:164-184): mean over GT-posed points of the
 * distance to the nearest EST-posed point; exact brute-force NN replaces
 * scipy.spatial.cKDTree.  out: [n_pairs] float64.                              */
int bop_adi_dev(const bop_model_dev* models, const int32_t* obj_idx,
                const double* pose_est, const double* pose_gt, int n_pairs,
                int max_pts, int gate, double* out, void* stream);

/* ---- K1: depth rasteriser ---------------------------------------------------
 * Renderer.render_object in 'depth' mode (renderer.py:62-74;
 * renderer_vispy.py:433-466,509-538; shaders :123-146).  One image per pose:
 * depth_out[b] is H x W float32, nearest-surface camera Z in mm at the pixel
 * centre (c+0.5, r+0.5), 0 = background.  cam: [n][4] float64 fx,fy,cx,cy.
 * max_geom: max over the models used of max(n_pts, (n_faces + 1) / 2); sizes the geometry scratch
 * (the assembled-triangle list: 2 * max_geom records of 36 bytes per pose).  A smaller value
 * is a contract violation and traps the kernel.
 * workspace: bop_render_workspace_bytes(n, H, W, max_geom) bytes of device memory, 256-byte
 * aligned.  Batches whose scratch would exceed 8 GiB are processed in several passes.  A
 * workspace holds the call's triangle lists and work counters until the call has finished on
 * the GPU: calls that share one must be stream-ordered (this holds for every workspace argument
 * in this header).   */
size_t bop_render_workspace_bytes(int n, int height, int width, int max_geom);
int bop_render_depth_dev(const bop_model_dev* models, const int32_t* obj_idx,
                         const double* pose, const double* cam, int n, int max_geom,
                         int height, int width, float* depth_out,
                         void* workspace, size_t workspace_bytes, void* stream);

/* ---- K2: fused depth->dist + visibility + per-tau cost ---------------------
 * pose_error.vsd after the two render calls (pose_error.py:48-91), i.e.
 * misc.depth_im_to_dist_im_fast (misc.py:143-163) x3,
 * visibility.estimate_visib_mask_gt/_est (visibility.py:45-75) and the tau sweep,
 * for pre-rendered depth images.  depth_est/depth_gt: [n_pairs][H][W] float32.
 * depth_test: [n_im][H][W] (prm->depth_dtype); im_idx: [n_pairs]; K: [n_im][9];
 * depth_scale: [n_im] float32 or NULL; diameters: [n_pairs] float64.
 * errors_out: [n_pairs][n_taus] float64.  counts_out (optional, may be NULL):
 * [n_pairs][2+n_taus] int64 = union, inter, per-tau cost counts (step cost).
 * workspace: bop_vsd_workspace_bytes(n_pairs, H, W, max_geom) bytes (max_geom = 1 suffices for
 * bop_vsd_from_depth_dev, which projects nothing).                              */
size_t bop_vsd_workspace_bytes(int n_pairs, int height, int width, int max_geom);
int bop_vsd_from_depth_dev(const void* depth_test, const float* depth_scale,
                           const int32_t* im_idx, const double* K,
                           const float* depth_est, const float* depth_gt,
                           const double* diameters, int n_pairs,
                           const bop_vsd_params* prm, double* errors_out,
                           int64_t* counts_out, void* workspace,
                           size_t workspace_bytes, void* stream);

/* bop_vsd_from_depth_dev without a workspace argument and without any allocation.  It has no
 * counterpart in the reference; it was added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same inputs, outputs and results as bop_vsd_from_depth_dev; the scratch (counters,
 * per-pair counts, tlinear sums: 272 B per pair) lives in a static __device__ array inside the
 * library, sized for BOP_STATIC_MAX_PAIRS pairs (1.06 MiB, held by every GPU that loads the
 * library until the process exits).  n_pairs > BOP_STATIC_MAX_PAIRS is rejected with an error
 * (split the batch, or use the workspace variant).  The function calls no cudaMalloc / cudaFree /
 * allocating routine; its only set-up is one cudaGetSymbolAddress per device on the first call
 * made with that device current (make that call once outside a graph capture).  All calls on a
 * device share the one array: each call clears the part it uses on `stream` first, and calls
 * that OVERLAP on the GPU (different streams without an ordering between them) race on it. */
#define BOP_STATIC_MAX_PAIRS 4096
int bop_vsd_from_depth_static_dev(const void* depth_test, const float* depth_scale,
                                  const int32_t* im_idx, const double* K,
                                  const float* depth_est, const float* depth_gt,
                                  const double* diameters, int n_pairs,
                                  const bop_vsd_params* prm, double* errors_out,
                                  int64_t* counts_out, void* stream);

/* ---- K1+K2 fused: pose_error.vsd for a batch of pose pairs -----------------
 * pose_error.vsd (pose_error.py:17-93) with the library's own rasteriser: both
 * poses are rendered into shared-memory z-tiles over the union ROI and consumed
 * by the pixel pass without ever being written to HBM.
 * diameters: optional [n_pairs] float64 overriding models[obj].diameter (the `diameter`
 * argument of pose_error.vsd); NULL = use the model table.
 * errors_out: [n_pairs][n_taus] float64; counts_out optional as above.        */
int bop_vsd_dev(const bop_model_dev* models, const int32_t* obj_idx,
                const double* pose_est, const double* pose_gt,
                const void* depth_test, const float* depth_scale,
                const int32_t* im_idx, const double* K,
                const double* diameters, int n_pairs, int max_geom,
                const bop_vsd_params* prm, double* errors_out,
                int64_t* counts_out, void* workspace, size_t workspace_bytes,
                void* stream);
/* bop_vsd_dev with its two passes on two streams: the geometry pass (projection, planning, sorted
 * triangle lists) is launched on `geometry_stream`, the raster + pixel pass and the final
 * reduction on `raster_stream`; events order each raster pass behind its geometry pass (and, for
 * batches that take several workspace chunks, the next geometry pass behind the raster pass that
 * still reads the workspace).  For callers that partition the SMs (CUDA green contexts,
 * bop_toolkit_b200/sm_partition.py): with HOST-resident depth the pixel pass is bound by the PCIe
 * link, and on a partition of its own it overlaps the geometry pass of the next batch.  The
 * inputs the geometry pass reads must be ready on geometry_stream, depth_test / depth_scale on
 * raster_stream (or earlier on geometry_stream); the results are ready on raster_stream.  A
 * workspace may be reused by a later call once that call's geometry_stream is ordered behind this
 * call's raster_stream.  Same results as bop_vsd_dev, bit for bit.  (Reference: the same
 * pose_error.vsd, :17-93; the reference has no streams.)                                      */
int bop_vsd_split_dev(const bop_model_dev* models, const int32_t* obj_idx,
                      const double* pose_est, const double* pose_gt, const void* depth_test,
                      const float* depth_scale, const int32_t* im_idx, const double* K,
                      const double* diameters, int n_pairs, int max_geom,
                      const bop_vsd_params* prm, double* errors_out, int64_t* counts_out,
                      void* workspace, size_t workspace_bytes, void* geometry_stream,
                      void* raster_stream);

/* bop_vsd_dev for pairs that share their GT pose: scripts/eval_calc_errors.py:281-289 pairs every
 * estimate of an object with every GT pose of it, so a GT pose is typically evaluated against
 * several estimates.  group_off_host (HOST array, [n_groups + 1], group_off_host[0] = 0,
 * group_off_host[n_groups] = n_pairs) cuts the batch into groups of 1 .. 32 consecutive pairs
 * with the same obj_idx, im_idx, pose_gt and diameter (the caller sorts, batch.group_pairs_by_gt
 * does; a group that mixes them is a contract violation and traps the kernel).
 * The GT pose of a group is projected, sorted and rasterised once; its z-tile stays in shared
 * memory while the estimates of the group take turns.  Results are bit-identical to bop_vsd_dev
 * on the same pairs (tests/test_gpu_parity.py::test_vsd_grouped_equals_ungrouped). */
int bop_vsd_grouped_dev(const bop_model_dev* models, const int32_t* obj_idx,
                        const double* pose_est, const double* pose_gt, const void* depth_test,
                        const float* depth_scale, const int32_t* im_idx, const double* K,
                        const double* diameters, int n_pairs, const int32_t* group_off_host,
                        int n_groups, int max_geom, const bop_vsd_params* prm, double* errors_out,
                        int64_t* counts_out, void* workspace, size_t workspace_bytes, void* stream);

/* ---- mask overlap: pose_error.cus / cou_bb_proj / cou_mask --------------------
 * pose_error.cus (pose_error.py:256-286): 1 - |mask_est & mask_gt| / |mask_est | mask_gt| of the
 * two rendered silhouettes (1.0 when the union is empty); pose_error.cou_bb_proj (:300-330):
 * 1 - IoU of the pixel bounding boxes of the two silhouettes (misc.calc_2d_bbox misc.py:202-218,
 * misc.iou misc.py:236-263; NaN when a silhouette is empty -- the reference raises there).
 * cam: [n_pairs][4] fx,fy,cx,cy.  gate != 0 applies the driver's sphere-projection gate
 * (scripts/eval_calc_errors.py:296-300,349-354) -> cus = 1.0.
 * Outputs (each may be NULL): cus_out, cou_bb_out [n_pairs] float64; counts_out [n_pairs][4]
 * int64 = union, inter, est, gt pixel counts; bbox_out [n_pairs][8] int32 = (x, y, w, h) of the
 * est then of the gt silhouette (w = xmax - xmin; w = h = -1 when empty).
 * workspace: bop_vsd_workspace_bytes(n_pairs, H, W, max_geom) bytes.                      */
int bop_mask_overlap_dev(const bop_model_dev* models, const int32_t* obj_idx,
                         const double* pose_est, const double* pose_gt, const double* cam,
                         int n_pairs, int max_geom, int height, int width, int gate,
                         double* cus_out, double* cou_bb_out, int64_t* counts_out,
                         int32_t* bbox_out, void* workspace, size_t workspace_bytes,
                         void* stream);
/* Same statistics for pre-rendered depth images or binary masks ([n_pairs][H][W] float32,
 * non-zero = inside): pose_error.cou_mask (pose_error.py:235-253) and the tail of cus /
 * cou_bb_proj when a foreign Renderer produced the images.
 * workspace: bop_mask_stats_workspace_bytes(n_pairs) bytes.                              */
size_t bop_mask_stats_workspace_bytes(int n_pairs);
int bop_mask_overlap_from_depth_dev(const float* depth_est, const float* depth_gt, int n_pairs,
                                    int height, int width, double* cus_out, double* cou_bb_out,
                                    int64_t* counts_out, int32_t* bbox_out, void* workspace,
                                    size_t workspace_bytes, void* stream);

/* ---- GT info / GT masks -----------------------------------------------------------
 * The per-instance body of scripts/calc_gt_info.py:113-175 and scripts/calc_gt_masks.py:110-133:
 * render the GT pose, visibility.estimate_visib_mask_gt (visibility.py:45-57, 'bop19') against
 * the sensor depth, pixel counts and boxes.  margin_x/margin_y: calc_gt_info renders on a 3x
 * canvas (margins = image width/height, calc_gt_info.py:76-79) so px_count_all / bbox_obj cover
 * the part of the object outside the image; calc_gt_masks uses margins 0.
 * depth_test [n_im][H][W] (depth_dtype), depth_scale [n_im] or NULL, im_idx [n], K [n_im][9].
 * counts_out [n][3] int64 = px_count_all, px_count_valid, px_count_visib; visib_fract_out [n];
 * bbox_out [n][8] int32 = bbox_obj (x,y,w,h) then bbox_visib, all -1 unless px_count_visib > 0;
 * mask_out / mask_visib_out (optional) [n][H][W] uint8, 255 inside (the PNG payloads).
 * workspace: bop_gt_info_workspace_bytes(...) bytes, 256-byte aligned.                     */
size_t bop_gt_info_workspace_bytes(int n, int height, int width, int margin_x, int margin_y,
                                   int max_geom);
int bop_gt_info_dev(const bop_model_dev* models, const int32_t* obj_idx, const double* pose_gt,
                    const void* depth_test, const float* depth_scale, const int32_t* im_idx,
                    const double* K, int n, int max_geom, int height, int width, int depth_dtype,
                    double delta, int margin_x, int margin_y, int64_t* counts_out,
                    double* visib_fract_out, int32_t* bbox_out, unsigned char* mask_out,
                    unsigned char* mask_visib_out, void* workspace, size_t workspace_bytes,
                    void* stream);

/* depth_test may also be a pointer into PINNED HOST memory (cudaHostAlloc / cudaHostRegister;
 * with unified addressing the host pointer is valid on the device): the fused kernel then pulls
 * only the depth samples it needs (the pixels under the two rendered silhouettes) across PCIe
 * instead of the caller copying whole images.  bop_vsd_depth_bytes_fetched reports how many
 * bytes of depth_test the most recent pass on `workspace` fetched (synchronises the device). */
int bop_vsd_depth_bytes_fetched(const void* workspace, int depth_dtype, uint64_t* bytes);
/* Test hook: the assembled-triangle list (geometry pass) that the most recent bop_vsd_dev call
 * with these sizes left in `workspace` for (pair, pose; 0 = est, 1 = gt): xy[k][6] = snapped
 * window positions (1/256 px) of the three vertices of entry k, at most `cap` entries; *count =
 * length of the list.  The list is sorted by the size of the candidate-pixel rectangle (rows, then
 * columns, largest first) so that the 32 lanes of a raster warp run the same loop trip counts. */
int bop_vsd_debug_triangle_list(const void* workspace, int n_pairs, int height, int width,
                                int max_geom, int pair, int pose, int32_t* xy, int cap, int* count);

/* ---- host-buffer convenience layer -----------------------------------------
 * Upload a model (host arrays, float64 pts as pose_error.* receives them) and get a
 * filled descriptor whose pointers are device memory owned by the library until
 * bop_model_free.  `desc_dev_out` receives a device copy of the descriptor that can
 * be used as a 1-row `models` table.                                           */
int bop_model_upload(const double* pts, int n_pts, const int32_t* faces, int n_faces,
                     const double* syms, int n_syms, double diameter,
                     bop_model_dev* desc_host_out, bop_model_dev** desc_dev_out);
int bop_model_free(bop_model_dev* desc_host, bop_model_dev* desc_dev);

/* Single-model batches with host buffers (what a ctypes binding of
 * pose_error.mssd/mspd/add/adi would call). */
int bop_mssd_host(const bop_model_dev* model_dev, int n_syms, const double* pose_est,
                  const double* pose_gt, int n_pairs, int gate, double* out);
int bop_mspd_host(const bop_model_dev* model_dev, int n_syms, const double* pose_est,
                  const double* pose_gt, const double* K, int n_pairs, double* out);
int bop_add_host(const bop_model_dev* model_dev, int n_pts, const double* pose_est,
                 const double* pose_gt, int n_pairs, int gate, double* out);
int bop_adi_host(const bop_model_dev* model_dev, int n_pts, const double* pose_est,
                 const double* pose_gt, int n_pairs, int gate, double* out);
int bop_render_depth_host(const bop_model_dev* model_dev, const double* pose,
                          const double* cam, int n, int height, int width,
                          float* depth_out);
int bop_vsd_host(const bop_model_dev* model_dev, const double* pose_est,
                 const double* pose_gt, const void* depth_test,
                 const float* depth_scale, const int32_t* im_idx, const double* K,
                 int n_im, int n_pairs, const bop_vsd_params* prm, double* errors_out);

/* Tunables of the fused VSD kernel (persist per process; used for on-GPU tuning).
 * tile_px: z-tile capacity in pixels per pose (shared memory = 8*tile_px bytes);
 * threads: CTA size; ctas_per_sm: persistent CTAs per SM.  0 keeps the current value. */
int bop_vsd_set_tuning(int tile_px, int threads, int ctas_per_sm, int small_tri_px);
/* Test hook: exact == 1 forces the float64 arithmetic of the reference for every pixel of the
 * step-cost pass; by default a float32 pre-test decides the comparisons that are not close to
 * their thresholds, and ascending taus are swept as a histogram (one bin per pixel instead of
 * one counter per tau); exact == 2 keeps the float32 pre-test but uses per-tau counters.
 * Bit-identical counts in all three modes; tests compare them. */
int bop_vsd_set_exact(int exact);
/* Test hook: exact != 0 makes bop_mssd_dev / bop_mspd_dev evaluate every vertex in float64 (the
 * reference's dtype).  By default models whose points are float32-representable are evaluated in
 * float32 AFTER the rigid parts were folded in float64 (dR p + dt; difference of projections as
 * (D.x n.w - n.x D.w) / (n.w (n.w + D.w))): no cancellation is left, results agree to ~1e-6
 * relative (bar of pose_error.mssd / mspd: 1e-5).  Tests compare the two. */
int bop_points_set_exact(int exact);


/* ---- depth-PNG input on the device -------------------------------------------------------
 * Replaces inout.load_depth (bop_toolkit_lib/inout.py:39-46: imageio.imread of a 16-bit
 * greyscale PNG) as called per image by scripts/eval_calc_errors.py:236-243.  The host only
 * reads the files and strips the PNG container; zlib inflate and the PNG row filters run on
 * the GPU, one warp per image, and leave uint16 samples in HBM in the layout the VSD entry
 * points take as BOP_DEPTH_U16.  The zlib Adler-32 is verified on the device; chunk CRCs are not.         */
#define BOP_PNG_IO_ERROR          (-101)  /* file missing / unreadable                         */
#define BOP_PNG_NOT_PNG           (-102)  /* bad signature or IHDR                             */
#define BOP_PNG_TRUNCATED         (-103)  /* chunk runs past the end of the file, no IDAT      */
#define BOP_PNG_UNSUPPORTED       (-104)  /* not 16-bit greyscale non-interlaced: decode it on
                                             the host (the caller's choice), fields are filled */
#define BOP_PNG_STAGING_TOO_SMALL (-105)  /* returned by the call; *needed_bytes tells the size */
typedef struct bop_png_info {
  int32_t width, height, bit_depth, color_type, interlace;
  int32_t status;            /* 0 or one of BOP_PNG_* for this file                             */
  int64_t idat_bytes;        /* length of the concatenated IDAT payloads (the zlib stream)      */
} bop_png_info;
/* Reads `n` files with `threads` host threads into `staging` (pinned memory recommended) and
 * compacts each file's IDAT payloads in place: afterwards staging + offsets[i] holds the zlib
 * stream of file i (comp_bytes[i] bytes, 16-byte aligned, zero padded).  Per-file problems are
 * reported in infos[i].status (comp_bytes[i] = 0).  If staging_bytes is too small nothing is
 * read, *needed_bytes is set and BOP_PNG_STAGING_TOO_SMALL is returned.                       */
int bop_png_stage_files_host(const char* const* paths, int n, void* staging, size_t staging_bytes,
                             int64_t* offsets, int32_t* comp_bytes, bop_png_info* infos,
                             int threads, size_t* needed_bytes);
/* Same container walk for one file already in memory (compacts in place in `file_bytes`). */
int bop_png_scan_host(void* file_bytes, size_t nbytes, bop_png_info* info, int64_t* idat_bytes);
size_t bop_png_decode_workspace_bytes(int n, int height, int width);
/* comp/offsets/comp_bytes: the staged streams in DEVICE memory (comp 16-byte aligned);
 * depth_out [n][height][width] uint16; status[n] int32: 0 or a negative decoder status
 * (bop_toolkit_b200/csrc/png_inflate.cuh) -- images with a non-zero status are not written. */
int bop_png_decode_dev(const void* comp, const int64_t* offsets, const int32_t* comp_bytes, int n,
                       int height, int width, uint16_t* depth_out, int32_t* status,
                       void* workspace, size_t workspace_bytes, void* stream);
/* The device decoder's source compiled for the host, for the CPU test-suite only (checks the
 * inflate core against zlib without a GPU).  No product path calls it.                       */
int bop_png_decode_selftest_host(const void* idat, int32_t comp_bytes, int height, int width,
                                 uint16_t* depth_out);


/* ---- error files -------------------------------------------------------------------------
 * The text inout.save_json (bop_toolkit_lib/inout.py:85-113) writes for the record list of
 * scripts/eval_calc_errors.py:380-418, from flat arrays: records r = 0..n_rec-1 with
 * (im_id, obj_id, est_id, score); pairs sorted by record, GT ids ascending within a record,
 * errs [n_pairs][n_cols].  col >= 0 writes the 1-element lists of one VSD tau file, col < 0
 * whole lists.  Floats are formatted like Python's float.__repr__ / json (Infinity, NaN).     */
int bop_write_error_file_host(const char* path, int64_t n_rec, const int64_t* im_id,
                              const int64_t* obj_id, const int64_t* est_id, const double* score,
                              int64_t n_pairs, const int64_t* pair_rec, const int64_t* pair_gt,
                              const double* errs, int n_cols, int col);
/* float.__repr__(v) as json.dumps writes it, NUL-terminated into out32 (>= 32 bytes). */
int bop_format_double_host(double v, char* out32);

/* The numbers of a BOP19 results CSV (inout.load_bop_results, bop_toolkit_lib/inout.py:221-261)
 * into flat arrays of `cap` rows: ids [n][3] = scene_id, im_id, obj_id; score [n]; R [n][9];
 * t [n][3]; time [n].  Returns 0 on success, 1 when a line is anything but plainly formatted
 * numbers (the caller then parses the file the reference's way and raises its errors).        */
int bop_parse_results_csv_host(const char* text, size_t nbytes, int64_t cap, int64_t* ids,
                               double* score, double* R, double* t, double* time, int64_t* n_out);

#ifdef __cplusplus
}
#endif
#endif  /* BOP_B200_H_ */
