/* gsw.h — C ABI of libgsw.so: B200-native (sm_100a) GSWorld observation path.
 *
 * One hot path: per-link / per-object rigid transform of the movable Gaussians of every environment,
 * followed by 3D-Gaussian-splatting forward rasterization of scene + robot for every (env, camera) frame.
 * Plain pointers and sizes only; no torch types.  All `const float*` / `int32_t*` arguments below are
 * DEVICE pointers unless the comment says "host".  The caller owns every input and output buffer; the
 * context owns only its grow-only workspace.  Nothing here synchronises the host except gsw_finish()
 * and calls that (re)allocate workspace because a shape grew.
 *
 * Every function returns 0 on success, a negative GSW_E_* code for an invalid argument or capacity
 * problem, or a positive cudaError_t value; gsw_last_error() gives the text (thread-local).
 *
 * What each entry point replaces in the reference (luccachiang/GSWorld @ 5329504):
 *   gsw_scene_set          <- the merged model held by GSWorldWrapper.initial_merger_robot
 *                             (gsworld/mani_skill/utils/wrappers/gs_world_wrapper.py:75-79) plus the
 *                             activations of gaussian_model.py:102-130 done once instead of per frame.
 *   gsw_render_frames      <- GSWorldWrapper.transform_gs_perlink (gs_world_wrapper.py:110-162,
 *                             Gaussian-level part = gs_utils.py:283-385) + _render_gsworld
 *                             (gs_world_wrapper.py:232-275): B*C deepcopies, isin write-backs and
 *                             serial render() calls become one batched launch set.
 *   gsw_rasterize_forward  <- diff_gaussian_rasterization._C.rasterize_gaussians as called by
 *                             GaussianRasterizer.forward (call site
 *                             submodules/gaussian-splatting/gaussian_renderer/__init__.py:36-52,102-110;
 *                             the extension source itself is an absent submodule, pinned at
 *                             graphdeco-inria/diff-gaussian-rasterization@9c5c2028, near plane 0.05 per
 *                             GSWorld README.md:33).
 *   gsw_mark_visible       <- GaussianRasterizer.markVisible (same absent extension).
 *   gsw_rig_set / gsw_compose_poses
 *                          <- the matrix algebra of transform_gs_perlink: link_trans = S @ M_link @
 *                             inv(M_scan) @ inv(S) (gs_world_wrapper.py:115-120), objects through
 *                             extract_rigid_transform (gs_world_wrapper.py:139-156, pcd_utils.py:224-252)
 *                             and matrix_to_quaternion (gs_utils.py:322).
 *   gsw_compose_cameras    <- GSWorldWrapper.cam_maniskill2gs (gs_world_wrapper.py:277-325) + Camera.__init__
 *                             (scene/cameras.py:80-89) + getWorld2View2/getProjectionMatrix
 *                             (utils/graphics_utils.py:38-71), without the .cpu() round trips.
 *   gsw_debug_*            <- no counterpart; parity taps (tiles_touched, sorted keys, tile ranges).
 */
#ifndef GSW_H_
#define GSW_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GSW_VERSION 100

/* error codes (negative); positive return values are cudaError_t */
#define GSW_OK 0
#define GSW_E_INVALID (-1)   /* bad argument */
#define GSW_E_STATE (-2)     /* call order (e.g. render before scene_set) */
#define GSW_E_CAPACITY (-3)  /* tile-entry workspace too small for the last launch; see gsw_finish */
#define GSW_E_NOMEM (-4)

/* floats per (env, slot) pose row: R[9] row-major, t[3], qR[4] wxyz, kappa, has_scale, pad[2] */
#define GSW_POSE_STRIDE 20
/* value of out_seg where no Gaussian was composited */
#define GSW_SEG_NONE (-2)

typedef struct gsw_ctx gsw_ctx;
typedef void* gsw_stream; /* cudaStream_t */

int gsw_version(void);
const char* gsw_last_error(void);

int gsw_create(int device, gsw_ctx** out);
int gsw_destroy(gsw_ctx* ctx);

/* Scene upload.  Arrays are used IN PLACE (zero copy) and must outlive the context or the next
 * gsw_scene_set.  Layouts are the reference's own:
 *   xyz[N,3]; scale_act[N,3] = exp(_scaling); quat_act[N,4] = normalize(_rotation) (wxyz);
 *   opacity_act[N] = sigmoid(_opacity); sh[N,16,3] = cat(_features_dc,_features_rest);
 *   log_scale_raw[N,3] = _scaling and quat_raw[N,4] = _rotation (read for movable Gaussians only);
 *   slot[N]: -1 = static background, else index into the per-env pose table (link or object).
 * A compact copy of the movable Gaussians (slot >= 0) is made here; this call synchronises. */
int gsw_scene_set(gsw_ctx* ctx, int64_t N, const float* xyz, const float* scale_act, const float* quat_act,
                  const float* opacity_act, const float* sh, const float* log_scale_raw, const float* quat_raw,
                  const int32_t* slot, int num_slots, gsw_stream stream);

/* Optional per-Gaussian integer labels (device int32[N], original order, e.g. the scene's `_semantics`), used
 * in place; needed only for gsw_frames_desc.out_seg.  Reset by gsw_scene_set. */
int gsw_scene_set_labels(gsw_ctx* ctx, const int32_t* labels);

typedef struct gsw_frames_desc {
  int32_t num_envs;   /* B */
  int32_t num_views;  /* V distinct cameras; static Gaussians are projected once per view */
  int32_t num_frames; /* F rendered images; frame f shows env frame_env[f] through view frame_view[f] */
  int32_t height, width;
  int32_t antialiasing;
  float scale_modifier;
  float near_plane; /* 0.05 = GSWorld (README.md:33); 0.2 = upstream */
  float bg[3];
  const float* pose;         /* [B, num_slots, GSW_POSE_STRIDE] */
  const float* viewmatrix;   /* [V,16] world_view_transform, transposed row-major (cameras.py:86) */
  const float* projmatrix;   /* [V,16] full_proj_transform, same convention (cameras.py:88) */
  const float* campos;       /* [V,3] */
  const float* tanfov;       /* [V,2] tan(FoVx/2), tan(FoVy/2) */
  const int32_t* frame_env;  /* [F] */
  const int32_t* frame_view; /* [F] */
  uint8_t* out_rgb_u8;       /* [F,H,W,3] or NULL: (clamp01(c)*255) truncated, gs_world_wrapper.py:268-270 */
  float* out_color;          /* [F,3,H,W] or NULL: un-clamped C + T*bg as the rasterizer returns it */
  float* out_invdepth;       /* [F,1,H,W] or NULL */
  int32_t* out_radii;        /* [F,N] or NULL */
  uint8_t* host_rgb_u8;      /* HOST pinned [F,H,W,3] or NULL: needs out_rgb_u8.  The frames are delivered device->host
                              * in a few groups of doubling size on the context's copy stream while later frames are still being
                              * composited: the blend stays one launch, the last CTA of a group publishes a flag and
                              * the copy stream waits on it with a stream memory operation (cuStreamWaitValue32).
                              * Where the driver refuses that operation, or with GSW_NO_WAIT_VALUE set in the
                              * environment at gsw_create, the frames are blended in up to 5 launches with an event per
                              * group instead.  `stream` waits for the last copy, so synchronising `stream` makes the
                              * host buffer valid. */
  int32_t* out_seg;          /* [F,H,W] or NULL: label (gsw_scene_set_labels) of the Gaussian with the largest blending
                              * weight alpha*T at each pixel, GSW_SEG_NONE where nothing was composited.  No counterpart
                              * in the reference (its segmentation comes from SAPIEN's Vulkan renderer, unavailable
                              * here): SURVEY.md §8(f) row 3. */
  /* Peer delivery (multi-GPU): HOST array of num_peer_rgb_u8 (<= 16) DEVICE pointers, each a [F,H,W,3] uint8 buffer
   * in this or another GPU's memory that is mapped into this process (peer access / symmetric memory), or NULL.
   * Needs out_rgb_u8.  Every finished frame group is pushed to each of them on the context's copy stream
   * (copy engines over NVLink / NVSwitch, no SMs), overlapped with the compositing of the later frames, exactly
   * like host_rgb_u8: this is the observation exchange of SURVEY.md §8(e) without a collective kernel.  `stream`
   * waits for the last push; cross-rank ordering (a barrier before the data is read on the other GPU) is the
   * caller's.  No counterpart in the reference (single GPU). */
  uint8_t* const* peer_rgb_u8;
  int32_t num_peer_rgb_u8;
  /* GSW_DELIVER_NO_JOIN: `stream` does NOT wait for the last host copy / peer push of this call, so the next step
   * can start while the delivery drains; the caller orders whoever consumes the delivered frames (and whoever
   * overwrites out_rgb_u8) with gsw_delivery_wait.  out_rgb_u8 must then stay untouched until that wait. */
  int32_t deliver_flags;
} gsw_frames_desc;

#define GSW_DELIVER_NO_JOIN 1
/* `stream` waits for the host copies of this call but NOT for its peer pushes: a host that reads its own frames after
 * every step then does not sit on the NVLink traffic and the cross-rank flow control behind it; whoever consumes the
 * pushed frames (and whoever overwrites out_rgb_u8) is ordered with gsw_delivery_wait as with GSW_DELIVER_NO_JOIN. */
#define GSW_DELIVER_NO_PEER_JOIN 2
/* Deliver AFTER the blend launch instead of group by group while it runs: no completion signals inside the blend
 * kernel (each of its CTAs otherwise ends with a release operation that waits for its stores: ~12 % of the kernel).
 * For fully asynchronous producers, where the delivery overlaps the NEXT step's kernels anyway; combine with
 * GSW_DELIVER_NO_JOIN.  A host that waits for every step wants the default (delivery overlapped with the blend). */
#define GSW_DELIVER_AFTER_BLEND 4

/* Enqueue transform + rasterization of all frames on `stream`.  Asynchronous.
 * Static-list cache: a leading run of views whose 37 camera words (viewmatrix, projmatrix, campos, tanfov) are bitwise
 * those of the previous gsw_render_frames call on this context keeps its static tile lists and static-only tiles
 * from that call (compared on the device; scene, sizes, switches, outputs, stream and workspace must be unchanged as
 * well, else everything is rebuilt).  Same bytes out either way; number world-fixed cameras first to profit from it.
 * GSW_NO_STATIC_CACHE=1 in the environment turns it off.  No counterpart in the reference, which rebuilds the
 * lists of every frame every step. */
int gsw_render_frames(gsw_ctx* ctx, const gsw_frames_desc* desc, gsw_stream stream);

/* Wait for the last launch on this context and report it.  Returns GSW_E_CAPACITY if the tile-entry
 * workspace overflowed (outputs are then invalid); *needed_entries (may be NULL) receives the entry
 * count that launch needed.  Call gsw_reserve_entries() and launch again.  stats (may be NULL):
 * [0] = D tile entries written (static per view + dynamic per frame), [1] = entries the blend staged
 * (<= sum of per-frame list lengths; early termination reads less), [2] = entry capacity, [3] = launches. */
int gsw_finish(gsw_ctx* ctx, int64_t* needed_entries, int64_t stats[4]);
/* Make `stream` wait for the delivery (host_rgb_u8 / peer_rgb_u8) of the last gsw_render_frames call; a no-op if
 * that call delivered nothing.  Needed only with GSW_DELIVER_NO_JOIN.  No counterpart in the reference. */
int gsw_delivery_wait(gsw_ctx* ctx, gsw_stream stream);
/* The reverse ordering: the peer pushes (peer_rgb_u8) of the NEXT gsw_render_frames call on this context start only
 * after everything enqueued on `stream` so far has completed (the compositing and the host copies are not held back).  A multi-GPU
 * exchange calls it with the stream that ran the previous step's cross-rank barrier, so that a peer's buffer is not
 * overwritten while that peer may still be reading it.  No counterpart in the reference (single GPU). */
int gsw_delivery_gate(gsw_ctx* ctx, gsw_stream stream);
int gsw_reserve_entries(gsw_ctx* ctx, int64_t entries);
/* depth sub-buckets per tile list (power of two, 1..64); 0 = automatic */
int gsw_set_depth_buckets(gsw_ctx* ctx, int nb);
/* Exact tile culling (default off = the reference's tile lists).  The reference duplicates a Gaussian into every
 * tile of the bounding square of its 3-sigma radius (rasterizer forward.cu getRect / duplicateWithKeys); its
 * blend loop then skips, without touching any state, every entry whose alpha stays below 1/255 on the pixel.
 * With culling on, a (Gaussian, tile) entry that provably stays below 1/255 on all 256 pixels of the tile is
 * not created: images, radii and the per-Gaussian tile count tap stay bit-identical, the per-tile lists
 * (gsw_debug_frame_lists, stats D) become order-preserving sub-lists of the reference's.  No counterpart in
 * the reference. */
int gsw_set_tile_culling(gsw_ctx* ctx, int on);

/* Single-frame forward with the upstream argument list (drop-in for GaussianRasterizer.forward).
 * Exactly one of shs / colors_precomp and exactly one of (scales, rotations) / cov3D_precomp.
 * shs is [P,M,3]; degree <= 3.  Outputs: out_color[3,H,W], out_radii[P], out_invdepth[1,H,W] (NULL ok).
 * bg, viewmatrix, projmatrix, campos are device pointers like upstream tensors. */
int gsw_rasterize_forward(gsw_ctx* ctx, int32_t P, int32_t sh_degree, int32_t M, const float* bg, int32_t width,
                          int32_t height, const float* means3D, const float* shs, const float* colors_precomp,
                          const float* opacities, const float* scales, float scale_modifier, const float* rotations,
                          const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                          const float* campos, float tan_fovx, float tan_fovy, int32_t prefiltered,
                          int32_t antialiasing, float near_plane, float* out_color, int32_t* out_radii,
                          float* out_invdepth, gsw_stream stream);

/* Single-frame forward that allocates nothing.  No counterpart in the reference (its rasterizer grows
 * torch buffers on demand); it exists for callers that cannot call cudaMalloc at render time: a
 * hard-real-time control loop, a plugin confined to a fixed GPU memory budget next to a simulator, or a
 * host runtime that owns the allocator.  Same arguments and results as gsw_rasterize_forward, minus the
 * context: all scratch lives in one fixed ~17 MB array inside libgsw.so (held on every GPU that loads the
 * library until exit), so sizes are capped: P <= 32768 Gaussians, ceil(W/16)*ceil(H/16) <= 1024 tiles,
 * <= 262144 (Gaussian, tile) entries.  Larger P / image: GSW_E_INVALID with a message, nothing launched.
 * Too many entries (data dependent): the frame is left as background and out_status says so.
 *   out_status: device int64[4] or NULL; [0] = entries the frame needed, [2] != 0 if they did not fit.
 * Runs on the CURRENT device.  The first call on a device also looks up the array's address; make it
 * before any stream capture.  All calls on a device share the array (cleared as needed at the start of
 * each call): calls whose GPU work overlaps race, so keep them on one stream or otherwise serialised. */
int gsw_rasterize_forward_static(int32_t P, int32_t sh_degree, int32_t M, const float* bg, int32_t width,
                                 int32_t height, const float* means3D, const float* shs, const float* colors_precomp,
                                 const float* opacities, const float* scales, float scale_modifier,
                                 const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                                 const float* projmatrix, const float* campos, float tan_fovx, float tan_fovy,
                                 int32_t antialiasing, float near_plane, float* out_color, int32_t* out_radii,
                                 float* out_invdepth, int64_t* out_status, gsw_stream stream);

int gsw_mark_visible(gsw_ctx* ctx, int32_t P, const float* means3D, const float* viewmatrix, float near_plane,
                     uint8_t* out_visible, gsw_stream stream);

/* Rig constants, HOST pointers, row-major 4x4: S = sim->GS similarity of the robot, S_inv its inverse,
 * scan_inv[L] = inverse link poses at the scan qpos, link_offset[3] (xarm: 0,0,0.05; NULL = none),
 * obj_S_inv[K] = inverse object similarities, obj_offset[K,3], obj_scale[K].  Synchronises. */
int gsw_rig_set(gsw_ctx* ctx, int L, int K, const float* S, const float* S_inv, const float* scan_inv,
                const float* link_offset, const float* obj_S_inv, const float* obj_offset, const float* obj_scale);
/* One launch: pose_out[B, L+K, GSW_POSE_STRIDE] from link_mats[B,L,16] (sim-world link poses) and
 * obj_pose[B,K,obj_stride] (p(3), q_wxyz(4), ...; obj_stride >= 7), all device pointers.
 * fp32 like the reference; the polar factor of object transforms is computed in fp64. */
int gsw_compose_poses(gsw_ctx* ctx, int B, const float* link_mats, const float* obj_pose, int obj_stride,
                      float* pose_out, gsw_stream stream);
/* One launch: n cameras from OpenCV extrinsics[n,3,4] and intrinsics[n,3,3] (device) to the transposed
 * view / full-projection matrices, camera centre and tan(fov/2) the rasterizer consumes.
 * rigid_sim2real: HOST 4x4 row-major.  Evaluated in fp64 and rounded once (the reference mixes fp32
 * torch and fp64 numpy): agrees with it to ~1 ulp of fp32, not bit-for-bit. */
int gsw_compose_cameras(gsw_ctx* ctx, int n, const float* extrinsic_cv, const float* intrinsic_cv, int height,
                        int width, const float* rigid_sim2real, float scale_sim2real, float* viewmatrix,
                        float* projmatrix, float* campos, float* tanfov, gsw_stream stream);

/* ---- measurement hooks (no counterpart in the reference) ---- */
/* Number of kernels this library has launched in the calling process so far. */
long long gsw_kernel_launches(void);
/* With profiling on, every launch records a CUDA event after each stage on the launch stream;
 * gsw_profile_read waits for the last launch and returns the stage durations in ms, in the order
 * preprocess_static, preprocess_dynamic, scan, scatter, sort (keys), gather, blend (incl. view pass), radii (n >= 8). */
int gsw_profile_enable(gsw_ctx* ctx, int on);
int gsw_profile_read(gsw_ctx* ctx, float* stage_ms, int n);
/* Work counters of the last launch (waits for it): out[0] = (splat, pixel) pair evaluations of the blend kernel, i.e.
 * how often the per-pixel alpha / compositing arithmetic ran (after tile culling, the per-warp footprint cull and
 * early termination; the reference evaluates every entry of a tile's list on all 256 pixels until the tile is done),
 * out[1] = tile entries the blend staged; with n >= 3, out[2] = (view, chunk) pairs of the static layer that
 * survived chunk-level frustum culling (0 when the work list is not in use); with n >= 5, out[3] / out[4] = visible
 * static Gaussian-views / movable Gaussian-frames (records the preprocess wrote); with n >= 6, out[5] = items per
 * chunk of the static layer in that launch (256 ... 2048, chosen from the number of views).  Views whose static lists
 * were kept from the previous launch contribute nothing to out[2] / out[3].  n >= 2.  Feeds bench.py's rooflines. */
int gsw_profile_counters(gsw_ctx* ctx, int64_t* out, int n);

/* ---- parity taps for the LAST gsw_render_frames / gsw_rasterize_forward launch (device outputs) ---- */
/* Per-Gaussian preprocess results of frame f, indexed by original Gaussian index; any pointer NULL ok. */
int gsw_debug_frame_attrs(gsw_ctx* ctx, int32_t frame, int32_t* radii, int32_t* tiles_touched, float* depth,
                          float* xy, float* conic_opacity, float* rgb, gsw_stream stream);
/* Tile lists of frame f in blend order: ranges[T,2] into keys/vals; keys = tile<<32 | depth bits,
 * vals = original Gaussian index (what the upstream sort produces).  Returns GSW_E_CAPACITY if cap is
 * too small; *count (host) receives D for the frame.  Synchronises. */
int gsw_debug_frame_lists(gsw_ctx* ctx, int32_t frame, int32_t* ranges, uint64_t* keys, uint32_t* vals,
                          int64_t cap, int64_t* count, gsw_stream stream);
/* Exclusive offsets of the (list, tile) bins of the last launch, copied to HOST memory: lists 0..V-1 are the
 * static lists of the views, V..V+F-1 the dynamic lists of the frames; dst[nbins] = total entries.  *nbins
 * receives the bin count; GSW_E_CAPACITY if max_bins is too small.  Tuning / test aid (bin-size histogram);
 * no counterpart in the reference.  Synchronises. */
int gsw_debug_bin_offsets(gsw_ctx* ctx, uint32_t* dst_host, int64_t max_bins, int64_t* nbins, gsw_stream stream);

#ifdef __cplusplus
}
#endif
#endif /* GSW_H_ */
