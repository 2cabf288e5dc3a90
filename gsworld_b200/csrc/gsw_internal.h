// gsw_internal.h — shared declarations of libgsw.so (not installed).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <string>

#include "../../include/gsw.h"

namespace gsw {

constexpr int kTile = 16;           // 16x16 pixel tiles, as the reference rasterizer (BLOCK_X/BLOCK_Y)
constexpr int kSortSmemKeys = 4096; // largest segment sorted in shared memory (2 x 32 KB ping-pong of u64 keys)

void set_error(const char* fmt, ...);
extern thread_local std::string g_last_error;
void count_launches(int n);  // bookkeeping for gsw_kernel_launches()

constexpr int kNumStages = 8;  // preprocess_static, preprocess_dynamic, scan, scatter, sort (keys), gather, blend, radii

#define GSW_CUDA(expr)                                                                   \
  do {                                                                                   \
    cudaError_t _e = (expr);                                                             \
    if (_e != cudaSuccess) {                                                             \
      ::gsw::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return (int)_e;                                                                    \
    }                                                                                    \
  } while (0)

// grow-only device buffer
struct DevBuf {
  void* p = nullptr;
  size_t bytes = 0;
  bool fixed = false;  // a slice of memory owned by someone else: never grown, never freed (GSW_E_NOMEM instead)
  uint32_t gen = 0;    // bumped by every (re)allocation: a new block at an old address still reads as a new block
  int reserve(size_t need);  // returns 0 / cudaError / GSW_E_NOMEM
  void release();
  template <typename T>
  T* as() const { return reinterpret_cast<T*>(p); }
};

// Stream record consumed by the blend kernel: 48 B, three float4.
//   a = {px, py, hx, hy}   pixel centre and half extents of the alpha >= 1/255 footprint (cull box)
//   b = {conic.x, conic.y, conic.z, opacity}
//   c = {r, g, b, 1/depth}
struct Splat {
  float4 a, b, c;
};

// Completion signals of the blend launch, for delivering finished frame groups to the host while later
// frames are still being composited (gsw_render_frames with host_rgb_u8).  The F frames are cut into n groups,
// group k = frames [cut(k), cut(k+1)); the last CTA of a group to finish publishes `epoch` in flag[group], on
// which the copy stream waits with a stream memory operation.  cnt == nullptr: no signalling.
// (No array member: a dynamically indexed kernel parameter is copied to local memory by every thread.)
constexpr int kMaxSigGroups = 16;
// Status block: the first 256 bytes are cleared at the start of every launch; int64 word kStatusPoison is not: a
// launch that overflowed sets it, and the next launch's cache_prefix_kernel reads and clears it (the static-list
// cache keeps nothing of an overflowed launch).
constexpr int kStatusBytes = 512;
constexpr int kStatusPoison = 32;
struct DoneSig {
  uint32_t* cnt = nullptr;   // [n] finished CTAs per group (zeroed before the launch)
  uint32_t* flag = nullptr;  // [n] last epoch whose group is complete
  uint32_t epoch = 0;
  int32_t n = 0, F = 0;
  unsigned long long* times = nullptr;  // tuning aid (GSW_SIG_TIMES): [n] globaltimer at a group's completion, [16] first CTA's start
  int32_t geo = 0;  // 1: group sizes double from group to group (first group small: delivery starts early, few groups)
  __host__ __device__ int32_t cut(int k) const {
    if (geo) return (int32_t)((int64_t)F * ((1 << k) - 1) / ((1 << n) - 1));
    return (int32_t)((int64_t)F * k / n);
  }
};

#ifdef __CUDACC__
// ---- TMA bulk copies (cp.async.bulk, global -> this CTA's shared memory) completed through an mbarrier ----
__device__ __forceinline__ void mbar_init(uint32_t mbar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mbar), "r"(count));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t mbar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t mbar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(mbar)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t mbar, uint32_t parity) {
  uint32_t ok;
  do {
    asm volatile(
        "{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
        : "=r"(ok)
        : "r"(mbar), "r"(parity)
        : "memory");
  } while (!ok);
}

// Preprocess record (64 B item slot {px,py,hx,hy | conic,opacity | r,g,b,depth}) -> blend record.  The conic is
// pre-scaled so the blend evaluates alpha = o * 2^(A' dx^2 + B' dx dy + C' dy^2) with A' = -0.5 log2(e) A,
// B' = -log2(e) B, C' = -0.5 log2(e) C.  Used by the gather kernel and by the blend's own (lazy) fetch, so both
// produce the same bits.
__device__ __forceinline__ Splat make_splat(const float4* __restrict__ rec, int64_t item) {
  constexpr float kLog2e = 1.4426950408889634f;
  const float4 a = rec[4 * item], b = rec[4 * item + 1], c = rec[4 * item + 2];
  Splat sp;
  sp.a = a;
  sp.b = make_float4(-0.5f * kLog2e * b.x, -kLog2e * b.y, -0.5f * kLog2e * b.z, b.w);
  sp.c = make_float4(c.x, c.y, c.z, 1 / c.w);
  return sp;
}
#endif

// Inputs of one preprocess "layer" launch.
struct GaussArrays {
  const float* xyz;            // [n,3]
  const float* scale_act;      // [n,3] or null (cov3d given)
  const float* quat_act;       // [n,4] or null
  const float* opacity;        // [n]
  const float* sh;             // [n,M,3] or null
  const float* colors_precomp; // [n,3] or null
  const float* cov3d_precomp;  // [n,6] or null
  const int32_t* slot;         // static layer: skip entries with slot >= 0 (null: none skipped)
  // dynamic layer only
  const float* log_scale_raw;  // [n,3]
  const float* quat_raw;       // [n,4]
  const int32_t* mov_slot;     // [n]
  // index maps of a layer whose items are a re-ordered subset of the scene (null = identity)
  const int32_t* key_idx;      // [n] original Gaussian index (goes into the sort key / radii)
  const int32_t* sh_idx;       // [n] row of `sh` / `colors_precomp` / `cov3d_precomp` to read
  const float* bounds;         // [ceil(n/256), 8] bounding data of each 256-item group (null = no group culling)
  int32_t sh_degree;
  int32_t M;
};

struct FrameGeom {
  int32_t H, W, gx, gy, T;  // tiles
  int32_t NB, nb_shift;     // depth sub-buckets per tile, shift applied to (depth_bits - kDepthLo)
  int32_t antialiasing;
  int32_t prefiltered;
  int32_t tile_cull;  // drop (Gaussian, tile) entries that provably stay below alpha 1/255 on the whole tile
  float scale_modifier;
  float near_plane;
};

constexpr uint32_t kDepthLo = 0x3D000000u;  // 2^-5; sub-buckets span [2^-5, 2^3) in float-bit space
constexpr int kDepthSpanLog2 = 26;

struct Workspace {
  // per-item preprocess results; items = [V*N static | F*Nm dynamic].  rec: 64 B per item =
  // {px,py,hx,hy | conic.xyz,opacity | r,g,b,depth | pad} so a gather touches two aligned 32 B sectors.
  DevBuf rec, rad;
  // bins: (list, tile, bucket); lists = [V static | F dynamic]
  DevBuf bin_count, bin_off, bin_cursor, scan_tmp;
  // tile entries
  DevBuf keys;    // u64 [cap_entries]   sorted (depth bits << 32 | index) per bin
  DevBuf splats;  // Splat [cap_entries] blend records in the same order
  DevBuf merge_cnt, merge_off;  // u32 [T + 1] scratch of the debug list dump
  DevBuf seg_queue;             // Segment[]: depth sub-ranges of bins too large for one shared-memory sort
  DevBuf view_cnt, st_u8, st_color, st_invd, st_seg;  // frames per view; static-only image of each shared view
  DevBuf lab_stream;            // int32 [cap_entries]: label of every stream entry (segmentation observation only)
  DevBuf cta_hist;  // u32: per preprocess CTA {live flag, per-bin entry counts}, consumed by the scatter pass
  // per (list, 2048-item chunk): mask of the 256-item groups that survived group culling ([V static lists | F
  // dynamic lists]); items of dead groups are never written, so every reader of `rad` looks here first
  DevBuf chunk_mask;
  DevBuf work;         // u32 (view << 16 | chunk): live (view, chunk) pairs of the static layer, cull_chunks_kernel
  DevBuf view_consts;  // 48 words per view: ViewSmem + ViewCull, written by cull_chunks_kernel
  // dynamic layer: weight (tile entries) of every (frame, chunk) CTA of the last launch and the dispatch order made
  // from it (order_chunks_kernel); the order outlives the launch: the next step's preprocess uses it
  DevBuf dyn_cost, dyn_order;
  // static-list cache across steps (cache_prefix_kernel): last launch's camera per view [V][40], frames per view,
  // "camera unchanged" flags of this launch, and m = leading views whose static lists are kept
  DevBuf view_prev, view_cnt_prev, view_same, cache_m;
  DevBuf status;  // device status words
  DevBuf cam;     // single-frame API staging: view/proj/campos/tanfov/frame maps
};

}  // namespace gsw

struct gsw_ctx {
  int device = 0;
  // scene
  int64_t N = 0;
  int32_t num_slots = 0;
  int64_t Nm = 0;  // movable count
  gsw::GaussArrays scene{};
  gsw::DevBuf mov_xyz, mov_scale_act, mov_opacity, mov_sh, mov_log_scale, mov_quat_raw, mov_slot, mov_idx, mov_bounds;
  // static layer: the Gaussians with slot < 0, re-ordered along a Morton curve so that 256-item groups are
  // spatially compact (per-view frustum culling of whole groups); compact copies of the small arrays
  int64_t Ns = 0;
  gsw::DevBuf st_xyz, st_scale, st_quat, st_opacity, st_idx, st_bounds;
  gsw::DevBuf pos_of;           // [N] original index -> position inside its layer (static or movable)
  const int32_t* slot_orig = nullptr;  // caller's slot array (original order)
  const int32_t* labels = nullptr;     // caller's per-Gaussian labels (original order), gsw_scene_set_labels
  gsw::GaussArrays stat_layer{};
  bool have_scene = false;
  // rig constants (gsw_rig_set): sim->GS similarity, its inverse, per-link scan-pose inverses, objects
  struct {
    bool valid = false;
    int L = 0, K = 0;
    float S[16] = {}, S_inv[16] = {}, link_off[3] = {};
    gsw::DevBuf dev;  // [L*16 scan_inv | K*16 obj_S_inv | K*3 obj_off | K obj_scale]
  } rig;
  // workspace
  gsw::Workspace ws;
  int64_t cap_entries = 0;  // capacity of [sorted lists | merged lists] combined
  int nb_user = 0;
  int tile_cull = 0;  // gsw_set_tile_culling
  // last launch
  struct {
    bool valid = false;
    bool single = false;  // gsw_rasterize_forward launch (scene arrays are the caller's)
    int32_t V = 0, F = 0;
    int64_t N = 0, Ns = 0, Nm = 0;        // original count, static-layer items per view, movable items per frame
    const int32_t* slot_orig = nullptr;   // null: every Gaussian is in the static layer
    const int32_t* pos_of = nullptr;      // null: identity
    gsw::FrameGeom g{};
    gsw::GaussArrays stat{};
    const int32_t* frame_env = nullptr;
    const int32_t* frame_view = nullptr;
    int64_t nbins = 0, cap = 0;
  } last;
  int64_t* h_status = nullptr;  // pinned: [0]=D, [1]=unused, [2]=overflow flag, [3]=entries staged by the blend
  bool profile = false;         // record an event after every stage (gsw_profile_*)
  cudaEvent_t stage_evt[gsw::kNumStages + 1] = {};
  cudaEvent_t done_evt = nullptr;
  cudaStream_t copy_stream = nullptr;  // device->host delivery of finished frame groups (host_rgb_u8)
  cudaEvent_t grp_evt[5] = {};
  void* wait_value32 = nullptr;  // cuStreamWaitValue32 (driver entry point; null: grouped blend launches instead)
  gsw::DevBuf sig;               // DoneSig counters [16] + flags [16]
  uint32_t sig_epoch = 0;
  cudaStream_t peer_stream = nullptr;  // pushes to peer GPUs (NVLink), so that they do not queue behind PCIe copies
  cudaEvent_t peer_evt = nullptr;
  cudaEvent_t gate_evt = nullptr;  // gsw_delivery_gate
  cudaEvent_t copy_evt = nullptr;
  bool delivered = false;  // copy_evt has been recorded by a delivery
  bool delivered_peer = false;  // ... and peer_evt (the last delivery pushed to peers)
  // static-layer kernels (preprocess, scatter) run on side_stream next to the dynamic-layer ones: they are
  // independent until the scan / the sort, and each leaves a good part of the SMs idle on its own
  cudaStream_t side_stream = nullptr;
  cudaStream_t side2_stream = nullptr;  // the dynamic lists' sort forks onto it while side_stream carries the static chain
  cudaEvent_t fork_evt = nullptr, join_evt = nullptr;
  cudaEvent_t order_evt[2] = {};  // the dispatch-order kernel of the dynamic layer runs beside the bin scan
  int64_t launches = 0;
  // static-list cache: cache_on = this launch runs the cache logic (scene layers with a chunk work list); cache_ok = and
  // nothing the kept lists depend on has changed on the host side since the previous launch (cache_sig)
  bool cache_on = false, cache_ok = false;
  int32_t ipb_static = 2048;  // items per chunk of the static layer in the last launch (pick_static_ipb)
  struct CacheSig {
    uint64_t scene_gen = 0;
    int64_t Ns = 0, cap = 0;
    int32_t V = 0, F = 0, H = 0, W = 0, NB = 0, tile_cull = 0, antialiasing = 0, outputs = 0;
    float scale_modifier = 0.f, near_plane = 0.f, bg[3] = {0.f, 0.f, 0.f};
    int32_t ipb_static = 0;  // (also: no padding bytes, the signature is compared with memcmp)
    const void* ptrs[16] = {};
    uint32_t gens[16] = {};
    const void* stream = nullptr;  // kept lists are only ordered before a launch on the same stream
  } cache_sig;
  uint64_t scene_gen = 0;      // bumped by gsw_scene_set / gsw_scene_set_labels
  int32_t view_pass_outputs = 0;  // set by gsw_render_frames before run_pipeline: which static-only images the view pass writes
  float view_pass_bg[3] = {0.f, 0.f, 0.f};
  bool side_pending = false;  // run_pipeline left the static chain running on side_stream (join_side closes it)
  int64_t dyn_order_n = 0;  // number of (frame, chunk) pairs ws.dyn_order is valid for (0: none yet)
};

namespace gsw {

// ---- kernels (host launchers) ----
// preprocess.cu  (compiled with -fmad=false: every fp32 op rounds once, like the CPU oracle)
int launch_preprocess(const gsw_ctx* ctx, bool dynamic, const GaussArrays& ga, int64_t n_items, int32_t n_lists,
                      int32_t list_base, int64_t item_base, const FrameGeom& g, const float* view, const float* proj,
                      const float* campos, const float* tanfov, const int32_t* frame_env, const int32_t* frame_view,
                      const float* pose, int32_t num_slots, uint32_t* cta_hist, uint32_t* chunk_mask, int* kept_hist,
                      cudaStream_t s, const uint32_t* order = nullptr, uint32_t* cost = nullptr);
// heaviest-first dispatch order of n (list, chunk) CTAs from their weights (tile entries produced)
int launch_order_chunks(const uint32_t* cost, int64_t n, uint32_t* order, cudaStream_t s);
int launch_cache_prefix(const gsw_ctx* ctx, int32_t V, cudaStream_t s);
size_t cta_hist_words(const gsw_ctx* ctx, int64_t n_items, int32_t n_lists, const FrameGeom& g, bool dynamic);
// chunks of a layer (one preprocess / scatter CTA round each); the static layer's chunk size is ctx->ipb_static
int64_t layer_chunks(const gsw_ctx* ctx, int64_t n_items, bool dynamic);
int pick_static_ipb(const GaussArrays& ga, int64_t n_items, int32_t n_lists, int32_t bins_per_list);
int launch_scatter(const gsw_ctx* ctx, bool dynamic, const GaussArrays& ga, int64_t n_items, int32_t n_lists,
                   int32_t list_base, int64_t item_base, const FrameGeom& g, int64_t cap, const uint32_t* cta_hist,
                   const uint32_t* chunk_mask, cudaStream_t s, const uint32_t* order = nullptr);
int launch_frame_attrs(const gsw_ctx* ctx, int32_t frame, int32_t* radii, int32_t* tiles_touched, float* depth,
                       float* xy, float* conic_opacity, float* rgb, cudaStream_t s);
int launch_radii(const gsw_ctx* ctx, int32_t* out_radii, cudaStream_t s);
int launch_mark_visible(int32_t P, const float* means3D, const float* viewmatrix, float near_plane, uint8_t* out,
                        cudaStream_t s);
int launch_gather_layers(const gsw_ctx* ctx, cudaStream_t s);  // compact static + movable copies, group bounds

// binsort.cu
int launch_scan_u32(const uint32_t* in, uint32_t* out_excl, int64_t n, uint32_t* tmp, cudaStream_t s);
int launch_sort_bins(const gsw_ctx* ctx, int64_t bin0, int64_t nb, int64_t cap, int set, cudaStream_t s,
                     cudaStream_t side, int32_t bins_per_list);
int launch_gather(const gsw_ctx* ctx, int64_t nbins, int32_t V, int64_t Ns, int64_t Nm, const FrameGeom& g, int64_t cap,
                  bool with_labels, cudaStream_t s);
int launch_frame_lists(const gsw_ctx* ctx, int32_t frame, int32_t* ranges, uint64_t* keys, uint32_t* vals,
                       int64_t cap, cudaStream_t s);

// pose.cu
int launch_compose_poses(const gsw_ctx* ctx, int B, const float* link_mats, const float* obj_pose, int obj_stride,
                         float* pose_out, cudaStream_t s);
int launch_compose_cameras(int n, const float* ext, const float* K, int H, int W, const float* rigid, float scale,
                           float* view, float* proj, float* campos, float* tanfov, cudaStream_t s);

// blend.cu
int launch_blend(const gsw_ctx* ctx, int32_t V, int32_t f_begin, int32_t F, int64_t nbins, const int32_t* frame_view, int64_t cap,
                 const FrameGeom& g, const float bg[3], const float* bg_dev, uint8_t* out_u8, float* out_color,
                 float* out_invdepth, int32_t* out_seg, bool share, cudaStream_t s, const DoneSig* sig = nullptr);
// per-view preprocess constants (ws.view_consts) + frames per view (ws.view_cnt): first kernel of a launch set
int launch_prepare_views(const gsw_ctx* ctx, int32_t V, int32_t F, const FrameGeom& g, const float* view,
                         const float* proj, const float* campos, const float* tanfov, const int32_t* frame_view,
                         cudaStream_t s);
int launch_blend_view_pass(const gsw_ctx* ctx, int32_t V, int32_t F, int64_t nbins, const int32_t* frame_view,
                           int64_t cap, const FrameGeom& g, const float bg[3], bool want_u8, bool want_color,
                           bool want_invd, bool want_seg, cudaStream_t s);

size_t scan_tmp_elems(int64_t n);
size_t seg_queue_bytes(int64_t cap, int64_t nbins);  // sort work queues (launch_sort_bins) for `cap` entries in `nbins` bins

}  // namespace gsw
