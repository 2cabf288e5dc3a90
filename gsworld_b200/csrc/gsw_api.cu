// gsw_api.cu — extern "C" entry points of libgsw.so and launch orchestration.  See include/gsw.h.
#include <cuda.h>  // types of the one driver entry point used (cuStreamWaitValue32); libcuda is not linked
#include <algorithm>
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <vector>

#include "gsw_internal.h"

namespace gsw {

thread_local std::string g_last_error;
static std::atomic<long long> g_launches{0};
void count_launches(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

void set_error(const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_last_error = buf;
}

int DevBuf::reserve(size_t need) {
  if (need <= bytes) return 0;
  if (fixed) {  // carved out of a fixed array (gsw_rasterize_forward_static): growing would mean allocating
    set_error("fixed workspace slice of %zu bytes cannot hold %zu bytes", bytes, need);
    return GSW_E_NOMEM;
  }
  if (p) {
    cudaError_t e = cudaFree(p);  // synchronises: no kernel still uses the old block afterwards
    p = nullptr;
    bytes = 0;
    if (e != cudaSuccess) {
      set_error("cudaFree failed: %s", cudaGetErrorString(e));
      return (int)e;
    }
  }
  const size_t want = (need + 255) & ~(size_t)255;
  cudaError_t e = cudaMalloc(&p, want);
  if (e != cudaSuccess) {
    p = nullptr;
    set_error("cudaMalloc(%zu bytes) failed: %s", want, cudaGetErrorString(e));
    return (int)e;
  }
  bytes = want;
  ++gen;
  return 0;
}

void DevBuf::release() {
  if (p && !fixed) cudaFree(p);
  p = nullptr;
  bytes = 0;
}

namespace {

// tan(fov/2) and the trivial frame maps of the single-frame entry points: [tanx, tany, env 0, view 0].  Written by
// a one-thread kernel from by-value arguments, so those calls read no host memory after they return.
__global__ void set_single_cam_kernel(float* cam, float tanx, float tany) {
  cam[0] = tanx;
  cam[1] = tany;
  reinterpret_cast<int32_t*>(cam)[2] = 0;
  reinterpret_cast<int32_t*>(cam)[3] = 0;
}

int stage_single_cam(DevBuf& cam, float tanx, float tany, cudaStream_t s) {
  set_single_cam_kernel<<<1, 1, 0, s>>>(cam.as<float>(), tanx, tany);
  count_launches(1);
  GSW_CUDA(cudaGetLastError());
  return 0;
}

int ilog2(int v) {
  int l = 0;
  while ((1 << l) < v) ++l;
  return l;
}

FrameGeom make_geom(int H, int W, int nb, int antialiasing, float scale_modifier, float near_plane) {
  FrameGeom g{};
  g.H = H;
  g.W = W;
  g.gx = (W + kTile - 1) / kTile;
  g.gy = (H + kTile - 1) / kTile;
  g.T = g.gx * g.gy;
  g.NB = nb;
  g.nb_shift = kDepthSpanLog2 - ilog2(nb);
  g.antialiasing = antialiasing;
  g.prefiltered = 0;
  g.tile_cull = 0;
  g.scale_modifier = scale_modifier;
  g.near_plane = near_plane;
  return g;
}

struct Guard {  // make ctx->device current for the duration of a call
  int prev = -1;
  bool ok = true;
  explicit Guard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    if (prev != dev) ok = cudaSetDevice(dev) == cudaSuccess;
  }
  ~Guard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

// Shared body of gsw_render_frames / gsw_rasterize_forward once the layer inputs are known.
int run_pipeline(gsw_ctx* ctx, const GaussArrays& stat, int64_t N, int64_t Ns, const int32_t* slot_orig,
                 const int32_t* pos_of, const GaussArrays* dyn, int64_t Nm, int32_t V, int32_t F, const FrameGeom& g, const float* view, const float* proj, const float* campos,
                 const float* tanfov, const int32_t* frame_env, const int32_t* frame_view, const float* pose,
                 int32_t num_slots, cudaStream_t s, bool with_labels = false) {
  auto& w = ctx->ws;
  const int64_t items = (int64_t)V * Ns + (int64_t)F * Nm;  // N = original count (outputs), Ns = static-layer items
  const int64_t nbins = (int64_t)(V + F) * g.T * g.NB;
  const int64_t FT = (int64_t)F * g.T;
  if (nbins >= (int64_t)1 << 31 || items >= (int64_t)1 << 40) {
    set_error("launch too large: %lld bins, %lld items (split the frames into several calls)", (long long)nbins,
              (long long)items);
    return GSW_E_INVALID;
  }
  if (ctx->cap_entries == 0) ctx->cap_entries = items + items / 2 > (4 << 20) ? items + items / 2 : (4 << 20);
  if (ctx->cap_entries >= ((int64_t)1 << 32) - 1) ctx->cap_entries = ((int64_t)1 << 32) - 2;
  const int64_t cap = ctx->cap_entries;
  int rc;
  const size_t it = (size_t)(items > 0 ? items : 1);
  if ((rc = w.rec.reserve(it * 64)) || (rc = w.rad.reserve(it * 4))) return rc;
  if ((rc = w.bin_count.reserve((size_t)(nbins + 1) * 4)) || (rc = w.bin_off.reserve((size_t)(nbins + 1) * 4)) ||
      (rc = w.bin_cursor.reserve((size_t)(nbins + 1) * 4)))
    return rc;
  const int64_t scan_n = nbins;
  (void)FT;
  if ((rc = w.scan_tmp.reserve(scan_tmp_elems(scan_n) * 4))) return rc;
  const int64_t mc = (int64_t)g.T + 1;
  if ((rc = w.merge_cnt.reserve((size_t)mc * 4)) || (rc = w.merge_off.reserve((size_t)mc * 4))) return rc;
  if ((rc = w.keys.reserve((size_t)cap * 8)) || (rc = w.splats.reserve((size_t)cap * sizeof(Splat)))) return rc;
  if ((rc = w.status.reserve(kStatusBytes))) return rc;
  ctx->ipb_static = pick_static_ipb(stat, Ns, V, g.T * g.NB);
  const size_t chs = cta_hist_words(ctx, Ns, V, g, false), chd = dyn ? cta_hist_words(ctx, Nm, F, g, true) : 0;
  if ((rc = w.cta_hist.reserve((chs + chd) * 4 + 16))) return rc;
  uint32_t* const ch_s = chs ? w.cta_hist.as<uint32_t>() : nullptr;
  uint32_t* const ch_d = chd ? w.cta_hist.as<uint32_t>() + chs : nullptr;
  const size_t cm_s = (size_t)V * (size_t)layer_chunks(ctx, Ns, false), cm_d = dyn ? (size_t)F * (size_t)layer_chunks(ctx, Nm, true) : 0;
  if ((rc = w.chunk_mask.reserve((cm_s + cm_d) * 4 + 16)) || (rc = w.work.reserve(cm_s * 4 + 16)) ||
      (rc = w.view_consts.reserve((size_t)V * 48 * 4)))
    return rc;
  uint32_t* const mask_s = w.chunk_mask.as<uint32_t>();
  uint32_t* const mask_d = w.chunk_mask.as<uint32_t>() + cm_s;
  int kept_s = 0, kept_d = 0;
  // medium and mid queue (half each): whole bins (<= nbins each) plus sub-segments of huge bins, of which
  // every one but the last of a bin pairs with a neighbour to more than kSortSmemKeys / 2 entries
  if ((rc = w.seg_queue.reserve(2 * seg_queue_bytes(cap, nbins)))) return rc;  // two independent queue sets

  // ---- static-list cache across steps (cache_prefix_kernel) ----
  // cache_on: the static layer goes through the chunk work list (scene layers).  cache_ok: everything the kept lists
  // depend on besides the cameras is what it was at the previous launch: scene, sizes, geometry, the view pass'
  // outputs and background, and the workspace blocks themselves (a reallocation loses their contents).
  const int64_t nbs0 = (int64_t)V * g.T * g.NB;
  ctx->cache_on = stat.bounds != nullptr && V <= 65535 && layer_chunks(ctx, Ns, false) <= 65535 && Ns > 0 &&
                  getenv("GSW_NO_WORK_LIST") == nullptr && getenv("GSW_NO_STATIC_CACHE") == nullptr;
  ctx->cache_ok = false;
  if (ctx->cache_on) {
    if ((rc = w.view_prev.reserve((size_t)V * 40 * 4)) || (rc = w.view_cnt_prev.reserve((size_t)V * 4 + 16)) ||
        (rc = w.view_same.reserve((size_t)V * 4 + 16)) || (rc = w.cache_m.reserve(16)))
      return rc;
    gsw_ctx::CacheSig sig;
    sig.scene_gen = ctx->scene_gen;
    sig.stream = (const void*)s;
    sig.Ns = Ns;
    sig.cap = cap;
    sig.V = V; sig.F = F; sig.H = g.H; sig.W = g.W; sig.NB = g.NB; sig.tile_cull = g.tile_cull;
    sig.antialiasing = g.antialiasing;
    sig.ipb_static = ctx->ipb_static;
    sig.outputs = ctx->view_pass_outputs | (with_labels ? 16 : 0);
    sig.scale_modifier = g.scale_modifier;
    sig.near_plane = g.near_plane;
    for (int k = 0; k < 3; ++k) sig.bg[k] = ctx->view_pass_bg[k];
    const DevBuf* bufs[16] = {&w.rec, &w.rad, &w.keys, &w.splats, &w.bin_count, &w.bin_off, &w.cta_hist,
                              &w.chunk_mask, &w.st_u8, &w.st_color, &w.st_invd, &w.st_seg, &w.lab_stream,
                              &w.view_prev, &w.view_cnt_prev, &w.view_same};
    for (int k = 0; k < 16; ++k) {
      sig.ptrs[k] = bufs[k]->p;
      sig.gens[k] = bufs[k]->gen;
    }
    ctx->cache_ok = memcmp(&sig, &ctx->cache_sig, sizeof sig) == 0;
    ctx->cache_sig = sig;
  } else {
    ctx->cache_sig = gsw_ctx::CacheSig{};
  }
  if (ctx->cache_on) {
    // the static bins are cleared on the device, view by view, for the views that are rebuilt (cull_chunks_kernel)
    GSW_CUDA(cudaMemsetAsync(w.bin_count.as<uint32_t>() + nbs0, 0, (size_t)(nbins - nbs0 + 1) * 4, s));
    GSW_CUDA(cudaMemsetAsync(w.bin_cursor.as<uint32_t>() + nbs0, 0, (size_t)(nbins - nbs0 + 1) * 4, s));
  } else {
    GSW_CUDA(cudaMemsetAsync(w.bin_count.p, 0, (size_t)(nbins + 1) * 4, s));
    GSW_CUDA(cudaMemsetAsync(w.bin_cursor.p, 0, (size_t)(nbins + 1) * 4, s));
  }
  GSW_CUDA(cudaMemsetAsync(w.status.p, 0, 256, s));
  auto mark = [&](int k) {
    if (ctx->profile) cudaEventRecord(ctx->stage_evt[k], s);
  };

  // The static and the dynamic layer are independent up to the scan and again up to the sort; each kernel alone
  // leaves SMs idle towards its end (ncu: SMs active 32-59 % of the static kernels' duration), so the static
  // ones run on a side stream.  With stage profiling on everything stays on `s` so that the stage events
  // bracket one layer each (the stage times then add up to more than the step).
  const bool two = ctx->side_stream != nullptr && !ctx->profile && dyn != nullptr && Nm > 0 && Ns > 0;
  cudaStream_t ss = two ? ctx->side_stream : s;
  auto fork = [&]() -> int {
    if (!two) return 0;
    GSW_CUDA(cudaEventRecord(ctx->fork_evt, s));
    GSW_CUDA(cudaStreamWaitEvent(ss, ctx->fork_evt, 0));
    return 0;
  };
  auto join = [&]() -> int {
    if (!two) return 0;
    GSW_CUDA(cudaEventRecord(ctx->join_evt, ss));
    GSW_CUDA(cudaStreamWaitEvent(s, ctx->join_evt, 0));
    return 0;
  };
  if ((rc = w.view_cnt.reserve((size_t)V * 4 + 16))) return rc;
  if ((rc = launch_prepare_views(ctx, V, F, g, view, proj, campos, tanfov, frame_view, s))) return rc;
  mark(0);
  if ((rc = fork())) return rc;
  if (ctx->cache_on && (rc = launch_cache_prefix(ctx, V, ss))) return rc;
  if ((rc = launch_preprocess(ctx, false, stat, Ns, V, 0, 0, g, view, proj, campos, tanfov, frame_env, frame_view,
                              pose, num_slots, ch_s, mask_s, &kept_s, ss)))
    return rc;
  mark(1);
  // dynamic layer: CTAs are dispatched heaviest chunk first.  The preprocess pass takes the order of the previous
  // step (same frame / chunk count), the scatter pass the order made from this step's own weights.
  const int64_t n_dyn_ctas = (dyn && Nm > 0) ? (int64_t)cm_d : 0;
  const bool use_order = n_dyn_ctas > 1 && n_dyn_ctas <= (1 << 20) && chd > 0 && getenv("GSW_NO_ORDER") == nullptr;
  if (use_order &&
      ((rc = w.dyn_cost.reserve((size_t)n_dyn_ctas * 4)) || (rc = w.dyn_order.reserve((size_t)n_dyn_ctas * 4))))
    return rc;
  const uint32_t* prev_order = (use_order && ctx->dyn_order_n == n_dyn_ctas) ? w.dyn_order.as<uint32_t>() : nullptr;
  bool have_order = false, order_aside = false;
  if (dyn && Nm > 0) {
    if ((rc = launch_preprocess(ctx, true, *dyn, Nm, F, V, (int64_t)V * Ns, g, view, proj, campos, tanfov, frame_env,
                                frame_view, pose, num_slots, ch_d, mask_d, &kept_d, s, prev_order,
                                use_order ? w.dyn_cost.as<uint32_t>() : nullptr)))
      return rc;
    if (use_order && kept_d) {
      // The order is first needed by the scatter.  In a small launch, where the step is a chain of launch gaps, the
      // kernel that makes it runs beside the bin scan instead of before it (8 envs x 2 cameras: 0.360 -> 0.350 ms per
      // step); in a large one that measured slower (64 envs: 1.008 -> 1.027 ms), so there it stays in line.
      static const bool inline_order = getenv("GSW_ORDER_INLINE") != nullptr;  // tuning aid
      cudaStream_t os = (ctx->side2_stream != nullptr && ctx->order_evt[1] != nullptr && !inline_order &&
                         !ctx->profile && n_dyn_ctas <= 1024)
                            ? ctx->side2_stream : s;
      if (os != s) {
        GSW_CUDA(cudaEventRecord(ctx->order_evt[0], s));
        GSW_CUDA(cudaStreamWaitEvent(os, ctx->order_evt[0], 0));
      }
      if ((rc = launch_order_chunks(w.dyn_cost.as<uint32_t>(), n_dyn_ctas, w.dyn_order.as<uint32_t>(), os))) return rc;
      if (os != s) {
        GSW_CUDA(cudaEventRecord(ctx->order_evt[1], os));
        order_aside = true;
      }
      ctx->dyn_order_n = n_dyn_ctas;
      have_order = true;
    } else {
      ctx->dyn_order_n = 0;
    }
  }
  if ((rc = join())) return rc;
  mark(2);
  if ((rc = launch_scan_u32(w.bin_count.as<uint32_t>(), w.bin_off.as<uint32_t>(), nbins, w.scan_tmp.as<uint32_t>(),
                            s)))
    return rc;
  mark(3);
  if ((rc = fork())) return rc;
  if ((rc = launch_scatter(ctx, false, stat, Ns, V, 0, 0, g, cap, kept_s ? ch_s : nullptr, mask_s, ss))) return rc;
  if (order_aside) GSW_CUDA(cudaStreamWaitEvent(s, ctx->order_evt[1], 0));
  if (dyn && Nm > 0)
    if ((rc = launch_scatter(ctx, true, *dyn, Nm, F, V, (int64_t)V * Ns, g, cap, kept_d ? ch_d : nullptr, mask_d, s,
                             have_order ? w.dyn_order.as<uint32_t>() : nullptr)))
      return rc;
  ctx->last.pos_of = pos_of;  // the gather reads it
  if (with_labels && (rc = w.lab_stream.reserve((size_t)cap * 4))) return rc;
  const int64_t nbs = (int64_t)V * g.T * g.NB;  // static bins come first
  if (two) {
    // From here on the two layers are independent until the blend.  Side stream: sort of the static lists, gather
    // of the shared ones and (gsw_render_frames) the view pass over them.  Main stream: sort of the dynamic lists,
    // which forks onto a third stream inside.  join_side() closes the fork before the blend.
    if ((rc = launch_sort_bins(ctx, 0, nbs, cap, 1, ss, ss, g.T * g.NB))) return rc;
    if ((rc = launch_gather(ctx, nbins, V, Ns, Nm, g, cap, with_labels, ss))) return rc;
    if ((rc = launch_sort_bins(ctx, nbs, nbins - nbs, cap, 0, s, ctx->side2_stream ? ctx->side2_stream : s,
                               g.T * g.NB)))
      return rc;
    ctx->side_pending = true;
  } else {
    if ((rc = join())) return rc;
    mark(4);
    if ((rc = launch_sort_bins(ctx, 0, nbins, cap, 0, s, ctx->side_stream ? ctx->side_stream : s, g.T * g.NB)))
      return rc;
    mark(5);
    if ((rc = launch_gather(ctx, nbins, V, Ns, Nm, g, cap, with_labels, s))) return rc;
    mark(6);
  }

  ctx->last.valid = true;
  ctx->last.V = V;
  ctx->last.F = F;
  ctx->last.N = N;
  ctx->last.Ns = Ns;
  ctx->last.Nm = Nm;
  ctx->last.slot_orig = slot_orig;
  ctx->last.g = g;
  ctx->last.stat = stat;
  ctx->last.frame_env = frame_env;
  ctx->last.frame_view = frame_view;
  ctx->last.nbins = nbins;
  ctx->last.cap = cap;
  return 0;
}

// Closes the fork run_pipeline left open (static chain on the side stream): `s` waits for it.
int join_side(gsw_ctx* ctx, cudaStream_t s) {
  if (!ctx->side_pending) return 0;
  ctx->side_pending = false;
  GSW_CUDA(cudaEventRecord(ctx->join_evt, ctx->side_stream));
  GSW_CUDA(cudaStreamWaitEvent(s, ctx->join_evt, 0));
  return 0;
}

int finish_launch(gsw_ctx* ctx, cudaStream_t s) {
  GSW_CUDA(cudaMemcpyAsync(ctx->h_status, ctx->ws.status.p, 18 * sizeof(int64_t), cudaMemcpyDeviceToHost, s));
  GSW_CUDA(cudaEventRecord(ctx->done_evt, s));
  ctx->launches++;
  return 0;
}

}  // namespace
}  // namespace gsw

using namespace gsw;

extern "C" {

int gsw_version(void) { return GSW_VERSION; }

const char* gsw_last_error(void) { return g_last_error.c_str(); }

int gsw_create(int device, gsw_ctx** out) {
  if (!out) {
    set_error("gsw_create: out is NULL");
    return GSW_E_INVALID;
  }
  *out = nullptr;
  int ndev = 0;
  GSW_CUDA(cudaGetDeviceCount(&ndev));
  if (device < 0 || device >= ndev) {
    set_error("gsw_create: device %d out of range (have %d)", device, ndev);
    return GSW_E_INVALID;
  }
  Guard gd(device);
  gsw_ctx* c = new gsw_ctx();
  c->device = device;
  cudaError_t e = cudaMallocHost((void**)&c->h_status, 32 * sizeof(int64_t));
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->done_evt, cudaEventDisableTiming);
  for (int k = 0; k <= kNumStages && e == cudaSuccess; ++k) e = cudaEventCreate(&c->stage_evt[k]);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking);
  for (int k = 0; k < 5 && e == cudaSuccess; ++k) e = cudaEventCreateWithFlags(&c->grp_evt[k], cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->copy_evt, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->peer_stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->peer_evt, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->gate_evt, cudaEventDisableTiming);
  if (e == cudaSuccess) {
    // stream memory operation used for in-kernel completion signals (no link-time dependency on libcuda)
    cudaDriverEntryPointQueryResult qr = cudaDriverEntryPointSymbolNotFound;
    void* fn = nullptr;
    if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &fn, cudaEnableDefault, &qr) == cudaSuccess &&
        qr == cudaDriverEntryPointSuccess && getenv("GSW_NO_WAIT_VALUE") == nullptr)
      c->wait_value32 = fn;
    else
      (void)cudaGetLastError();
    // counters [16] + flags [16] + (tuning aid) completion times [17] as 64-bit words
    if (c->sig.reserve(2 * kMaxSigGroups * sizeof(uint32_t) + 32 * 8) != 0) c->wait_value32 = nullptr;
    else e = cudaMemset(c->sig.p, 0, 2 * kMaxSigGroups * sizeof(uint32_t) + 32 * 8);
  }
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->side_stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->side2_stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->fork_evt, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->join_evt, cudaEventDisableTiming);
  for (int k = 0; k < 2 && e == cudaSuccess; ++k) e = cudaEventCreateWithFlags(&c->order_evt[k], cudaEventDisableTiming);
  if (e != cudaSuccess) {
    set_error("gsw_create: %s", cudaGetErrorString(e));
    delete c;
    return (int)e;
  }
  memset(c->h_status, 0, 32 * sizeof(int64_t));
  *out = c;
  return GSW_OK;
}

int gsw_destroy(gsw_ctx* c) {
  if (!c) return GSW_OK;
  Guard gd(c->device);
  cudaDeviceSynchronize();
  auto& w = c->ws;
  DevBuf* bufs[] = {&w.rec, &w.rad, &w.bin_count, &w.bin_off, &w.bin_cursor, &w.scan_tmp,
                    &w.keys, &w.splats, &w.seg_queue, &w.cta_hist, &w.chunk_mask, &w.work, &w.view_consts, &w.dyn_cost, &w.dyn_order, &w.view_prev, &w.view_cnt_prev, &w.view_same, &w.cache_m, &c->sig, &w.view_cnt, &w.st_u8, &w.st_color, &w.st_invd, &w.st_seg, &w.lab_stream, &w.merge_cnt, &w.merge_off, &w.status, &w.cam,
                    &c->mov_xyz, &c->mov_scale_act, &c->mov_opacity, &c->mov_sh, &c->mov_log_scale,
                    &c->mov_quat_raw, &c->mov_slot, &c->mov_idx, &c->mov_bounds, &c->pos_of, &c->st_xyz, &c->st_scale,
                    &c->st_quat, &c->st_opacity, &c->st_idx, &c->st_bounds, &c->rig.dev};
  for (DevBuf* b : bufs) b->release();
  if (c->h_status) cudaFreeHost(c->h_status);
  if (c->done_evt) cudaEventDestroy(c->done_evt);
  for (int k = 0; k < 5; ++k)
    if (c->grp_evt[k]) cudaEventDestroy(c->grp_evt[k]);
  if (c->copy_evt) cudaEventDestroy(c->copy_evt);
  if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
  if (c->peer_evt) cudaEventDestroy(c->peer_evt);
  if (c->gate_evt) cudaEventDestroy(c->gate_evt);
  if (c->peer_stream) cudaStreamDestroy(c->peer_stream);
  if (c->fork_evt) cudaEventDestroy(c->fork_evt);
  if (c->join_evt) cudaEventDestroy(c->join_evt);
  for (int k = 0; k < 2; ++k)
    if (c->order_evt[k]) cudaEventDestroy(c->order_evt[k]);
  if (c->side_stream) cudaStreamDestroy(c->side_stream);
  if (c->side2_stream) cudaStreamDestroy(c->side2_stream);
  for (int k = 0; k <= kNumStages; ++k)
    if (c->stage_evt[k]) cudaEventDestroy(c->stage_evt[k]);
  delete c;
  return GSW_OK;
}

int gsw_scene_set(gsw_ctx* c, int64_t N, const float* xyz, const float* scale_act, const float* quat_act,
                  const float* opacity_act, const float* sh, const float* log_scale_raw, const float* quat_raw,
                  const int32_t* slot, int num_slots, gsw_stream stream) {
  if (!c || N < 0 || N >= ((int64_t)1 << 31) || (N > 0 && (!xyz || !scale_act || !quat_act || !opacity_act || !sh))) {
    set_error("gsw_scene_set: invalid argument");
    return GSW_E_INVALID;
  }
  if (num_slots < 0 || num_slots > 512) {
    set_error("gsw_scene_set: num_slots %d out of range [0,512]", num_slots);
    return GSW_E_INVALID;
  }
  if (slot && num_slots > 0 && N > 0 && (!log_scale_raw || !quat_raw)) {
    set_error("gsw_scene_set: log_scale_raw/quat_raw are required when slots are given");
    return GSW_E_INVALID;
  }
  Guard gd(c->device);
  cudaStream_t s = (cudaStream_t)stream;
  c->have_scene = false;
  c->dyn_order_n = 0;
  c->scene_gen++;
  c->last.valid = false;
  c->N = N;
  c->num_slots = num_slots;
  c->scene = GaussArrays{};
  c->scene.xyz = xyz;
  c->scene.scale_act = scale_act;
  c->scene.quat_act = quat_act;
  c->scene.opacity = opacity_act;
  c->scene.sh = sh;
  c->scene.log_scale_raw = log_scale_raw;
  c->scene.quat_raw = quat_raw;
  c->scene.slot = (slot && num_slots > 0) ? slot : nullptr;
  c->scene.sh_degree = 3;
  c->scene.M = 16;
  c->slot_orig = c->scene.slot;
  c->labels = nullptr;
  c->Nm = 0;
  c->Ns = 0;
  if (N == 0) {
    c->stat_layer = c->scene;
    c->have_scene = true;
    return GSW_OK;
  }
  // ---- host side, once per scene: split into layers and order each along a Morton curve ----
  std::vector<int32_t> hslot;
  if (c->scene.slot) {
    hslot.resize((size_t)N);
    GSW_CUDA(cudaMemcpyAsync(hslot.data(), slot, (size_t)N * 4, cudaMemcpyDeviceToHost, s));
  }
  std::vector<float> hxyz((size_t)N * 3);
  GSW_CUDA(cudaMemcpyAsync(hxyz.data(), xyz, (size_t)N * 12, cudaMemcpyDeviceToHost, s));
  GSW_CUDA(cudaStreamSynchronize(s));
  float lo[3] = {3.0e38f, 3.0e38f, 3.0e38f}, hi[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
  for (int64_t i = 0; i < N; ++i) {
    if (c->scene.slot && hslot[i] >= num_slots) {
      set_error("gsw_scene_set: slot[%lld] = %d >= num_slots %d", (long long)i, hslot[i], num_slots);
      return GSW_E_INVALID;
    }
    for (int k = 0; k < 3; ++k) {
      const float v = hxyz[3 * i + k];
      if (v == v) {
        lo[k] = v < lo[k] ? v : lo[k];
        hi[k] = v > hi[k] ? v : hi[k];
      }
    }
  }
  auto spread = [](uint32_t v) {  // 10 bits -> every third bit
    v &= 1023u;
    v = (v | (v << 16)) & 0x030000FFu;
    v = (v | (v << 8)) & 0x0300F00Fu;
    v = (v | (v << 4)) & 0x030C30C3u;
    v = (v | (v << 2)) & 0x09249249u;
    return v;
  };
  auto morton = [&](int64_t i) -> uint32_t {
    uint32_t q[3];
    for (int k = 0; k < 3; ++k) {
      const float ext = hi[k] - lo[k];
      float t = ext > 0.f ? (hxyz[3 * i + k] - lo[k]) / ext : 0.f;
      if (!(t >= 0.f)) t = 0.f;
      if (t > 1.f) t = 1.f;
      q[k] = (uint32_t)(t * 1023.0f);
    }
    return spread(q[0]) | (spread(q[1]) << 1) | (spread(q[2]) << 2);
  };
  std::vector<std::pair<uint64_t, int32_t>> ks, km;  // (sort key, original index)
  ks.reserve((size_t)N);
  for (int64_t i = 0; i < N; ++i) {
    const int32_t sl = c->scene.slot ? hslot[i] : -1;
    if (sl < 0)
      ks.emplace_back((uint64_t)morton(i), (int32_t)i);
    else
      km.emplace_back(((uint64_t)(uint32_t)sl << 32) | morton(i), (int32_t)i);  // by link first, then by position
  }
  std::sort(ks.begin(), ks.end());
  std::sort(km.begin(), km.end());
  const int64_t Ns = (int64_t)ks.size(), Nm = (int64_t)km.size();
  std::vector<int32_t> sidx((size_t)Ns), midx((size_t)Nm), pos((size_t)N);
  for (int64_t j = 0; j < Ns; ++j) {
    sidx[j] = ks[j].second;
    pos[ks[j].second] = (int32_t)j;
  }
  for (int64_t m = 0; m < Nm; ++m) {
    midx[m] = km[m].second;
    pos[km[m].second] = (int32_t)m;
  }
  int rc;
  if ((rc = c->pos_of.reserve((size_t)N * 4))) return rc;
  GSW_CUDA(cudaMemcpyAsync(c->pos_of.p, pos.data(), (size_t)N * 4, cudaMemcpyHostToDevice, s));
  if (Ns > 0) {
    const size_t n = (size_t)Ns, ng = (size_t)((Ns + 255) / 256);
    if ((rc = c->st_idx.reserve(n * 4)) || (rc = c->st_xyz.reserve(n * 12)) || (rc = c->st_scale.reserve(n * 12)) ||
        (rc = c->st_quat.reserve(n * 16)) || (rc = c->st_opacity.reserve(n * 4)) ||
        (rc = c->st_bounds.reserve(ng * 32)))
      return rc;
    GSW_CUDA(cudaMemcpyAsync(c->st_idx.p, sidx.data(), n * 4, cudaMemcpyHostToDevice, s));
  }
  if (Nm > 0) {
    const size_t m = (size_t)Nm, ng = (size_t)((Nm + 255) / 256);
    if ((rc = c->mov_idx.reserve(m * 4)) || (rc = c->mov_xyz.reserve(m * 12)) ||
        (rc = c->mov_scale_act.reserve(m * 12)) || (rc = c->mov_opacity.reserve(m * 4)) ||
        (rc = c->mov_sh.reserve(m * 192)) || (rc = c->mov_log_scale.reserve(m * 12)) ||
        (rc = c->mov_quat_raw.reserve(m * 16)) || (rc = c->mov_slot.reserve(m * 4)) ||
        (rc = c->mov_bounds.reserve(ng * 32)))
      return rc;
    GSW_CUDA(cudaMemcpyAsync(c->mov_idx.p, midx.data(), m * 4, cudaMemcpyHostToDevice, s));
  }
  c->Ns = Ns;
  c->Nm = Nm;
  if ((rc = launch_gather_layers(c, s))) return rc;
  GSW_CUDA(cudaStreamSynchronize(s));  // host staging vectors go out of scope
  // the static layer the per-step launches consume: compact small arrays in Morton order, SH rows by original index
  GaussArrays st{};
  st.xyz = c->st_xyz.as<float>();
  st.scale_act = c->st_scale.as<float>();
  st.quat_act = c->st_quat.as<float>();
  st.opacity = c->st_opacity.as<float>();
  st.sh = sh;
  st.key_idx = c->st_idx.as<int32_t>();
  st.sh_idx = c->st_idx.as<int32_t>();
  st.bounds = c->st_bounds.as<float>();
  st.sh_degree = 3;
  st.M = 16;
  c->stat_layer = st;
  c->have_scene = true;
  return GSW_OK;
}

int gsw_scene_set_labels(gsw_ctx* c, const int32_t* labels) {
  if (!c) {
    set_error("gsw_scene_set_labels: NULL ctx");
    return GSW_E_INVALID;
  }
  c->labels = labels;
  c->scene_gen++;
  return GSW_OK;
}

int gsw_set_depth_buckets(gsw_ctx* c, int nb) {
  if (!c || nb < 0 || nb > 64 || (nb & (nb - 1))) {
    set_error("gsw_set_depth_buckets: nb must be 0 or a power of two <= 64");
    return GSW_E_INVALID;
  }
  c->nb_user = nb;
  return GSW_OK;
}

int gsw_set_tile_culling(gsw_ctx* c, int on) {
  if (!c) {
    set_error("gsw_set_tile_culling: NULL ctx");
    return GSW_E_INVALID;
  }
  c->tile_cull = on ? 1 : 0;
  return GSW_OK;
}

int gsw_reserve_entries(gsw_ctx* c, int64_t entries) {
  if (!c || entries < 0) {
    set_error("gsw_reserve_entries: invalid argument");
    return GSW_E_INVALID;
  }
  if (entries > c->cap_entries) c->cap_entries = entries;
  return GSW_OK;
}

int gsw_render_frames(gsw_ctx* c, const gsw_frames_desc* d, gsw_stream stream) {
  if (!c || !d) {
    set_error("gsw_render_frames: NULL argument");
    return GSW_E_INVALID;
  }
  if (!c->have_scene) {
    set_error("gsw_render_frames: call gsw_scene_set first");
    return GSW_E_STATE;
  }
  if (d->out_seg && !c->labels) {
    set_error("gsw_render_frames: out_seg needs per-Gaussian labels (gsw_scene_set_labels)");
    return GSW_E_STATE;
  }
  if (d->num_frames <= 0 || d->num_views <= 0 || d->num_envs <= 0 || d->height <= 0 || d->width <= 0 ||
      d->num_frames > 65535 || d->num_views > 65535) {
    set_error("gsw_render_frames: bad sizes (B=%d V=%d F=%d H=%d W=%d)", d->num_envs, d->num_views, d->num_frames,
              d->height, d->width);
    return GSW_E_INVALID;
  }
  if (!d->viewmatrix || !d->projmatrix || !d->campos || !d->tanfov || !d->frame_env || !d->frame_view ||
      (c->Nm > 0 && !d->pose)) {
    set_error("gsw_render_frames: missing camera / frame map / pose pointer");
    return GSW_E_INVALID;
  }
  // delivery arguments are checked before anything is enqueued
  const int n_peer = d->peer_rgb_u8 ? d->num_peer_rgb_u8 : 0;
  if (n_peer < 0 || n_peer > 16 || ((d->host_rgb_u8 || n_peer > 0) && !d->out_rgb_u8)) {
    set_error("gsw_render_frames: host_rgb_u8 / peer_rgb_u8 need out_rgb_u8; at most 16 peers");
    return GSW_E_INVALID;
  }
  Guard gd(c->device);
  cudaStream_t s = (cudaStream_t)stream;
  const int nb = c->nb_user > 0 ? c->nb_user : 1;
  FrameGeom g = make_geom(d->height, d->width, nb, d->antialiasing, d->scale_modifier, d->near_plane);
  g.tile_cull = c->tile_cull;

  GaussArrays dyn{};
  dyn.xyz = c->mov_xyz.as<float>();
  dyn.scale_act = c->mov_scale_act.as<float>();
  dyn.opacity = c->mov_opacity.as<float>();
  dyn.sh = c->mov_sh.as<float>();
  dyn.log_scale_raw = c->mov_log_scale.as<float>();
  dyn.quat_raw = c->mov_quat_raw.as<float>();
  dyn.mov_slot = c->mov_slot.as<int32_t>();
  dyn.key_idx = c->mov_idx.as<int32_t>();
  dyn.bounds = c->mov_bounds.as<float>();
  dyn.sh_degree = 3;
  dyn.M = 16;

  // static-tile sharing: when some view is looked through by several frames, its static-only tiles are
  // composited once (view pass) and copied into every frame that has no movable Gaussian in that tile.  Its
  // images are reserved before the pipeline runs: the static-list cache keys on where they live.
  const bool share = d->num_frames > d->num_views;
  int rc;
  if (share) {
    auto& w = c->ws;
    const size_t HW = (size_t)d->height * d->width, Vn = (size_t)d->num_views;
    if ((rc = w.st_u8.reserve(Vn * HW * 3)) ||
        (rc = w.st_color.reserve(d->out_color ? Vn * HW * 12 : 16)) ||
        (rc = w.st_invd.reserve(d->out_invdepth ? Vn * HW * 4 : 16)) ||
        (rc = w.st_seg.reserve(d->out_seg ? Vn * HW * 4 : 16)))
      return rc;
  }
  c->view_pass_outputs = (share ? 32 : 0) | (d->out_rgb_u8 ? 1 : 0) | (d->out_color ? 2 : 0) |
                         (d->out_invdepth ? 4 : 0) | (d->out_seg ? 8 : 0);
  for (int k = 0; k < 3; ++k) c->view_pass_bg[k] = d->bg[k];

  rc = run_pipeline(c, c->stat_layer, c->N, c->Ns, c->slot_orig, c->pos_of.as<int32_t>(), &dyn, c->Nm,
                        d->num_views, d->num_frames, g, d->viewmatrix,
                        d->projmatrix, d->campos, d->tanfov, d->frame_env, d->frame_view, d->pose, c->num_slots, s,
                        d->out_seg != nullptr);
  if (rc) return rc;
  c->last.single = false;
  if (share) {
    if ((rc = launch_blend_view_pass(c, d->num_views, d->num_frames, c->last.nbins, d->frame_view, c->last.cap, g,
                                     d->bg, d->out_rgb_u8 != nullptr, d->out_color != nullptr,
                                     d->out_invdepth != nullptr, d->out_seg != nullptr,
                                     c->side_pending ? c->side_stream : s)))
      return rc;
  }
  if ((rc = join_side(c, s))) return rc;
  const bool deliver = d->host_rgb_u8 != nullptr || n_peer > 0;
  // one group of finished frames to every destination (pinned host and / or peer GPUs), on the copy stream
  auto send_group = [&](int f0, int f1) -> int {
    const size_t frame_bytes = (size_t)d->height * d->width * 3;
    const size_t off = (size_t)f0 * frame_bytes, bytes = (size_t)(f1 - f0) * frame_bytes;
    if (d->host_rgb_u8)
      GSW_CUDA(cudaMemcpyAsync(d->host_rgb_u8 + off, d->out_rgb_u8 + off, bytes, cudaMemcpyDeviceToHost,
                               c->copy_stream));
    for (int p = 0; p < n_peer; ++p)
      if (d->peer_rgb_u8[p] != nullptr && d->peer_rgb_u8[p] != d->out_rgb_u8)
        GSW_CUDA(cudaMemcpyAsync(d->peer_rgb_u8[p] + off, d->out_rgb_u8 + off, bytes, cudaMemcpyDeviceToDevice,
                                 c->peer_stream));
    return 0;
  };
  // the copy stream's completion event covers both delivery streams
  auto close_delivery = [&]() -> int {
    c->delivered_peer = n_peer > 0;
    if (n_peer > 0) GSW_CUDA(cudaEventRecord(c->peer_evt, c->peer_stream));
    GSW_CUDA(cudaEventRecord(c->copy_evt, c->copy_stream));
    c->delivered = true;
    if (!(d->deliver_flags & GSW_DELIVER_NO_JOIN)) {  // `stream` order covers the delivery
      GSW_CUDA(cudaStreamWaitEvent(s, c->copy_evt, 0));
      if (n_peer > 0 && !(d->deliver_flags & GSW_DELIVER_NO_PEER_JOIN)) GSW_CUDA(cudaStreamWaitEvent(s, c->peer_evt, 0));
    }
    return 0;
  };
  if (deliver && (d->deliver_flags & GSW_DELIVER_AFTER_BLEND)) {
    // plain blend launch, then everything at once on the delivery streams (which run beside the next step's kernels)
    if ((rc = launch_blend(c, d->num_views, 0, d->num_frames, c->last.nbins, d->frame_view, c->last.cap, g, d->bg,
                           nullptr, d->out_rgb_u8, d->out_color, d->out_invdepth, d->out_seg, share, s)))
      return rc;
    GSW_CUDA(cudaEventRecord(c->grp_evt[0], s));
    GSW_CUDA(cudaStreamWaitEvent(c->copy_stream, c->grp_evt[0], 0));
    if (n_peer > 0) GSW_CUDA(cudaStreamWaitEvent(c->peer_stream, c->grp_evt[0], 0));
    if ((rc = send_group(0, d->num_frames))) return rc;
    if ((rc = close_delivery())) return rc;
  } else if (deliver && c->wait_value32 != nullptr) {
    // ONE blend launch; the frames are cut into groups, the last CTA of a group to finish publishes the launch
    // epoch in the group's flag, and the copy stream waits on that flag (stream memory operation) before it
    // sends the group's uint8 frames to the pinned host buffer and / or the peer GPUs.  Delivery overlaps the compositing of the later
    // frames without splitting the launch (every extra launch costs a tail of half-empty SMs).
    typedef CUresult (*wait_fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
    const wait_fn wait32 = reinterpret_cast<wait_fn>(c->wait_value32);
    DoneSig sg;
    {  // GSW_SIG_GROUPS: "u<N>" = N equal groups, "g<N>" = N groups of doubling size (tuning aid; default g5: 1.61 vs 1.63 ms e2e step for 8 equal groups)
      const char* e = getenv("GSW_SIG_GROUPS");
      int n = 5, geo = 1;
      if (e != nullptr && (e[0] == 'u' || e[0] == 'g') && atoi(e + 1) > 0) {
        geo = e[0] == 'g';
        n = atoi(e + 1);
      }
      if (n > kMaxSigGroups) n = kMaxSigGroups;
      if (geo && n > 8) n = 8;
      // every group must hold at least one frame (an empty group's flag would never be published)
      while (n > 1 && (geo ? (int64_t)d->num_frames < ((int64_t)1 << n) - 1 : d->num_frames < n)) --n;
      sg.n = n;
      sg.geo = geo && n > 1;
    }
    sg.cnt = c->sig.as<uint32_t>();
    sg.flag = c->sig.as<uint32_t>() + kMaxSigGroups;
    sg.epoch = ++c->sig_epoch;
    sg.F = d->num_frames;
    static const bool sig_times = getenv("GSW_SIG_TIMES") != nullptr;
    if (sig_times) sg.times = reinterpret_cast<unsigned long long*>(c->sig.as<uint32_t>() + 2 * kMaxSigGroups);
    GSW_CUDA(cudaMemsetAsync(sg.cnt, 0, kMaxSigGroups * sizeof(uint32_t), s));
    if ((rc = launch_blend(c, d->num_views, 0, d->num_frames, c->last.nbins, d->frame_view, c->last.cap, g, d->bg,
                           nullptr, d->out_rgb_u8, d->out_color, d->out_invdepth, d->out_seg, share, s, &sg)))
      return rc;
    for (int k = 0; k < sg.n; ++k) {
      if (sg.cut(k + 1) <= sg.cut(k)) continue;  // (cannot happen with the group counts chosen above)
      CUresult wr = CUDA_SUCCESS;
      if (c->wait_value32 != nullptr) {
        wr = wait32((CUstream)c->copy_stream, (CUdeviceptr)(uintptr_t)(sg.flag + k), sg.epoch,
                    CU_STREAM_WAIT_VALUE_GEQ);
        if (wr == CUDA_SUCCESS && n_peer > 0)
          wr = wait32((CUstream)c->peer_stream, (CUdeviceptr)(uintptr_t)(sg.flag + k), sg.epoch,
                      CU_STREAM_WAIT_VALUE_GEQ);
      }
      if (wr != CUDA_SUCCESS && c->wait_value32 != nullptr) {
        // not supported here (e.g. a restricted driver mode): order the remaining copies after the whole blend
        // launch instead, and use the grouped launches from the next call on
        c->wait_value32 = nullptr;
        GSW_CUDA(cudaEventRecord(c->grp_evt[0], s));
        GSW_CUDA(cudaStreamWaitEvent(c->copy_stream, c->grp_evt[0], 0));
        GSW_CUDA(cudaStreamWaitEvent(c->peer_stream, c->grp_evt[0], 0));
      }
      if ((rc = send_group(sg.cut(k), sg.cut(k + 1)))) return rc;
    }
    if ((rc = close_delivery())) return rc;
  } else if (deliver) {
    // blend in up to 5 frame groups; each group's uint8 frames go to the pinned host buffer on the copy stream
    // as soon as that group is done, while the next group is being composited.  The last groups are small:
    // only the final group's copy is not hidden behind a blend launch.
    static const int kCut[6] = {0, 32, 64, 96, 116, 128};  // of 128
    const int G = d->num_frames < 5 ? d->num_frames : 5;
    for (int k = 0; k < G; ++k) {
      const int f0 = G < 5 ? k : (int)((int64_t)d->num_frames * kCut[k] / 128);
      const int f1 = G < 5 ? k + 1 : (int)((int64_t)d->num_frames * kCut[k + 1] / 128);
      if (f1 <= f0) continue;
      if ((rc = launch_blend(c, d->num_views, f0, f1 - f0, c->last.nbins, d->frame_view, c->last.cap, g, d->bg,
                             nullptr, d->out_rgb_u8, d->out_color, d->out_invdepth, d->out_seg, share, s)))
        return rc;
      GSW_CUDA(cudaEventRecord(c->grp_evt[k], s));
      GSW_CUDA(cudaStreamWaitEvent(c->copy_stream, c->grp_evt[k], 0));
      if (n_peer > 0) GSW_CUDA(cudaStreamWaitEvent(c->peer_stream, c->grp_evt[k], 0));
      if ((rc = send_group(f0, f1))) return rc;
    }
    if ((rc = close_delivery())) return rc;
  } else if ((rc = launch_blend(c, d->num_views, 0, d->num_frames, c->last.nbins, d->frame_view, c->last.cap, g, d->bg,
                                nullptr, d->out_rgb_u8, d->out_color, d->out_invdepth, d->out_seg, share, s))) {
    return rc;
  }
  if (c->profile) cudaEventRecord(c->stage_evt[7], s);
  if (d->out_radii)
    if ((rc = launch_radii(c, d->out_radii, s))) return rc;
  if (c->profile) cudaEventRecord(c->stage_evt[8], s);
  return finish_launch(c, s);
}

int gsw_delivery_wait(gsw_ctx* c, gsw_stream stream) {
  if (!c) {
    set_error("gsw_delivery_wait: NULL ctx");
    return GSW_E_INVALID;
  }
  if (!c->delivered) return GSW_OK;
  Guard gd(c->device);
  GSW_CUDA(cudaStreamWaitEvent((cudaStream_t)stream, c->copy_evt, 0));
  if (c->delivered_peer) GSW_CUDA(cudaStreamWaitEvent((cudaStream_t)stream, c->peer_evt, 0));
  return GSW_OK;
}

int gsw_delivery_gate(gsw_ctx* c, gsw_stream stream) {
  if (!c) {
    set_error("gsw_delivery_gate: NULL ctx");
    return GSW_E_INVALID;
  }
  Guard gd(c->device);
  GSW_CUDA(cudaEventRecord(c->gate_evt, (cudaStream_t)stream));
  GSW_CUDA(cudaStreamWaitEvent(c->peer_stream, c->gate_evt, 0));  // (copies to this process' host memory need no gate)
  return GSW_OK;
}

int gsw_finish(gsw_ctx* c, int64_t* needed_entries, int64_t stats[4]) {
  if (!c) {
    set_error("gsw_finish: NULL ctx");
    return GSW_E_INVALID;
  }
  Guard gd(c->device);
  GSW_CUDA(cudaEventSynchronize(c->done_evt));
  if (getenv("GSW_SIG_TIMES") != nullptr && c->sig.p != nullptr) {  // tuning aid: when each delivery group completed
    unsigned long long t[17];
    if (cudaMemcpy(t, c->sig.as<uint32_t>() + 2 * kMaxSigGroups, sizeof t, cudaMemcpyDeviceToHost) == cudaSuccess) {
      fprintf(stderr, "[gsw] delivery groups completed at (us after the blend's first CTA):");
      for (int k = 0; k < 16 && t[k] >= t[16] && t[k] != 0; ++k) fprintf(stderr, " %.0f", (double)(t[k] - t[16]) * 1e-3);
      fprintf(stderr, "\n");
    }
  }
  const int64_t D = c->h_status[0], M = 0, ovf = c->h_status[2];
  if (needed_entries) *needed_entries = D + M;
  if (stats) {
    stats[0] = D;
    stats[1] = c->h_status[3];
    stats[2] = c->cap_entries;
    stats[3] = c->launches;
  }
  if (ovf) {
    set_error("tile-entry workspace overflow: need %lld entries, capacity %lld", (long long)(D + M),
              (long long)c->cap_entries);
    return GSW_E_CAPACITY;
  }
  return GSW_OK;
}

int gsw_rig_set(gsw_ctx* c, int L, int K, const float* S, const float* S_inv, const float* scan_inv,
                const float* link_offset, const float* obj_S_inv, const float* obj_offset, const float* obj_scale) {
  if (!c || L < 0 || K < 0 || L + K <= 0 || L + K > 512 || !S || !S_inv || (L > 0 && !scan_inv) ||
      (K > 0 && (!obj_S_inv || !obj_offset || !obj_scale))) {
    set_error("gsw_rig_set: invalid argument");
    return GSW_E_INVALID;
  }
  Guard gd(c->device);
  memcpy(c->rig.S, S, 64);
  memcpy(c->rig.S_inv, S_inv, 64);
  for (int k = 0; k < 3; ++k) c->rig.link_off[k] = link_offset ? link_offset[k] : 0.f;
  c->rig.L = L;
  c->rig.K = K;
  std::vector<float> h((size_t)L * 16 + (size_t)K * 20);
  if (L) memcpy(h.data(), scan_inv, (size_t)L * 64);
  if (K) {
    memcpy(h.data() + (size_t)L * 16, obj_S_inv, (size_t)K * 64);
    memcpy(h.data() + (size_t)L * 16 + (size_t)K * 16, obj_offset, (size_t)K * 12);
    memcpy(h.data() + (size_t)L * 16 + (size_t)K * 19, obj_scale, (size_t)K * 4);
  }
  int rc = c->rig.dev.reserve(h.size() * 4);
  if (rc) return rc;
  GSW_CUDA(cudaMemcpy(c->rig.dev.p, h.data(), h.size() * 4, cudaMemcpyHostToDevice));
  c->rig.valid = true;
  return GSW_OK;
}

int gsw_compose_poses(gsw_ctx* c, int B, const float* link_mats, const float* obj_pose, int obj_stride,
                      float* pose_out, gsw_stream stream) {
  if (!c || !c->rig.valid) {
    set_error("gsw_compose_poses: call gsw_rig_set first");
    return GSW_E_STATE;
  }
  if (B <= 0 || !pose_out || (c->rig.L > 0 && !link_mats) || (c->rig.K > 0 && (!obj_pose || obj_stride < 7))) {
    set_error("gsw_compose_poses: invalid argument");
    return GSW_E_INVALID;
  }
  Guard gd(c->device);
  return launch_compose_poses(c, B, link_mats, obj_pose, obj_stride, pose_out, (cudaStream_t)stream);
}

int gsw_compose_cameras(gsw_ctx* c, int n, const float* extrinsic_cv, const float* intrinsic_cv, int height,
                        int width, const float* rigid_sim2real, float scale_sim2real, float* viewmatrix,
                        float* projmatrix, float* campos, float* tanfov, gsw_stream stream) {
  if (!c || n < 0 || height <= 0 || width <= 0 || !rigid_sim2real ||
      (n > 0 && (!extrinsic_cv || !intrinsic_cv || !viewmatrix || !projmatrix || !campos || !tanfov))) {
    set_error("gsw_compose_cameras: invalid argument");
    return GSW_E_INVALID;
  }
  Guard gd(c->device);
  return launch_compose_cameras(n, extrinsic_cv, intrinsic_cv, height, width, rigid_sim2real, scale_sim2real,
                                viewmatrix, projmatrix, campos, tanfov, (cudaStream_t)stream);
}

long long gsw_kernel_launches(void) { return g_launches.load(std::memory_order_relaxed); }

int gsw_profile_enable(gsw_ctx* c, int on) {
  if (!c) {
    set_error("gsw_profile_enable: NULL ctx");
    return GSW_E_INVALID;
  }
  c->profile = on != 0;
  return GSW_OK;
}

int gsw_profile_counters(gsw_ctx* c, int64_t* out, int n) {
  if (!c || !out || n < 2) {
    set_error("gsw_profile_counters: need room for 2 counters");
    return GSW_E_INVALID;
  }
  Guard gd(c->device);
  GSW_CUDA(cudaEventSynchronize(c->done_evt));
  out[0] = c->h_status[4];  // (record, pixel) pair evaluations of the blend (incl. the view pass)
  out[1] = c->h_status[3];  // entries staged by the blend
  if (n >= 3) out[2] = (int64_t)(uint32_t)c->h_status[10];  // live (view, chunk) pairs of the static layer's work list
  if (n >= 5) {
    out[3] = c->h_status[16];  // visible static Gaussian-views (records written by the static preprocess)
    out[4] = c->h_status[17];  // visible movable Gaussian-frames
  }
  if (n >= 6) out[5] = c->ipb_static;  // items per chunk of the static layer in that launch
  return GSW_OK;
}

int gsw_profile_read(gsw_ctx* c, float* stage_ms, int n) {
  if (!c || !stage_ms || n < kNumStages) {
    set_error("gsw_profile_read: need room for %d stages", kNumStages);
    return GSW_E_INVALID;
  }
  if (!c->profile) {
    set_error("gsw_profile_read: profiling is off (gsw_profile_enable)");
    return GSW_E_STATE;
  }
  Guard gd(c->device);
  GSW_CUDA(cudaEventSynchronize(c->stage_evt[kNumStages]));
  for (int k = 0; k < kNumStages; ++k) GSW_CUDA(cudaEventElapsedTime(&stage_ms[k], c->stage_evt[k], c->stage_evt[k + 1]));
  return GSW_OK;
}

int gsw_rasterize_forward(gsw_ctx* c, int32_t P, int32_t sh_degree, int32_t M, const float* bg, int32_t width,
                          int32_t height, const float* means3D, const float* shs, const float* colors_precomp,
                          const float* opacities, const float* scales, float scale_modifier, const float* rotations,
                          const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                          const float* campos, float tan_fovx, float tan_fovy, int32_t prefiltered,
                          int32_t antialiasing, float near_plane, float* out_color, int32_t* out_radii,
                          float* out_invdepth, gsw_stream stream) {
  if (!c || P < 0 || width <= 0 || height <= 0 || !bg || !viewmatrix || !projmatrix || !campos) {
    set_error("gsw_rasterize_forward: invalid argument");
    return GSW_E_INVALID;
  }
  if (P == 0) {  // upstream leaves the zero-initialised outputs untouched when there is nothing to draw
    Guard gd0(c->device);
    const size_t HW0 = (size_t)height * width;
    if (out_color) GSW_CUDA(cudaMemsetAsync(out_color, 0, 3 * HW0 * 4, (cudaStream_t)stream));
    if (out_invdepth) GSW_CUDA(cudaMemsetAsync(out_invdepth, 0, HW0 * 4, (cudaStream_t)stream));
    c->last.valid = false;
    return GSW_OK;
  }
  if ((shs == nullptr) == (colors_precomp == nullptr)) {
    set_error("Please provide excatly one of either SHs or precomputed colors!");
    return GSW_E_INVALID;
  }
  if (((scales == nullptr || rotations == nullptr) && cov3D_precomp == nullptr) ||
      ((scales != nullptr || rotations != nullptr) && cov3D_precomp != nullptr)) {
    set_error("Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!");
    return GSW_E_INVALID;
  }
  if (shs && (sh_degree < 0 || sh_degree > 3 || M < (sh_degree + 1) * (sh_degree + 1))) {
    set_error("gsw_rasterize_forward: sh_degree %d needs M >= %d (got %d); degree <= 3 supported", sh_degree,
              (sh_degree + 1) * (sh_degree + 1), M);
    return GSW_E_INVALID;
  }
  (void)prefiltered;
  Guard gd(c->device);
  cudaStream_t s = (cudaStream_t)stream;
  const int nb = c->nb_user > 0 ? c->nb_user : 1;
  FrameGeom g = make_geom(height, width, nb, antialiasing, scale_modifier, near_plane);
  g.tile_cull = c->tile_cull;
  GaussArrays st{};
  st.xyz = means3D;
  st.scale_act = scales;
  st.quat_act = rotations;
  st.opacity = opacities;
  st.sh = shs;
  st.colors_precomp = colors_precomp;
  st.cov3d_precomp = cov3D_precomp;
  st.slot = nullptr;
  st.sh_degree = sh_degree;
  st.M = M;
  // tan(fov) and the trivial frame maps live in a small staging block: [tanx, tany, 0(int), 0(int)]
  int rc;
  if ((rc = c->ws.cam.reserve(64))) return rc;
  if ((rc = stage_single_cam(c->ws.cam, tan_fovx, tan_fovy, s))) return rc;
  const float* tanfov = c->ws.cam.as<float>();
  const int32_t* fe = c->ws.cam.as<int32_t>() + 2;
  const int32_t* fv = c->ws.cam.as<int32_t>() + 3;
  rc = run_pipeline(c, st, P, P, nullptr, nullptr, nullptr, 0, 1, 1, g, viewmatrix, projmatrix, campos, tanfov, fe, fv,
                    nullptr, 0, s);
  if (rc) return rc;
  c->last.single = true;
  if ((rc = launch_blend(c, 1, 0, 1, c->last.nbins, fv, c->last.cap, g, nullptr, bg, nullptr, out_color, out_invdepth, nullptr, false, s)))
    return rc;
  if (c->profile) cudaEventRecord(c->stage_evt[7], s);
  if (out_radii)
    if ((rc = launch_radii(c, out_radii, s))) return rc;
  if (c->profile) cudaEventRecord(c->stage_evt[8], s);
  return finish_launch(c, s);
}

// ---------------------------------------------------------------------------------------------------
// gsw_rasterize_forward_static: the single-frame forward for callers that must not allocate.
//
// Who needs this: a render call made from inside a control loop with a hard latency bound (cudaMalloc /
// cudaFree can stall behind every other stream of the process), or from a plugin that shares the GPU
// with a simulator and has been handed a fixed memory budget it may not grow, or from a process whose
// allocator is owned by someone else (an embedding runtime that forbids raw cudaMalloc).  Those callers
// cannot use gsw_rasterize_forward, because that entry point sizes and (re)allocates a workspace in its
// context whenever a shape grows.  Here every byte of scratch (projected records, tile bins, sort keys,
// blend stream) lives in ONE fixed array compiled into the library, so the call allocates nothing and
// frees nothing; the price is a hard size limit and ~17 MB of device memory held on every GPU that loads
// libgsw.so, until the process exits.
//
// Limits (rejected with GSW_E_INVALID and a message, nothing is launched): P <= 32768 Gaussians,
// ceil(W/16)*ceil(H/16) <= 1024 tiles (e.g. 512x512), and at most 262144 (Gaussian, tile) entries; the
// entry count depends on the data, so it cannot be checked on the host: if a frame needs more, the
// image is left as background and *out_status (a caller-provided device int64[4], may be NULL) reports
// it: [0] = entries needed, [2] != 0 on overflow.
//
// Sharing: all calls on one device use the same array and it keeps whatever the previous call left.
// Everything that must start at zero (bin counters, cursors, status words, queue counters) is cleared on
// the stream at the start of every call.  Two calls whose GPU work overlaps (different streams, or
// different host threads on the same device) would race on the array: serialise them.
//
// The first call on each device looks the array's device address up (cudaGetSymbolAddress needs that
// device to be current; it is, this function uses the current device) and caches it.  Make that first
// call before starting any stream capture.
namespace {

constexpr int32_t kNaMaxP = 32768;
constexpr int32_t kNaMaxTiles = 1024;
constexpr int64_t kNaMaxEntries = 262144;
constexpr size_t kNaBytes = (size_t)17 << 20;

// In host code this name is only a handle for cudaGetSymbolAddress, never a device pointer.
static __device__ __align__(256) unsigned char g_na_scratch[kNaBytes];

constexpr int kNaMaxDevices = 64;
struct NaDevice {
  bool ready = false;
  gsw_ctx ctx;  // workspace DevBufs point into g_na_scratch; nothing in it is ever allocated or freed
};
NaDevice g_na_dev[kNaMaxDevices];
std::mutex g_na_mutex;

// Carve the fixed array into the workspace slices run_pipeline expects, sized for the maxima above.
int na_setup(int device, NaDevice& nd) {
  void* basep = nullptr;
  GSW_CUDA(cudaGetSymbolAddress(&basep, g_na_scratch));  // device `device` is current (checked by the caller)
  unsigned char* base = static_cast<unsigned char*>(basep);
  size_t off = 0;
  auto carve = [&](DevBuf& b, size_t bytes) {
    bytes = (bytes + 255) & ~(size_t)255;
    b.p = base + off;
    b.bytes = bytes;
    b.fixed = true;  // reserve() beyond the slice fails with GSW_E_NOMEM instead of freeing / allocating
    off += bytes;
  };
  gsw_ctx& c = nd.ctx;
  c.device = device;
  const int64_t nbins = 2 * (int64_t)kNaMaxTiles;  // one static + one (empty) dynamic list, NB = 1
  auto& w = c.ws;
  carve(w.rec, (size_t)kNaMaxP * 64);
  carve(w.rad, (size_t)kNaMaxP * 4);
  carve(w.bin_count, (size_t)(nbins + 1) * 4);
  carve(w.bin_off, (size_t)(nbins + 1) * 4);
  carve(w.bin_cursor, (size_t)(nbins + 1) * 4);
  carve(w.scan_tmp, scan_tmp_elems(nbins) * 4);
  carve(w.merge_cnt, (size_t)(kNaMaxTiles + 1) * 4);
  carve(w.merge_off, (size_t)(kNaMaxTiles + 1) * 4);
  carve(w.keys, (size_t)kNaMaxEntries * 8);
  carve(w.splats, (size_t)kNaMaxEntries * sizeof(Splat));
  carve(w.status, kStatusBytes);
  carve(w.seg_queue, 2 * seg_queue_bytes(kNaMaxEntries, nbins));
  carve(w.cam, 64);
  carve(w.view_cnt, 64);  // one view
  {  // the largest per-CTA histogram block run_pipeline can ask for within the limits: kNaMaxP items, kNaMaxTiles tiles
    FrameGeom gmax{};
    gmax.T = kNaMaxTiles;
    gmax.NB = 1;
    carve(w.cta_hist, cta_hist_words(&nd.ctx, kNaMaxP, 1, gmax, false) * 4 + 16);  // (2048-item chunks: no work list here)
  }
  carve(w.chunk_mask, (size_t)layer_chunks(&nd.ctx, kNaMaxP, false) * 4 + 16);
  carve(w.work, (size_t)layer_chunks(&nd.ctx, kNaMaxP, false) * 4 + 16);
  carve(w.view_consts, 48 * 4);
  if (off > kNaBytes) {
    set_error("gsw_rasterize_forward_static: internal layout (%zu B) exceeds the fixed array (%zu B)", off, kNaBytes);
    return GSW_E_NOMEM;
  }
  c.cap_entries = kNaMaxEntries;
  c.nb_user = 1;
  nd.ready = true;
  return GSW_OK;
}

}  // namespace

int gsw_rasterize_forward_static(int32_t P, int32_t sh_degree, int32_t M, const float* bg, int32_t width,
                                 int32_t height, const float* means3D, const float* shs, const float* colors_precomp,
                                 const float* opacities, const float* scales, float scale_modifier,
                                 const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                                 const float* projmatrix, const float* campos, float tan_fovx, float tan_fovy,
                                 int32_t antialiasing, float near_plane, float* out_color, int32_t* out_radii,
                                 float* out_invdepth, int64_t* out_status, gsw_stream stream) {
  if (P < 0 || width <= 0 || height <= 0 || !bg || !viewmatrix || !projmatrix || !campos) {
    set_error("gsw_rasterize_forward_static: invalid argument");
    return GSW_E_INVALID;
  }
  const int64_t tiles = (int64_t)((width + kTile - 1) / kTile) * ((height + kTile - 1) / kTile);
  if (P > kNaMaxP || tiles > kNaMaxTiles) {
    set_error("gsw_rasterize_forward_static: P=%d (max %d) or %lld tiles for %dx%d (max %d) exceeds the fixed "
              "scratch array; use gsw_rasterize_forward for larger frames", P, kNaMaxP, (long long)tiles, width,
              height, kNaMaxTiles);
    return GSW_E_INVALID;
  }
  cudaStream_t s = (cudaStream_t)stream;
  const size_t HW = (size_t)height * width;
  if (P == 0) {  // same convention as gsw_rasterize_forward: nothing to draw leaves zeros
    if (out_color) GSW_CUDA(cudaMemsetAsync(out_color, 0, 3 * HW * 4, s));
    if (out_invdepth) GSW_CUDA(cudaMemsetAsync(out_invdepth, 0, HW * 4, s));
    if (out_status) GSW_CUDA(cudaMemsetAsync(out_status, 0, 4 * sizeof(int64_t), s));
    return GSW_OK;
  }
  if ((shs == nullptr) == (colors_precomp == nullptr)) {
    set_error("Please provide excatly one of either SHs or precomputed colors!");
    return GSW_E_INVALID;
  }
  if (((scales == nullptr || rotations == nullptr) && cov3D_precomp == nullptr) ||
      ((scales != nullptr || rotations != nullptr) && cov3D_precomp != nullptr)) {
    set_error("Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!");
    return GSW_E_INVALID;
  }
  if (shs && (sh_degree < 0 || sh_degree > 3 || M < (sh_degree + 1) * (sh_degree + 1))) {
    set_error("gsw_rasterize_forward_static: sh_degree %d needs M >= %d (got %d); degree <= 3 supported", sh_degree,
              (sh_degree + 1) * (sh_degree + 1), M);
    return GSW_E_INVALID;
  }
  int device = -1;
  GSW_CUDA(cudaGetDevice(&device));
  if (device < 0 || device >= kNaMaxDevices) {
    set_error("gsw_rasterize_forward_static: device ordinal %d not supported (max %d)", device, kNaMaxDevices - 1);
    return GSW_E_INVALID;
  }
  NaDevice& nd = g_na_dev[device];
  {
    std::lock_guard<std::mutex> lock(g_na_mutex);
    if (!nd.ready) {  // once per device, with that device current
      int rc0 = na_setup(device, nd);
      if (rc0) return rc0;
    }
  }
  gsw_ctx* c = &nd.ctx;
  const FrameGeom g = make_geom(height, width, 1, antialiasing, scale_modifier, near_plane);
  GaussArrays st{};
  st.xyz = means3D;
  st.scale_act = scales;
  st.quat_act = rotations;
  st.opacity = opacities;
  st.sh = shs;
  st.colors_precomp = colors_precomp;
  st.cov3d_precomp = cov3D_precomp;
  st.slot = nullptr;
  st.sh_degree = sh_degree;
  st.M = M;
  int rc;
  if ((rc = stage_single_cam(c->ws.cam, tan_fovx, tan_fovy, s))) return rc;
  const float* tanfov = c->ws.cam.as<float>();
  const int32_t* fe = c->ws.cam.as<int32_t>() + 2;
  const int32_t* fv = c->ws.cam.as<int32_t>() + 3;
  // Every workspace request below fits its slice (sizes were checked against the maxima above), so the
  // grow-only buffers never reach their allocation branch; the zero-initialised state is set by
  // run_pipeline's memsets on the stream.
  rc = run_pipeline(c, st, P, P, nullptr, nullptr, nullptr, 0, 1, 1, g, viewmatrix, projmatrix, campos, tanfov, fe, fv,
                    nullptr, 0, s);
  if (rc) return rc;
  c->last.single = true;
  if ((rc = launch_blend(c, 1, 0, 1, c->last.nbins, fv, c->last.cap, g, nullptr, bg, nullptr, out_color, out_invdepth, nullptr, false, s)))
    return rc;
  if (out_radii)
    if ((rc = launch_radii(c, out_radii, s))) return rc;
  if (out_status)
    GSW_CUDA(cudaMemcpyAsync(out_status, c->ws.status.p, 4 * sizeof(int64_t), cudaMemcpyDeviceToDevice, s));
  return GSW_OK;
}

int gsw_mark_visible(gsw_ctx* c, int32_t P, const float* means3D, const float* viewmatrix, float near_plane,
                     uint8_t* out_visible, gsw_stream stream) {
  if (!c || P < 0 || (P > 0 && (!means3D || !out_visible)) || !viewmatrix) {
    set_error("gsw_mark_visible: invalid argument");
    return GSW_E_INVALID;
  }
  Guard gd(c->device);
  return launch_mark_visible(P, means3D, viewmatrix, near_plane, out_visible, (cudaStream_t)stream);
}

int gsw_debug_frame_attrs(gsw_ctx* c, int32_t frame, int32_t* radii, int32_t* tiles_touched, float* depth, float* xy,
                          float* conic_opacity, float* rgb, gsw_stream stream) {
  if (!c || !c->last.valid || frame < 0 || frame >= c->last.F) {
    set_error("gsw_debug_frame_attrs: no valid launch / frame out of range");
    return GSW_E_STATE;
  }
  Guard gd(c->device);
  return launch_frame_attrs(c, frame, radii, tiles_touched, depth, xy, conic_opacity, rgb, (cudaStream_t)stream);
}

int gsw_debug_frame_lists(gsw_ctx* c, int32_t frame, int32_t* ranges, uint64_t* keys, uint32_t* vals, int64_t cap,
                          int64_t* count, gsw_stream stream) {
  if (!c || !c->last.valid || frame < 0 || frame >= c->last.F || !ranges || !keys || !vals) {
    set_error("gsw_debug_frame_lists: no valid launch / bad argument");
    return GSW_E_STATE;
  }
  Guard gd(c->device);
  cudaStream_t s = (cudaStream_t)stream;
  int rc = launch_frame_lists(c, frame, ranges, keys, vals, cap, s);
  if (rc) return rc;
  uint32_t total = 0;
  GSW_CUDA(cudaMemcpyAsync(&total, c->ws.merge_off.as<uint32_t>() + c->last.g.T, 4, cudaMemcpyDeviceToHost, s));
  GSW_CUDA(cudaStreamSynchronize(s));
  if (count) *count = total;
  if ((int64_t)total > cap) {
    set_error("gsw_debug_frame_lists: need %u entries, cap %lld", total, (long long)cap);
    return GSW_E_CAPACITY;
  }
  return GSW_OK;
}

int gsw_debug_bin_offsets(gsw_ctx* c, uint32_t* dst_host, int64_t max_bins, int64_t* nbins, gsw_stream stream) {
  if (!c || !c->last.valid || !dst_host) {
    set_error("gsw_debug_bin_offsets: no valid launch / bad argument");
    return GSW_E_STATE;
  }
  if (nbins) *nbins = c->last.nbins;
  if (max_bins < c->last.nbins) {
    set_error("gsw_debug_bin_offsets: need room for %lld bins", (long long)c->last.nbins);
    return GSW_E_CAPACITY;
  }
  Guard gd(c->device);
  cudaStream_t s = (cudaStream_t)stream;
  GSW_CUDA(cudaMemcpyAsync(dst_host, c->ws.bin_off.p, (size_t)(c->last.nbins + 1) * 4, cudaMemcpyDeviceToHost, s));
  GSW_CUDA(cudaStreamSynchronize(s));
  return GSW_OK;
}

}  // extern "C"
