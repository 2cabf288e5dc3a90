// blend.cu — 16x16-tile front-to-back alpha compositing over the sorted splat streams.  sm_100a.
//
// One CTA per (frame, tile); 8 warps, each owning an 8x4 pixel block.  A frame's tile list is the merge
// of two sorted streams: the STATIC list of the frame's view (shared by every env that looks through
// that camera) and the frame's own DYNAMIC list (robot links / objects of that env).  Chunks of up to
// 256 records are staged in shared memory; where both streams are non-empty the chunk is produced by a
// merge-path step over two 256-key windows (keys = depth bits << 32 | Gaussian index, unique), so the
// blend order is exactly the reference's single sorted list.  Every warp then culls 32 staged records at
// a time against its own pixel block (one record per lane, one ballot) and runs the per-pixel blend only
// for the survivors, in list order.  Skipped records are exactly those whose alpha would be < 1/255 on
// all 32 pixels, which the reference skips too.  Warps whose 32 pixels have all saturated stop
// evaluating; the CTA leaves when every warp has.
//
// Reference arithmetic: diff-gaussian-rasterization@9c5c2028 forward.cu renderCUDA (absent submodule):
//   power = -0.5(A dx^2 + C dy^2) - B dx dy; skip if > 0; alpha = min(0.99, o*exp(power)); skip if
//   < 1/255; stop when T(1-alpha) < 1e-4; C += rgb*alpha*T; invD += alpha*T/depth; out = C + T*bg.
// Here exp(power) is evaluated as 2^(power*log2 e) with the conic pre-scaled at gather time (binsort.cu);
// colour / inverse depth are tolerance quantities (max-abs 1e-3).
// uint8 epilogue: gs_world_wrapper.py:268-270 after render()'s clamp(0,1) (gaussian_renderer/__init__.py:119).
#include <cstdlib>

#include "gsw_internal.h"

namespace gsw {
namespace {

constexpr int kChunk = 256;

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float4 lds128(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ float2 lds64(uint32_t addr) {
  float2 v;
  asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(addr));
  return v;
}

// keeps a shared-memory byte address in a register (the compiler otherwise re-derives it every use)
__device__ __forceinline__ uint32_t pin(uint32_t x) {
  asm volatile("" : "+r"(x));
  return x;
}
__device__ __forceinline__ uint32_t lds_u8(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts_u8(uint32_t addr, uint32_t v) {
  asm volatile("st.shared.u8 [%0], %1;" ::"r"(addr), "r"(v));
}

// ---- packed fp32x2 arithmetic (sm_100 FFMA2 / FMUL2 / FADD2): one issue slot for two pixels.  An operand built with
// bc() (both halves the same scalar) costs no extra instruction: ptxas uses the single-register broadcast form.
typedef unsigned long long f2;
__device__ __forceinline__ f2 pk(float lo, float hi) {
  f2 r;
  asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ f2 bc(float v) { return pk(v, v); }
__device__ __forceinline__ float lo_of(f2 v) {
  float a;
  asm("mov.b64 {%0,_}, %1;" : "=f"(a) : "l"(v));
  return a;
}
__device__ __forceinline__ float hi_of(f2 v) {
  float b;
  asm("mov.b64 {_,%0}, %1;" : "=f"(b) : "l"(v));
  return b;
}
__device__ __forceinline__ f2 fma2(f2 a, f2 b, f2 c) {
  f2 r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ f2 mul2(f2 a, f2 b) {
  f2 r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f2 add2(f2 a, f2 b) {
  f2 r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}

// alpha if the reference would composite it (power <= 0 and alpha >= 1/255, NaNs as the reference's comparisons treat
// them), else 0: two compares chained through the predicate and one select
__device__ __forceinline__ float keep_alpha(float alpha, float power) {
  float r;
  asm("{ .reg .pred q, p;\n\t"
      "setp.gt.f32 q, %2, 0f00000000;\n\t"
      "setp.geu.and.f32 p, %1, 0f3B808081, !q;\n\t"
      "selp.f32 %0, %1, 0f00000000, p; }"
      : "=f"(r)
      : "f"(alpha), "f"(power));
  return r;
}

// PPT = pixels per thread.  1: 256 threads, a warp owns an 8x4 pixel block.  2: 128 threads, a warp owns an 8x8
// block and every lane the pixels (x, y) and (x, y + 4); the per-pixel arithmetic then runs on packed fp32x2
// instructions and the record loads / queue reads are paid once for two pixels.
template <bool SEG, int PPT>
__device__ __forceinline__ void
blend_body(const uint32_t* __restrict__ bin_off, int64_t nbins, const int32_t* __restrict__ frame_view, int32_t V,
             int32_t f_begin,
             const unsigned long long* __restrict__ keys, const Splat* __restrict__ splats, int64_t cap, FrameGeom g,
             float bg0, float bg1, float bg2, const float* __restrict__ bg_dev, uint8_t* __restrict__ out_u8,
             float* __restrict__ out_color, float* __restrict__ out_invd, long long* __restrict__ status,
             int mode, const uint32_t* __restrict__ view_cnt, const uint8_t* __restrict__ st_u8,
             const float* __restrict__ st_color, const float* __restrict__ st_invd,
             const int32_t* __restrict__ lab_stream, int32_t* __restrict__ out_seg, const int32_t* __restrict__ st_seg,
             int use_view_images, const float4* __restrict__ rec, const int32_t* __restrict__ pos_of, int64_t Ns,
             const int32_t* __restrict__ labels, int64_t Nm, int tma_stage, const uint32_t* __restrict__ cache_m) {
  // Only static lists of views shared by several frames are materialised as a splat stream
  // (gather_splats_kernel); for every other list this kernel fetches the records of the entries it actually
  // reaches (about 70 % of them at cfg3, a quarter with per-env cameras) from the per-item records through
  // the sorted keys: key -> Gaussian index -> position in its layer -> 64 B record -> make_splat().
  // SEG: additionally reports, per pixel, the label of the Gaussian with the largest blending weight alpha*T
  // (GSW_SEG_NONE where nothing was composited); lab_stream holds the label of every stream entry.
  // mode 0: frames.  mode 1: "view pass": blockIdx enumerates (view, tile); renders the static list of every
  // view that is shared by >= 2 frames into the st_* images (passed as out_*), once.  In mode 0 a tile whose
  // frame has no dynamic entries there is then a plain copy of the view's tile: same list, same code, same
  // bits, computed once instead of once per env.
  __shared__ __align__(128) float4 sRec[kChunk * 3];   // staged records, 48 B each: {a, b, c}
  __shared__ __align__(8) unsigned long long mbar_s;   // completion of the bulk copies (tma_stage)
  __shared__ unsigned long long kS[kChunk], kD[kChunk];  // key windows of the two streams (merged tiles only)
  constexpr int NT = 256 / PPT, NW = 8 / PPT;          // threads, warps per CTA
  __shared__ uint8_t sQ[NW][kChunk];                   // per-warp queue of staged records that touch its pixels
  __shared__ int32_t sLab[SEG ? kChunk : 1];
  __shared__ uint32_t sPairs[8];  // per warp: records it ran the per-pixel blend for (x 32 pixels = pair evaluations)
  // 32-bit unsigned division (the grid is f-major: blockIdx.x = (f - f_begin) * T + tile)
  const uint32_t fl = blockIdx.x / (uint32_t)g.T;
  const int f = f_begin + (int)fl, tile = (int)(blockIdx.x - fl * (uint32_t)g.T);
  const int tx = tile % g.gx, ty = tile / g.gx;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int wx = tx * kTile + (w & 1) * 8, wy = ty * kTile + (w >> 1) * 4 * PPT;
  const int px = wx + (lane & 7), py = wy + (lane >> 3);  // PPT == 2: second pixel at (px, py + 4)
  const bool inside = px < g.W && py < g.H;
  const bool inside1 = PPT == 2 && px < g.W && py + 4 < g.H;
  const float pxf = (float)px, pyf = (float)py;
  const float bx0 = (float)wx, bx1 = (float)(wx + 7), by0 = (float)wy, by1 = (float)(wy + 4 * PPT - 1);

  const int v = mode == 1 ? f : (frame_view ? frame_view[f] : 0);
  if (mode == 1 && view_cnt[v] < 2u) return;  // view used by a single frame: nothing to share
  if (mode == 1 && cache_m != nullptr && (uint32_t)v < *cache_m) return;  // static-only tiles kept from the last step
  // bin indices fit 32 bits (run_pipeline rejects launches with 2^31 bins or more)
  const uint32_t sb = ((uint32_t)v * (uint32_t)g.T + (uint32_t)tile) * (uint32_t)g.NB;
  const uint32_t db = ((uint32_t)(V + f) * (uint32_t)g.T + (uint32_t)tile) * (uint32_t)g.NB;
  const uint32_t s0 = bin_off[sb], d0 = mode == 1 ? 0u : bin_off[db];
  uint32_t ns = bin_off[sb + g.NB] - s0, nd = mode == 1 ? 0u : bin_off[db + g.NB] - d0;
  const uint32_t D = bin_off[nbins];
  if ((int64_t)D > cap) ns = nd = 0;  // workspace overflow: lists are incomplete; the launch is flagged invalid
  if (mode == 0 && f == 0 && tile == 0 && threadIdx.x == 0) {
    status[0] = (long long)D;
    if ((int64_t)D > cap) {
      status[2] = 1;
      status[kStatusPoison] = 1;  // the static-list cache must not keep anything of this launch
    }
  }
  const bool lazyA = mode == 0 && view_cnt[v] < 2u;
  constexpr bool lazyD = true;  // dynamic lists are never materialised
  const int64_t itemA = (int64_t)v * Ns, itemD = (int64_t)V * Ns + (int64_t)f * Nm;
  if (mode == 0 && use_view_images && nd == 0 && view_cnt[v] >= 2u) {  // static-only tile of a shared view
#pragma unroll
    for (int q = 0; q < PPT; ++q)
    if (q == 0 ? inside : inside1) {
      const size_t HW = (size_t)g.H * g.W, pid = (size_t)(py + 4 * q) * g.W + px;
      if (out_u8) {
        const uint8_t* src = st_u8 + ((size_t)v * HW + pid) * 3;
        uint8_t* o = out_u8 + ((size_t)f * HW + pid) * 3;
        o[0] = src[0];
        o[1] = src[1];
        o[2] = src[2];
      }
      if (out_color) {
        const float* src = st_color + (size_t)v * 3 * HW + pid;
        float* o = out_color + (size_t)f * 3 * HW + pid;
        o[0] = src[0];
        o[HW] = src[HW];
        o[2 * HW] = src[2 * HW];
      }
      if (out_invd) out_invd[(size_t)f * HW + pid] = st_invd[(size_t)v * HW + pid];
      if (SEG && out_seg) out_seg[(size_t)f * HW + pid] = st_seg[(size_t)v * HW + pid];
    }
    return;
  }
  const uint32_t total = ns + nd;
  uint32_t i0 = 0, j0 = 0;  // consumed from the static / dynamic stream (block-uniform)

  const uint32_t aRec = pin((uint32_t)__cvta_generic_to_shared(sRec));
  const uint32_t aQ = pin((uint32_t)__cvta_generic_to_shared(&sQ[w][0]));

  // Tacc is the transmittance used for accumulation; it is forced to 0 once the pixel has saturated so
  // that every later record fails the T test at once.  Tout keeps the value the reference writes out.
  float Tacc = inside ? 1.0f : 0.0f, Tout = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f, invd = 0.f;
  float best_w = 0.f;
  int32_t best_lab = GSW_SEG_NONE;
  // PPT == 2: the same state for both pixels, packed {pixel 0, pixel 1}
  // Its transmittance is sign-coded: T > 0 while the pixel is live, -T (the value the reference writes out) once it
  // has saturated, so that T (1 - alpha) < 1e-4 holds for every later record without a second state variable.
  f2 T2 = pk(inside ? 1.0f : -1.0f, inside1 ? 1.0f : -1.0f), C0_2 = 0ull, C1_2 = 0ull, C2_2 = 0ull, invd2 = 0ull;
  const f2 npy2 = pk(-pyf, -(pyf + 4.0f));
  float best_w1 = 0.f;
  int32_t best_lab1 = GSW_SEG_NONE;
  auto all_done = [&]() -> bool {
    return PPT == 1 ? Tacc == 0.0f : (T2 & 0x8000000080000000ull) == 0x8000000080000000ull;  // both saturated
  };
  uint32_t base = 0;
  if (threadIdx.x < 8) sPairs[threadIdx.x] = 0u;  // (the first barrier of the loop orders this before any update)
  // tma_stage: chunks that come from ONE materialised stream (the shared static list of a view: contiguous 48 B
  // splats) are pulled in by one elected thread with a single bulk copy (cp.async.bulk + mbarrier) instead of three
  // 16 B loads and stores per thread.  Chunks that are merged from two streams or fetched through keys -> records are
  // gathers and stay as they are.
  const uint32_t mb = (uint32_t)__cvta_generic_to_shared(&mbar_s);
  uint32_t tma_phase = 0;
  if (tma_stage && threadIdx.x == 0) mbar_init(mb, 1);  // visible to the CTA after the loop's first barrier
  for (; base < total; base += kChunk) {
    if (__syncthreads_and(all_done())) break;  // also protects the staging buffers of the previous round
    const uint32_t n = min((uint32_t)kChunk, total - base);
    const uint32_t rs = ns - i0, rd = nd - j0;
    if (rs == 0 || rd == 0) {  // single-stream chunk
      const uint32_t src = rd == 0 ? s0 + i0 : d0 + j0;
      if (!SEG && tma_stage && rd == 0 && !lazyA) {  // CTA-uniform
        if (threadIdx.x == 0) {
          mbar_expect_tx(mb, n * 48u);
          bulk_g2s(aRec, splats + src, n * 48u, mb);
        }
        mbar_wait(mb, tma_phase);
        tma_phase ^= 1u;
      } else
      for (uint32_t t = threadIdx.x; t < n; t += NT) {
        if (rd == 0 ? lazyA : lazyD) {
          const uint32_t idx = (uint32_t)(keys[src + t] & 0xffffffffull);
          const Splat sp = make_splat(rec, (rd == 0 ? itemA : itemD) + (pos_of ? (int64_t)pos_of[idx] : (int64_t)idx));
          sRec[3 * t] = sp.a;
          sRec[3 * t + 1] = sp.b;
          sRec[3 * t + 2] = sp.c;
          if (SEG) sLab[t] = labels[idx];
        } else {
          const Splat* sp = splats + src + t;
          sRec[3 * t] = __ldg(&sp->a);
          sRec[3 * t + 1] = __ldg(&sp->b);
          sRec[3 * t + 2] = __ldg(&sp->c);
          if (SEG) sLab[t] = lab_stream[src + t];
        }
      }
      if (rd == 0) i0 += n; else j0 += n;
    } else {  // merge-path over two key windows
      const uint32_t ws = min((uint32_t)kChunk, rs), wd = min((uint32_t)kChunk, rd);
      for (uint32_t t = threadIdx.x; t < ws; t += NT) kS[t] = keys[s0 + i0 + t];
      for (uint32_t t = threadIdx.x; t < wd; t += NT) kD[t] = keys[d0 + j0 + t];
      __syncthreads();
      uint32_t took_s = 0;
      for (uint32_t t = threadIdx.x; t < n; t += NT) {
        uint32_t lo = t > wd ? t - wd : 0u, hi = min(t, ws);
        while (lo < hi) {
          const uint32_t mid = (lo + hi) >> 1;
          if (kS[mid] < kD[t - 1 - mid]) lo = mid + 1; else hi = mid;
        }
        const uint32_t a = lo, b = t - lo;
        const bool take_s = a < ws && (b >= wd || kS[a] < kD[b]);
        took_s += take_s ? 1u : 0u;
        const uint32_t pos = take_s ? s0 + i0 + a : d0 + j0 + b;
        if (take_s ? lazyA : lazyD) {
          const uint32_t idx = (uint32_t)((take_s ? kS[a] : kD[b]) & 0xffffffffull);
          const Splat sp = make_splat(rec, (take_s ? itemA : itemD) + (pos_of ? (int64_t)pos_of[idx] : (int64_t)idx));
          sRec[3 * t] = sp.a;
          sRec[3 * t + 1] = sp.b;
          sRec[3 * t + 2] = sp.c;
          if (SEG) sLab[t] = labels[idx];
        } else {
          const Splat* sp = splats + pos;
          sRec[3 * t] = __ldg(&sp->a);
          sRec[3 * t + 1] = __ldg(&sp->b);
          sRec[3 * t + 2] = __ldg(&sp->c);
          if (SEG) sLab[t] = lab_stream[pos];
        }
      }
      uint32_t cs = (uint32_t)__syncthreads_count(took_s >= 1u);
      if (PPT == 2) cs += (uint32_t)__syncthreads_count(took_s >= 2u);
      i0 += cs;
      j0 += n - cs;
    }
    __syncthreads();
    if (__all_sync(0xffffffffu, all_done())) continue;

    // ---- cull: one staged record per lane against this warp's 8x4 pixel block; hits go to the warp's queue.
    // Stage 1: bounding box of the alpha >= 1/255 footprint.  Stage 2 (box hits only): the exact maximum of
    // the (concave) exponent over the pixel block, reached either at the splat centre or on one of the two
    // block edges facing it; the record is dropped when even that maximum gives alpha < 1/255.
    uint32_t cnt = 0;
    for (uint32_t j0w = 0; j0w < n; j0w += 32) {
      const uint32_t j = j0w + lane;
      bool hit = false;
      if (j < n) {
        const float4 a = lds128(aRec + j * 48);
        hit = a.z >= 0.f && !(a.x + a.z < bx0 || a.x - a.z > bx1 || a.y + a.w < by0 || a.y - a.w > by1);
        if (hit && a.z < 1.0e30f) {
          const float4 b = lds128(aRec + j * 48 + 16);
          const float x0 = bx0 - a.x, x1 = bx1 - a.x, y0 = by0 - a.y, y1 = by1 - a.y;
          const float xe = fminf(fmaxf(0.f, x0), x1), ye = fminf(fmaxf(0.f, y0), y1);
          const float ys = fminf(fmaxf(__fdividef(-0.5f * b.y * xe, b.z), y0), y1);
          const float xs = fminf(fmaxf(__fdividef(-0.5f * b.y * ye, b.x), x0), x1);
          const float p1 = xe * (b.x * xe + b.y * ys) + b.z * ys * ys;
          const float p2 = xs * (b.x * xs + b.y * ye) + b.z * ye * ye;
          const float t2 = __log2f(255.0f * b.w) * 1.002f + 0.02f;  // log2 of the largest useful attenuation
          hit = fmaxf(p1, p2) >= -t2;
        }
      }
      const uint32_t mask = __ballot_sync(0xffffffffu, hit);
      if (hit) sts_u8(aQ + cnt + __popc(mask & ((1u << lane) - 1u)), j);
      cnt += __popc(mask);
    }
    __syncwarp();

    // ---- blend the queued records in list order ----
    uint32_t evaluated = 0;  // records of this chunk the warp really blended (it may stop early)
    for (uint32_t k0 = 0; k0 < cnt; k0 += 16) {
      if (k0 && __all_sync(0xffffffffu, all_done())) break;
      const uint32_t kend = min(cnt, k0 + 16);
      evaluated = kend;
#pragma unroll 4
      for (uint32_t k = k0; k < kend; ++k) {
        const uint32_t qi = lds_u8(aQ + k);
        const uint32_t ra = aRec + qi * 48;
        const float2 xy = lds64(ra);
        const float4 b = lds128(ra + 16);
        if (PPT == 2) {
          // two pixels of one column: dx is shared, everything that depends on dy is packed
          const float4 c = lds128(ra + 32);
          const float dx = xy.x - pxf;
          const f2 dy = add2(bc(xy.y), npy2);
          const f2 t = fma2(bc(b.y), dy, bc(b.x * dx));
          const f2 u = mul2(mul2(bc(b.z), dy), dy);
          const f2 p2 = fma2(bc(dx), t, u);  // power * log2(e)
          const float p20 = lo_of(p2), p21 = hi_of(p2);
          f2 al = mul2(bc(b.w), pk(ex2_approx(p20), ex2_approx(p21)));
          float a0 = fminf(0.99f, lo_of(al)), a1 = fminf(0.99f, hi_of(al));
          a0 = keep_alpha(a0, p20);  // skipped: alpha 0 leaves T unchanged bit for bit
          a1 = keep_alpha(a1, p21);
          al = pk(a0, a1);
          const f2 tt = fma2(pk(-a0, -a1), T2, T2);  // T (1 - alpha) = T - alpha T
          const float tt0 = lo_of(tt), tt1 = hi_of(tt);
          const bool sat0 = tt0 < 0.0001f, sat1 = tt1 < 0.0001f;  // saturated now, or already (T < 0)
          f2 wg = mul2(al, T2);
          const float w0 = sat0 ? 0.0f : lo_of(wg), w1 = sat1 ? 0.0f : hi_of(wg);
          wg = pk(w0, w1);
          C0_2 = fma2(bc(c.x), wg, C0_2);
          C1_2 = fma2(bc(c.y), wg, C1_2);
          C2_2 = fma2(bc(c.z), wg, C2_2);
          invd2 = fma2(bc(c.w), wg, invd2);
          if (SEG) {
            if (w0 > best_w) {
              best_w = w0;
              best_lab = sLab[qi];
            }
            if (w1 > best_w1) {
              best_w1 = w1;
              best_lab1 = sLab[qi];
            }
          }
          // (a skipped record has alpha 0, hence tt == T bit for bit)
          T2 = pk(sat0 ? -fabsf(lo_of(T2)) : tt0, sat1 ? -fabsf(hi_of(T2)) : tt1);
          continue;
        }
        const float dx = xy.x - pxf, dy = xy.y - pyf;
        const float p2 = dx * (b.x * dx + b.y * dy) + (b.z * dy) * dy;  // power * log2(e)
        // Branch-free form of the reference's three tests (power > 0 -> skip; alpha < 1/255 -> skip;
        // T(1-alpha) < 1e-4 -> done): a skipped record takes alpha = 0, which leaves T unchanged bit for bit,
        // and the accumulation is predicated.  The warp executes the union of its lanes' paths either way, so
        // the divergent branches only cost issue slots (ncu source view: 16 % of the kernel's instructions).
        const float4 c = lds128(ra + 32);
        float alpha = fminf(0.99f, b.w * ex2_approx(p2));
        const bool skip = (p2 > 0.0f) || (alpha < 1.0f / 255.0f);
        alpha = skip ? 0.0f : alpha;
        const float test_T = Tacc * (1 - alpha);
        const bool sat = test_T < 0.0001f;  // saturated now, or already (Tacc == 0)
        const float wgt = alpha * Tacc;
        if (!skip && !sat) {
          C0 += c.x * wgt;
          C1 += c.y * wgt;
          C2 += c.z * wgt;
          invd += c.w * wgt;
          if (SEG && wgt > best_w) {
            best_w = wgt;
            best_lab = sLab[qi];
          }
          Tout = test_T;
        }
        Tacc = sat ? 0.0f : test_T;
      }
    }
    if (lane == 0) sPairs[w] += evaluated * PPT;
  }
  // measurement counters: entries staged, and (record, pixel) pair evaluations = 32 x records blended per warp
  __syncthreads();
  if (threadIdx.x == 0) {
    atomicAdd((unsigned long long*)&status[3], (unsigned long long)min(base, total));
    uint32_t ev = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) ev += sPairs[k];
    atomicAdd((unsigned long long*)&status[4], 32ull * ev);
  }

  if (bg_dev) {
    bg0 = bg_dev[0];
    bg1 = bg_dev[1];
    bg2 = bg_dev[2];
  }
#pragma unroll
  for (int q = 0; q < PPT; ++q) {
    if (!(q == 0 ? inside : inside1)) continue;
    float T = Tout, c0 = C0, c1 = C1, c2 = C2, idp = invd;
    int32_t lab = best_lab;
    if (PPT == 2) {
      T = fabsf(q ? hi_of(T2) : lo_of(T2));
      c0 = q ? hi_of(C0_2) : lo_of(C0_2);
      c1 = q ? hi_of(C1_2) : lo_of(C1_2);
      c2 = q ? hi_of(C2_2) : lo_of(C2_2);
      idp = q ? hi_of(invd2) : lo_of(invd2);
      lab = q ? best_lab1 : best_lab;
    }
    const float r = c0 + T * bg0, gr = c1 + T * bg1, b = c2 + T * bg2;
    const size_t HW = (size_t)g.H * g.W, pid = (size_t)(py + 4 * q) * g.W + px;
    if (out_color) {
      float* o = out_color + (size_t)f * 3 * HW + pid;
      o[0] = r;
      o[HW] = gr;
      o[2 * HW] = b;
    }
    if (out_invd) out_invd[(size_t)f * HW + pid] = idp;
    if (SEG && out_seg) out_seg[(size_t)f * HW + pid] = lab;
    if (out_u8) {
      uint8_t* o = out_u8 + ((size_t)f * HW + pid) * 3;
      o[0] = (uint8_t)(fminf(fmaxf(r, 0.f), 1.f) * 255.f);
      o[1] = (uint8_t)(fminf(fmaxf(gr, 0.f), 1.f) * 255.f);
      o[2] = (uint8_t)(fminf(fmaxf(b, 0.f), 1.f) * 255.f);
    }
  }
}

// The kernel proper: blend_body, then (delivery to the host only) the group-completion signal.
#ifndef GSW_BLEND_CTAS
#define GSW_BLEND_CTAS 8
#endif
template <bool SEG, int PPT>
__global__ void __launch_bounds__(256 / PPT, PPT == 1 ? 5 : GSW_BLEND_CTAS)  // 48 / 64 registers: 40 / 32 warps per SM
blend_kernel(const uint32_t* __restrict__ bin_off, int64_t nbins, const int32_t* __restrict__ frame_view, int32_t V,
             int32_t f_begin, const unsigned long long* __restrict__ keys, const Splat* __restrict__ splats,
             int64_t cap, FrameGeom g, float bg0, float bg1, float bg2, const float* __restrict__ bg_dev,
             uint8_t* __restrict__ out_u8, float* __restrict__ out_color, float* __restrict__ out_invd,
             long long* __restrict__ status, int mode, const uint32_t* __restrict__ view_cnt,
             const uint8_t* __restrict__ st_u8, const float* __restrict__ st_color,
             const float* __restrict__ st_invd, const int32_t* __restrict__ lab_stream, int32_t* __restrict__ out_seg,
             const int32_t* __restrict__ st_seg, int use_view_images, const float4* __restrict__ rec,
             const int32_t* __restrict__ pos_of, int64_t Ns, const int32_t* __restrict__ labels, int64_t Nm,
             DoneSig sig, int tma_stage, const uint32_t* __restrict__ cache_m) {
  if (sig.times != nullptr && blockIdx.x == 0 && threadIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    sig.times[16] = t;
  }
  blend_body<SEG, PPT>(bin_off, nbins, frame_view, V, f_begin, keys, splats, cap, g, bg0, bg1, bg2, bg_dev, out_u8,
                  out_color, out_invd, status, mode, view_cnt, st_u8, st_color, st_invd, lab_stream, out_seg, st_seg,
                  use_view_images, rec, pos_of, Ns, labels, Nm, tma_stage, cache_m);
  if (sig.cnt != nullptr) {
    // every thread's pixel stores precede the barrier; thread 0's release makes them visible device-wide (and to
    // the copy engine) before the counter moves; the CTA that completes a group publishes the epoch
    __syncthreads();
    if (threadIdx.x == 0) {
      const int f = f_begin + (int)(blockIdx.x / (uint32_t)g.T);
      int k = 0;
      while (k + 1 < sig.n && f >= sig.cut(k + 1)) ++k;
      const uint32_t ctas = (uint32_t)(sig.cut(k + 1) - sig.cut(k)) * (uint32_t)g.T;
      // Release-only ordering: the counter increment is a gpu-scope release (cumulative over the CTA's stores
      // through the barrier above), the flag store a system-scope release.  __threadfence() would do, but it also
      // carries acquire semantics, for which ptxas emits CCTL.IVALL: every one of the F*T CTAs would invalidate
      // its SM's L1 under the co-resident CTAs' record gathers (measured: +9 % blend time with signalling on).
      uint32_t old;
      asm volatile("atom.add.release.gpu.global.u32 %0, [%1], 1;" : "=r"(old) : "l"(sig.cnt + k) : "memory");
      if (old == ctas - 1u) {
        if (sig.times != nullptr) {
          unsigned long long t;
          asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
          sig.times[k] = t;
        }
        __threadfence();  // acquire side, once per group: the other CTAs' releases on the counter happen-before the flag
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(sig.flag + k), "r"(sig.epoch) : "memory");
      }
    }
  }
}


}  // namespace

// Pixels per thread of the blend kernel (see blend_body).  GSW_BLEND_PPT=1 selects the one-pixel form (A/B aid); the
// view pass and the frame pass of a launch set always use the same form, so shared static tiles stay bit-identical.
static int blend_ppt() {
  static const int v = [] {
    const char* e = getenv("GSW_BLEND_PPT");
    return (e != nullptr && e[0] == '1') ? 1 : 2;
  }();
  return v;
}

// Bulk-copy staging of single-stream chunks of materialised lists (see blend_body).  GSW_BLEND_TMA=0 turns it off.
static int blend_tma() {
  static const int v = [] {
    const char* e = getenv("GSW_BLEND_TMA");
    return (e != nullptr && e[0] == '0') ? 0 : 1;
  }();
  return v;
}

// Static-tile sharing: renders the static-only image of every view that is
// shared by >= 2 frames (mode 1).  Buffers live in the workspace; call before launch_blend when F > V.
int launch_blend_view_pass(const gsw_ctx* ctx, int32_t V, int32_t F, int64_t nbins, const int32_t* frame_view,
                           int64_t cap, const FrameGeom& g, const float bg[3], bool want_u8, bool want_color,
                           bool want_invd, bool want_seg, cudaStream_t s) {
  auto& w = ctx->ws;
  (void)F;
#define GSW_VIEW_PASS(SEG_, PPT_)                                                                                \
  blend_kernel<SEG_, PPT_><<<(unsigned)((int64_t)V * g.T), 256 / PPT_, 0, s>>>(                                    \
      w.bin_off.as<uint32_t>(), nbins, frame_view, V, 0, w.keys.as<unsigned long long>(), w.splats.as<Splat>(), cap, g, \
      bg[0], bg[1], bg[2], nullptr, want_u8 ? w.st_u8.as<uint8_t>() : nullptr,                                      \
      want_color ? w.st_color.as<float>() : nullptr, want_invd ? w.st_invd.as<float>() : nullptr,                   \
      w.status.as<long long>(), 1, w.view_cnt.as<uint32_t>(), nullptr, nullptr, nullptr, w.lab_stream.as<int32_t>(), \
      want_seg ? w.st_seg.as<int32_t>() : nullptr, nullptr, 0, w.rec.as<float4>(), ctx->last.pos_of, ctx->last.Ns,    \
      ctx->labels, ctx->last.Nm, DoneSig{}, blend_tma(), ctx->cache_on ? w.cache_m.as<uint32_t>() : nullptr)
  if (blend_ppt() == 2) {
    if (want_seg) GSW_VIEW_PASS(true, 2); else GSW_VIEW_PASS(false, 2);
  } else {
    if (want_seg) GSW_VIEW_PASS(true, 1); else GSW_VIEW_PASS(false, 1);
  }
#undef GSW_VIEW_PASS
  count_launches(1);
  GSW_CUDA(cudaGetLastError());
  return 0;
}

// blends frames [f_begin, f_begin + F)
// `share`: the view pass ran for this launch set, static-only tiles of shared views are copied from it
int launch_blend(const gsw_ctx* ctx, int32_t V, int32_t f_begin, int32_t F, int64_t nbins, const int32_t* frame_view,
                 int64_t cap, const FrameGeom& g, const float bg[3], const float* bg_dev, uint8_t* out_u8,
                 float* out_color, float* out_invdepth, int32_t* out_seg, bool share, cudaStream_t s,
                 const DoneSig* sig) {
  const int64_t FT = (int64_t)F * g.T;
  if (FT <= 0) return 0;
  auto& w = ctx->ws;
  const DoneSig sg = sig ? *sig : DoneSig{};
#define GSW_BLEND(SEG_, PPT_)                                                                                      \
  blend_kernel<SEG_, PPT_><<<(unsigned)FT, 256 / PPT_, 0, s>>>(                                                      \
      w.bin_off.as<uint32_t>(), nbins, frame_view, V, f_begin, w.keys.as<unsigned long long>(), w.splats.as<Splat>(), \
      cap, g, bg ? bg[0] : 0.f, bg ? bg[1] : 0.f, bg ? bg[2] : 0.f, bg_dev, out_u8, out_color, out_invdepth,          \
      w.status.as<long long>(), 0, w.view_cnt.as<uint32_t>(), w.st_u8.as<uint8_t>(), w.st_color.as<float>(),          \
      w.st_invd.as<float>(), w.lab_stream.as<int32_t>(), out_seg, w.st_seg.as<int32_t>(), share ? 1 : 0,              \
      w.rec.as<float4>(), ctx->last.pos_of, ctx->last.Ns, ctx->labels, ctx->last.Nm, sg, blend_tma(), nullptr)
  if (blend_ppt() == 2) {
    if (out_seg) GSW_BLEND(true, 2); else GSW_BLEND(false, 2);
  } else {
    if (out_seg) GSW_BLEND(true, 1); else GSW_BLEND(false, 1);
  }
#undef GSW_BLEND
  count_launches(1);
  GSW_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace gsw
