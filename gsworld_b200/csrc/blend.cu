// blend.cu — 16x16-tile front-to-back alpha compositing over the sorted splat stream.  sm_100a.
//
// One CTA per (frame, tile); 8 warps, each owning an 8x4 pixel block.  A chunk of 256 splat records is
// staged in shared memory; every warp then culls 32 staged records at a time against its own pixel
// block (one record per lane, one ballot) and runs the per-pixel blend only for the survivors, in list
// order.  Skipped records are exactly those whose alpha would be < 1/255 on all 32 pixels, which the
// reference skips too, so the result is unchanged.  Warps whose 32 pixels have all saturated stop
// evaluating; the CTA leaves when every warp has.
//
// Reference arithmetic: diff-gaussian-rasterization@9c5c2028 forward.cu renderCUDA (absent submodule):
//   power = -0.5(A dx^2 + C dy^2) - B dx dy; skip if > 0; alpha = min(0.99, o*exp(power)); skip if
//   < 1/255; stop when T(1-alpha) < 1e-4; C += rgb*alpha*T; invD += alpha*T/depth; out = C + T*bg.
// uint8 epilogue: gs_world_wrapper.py:268-270 after render()'s clamp(0,1) (gaussian_renderer/__init__.py:119).
#include "gsw_internal.h"

namespace gsw {
namespace {

constexpr int kChunk = 256;

__global__ void __launch_bounds__(256)
blend_kernel(const TileDesc* __restrict__ tdesc, const Splat* __restrict__ splats, FrameGeom g, float bg0, float bg1,
             float bg2, const float* __restrict__ bg_dev, uint8_t* __restrict__ out_u8, float* __restrict__ out_color,
             float* __restrict__ out_invd) {
  __shared__ float4 sA[kChunk], sB[kChunk], sC[kChunk];
  const int64_t ft = blockIdx.x;
  const int f = (int)(ft / g.T), tile = (int)(ft % g.T);
  const int tx = tile % g.gx, ty = tile / g.gx;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int wx = tx * kTile + (w & 1) * 8, wy = ty * kTile + (w >> 1) * 4;
  const int px = wx + (lane & 7), py = wy + (lane >> 3);
  const bool inside = px < g.W && py < g.H;
  const float pxf = (float)px, pyf = (float)py;
  const float bx0 = (float)wx, bx1 = (float)(wx + 7), by0 = (float)wy, by1 = (float)(wy + 3);
  const TileDesc td = tdesc[ft];

  float T = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f, invd = 0.f;
  bool done = !inside;

  for (uint32_t base = 0; base < td.count; base += kChunk) {
    if (__syncthreads_and(done)) break;  // also protects the staging buffers of the previous round
    const uint32_t n = min((uint32_t)kChunk, td.count - base);
    if (threadIdx.x < n) {
      const Splat* sp = splats + td.offset + base + threadIdx.x;
      sA[threadIdx.x] = __ldg(&sp->a);
      sB[threadIdx.x] = __ldg(&sp->b);
      sC[threadIdx.x] = __ldg(&sp->c);
    }
    __syncthreads();
    if (__all_sync(0xffffffffu, done)) continue;
    for (uint32_t j0 = 0; j0 < n; j0 += 32) {
      const uint32_t j = j0 + lane;
      bool hit = false;
      if (j < n) {
        const float4 a = sA[j];
        hit = a.z >= 0.f && !(a.x + a.z < bx0 || a.x - a.z > bx1 || a.y + a.w < by0 || a.y - a.w > by1);
      }
      uint32_t mask = __ballot_sync(0xffffffffu, hit);
      while (mask) {
        const uint32_t jj = j0 + (uint32_t)(__ffs(mask) - 1);
        mask &= mask - 1;
        if (done) continue;
        const float4 a = sA[jj];
        const float4 b = sB[jj];
        const float dx = a.x - pxf, dy = a.y - pyf;
        const float power = -0.5f * (b.x * dx * dx + b.z * dy * dy) - b.y * dx * dy;
        if (power > 0.0f) continue;
        const float alpha = fminf(0.99f, b.w * __expf(power));
        if (alpha < 1.0f / 255.0f) continue;
        const float test_T = T * (1 - alpha);
        if (test_T < 0.0001f) {
          done = true;
          continue;
        }
        const float4 c = sC[jj];
        const float wgt = alpha * T;
        C0 += c.x * wgt;
        C1 += c.y * wgt;
        C2 += c.z * wgt;
        invd += c.w * wgt;
        T = test_T;
      }
      if (__all_sync(0xffffffffu, done)) break;
    }
  }

  if (inside) {
    if (bg_dev) {
      bg0 = bg_dev[0];
      bg1 = bg_dev[1];
      bg2 = bg_dev[2];
    }
    const float r = C0 + T * bg0, gr = C1 + T * bg1, b = C2 + T * bg2;
    const size_t HW = (size_t)g.H * g.W, pid = (size_t)py * g.W + px;
    if (out_color) {
      float* o = out_color + (size_t)f * 3 * HW + pid;
      o[0] = r;
      o[HW] = gr;
      o[2 * HW] = b;
    }
    if (out_invd) out_invd[(size_t)f * HW + pid] = invd;
    if (out_u8) {
      uint8_t* o = out_u8 + ((size_t)f * HW + pid) * 3;
      o[0] = (uint8_t)(fminf(fmaxf(r, 0.f), 1.f) * 255.f);
      o[1] = (uint8_t)(fminf(fmaxf(gr, 0.f), 1.f) * 255.f);
      o[2] = (uint8_t)(fminf(fmaxf(b, 0.f), 1.f) * 255.f);
    }
  }
}

}  // namespace

int launch_blend(const gsw_ctx* ctx, int32_t F, const FrameGeom& g, const float bg[3], uint8_t* out_u8,
                 float* out_color, float* out_invdepth, cudaStream_t s) {
  const int64_t FT = (int64_t)F * g.T;
  if (FT <= 0) return 0;
  auto& w = ctx->ws;
  blend_kernel<<<(unsigned)FT, 256, 0, s>>>(w.tdesc.as<TileDesc>(), w.splats.as<Splat>(), g, bg[0], bg[1], bg[2],
                                            nullptr, out_u8, out_color, out_invdepth);
  count_launches(1);
  GSW_CUDA(cudaGetLastError());
  return 0;
}

int launch_blend_bgdev(const gsw_ctx* ctx, const FrameGeom& g, const float* bg_dev, float* out_color,
                       float* out_invdepth, cudaStream_t s) {
  auto& w = ctx->ws;
  blend_kernel<<<(unsigned)g.T, 256, 0, s>>>(w.tdesc.as<TileDesc>(), w.splats.as<Splat>(), g, 0.f, 0.f, 0.f, bg_dev,
                                             nullptr, out_color, out_invdepth);
  count_launches(1);
  GSW_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace gsw
