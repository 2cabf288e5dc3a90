// preprocess.cu — per-Gaussian stage: (optional) rigid transform, EWA projection, cull, SH colour,
// tile-bin counting and key scatter.  sm_100a.
//
// COMPILED WITH -fmad=false.  radii, tile counts, depth bits and pixel positions are integer/bit-exact
// parity quantities, so every fp32 operation here must round once, in the order the CPU oracle
// (oracle/gsw_oracle.cpp) and the published rasterizer write it.  The stage is HBM-bound (45 B read per
// Gaussian, 236 B more per visible one), so the un-fused multiplies cost nothing measurable.
//
// Reference for the arithmetic: diff-gaussian-rasterization@9c5c2028 forward.cu preprocessCUDA /
// computeCov3D / computeCov2D / computeColorFromSH and auxiliary.h (absent submodule; see include/gsw.h),
// near plane per GSWorld README.md:33; transform per gsworld/utils/gs_utils.py:283-385.
#include "gsw_internal.h"

namespace gsw {
namespace {

__device__ __constant__ float kC0 = 0.28209479177387814f;
__device__ __constant__ float kC1 = 0.4886025119029199f;
__device__ __constant__ float kC2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                        -1.0925484305920792f, 0.5462742152960396f};
__device__ __constant__ float kC3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                        0.3731763325901154f,  -0.4570457994644658f, 1.445305721320277f,
                                        -0.5900435899266435f};

struct M3 {  // m[c][r], glm-style column major
  float m[3][3];
};

__device__ __forceinline__ M3 mul3(const M3& a, const M3& b) {
  M3 o;
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      float s = a.m[0][r] * b.m[c][0];
      s = s + a.m[1][r] * b.m[c][1];
      s = s + a.m[2][r] * b.m[c][2];
      o.m[c][r] = s;
    }
  return o;
}

__device__ __forceinline__ M3 tr3(const M3& a) {
  M3 o;
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int r = 0; r < 3; ++r) o.m[c][r] = a.m[r][c];
  return o;
}

__device__ __forceinline__ float ndc2pix(float v, int S) {
  return (float)((((double)v + 1.0) * (double)S - 1.0) * 0.5);
}

struct Rect {
  int x0, y0, x1, y1;
};

__device__ __forceinline__ Rect tile_rect(float px, float py, int radius, int gx, int gy) {
  const float r = (float)radius;
  Rect q;
  q.x0 = min(gx, max(0, (int)((px - r) / (float)kTile)));
  q.y0 = min(gy, max(0, (int)((py - r) / (float)kTile)));
  q.x1 = min(gx, max(0, (int)((px + r + (float)(kTile - 1)) / (float)kTile)));
  q.y1 = min(gy, max(0, (int)((py + r + (float)(kTile - 1)) / (float)kTile)));
  return q;
}

// ---- exact tile culling (FrameGeom::tile_cull) ----
// The reference duplicates a Gaussian into every tile of the bounding square of its 3-sigma radius.  Many of
// those (Gaussian, tile) entries never do anything: on all 256 pixels of the tile alpha stays below 1/255 (or
// power > 0), and the reference's blend loop `continue`s past them without touching any state.  Dropping such
// an entry before it is binned therefore leaves every pixel bit-identical, and saves its scatter, sort,
// staging and cull work.  The test is the one the blend applies per warp (blend.cu), on the whole tile:
// the largest power over the tile's pixel centres is reached at the centre if it is inside, else on one of
// the two tile edges facing it (moving any tile point towards the centre along its ray lowers the convex
// form and stays in the tile until one of those edges is hit).  Margins (0.2 % + 0.02 on the log side) keep
// it conservative against ex2.approx / FMA differences in the blend; any NaN keeps the entry.
struct Foot {
  float px, py, hx, hy;  // centre (pixels), half extents of the alpha >= 1/255 box (-1: never visible, 3e38: unknown)
  float A, B, C;         // pre-scaled conic (-0.5 A, -B, -0.5 C): power = A dx^2 + B dx dy + C dy^2
  float kyx, kxy;        // -B/C, -B/A: the y (x) that maximises the power on a line of fixed x (y)
  float t;               // ln(255 opacity) with margin: entries with max power < -t are dead
};

// (make_foot is evaluated with identical inputs and code by the preprocess and the scatter pass.)
__device__ __forceinline__ Foot make_foot(float px, float py, float hx, float hy, float A, float B, float C,
                                          float opac) {
  Foot f;
  f.px = px; f.py = py; f.hx = hx; f.hy = hy; f.A = -0.5f * A; f.B = -B; f.C = -0.5f * C;
  f.kyx = -B / C;
  f.kxy = -B / A;
  f.t = logf(255.0f * opac) * 1.002f + 0.02f;
  return f;
}

// For tiles inside alive_rect(f, .) only: the box test against the alpha >= 1/255 extents is done there, once per
// Gaussian, by shrinking the candidate rectangle.
__device__ __forceinline__ bool tile_alive(const Foot& f, int tx, int ty) {
  if (f.hx >= 1.0e30f) return true;
  const float x0 = (float)(tx * kTile) - f.px, x1 = x0 + (float)(kTile - 1);
  const float y0 = (float)(ty * kTile) - f.py, y1 = y0 + (float)(kTile - 1);
  const float xe = fminf(fmaxf(0.f, x0), x1), ye = fminf(fmaxf(0.f, y0), y1);
  const float ys = fminf(fmaxf(f.kyx * xe, y0), y1);
  const float xs = fminf(fmaxf(f.kxy * ye, x0), x1);
  // (not a parity quantity: explicit FMAs although the file is built with -fmad=false; the preprocess and the
  // scatter pass run this very code on the same bits, so they always agree)
  const float p1 = fmaf(xe, fmaf(f.A, xe, f.B * ys), (f.C * ys) * ys);
  const float p2 = fmaf(xs, fmaf(f.A, xs, f.B * ye), (f.C * ye) * ye);
  return !(fmaxf(p1, p2) < -f.t);
}

// Tiles of rc that overlap the alpha >= 1/255 box of the Gaussian (usually much smaller than the
// bounding square of the 3-sigma radius that rc comes from): the loops below only visit these.  Part of the
// decision (both passes call it), not just a shortcut.
__device__ __forceinline__ Rect alive_rect(const Foot& f, const Rect rc) {
  Rect q = rc;
  if (f.hx < 0.f) {
    q.x1 = q.x0;
    q.y1 = q.y0;
  } else if (f.hx < 1.0e30f) {
    const float s = 1.0f / (float)kTile;  // exact scaling
    q.x0 = max(rc.x0, (int)ceilf((f.px - f.hx - (float)(kTile - 1)) * s));
    q.x1 = min(rc.x1, (int)floorf((f.px + f.hx) * s) + 1);
    q.y0 = max(rc.y0, (int)ceilf((f.py - f.hy - (float)(kTile - 1)) * s));
    q.y1 = min(rc.y1, (int)floorf((f.py + f.hy) * s) + 1);
    if (q.x1 < q.x0) q.x1 = q.x0;
    if (q.y1 < q.y0) q.y1 = q.y0;
  }
  return q;
}

__device__ __forceinline__ uint32_t depth_bucket(float depth, const FrameGeom& g) {
  const uint32_t bits = __float_as_uint(depth);
  if (bits <= kDepthLo) return 0u;
  return min((uint32_t)(g.NB - 1), (bits - kDepthLo) >> g.nb_shift);
}

struct ViewSmem {
  float V[16], P[16], cam[3], tanx, tany;
  float wnorm;  // upper bound on the squared spectral norm of the view matrix's 3x3 block (1 for a rigid camera)
  float focal_x, focal_y;  // W / (2 tanx), H / (2 tany): the same expressions the per-item code used to evaluate
};

// One colour channel of the degree<=3 SH polynomial, same association as the published kernel.
template <typename F>
__device__ __forceinline__ float sh_channel(int deg, float x, float y, float z, F S) {
  float res = kC0 * S(0);
  if (deg > 0) {
    res = res - kC1 * y * S(1) + kC1 * z * S(2) - kC1 * x * S(3);
    if (deg > 1) {
      const float xx = x * x, yy = y * y, zz = z * z;
      const float xy = x * y, yz = y * z, xz = x * z;
      res = res + kC2[0] * xy * S(4) + kC2[1] * yz * S(5) + kC2[2] * (2.0f * zz - xx - yy) * S(6) +
            kC2[3] * xz * S(7) + kC2[4] * (xx - yy) * S(8);
      if (deg > 2) {
        res = res + kC3[0] * y * (3.0f * xx - yy) * S(9) + kC3[1] * xy * z * S(10) +
              kC3[2] * y * (4.0f * zz - xx - yy) * S(11) +
              kC3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * S(12) +
              kC3[4] * x * (4.0f * zz - xx - yy) * S(13) + kC3[5] * z * (xx - yy) * S(14) +
              kC3[6] * x * (xx - 3.0f * yy) * S(15);
      }
    }
  }
  res += 0.5f;
  return fmaxf(res, 0.0f);
}

// Items per CTA round ("chunk") of each layer: 256 threads x chunk / 256 consecutive items of one list.  kIpbStatic is
// the static layer's LARGEST chunk (and the only one of the forms without a work list); with a work list the launch
// picks 256 ... kIpbStatic (pick_static_ipb).  The dynamic layer's is fixed (measured at 64 envs, step in ms: 512 items
// 1.10, 1024 items 1.02, 2048 items 1.01, 4096 items 1.05; 512 is no faster at 8 envs once the static chunk is small).
#ifndef GSW_IPB_DYN
#define GSW_IPB_DYN 2048
#endif
#ifndef GSW_IPB_STAT
#define GSW_IPB_STAT 2048
#endif
constexpr int kIpbStatic = GSW_IPB_STAT, kIpbDynamic = GSW_IPB_DYN;
constexpr size_t kMaxHistBytes = 160 * 1024;         // per-CTA shared histogram budget; above it: global atomics

// Per-item body.  `hist` (shared, one counter per (tile, bucket) of this list) may be null, in which case
// the global bin counters are hit directly.
template <bool DYN, bool CULL>
__device__ __forceinline__ void preprocess_item(const GaussArrays& ga, const GaussArrays& la, int64_t i, int64_t il, int64_t item, int list_global,
                                                const FrameGeom& g, const ViewSmem& vs, const float* pose_s,
                                                float4* __restrict__ recA, float4* __restrict__ recB,
                                                float4* __restrict__ recC, int32_t* __restrict__ rad,
                                                uint32_t* __restrict__ bin_count, uint32_t* hist, Rect& out_rc,
                                                uint32_t& out_bk, Foot& out_ft, uint32_t* s_vis) {
  // CULL: the caller counts the Gaussian into its bins (warp-cooperatively, with the exact tile test) from
  // out_rc / out_bk / out_ft, which stay empty on every early return
  if (!DYN && la.slot != nullptr && la.slot[il] >= 0) {  // movable: handled by the dynamic layer
    rad[item] = 0;
    return;
  }

  float p[3] = {la.xyz[3 * il], la.xyz[3 * il + 1], la.xyz[3 * il + 2]};
  float sc[3] = {0.f, 0.f, 0.f};
  float q[4] = {1.f, 0.f, 0.f, 0.f};
  const bool have_cov = (!DYN) && ga.cov3d_precomp != nullptr;
  const int64_t big = ga.sh_idx ? (int64_t)ga.sh_idx[i] : i;  // row of the large per-Gaussian arrays

  if (DYN) {
    // ---- rigid transform of a movable Gaussian (gs_utils.py:283-385 order: scale, rotate, translate) ----
    const float* ps = pose_s + la.mov_slot[il] * GSW_POSE_STRIDE;
    if (ps[17] != 0.f) {  // object: positions scaled by kappa
      const float kappa = ps[16];
#pragma unroll
      for (int k = 0; k < 3; ++k) p[k] = p[k] * kappa;
    }
    float rp[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) rp[r] = ps[3 * r] * p[0] + ps[3 * r + 1] * p[1] + ps[3 * r + 2] * p[2];
#pragma unroll
    for (int k = 0; k < 3; ++k) p[k] = rp[k] + ps[9 + k];
    // (scales and the rotation part of the transform follow the early culls below: most movable Gaussians of a frame
    // end there)
  }

  // ---- near-plane cull on view-space depth ----
  const float* V = vs.V;
  const float pvx = V[0] * p[0] + V[4] * p[1] + V[8] * p[2] + V[12];
  const float pvy = V[1] * p[0] + V[5] * p[1] + V[9] * p[2] + V[13];
  const float pvz = V[2] * p[0] + V[6] * p[1] + V[10] * p[2] + V[14];
  if (pvz <= g.near_plane) {
    rad[item] = 0;
    return;
  }
  const float* P = vs.P;
  const float phx = P[0] * p[0] + P[4] * p[1] + P[8] * p[2] + P[12];
  const float phy = P[1] * p[0] + P[5] * p[1] + P[9] * p[2] + P[13];
  const float phw = P[3] * p[0] + P[7] * p[1] + P[11] * p[2] + P[15];
  const float pw = 1.0f / (phw + 0.0000001f);
  const float ppx = phx * pw, ppy = phy * pw;
  const float px = ndc2pix(ppx, g.W), py = ndc2pix(ppy, g.H);

  // ---- Jacobian of the projection (EWA) ----
  const float focal_y = vs.focal_y, focal_x = vs.focal_x;
  float tx = pvx, ty = pvy;
  const float tz = pvz;
  {
    const float limx = 1.3f * vs.tanx, limy = 1.3f * vs.tany;
    const float txtz = tx / tz, tytz = ty / tz;
    tx = fminf(limx, fmaxf(-limx, txtz)) * tz;
    ty = fminf(limy, fmaxf(-limy, tytz)) * tz;
  }
  M3 J;
  J.m[0][0] = focal_x / tz;  J.m[0][1] = 0.f;            J.m[0][2] = -(focal_x * tx) / (tz * tz);
  J.m[1][0] = 0.f;           J.m[1][1] = focal_y / tz;   J.m[1][2] = -(focal_y * ty) / (tz * tz);
  J.m[2][0] = 0.f;           J.m[2][1] = 0.f;            J.m[2][2] = 0.f;

  // ---- conservative frustum cull (not in the reference; it only skips work whose result is "no tile") ----
  // trace(cov2D) <= lambda_max(Sigma) * |J|_F^2 * |W|_2^2 and lambda_1 <= trace(cov2D) + 0.92, so the true
  // radius ceil(3 sqrt(lambda_1)) is <= r_b; if even a splat of radius r_b around (px,py) cannot reach the
  // tile grid, the exact path below would end in the same empty tile rectangle.  NaNs fail every test.
  // lambda_max(Sigma) <= trace(Sigma) = sum_k s_k^2 |row_k(R)|^2, and the rows of R (built from a quaternion
  // normalised to a few ulp) have length 1 to ~1e-6: the scales alone give the bound, before any of the
  // rotation algebra (quaternion product, R, S R) is spent on a Gaussian that ends here.
  float c3[6];
  float tr_sigma;  // >= trace of the 3D covariance >= its largest eigenvalue
  if (have_cov) {
#pragma unroll
    for (int k = 0; k < 6; ++k) c3[k] = ga.cov3d_precomp[6 * big + k];
    tr_sigma = c3[0] + c3[3] + c3[5];
  } else {
    bool raw_scale = false;
    if (DYN) {
      const float* ps = pose_s + la.mov_slot[il] * GSW_POSE_STRIDE;
      if (ps[17] != 0.f) {  // object: scales go through exp -> * kappa -> inverse_sigmoid -> exp (gs_utils.py:310)
        raw_scale = true;
        const float kappa = ps[16];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          const float u = expf(la.log_scale_raw[3 * il + k]) * kappa;
          const float ls = logf(u / (1.0f - u));  // inverse_sigmoid, the reference behaviour
          sc[k] = expf(ls);
        }
      }
    }
    if (!raw_scale) {
      sc[0] = la.scale_act[3 * il];
      sc[1] = la.scale_act[3 * il + 1];
      sc[2] = la.scale_act[3 * il + 2];
    }
    const float s0 = g.scale_modifier * sc[0], s1 = g.scale_modifier * sc[1], s2 = g.scale_modifier * sc[2];
    tr_sigma = (s0 * s0 + s1 * s1 + s2 * s2) * 1.001f;
  }
  {
    const float jf2 = J.m[0][0] * J.m[0][0] + J.m[0][2] * J.m[0][2] + J.m[1][1] * J.m[1][1] + J.m[1][2] * J.m[1][2];
    const float lam_b = tr_sigma * jf2 * vs.wnorm * 1.02f + 1.0f;
    const float r_b = 3.0f * sqrtf(lam_b) + 2.0f;
    if (tr_sigma >= 0.f &&
        (px + r_b < 0.5f || px - r_b > (float)(g.gx * kTile) + 0.5f || py + r_b < 0.5f ||
         py - r_b > (float)(g.gy * kTile) + 0.5f)) {
      rad[item] = 0;
      return;
    }
  }

  // ---- 3D covariance ----
  if (!have_cov) {
    if (DYN) {
      // rotation part of the rigid transform (gs_utils.py:283-385): q = normalize(pose_q * normalize(q_raw) * |q_raw|)
      const float* ps = pose_s + la.mov_slot[il] * GSW_POSE_STRIDE;
      const float4 qr = reinterpret_cast<const float4*>(la.quat_raw)[il];
      const float nrm = sqrtf(qr.x * qr.x + qr.y * qr.y + qr.z * qr.z + qr.w * qr.w);
      const float bw = qr.x / nrm, bx = qr.y / nrm, by = qr.z / nrm, bz = qr.w / nrm;
      const float aw = ps[12], ax = ps[13], ay = ps[14], az = ps[15];
      float ow = aw * bw - ax * bx - ay * by - az * bz;
      float ox = aw * bx + ax * bw + ay * bz - az * by;
      float oy = aw * by - ax * bz + ay * bw + az * bx;
      float oz = aw * bz + ax * by - ay * bx + az * bw;
      if (ow < 0.f) {
        ow = -ow; ox = -ox; oy = -oy; oz = -oz;
      }
      const float q2[4] = {ow * nrm, ox * nrm, oy * nrm, oz * nrm};
      const float n2 = sqrtf(q2[0] * q2[0] + q2[1] * q2[1] + q2[2] * q2[2] + q2[3] * q2[3]);
      const float den = fmaxf(n2, 1e-12f);  // F.normalize eps (gaussian_model.py:47)
#pragma unroll
      for (int k = 0; k < 4; ++k) q[k] = q2[k] / den;
    } else {
      const float4 qq = reinterpret_cast<const float4*>(la.quat_act)[il];
      q[0] = qq.x; q[1] = qq.y; q[2] = qq.z; q[3] = qq.w;
    }
    M3 M;
    M3 S;
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
      for (int r = 0; r < 3; ++r) S.m[c][r] = 0.f;
    S.m[0][0] = g.scale_modifier * sc[0];
    S.m[1][1] = g.scale_modifier * sc[1];
    S.m[2][2] = g.scale_modifier * sc[2];
    const float r = q[0], x = q[1], y = q[2], z = q[3];
    M3 R;
    R.m[0][0] = 1.f - 2.f * (y * y + z * z);
    R.m[0][1] = 2.f * (x * y - r * z);
    R.m[0][2] = 2.f * (x * z + r * y);
    R.m[1][0] = 2.f * (x * y + r * z);
    R.m[1][1] = 1.f - 2.f * (x * x + z * z);
    R.m[1][2] = 2.f * (y * z - r * x);
    R.m[2][0] = 2.f * (x * z - r * y);
    R.m[2][1] = 2.f * (y * z + r * x);
    R.m[2][2] = 1.f - 2.f * (x * x + y * y);
    M = mul3(S, R);
    const M3 Sig = mul3(tr3(M), M);
    c3[0] = Sig.m[0][0]; c3[1] = Sig.m[0][1]; c3[2] = Sig.m[0][2];
    c3[3] = Sig.m[1][1]; c3[4] = Sig.m[1][2]; c3[5] = Sig.m[2][2];
  }

  // ---- EWA 2D covariance ----
  M3 Wm;
  Wm.m[0][0] = V[0]; Wm.m[0][1] = V[4]; Wm.m[0][2] = V[8];
  Wm.m[1][0] = V[1]; Wm.m[1][1] = V[5]; Wm.m[1][2] = V[9];
  Wm.m[2][0] = V[2]; Wm.m[2][1] = V[6]; Wm.m[2][2] = V[10];
  const M3 Tm = mul3(Wm, J);
  M3 Vrk;
  Vrk.m[0][0] = c3[0]; Vrk.m[0][1] = c3[1]; Vrk.m[0][2] = c3[2];
  Vrk.m[1][0] = c3[1]; Vrk.m[1][1] = c3[3]; Vrk.m[1][2] = c3[4];
  Vrk.m[2][0] = c3[2]; Vrk.m[2][1] = c3[4]; Vrk.m[2][2] = c3[5];
  const M3 cov = mul3(mul3(tr3(Tm), tr3(Vrk)), Tm);
  float cx = cov.m[0][0];
  const float cy = cov.m[0][1];
  float cz = cov.m[1][1];

  const float det_cov = cx * cz - cy * cy;
  cx += 0.3f;
  cz += 0.3f;
  const float det = cx * cz - cy * cy;
  float h_scale = 1.0f;
  if (g.antialiasing) h_scale = sqrtf(fmaxf(0.000025f, det_cov / det));
  if (det == 0.0f) {
    rad[item] = 0;
    return;
  }
  const float det_inv = 1.f / det;
  const float conx = cz * det_inv, cony = -cy * det_inv, conz = cx * det_inv;
  const float mid = 0.5f * (cx + cz);
  const float lam1 = mid + sqrtf(fmaxf(0.1f, mid * mid - det));
  const float lam2 = mid - sqrtf(fmaxf(0.1f, mid * mid - det));
  const float my_radius = ceilf(3.f * sqrtf(fmaxf(lam1, lam2)));
  const int radius = (int)my_radius;
  const Rect rc = tile_rect(px, py, radius, g.gx, g.gy);
  if ((rc.x1 - rc.x0) * (rc.y1 - rc.y0) == 0) {
    rad[item] = 0;
    return;
  }

  // ---- colour ----
  float rgb[3];
  if (!DYN && ga.colors_precomp != nullptr) {
    rgb[0] = ga.colors_precomp[3 * big];
    rgb[1] = ga.colors_precomp[3 * big + 1];
    rgb[2] = ga.colors_precomp[3 * big + 2];
  } else {
    float dx = p[0] - vs.cam[0], dy = p[1] - vs.cam[1], dz = p[2] - vs.cam[2];
    const float len = sqrtf(dx * dx + dy * dy + dz * dz);
    dx = dx / len;
    dy = dy / len;
    dz = dz / len;
    const float* shp = ga.sh + (size_t)big * ga.M * 3;
    if (ga.M == 16 && ((reinterpret_cast<uintptr_t>(shp) & 15) == 0)) {
      float c[48];
      const float4* s4 = reinterpret_cast<const float4*>(shp);
#pragma unroll
      for (int k = 0; k < 12; ++k) {
        const float4 v = __ldg(s4 + k);
        c[4 * k] = v.x; c[4 * k + 1] = v.y; c[4 * k + 2] = v.z; c[4 * k + 3] = v.w;
      }
#pragma unroll
      for (int ch = 0; ch < 3; ++ch)
        rgb[ch] = sh_channel(ga.sh_degree, dx, dy, dz, [&](int k) { return c[3 * k + ch]; });
    } else {
#pragma unroll
      for (int ch = 0; ch < 3; ++ch)
        rgb[ch] = sh_channel(ga.sh_degree, dx, dy, dz, [&](int k) { return __ldg(shp + 3 * k + ch); });
    }
  }

  // ---- footprint of alpha >= 1/255 (cull box for the blend; conservative, not a parity quantity) ----
  const float opac = la.opacity[il] * h_scale;
  float hx, hy;
  if (!(opac * 255.0f >= 0.999f)) {
    hx = hy = -1.0f;  // can never reach alpha 1/255
  } else if (!(det > 1e-4f * (cx * cz))) {
    hx = hy = 3.0e38f;  // ill-conditioned conic: do not cull
  } else {
    const float tau = fmaxf(logf(opac * 255.0f), 0.f) * 1.02f + 0.05f;
    hx = sqrtf(2.f * tau * cx) * 1.001f + 0.01f;
    hy = sqrtf(2.f * tau * cz) * 1.001f + 0.01f;
  }

  {  // measurement counter: one shared-memory atomic per warp for the lanes that reach this point together
    const unsigned m = __activemask();
    if ((threadIdx.x & 31) == __ffs(m) - 1) atomicAdd(s_vis, (uint32_t)__popc(m));
  }
  recA[4 * item] = make_float4(px, py, hx, hy);
  recB[4 * item] = make_float4(conx, cony, conz, opac);
  recC[4 * item] = make_float4(rgb[0], rgb[1], rgb[2], pvz);
  rad[item] = radius;

  // ---- count this Gaussian into every (tile, depth bucket) bin it touches ----
  // (per-thread loop: here the imbalance between one-tile and hundred-tile Gaussians hides behind the other
  // warps' arithmetic; the warp-cooperative form measured slower in this kernel, faster in the scatter)
  const uint32_t bk = depth_bucket(pvz, g);
  if (CULL) {
    const Foot ft = make_foot(px, py, hx, hy, conx, cony, conz, opac);
    out_rc = rc;
    out_bk = bk;
    out_ft = ft;
    return;
  }
  if (hist != nullptr) {
    for (int y = rc.y0; y < rc.y1; ++y)
      for (int x = rc.x0; x < rc.x1; ++x) atomicAdd(&hist[(uint32_t)(y * g.gx + x) * (uint32_t)g.NB + bk], 1u);
  } else {
    const uint32_t lbase = (uint32_t)list_global * (uint32_t)g.T;
    for (int y = rc.y0; y < rc.y1; ++y)
      for (int x = rc.x0; x < rc.x1; ++x)
        atomicAdd(&bin_count[(lbase + (uint32_t)(y * g.gx + x)) * (uint32_t)g.NB + bk], 1u);
  }
}

// Tiles per Gaussian above which the whole warp expands its rectangle.  Close to a camera (wrist view) one
// Gaussian covers a hundred tiles and more while its neighbours cover one or two: a per-thread tile loop
// leaves 31 lanes waiting, and whole CTAs trail the kernel (ncu: SMs active 30 % of the scatter's duration).
// Used by the scatter pass.
constexpr int kCoopTiles = 8;
constexpr int kCoopTilesCull = 32;  // exact-tile-culling count pass
constexpr int kCoopBitsCull = 16;   // exact-tile-culling scatter pass (alive tiles of a mask)

// Calls f(bin_in_list, p0, p1) once per (tile, bucket) bin of every lane's rectangle; p0/p1 are the owning
// lane's payload words (broadcast when the warp expands the rectangle).  Must be reached by all 32 lanes.
template <class F>
__device__ __forceinline__ void warp_for_each_bin(const Rect rc, uint32_t bk, uint32_t p0, uint32_t p1, int gx, int NB,
                                                  F&& f) {
  const int w = rc.x1 - rc.x0, n = w * (rc.y1 - rc.y0);
  const bool coop = n > kCoopTiles;
  if (!coop)
    for (int y = rc.y0; y < rc.y1; ++y)
      for (int x = rc.x0; x < rc.x1; ++x) f((uint32_t)(y * gx + x) * (uint32_t)NB + bk, p0, p1);
  unsigned m = __ballot_sync(0xffffffffu, coop);
  const int lane = threadIdx.x & 31;
  while (m) {
    const int src = __ffs(m) - 1;
    m &= m - 1;
    const int x0 = __shfl_sync(0xffffffffu, rc.x0, src), y0 = __shfl_sync(0xffffffffu, rc.y0, src);
    const int ww = __shfl_sync(0xffffffffu, w, src), nn = __shfl_sync(0xffffffffu, n, src);
    const uint32_t b = __shfl_sync(0xffffffffu, bk, src);
    const uint32_t q0 = __shfl_sync(0xffffffffu, p0, src), q1 = __shfl_sync(0xffffffffu, p1, src);
    for (int t = lane; t < nn; t += 32) {
      const int dy = t / ww;
      f((uint32_t)((y0 + dy) * gx + x0 + (t - dy * ww)) * (uint32_t)NB + b, q0, q1);
    }
  }
}

// Exact-tile-culling form of the count (preprocess pass): calls f(bin_in_list) for every tile of the lane's
// rectangle that passes alive_rect + tile_alive and returns the lane's mask of alive tiles (bit = row-major
// position inside the rectangle, first 64 tiles only: larger rectangles are tested again by the scatter pass).
// Candidate sets of more than kCoopTilesCull tiles are expanded by the whole warp.  Must be reached by all 32
// lanes.
template <class F>
__device__ __forceinline__ unsigned long long warp_cull_count(const Rect rc, uint32_t bk, const Foot& ft, int gx,
                                                              int NB, F&& f) {
  const int w = rc.x1 - rc.x0;
  const Rect q = alive_rect(ft, rc);
  const int qw = q.x1 - q.x0, n = qw * (q.y1 - q.y0);
  uint32_t mlo = 0u, mhi = 0u;  // the 64-bit mask as two words (64-bit shifts are several instructions each)
  auto set_bit = [](uint32_t& lo, uint32_t& hi, int bit) {
    if (bit < 32) lo |= 1u << bit;
    else if (bit < 64) hi |= 1u << (bit - 32);
  };
  // neighbours in Morton order have similar footprints, so per-thread loops stay balanced up to a full warp's
  // worth of tiles; the broadcast of a whole Foot only pays beyond that
  const bool coop = n > kCoopTilesCull;
  if (!coop) {
    for (int y = q.y0; y < q.y1; ++y)
      for (int x = q.x0; x < q.x1; ++x)
        if (tile_alive(ft, x, y)) {
          f((uint32_t)(y * gx + x) * (uint32_t)NB + bk);
          const int bit = (y - rc.y0) * w + (x - rc.x0);
          set_bit(mlo, mhi, bit);
        }
  }
  unsigned m = __ballot_sync(0xffffffffu, coop);
  const int lane = threadIdx.x & 31;
  while (m) {
    const int src = __ffs(m) - 1;
    m &= m - 1;
    const int x0 = __shfl_sync(0xffffffffu, q.x0, src), y0 = __shfl_sync(0xffffffffu, q.y0, src);
    const int ww = __shfl_sync(0xffffffffu, qw, src), nn = __shfl_sync(0xffffffffu, n, src);
    const int rx0 = __shfl_sync(0xffffffffu, rc.x0, src), ry0 = __shfl_sync(0xffffffffu, rc.y0, src);
    const int rw = __shfl_sync(0xffffffffu, w, src);
    const uint32_t b = __shfl_sync(0xffffffffu, bk, src);
    Foot fs;
    fs.px = __shfl_sync(0xffffffffu, ft.px, src); fs.py = __shfl_sync(0xffffffffu, ft.py, src);
    fs.hx = __shfl_sync(0xffffffffu, ft.hx, src); fs.hy = __shfl_sync(0xffffffffu, ft.hy, src);
    fs.A = __shfl_sync(0xffffffffu, ft.A, src);   fs.B = __shfl_sync(0xffffffffu, ft.B, src);
    fs.C = __shfl_sync(0xffffffffu, ft.C, src);   fs.t = __shfl_sync(0xffffffffu, ft.t, src);
    fs.kyx = __shfl_sync(0xffffffffu, ft.kyx, src); fs.kxy = __shfl_sync(0xffffffffu, ft.kxy, src);
    uint32_t lo = 0u, hi = 0u;  // alive tiles this lane found, as bits of the owner's mask
    for (int t = lane; t < nn; t += 32) {
      const int dy = t / ww;
      const int x = x0 + (t - dy * ww), y = y0 + dy;
      if (tile_alive(fs, x, y)) {
        f((uint32_t)(y * gx + x) * (uint32_t)NB + b);
        const int bit = (y - ry0) * rw + (x - rx0);
        set_bit(lo, hi, bit);
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      lo |= __shfl_xor_sync(0xffffffffu, lo, o);
      hi |= __shfl_xor_sync(0xffffffffu, hi, o);
    }
    if (lane == src) {
      mlo = lo;
      mhi = hi;
    }
  }
  return ((unsigned long long)mhi << 32) | mlo;
}

// ---- group-level frustum culling ----
// Items are stored so that 256 consecutive ones are spatially compact (Morton order, see gsw_scene_set);
// `b` = {cx, cy, cz, rho, sigma, smax, slot, -}: bounding sphere of the group's centres (scene frame, scan
// pose for movable ones), sigma = max sqrt(trace(Sigma)) at scale_modifier 1, smax = largest raw scale,
// slot = the group's single pose slot or -1.  Returns true when every item of the group is certain to end
// with "no tile" in the exact path: either behind the near plane, or even the item-level radius bound
// (see the conservative frustum cull in preprocess_item) cannot reach the tile grid.  The side-plane tests
// assume the standard projection (w_clip = view depth, x_clip = x_view / tan); `proj_std` says whether the
// view's matrices have that form.
struct ViewCull {
  float fx, fy, a, nzl, nzr, nzt, nzb;
  int proj_std;
};

template <bool DYN>
__device__ bool group_culled(const float* __restrict__ b, const ViewSmem& vs, const ViewCull& vc, const FrameGeom& g,
                             const float* pose_s) {
  float cx = b[0], cy = b[1], cz = b[2], rho = b[3], sigma = b[4];
  if (DYN) {
    const int slot = (int)b[6];
    if (slot < 0) return false;  // group spans several links: no common transform
    const float* ps = pose_s + slot * GSW_POSE_STRIDE;
    float k = 1.0f;
    if (ps[17] != 0.f) {  // object: positions scaled by kappa, scales go through u / (1 - u)
      k = ps[16];
      const float u = k * b[5];
      if (!(u < 0.5f) || !(k > 0.f)) return false;
      sigma = sigma * k / (1.0f - u);
    }
    const float x = k * cx, y = k * cy, z = k * cz;
    cx = ps[0] * x + ps[1] * y + ps[2] * z + ps[9];
    cy = ps[3] * x + ps[4] * y + ps[5] * z + ps[10];
    cz = ps[6] * x + ps[7] * y + ps[8] * z + ps[11];
    // |R|_2^2 <= Gershgorin bound of R^T R (1 for a rotation)
    float rn = 0.f;
    for (int a = 0; a < 3; ++a) {
      float row = 0.f;
      for (int c = 0; c < 3; ++c) row += fabsf(ps[a] * ps[c] + ps[3 + a] * ps[3 + c] + ps[6 + a] * ps[6 + c]);
      rn = fmaxf(rn, row);
    }
    rho = rho * fabsf(k) * sqrtf(rn) * 1.001f;
    sigma = sigma * sqrtf(rn) * 1.001f;
  }
  const float* V = vs.V;
  const float vx = V[0] * cx + V[4] * cy + V[8] * cz + V[12];
  const float vy = V[1] * cx + V[5] * cy + V[9] * cz + V[13];
  const float vz = V[2] * cx + V[6] * cy + V[10] * cz + V[14];
  const float wn = sqrtf(vs.wnorm);
  const float r = rho * wn * 1.001f + 1e-4f * (1.0f + fabsf(vx) + fabsf(vy) + fabsf(vz));
  if (vz + r <= g.near_plane) return true;  // whole sphere at or behind the near plane
  if (!vc.proj_std) return false;
  const float s = sigma * g.scale_modifier * vc.a * 1.01f;
  if (!(s >= 0.f)) return false;
  // left / right / top / bottom: linear in (x, z) resp. (y, z), maximised over the sphere
  if (vc.fx * vx + vc.nzl * vz + r * sqrtf(vc.fx * vc.fx + vc.nzl * vc.nzl) + s < 0.f) return true;
  if (vc.fx * vx - vc.nzr * vz - r * sqrtf(vc.fx * vc.fx + vc.nzr * vc.nzr) - s > 0.f) return true;
  if (vc.fy * vy + vc.nzt * vz + r * sqrtf(vc.fy * vc.fy + vc.nzt * vc.nzt) + s < 0.f) return true;
  if (vc.fy * vy - vc.nzb * vz - r * sqrtf(vc.fy * vc.fy + vc.nzb * vc.nzb) - s > 0.f) return true;
  return false;
}

// ---- TMA bulk staging (cp.async.bulk + mbarrier): one elected thread pulls the CTA's contiguous slices of
// the small per-Gaussian arrays into shared memory in a few large asynchronous copies, so the per-item
// code never waits on a dependent global load for them (the kernel was latency-bound on exactly those).
// (helpers: gsw_internal.h)

// Per-view constants of the preprocess pass: the view itself plus the numbers of the group-level frustum tests.
// 48 words; computed by a few designated threads of a 256-thread CTA (compute_view_consts) and, for the static
// layer's work-list form, kept in global memory by the chunk-cull kernel so that a preprocess CTA only copies them.
struct ViewConsts {
  ViewSmem vs;
  ViewCull vc;
};
static_assert(sizeof(ViewConsts) == 48 * 4, "ViewConsts: 48 words");
constexpr int kViewConstWords = 48;

// Must be called by all threads of a CTA of >= 96 threads; the caller synchronises afterwards.
__device__ __forceinline__ void compute_view_consts(int view, const FrameGeom& g, const float* __restrict__ views,
                                                    const float* __restrict__ projs,
                                                    const float* __restrict__ campos,
                                                    const float* __restrict__ tanfov, ViewSmem& vs, ViewCull& vc) {
  if (threadIdx.x < 16) {
    vs.V[threadIdx.x] = views[view * 16 + threadIdx.x];
    vs.P[threadIdx.x] = projs[view * 16 + threadIdx.x];
  } else if (threadIdx.x < 19) {
    vs.cam[threadIdx.x - 16] = campos[view * 3 + threadIdx.x - 16];
  } else if (threadIdx.x == 19) {
    vs.tanx = tanfov[view * 2];
  } else if (threadIdx.x == 20) {
    vs.tany = tanfov[view * 2 + 1];
  } else if (threadIdx.x == 21) {
    vs.focal_x = (float)g.W / (2.0f * tanfov[view * 2]);
  } else if (threadIdx.x == 22) {
    vs.focal_y = (float)g.H / (2.0f * tanfov[view * 2 + 1]);
  } else if (threadIdx.x == 32) {  // Gershgorin bound on the largest eigenvalue of W^T W (W = view 3x3 block)
    const float* Vg = views + view * 16;
    float w[3][3];
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c) w[r][c] = Vg[4 * c + r];
    float best = 0.f;
    for (int a = 0; a < 3; ++a) {
      float row = 0.f;
      for (int b = 0; b < 3; ++b) row += fabsf(w[0][a] * w[0][b] + w[1][a] * w[1][b] + w[2][a] * w[2][b]);
      best = fmaxf(best, row);
    }
    vs.wnorm = best * 1.001f;
  } else if (threadIdx.x == 64) {  // constants of the group-level side-plane tests
    const float* Vg = views + view * 16;
    const float* Pg = projs + view * 16;
    const float tanx = tanfov[view * 2], tany = tanfov[view * 2 + 1];
    const float fx = (float)g.W / (2.0f * tanx), fy = (float)g.H / (2.0f * tany);
    // standard projection: clip.w = view.z, clip.x = view.x / tanx, clip.y = view.y / tany  (transposed storage)
    bool ok = true;
    for (int k = 0; k < 4; ++k) {
      const float tol = 1e-4f * (1.0f + fabsf(Vg[4 * k + 2]) + fabsf(Vg[4 * k]) + fabsf(Vg[4 * k + 1]));
      ok = ok && fabsf(Pg[4 * k + 3] - Vg[4 * k + 2]) <= tol && fabsf(Pg[4 * k] * tanx - Vg[4 * k]) <= tol &&
           fabsf(Pg[4 * k + 1] * tany - Vg[4 * k + 1]) <= tol;
    }
    float wn = 0.f;
    {
      float w[3][3];
      for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) w[r][c] = Vg[4 * c + r];
      for (int a = 0; a < 3; ++a) {
        float row = 0.f;
        for (int b = 0; b < 3; ++b) row += fabsf(w[0][a] * w[0][b] + w[1][a] * w[1][b] + w[2][a] * w[2][b]);
        wn = fmaxf(wn, row);
      }
    }
    const float kj = fx * fx * (1.0f + 1.69f * tanx * tanx) + fy * fy * (1.0f + 1.69f * tany * tany);
    vc.fx = fx;
    vc.fy = fy;
    vc.a = 3.0f * sqrtf(1.02f * kj * wn * 1.001f);
    vc.nzl = 0.5f * (float)g.W + 6.0f;
    vc.nzr = (float)(g.gx * kTile) - 0.5f * (float)g.W + 7.0f;
    vc.nzt = 0.5f * (float)g.H + 6.0f;
    vc.nzb = (float)(g.gy * kTile) - 0.5f * (float)g.H + 7.0f;
    vc.proj_std = ok ? 1 : 0;
  }
}

// First kernel of a launch set: the per-view constants of every view (block v < V; both layers' preprocess CTAs then
// only copy 48 words instead of each deriving them with two serial threads) and the number of frames that look
// through each view (view_cnt, zeroed before; decides which static lists are shared).  grid max(V, ceil(F / 256)).
__global__ void __launch_bounds__(256)
prepare_views_kernel(int32_t V, int32_t F, FrameGeom g, const float* __restrict__ views,
                     const float* __restrict__ projs, const float* __restrict__ campos,
                     const float* __restrict__ tanfov, const int32_t* __restrict__ frame_view,
                     uint32_t* __restrict__ view_consts, uint32_t* __restrict__ view_cnt, int cache_ok,
                     uint32_t* __restrict__ view_prev, uint32_t* __restrict__ view_same) {
  __shared__ ViewConsts c;
  const int f = blockIdx.x * 256 + threadIdx.x;
  if (f < F) atomicAdd(&view_cnt[frame_view ? frame_view[f] : 0], 1u);
  if ((int)blockIdx.x >= V) return;  // block-uniform
  if (view_prev != nullptr) {
    // Static-list cache (see cache_prefix_kernel): is this view's camera bitwise the one of the previous launch?
    // 37 words: view 16, projection 16, camera centre 3, tan(fov/2) 2; kept in view_prev[v][40].
    const int v = blockIdx.x;
    uint32_t now = 0u;
    const int t = threadIdx.x;
    if (t < 16) now = __float_as_uint(views[v * 16 + t]);
    else if (t < 32) now = __float_as_uint(projs[v * 16 + t - 16]);
    else if (t < 35) now = __float_as_uint(campos[v * 3 + t - 32]);
    else if (t < 37) now = __float_as_uint(tanfov[v * 2 + t - 35]);
    const bool diff = t < 37 && view_prev[v * 40 + t] != now;
    const int any_diff = __syncthreads_or(diff ? 1 : 0);
    if (t < 37) view_prev[v * 40 + t] = now;
    if (t == 0) view_same[v] = (cache_ok && !any_diff) ? 1u : 0u;
  }
  compute_view_consts(blockIdx.x, g, views, projs, campos, tanfov, c.vs, c.vc);
  __syncthreads();
  if (threadIdx.x < kViewConstWords)
    view_consts[(size_t)blockIdx.x * kViewConstWords + threadIdx.x] = reinterpret_cast<const uint32_t*>(&c)[threadIdx.x];
}

// Static-list cache across steps.  A view whose camera, scene, geometry and workspace are exactly those of the previous
// launch (a world-fixed camera: SURVEY.md section 8(f) row 4, "or once ever") would get bit for bit the same static
// records, bins, sorted keys, gathered splats and static-only tiles again, and all of those still sit in the workspace.
// Because the static lists are laid out in view order, only a PREFIX of unchanged views can be kept: m = number of
// leading views that are unchanged and whose shared / private status is the same as before.  Every static-chain kernel
// skips views below m (the work list simply holds no chunk of them); their bin counters are not cleared.  One warp.
__global__ void cache_prefix_kernel(int32_t V, const uint32_t* __restrict__ view_same,
                                    const uint32_t* __restrict__ view_cnt, uint32_t* __restrict__ view_cnt_prev,
                                    uint32_t* __restrict__ cache_m, long long* __restrict__ status) {
  const int lane = threadIdx.x;
  int m = V;
  if (status[kStatusPoison] != 0) m = 0;  // the previous launch overflowed: its lists are incomplete
  __syncwarp();
  if (lane == 0) status[kStatusPoison] = 0;
  for (int v0 = 0; v0 < V && m == V; v0 += 32) {
    const int v = v0 + lane;
    bool keep = true;
    if (v < V) keep = view_same[v] != 0u && (view_cnt[v] >= 2u) == (view_cnt_prev[v] >= 2u);
    const unsigned bad = __ballot_sync(0xffffffffu, !keep);
    if (bad) m = v0 + __ffs(bad) - 1;
  }
  for (int v = lane; v < V; v += 32) view_cnt_prev[v] = view_cnt[v];
  if (lane == 0) *cache_m = (uint32_t)m;
}

// ---- chunk-level culling of the static layer, ahead of the preprocess pass ----
// One thread per (view, 2048-item chunk): tests the chunk's 8 group spheres against the view and leaves the mask
// of live groups in chunk_mask[view * nchunks + chunk]; chunks with at least one live group are appended to a
// compact work list (view << 16 | chunk) that the preprocess and scatter passes of the static layer walk with
// persistent CTAs.  A wrist camera sees a sliver of the room: with one view per env most (view, chunk) pairs are
// dead, and launching a CTA for each only to find that out (and to write "radius 0" for its 2048 items) cost more
// than the live ones' arithmetic.  Items of dead groups are never written at all now; every reader of `rad`
// consults the mask.  grid (ceil(nchunks / 256), V).
__global__ void __launch_bounds__(256)
cull_chunks_kernel(const float* __restrict__ bounds, int64_t n_items, int32_t nchunks, FrameGeom g,
                   const float* __restrict__ views, const float* __restrict__ projs,
                   const float* __restrict__ campos, const float* __restrict__ tanfov,
                   uint32_t* __restrict__ chunk_mask, uint32_t* __restrict__ work, uint32_t* __restrict__ work_count,
                   const uint32_t* __restrict__ view_consts, const uint32_t* __restrict__ cache_m,
                   uint32_t* __restrict__ bin_count, uint32_t* __restrict__ bin_cursor, int32_t nbl, int32_t ipb) {
  __shared__ ViewConsts c;
  const int view = blockIdx.y;
  if (cache_m != nullptr) {
    if ((uint32_t)view < *cache_m) return;  // static lists of this view are kept from the previous step
    // this view's bin counters were left alone by the host (they may belong to a kept view): clear them here, before
    // the preprocess pass that follows on this stream counts into them
    for (int k = blockIdx.x * 256 + threadIdx.x; k < nbl; k += gridDim.x * 256) {
      bin_count[(size_t)view * nbl + k] = 0u;
      bin_cursor[(size_t)view * nbl + k] = 0u;
    }
  }
  if (threadIdx.x < kViewConstWords)  // written by prepare_views_kernel
    reinterpret_cast<uint32_t*>(&c)[threadIdx.x] = view_consts[(size_t)view * kViewConstWords + threadIdx.x];
  __syncthreads();
  const int chunk = blockIdx.x * 256 + threadIdx.x;
  uint32_t mask = 0;
  if (chunk < nchunks) {
    const int groups = ipb >> 8;  // the static layer's chunk is chosen per launch (pick_static_ipb)
#pragma unroll 1
    for (int k = 0; k < groups; ++k) {
      const int64_t g0 = (int64_t)chunk * ipb + k * 256;
      if (g0 < n_items && !group_culled<false>(bounds + (g0 >> 8) * 8, c.vs, c.vc, g, nullptr)) mask |= 1u << k;
    }
    chunk_mask[(size_t)view * nchunks + chunk] = mask;
  }
  const unsigned live = __ballot_sync(0xffffffffu, mask != 0);
  if (live) {
    const int lane = threadIdx.x & 31;
    uint32_t base = 0;
    if (lane == 0) base = atomicAdd(work_count, (uint32_t)__popc(live));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (mask) work[base + __popc(live & ((1u << lane) - 1u))] = ((uint32_t)view << 16) | (uint32_t)chunk;
  }
}

// grid: (ceil(n_items / IPB), n_lists); list = view (static layer) or frame (dynamic layer).  WL (static layer
// only): persistent CTAs pull (view, chunk) pairs from the work list of cull_chunks_kernel with a ticket counter.
// Dynamic shared memory: [staged arrays (if use_tma)] [hist: T*NB u32 (if use_hist)] [pose table (DYN)].
template <bool DYN, int IPB, bool TMA, bool CULL, bool WL>
__global__ void __launch_bounds__(256, TMA ? 2 : 3)
preprocess_kernel(GaussArrays ga, int64_t n_items, int32_t list_base, int64_t item_base, FrameGeom g,
                  const float* __restrict__ views, const float* __restrict__ projs,
                  const float* __restrict__ campos, const float* __restrict__ tanfov,
                  const int32_t* __restrict__ frame_env, const int32_t* __restrict__ frame_view,
                  const float* __restrict__ pose, int32_t num_slots, float4* __restrict__ recA,
                  float4* __restrict__ recB, float4* __restrict__ recC, int32_t* __restrict__ rad,
                  uint32_t* __restrict__ bin_count, int use_hist, uint32_t* __restrict__ cta_hist,
                  uint32_t* __restrict__ chunk_mask, int32_t nchunks, const uint32_t* __restrict__ work,
                  const uint32_t* __restrict__ work_count, uint32_t* __restrict__ ticket,
                  const uint32_t* __restrict__ view_consts, const uint32_t* __restrict__ order,
                  uint32_t* __restrict__ cost, unsigned long long* __restrict__ vis_count) {
  static_assert(!(WL && DYN), "the work-list form is the static layer's");
  constexpr int use_tma = TMA ? 1 : 0;
  __shared__ ViewConsts vcs;
  ViewSmem& vs = vcs.vs;
  ViewCull& vc = vcs.vc;
  __shared__ __align__(8) unsigned long long mbar_s;
  __shared__ uint32_t s_work;
  __shared__ uint32_t s_cost, s_vis;
  if (threadIdx.x == 0) s_vis = 0u;  // visible items of this CTA (measurement counter, flushed once at exit)
  if (threadIdx.x == 0) s_cost = 0u;  // (cost is only kept by the one-round forms; a barrier follows before its use)
  extern __shared__ __align__(128) unsigned char dsm_raw[];
  constexpr int IPT = IPB / 256;
  const int nbl = g.T * g.NB;
  // staged layout (bytes per item): static 12 xyz + 12 scale + 16 quat + 4 opacity + 4 slot = 48;
  //                                 dynamic 12 xyz + 12 scale + 12 log-scale + 16 quat_raw + 4 opacity + 4 slot = 60
  constexpr size_t kStage = (size_t)IPB * (DYN ? 60 : 48);
  unsigned char* stage = dsm_raw;
  uint32_t* hist_base = reinterpret_cast<uint32_t*>(dsm_raw + (use_tma ? kStage : 0));
  uint32_t* hist = use_hist ? hist_base : nullptr;
  float* pose_s = reinterpret_cast<float*>(hist_base + (use_hist ? nbl : 0));
  float* xyz_s = reinterpret_cast<float*>(stage);
  float* sc_s = xyz_s + 3 * IPB;
  float* ls_s = sc_s + 3 * IPB;                                // DYN only
  float* q_s = DYN ? ls_s + 3 * IPB : sc_s + 3 * IPB;         // 16 B aligned: IPB multiple of 4
  float* op_s = q_s + 4 * IPB;
  int32_t* slot_s = reinterpret_cast<int32_t*>(op_s + IPB);
  __shared__ int grp_culled[IPT];
  const uint32_t mb = (uint32_t)__cvta_generic_to_shared(&mbar_s);
  if (use_tma && threadIdx.x == 0) mbar_init(mb, 1);  // made visible by the first barrier below
  uint32_t tma_phase = 0;
#pragma unroll 1
  for (;;) {
    int list, chunk;
    if (WL) {
      // (the barrier after the view set-up separates this write from the previous round's reads of s_work)
      if (threadIdx.x == 0) {
        const uint32_t t = atomicAdd(ticket, 1u);
        s_work = t < *work_count ? work[t] : 0xffffffffu;
      }
      __syncthreads();
      const uint32_t wv = s_work;
      if (wv == 0xffffffffu) break;  // CTA-uniform
      list = (int)(wv >> 16);
      chunk = (int)(wv & 0xffffu);
    } else {
      // DYN: dispatch in the cost order of order_chunks_kernel when there is one (heaviest chunks first)
      const uint32_t lin = blockIdx.y * gridDim.x + blockIdx.x;
      const uint32_t id = (DYN && order != nullptr) ? order[lin] : lin;
      list = (int)(id / (uint32_t)nchunks);
      chunk = (int)(id - (uint32_t)list * (uint32_t)nchunks);
    }
    int view = list, env = 0;
    if (DYN) {
      view = frame_view[list];
      env = frame_env[list];
    }
    const int64_t base = (int64_t)chunk * IPB;
    const bool full = base + IPB <= n_items;
    const bool staged = use_tma && full;  // partial tail chunk: plain global loads
    const size_t chunk_id = (size_t)list * (size_t)nchunks + (size_t)chunk;
    if (view_consts != nullptr) {  // prepared once per view by prepare_views_kernel
      if (threadIdx.x < kViewConstWords)
        reinterpret_cast<uint32_t*>(&vcs)[threadIdx.x] = view_consts[(size_t)view * kViewConstWords + threadIdx.x];
    } else {
      compute_view_consts(view, g, views, projs, campos, tanfov, vs, vc);
    }
    if (use_hist)
      for (int k = threadIdx.x; k < nbl; k += blockDim.x) hist[k] = 0u;
    if (DYN) {
      for (int k = threadIdx.x; k < num_slots * GSW_POSE_STRIDE; k += blockDim.x)
        pose_s[k] = pose[(size_t)env * num_slots * GSW_POSE_STRIDE + k];
    }
    __syncthreads();  // view / pose / histogram ready
    if (threadIdx.x < IPT) {
      const int64_t g0 = base + (int64_t)threadIdx.x * 256;
      int c = 0;
      if (g0 >= n_items) c = 1;
      else if (WL) c = ((chunk_mask[chunk_id] >> threadIdx.x) & 1u) ? 0 : 1;
      else if (ga.bounds != nullptr) c = group_culled<DYN>(ga.bounds + (g0 >> 8) * 8, vs, vc, g, pose_s) ? 1 : 0;
      grp_culled[threadIdx.x] = c;
    }
    __syncthreads();
    uint32_t live_mask = 0;
#pragma unroll
    for (int it = 0; it < IPT; ++it) live_mask |= grp_culled[it] == 0 ? 1u << it : 0u;
    // items of dead groups are never written: every reader of `rad` consults this mask first
    if (!WL && chunk_mask != nullptr && threadIdx.x == 0) chunk_mask[chunk_id] = live_mask;
    if (live_mask == 0) {  // nothing of this chunk can reach the image
      if (cta_hist != nullptr && threadIdx.x == 0) cta_hist[chunk_id * (size_t)(nbl + 1)] = 0u;  // scatter skips the chunk
      if (cost != nullptr && threadIdx.x == 0) cost[chunk_id] = 0u;
      if (WL) continue;
      return;
    }
    if (staged && threadIdx.x == 0) {
      const bool want_slot = DYN || ga.slot != nullptr;
      const uint32_t bytes = (uint32_t)IPB * ((DYN ? 56u : 44u) + (want_slot ? 4u : 0u));
      mbar_expect_tx(mb, bytes);
      auto sa = [](const void* p) { return (uint32_t)__cvta_generic_to_shared(p); };
      bulk_g2s(sa(xyz_s), ga.xyz + 3 * base, IPB * 12, mb);
      bulk_g2s(sa(sc_s), ga.scale_act + 3 * base, IPB * 12, mb);
      if (DYN) {
        bulk_g2s(sa(ls_s), ga.log_scale_raw + 3 * base, IPB * 12, mb);
        bulk_g2s(sa(q_s), ga.quat_raw + 4 * base, IPB * 16, mb);
        bulk_g2s(sa(slot_s), ga.mov_slot + base, IPB * 4, mb);
      } else {
        bulk_g2s(sa(q_s), ga.quat_act + 4 * base, IPB * 16, mb);
        if (want_slot) bulk_g2s(sa(slot_s), ga.slot + base, IPB * 4, mb);
      }
      bulk_g2s(sa(op_s), ga.opacity + base, IPB * 4, mb);
    }
    GaussArrays la = ga;
    if (staged) {
      mbar_wait(mb, tma_phase);
      tma_phase ^= 1u;
      la.xyz = xyz_s;
      la.scale_act = sc_s;
      la.opacity = op_s;
      if (DYN) {
        la.log_scale_raw = ls_s;
        la.quat_raw = q_s;
        la.mov_slot = slot_s;
      } else {
        la.quat_act = q_s;
        if (ga.slot != nullptr) la.slot = slot_s;
      }
    }
    const uint32_t lbase = (uint32_t)(list_base + list) * (uint32_t)nbl;
#pragma unroll 1
    for (int it = 0; it < IPT; ++it) {
      const int il = it * 256 + threadIdx.x;
      const int64_t i = base + il;
      Rect rc{0, 0, 0, 0};
      uint32_t bk = 0;
      Foot ft{};
      if (i < n_items && !grp_culled[it]) {
        const int64_t item = item_base + (int64_t)list * n_items + i;
        if (TMA) {
          preprocess_item<DYN, CULL>(ga, la, i, staged ? (int64_t)il : i, item, list_base + list, g, vs, pose_s, recA,
                                     recB, recC, rad, bin_count, hist, rc, bk, ft, &s_vis);
        } else {  // same object for both roles: every load stays a read-only global load
          preprocess_item<DYN, CULL>(ga, ga, i, i, item, list_base + list, g, vs, pose_s, recA, recB, recC, rad,
                                     bin_count, hist, rc, bk, ft, &s_vis);
        }
      }
      if (CULL) {  // count into the bins of the tiles that pass the exact test (all 32 lanes arrive here)
        unsigned long long alive;
        if (use_hist)
          alive = warp_cull_count(rc, bk, ft, g.gx, g.NB, [&](uint32_t b) { atomicAdd(&hist[b], 1u); });
        else
          alive = warp_cull_count(rc, bk, ft, g.gx, g.NB, [&](uint32_t b) { atomicAdd(&bin_count[lbase + b], 1u); });
        // recD (4th float4 of the 64 B slot): mask of the alive tiles among the first 64 of the rectangle,
        // row-major; the scatter pass walks the set bits instead of repeating the test
        if (rc.x1 > rc.x0)
          recA[4 * (item_base + (int64_t)list * n_items + i) + 3] =
              make_float4(__uint_as_float((uint32_t)alive), __uint_as_float((uint32_t)(alive >> 32)), 0.f, 0.f);
      }
    }
    if (use_hist) {
      __syncthreads();
      const uint32_t gb = (uint32_t)(list_base + list) * (uint32_t)nbl;
      // the per-CTA counts are kept for the scatter pass, which walks the same chunk: word 0 = chunk is live
      uint32_t* keep = cta_hist ? cta_hist + chunk_id * (size_t)(nbl + 1) : nullptr;
      if (keep != nullptr && threadIdx.x == 0) keep[0] = 1u;
      uint32_t mine = 0;
      for (int k = threadIdx.x; k < nbl; k += blockDim.x) {
        const uint32_t c = hist[k];
        if (c) atomicAdd(&bin_count[gb + k], c);
        if (keep != nullptr) keep[1 + k] = c;
        mine += c;
      }
      if (cost != nullptr) {  // tile entries this chunk produced: its weight in the dispatch order (order_chunks_kernel)
        for (int o = 16; o > 0; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);
        if ((threadIdx.x & 31) == 0 && mine) atomicAdd(&s_cost, mine);
        __syncthreads();
        if (threadIdx.x == 0) cost[chunk_id] = s_cost;
      }
    }
    if (!WL) break;
    __syncthreads();  // staged arrays, histogram and view constants are rewritten by the next round
  }
  __syncthreads();
  if (threadIdx.x == 0 && s_vis) atomicAdd(vis_count, (unsigned long long)s_vis);
}

// Second pass: write (depth bits << 32 | original index) into the bin's slice.  Order inside a bin is
// arbitrary here; the per-bin sort orders by the full 64-bit key, which reproduces the reference's
// stable (tile, depth) radix sort exactly (ties on depth fall back to ascending Gaussian index).
// With use_hist the CTA takes its own per-bin entry counts (kept by the preprocess CTA of the same chunk in
// cta_hist; counted again here only if that is unavailable), reserves one contiguous range per non-empty bin
// with a single global atomic, then hands out slots with shared-memory atomics.
// Item handled by thread t in iteration it of a 2048-item chunk.  Items are spatially ordered, so the 256
// items of a group fall into very few tile bins; letting the 32 lanes of a warp take 4 items from each of
// the chunk's 8 groups (instead of 32 neighbours) cuts same-address shared-memory atomic conflicts 8-fold.
// (G = groups per chunk, NT = threads per CTA: NT / G threads share a group, each takes 256 G / NT of its items)
template <int G, int NT>
__device__ __forceinline__ int scatter_slot(int it, int t) {
  return ((t & (G - 1)) << 8) + it * (NT / G) + t / G;
}
// Threads per scatter CTA.  The dynamic layer's heaviest chunks (a gripper close-up: every Gaussian covers dozens of
// tiles) bound the pass; more warps per chunk shorten exactly those.
#ifndef GSW_SCATTER_NT_DYN
#define GSW_SCATTER_NT_DYN 256
#endif
constexpr int kScatterThreadsStatic = 256, kScatterThreadsDyn = GSW_SCATTER_NT_DYN;

// (list, chunk) of a scatter CTA's next round.  Plain form: one round, the block indices.  Work-list form (static
// layer): persistent CTAs take (view, chunk) pairs of cull_chunks_kernel's list by ticket.  CTA-uniform result.
template <bool WL>
__device__ __forceinline__ bool next_chunk(int& list, int& chunk, bool first, const uint32_t* __restrict__ work,
                                           const uint32_t* __restrict__ work_count, uint32_t* __restrict__ ticket,
                                           uint32_t* s_work, const uint32_t* __restrict__ order) {
  if (!WL) {
    const uint32_t lin = blockIdx.y * gridDim.x + blockIdx.x;
    const uint32_t id = order != nullptr ? order[lin] : lin;
    list = (int)(id / gridDim.x);
    chunk = (int)(id - (uint32_t)list * gridDim.x);
    return first;
  }
  __syncthreads();  // the previous round is done with the shared counters and with s_work
  if (threadIdx.x == 0) {
    const uint32_t t = atomicAdd(ticket, 1u);
    *s_work = t < *work_count ? work[t] : 0xffffffffu;
  }
  __syncthreads();
  const uint32_t wv = *s_work;
  list = (int)(wv >> 16);
  chunk = (int)(wv & 0xffffu);
  return wv != 0xffffffffu;
}

template <bool DYN, bool WL, int IPBS = kIpbStatic>
__global__ void __launch_bounds__(DYN ? kScatterThreadsDyn : kScatterThreadsStatic)
scatter_kernel(GaussArrays ga, int64_t n_items, int32_t list_base, int64_t item_base, FrameGeom g,
               const float4* __restrict__ recA, const float4* __restrict__ /*recB*/,
               const float4* __restrict__ recC, const int32_t* __restrict__ rad,
               const uint32_t* __restrict__ bin_off, uint32_t* __restrict__ bin_cursor,
               unsigned long long* __restrict__ keys, int64_t cap, int use_hist,
               const uint32_t* __restrict__ cta_hist, const uint32_t* __restrict__ chunk_mask, int32_t nchunks,
               const uint32_t* __restrict__ work, const uint32_t* __restrict__ work_count,
               uint32_t* __restrict__ ticket, const uint32_t* __restrict__ order) {
  constexpr int kItemsPerBlock = DYN ? kIpbDynamic : IPBS;  // this layer's chunk
  constexpr int NT = DYN ? kScatterThreadsDyn : kScatterThreadsStatic, kGroups = kItemsPerBlock / 256;
  constexpr int kItemsPerThread = kItemsPerBlock / NT;
  extern __shared__ uint32_t dsm[];
  __shared__ uint32_t s_work;
  const int nbl = g.T * g.NB;
  uint32_t* hist = dsm;         // counts, then running slot counters
  uint32_t* hbase = dsm + nbl;  // absolute start of this CTA's range in each bin
  int list, chunk;
  bool first = true;
#pragma unroll 1
  for (; next_chunk<WL>(list, chunk, first, work, work_count, ticket, &s_work, order); first = false) {
  const uint32_t gb = (uint32_t)(list_base + list) * (uint32_t)nbl;
  const int64_t base = (int64_t)chunk * kItemsPerBlock;
  const size_t chunk_id = (size_t)list * (size_t)nchunks + (size_t)chunk;
  // groups of the chunk that the preprocess pass found dead were never written: their items count as radius 0.
  // scatter_slot() gives every thread items of ONE group, threadIdx.x & (groups - 1).
  const bool glive = (chunk_mask[chunk_id] >> (threadIdx.x & (kGroups - 1))) & 1u;
  if (chunk_mask[chunk_id] == 0u) continue;  // whole chunk culled (CTA-uniform)
  if (use_hist && cta_hist != nullptr) {
    // the preprocess CTA that walked this chunk left its per-bin counts: no second counting pass
    const uint32_t* keep = cta_hist + chunk_id * (size_t)(nbl + 1);
    for (int k = threadIdx.x; k < nbl; k += blockDim.x) {
      const uint32_t c = keep[1 + k];
      if (c) hbase[k] = bin_off[gb + k] + atomicAdd(&bin_cursor[gb + k], c);
      hist[k] = 0u;
    }
    __syncthreads();
  } else if (use_hist) {
    for (int k = threadIdx.x; k < nbl; k += blockDim.x) hist[k] = 0u;
    __syncthreads();
#pragma unroll 1
    for (int it = 0; it < kItemsPerThread; ++it) {
      const int64_t i = base + scatter_slot<kGroups, NT>(it, threadIdx.x);
      if (i >= n_items || !glive) continue;
      const int64_t item = item_base + (int64_t)list * n_items + i;
      const int radius = rad[item];
      if (radius <= 0) continue;
      const float4 a = recA[4 * item];
      const float depth = recC[4 * item].w;
      const Rect rc = tile_rect(a.x, a.y, radius, g.gx, g.gy);
      const uint32_t bk = depth_bucket(depth, g);
      for (int y = rc.y0; y < rc.y1; ++y)
        for (int x = rc.x0; x < rc.x1; ++x) atomicAdd(&hist[(uint32_t)(y * g.gx + x) * (uint32_t)g.NB + bk], 1u);
    }  // (per-thread loops: this pass only runs when the preprocess kept no counts)
    __syncthreads();
    for (int k = threadIdx.x; k < nbl; k += blockDim.x) {
      const uint32_t c = hist[k];
      if (c) hbase[k] = bin_off[gb + k] + atomicAdd(&bin_cursor[gb + k], c);
      hist[k] = 0u;
    }
    __syncthreads();
  }
  // all loads of the thread's 8 items are issued before the first dependent use (8 independent DRAM round trips
  // in flight per thread instead of one), then the tile loops run from registers
  int radius[kItemsPerThread];
  float ax[kItemsPerThread], ay[kItemsPerThread], dep[kItemsPerThread];
#pragma unroll
  for (int it = 0; it < kItemsPerThread; ++it) {
    const int64_t i = base + scatter_slot<kGroups, NT>(it, threadIdx.x);
    radius[it] = (i < n_items && glive) ? rad[item_base + (int64_t)list * n_items + i] : 0;
  }
#pragma unroll
  for (int it = 0; it < kItemsPerThread; ++it) {
    ax[it] = ay[it] = dep[it] = 0.f;
    if (radius[it] > 0) {
      const int64_t item = item_base + (int64_t)list * n_items + base + scatter_slot<kGroups, NT>(it, threadIdx.x);
      const float4 a = recA[4 * item];
      ax[it] = a.x;
      ay[it] = a.y;
      dep[it] = recC[4 * item].w;
    }
  }
#pragma unroll
  for (int it = 0; it < kItemsPerThread; ++it) {
    Rect rc{0, 0, 0, 0};
    uint32_t bk = 0, klo = 0, khi = 0;
    if (radius[it] > 0) {
      const int64_t i = base + scatter_slot<kGroups, NT>(it, threadIdx.x);
      rc = tile_rect(ax[it], ay[it], radius[it], g.gx, g.gy);
      klo = ga.key_idx ? (uint32_t)ga.key_idx[i] : (uint32_t)i;
      khi = __float_as_uint(dep[it]);
      bk = depth_bucket(dep[it], g);
    }
    warp_for_each_bin(rc, bk, klo, khi, g.gx, g.NB, [&](uint32_t b, uint32_t lo, uint32_t hi) {
      uint32_t pos;
      if (use_hist)
        pos = hbase[b] + atomicAdd(&hist[b], 1u);
      else
        pos = bin_off[gb + b] + atomicAdd(&bin_cursor[gb + b], 1u);
      if ((int64_t)pos < cap) keys[pos] = ((unsigned long long)hi << 32) | lo;
    });
  }
  }  // rounds
}

// Scatter with exact tile culling: the preprocess pass left, per visible Gaussian, the mask of its alive tiles
// (first 64 tiles of the rectangle, recD); this pass walks the set bits.  Masks with more than kCoopTiles bits
// (kCoopBitsCull) and rectangles of more than 64 tiles (tested again, same make_foot / tile_alive on the same bits) are
// expanded by the whole warp.  Needs the preprocess CTA's bin counts (cta_hist) for the shared-memory slot
// counters; without them it falls back to global cursors.
template <bool DYN, bool WL, int IPBS = kIpbStatic>
__global__ void __launch_bounds__(DYN ? kScatterThreadsDyn : kScatterThreadsStatic)
scatter_cull_kernel(GaussArrays ga, int64_t n_items, int32_t list_base, int64_t item_base, FrameGeom g,
                    const float4* __restrict__ recA, const float4* __restrict__ recB,
                    const float4* __restrict__ recC, const int32_t* __restrict__ rad,
                    const uint32_t* __restrict__ bin_off, uint32_t* __restrict__ bin_cursor,
                    unsigned long long* __restrict__ keys, int64_t cap, int use_hist,
                    const uint32_t* __restrict__ cta_hist, const uint32_t* __restrict__ chunk_mask, int32_t nchunks,
                    const uint32_t* __restrict__ work, const uint32_t* __restrict__ work_count,
                    uint32_t* __restrict__ ticket, const uint32_t* __restrict__ order) {
  constexpr int kItemsPerBlock = DYN ? kIpbDynamic : IPBS;  // this layer's chunk
  constexpr int NT = DYN ? kScatterThreadsDyn : kScatterThreadsStatic, kGroups = kItemsPerBlock / 256;
  constexpr int kItemsPerThread = kItemsPerBlock / NT;
  extern __shared__ uint32_t dsm[];
  __shared__ uint32_t s_work;
  const int nbl = g.T * g.NB;
  uint32_t* hist = dsm;         // running slot counters
  uint32_t* hbase = dsm + nbl;  // absolute start of this CTA's range in each bin
  int list, chunk;
  bool first = true;
#pragma unroll 1
  for (; next_chunk<WL>(list, chunk, first, work, work_count, ticket, &s_work, order); first = false) {
  const uint32_t gb = (uint32_t)(list_base + list) * (uint32_t)nbl;
  const int64_t base = (int64_t)chunk * kItemsPerBlock;
  const size_t chunk_id = (size_t)list * (size_t)nchunks + (size_t)chunk;
  const bool glive = (chunk_mask[chunk_id] >> (threadIdx.x & (kGroups - 1))) & 1u;  // see scatter_kernel
  if (chunk_mask[chunk_id] == 0u) continue;  // whole chunk culled (CTA-uniform)
  if (use_hist) {
    const uint32_t* keep = cta_hist + chunk_id * (size_t)(nbl + 1);
    for (int k = threadIdx.x; k < nbl; k += blockDim.x) {
      const uint32_t c = keep[1 + k];
      if (c) hbase[k] = bin_off[gb + k] + atomicAdd(&bin_cursor[gb + k], c);
      hist[k] = 0u;
    }
    __syncthreads();
  }
  auto emit = [&](uint32_t b, uint32_t lo, uint32_t hi) {
    uint32_t pos;
    if (use_hist)
      pos = hbase[b] + atomicAdd(&hist[b], 1u);
    else
      pos = bin_off[gb + b] + atomicAdd(&bin_cursor[gb + b], 1u);
    if ((int64_t)pos < cap) keys[pos] = ((unsigned long long)hi << 32) | lo;
  };
  const int lane = threadIdx.x & 31;
  int radius[kItemsPerThread];
#pragma unroll
  for (int it = 0; it < kItemsPerThread; ++it) {
    const int64_t i = base + scatter_slot<kGroups, NT>(it, threadIdx.x);
    radius[it] = (i < n_items && glive) ? rad[item_base + (int64_t)list * n_items + i] : 0;
  }
  // The kernel is bound by the latency of its dependent loads (radius -> record -> key store), not by issue: the
  // records of item it + 1 are requested before item it is expanded (one item of look-ahead in registers; loading
  // all 8 ahead, as the plain scatter does, measured slower here: the loop is not unrolled, the batch went to
  // local memory).
  struct Pre {
    float4 a;
    float2 m;
    float depth;
    uint32_t key;
  };
  auto fetch = [&](int it) -> Pre {
    Pre p;
    p.a = make_float4(0.f, 0.f, 0.f, 0.f);
    p.m = make_float2(0.f, 0.f);
    p.depth = 0.f;
    p.key = 0u;
    if (it < kItemsPerThread && radius[it] > 0) {
      const int64_t i = base + scatter_slot<kGroups, NT>(it, threadIdx.x);
      const int64_t item = item_base + (int64_t)list * n_items + i;
      p.a = recA[4 * item];
      p.m = *reinterpret_cast<const float2*>(recA + 4 * item + 3);
      p.depth = recC[4 * item].w;
      p.key = ga.key_idx ? (uint32_t)ga.key_idx[i] : (uint32_t)i;
    }
    return p;
  };
  Pre nxt = fetch(0);
#pragma unroll 1
  for (int it = 0; it < kItemsPerThread; ++it) {
    const int64_t i = base + scatter_slot<kGroups, NT>(it, threadIdx.x);
    const int64_t item = item_base + (int64_t)list * n_items + i;
    const Pre cur = nxt;
    nxt = fetch(it + 1);
    Rect rc{0, 0, 0, 0};
    uint32_t bk = 0, klo = 0, khi = 0, mlo = 0, mhi = 0;
    const float4 a = cur.a;
    if (radius[it] > 0) {
      const float depth = cur.depth;
      rc = tile_rect(a.x, a.y, radius[it], g.gx, g.gy);
      klo = cur.key;
      khi = __float_as_uint(depth);
      bk = depth_bucket(depth, g);
      mlo = __float_as_uint(cur.m.x);
      mhi = __float_as_uint(cur.m.y);
    }
    const int w = rc.x1 - rc.x0, n = w * (rc.y1 - rc.y0);
    const bool huge = n > 64;                                          // test again, whole warp
    const bool wide = !huge && __popc(mlo) + __popc(mhi) > kCoopBitsCull;  // walk the mask, whole warp
    if (!huge && !wide) {
      if (w < 32) {
        // row by row: row dy of the rectangle holds bits [dy w, dy w + w) of the mask; the 64-bit mask is kept as a
        // two-word window that slides down by w per row (no 64-bit shifts, no division per bit)
        uint32_t lo = mlo, hi = mhi;
        const uint32_t rowmask = (1u << w) - 1u;
        uint32_t tile_row = (uint32_t)(rc.y0 * g.gx + rc.x0);
        while (lo | hi) {
          uint32_t row = lo & rowmask;
          while (row) {
            const uint32_t dx = (uint32_t)__ffs((int)row) - 1u;
            row &= row - 1u;
            emit((tile_row + dx) * (uint32_t)g.NB + bk, klo, khi);
          }
          lo = __funnelshift_r(lo, hi, (uint32_t)w);
          hi >>= w;
          tile_row += (uint32_t)g.gx;
        }
      } else {
        unsigned long long mask = ((unsigned long long)mhi << 32) | mlo;
        while (mask) {
          const int bit = __ffsll((long long)mask) - 1;
          mask &= mask - 1;
          const int dy = bit / w;
          emit((uint32_t)((rc.y0 + dy) * g.gx + rc.x0 + (bit - dy * w)) * (uint32_t)g.NB + bk, klo, khi);
        }
      }
    }
    unsigned m = __ballot_sync(0xffffffffu, huge || wide);
    while (m) {
      const int src = __ffs(m) - 1;
      m &= m - 1;
      const int x0 = __shfl_sync(0xffffffffu, rc.x0, src), y0 = __shfl_sync(0xffffffffu, rc.y0, src);
      const int ww = __shfl_sync(0xffffffffu, w, src), nn = __shfl_sync(0xffffffffu, n, src);
      const uint32_t b = __shfl_sync(0xffffffffu, bk, src);
      const uint32_t lo = __shfl_sync(0xffffffffu, klo, src), hi = __shfl_sync(0xffffffffu, khi, src);
      if (nn <= 64) {  // warp-uniform
        const uint32_t q0 = __shfl_sync(0xffffffffu, mlo, src), q1 = __shfl_sync(0xffffffffu, mhi, src);
        for (int t0 = 0; t0 < nn; t0 += 32) {
          const int t = t0 + lane;
          if (t < nn && (((t0 ? q1 : q0) >> lane) & 1u)) {
            const int dy = t / ww;
            emit((uint32_t)((y0 + dy) * g.gx + x0 + (t - dy * ww)) * (uint32_t)g.NB + b, lo, hi);
          }
        }
      } else {
        Foot fo{};
        if (lane == src) {
          const float4 cb = recB[4 * item];
          fo = make_foot(a.x, a.y, a.z, a.w, cb.x, cb.y, cb.z, cb.w);
        }
        Foot fs;
        fs.px = __shfl_sync(0xffffffffu, fo.px, src); fs.py = __shfl_sync(0xffffffffu, fo.py, src);
        fs.hx = __shfl_sync(0xffffffffu, fo.hx, src); fs.hy = __shfl_sync(0xffffffffu, fo.hy, src);
        fs.A = __shfl_sync(0xffffffffu, fo.A, src);   fs.B = __shfl_sync(0xffffffffu, fo.B, src);
        fs.C = __shfl_sync(0xffffffffu, fo.C, src);   fs.t = __shfl_sync(0xffffffffu, fo.t, src);
        fs.kyx = __shfl_sync(0xffffffffu, fo.kyx, src); fs.kxy = __shfl_sync(0xffffffffu, fo.kxy, src);
        Rect full;
        full.x0 = x0; full.y0 = y0; full.x1 = x0 + ww; full.y1 = y0 + nn / ww;
        const Rect q = alive_rect(fs, full);
        const int qw = q.x1 - q.x0, qn = qw * (q.y1 - q.y0);
        for (int t = lane; t < qn; t += 32) {
          const int dy = t / qw;
          const int x = q.x0 + (t - dy * qw), y = q.y0 + dy;
          if (tile_alive(fs, x, y)) emit((uint32_t)(y * g.gx + x) * (uint32_t)g.NB + b, lo, hi);
        }
      }
    }
  }
  }  // rounds
}

// Radius of original Gaussian i in frame f (view v), 0 for items of groups the preprocess pass found dead (those
// are never written; chunk_mask = live groups per (list, chunk): V static lists, then F dynamic lists).
__device__ __forceinline__ int radius_of(int64_t i, int32_t f, int32_t v, int32_t V, int64_t Ns, int64_t Nm,
                                         const int32_t* slot, const int32_t* pos_of, const int32_t* rad,
                                         const uint32_t* chunk_mask, int64_t* item_out, int32_t ipb_s) {
  const int64_t pos = pos_of ? (int64_t)pos_of[i] : i;
  const bool dyn = slot != nullptr && slot[i] >= 0;
  const int64_t nch_s = (Ns + ipb_s - 1) / ipb_s, nch_d = (Nm + kIpbDynamic - 1) / kIpbDynamic;
  const int64_t item = dyn ? (int64_t)V * Ns + (int64_t)f * Nm + pos : (int64_t)v * Ns + pos;
  const int64_t ipb = dyn ? kIpbDynamic : ipb_s;
  const int64_t cid = (dyn ? (int64_t)V * nch_s + (int64_t)f * nch_d : (int64_t)v * nch_s) + pos / ipb;
  *item_out = item;
  if (chunk_mask != nullptr && !((chunk_mask[cid] >> ((pos % ipb) >> 8)) & 1u)) return 0;
  return rad[item];
}

// Per-Gaussian taps of one frame, addressed by original index (parity tests / radii output).
__global__ void frame_attrs_kernel(int64_t N, int64_t Ns, int64_t Nm, int32_t V, int32_t frame,
                                   const int32_t* frame_view, const int32_t* slot, const int32_t* pos_of, FrameGeom g,
                                   const float4* recA, const float4* recB, const float4* recC, const int32_t* rad,
                                   const uint32_t* chunk_mask, int32_t* radii, int32_t* tiles, float* depth, float* xy,
                                   float* conic_o, float* rgb, int32_t ipb_s) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  const int view = frame_view ? frame_view[frame] : 0;
  int64_t item;
  const int r = radius_of(i, frame, view, V, Ns, Nm, slot, pos_of, rad, chunk_mask, &item, ipb_s);
  float4 a = make_float4(0, 0, 0, 0), b = a, c = a;
  int tt = 0;
  if (r > 0) {
    a = recA[4 * item];
    b = recB[4 * item];
    c = recC[4 * item];
    const Rect rc = tile_rect(a.x, a.y, r, g.gx, g.gy);
    tt = (rc.y1 - rc.y0) * (rc.x1 - rc.x0);
  }
  if (radii) radii[i] = r;
  if (tiles) tiles[i] = tt;
  if (depth) depth[i] = c.w;
  if (xy) {
    xy[2 * i] = a.x;
    xy[2 * i + 1] = a.y;
  }
  if (conic_o) {
    conic_o[4 * i] = b.x; conic_o[4 * i + 1] = b.y; conic_o[4 * i + 2] = b.z; conic_o[4 * i + 3] = b.w;
  }
  if (rgb) {
    rgb[3 * i] = c.x; rgb[3 * i + 1] = c.y; rgb[3 * i + 2] = c.z;
  }
}

// radii of every frame: out[f, i]
__global__ void radii_kernel(int64_t N, int64_t Ns, int64_t Nm, int32_t V, const int32_t* frame_view,
                             const int32_t* slot, const int32_t* pos_of, const int32_t* rad,
                             const uint32_t* chunk_mask, int32_t* out, int32_t ipb_s) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  const int f = blockIdx.y;
  const int view = frame_view ? frame_view[f] : 0;
  int64_t item;
  out[(int64_t)f * N + i] = radius_of(i, f, view, V, Ns, Nm, slot, pos_of, rad, chunk_mask, &item, ipb_s);
}

__global__ void mark_visible_kernel(int32_t P, const float* xyz, const float* V, float near_plane, uint8_t* out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const float x = xyz[3 * i], y = xyz[3 * i + 1], z = xyz[3 * i + 2];
  const float pvz = V[2] * x + V[6] * y + V[10] * z + V[14];
  out[i] = pvz > near_plane ? 1 : 0;
}

// compact, re-ordered copies of the two layers, made once per scene upload
__global__ void gather_static_kernel(int64_t Ns, const int32_t* idx, GaussArrays sc, float* xyz, float* scale,
                                     float* quat, float* opac) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= Ns) return;
  const int64_t i = idx[j];
  for (int k = 0; k < 3; ++k) {
    xyz[3 * j + k] = sc.xyz[3 * i + k];
    scale[3 * j + k] = sc.scale_act[3 * i + k];
  }
  for (int k = 0; k < 4; ++k) quat[4 * j + k] = sc.quat_act[4 * i + k];
  opac[j] = sc.opacity[i];
}

__global__ void gather_movable_kernel(int64_t Nm, const int32_t* mov_idx, GaussArrays sc, float* xyz, float* scale,
                                      float* opac, float* sh, float* ls, float* qr, int32_t* mslot) {
  const int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (m >= Nm) return;
  const int64_t i = mov_idx[m];
  for (int k = 0; k < 3; ++k) {
    xyz[3 * m + k] = sc.xyz[3 * i + k];
    scale[3 * m + k] = sc.scale_act[3 * i + k];
    ls[3 * m + k] = sc.log_scale_raw[3 * i + k];
  }
  for (int k = 0; k < 4; ++k) qr[4 * m + k] = sc.quat_raw[4 * i + k];
  opac[m] = sc.opacity[i];
  for (int k = 0; k < 48; ++k) sh[48 * m + k] = sc.sh[48 * i + k];
  mslot[m] = sc.slot[i];
}

// One CTA (256 threads) per 256-item group of a layer's compact arrays: bounding sphere of the centres,
// sigma = max sqrt(trace Sigma) (static: from the activated scale and quaternion; movable: sqrt(sum s^2),
// the transformed quaternion is re-normalised), smax = largest scale, slot = the group's slot or -1.
__global__ void __launch_bounds__(256)
group_bounds_kernel(int64_t n, const float* __restrict__ xyz, const float* __restrict__ scale,
                    const float* __restrict__ quat /*static only*/, const int32_t* __restrict__ mslot /*movable only*/,
                    float* __restrict__ out) {
  __shared__ float red[8][8];
  __shared__ float cen[3];
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  const bool ok = i < n;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  float p[3] = {0.f, 0.f, 0.f}, sig = 0.f, smax = 0.f;
  int sl = -2;
  if (ok) {
    for (int k = 0; k < 3; ++k) p[k] = xyz[3 * i + k];
    const float s0 = scale[3 * i], s1 = scale[3 * i + 1], s2 = scale[3 * i + 2];
    smax = fmaxf(s0, fmaxf(s1, s2));
    if (quat != nullptr) {
      const float r = quat[4 * i], x = quat[4 * i + 1], y = quat[4 * i + 2], z = quat[4 * i + 3];
      const float R[3][3] = {{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
                             {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
                             {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}};
      const float sc[3] = {s0, s1, s2};
      float tr = 0.f;
      for (int c = 0; c < 3; ++c)
        for (int rr = 0; rr < 3; ++rr) tr += (sc[rr] * R[c][rr]) * (sc[rr] * R[c][rr]);
      sig = sqrtf(tr);
    } else {
      sig = sqrtf(s0 * s0 + s1 * s1 + s2 * s2);
    }
    if (mslot != nullptr) sl = mslot[i];
  }
  // AABB -> centre
  float mn[3], mx[3];
  for (int k = 0; k < 3; ++k) {
    mn[k] = ok ? p[k] : 3.0e38f;
    mx[k] = ok ? p[k] : -3.0e38f;
    for (int o = 16; o > 0; o >>= 1) {
      mn[k] = fminf(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], o));
      mx[k] = fmaxf(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], o));
    }
  }
  if (lane == 0)
    for (int k = 0; k < 3; ++k) {
      red[w][k] = mn[k];
      red[w][3 + k] = mx[k];
    }
  __syncthreads();
  if (threadIdx.x < 3) {
    float a = 3.0e38f, b = -3.0e38f;
    for (int q = 0; q < 8; ++q) {
      a = fminf(a, red[q][threadIdx.x]);
      b = fmaxf(b, red[q][3 + threadIdx.x]);
    }
    cen[threadIdx.x] = 0.5f * (a + b);
  }
  __syncthreads();
  float d = 0.f;
  if (ok) {
    const float dx = p[0] - cen[0], dy = p[1] - cen[1], dz = p[2] - cen[2];
    d = sqrtf(dx * dx + dy * dy + dz * dz);
  }
  int smin = ok ? sl : 0x7fffffff, smaxi = ok ? sl : -0x7fffffff;
  for (int o = 16; o > 0; o >>= 1) {
    d = fmaxf(d, __shfl_xor_sync(0xffffffffu, d, o));
    sig = fmaxf(sig, __shfl_xor_sync(0xffffffffu, sig, o));
    smax = fmaxf(smax, __shfl_xor_sync(0xffffffffu, smax, o));
    smin = min(smin, __shfl_xor_sync(0xffffffffu, smin, o));
    smaxi = max(smaxi, __shfl_xor_sync(0xffffffffu, smaxi, o));
  }
  __syncthreads();
  if (lane == 0) {
    red[w][0] = d;
    red[w][1] = sig;
    red[w][2] = smax;
    red[w][3] = __int_as_float(smin);
    red[w][4] = __int_as_float(smaxi);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float D = 0.f, S = 0.f, M = 0.f;
    int a = 0x7fffffff, b = -0x7fffffff;
    for (int q = 0; q < 8; ++q) {
      D = fmaxf(D, red[q][0]);
      S = fmaxf(S, red[q][1]);
      M = fmaxf(M, red[q][2]);
      a = min(a, __float_as_int(red[q][3]));
      b = max(b, __float_as_int(red[q][4]));
    }
    float* o = out + (int64_t)blockIdx.x * 8;
    o[0] = cen[0];
    o[1] = cen[1];
    o[2] = cen[2];
    o[3] = D * 1.0001f + 1e-6f;
    o[4] = S * 1.0001f;
    o[5] = M * 1.0001f;
    o[6] = (mslot != nullptr && a == b) ? (float)a : -1.0f;
    o[7] = 0.f;
  }
}

}  // namespace

namespace {
// Dispatch order of a layer's (list, chunk) CTAs: heaviest first.  A chunk's weight is the number of tile entries it
// produced (written by its preprocess CTA); a close-up of the gripper in a wrist camera makes single chunks hundreds
// of times heavier than the rest, and with CTAs dispatched in index order those happened to come last: the SMs sat
// idle for a quarter (preprocess) to a half (scatter) of the kernels' duration (ncu: sm cycles active / elapsed).
// Counting sort by weight class (two classes per power of two, descending), one CTA; order inside a class is
// arbitrary, which is harmless: the order only decides WHEN a chunk is processed.  The scatter pass of the same step
// uses it, and so does the preprocess pass of the NEXT step (poses change little from step to step).
constexpr int kCostClasses = 66;
__global__ void __launch_bounds__(1024)
order_chunks_kernel(const uint32_t* __restrict__ cost, uint32_t n, uint32_t* __restrict__ order) {
  __shared__ uint32_t cnt[kCostClasses], start[kCostClasses];
  auto cls = [](uint32_t c) -> int {
    if (c == 0u) return 0;
    const int l = 31 - __clz(c);                          // floor(log2 c)
    const int half = l > 0 ? (int)((c >> (l - 1)) & 1u) : 0;
    return 1 + 2 * l + half;                              // 1 .. 64
  };
  for (int k = threadIdx.x; k < kCostClasses; k += blockDim.x) cnt[k] = 0u;
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) atomicAdd(&cnt[cls(cost[i])], 1u);
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t run = 0;
    for (int k = kCostClasses - 1; k >= 0; --k) {  // heaviest class first
      start[k] = run;
      run += cnt[k];
    }
  }
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) order[atomicAdd(&start[cls(cost[i])], 1u)] = i;
}

template <auto Kernel>
int ensure_smem(size_t bytes, int device) {
  static size_t granted[64] = {};  // per kernel and device: raise the opt-in shared-memory limit only when it grows
  size_t& g = granted[device & 63];
  if (bytes > 48 * 1024 && bytes > g) {
    GSW_CUDA(cudaFuncSetAttribute(Kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    g = bytes;
  }
  return 0;
}
}  // namespace

// The dynamic layer is arithmetic-bound (transform + SH for nearly every item), not load-latency-bound: it
// keeps plain loads and the larger chunk; only the static layer stages through TMA.
// Words of Workspace::status (u32 view) used by the static layer's work list: entries, and one ticket per pass.
constexpr int kWorkCountWord = 20, kPreTicketWord = 21, kScatterTicketWord = 22;

// tuning aid: persistent CTAs per SM of the static layer's work-list preprocess (GSW_WL_CTAS; 0 = as many as fit)
static int wl_ctas_per_sm() {
  static const int v = [] {
    const char* e = getenv("GSW_WL_CTAS");
    return e ? atoi(e) : 0;
  }();
  return v;
}

int64_t layer_chunks(const gsw_ctx* ctx, int64_t n_items, bool dynamic) {
  const int64_t ipb = dynamic ? kIpbDynamic : ctx->ipb_static;
  return (n_items + ipb - 1) / ipb;
}

// The static layer goes through the chunk-cull work list when its groups have bounds (scene layers; the single-frame
// entry points have none) and (view, chunk) fits the 16 + 16 bit packing.
bool use_work_list(const gsw_ctx* ctx, bool dynamic, const GaussArrays& ga, int64_t n_items, int32_t n_lists) {
  return !dynamic && ga.bounds != nullptr && n_lists <= 65535 && layer_chunks(ctx, n_items, false) <= 65535 &&
         getenv("GSW_NO_WORK_LIST") == nullptr;
}

// Items per chunk of the static layer for this launch.  A chunk is the unit the persistent preprocess / scatter CTAs
// take from the work list, and one CTA walks it 256 items at a time: with two views a 2048-item chunk leaves a few
// hundred CTAs each walking eight dependent rounds (measured, 1M Gaussians: static preprocess 0.10 ms at 2048 items
// per chunk, 0.056 ms at 256; step 1.070 -> 1.017 ms at 64 envs and 0.394 -> 0.362 ms at 8 envs), while with 65 views
// small chunks only add histogram flushes and per-chunk count rows (per-env cameras 2.95 ms at 2048, 3.40 ms at 256).
// So: the largest chunk that still yields ~7000 (view, chunk) pairs, 256 at least; always 2048 without a work list
// (the staged and the plain forms are only instantiated for it).
int pick_static_ipb(const GaussArrays& ga, int64_t n_items, int32_t n_lists, int32_t bins_per_list) {
  if (ga.bounds == nullptr || n_lists > 65535 || getenv("GSW_NO_WORK_LIST") != nullptr ||
      getenv("GSW_WL_TMA") != nullptr)
    return kIpbStatic;
  if (const char* e = getenv("GSW_IPB_STATIC")) {  // tuning aid
    const int v = atoi(e);
    if ((v == 256 || v == 512 || v == 1024 || v == 2048) && (n_items + v - 1) / v <= 65535) return v;
  }
  int ipb = 256;
  for (int c = kIpbStatic; c > 256; c >>= 1)
    if ((int64_t)n_lists * ((n_items + c - 1) / c) >= 7000) {
      ipb = c;
      break;
    }
  // 16-bit chunk field of the work list; and every chunk keeps a row of per-bin counts for the scatter pass (cta_hist):
  // small chunks on a large tile grid must not turn that into gigabytes
  while (ipb < kIpbStatic && ((n_items + ipb - 1) / ipb > 65535 ||
                              (int64_t)n_lists * ((n_items + ipb - 1) / ipb) * (bins_per_list + 1) * 4 > ((int64_t)64 << 20)))
    ipb <<= 1;
  return ipb;
}

int launch_preprocess(const gsw_ctx* ctx, bool dynamic, const GaussArrays& ga, int64_t n_items, int32_t n_lists,
                      int32_t list_base, int64_t item_base, const FrameGeom& g, const float* view, const float* proj,
                      const float* campos, const float* tanfov, const int32_t* frame_env, const int32_t* frame_view,
                      const float* pose, int32_t num_slots, uint32_t* cta_hist, uint32_t* chunk_mask, int* kept_hist,
                      cudaStream_t s, const uint32_t* order, uint32_t* cost) {
  if (kept_hist) *kept_hist = 0;
  if (n_items <= 0 || n_lists <= 0) return 0;
  auto& w = ctx->ws;
  const size_t hist_b = (size_t)g.T * g.NB * 4;
  const size_t pose_b = dynamic ? (size_t)num_slots * GSW_POSE_STRIDE * sizeof(float) : 0;
  auto aligned16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  // TMA staging needs 16 B aligned sources and the scale/rotation input form
  bool tma_ok = aligned16(ga.xyz) && aligned16(ga.scale_act) && aligned16(ga.opacity) && ga.scale_act != nullptr;
  if (dynamic)
    tma_ok = tma_ok && aligned16(ga.log_scale_raw) && aligned16(ga.quat_raw) && aligned16(ga.mov_slot);
  else
    tma_ok = tma_ok && ga.quat_act != nullptr && aligned16(ga.quat_act) && (ga.slot == nullptr || aligned16(ga.slot));
  const size_t stage_b = dynamic ? (size_t)kIpbDynamic * 60 : (size_t)ctx->ipb_static * 48;
  const size_t budget = 200 * 1024;
  // The bulk staging pays where a CTA's arithmetic is short next to its load latency (few visible items per chunk).
  // The work-list form only visits live chunks; there the 96 KB stage costs a third of the resident warps and
  // measured slower (per-env cameras: 1.08 -> 0.90 ms), so it keeps plain loads.  GSW_WL_TMA=1 / GSW_NO_TMA=1 override.
  const bool wl_pre = use_work_list(ctx, dynamic, ga, n_items, n_lists);
  int use_tma = !dynamic && tma_ok && stage_b + pose_b + 1024 <= budget && getenv("GSW_NO_TMA") == nullptr &&
                (!wl_pre || getenv("GSW_WL_TMA") != nullptr);
  int use_hist = hist_b <= kMaxHistBytes && (use_tma ? stage_b : 0) + hist_b + pose_b <= budget;
  if (!use_hist && use_tma && hist_b <= kMaxHistBytes) {  // prefer the histogram over staging when both do not fit
    use_tma = 0;
    use_hist = 1;
  }
  const size_t sm = (use_tma ? stage_b : 0) + (use_hist ? hist_b : 0) + pose_b;
  const int32_t nchunks = (int32_t)layer_chunks(ctx, n_items, dynamic);
  const bool wl = use_work_list(ctx, dynamic, ga, n_items, n_lists);
  uint32_t* const st32 = w.status.as<uint32_t>();
  if (wl) {
    dim3 cg((unsigned)((nchunks + 255) / 256), (unsigned)n_lists);
    cull_chunks_kernel<<<cg, 256, 0, s>>>(ga.bounds, n_items, nchunks, g, view, proj, campos, tanfov, chunk_mask,
                                          w.work.as<uint32_t>(), st32 + kWorkCountWord,
                                          w.view_consts.as<uint32_t>(),
                                          ctx->cache_on ? w.cache_m.as<uint32_t>() : nullptr,
                                          w.bin_count.as<uint32_t>(), w.bin_cursor.as<uint32_t>(), g.T * g.NB,
                                          ctx->ipb_static);
    count_launches(1);
  }
  int rc;
#define GSW_LAUNCH_PRE(DYN_, IPB_, TMA_, CULL_, WL_)                                                              \
  do {                                                                                                            \
    const int64_t pairs = (int64_t)nchunks * n_lists;                                                             \
    const int64_t persistent = 148 * (wl_ctas_per_sm() > 0 ? wl_ctas_per_sm() : ((TMA_) ? 2 : 3));               \
    dim3 grid((unsigned)nchunks, (unsigned)n_lists);                                                              \
    if (WL_) grid = dim3((unsigned)(pairs < persistent ? pairs : persistent), 1u);                               \
    if ((rc = ensure_smem<preprocess_kernel<DYN_, IPB_, TMA_, CULL_, WL_>>(sm, ctx->device))) return rc;          \
    preprocess_kernel<DYN_, IPB_, TMA_, CULL_, WL_><<<grid, 256, sm, s>>>(                                        \
        ga, n_items, list_base, item_base, g, view, proj, campos, tanfov, frame_env, frame_view, pose, num_slots, \
        w.rec.as<float4>(), w.rec.as<float4>() + 1, w.rec.as<float4>() + 2, w.rad.as<int32_t>(),                  \
        w.bin_count.as<uint32_t>(), use_hist, use_hist ? cta_hist : nullptr, chunk_mask, nchunks,                 \
        w.work.as<uint32_t>(), st32 + kWorkCountWord, st32 + kPreTicketWord, w.view_consts.as<uint32_t>(),       \
        order, (use_hist && cta_hist != nullptr) ? cost : nullptr,                                                 \
        w.status.as<unsigned long long>() + (dynamic ? 17 : 16));                                                  \
  } while (0)
  const bool cull = g.tile_cull != 0;
  if (dynamic) {
    if (cull) GSW_LAUNCH_PRE(true, kIpbDynamic, false, true, false);
    else GSW_LAUNCH_PRE(true, kIpbDynamic, false, false, false);
  } else if (use_tma) {
    if (wl) {
      if (cull) GSW_LAUNCH_PRE(false, kIpbStatic, true, true, true);
      else GSW_LAUNCH_PRE(false, kIpbStatic, true, false, true);
    } else {
      if (cull) GSW_LAUNCH_PRE(false, kIpbStatic, true, true, false);
      else GSW_LAUNCH_PRE(false, kIpbStatic, true, false, false);
    }
  } else {
    if (wl) {  // the form whose chunk size follows the launch (pick_static_ipb)
      switch (ctx->ipb_static) {
        case 256:
          if (cull) GSW_LAUNCH_PRE(false, 256, false, true, true);
          else GSW_LAUNCH_PRE(false, 256, false, false, true);
          break;
        case 512:
          if (cull) GSW_LAUNCH_PRE(false, 512, false, true, true);
          else GSW_LAUNCH_PRE(false, 512, false, false, true);
          break;
        case 1024:
          if (cull) GSW_LAUNCH_PRE(false, 1024, false, true, true);
          else GSW_LAUNCH_PRE(false, 1024, false, false, true);
          break;
        default:
          if (cull) GSW_LAUNCH_PRE(false, kIpbStatic, false, true, true);
          else GSW_LAUNCH_PRE(false, kIpbStatic, false, false, true);
      }
    } else {
      if (cull) GSW_LAUNCH_PRE(false, kIpbStatic, false, true, false);
      else GSW_LAUNCH_PRE(false, kIpbStatic, false, false, false);
    }
  }
#undef GSW_LAUNCH_PRE
  if (kept_hist) *kept_hist = use_hist && cta_hist != nullptr;
  count_launches(1);
  GSW_CUDA(cudaGetLastError());
  return 0;
}

// words of cta_hist one layer needs (0: the bins do not fit a shared-memory histogram, nothing is kept)
size_t cta_hist_words(const gsw_ctx* ctx, int64_t n_items, int32_t n_lists, const FrameGeom& g, bool dynamic) {
  if (n_items <= 0 || n_lists <= 0 || (size_t)g.T * g.NB * 4 > kMaxHistBytes) return 0;
  return (size_t)n_lists * (size_t)layer_chunks(ctx, n_items, dynamic) * (size_t)(g.T * g.NB + 1);
}

int launch_scatter(const gsw_ctx* ctx, bool dynamic, const GaussArrays& ga, int64_t n_items, int32_t n_lists,
                   int32_t list_base, int64_t item_base, const FrameGeom& g, int64_t cap, const uint32_t* cta_hist,
                   const uint32_t* chunk_mask, cudaStream_t s, const uint32_t* order) {
  if (n_items <= 0 || n_lists <= 0) return 0;
  const int32_t nchunks = (int32_t)layer_chunks(ctx, n_items, dynamic);
  const bool wl = use_work_list(ctx, dynamic, ga, n_items, n_lists);
  const int64_t pairs = (int64_t)nchunks * n_lists, persistent = 148 * 4;
  dim3 grid((unsigned)nchunks, (unsigned)n_lists);
  if (wl) grid = dim3((unsigned)(pairs < persistent ? pairs : persistent), 1u);
  auto& w = ctx->ws;
  uint32_t* const st32 = w.status.as<uint32_t>();
  const size_t hist_b = (size_t)g.T * g.NB * 8;
  const int use_hist = hist_b <= kMaxHistBytes;
  const int ipbs = ctx->ipb_static;
  int rc;
#define GSW_LAUNCH_SCATTER(KERNEL_, HIST_, CH_)                                                                     \
  do {                                                                                                            \
    if ((rc = ensure_smem<KERNEL_>((HIST_) ? hist_b : 0, ctx->device))) return rc;                                \
    KERNEL_<<<grid, dynamic ? kScatterThreadsDyn : kScatterThreadsStatic, (HIST_) ? hist_b : 0, s>>>(                                                              \
        ga, n_items, list_base, item_base, g, w.rec.as<float4>(), w.rec.as<float4>() + 1, w.rec.as<float4>() + 2, \
        w.rad.as<int32_t>(), w.bin_off.as<uint32_t>(), w.bin_cursor.as<uint32_t>(),                               \
        w.keys.as<unsigned long long>(), cap, (HIST_) ? 1 : 0, CH_, chunk_mask, nchunks, w.work.as<uint32_t>(),   \
        st32 + kWorkCountWord, st32 + kScatterTicketWord, order);                                                 \
  } while (0)
#define GSW_COMMA ,
  if (g.tile_cull != 0) {  // the culling form has no counting pass of its own: shared counters need cta_hist
    const bool h = use_hist && cta_hist != nullptr;
    if (dynamic) GSW_LAUNCH_SCATTER(scatter_cull_kernel<true GSW_COMMA false>, h, cta_hist);
    else if (wl && ipbs == 256) GSW_LAUNCH_SCATTER(scatter_cull_kernel<false GSW_COMMA true GSW_COMMA 256>, h, cta_hist);
    else if (wl && ipbs == 512) GSW_LAUNCH_SCATTER(scatter_cull_kernel<false GSW_COMMA true GSW_COMMA 512>, h, cta_hist);
    else if (wl && ipbs == 1024) GSW_LAUNCH_SCATTER(scatter_cull_kernel<false GSW_COMMA true GSW_COMMA 1024>, h, cta_hist);
    else if (wl) GSW_LAUNCH_SCATTER(scatter_cull_kernel<false GSW_COMMA true>, h, cta_hist);
    else GSW_LAUNCH_SCATTER(scatter_cull_kernel<false GSW_COMMA false>, h, cta_hist);
  } else {
    if (dynamic) GSW_LAUNCH_SCATTER(scatter_kernel<true GSW_COMMA false>, use_hist, cta_hist);
    else if (wl && ipbs == 256) GSW_LAUNCH_SCATTER(scatter_kernel<false GSW_COMMA true GSW_COMMA 256>, use_hist, cta_hist);
    else if (wl && ipbs == 512) GSW_LAUNCH_SCATTER(scatter_kernel<false GSW_COMMA true GSW_COMMA 512>, use_hist, cta_hist);
    else if (wl && ipbs == 1024) GSW_LAUNCH_SCATTER(scatter_kernel<false GSW_COMMA true GSW_COMMA 1024>, use_hist, cta_hist);
    else if (wl) GSW_LAUNCH_SCATTER(scatter_kernel<false GSW_COMMA true>, use_hist, cta_hist);
    else GSW_LAUNCH_SCATTER(scatter_kernel<false GSW_COMMA false>, use_hist, cta_hist);
  }
#undef GSW_COMMA
#undef GSW_LAUNCH_SCATTER
  count_launches(1);
  GSW_CUDA(cudaGetLastError());
  return 0;
}

int launch_prepare_views(const gsw_ctx* ctx, int32_t V, int32_t F, const FrameGeom& g, const float* view,
                         const float* proj, const float* campos, const float* tanfov, const int32_t* frame_view,
                         cudaStream_t s) {
  auto& w = ctx->ws;
  GSW_CUDA(cudaMemsetAsync(w.view_cnt.p, 0, (size_t)V * 4, s));
  const int blocks = V > (F + 255) / 256 ? V : (F + 255) / 256;
  prepare_views_kernel<<<blocks, 256, 0, s>>>(V, F, g, view, proj, campos, tanfov, frame_view,
                                              w.view_consts.as<uint32_t>(), w.view_cnt.as<uint32_t>(),
                                              ctx->cache_ok ? 1 : 0,
                                              ctx->cache_on ? w.view_prev.as<uint32_t>() : nullptr,
                                              w.view_same.as<uint32_t>());
  count_launches(1);
  GSW_CUDA(cudaGetLastError());
  return 0;
}

// m = number of leading views whose static lists are kept from the previous launch (ws.cache_m); after prepare_views
int launch_cache_prefix(const gsw_ctx* ctx, int32_t V, cudaStream_t s) {
  auto& w = ctx->ws;
  cache_prefix_kernel<<<1, 32, 0, s>>>(V, w.view_same.as<uint32_t>(), w.view_cnt.as<uint32_t>(),
                                       w.view_cnt_prev.as<uint32_t>(), w.cache_m.as<uint32_t>(),
                                       w.status.as<long long>());
  count_launches(1);
  GSW_CUDA(cudaGetLastError());
  return 0;
}

int launch_order_chunks(const uint32_t* cost, int64_t n, uint32_t* order, cudaStream_t s) {
  if (n <= 0) return 0;
  order_chunks_kernel<<<1, 1024, 0, s>>>(cost, (uint32_t)n, order);
  count_launches(1);
  GSW_CUDA(cudaGetLastError());
  return 0;
}

int launch_frame_attrs(const gsw_ctx* ctx, int32_t frame, int32_t* radii, int32_t* tiles_touched, float* depth,
                       float* xy, float* conic_opacity, float* rgb, cudaStream_t s) {
  const auto& L = ctx->last;
  auto& w = ctx->ws;
  if (L.N <= 0) return 0;
  frame_attrs_kernel<<<(unsigned)((L.N + 255) / 256), 256, 0, s>>>(
      L.N, L.Ns, L.Nm, L.V, frame, L.frame_view, L.slot_orig, L.pos_of, L.g, w.rec.as<float4>(),
      w.rec.as<float4>() + 1, w.rec.as<float4>() + 2, w.rad.as<int32_t>(), w.chunk_mask.as<uint32_t>(), radii,
      tiles_touched, depth, xy, conic_opacity, rgb, ctx->ipb_static);
  count_launches(1);
  GSW_CUDA(cudaGetLastError());
  return 0;
}

int launch_radii(const gsw_ctx* ctx, int32_t* out_radii, cudaStream_t s) {
  const auto& L = ctx->last;
  auto& w = ctx->ws;
  if (L.N <= 0 || L.F <= 0) return 0;
  dim3 grid((unsigned)((L.N + 255) / 256), (unsigned)L.F);
  radii_kernel<<<grid, 256, 0, s>>>(L.N, L.Ns, L.Nm, L.V, L.frame_view, L.slot_orig, L.pos_of, w.rad.as<int32_t>(),
                                    w.chunk_mask.as<uint32_t>(), out_radii, ctx->ipb_static);
  count_launches(1);
  GSW_CUDA(cudaGetLastError());
  return 0;
}

int launch_mark_visible(int32_t P, const float* means3D, const float* viewmatrix, float near_plane, uint8_t* out,
                        cudaStream_t s) {
  if (P <= 0) return 0;
  mark_visible_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, means3D, viewmatrix, near_plane, out);
  count_launches(1);
  GSW_CUDA(cudaGetLastError());
  return 0;
}

// `src` = the caller's scene arrays in original order (xyz, scale_act, quat_act, opacity, sh, raw arrays, slot)
int launch_gather_layers(const gsw_ctx* ctx, cudaStream_t s) {
  const GaussArrays& src = ctx->scene;
  if (ctx->Ns > 0) {
    const unsigned nb = (unsigned)((ctx->Ns + 255) / 256);
    gather_static_kernel<<<nb, 256, 0, s>>>(ctx->Ns, ctx->st_idx.as<int32_t>(), src, ctx->st_xyz.as<float>(),
                                            ctx->st_scale.as<float>(), ctx->st_quat.as<float>(),
                                            ctx->st_opacity.as<float>());
    group_bounds_kernel<<<nb, 256, 0, s>>>(ctx->Ns, ctx->st_xyz.as<float>(), ctx->st_scale.as<float>(),
                                           ctx->st_quat.as<float>(), nullptr, ctx->st_bounds.as<float>());
    count_launches(2);
  }
  if (ctx->Nm > 0) {
    const unsigned nb = (unsigned)((ctx->Nm + 255) / 256);
    gather_movable_kernel<<<nb, 256, 0, s>>>(ctx->Nm, ctx->mov_idx.as<int32_t>(), src, ctx->mov_xyz.as<float>(),
                                             ctx->mov_scale_act.as<float>(), ctx->mov_opacity.as<float>(),
                                             ctx->mov_sh.as<float>(), ctx->mov_log_scale.as<float>(),
                                             ctx->mov_quat_raw.as<float>(), ctx->mov_slot.as<int32_t>());
    group_bounds_kernel<<<nb, 256, 0, s>>>(ctx->Nm, ctx->mov_xyz.as<float>(), ctx->mov_scale_act.as<float>(), nullptr,
                                           ctx->mov_slot.as<int32_t>(), ctx->mov_bounds.as<float>());
    count_launches(2);
  }
  GSW_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace gsw
