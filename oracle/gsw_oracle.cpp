// gsw_oracle.cpp — CPU ORACLE (TEST INFRASTRUCTURE, NOT PRODUCT CODE).
//
// A scalar fp32 restatement of the GSWorld per-step observation path:
//   (1) per-link / per-object rigid transform of selected Gaussians,
//   (2) 3DGS forward rasterization (preprocess -> duplicate -> stable sort ->
//       tile ranges -> front-to-back blend).
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may load this library.  The product (gsworld_b200/)
// never links, imports or calls it.
//
// PARITY STATUS
//   * transform stage: follows /root/reference gsworld/utils/gs_utils.py:283-385
//     (transform_gaussians), :225-231 (quaternion re-scaling), :169-170
//     (inverse_sigmoid) and the activations of
//     submodules/gaussian-splatting/scene/gaussian_model.py:39-47,102-130.
//     Pinned by tests/golden/transform_*.npz, which were produced by importing
//     those reference functions (tests/golden/make_golden.py).
//   * SH colour and cov3D follow submodules/gaussian-splatting/utils/sh_utils.py:57-112
//     and utils/general_utils.py:78-110; pinned by tests/golden/sh_cov_*.npz.
//   * projection / EWA / tiling / sort / blend: the algorithm lives in the
//     third-party rasterizer graphdeco-inria/diff-gaussian-rasterization @
//     9c5c2028f6fbee2be239bc4c9421ff894fe4fbe0 (branch dr_aa), whose source is
//     ABSENT from /root/reference (empty submodule dir; .SUBMODULES.json:24-28),
//     with GSWorld's hand patch of the near plane 0.2 -> 0.05 (README.md:33).
//     Those stages restate its published algorithm (expression order of
//     forward.cu / auxiliary.h / rasterizer_impl.cu as published) anchored on the
//     reference call site gaussian_renderer/__init__.py:36-52,102-110.
//     No in-tree test or golden vector exists for them: PARITY UNPINNED.
//
// Build: g++ -O2 -ffp-contract=off -fopenmp -shared -fPIC (see oracle/Makefile).
// -ffp-contract=off is REQUIRED: every fp32 operation below is meant to round
// once, exactly where written.

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <vector>

#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

constexpr int kTile = 16;  // BLOCK_X == BLOCK_Y == 16 in the published rasterizer

constexpr float SH_C0 = 0.28209479177387814f;
constexpr float SH_C1 = 0.4886025119029199f;
constexpr float SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                            -1.0925484305920792f, 0.5462742152960396f};
constexpr float SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                            0.3731763325901154f,  -0.4570457994644658f, 1.445305721320277f,
                            -0.5900435899266435f};

// CUDA's float->int conversion (cvt.rzi.s32.f32): truncates, saturates, NaN -> 0.
inline int f2i_rz_sat(float v) {
  if (std::isnan(v)) return 0;
  if (v >= 2147483648.0f) return std::numeric_limits<int>::max();
  if (v <= -2147483648.0f) return std::numeric_limits<int>::min();
  return (int)v;
}

// CUDA min/max on floats follow fminf/fmaxf (a NaN operand is dropped).
inline float cmax(float a, float b) { return fmaxf(a, b); }
inline float cmin(float a, float b) { return fminf(a, b); }

struct M3 {  // glm-style: m[c][r], column major
  float m[3][3];
};

inline M3 glm_mul(const M3& a, const M3& b) {  // result[c][r] = sum_k a[k][r]*b[c][k], k ascending
  M3 o;
  for (int c = 0; c < 3; ++c)
    for (int r = 0; r < 3; ++r) {
      float s = a.m[0][r] * b.m[c][0];
      s = s + a.m[1][r] * b.m[c][1];
      s = s + a.m[2][r] * b.m[c][2];
      o.m[c][r] = s;
    }
  return o;
}

inline M3 glm_transpose(const M3& a) {
  M3 o;
  for (int c = 0; c < 3; ++c)
    for (int r = 0; r < 3; ++r) o.m[c][r] = a.m[r][c];
  return o;
}

inline void xform4x3(const float p[3], const float* m, float o[3]) {
  o[0] = m[0] * p[0] + m[4] * p[1] + m[8] * p[2] + m[12];
  o[1] = m[1] * p[0] + m[5] * p[1] + m[9] * p[2] + m[13];
  o[2] = m[2] * p[0] + m[6] * p[1] + m[10] * p[2] + m[14];
}

inline void xform4x4(const float p[3], const float* m, float o[4]) {
  o[0] = m[0] * p[0] + m[4] * p[1] + m[8] * p[2] + m[12];
  o[1] = m[1] * p[0] + m[5] * p[1] + m[9] * p[2] + m[13];
  o[2] = m[2] * p[0] + m[6] * p[1] + m[10] * p[2] + m[14];
  o[3] = m[3] * p[0] + m[7] * p[1] + m[11] * p[2] + m[15];
}

// published: ((v + 1.0) * S - 1.0) * 0.5 with double literals, result narrowed to float
inline float ndc2pix(float v, int S) { return (float)((((double)v + 1.0) * (double)S - 1.0) * 0.5); }

inline void get_rect(const float p[2], int max_radius, int gx, int gy, int rmin[2], int rmax[2]) {
  const float r = (float)max_radius;
  rmin[0] = std::min(gx, std::max(0, f2i_rz_sat((p[0] - r) / (float)kTile)));
  rmin[1] = std::min(gy, std::max(0, f2i_rz_sat((p[1] - r) / (float)kTile)));
  rmax[0] = std::min(gx, std::max(0, f2i_rz_sat((p[0] + r + (float)(kTile - 1)) / (float)kTile)));
  rmax[1] = std::min(gy, std::max(0, f2i_rz_sat((p[1] + r + (float)(kTile - 1)) / (float)kTile)));
}

// cov3D from (already activated) scale and quaternion; the quaternion is used as given.
// Same numbers as build_scaling_rotation/strip_symmetric (general_utils.py:64-110) up to
// summation order, which here follows the published kernel (M = S*R; Sigma = M^T * M in glm).
inline void cov3d_from_scale_rot(const float s[3], float mod, const float q[4], float cov[6]) {
  M3 S{};
  S.m[0][0] = mod * s[0];
  S.m[1][1] = mod * s[1];
  S.m[2][2] = mod * s[2];
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
  M3 M = glm_mul(S, R);
  M3 Sig = glm_mul(glm_transpose(M), M);
  cov[0] = Sig.m[0][0];
  cov[1] = Sig.m[0][1];
  cov[2] = Sig.m[0][2];
  cov[3] = Sig.m[1][1];
  cov[4] = Sig.m[1][2];
  cov[5] = Sig.m[2][2];
}

inline void cov2d(const float mean[3], float fx, float fy, float tanx, float tany, const float* c3,
                  const float* V, float out[3]) {
  float t[3];
  xform4x3(mean, V, t);
  const float limx = 1.3f * tanx;
  const float limy = 1.3f * tany;
  const float txtz = t[0] / t[2];
  const float tytz = t[1] / t[2];
  t[0] = cmin(limx, cmax(-limx, txtz)) * t[2];
  t[1] = cmin(limy, cmax(-limy, tytz)) * t[2];

  M3 J{};
  J.m[0][0] = fx / t[2];
  J.m[0][1] = 0.f;
  J.m[0][2] = -(fx * t[0]) / (t[2] * t[2]);
  J.m[1][0] = 0.f;
  J.m[1][1] = fy / t[2];
  J.m[1][2] = -(fy * t[1]) / (t[2] * t[2]);
  M3 W;
  W.m[0][0] = V[0]; W.m[0][1] = V[4]; W.m[0][2] = V[8];
  W.m[1][0] = V[1]; W.m[1][1] = V[5]; W.m[1][2] = V[9];
  W.m[2][0] = V[2]; W.m[2][1] = V[6]; W.m[2][2] = V[10];
  M3 T = glm_mul(W, J);
  M3 Vrk;
  Vrk.m[0][0] = c3[0]; Vrk.m[0][1] = c3[1]; Vrk.m[0][2] = c3[2];
  Vrk.m[1][0] = c3[1]; Vrk.m[1][1] = c3[3]; Vrk.m[1][2] = c3[4];
  Vrk.m[2][0] = c3[2]; Vrk.m[2][1] = c3[4]; Vrk.m[2][2] = c3[5];
  M3 cov = glm_mul(glm_mul(glm_transpose(T), glm_transpose(Vrk)), T);
  out[0] = cov.m[0][0];
  out[1] = cov.m[0][1];
  out[2] = cov.m[1][1];
}

// Degree <= 3 SH -> RGB, same polynomial as sh_utils.py:57-112, +0.5, clamp at 0.
inline void sh_to_rgb(int deg, int max_coeffs, const float pos[3], const float cam[3],
                      const float* sh /* [max_coeffs][3] */, float rgb[3]) {
  float d[3] = {pos[0] - cam[0], pos[1] - cam[1], pos[2] - cam[2]};
  const float len = sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
  d[0] = d[0] / len;
  d[1] = d[1] / len;
  d[2] = d[2] / len;
  (void)max_coeffs;
  for (int c = 0; c < 3; ++c) {
    auto S = [&](int k) { return sh[3 * k + c]; };
    float res = SH_C0 * S(0);
    if (deg > 0) {
      const float x = d[0], y = d[1], z = d[2];
      res = res - SH_C1 * y * S(1) + SH_C1 * z * S(2) - SH_C1 * x * S(3);
      if (deg > 1) {
        const float xx = x * x, yy = y * y, zz = z * z;
        const float xy = x * y, yz = y * z, xz = x * z;
        res = res + SH_C2[0] * xy * S(4) + SH_C2[1] * yz * S(5) +
              SH_C2[2] * (2.0f * zz - xx - yy) * S(6) + SH_C2[3] * xz * S(7) +
              SH_C2[4] * (xx - yy) * S(8);
        if (deg > 2) {
          res = res + SH_C3[0] * y * (3.0f * xx - yy) * S(9) + SH_C3[1] * xy * z * S(10) +
                SH_C3[2] * y * (4.0f * zz - xx - yy) * S(11) +
                SH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * S(12) +
                SH_C3[4] * x * (4.0f * zz - xx - yy) * S(13) + SH_C3[5] * z * (xx - yy) * S(14) +
                SH_C3[6] * x * (xx - 3.0f * yy) * S(15);
        }
      }
    }
    res += 0.5f;
    rgb[c] = cmax(res, 0.0f);
  }
}

inline uint32_t fbits(float f) {
  uint32_t u;
  std::memcpy(&u, &f, 4);
  return u;
}

}  // namespace

extern "C" {

// ---------------------------------------------------------------------------
// Stage A: Gaussian-level rigid transform for ONE (env, link/object) pair.
// Follows gs_utils.py:283-385 (order scale -> rotate -> translate) and the getters of
// gaussian_model.py:102-130 that the renderer applies afterwards.
//   xyz[n,3], log_scale[n,3], quat_raw[n,4] : the selected rows of the loaded model
//   R[9] row-major, t[3], qR[4] (= matrix_to_quaternion(R), wxyz), kappa (<=0: no scale step)
// Outputs: raw transformed xyz / log-scale / quaternion as transform_gaussians returns them,
// and the activated scale (exp) and rotation (F.normalize, eps 1e-12) the rasterizer sees.
// ---------------------------------------------------------------------------
void orc_transform_gaussians(int n, const float* xyz, const float* log_scale, const float* quat_raw,
                             const float* R, const float* t, const float* qR, int has_scale, float kappa,
                             float* out_xyz, float* out_log_scale, float* out_quat, float* out_scale_act,
                             float* out_quat_act) {
#pragma omp parallel for schedule(static)
  for (int i = 0; i < n; ++i) {
    float p[3] = {xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]};
    float ls[3] = {log_scale[3 * i], log_scale[3 * i + 1], log_scale[3 * i + 2]};
    if (has_scale) {  // gs_utils.py:296-316 ; inverse_sigmoid (not log) is the reference behaviour
      for (int k = 0; k < 3; ++k) {
        p[k] = p[k] * kappa;
        const float u = expf(ls[k]) * kappa;
        ls[k] = logf(u / (1.0f - u));
      }
    }
    // rotate points: matmul(R, p)   gs_utils.py:324-339
    float rp[3];
    for (int r = 0; r < 3; ++r) rp[r] = R[3 * r] * p[0] + R[3 * r + 1] * p[1] + R[3 * r + 2] * p[2];
    // rotate quaternions: standardize(qR (x) q/|q|) * |q|   gs_utils.py:225-231, 342-354
    const float qw = quat_raw[4 * i], qx = quat_raw[4 * i + 1], qy = quat_raw[4 * i + 2], qz = quat_raw[4 * i + 3];
    const float nrm = sqrtf(qw * qw + qx * qx + qy * qy + qz * qz);
    const float bw = qw / nrm, bx = qx / nrm, by = qy / nrm, bz = qz / nrm;
    const float aw = qR[0], ax = qR[1], ay = qR[2], az = qR[3];
    float ow = aw * bw - ax * bx - ay * by - az * bz;
    float ox = aw * bx + ax * bw + ay * bz - az * by;
    float oy = aw * by - ax * bz + ay * bw + az * bx;
    float oz = aw * bz + ax * by - ay * bx + az * bw;
    if (ow < 0.f) {
      ow = -ow; ox = -ox; oy = -oy; oz = -oz;
    }
    const float q2[4] = {ow * nrm, ox * nrm, oy * nrm, oz * nrm};
    // translate   gs_utils.py:357-367
    for (int k = 0; k < 3; ++k) out_xyz[3 * i + k] = rp[k] + t[k];
    for (int k = 0; k < 3; ++k) out_log_scale[3 * i + k] = ls[k];
    for (int k = 0; k < 4; ++k) out_quat[4 * i + k] = q2[k];
    // getters: exp / F.normalize(eps=1e-12)   gaussian_model.py:39-47,102-108
    for (int k = 0; k < 3; ++k) out_scale_act[3 * i + k] = expf(ls[k]);
    const float n2 = sqrtf(q2[0] * q2[0] + q2[1] * q2[1] + q2[2] * q2[2] + q2[3] * q2[3]);
    const float den = cmax(n2, 1e-12f);
    for (int k = 0; k < 4; ++k) out_quat_act[4 * i + k] = q2[k] / den;
  }
}

// ---------------------------------------------------------------------------
// Stage B: forward rasterization of one frame.
// Argument list mirrors the published RasterizeGaussiansCUDA / Rasterizer::forward, as pinned by
// the call site gaussian_renderer/__init__.py:36-52,102-110.
//   means3D[P,3]; shs[P,M,3] or NULL; colors_precomp[P,3] or NULL; opacities[P];
//   scales[P,3] + rotations[P,4] (already activated) or cov3D_precomp[P,6];
//   viewmatrix/projmatrix: the TRANSPOSED row-major 4x4 tensors of scene/cameras.py:86-88.
// Outputs (any may be NULL): out_color[3,H,W], out_radii[P], out_invdepth[H,W]; debug taps:
//   dbg_tiles_touched[P], dbg_depth[P], dbg_xy[P,2], dbg_conic_opacity[P,4], dbg_rgb[P,3],
//   dbg_keys[cap] / dbg_vals[cap] (sorted, first D valid), dbg_ranges[T,2], dbg_ncontrib[H,W],
//   dbg_final_T[H,W], dbg_fragile[H,W] (1 where a threshold decision was within `fragile_eps`
//   relative distance of flipping: alpha vs 1/255, T vs 1e-4, power vs 0), dbg_pairs (count of
//   blend-loop iterations = sum over pixels of entries visited), dbg_best_id[H,W] / dbg_best_margin[H,W]
//   (index of the Gaussian with the largest blending weight alpha*T, -1 if none, and how far ahead of the
//   runner-up it is, relative; for the segmentation observation, which has no counterpart in the reference).
// Returns D (number of tile duplicates), or -1 if dbg_keys capacity was too small.
// near_plane: 0.05 for GSWorld (README.md:33); 0.2 upstream.
// ---------------------------------------------------------------------------
long long orc_rasterize_forward(int P, int D_sh, int M, const float* bg, int W, int H, const float* means3D,
                                const float* shs, const float* colors_precomp, const float* opacities,
                                const float* scales, float scale_modifier, const float* rotations,
                                const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                                const float* campos, float tan_fovx, float tan_fovy, int antialiasing,
                                float near_plane, float* out_color, int* out_radii, float* out_invdepth,
                                int* dbg_tiles_touched, float* dbg_depth, float* dbg_xy, float* dbg_conic_opacity,
                                float* dbg_rgb, unsigned long long* dbg_keys, unsigned int* dbg_vals,
                                long long dbg_cap, int* dbg_ranges, int* dbg_ncontrib, float* dbg_final_T,
                                unsigned char* dbg_fragile, float fragile_eps, long long* dbg_pairs,
                                int global_stable_sort, int* dbg_best_id, float* dbg_best_margin) {
  const int gx = (W + kTile - 1) / kTile, gy = (H + kTile - 1) / kTile;
  const int T = gx * gy;
  const float focal_y = (float)H / (2.0f * tan_fovy);
  const float focal_x = (float)W / (2.0f * tan_fovx);

  std::vector<int> radii(P, 0), touched(P, 0);
  std::vector<float> depth(P, 0.f), xy(2 * (size_t)P, 0.f), co(4 * (size_t)P, 0.f), rgb(3 * (size_t)P, 0.f);

  if (out_color) std::fill(out_color, out_color + 3 * (size_t)H * W, 0.f);
  if (out_invdepth) std::fill(out_invdepth, out_invdepth + (size_t)H * W, 0.f);
  if (P == 0) {  // published wrapper skips the whole forward when P == 0: outputs stay zero
    if (dbg_ranges) std::fill(dbg_ranges, dbg_ranges + 2 * (size_t)T, 0);
    if (dbg_pairs) *dbg_pairs = 0;
    return 0;
  }

  // ---- preprocess ----
#pragma omp parallel for schedule(static)
  for (int i = 0; i < P; ++i) {
    const float p[3] = {means3D[3 * i], means3D[3 * i + 1], means3D[3 * i + 2]};
    float pv[3];
    xform4x3(p, viewmatrix, pv);
    if (pv[2] <= near_plane) continue;
    float ph[4];
    xform4x4(p, projmatrix, ph);
    const float pw = 1.0f / (ph[3] + 0.0000001f);
    const float pp[2] = {ph[0] * pw, ph[1] * pw};
    float c3buf[6];
    const float* c3;
    if (cov3D_precomp) {
      c3 = cov3D_precomp + 6 * (size_t)i;
    } else {
      cov3d_from_scale_rot(scales + 3 * (size_t)i, scale_modifier, rotations + 4 * (size_t)i, c3buf);
      c3 = c3buf;
    }
    float cov[3];
    cov2d(p, focal_x, focal_y, tan_fovx, tan_fovy, c3, viewmatrix, cov);
    const float h_var = 0.3f;
    const float det_cov = cov[0] * cov[2] - cov[1] * cov[1];
    cov[0] += h_var;
    cov[2] += h_var;
    const float det_plus = cov[0] * cov[2] - cov[1] * cov[1];
    float h_scale = 1.0f;
    if (antialiasing) h_scale = sqrtf(cmax(0.000025f, det_cov / det_plus));
    const float det = det_plus;
    if (det == 0.0f) continue;
    const float det_inv = 1.f / det;
    const float conic[3] = {cov[2] * det_inv, -cov[1] * det_inv, cov[0] * det_inv};
    const float mid = 0.5f * (cov[0] + cov[2]);
    const float l1 = mid + sqrtf(cmax(0.1f, mid * mid - det));
    const float l2 = mid - sqrtf(cmax(0.1f, mid * mid - det));
    const float my_radius = ceilf(3.f * sqrtf(cmax(l1, l2)));
    const float pim[2] = {ndc2pix(pp[0], W), ndc2pix(pp[1], H)};
    int rmin[2], rmax[2];
    get_rect(pim, f2i_rz_sat(my_radius), gx, gy, rmin, rmax);
    if ((rmax[0] - rmin[0]) * (rmax[1] - rmin[1]) == 0) continue;
    if (colors_precomp) {
      for (int k = 0; k < 3; ++k) rgb[3 * (size_t)i + k] = colors_precomp[3 * (size_t)i + k];
    } else {
      sh_to_rgb(D_sh, M, p, campos, shs + 3 * (size_t)M * i, &rgb[3 * (size_t)i]);
    }
    depth[i] = pv[2];
    radii[i] = f2i_rz_sat(my_radius);
    xy[2 * (size_t)i] = pim[0];
    xy[2 * (size_t)i + 1] = pim[1];
    co[4 * (size_t)i + 0] = conic[0];
    co[4 * (size_t)i + 1] = conic[1];
    co[4 * (size_t)i + 2] = conic[2];
    co[4 * (size_t)i + 3] = opacities[i] * h_scale;
    touched[i] = (rmax[1] - rmin[1]) * (rmax[0] - rmin[0]);
  }

  // ---- inclusive scan + duplicateWithKeys ----
  std::vector<long long> offs(P);
  long long D = 0;
  for (int i = 0; i < P; ++i) {
    D += touched[i];
    offs[i] = D;
  }
  std::vector<uint64_t> keys((size_t)D);
  std::vector<uint32_t> vals((size_t)D);
#pragma omp parallel for schedule(static)
  for (int i = 0; i < P; ++i) {
    if (radii[i] > 0) {
      long long off = (i == 0) ? 0 : offs[i - 1];
      int rmin[2], rmax[2];
      get_rect(&xy[2 * (size_t)i], radii[i], gx, gy, rmin, rmax);
      for (int y = rmin[1]; y < rmax[1]; ++y)
        for (int x = rmin[0]; x < rmax[0]; ++x) {
          uint64_t key = (uint64_t)(y * gx + x);
          key <<= 32;
          key |= fbits(depth[i]);
          keys[off] = key;
          vals[off] = (uint32_t)i;
          ++off;
        }
    }
  }

  // ---- stable sort by key (LSB radix sort in the published code == stable) ----
  std::vector<uint64_t> skeys((size_t)D);
  std::vector<uint32_t> svals((size_t)D);
  if (global_stable_sort) {
    std::vector<uint32_t> perm((size_t)D);
    for (long long k = 0; k < D; ++k) perm[k] = (uint32_t)k;
    std::stable_sort(perm.begin(), perm.end(), [&](uint32_t a, uint32_t b) { return keys[a] < keys[b]; });
    for (long long k = 0; k < D; ++k) {
      skeys[k] = keys[perm[k]];
      svals[k] = vals[perm[k]];
    }
  } else {
    // same ordering, parallel: stable counting split by tile, then per-tile stable sort by depth bits
    std::vector<long long> cnt(T + 1, 0);
    for (long long k = 0; k < D; ++k) cnt[(keys[k] >> 32) + 1]++;
    for (int t = 0; t < T; ++t) cnt[t + 1] += cnt[t];
    std::vector<long long> cur(cnt.begin(), cnt.end() - 1);
    for (long long k = 0; k < D; ++k) {
      const long long d = cur[keys[k] >> 32]++;
      skeys[d] = keys[k];
      svals[d] = vals[k];
    }
#pragma omp parallel for schedule(dynamic, 1)
    for (int t = 0; t < T; ++t) {
      const long long a = cnt[t], b = cnt[t + 1];
      if (b - a < 2) continue;
      std::vector<std::pair<uint64_t, uint32_t>> tmp((size_t)(b - a));
      for (long long k = a; k < b; ++k) tmp[k - a] = {skeys[k], svals[k]};
      std::stable_sort(tmp.begin(), tmp.end(),
                       [](const std::pair<uint64_t, uint32_t>& x, const std::pair<uint64_t, uint32_t>& y) {
                         return x.first < y.first;
                       });
      for (long long k = a; k < b; ++k) {
        skeys[k] = tmp[k - a].first;
        svals[k] = tmp[k - a].second;
      }
    }
  }

  // ---- identifyTileRanges ----
  std::vector<int> ranges(2 * (size_t)T, 0);
  for (long long k = 0; k < D; ++k) {
    const uint32_t cur = (uint32_t)(skeys[k] >> 32);
    if (k == 0)
      ranges[2 * (size_t)cur] = 0;
    else {
      const uint32_t prev = (uint32_t)(skeys[k - 1] >> 32);
      if (cur != prev) {
        ranges[2 * (size_t)prev + 1] = (int)k;
        ranges[2 * (size_t)cur] = (int)k;
      }
    }
    if (k == D - 1) ranges[2 * (size_t)cur + 1] = (int)D;
  }

  // ---- render ----
  long long pairs_total = 0;
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : pairs_total)
  for (int tile = 0; tile < T; ++tile) {
    const int tx = tile % gx, ty = tile / gx;
    const int r0 = ranges[2 * (size_t)tile], r1 = ranges[2 * (size_t)tile + 1];
    for (int ly = 0; ly < kTile; ++ly)
      for (int lx = 0; lx < kTile; ++lx) {
        const int px = tx * kTile + lx, py = ty * kTile + ly;
        if (px >= W || py >= H) continue;
        const size_t pid = (size_t)W * py + px;
        const float pxf = (float)px, pyf = (float)py;
        float Tr = 1.0f;
        float C[3] = {0.f, 0.f, 0.f};
        float invd = 0.f;
        uint32_t contributor = 0, last = 0;
        unsigned char fragile = 0;
        // segmentation tap (no counterpart in the reference): Gaussian with the largest blending weight alpha*T
        float best_w = 0.f, second_w = 0.f;
        int best_id = -1;
        for (int k = r0; k < r1; ++k) {
          ++contributor;
          const uint32_t id = svals[k];
          const float dx = xy[2 * (size_t)id] - pxf, dy = xy[2 * (size_t)id + 1] - pyf;
          const float* c = &co[4 * (size_t)id];
          const float power = -0.5f * (c[0] * dx * dx + c[2] * dy * dy) - c[1] * dx * dy;
          if (fabsf(power) <= fragile_eps * 1e-3f) fragile = 1;
          if (power > 0.0f) continue;
          const float alpha = cmin(0.99f, c[3] * expf(power));
          if (fabsf(alpha - 1.0f / 255.0f) <= fragile_eps * (1.0f / 255.0f)) fragile = 1;
          if (alpha < 1.0f / 255.0f) continue;
          const float test_T = Tr * (1 - alpha);
          if (fabsf(test_T - 0.0001f) <= fragile_eps * 0.0001f) fragile = 1;
          if (test_T < 0.0001f) break;  // `done = true` in the published kernel
          for (int ch = 0; ch < 3; ++ch) C[ch] += rgb[3 * (size_t)id + ch] * alpha * Tr;
          invd += (1 / depth[id]) * alpha * Tr;
          {
            const float wgt = alpha * Tr;
            if (wgt > best_w) {
              second_w = best_w;
              best_w = wgt;
              best_id = (int)id;
            } else if (wgt > second_w) {
              second_w = wgt;
            }
          }
          Tr = test_T;
          last = contributor;
        }
        pairs_total += contributor;
        if (out_color)
          for (int ch = 0; ch < 3; ++ch) out_color[(size_t)ch * H * W + pid] = C[ch] + Tr * bg[ch];
        if (out_invdepth) out_invdepth[pid] = invd;
        if (dbg_ncontrib) dbg_ncontrib[pid] = (int)last;
        if (dbg_final_T) dbg_final_T[pid] = Tr;
        if (dbg_fragile) dbg_fragile[pid] = fragile;
        if (dbg_best_id) dbg_best_id[pid] = best_id;
        if (dbg_best_margin) dbg_best_margin[pid] = best_w > 0.f ? (best_w - second_w) / best_w : 1.0f;
      }
  }

  if (out_radii) std::copy(radii.begin(), radii.end(), out_radii);
  if (dbg_tiles_touched) std::copy(touched.begin(), touched.end(), dbg_tiles_touched);
  if (dbg_depth) std::copy(depth.begin(), depth.end(), dbg_depth);
  if (dbg_xy) std::copy(xy.begin(), xy.end(), dbg_xy);
  if (dbg_conic_opacity) std::copy(co.begin(), co.end(), dbg_conic_opacity);
  if (dbg_rgb) std::copy(rgb.begin(), rgb.end(), dbg_rgb);
  if (dbg_ranges) std::copy(ranges.begin(), ranges.end(), dbg_ranges);
  if (dbg_pairs) *dbg_pairs = pairs_total;
  if (dbg_keys || dbg_vals) {
    if (D > dbg_cap) return -1;
    if (dbg_keys) std::copy(skeys.begin(), skeys.end(), dbg_keys);
    if (dbg_vals) std::copy(svals.begin(), svals.end(), dbg_vals);
  }
  return D;
}

// markVisible of the published rasterizer: near-plane test only.
void orc_mark_visible(int P, const float* means3D, const float* viewmatrix, float near_plane, unsigned char* out) {
  for (int i = 0; i < P; ++i) {
    float pv[3];
    xform4x3(means3D + 3 * (size_t)i, viewmatrix, pv);
    out[i] = pv[2] > near_plane ? 1 : 0;
  }
}

// Python-side cross-check helpers (sh_utils.py / general_utils.py restated in the kernel's op order)
void orc_sh_to_rgb(int P, int deg, int M, const float* means3D, const float* campos, const float* shs, float* out) {
  for (int i = 0; i < P; ++i) sh_to_rgb(deg, M, means3D + 3 * (size_t)i, campos, shs + 3 * (size_t)M * i, out + 3 * (size_t)i);
}
void orc_cov3d(int P, const float* scales, float mod, const float* rots, float* out) {
  for (int i = 0; i < P; ++i) cov3d_from_scale_rot(scales + 3 * (size_t)i, mod, rots + 4 * (size_t)i, out + 6 * (size_t)i);
}

// uint8 epilogue of gs_world_wrapper.py:268-270 applied to render()'s clamp(0,1) output:
// (clamp01(c) * 255).clamp(0,255).to(uint8)  => truncation.  in: [3,H,W] f32, out: [H,W,3] u8.
void orc_to_uint8_hwc(int H, int W, const float* chw, unsigned char* hwc) {
  for (int y = 0; y < H; ++y)
    for (int x = 0; x < W; ++x)
      for (int c = 0; c < 3; ++c) {
        float v = chw[(size_t)c * H * W + (size_t)y * W + x];
        v = v < 0.f ? 0.f : (v > 1.f ? 1.f : v);  // torch.clamp(0,1); NaN propagates in torch, maps to 0 on cast
        float s = v * 255.f;
        s = s < 0.f ? 0.f : (s > 255.f ? 255.f : s);
        hwc[((size_t)y * W + x) * 3 + c] = std::isnan(s) ? 0 : (unsigned char)s;
      }
}

int orc_num_threads() {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

}  // extern "C"
