// pose.cu — per-step pose algebra in one launch each: (env, link/object) rigid transforms and camera
// matrices.  Replaces ~60 tiny torch launches and four .cpu() syncs per step of the reference
// (gs_world_wrapper.py:115-120 link_trans, :139-156 objects via extract_rigid_transform =
// pcd_utils.py:224-252, :277-325 cam_maniskill2gs + scene/cameras.py:80-89 + graphics_utils.py:38-71).
// These are float-tolerance quantities (the reference itself goes through LAPACK inverses / SVD);
// parity bar: 1e-5 absolute on every output (tests/test_gpu_pose_compose.py).
#include <cstring>

#include "gsw_internal.h"

namespace gsw {
namespace {

struct Mat4 {
  float m[16];
};

__device__ __forceinline__ void mul4(const float* a, const float* b, float* o) {  // row-major 4x4
#pragma unroll
  for (int r = 0; r < 4; ++r)
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      float s = a[4 * r] * b[c];
      s += a[4 * r + 1] * b[4 + c];
      s += a[4 * r + 2] * b[8 + c];
      s += a[4 * r + 3] * b[12 + c];
      o[4 * r + c] = s;
    }
}

// pytorch3d / ManiSkill matrix_to_quaternion: candidate with the largest |component| (wxyz)
__device__ void mat_to_quat(const float* R /*row-major 3x3*/, float* q) {
  const float m00 = R[0], m01 = R[1], m02 = R[2], m10 = R[3], m11 = R[4], m12 = R[5], m20 = R[6], m21 = R[7],
              m22 = R[8];
  float t[4] = {1.0f + m00 + m11 + m22, 1.0f + m00 - m11 - m22, 1.0f - m00 + m11 - m22, 1.0f - m00 - m11 + m22};
  float qa[4];
  int best = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    qa[k] = t[k] > 0.f ? sqrtf(t[k]) : 0.f;
    if (qa[k] > qa[best]) best = k;
  }
  float c[4];
  if (best == 0) {
    c[0] = qa[0] * qa[0]; c[1] = m21 - m12; c[2] = m02 - m20; c[3] = m10 - m01;
  } else if (best == 1) {
    c[0] = m21 - m12; c[1] = qa[1] * qa[1]; c[2] = m10 + m01; c[3] = m02 + m20;
  } else if (best == 2) {
    c[0] = m02 - m20; c[1] = m10 + m01; c[2] = qa[2] * qa[2]; c[3] = m12 + m21;
  } else {
    c[0] = m10 - m01; c[1] = m20 + m02; c[2] = m21 + m12; c[3] = qa[3] * qa[3];
  }
  const float den = 2.0f * fmaxf(qa[best], 0.1f);
#pragma unroll
  for (int k = 0; k < 4; ++k) q[k] = c[k] / den;
}

// Polar factor Q = U V^T and mean singular value of a 3x3 (double, cyclic Jacobi on A^T A).
__device__ void polar3(const double A[9], double Q[9], double* mean_sigma) {
  double S[9], V[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      double s = 0;
      for (int k = 0; k < 3; ++k) s += A[3 * k + i] * A[3 * k + j];
      S[3 * i + j] = s;
    }
  for (int sweep = 0; sweep < 12; ++sweep) {
    const double off = fabs(S[1]) + fabs(S[2]) + fabs(S[5]);
    if (off < 1e-30) break;
    for (int p = 0; p < 2; ++p)
      for (int q = p + 1; q < 3; ++q) {
        const double apq = S[3 * p + q];
        if (fabs(apq) < 1e-300) continue;
        const double theta = (S[3 * q + q] - S[3 * p + p]) / (2.0 * apq);
        const double t = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
        const double c = 1.0 / sqrt(t * t + 1.0), s = t * c;
        for (int k = 0; k < 3; ++k) {  // S <- S J
          const double skp = S[3 * k + p], skq = S[3 * k + q];
          S[3 * k + p] = c * skp - s * skq;
          S[3 * k + q] = s * skp + c * skq;
        }
        for (int k = 0; k < 3; ++k) {  // S <- J^T S
          const double spk = S[3 * p + k], sqk = S[3 * q + k];
          S[3 * p + k] = c * spk - s * sqk;
          S[3 * q + k] = s * spk + c * sqk;
        }
        for (int k = 0; k < 3; ++k) {  // V <- V J
          const double vkp = V[3 * k + p], vkq = V[3 * k + q];
          V[3 * k + p] = c * vkp - s * vkq;
          V[3 * k + q] = s * vkp + c * vkq;
        }
      }
  }
  double sig[3], inv[3];
  for (int k = 0; k < 3; ++k) {
    sig[k] = sqrt(fmax(S[4 * k], 0.0));
    inv[k] = sig[k] > 0 ? 1.0 / sig[k] : 0.0;
  }
  *mean_sigma = (sig[0] + sig[1] + sig[2]) / 3.0;
  // Q = A V diag(1/sigma) V^T
  double AV[9];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      double s = 0;
      for (int k = 0; k < 3; ++k) s += A[3 * i + k] * V[3 * k + j];
      AV[3 * i + j] = s * inv[j];
    }
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      double s = 0;
      for (int k = 0; k < 3; ++k) s += AV[3 * i + k] * V[3 * j + k];
      Q[3 * i + j] = s;
    }
}

struct RigConst {
  Mat4 S, S_inv;
  float link_off[3];
  int L, K;
};

// one thread per (env, slot)
__global__ void compose_poses_kernel(RigConst rc, int B, const float* __restrict__ link_mats,
                                     const float* __restrict__ scan_inv, const float* __restrict__ obj_pose,
                                     int obj_stride, const float* __restrict__ obj_S_inv,
                                     const float* __restrict__ obj_off, const float* __restrict__ obj_scale,
                                     float* __restrict__ out) {
  const int S = rc.L + rc.K;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * S) return;
  const int e = idx / S, s = idx % S;
  float* o = out + (size_t)idx * GSW_POSE_STRIDE;
  float R[9], t[3], kappa = 0.f, has_scale = 0.f;
  if (s < rc.L) {
    float M[16], A[16], Bm[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) M[k] = link_mats[((size_t)e * rc.L + s) * 16 + k];
    M[3] += rc.link_off[0];
    M[7] += rc.link_off[1];
    M[11] += rc.link_off[2];
    mul4(rc.S.m, M, A);
    mul4(A, scan_inv + (size_t)s * 16, Bm);
    mul4(Bm, rc.S_inv.m, A);
#pragma unroll
    for (int r = 0; r < 3; ++r) {
#pragma unroll
      for (int c = 0; c < 3; ++c) R[3 * r + c] = A[4 * r + c];
      t[r] = A[4 * r + 3];
    }
  } else {
    const int k = s - rc.L;
    const float* p = obj_pose + ((size_t)e * rc.K + k) * obj_stride;
    float qw = p[3], qx = p[4], qy = p[5], qz = p[6];
    const float n = sqrtf(qw * qw + qx * qx + qy * qy + qz * qz);
    qw /= n; qx /= n; qy /= n; qz /= n;
    float M[16] = {1 - 2 * (qy * qy + qz * qz), 2 * (qx * qy - qw * qz), 2 * (qx * qz + qw * qy), p[0] + obj_off[3 * k],
                   2 * (qx * qy + qw * qz), 1 - 2 * (qx * qx + qz * qz), 2 * (qy * qz - qw * qx), p[1] + obj_off[3 * k + 1],
                   2 * (qx * qz - qw * qy), 2 * (qy * qz + qw * qx), 1 - 2 * (qx * qx + qy * qy), p[2] + obj_off[3 * k + 2],
                   0.f, 0.f, 0.f, 1.f};
    float A[16], F[16];
    mul4(rc.S.m, M, A);
    mul4(A, obj_S_inv + (size_t)k * 16, F);
    double Ad[9], Q[9], ms;
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c) Ad[3 * r + c] = (double)F[4 * r + c];
    polar3(Ad, Q, &ms);
    for (int j = 0; j < 9; ++j) R[j] = (float)Q[j];
    for (int r = 0; r < 3; ++r) t[r] = F[4 * r + 3];
    kappa = (float)ms * obj_scale[k];
    has_scale = 1.f;
  }
  float q[4];
  mat_to_quat(R, q);
#pragma unroll
  for (int j = 0; j < 9; ++j) o[j] = R[j];
  o[9] = t[0]; o[10] = t[1]; o[11] = t[2];
  o[12] = q[0]; o[13] = q[1]; o[14] = q[2]; o[15] = q[3];
  o[16] = kappa; o[17] = has_scale; o[18] = 0.f; o[19] = 0.f;
}

__device__ void inv4d(const double* m, double* inv) {  // general 4x4 inverse (cofactors)
  double a[16];
  a[0] = m[5] * m[10] * m[15] - m[5] * m[11] * m[14] - m[9] * m[6] * m[15] + m[9] * m[7] * m[14] + m[13] * m[6] * m[11] - m[13] * m[7] * m[10];
  a[4] = -m[4] * m[10] * m[15] + m[4] * m[11] * m[14] + m[8] * m[6] * m[15] - m[8] * m[7] * m[14] - m[12] * m[6] * m[11] + m[12] * m[7] * m[10];
  a[8] = m[4] * m[9] * m[15] - m[4] * m[11] * m[13] - m[8] * m[5] * m[15] + m[8] * m[7] * m[13] + m[12] * m[5] * m[11] - m[12] * m[7] * m[9];
  a[12] = -m[4] * m[9] * m[14] + m[4] * m[10] * m[13] + m[8] * m[5] * m[14] - m[8] * m[6] * m[13] - m[12] * m[5] * m[10] + m[12] * m[6] * m[9];
  a[1] = -m[1] * m[10] * m[15] + m[1] * m[11] * m[14] + m[9] * m[2] * m[15] - m[9] * m[3] * m[14] - m[13] * m[2] * m[11] + m[13] * m[3] * m[10];
  a[5] = m[0] * m[10] * m[15] - m[0] * m[11] * m[14] - m[8] * m[2] * m[15] + m[8] * m[3] * m[14] + m[12] * m[2] * m[11] - m[12] * m[3] * m[10];
  a[9] = -m[0] * m[9] * m[15] + m[0] * m[11] * m[13] + m[8] * m[1] * m[15] - m[8] * m[3] * m[13] - m[12] * m[1] * m[11] + m[12] * m[3] * m[9];
  a[13] = m[0] * m[9] * m[14] - m[0] * m[10] * m[13] - m[8] * m[1] * m[14] + m[8] * m[2] * m[13] + m[12] * m[1] * m[10] - m[12] * m[2] * m[9];
  a[2] = m[1] * m[6] * m[15] - m[1] * m[7] * m[14] - m[5] * m[2] * m[15] + m[5] * m[3] * m[14] + m[13] * m[2] * m[7] - m[13] * m[3] * m[6];
  a[6] = -m[0] * m[6] * m[15] + m[0] * m[7] * m[14] + m[4] * m[2] * m[15] - m[4] * m[3] * m[14] - m[12] * m[2] * m[7] + m[12] * m[3] * m[6];
  a[10] = m[0] * m[5] * m[15] - m[0] * m[7] * m[13] - m[4] * m[1] * m[15] + m[4] * m[3] * m[13] + m[12] * m[1] * m[7] - m[12] * m[3] * m[5];
  a[14] = -m[0] * m[5] * m[14] + m[0] * m[6] * m[13] + m[4] * m[1] * m[14] - m[4] * m[2] * m[13] - m[12] * m[1] * m[6] + m[12] * m[2] * m[5];
  a[3] = -m[1] * m[6] * m[11] + m[1] * m[7] * m[10] + m[5] * m[2] * m[11] - m[5] * m[3] * m[10] - m[9] * m[2] * m[7] + m[9] * m[3] * m[6];
  a[7] = m[0] * m[6] * m[11] - m[0] * m[7] * m[10] - m[4] * m[2] * m[11] + m[4] * m[3] * m[10] + m[8] * m[2] * m[7] - m[8] * m[3] * m[6];
  a[11] = -m[0] * m[5] * m[11] + m[0] * m[7] * m[9] + m[4] * m[1] * m[11] - m[4] * m[3] * m[9] - m[8] * m[1] * m[7] + m[8] * m[3] * m[5];
  a[15] = m[0] * m[5] * m[10] - m[0] * m[6] * m[9] - m[4] * m[1] * m[10] + m[4] * m[2] * m[9] + m[8] * m[1] * m[6] - m[8] * m[2] * m[5];
  const double det = m[0] * a[0] + m[1] * a[4] + m[2] * a[8] + m[3] * a[12];
  const double id = 1.0 / det;
  for (int i = 0; i < 16; ++i) inv[i] = a[i] * id;
}

struct CamConst {
  Mat4 rigid;  // rigid_sim2real (row-major)
  float scale;
  int H, W;
  double znear, zfar;
};

// one thread per camera: cam_maniskill2gs (env row given) + Camera.__init__, in double, rounded once at the end
__global__ void compose_cameras_kernel(CamConst cc, int n, const float* __restrict__ ext, const float* __restrict__ Kin,
                                       float* __restrict__ view, float* __restrict__ proj, float* __restrict__ campos,
                                       float* __restrict__ tanfov) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double w2c[16], c2w[16], tmp[16], rw2c[16];
  for (int k = 0; k < 12; ++k) w2c[k] = (double)ext[(size_t)i * 12 + k];
  w2c[12] = w2c[13] = w2c[14] = 0.0;
  w2c[15] = 1.0;
  inv4d(w2c, c2w);
  c2w[3] *= (double)cc.scale;
  c2w[7] *= (double)cc.scale;
  c2w[11] *= (double)cc.scale;
  for (int r = 0; r < 4; ++r)
    for (int c = 0; c < 4; ++c) {
      double s = 0;
      for (int k = 0; k < 4; ++k) s += (double)cc.rigid.m[4 * r + k] * c2w[4 * k + c];
      tmp[4 * r + c] = s;
    }
  inv4d(tmp, rw2c);  // real_world2cam == the W2V matrix getWorld2View2 rebuilds (trans 0, scale 1)
  float Vt[16];      // world_view_transform = W2V^T, stored row-major
  for (int r = 0; r < 4; ++r)
    for (int c = 0; c < 4; ++c) Vt[4 * r + c] = (float)rw2c[4 * c + r];
  const double fx = (double)Kin[(size_t)i * 9], fy = (double)Kin[(size_t)i * 9 + 4];
  const double tanx = (double)cc.W / (2.0 * fx), tany = (double)cc.H / (2.0 * fy);  // tan(atan(.)) collapsed
  // projection_matrix^T (graphics_utils.py:51-71), fp32 entries as torch builds them
  float Pt[16] = {0};
  Pt[0] = (float)(1.0 / tanx);
  Pt[5] = (float)(1.0 / tany);
  Pt[10] = (float)(cc.zfar / (cc.zfar - cc.znear));
  Pt[11] = 1.0f;
  Pt[14] = (float)(-(cc.zfar * cc.znear) / (cc.zfar - cc.znear));
  float full[16];
  mul4(Vt, Pt, full);
  // camera_center = inverse(world_view_transform)[3,:3] = translation of C2W
  double c2w2[16];
  inv4d(rw2c, c2w2);
  for (int k = 0; k < 16; ++k) {
    view[(size_t)i * 16 + k] = Vt[k];
    proj[(size_t)i * 16 + k] = full[k];
  }
  campos[(size_t)i * 3] = (float)c2w2[3];
  campos[(size_t)i * 3 + 1] = (float)c2w2[7];
  campos[(size_t)i * 3 + 2] = (float)c2w2[11];
  tanfov[(size_t)i * 2] = (float)tanx;
  tanfov[(size_t)i * 2 + 1] = (float)tany;
}

}  // namespace

int launch_compose_poses(const gsw_ctx* ctx, int B, const float* link_mats, const float* obj_pose, int obj_stride,
                         float* pose_out, cudaStream_t s) {
  RigConst rc;
  memcpy(rc.S.m, ctx->rig.S, sizeof rc.S.m);
  memcpy(rc.S_inv.m, ctx->rig.S_inv, sizeof rc.S_inv.m);
  memcpy(rc.link_off, ctx->rig.link_off, sizeof rc.link_off);
  rc.L = ctx->rig.L;
  rc.K = ctx->rig.K;
  const int total = B * (rc.L + rc.K);
  if (total <= 0) return 0;
  const float* d = ctx->rig.dev.as<float>();
  const float* scan_inv = d;
  const float* obj_S_inv = d + (size_t)rc.L * 16;
  const float* obj_off = obj_S_inv + (size_t)rc.K * 16;
  const float* obj_scale = obj_off + (size_t)rc.K * 3;
  compose_poses_kernel<<<(total + 127) / 128, 128, 0, s>>>(rc, B, link_mats, scan_inv, obj_pose, obj_stride, obj_S_inv,
                                                         obj_off, obj_scale, pose_out);
  count_launches(1);
  GSW_CUDA(cudaGetLastError());
  return 0;
}

int launch_compose_cameras(int n, const float* ext, const float* K, int H, int W, const float* rigid, float scale,
                           float* view, float* proj, float* campos, float* tanfov, cudaStream_t s) {
  if (n <= 0) return 0;
  CamConst cc;
  memcpy(cc.rigid.m, rigid, sizeof cc.rigid.m);
  cc.scale = scale;
  cc.H = H;
  cc.W = W;
  cc.znear = 0.01;
  cc.zfar = 100.0;
  compose_cameras_kernel<<<(n + 63) / 64, 64, 0, s>>>(cc, n, ext, K, view, proj, campos, tanfov);
  count_launches(1);
  GSW_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace gsw
