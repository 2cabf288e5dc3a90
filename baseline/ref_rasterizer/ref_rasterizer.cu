// ref_rasterizer.cu — COMPARISON POINT, NOT PRODUCT CODE, NOT THE REFERENCE'S SOURCE.
//
// The reference's CUDA rasterizer (graphdeco-inria/diff-gaussian-rasterization @ 9c5c2028, near plane 0.05 per
// GSWorld README.md:33) is an absent submodule of /root/reference, so it cannot be built or timed here.  This
// file RESTATES its published forward structure so that bench.py can time a "reference-style" pipeline on the
// same B200: one frame per call, AoS inputs read one float at a time, one thread per Gaussian preprocess,
// cub::DeviceScan::InclusiveSum, a BLOCKING device->host copy of the duplicate count every frame,
// duplicateWithKeys, cub::DeviceRadixSort::SortPairs on (tile << 32 | depth) keys over bits [0, 32 + msb),
// identifyTileRanges, and the 16x16-thread render kernel that stages 256 (id, xy, conic/opacity) entries per
// round in shared memory and leaves when __syncthreads_count(done) == 256.  Compiled with nvcc defaults
// (FMA contraction on), as upstream's setup.py does.  Every number derived from it is labelled
// "restated reference" (BASELINE.md §3.2 step 2).  Nothing in gsworld_b200/ links or loads this.
#include <cub/cub.cuh>
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdio>

namespace {

constexpr int BX = 16, BY = 16, BS = BX * BY;
__device__ const float C0 = 0.28209479177387814f, C1 = 0.4886025119029199f;
__device__ const float C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                -1.0925484305920792f, 0.5462742152960396f};
__device__ const float C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                0.3731763325901154f,  -0.4570457994644658f, 1.445305721320277f,
                                -0.5900435899266435f};

__device__ inline float ndc2Pix(float v, int S) { return ((v + 1.0) * S - 1.0) * 0.5; }

__device__ inline void getRect(float2 p, int r, uint2& mn, uint2& mx, dim3 grid) {
  mn = {min(grid.x, (unsigned)max(0, (int)((p.x - r) / BX))), min(grid.y, (unsigned)max(0, (int)((p.y - r) / BY)))};
  mx = {min(grid.x, (unsigned)max(0, (int)((p.x + r + BX - 1) / BX))),
        min(grid.y, (unsigned)max(0, (int)((p.y + r + BY - 1) / BY)))};
}

struct M3 { float m[3][3]; };  // m[c][r]
__device__ inline M3 mul(const M3& a, const M3& b) {
  M3 o;
  for (int c = 0; c < 3; ++c)
    for (int r = 0; r < 3; ++r) o.m[c][r] = a.m[0][r] * b.m[c][0] + a.m[1][r] * b.m[c][1] + a.m[2][r] * b.m[c][2];
  return o;
}
__device__ inline M3 tr(const M3& a) {
  M3 o;
  for (int c = 0; c < 3; ++c)
    for (int r = 0; r < 3; ++r) o.m[c][r] = a.m[r][c];
  return o;
}

__global__ void preprocessK(int P, int D, int M, const float* means, const float* scales, float mod, const float* rots,
                            const float* opac, const float* shs, const float* V, const float* PM, const float* cam,
                            int W, int H, float tanx, float tany, float fx, float fy, float near_plane, int* radii,
                            float2* xy, float* depths, float* rgb, float4* conic_o, dim3 grid, uint32_t* touched) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  radii[idx] = 0;
  touched[idx] = 0;
  const float3 p = {means[3 * idx], means[3 * idx + 1], means[3 * idx + 2]};
  const float3 pv = {V[0] * p.x + V[4] * p.y + V[8] * p.z + V[12], V[1] * p.x + V[5] * p.y + V[9] * p.z + V[13],
                     V[2] * p.x + V[6] * p.y + V[10] * p.z + V[14]};
  if (pv.z <= near_plane) return;
  const float hx = PM[0] * p.x + PM[4] * p.y + PM[8] * p.z + PM[12];
  const float hy = PM[1] * p.x + PM[5] * p.y + PM[9] * p.z + PM[13];
  const float hw = PM[3] * p.x + PM[7] * p.y + PM[11] * p.z + PM[15];
  const float pw = 1.0f / (hw + 0.0000001f);
  const float3 pp = {hx * pw, hy * pw, 0.f};
  M3 S{};
  S.m[0][0] = mod * scales[3 * idx];
  S.m[1][1] = mod * scales[3 * idx + 1];
  S.m[2][2] = mod * scales[3 * idx + 2];
  const float r = rots[4 * idx], x = rots[4 * idx + 1], y = rots[4 * idx + 2], z = rots[4 * idx + 3];
  M3 R;
  R.m[0][0] = 1.f - 2.f * (y * y + z * z); R.m[0][1] = 2.f * (x * y - r * z); R.m[0][2] = 2.f * (x * z + r * y);
  R.m[1][0] = 2.f * (x * y + r * z); R.m[1][1] = 1.f - 2.f * (x * x + z * z); R.m[1][2] = 2.f * (y * z - r * x);
  R.m[2][0] = 2.f * (x * z - r * y); R.m[2][1] = 2.f * (y * z + r * x); R.m[2][2] = 1.f - 2.f * (x * x + y * y);
  const M3 Mm = mul(S, R);
  const M3 Sig = mul(tr(Mm), Mm);
  float3 t = pv;
  const float limx = 1.3f * tanx, limy = 1.3f * tany;
  t.x = min(limx, max(-limx, t.x / t.z)) * t.z;
  t.y = min(limy, max(-limy, t.y / t.z)) * t.z;
  M3 J{};
  J.m[0][0] = fx / t.z; J.m[0][2] = -(fx * t.x) / (t.z * t.z);
  J.m[1][1] = fy / t.z; J.m[1][2] = -(fy * t.y) / (t.z * t.z);
  M3 Wm;
  Wm.m[0][0] = V[0]; Wm.m[0][1] = V[4]; Wm.m[0][2] = V[8];
  Wm.m[1][0] = V[1]; Wm.m[1][1] = V[5]; Wm.m[1][2] = V[9];
  Wm.m[2][0] = V[2]; Wm.m[2][1] = V[6]; Wm.m[2][2] = V[10];
  const M3 T = mul(Wm, J);
  M3 Vrk;
  Vrk.m[0][0] = Sig.m[0][0]; Vrk.m[0][1] = Sig.m[0][1]; Vrk.m[0][2] = Sig.m[0][2];
  Vrk.m[1][0] = Sig.m[0][1]; Vrk.m[1][1] = Sig.m[1][1]; Vrk.m[1][2] = Sig.m[1][2];
  Vrk.m[2][0] = Sig.m[0][2]; Vrk.m[2][1] = Sig.m[1][2]; Vrk.m[2][2] = Sig.m[2][2];
  const M3 cov = mul(mul(tr(T), tr(Vrk)), T);
  float cx = cov.m[0][0] + 0.3f, cy = cov.m[0][1], cz = cov.m[1][1] + 0.3f;
  const float det = cx * cz - cy * cy;
  if (det == 0.0f) return;
  const float di = 1.f / det;
  const float3 conic = {cz * di, -cy * di, cx * di};
  const float mid = 0.5f * (cx + cz);
  const float l1 = mid + sqrt(max(0.1f, mid * mid - det)), l2 = mid - sqrt(max(0.1f, mid * mid - det));
  const float my_radius = ceil(3.f * sqrt(max(l1, l2)));
  const float2 pim = {ndc2Pix(pp.x, W), ndc2Pix(pp.y, H)};
  uint2 mn, mx;
  getRect(pim, (int)my_radius, mn, mx, grid);
  if ((mx.x - mn.x) * (mx.y - mn.y) == 0) return;
  float3 dir = {p.x - cam[0], p.y - cam[1], p.z - cam[2]};
  const float len = sqrtf(dir.x * dir.x + dir.y * dir.y + dir.z * dir.z);
  dir.x /= len; dir.y /= len; dir.z /= len;
  const float* sh = shs + (size_t)idx * M * 3;
  for (int ch = 0; ch < 3; ++ch) {
    float res = C0 * sh[ch];
    if (D > 0) {
      const float X = dir.x, Y = dir.y, Z = dir.z;
      res = res - C1 * Y * sh[3 + ch] + C1 * Z * sh[6 + ch] - C1 * X * sh[9 + ch];
      if (D > 1) {
        const float xx = X * X, yy = Y * Y, zz = Z * Z, xy_ = X * Y, yz = Y * Z, xz = X * Z;
        res = res + C2[0] * xy_ * sh[12 + ch] + C2[1] * yz * sh[15 + ch] + C2[2] * (2.0f * zz - xx - yy) * sh[18 + ch] +
              C2[3] * xz * sh[21 + ch] + C2[4] * (xx - yy) * sh[24 + ch];
        if (D > 2)
          res = res + C3[0] * Y * (3.0f * xx - yy) * sh[27 + ch] + C3[1] * xy_ * Z * sh[30 + ch] +
                C3[2] * Y * (4.0f * zz - xx - yy) * sh[33 + ch] + C3[3] * Z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * sh[36 + ch] +
                C3[4] * X * (4.0f * zz - xx - yy) * sh[39 + ch] + C3[5] * Z * (xx - yy) * sh[42 + ch] +
                C3[6] * X * (xx - 3.0f * yy) * sh[45 + ch];
      }
    }
    rgb[3 * idx + ch] = max(res + 0.5f, 0.0f);
  }
  depths[idx] = pv.z;
  radii[idx] = (int)my_radius;
  xy[idx] = pim;
  conic_o[idx] = {conic.x, conic.y, conic.z, opac[idx]};
  touched[idx] = (mx.y - mn.y) * (mx.x - mn.x);
}

__global__ void duplicateK(int P, const float2* xy, const float* depths, const uint32_t* offsets, uint64_t* keys,
                           uint32_t* vals, const int* radii, dim3 grid) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P || radii[idx] <= 0) return;
  uint32_t off = idx == 0 ? 0 : offsets[idx - 1];
  uint2 mn, mx;
  getRect(xy[idx], radii[idx], mn, mx, grid);
  for (unsigned y = mn.y; y < mx.y; y++)
    for (unsigned x = mn.x; x < mx.x; x++) {
      uint64_t key = y * grid.x + x;
      key <<= 32;
      key |= *((const uint32_t*)&depths[idx]);
      keys[off] = key;
      vals[off] = idx;
      off++;
    }
}

__global__ void rangesK(int L, const uint64_t* keys, uint2* ranges) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= L) return;
  const uint32_t cur = keys[idx] >> 32;
  if (idx == 0) ranges[cur].x = 0;
  else {
    const uint32_t prev = keys[idx - 1] >> 32;
    if (cur != prev) {
      ranges[prev].y = idx;
      ranges[cur].x = idx;
    }
  }
  if (idx == L - 1) ranges[cur].y = L;
}

__global__ void __launch_bounds__(BS)
renderK(const uint2* ranges, const uint32_t* list, int W, int H, const float2* xy, const float* feat,
        const float4* conic_o, const float* bg, float* out, const float* depths, float* invdepth) {
  const uint32_t hb = (W + BX - 1) / BX;
  const uint2 pmin = {blockIdx.x * BX, blockIdx.y * BY};
  const uint2 pix = {pmin.x + threadIdx.x, pmin.y + threadIdx.y};
  const uint32_t pid = W * pix.y + pix.x;
  const float2 pf = {(float)pix.x, (float)pix.y};
  const bool inside = pix.x < W && pix.y < H;
  bool done = !inside;
  const uint2 range = ranges[blockIdx.y * hb + blockIdx.x];
  const int rounds = (range.y - range.x + BS - 1) / BS;
  int toDo = range.y - range.x;
  __shared__ int cid[BS];
  __shared__ float2 cxy[BS];
  __shared__ float4 cco[BS];
  float T = 1.0f, C[3] = {0, 0, 0}, ed = 0.f;
  const int rank = threadIdx.y * BX + threadIdx.x;
  for (int i = 0; i < rounds; i++, toDo -= BS) {
    if (__syncthreads_count(done) == BS) break;
    const int prog = i * BS + rank;
    if (range.x + prog < range.y) {
      const int c = list[range.x + prog];
      cid[rank] = c;
      cxy[rank] = xy[c];
      cco[rank] = conic_o[c];
    }
    __syncthreads();
    for (int j = 0; !done && j < min(BS, toDo); j++) {
      const float2 d = {cxy[j].x - pf.x, cxy[j].y - pf.y};
      const float4 co = cco[j];
      const float power = -0.5f * (co.x * d.x * d.x + co.z * d.y * d.y) - co.y * d.x * d.y;
      if (power > 0.0f) continue;
      const float alpha = min(0.99f, co.w * exp(power));
      if (alpha < 1.0f / 255.0f) continue;
      const float tT = T * (1 - alpha);
      if (tT < 0.0001f) { done = true; continue; }
      for (int ch = 0; ch < 3; ch++) C[ch] += feat[cid[j] * 3 + ch] * alpha * T;
      ed += (1 / depths[cid[j]]) * alpha * T;
      T = tT;
    }
  }
  if (inside) {
    for (int ch = 0; ch < 3; ch++) out[ch * H * W + pid] = C[ch] + T * bg[ch];
    invdepth[pid] = ed;
  }
}

struct Bufs {
  void* p[16] = {};
  size_t n[16] = {};
  void* get(int k, size_t bytes) {
    if (bytes > n[k]) {
      if (p[k]) cudaFree(p[k]);
      cudaMalloc(&p[k], bytes);
      n[k] = bytes;
    }
    return p[k];
  }
} g_b;

}  // namespace

// One frame.  All pointers are device pointers.  Returns the number of duplicates (tile entries) or -1.
extern "C" long long refgpu_rasterize_forward(int P, int D, int M, const float* bg, int W, int H, const float* means,
                                              const float* shs, const float* opac, const float* scales, float mod,
                                              const float* rots, const float* V, const float* PM, const float* cam,
                                              float tanx, float tany, float near_plane, float* out_color,
                                              int* out_radii, float* out_invdepth, cudaStream_t s) {
  const float fy = H / (2.0f * tany), fx = W / (2.0f * tanx);
  const dim3 grid((W + BX - 1) / BX, (H + BY - 1) / BY, 1), block(BX, BY, 1);
  cudaMemsetAsync(out_color, 0, sizeof(float) * 3 * H * W, s);
  cudaMemsetAsync(out_invdepth, 0, sizeof(float) * H * W, s);
  if (P == 0) return 0;
  float2* xy = (float2*)g_b.get(0, sizeof(float2) * P);
  float* depths = (float*)g_b.get(1, 4 * (size_t)P);
  float* rgb = (float*)g_b.get(2, 12 * (size_t)P);
  float4* co = (float4*)g_b.get(3, 16 * (size_t)P);
  uint32_t* touched = (uint32_t*)g_b.get(4, 4 * (size_t)P);
  uint32_t* offs = (uint32_t*)g_b.get(5, 4 * (size_t)P);
  preprocessK<<<(P + 255) / 256, 256, 0, s>>>(P, D, M, means, scales, mod, rots, opac, shs, V, PM, cam, W, H, tanx, tany,
                                              fx, fy, near_plane, out_radii, xy, depths, rgb, co, grid, touched);
  size_t tb = 0;
  cub::DeviceScan::InclusiveSum(nullptr, tb, touched, offs, P, s);
  void* tmp = g_b.get(6, tb);
  cub::DeviceScan::InclusiveSum(tmp, tb, touched, offs, P, s);
  uint32_t num = 0;
  cudaMemcpyAsync(&num, offs + P - 1, 4, cudaMemcpyDeviceToHost, s);
  cudaStreamSynchronize(s);  // the reference blocks here every frame
  uint2* ranges = (uint2*)g_b.get(7, sizeof(uint2) * grid.x * grid.y);
  cudaMemsetAsync(ranges, 0, sizeof(uint2) * grid.x * grid.y, s);
  if (num > 0) {
    uint64_t* k0 = (uint64_t*)g_b.get(8, 8 * (size_t)num);
    uint64_t* k1 = (uint64_t*)g_b.get(9, 8 * (size_t)num);
    uint32_t* v0 = (uint32_t*)g_b.get(10, 4 * (size_t)num);
    uint32_t* v1 = (uint32_t*)g_b.get(11, 4 * (size_t)num);
    duplicateK<<<(P + 255) / 256, 256, 0, s>>>(P, xy, depths, offs, k0, v0, out_radii, grid);
    int bit = 0;
    for (uint32_t t = grid.x * grid.y; t; t >>= 1) ++bit;
    size_t sb = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, sb, k0, k1, v0, v1, (int)num, 0, 32 + bit, s);
    void* st = g_b.get(12, sb);
    cub::DeviceRadixSort::SortPairs(st, sb, k0, k1, v0, v1, (int)num, 0, 32 + bit, s);
    rangesK<<<(num + 255) / 256, 256, 0, s>>>((int)num, k1, ranges);
    renderK<<<grid, block, 0, s>>>(ranges, v1, W, H, xy, rgb, co, bg, out_color, depths, out_invdepth);
  } else {
    renderK<<<grid, block, 0, s>>>(ranges, nullptr, W, H, xy, rgb, co, bg, out_color, depths, out_invdepth);
  }
  return cudaGetLastError() == cudaSuccess ? (long long)num : -1;
}
