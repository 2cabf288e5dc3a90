// binsort.cu — tile-list construction: exclusive scan of bin counts, per-bin sort by the full
// (depth bits, Gaussian index) key with record gather, and the static/dynamic list merge.  sm_100a.
//
// Replaces the reference pipeline InclusiveSum -> blocking D2H -> duplicateWithKeys -> global 64-bit
// radix sort -> identifyTileRanges (diff-gaussian-rasterization rasterizer_impl.cu, absent submodule):
// entries are binned by (list, tile[, depth bucket]) with counting + scatter, then each bin is sorted
// independently in shared memory.  Sorting by (depth bits << 32 | original index) gives exactly the
// order of the reference's stable radix sort on (tile << 32 | depth bits) over index-ordered input.
#include "gsw_internal.h"

namespace gsw {
namespace {

constexpr int kScanBlock = 1024;
constexpr int kScanItems = 4;  // per thread => 4096 per block

// ---- 3-kernel exclusive scan (u32) ----
__global__ void __launch_bounds__(kScanBlock)
scan_local_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, int64_t n,
                  uint32_t* __restrict__ block_sums) {
  __shared__ uint32_t warp_sums[32];
  const int64_t base = ((int64_t)blockIdx.x * kScanBlock + threadIdx.x) * kScanItems;
  uint32_t v[kScanItems];
  uint32_t tsum = 0;
#pragma unroll
  for (int k = 0; k < kScanItems; ++k) {
    v[k] = (base + k < n) ? in[base + k] : 0u;
    tsum += v[k];
  }
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  uint32_t incl = tsum;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += t;
  }
  if (lane == 31) warp_sums[wid] = incl;
  __syncthreads();
  if (wid == 0) {
    uint32_t w = warp_sums[lane];
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, w, d);
      if (lane >= d) w += t;
    }
    warp_sums[lane] = w;  // inclusive over warps
  }
  __syncthreads();
  uint32_t run = incl - tsum + (wid > 0 ? warp_sums[wid - 1] : 0u);
#pragma unroll
  for (int k = 0; k < kScanItems; ++k) {
    if (base + k < n) out[base + k] = run;
    run += v[k];
  }
  if (threadIdx.x == kScanBlock - 1) block_sums[blockIdx.x] = run;
}

// single block: exclusive scan of block sums in place (nb <= 1024*1024 handled by looping with carry)
__global__ void __launch_bounds__(kScanBlock)
scan_sums_kernel(uint32_t* __restrict__ sums, int64_t nb, uint32_t* __restrict__ total) {
  __shared__ uint32_t warp_sums[32];
  __shared__ uint32_t carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  for (int64_t base = 0; base < nb; base += kScanBlock) {
    const int64_t i = base + threadIdx.x;
    const uint32_t v = i < nb ? sums[i] : 0u;
    uint32_t incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += t;
    }
    if (lane == 31) warp_sums[wid] = incl;
    __syncthreads();
    if (wid == 0) {
      uint32_t w = warp_sums[lane];
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, w, d);
        if (lane >= d) w += t;
      }
      warp_sums[lane] = w;
    }
    __syncthreads();
    const uint32_t carry = carry_s;
    const uint32_t excl = carry + incl - v + (wid > 0 ? warp_sums[wid - 1] : 0u);
    if (i < nb) sums[i] = excl;
    __syncthreads();
    if (threadIdx.x == kScanBlock - 1) carry_s = excl + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) *total = carry_s;
}

// adds block offsets; also writes out[n] = total (one-past-the-end sentinel)
__global__ void __launch_bounds__(kScanBlock)
scan_add_kernel(uint32_t* __restrict__ out, int64_t n, const uint32_t* __restrict__ sums,
                const uint32_t* __restrict__ total) {
  const int64_t base = ((int64_t)blockIdx.x * kScanBlock + threadIdx.x) * kScanItems;
  const uint32_t add = sums[blockIdx.x];
#pragma unroll
  for (int k = 0; k < kScanItems; ++k)
    if (base + k < n) out[base + k] += add;
  if (blockIdx.x == 0 && threadIdx.x == 0) out[n] = *total;
}

// ---- bitonic network with every comparator oriented ascending (first step of each merge mirrors) ----
// Works in place for any n: slots >= n are treated as +inf and never move.
template <typename KeyPtr>
__device__ __forceinline__ void bitonic_sort(KeyPtr a, uint32_t n, uint32_t n2) {
  for (uint32_t k = 2; k <= n2; k <<= 1) {
    const uint32_t half = k >> 1;
    for (uint32_t t = threadIdx.x; t < (n2 >> 1); t += blockDim.x) {
      const uint32_t blk = (t / half) * k, off = t % half;
      const uint32_t i = blk + off, j = blk + k - 1 - off;
      if (j < n) {
        const unsigned long long x = a[i], y = a[j];
        if (x > y) {
          a[i] = y;
          a[j] = x;
        }
      }
    }
    __syncthreads();
    for (uint32_t s = half >> 1; s >= 1; s >>= 1) {
      for (uint32_t t = threadIdx.x; t < (n2 >> 1); t += blockDim.x) {
        const uint32_t i = 2 * s * (t / s) + (t % s), j = i + s;
        if (j < n) {
          const unsigned long long x = a[i], y = a[j];
          if (x > y) {
            a[i] = y;
            a[j] = x;
          }
        }
      }
      __syncthreads();
    }
  }
}

__device__ __forceinline__ uint32_t pow2ceil(uint32_t n) {
  return n <= 1 ? 1u : 1u << (32 - __clz(n - 1));
}

// BIG=false: one CTA per bin, bins with n <= kSortSmemKeys sorted in shared memory.
// BIG=true: a small persistent grid strides over all bins and sorts the (rare) larger ones directly in
// global memory (L2) with the same network, one CTA per bin.
template <bool BIG>
__global__ void __launch_bounds__(512)
sort_bins_kernel(const uint32_t* __restrict__ bin_off, int64_t nbins, int32_t V, int64_t N, int64_t Nm,
                 int32_t bins_per_list, const int32_t* __restrict__ mov_pos, unsigned long long* __restrict__ keys,
                 const float4* __restrict__ recA, const float4* __restrict__ recB, const float4* __restrict__ recC,
                 Splat* __restrict__ splats, int64_t cap) {
  extern __shared__ unsigned long long skeys[];
  for (int64_t bin = blockIdx.x; bin < nbins; bin += gridDim.x) {
    const uint32_t start = bin_off[bin];
    const uint32_t end = bin_off[bin + 1];
    if (end <= start || (int64_t)end > cap) continue;
    const uint32_t n = end - start;
    if (BIG != (n > (uint32_t)kSortSmemKeys)) continue;
    const uint32_t n2 = pow2ceil(n);
    unsigned long long* g = keys + start;
    if (!BIG) {
      for (uint32_t t = threadIdx.x; t < n; t += blockDim.x) skeys[t] = g[t];
      __syncthreads();
      if (n > 1) bitonic_sort(skeys, n, n2);
    } else {
      __syncthreads();
      bitonic_sort(g, n, n2);
    }
    // gather the sorted records into the blend stream.  The conic is pre-scaled so the blend evaluates
    // alpha = o * 2^(A' dx^2 + B' dx dy + C' dy^2) with A' = -0.5 log2(e) A, B' = -log2(e) B, C' = -0.5 log2(e) C.
    const int64_t list = bin / bins_per_list;
    const bool dyn = list >= V;
    const int64_t item_base = dyn ? (int64_t)V * N + (list - V) * Nm : list * N;
    constexpr float kLog2e = 1.4426950408889634f;
    for (uint32_t t = threadIdx.x; t < n; t += blockDim.x) {
      const unsigned long long key = BIG ? g[t] : skeys[t];
      if (!BIG) g[t] = key;
      const uint32_t idx = (uint32_t)(key & 0xffffffffull);
      const int64_t item = item_base + (dyn ? (int64_t)mov_pos[idx] : (int64_t)idx);
      const float4 a = recA[item], b = recB[item], c = recC[item];
      Splat sp;
      sp.a = a;
      sp.b = make_float4(-0.5f * kLog2e * b.x, -kLog2e * b.y, -0.5f * kLog2e * b.z, b.w);
      sp.c = make_float4(c.x, c.y, c.z, 1 / c.w);
      splats[start + t] = sp;
    }
    __syncthreads();  // skeys is reused by the next bin of this CTA (BIG grid only loops)
  }
}

// Debug: lists of one frame in blend order (static view list merged with the frame's dynamic list), in the
// reference's key format (tile << 32 | depth bits) + Gaussian index.
__global__ void frame_counts_kernel(int32_t V, int32_t T, int32_t NB, int32_t frame, const int32_t* frame_view,
                                    const uint32_t* bin_off, uint32_t* cnt) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= T) return;
  const int v = frame_view ? frame_view[frame] : 0;
  const int64_t sb = ((int64_t)v * T + t) * NB, db = ((int64_t)(V + frame) * T + t) * NB;
  cnt[t] = (bin_off[sb + NB] - bin_off[sb]) + (bin_off[db + NB] - bin_off[db]);
}
__global__ void frame_lists_kernel(int32_t V, int32_t T, int32_t NB, int32_t frame, const int32_t* frame_view,
                                   const uint32_t* bin_off, const uint32_t* off, const unsigned long long* keys,
                                   int32_t* ranges, unsigned long long* okeys, uint32_t* ovals, int64_t cap) {
  const int tile = blockIdx.x;
  const int v = frame_view ? frame_view[frame] : 0;
  const int64_t sb = ((int64_t)v * T + tile) * NB, db = ((int64_t)(V + frame) * T + tile) * NB;
  const uint32_t s0 = bin_off[sb], ns = bin_off[sb + NB] - s0;
  const uint32_t d0 = bin_off[db], nd = bin_off[db + NB] - d0;
  const uint32_t o = off[tile], n = ns + nd;
  if (threadIdx.x == 0) {
    // empty tiles keep (0,0) like the reference's memset + identifyTileRanges
    ranges[2 * tile] = n ? (int32_t)o : 0;
    ranges[2 * tile + 1] = n ? (int32_t)(o + n) : 0;
  }
  const unsigned long long* ks = keys + s0;
  const unsigned long long* kd = keys + d0;
  for (uint32_t t = threadIdx.x; t < n; t += blockDim.x) {
    const bool from_s = t < ns;
    const uint32_t k = from_s ? t : t - ns;
    const unsigned long long key = from_s ? ks[k] : kd[k];
    const unsigned long long* other = from_s ? kd : ks;
    uint32_t lo = 0, hi = from_s ? nd : ns;
    while (lo < hi) {  // number of keys in the other list smaller than key (keys are unique)
      const uint32_t mid = (lo + hi) >> 1;
      if (other[mid] < key) lo = mid + 1; else hi = mid;
    }
    const int64_t dst = (int64_t)o + k + lo;
    if (dst >= cap) continue;
    okeys[dst] = ((unsigned long long)tile << 32) | (key >> 32);
    ovals[dst] = (uint32_t)(key & 0xffffffffull);
  }
}

}  // namespace

size_t scan_tmp_elems(int64_t n) {
  const int64_t per = (int64_t)kScanBlock * kScanItems;
  return (size_t)((n + per - 1) / per) + 2;
}

// out_excl must hold n+1 elements (out[n] = total).  tmp: scan_tmp_elems(n) u32.
int launch_scan_u32(const uint32_t* in, uint32_t* out_excl, int64_t n, uint32_t* tmp, cudaStream_t s) {
  if (n <= 0) {
    GSW_CUDA(cudaMemsetAsync(out_excl, 0, sizeof(uint32_t), s));
    return 0;
  }
  const int64_t per = (int64_t)kScanBlock * kScanItems;
  const int64_t nb = (n + per - 1) / per;
  uint32_t* sums = tmp;
  uint32_t* total = tmp + nb;
  scan_local_kernel<<<(unsigned)nb, kScanBlock, 0, s>>>(in, out_excl, n, sums);
  scan_sums_kernel<<<1, kScanBlock, 0, s>>>(sums, nb, total);
  scan_add_kernel<<<(unsigned)nb, kScanBlock, 0, s>>>(out_excl, n, sums, total);
  count_launches(3);
  GSW_CUDA(cudaGetLastError());
  return 0;
}

int launch_sort_bins(const gsw_ctx* ctx, int64_t nbins, int32_t V, int64_t N, int64_t Nm, const FrameGeom& g,
                     int64_t cap, cudaStream_t s) {
  if (nbins <= 0) return 0;
  auto& w = ctx->ws;
  static bool attr_set = false;
  const size_t smem = (size_t)kSortSmemKeys * sizeof(unsigned long long);
  if (!attr_set) {
    GSW_CUDA(cudaFuncSetAttribute(sort_bins_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_set = true;
  }
  const int32_t bpl = g.T * g.NB;
  sort_bins_kernel<false><<<(unsigned)nbins, 512, smem, s>>>(
      w.bin_off.as<uint32_t>(), nbins, V, N, Nm, bpl, ctx->mov_pos.as<int32_t>(), w.keys.as<unsigned long long>(),
      w.recA.as<float4>(), w.recB.as<float4>(), w.recC.as<float4>(), w.splats.as<Splat>(), cap);
  const unsigned big_grid = (unsigned)(nbins < 296 ? nbins : 296);
  sort_bins_kernel<true><<<big_grid, 512, 0, s>>>(
      w.bin_off.as<uint32_t>(), nbins, V, N, Nm, bpl, ctx->mov_pos.as<int32_t>(), w.keys.as<unsigned long long>(),
      w.recA.as<float4>(), w.recB.as<float4>(), w.recC.as<float4>(), w.splats.as<Splat>(), cap);
  count_launches(2);
  GSW_CUDA(cudaGetLastError());
  return 0;
}

int launch_frame_lists(const gsw_ctx* ctx, int32_t frame, int32_t* ranges, uint64_t* keys, uint32_t* vals,
                       int64_t cap, cudaStream_t s) {
  auto& w = ctx->ws;
  const auto& L = ctx->last;
  const int32_t T = L.g.T;
  // per-tile counts -> exclusive offsets (scratch: merge_cnt / merge_off hold >= T+1 elements)
  frame_counts_kernel<<<(T + 255) / 256, 256, 0, s>>>(L.V, T, L.g.NB, frame, L.frame_view, w.bin_off.as<uint32_t>(),
                                                      w.merge_cnt.as<uint32_t>());
  int rc = launch_scan_u32(w.merge_cnt.as<uint32_t>(), w.merge_off.as<uint32_t>(), T, w.scan_tmp.as<uint32_t>(), s);
  if (rc) return rc;
  frame_lists_kernel<<<T, 256, 0, s>>>(L.V, T, L.g.NB, frame, L.frame_view, w.bin_off.as<uint32_t>(),
                                       w.merge_off.as<uint32_t>(), w.keys.as<unsigned long long>(), ranges,
                                       reinterpret_cast<unsigned long long*>(keys), vals, cap);
  count_launches(2);
  GSW_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace gsw
