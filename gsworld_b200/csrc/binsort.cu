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

// ---- single-CTA exclusive scan: the bin counts of a small step are a few thousand words, for
// which three dependent launches cost more in launch gaps than in work.  One CTA walks the array in tiles of 4096 with
// a running carry.  out[n] = total.  Used up to kScanSingleMax elements.
__global__ void __launch_bounds__(kScanBlock)
scan_single_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, int64_t n) {
  __shared__ uint32_t warp_sums[32];
  __shared__ uint32_t carry_s;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  for (int64_t tile = 0; tile < n; tile += (int64_t)kScanBlock * kScanItems) {
    const int64_t base = tile + (int64_t)threadIdx.x * kScanItems;
    uint32_t v[kScanItems];
    uint32_t tsum = 0;
#pragma unroll
    for (int k = 0; k < kScanItems; ++k) {
      v[k] = (base + k < n) ? in[base + k] : 0u;
      tsum += v[k];
    }
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
    const uint32_t carry = carry_s;
    uint32_t run = carry + incl - tsum + (wid > 0 ? warp_sums[wid - 1] : 0u);
#pragma unroll
    for (int k = 0; k < kScanItems; ++k) {
      if (base + k < n) out[base + k] = run;
      run += v[k];
    }
    __syncthreads();  // everybody has read carry_s and warp_sums
    if (threadIdx.x == kScanBlock - 1) carry_s = run;
    __syncthreads();
  }
  if (threadIdx.x == 0) out[n] = carry_s;
}
// Measured (step, ms): 4.6 k bins (8 envs x 2 cameras) 0.360 -> 0.349 with the single CTA; 33 k bins (64 envs) 1.008 ->
// 1.018, where nine dependent tiles take longer than the three launches.  Four tiles at most.
constexpr int64_t kScanSingleMax = 4 * (int64_t)kScanBlock * kScanItems;

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

// ---- shared-memory merge sort of one segment (n <= kSortSmemKeys), 16 keys per thread step ----
typedef unsigned long long u64;

__device__ __forceinline__ void cswap(u64& a, u64& b) {
  const u64 lo = a < b ? a : b, hi = a < b ? b : a;
  a = lo;
  b = hi;
}

// 16-key sorting network with 60 comparators in 10 layers (Green's network; checked exhaustively with the
// 0-1 principle by tests/test_sort_network.py)
__device__ __forceinline__ void sort16(u64 (&k)[16]) {
  cswap(k[0], k[13]); cswap(k[1], k[12]); cswap(k[2], k[15]); cswap(k[3], k[14]); cswap(k[4], k[8]); cswap(k[5], k[6]); cswap(k[7], k[11]); cswap(k[9], k[10]);
  cswap(k[0], k[5]); cswap(k[1], k[7]); cswap(k[2], k[9]); cswap(k[3], k[4]); cswap(k[6], k[13]); cswap(k[8], k[14]); cswap(k[10], k[15]); cswap(k[11], k[12]);
  cswap(k[0], k[1]); cswap(k[2], k[3]); cswap(k[4], k[5]); cswap(k[6], k[8]); cswap(k[7], k[9]); cswap(k[10], k[11]); cswap(k[12], k[13]); cswap(k[14], k[15]);
  cswap(k[0], k[2]); cswap(k[1], k[3]); cswap(k[4], k[10]); cswap(k[5], k[11]); cswap(k[6], k[7]); cswap(k[8], k[9]); cswap(k[12], k[14]); cswap(k[13], k[15]);
  cswap(k[1], k[2]); cswap(k[3], k[12]); cswap(k[4], k[6]); cswap(k[5], k[7]); cswap(k[8], k[10]); cswap(k[9], k[11]); cswap(k[13], k[14]);
  cswap(k[1], k[4]); cswap(k[2], k[6]); cswap(k[5], k[8]); cswap(k[7], k[10]); cswap(k[9], k[13]); cswap(k[11], k[14]);
  cswap(k[2], k[4]); cswap(k[3], k[6]); cswap(k[9], k[12]); cswap(k[11], k[13]);
  cswap(k[3], k[5]); cswap(k[6], k[8]); cswap(k[7], k[9]); cswap(k[10], k[12]);
  cswap(k[3], k[4]); cswap(k[5], k[6]); cswap(k[7], k[8]); cswap(k[9], k[10]); cswap(k[11], k[12]);
  cswap(k[6], k[7]); cswap(k[8], k[9]);
}

// Shared-memory layout of the sort buffer: logical element e lives at e ^ ((e >> 4) & 15), a permutation
// inside every aligned block of 16 keys.  With the plain layout a warp whose lanes each own 16 consecutive
// 64-bit keys (the chunk loads/stores below, lane stride 128 B) would hit ONE bank pair -> 32 wavefronts per
// access; with the swizzle a half-warp covers all 32 banks once -> 2 wavefronts, the minimum for 256 B.
// The merge-path reads (lane stride ~8 keys) spread the same way.
__device__ __forceinline__ uint32_t swz(uint32_t e) { return e ^ ((e >> 4) & 15u); }

// Sorts buf[0..n) in place (logically padded to nP = roundup(n,16) with +inf; buf holds the swizzled layout).
// Registers first: every thread sorts 16 keys with the network.  Then log2(nP/16) merge rounds: a thread finds
// the merge-path split of its 16 outputs by binary search and merges sequentially, keeping the two run heads
// in registers; all threads write back after a barrier, so one buffer suffices.  Instruction count is what
// bounds this kernel (ncu: issue-bound once the bank conflicts are gone), hence 16 keys per search and the
// branch-free inner step.  CPT = max 16-key chunks per thread = capacity / (16 * THREADS).
template <int THREADS, int CPT>
__device__ __forceinline__ void smem_merge_sort(u64* buf, uint32_t n) {
  const uint32_t nP = (n + 15u) & ~15u;
  const uint32_t chunks = nP >> 4;
#pragma unroll
  for (int q = 0; q < CPT; ++q) {
    const uint32_t c = threadIdx.x + q * THREADS;
    if (c < chunks) {
      const uint32_t x = c & 15u;  // swz(16c + i) = 16c + (i ^ x)
      u64 k[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) k[i] = (16 * c + i < n) ? buf[16 * c + (i ^ x)] : ~0ull;
      sort16(k);
#pragma unroll
      for (int i = 0; i < 16; ++i) buf[16 * c + (i ^ x)] = k[i];
    }
  }
  __syncthreads();
  for (uint32_t L = 16; L < nP; L <<= 1) {
    u64 out[CPT][16];
#pragma unroll
    for (int q = 0; q < CPT; ++q) {
      const uint32_t c = threadIdx.x + q * THREADS;
      if (c < chunks) {
        const uint32_t o = 16 * c;
        const uint32_t A = o & ~(2 * L - 1), B = A + L;  // logical starts of the two runs
        const uint32_t lenA = min(L, nP - A);
        const uint32_t lenB = (nP - A > L) ? min(L, nP - B) : 0u;
        const uint32_t d = o - A;
        uint32_t lo = d > lenB ? d - lenB : 0u, hi = min(d, lenA);
        while (lo < hi) {
          const uint32_t mid = (lo + hi) >> 1;
          if (buf[swz(A + mid)] < buf[swz(B + d - 1 - mid)]) lo = mid + 1; else hi = mid;
        }
        uint32_t ia = A + lo, ib = B + (d - lo);
        const uint32_t endA = A + lenA, endB = B + lenB;
        // exhausted runs read as +inf (no real key is: depth bits of a finite positive float), so the head
        // comparison alone decides
        u64 ka = ia < endA ? buf[swz(ia)] : ~0ull, kb = ib < endB ? buf[swz(ib)] : ~0ull;
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const bool ta = ka <= kb;
          out[q][i] = ta ? ka : kb;
          if (i < 15) {
            ia += ta ? 1u : 0u;
            ib += ta ? 0u : 1u;
            const uint32_t p = ta ? ia : ib;
            const u64 v = p < (ta ? endA : endB) ? buf[swz(p)] : ~0ull;
            ka = ta ? v : ka;
            kb = ta ? kb : v;
          }
        }
      }
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < CPT; ++q) {
      const uint32_t c = threadIdx.x + q * THREADS;
      if (c < chunks) {
        const uint32_t x = c & 15u;
#pragma unroll
        for (int i = 0; i < 16; ++i) buf[16 * c + (i ^ x)] = out[q][i];
      }
    }
    __syncthreads();
  }
}

struct Segment {  // a slice of the key array that is sorted independently (a whole bin or a depth sub-range of one)
  uint32_t start, count, bin, pad;
};

// One thread per tile entry of a SHARED view's static list (blockIdx.y = view): fully parallel random gather (two
// aligned 32 B sectors per entry), coalesced key reads and splat writes.
// Only static lists of views shared by several frames are materialised (their records are re-read by every
// frame of the view and by the view pass).  Static lists private to one frame and all dynamic lists are not:
// part of their entries is never reached before the pixels saturate, so the blend fetches those records itself.
// CTAs of unshared views leave at once (with one view per env that is all but one of them).
__global__ void __launch_bounds__(256)
gather_splats_kernel(const uint32_t* __restrict__ bin_off, int64_t nbins, int32_t V, int64_t Ns,
                     int32_t bins_per_list, const int32_t* __restrict__ pos_of,
                     const u64* __restrict__ keys, const float4* __restrict__ rec, Splat* __restrict__ splats,
                     int64_t cap, const int32_t* __restrict__ labels, int32_t* __restrict__ lab_stream,
                     const uint32_t* __restrict__ view_cnt, const uint32_t* __restrict__ cache_m) {
  const int v = blockIdx.y;
  if (view_cnt[v] < 2u) return;  // fetched lazily by the blend
  if (cache_m != nullptr && (uint32_t)v < *cache_m) return;  // gathered in an earlier step, still in place
  if ((int64_t)bin_off[nbins] > cap) return;
  const uint32_t p0 = bin_off[(int64_t)v * bins_per_list], p1 = bin_off[(int64_t)(v + 1) * bins_per_list];
  for (uint32_t p = p0 + blockIdx.x * blockDim.x + threadIdx.x; p < p1; p += gridDim.x * blockDim.x) {
    const u64 key = keys[p];
    const uint32_t idx = (uint32_t)(key & 0xffffffffull);
    const int64_t pos = pos_of ? (int64_t)pos_of[idx] : (int64_t)idx;  // position inside the static layer
    const int64_t item = (int64_t)v * Ns + pos;
    splats[p] = make_splat(rec, item);
    if (lab_stream != nullptr) lab_stream[p] = labels[idx];  // segmentation observation only
  }
}

// Work classification: every non-empty bin goes to one of four queues by size, so that the sort CTAs
// are sized for their segments (a thread sorts 16 keys; idle threads only wait at the barriers of the merge
// rounds and hold warp slots) and no CTA is launched for an empty bin.
//   queue 0 (small):  n <= kSmallKeys (1024)   ->  64-thread CTAs,  8 KB shared memory
//   queue 3 (mid):    n <= kMidKeys (2048)     -> 128-thread CTAs, 16 KB
//   queue 1 (medium): n <= kSortSmemKeys       -> 256-thread CTAs, 32 KB
//   queue 2 (huge):   larger                   -> refine_big_bins_kernel splits by depth into mid/medium segments
constexpr int kSmallKeys = 1024;
constexpr int kMidKeys = 2048;
struct Queues {
  Segment* q[4];
  uint32_t* count;  // [4]
  uint32_t cap[4];
};

__global__ void classify_bins_kernel(const uint32_t* __restrict__ bin_off, int64_t nbins, int64_t cap, Queues Q,
                                     long long* __restrict__ status, const uint32_t* __restrict__ cache_m,
                                     int32_t bins_per_list) {
  const int64_t bin = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (bin >= nbins) return;
  if (cache_m != nullptr && bin < (int64_t)*cache_m * bins_per_list) return;  // list kept (and sorted) from the last step
  const uint32_t start = bin_off[bin], end = bin_off[bin + 1];
  if (end <= start || (int64_t)end > cap) return;
  const uint32_t n = end - start;
  const int k = n <= (uint32_t)kSmallKeys ? 0 : (n <= (uint32_t)kMidKeys ? 3 : (n <= (uint32_t)kSortSmemKeys ? 1 : 2));
  const uint32_t slot = atomicAdd(&Q.count[k], 1u);
  if (slot < Q.cap[k]) {
    Segment sg;
    sg.start = start;
    sg.count = n;
    sg.bin = (uint32_t)bin;
    sg.pad = 0;
    Q.q[k][slot] = sg;
  } else {
    status[2] = 3;
    status[kStatusPoison] = 1;
  }
}

// Persistent CTAs pull segments of one size class (ticket counter: segment sizes vary by 4x inside a class)
// and sort their keys in shared memory.
// Compiled for 1024 resident threads per SM (64 registers, ~80 B of spills) rather than for the 93 registers the
// merge step would like: the kernels wait on barriers and shared-memory latency, more resident warps hide it
// (per-env cameras: sort 0.66 -> 0.60 ms).  GSW_SORT_MINB=0 builds the uncapped form.
#ifndef GSW_SORT_MINB
#define GSW_SORT_MINB 1
#endif
template <int CAP, int THREADS>
__global__ void __launch_bounds__(THREADS, GSW_SORT_MINB ? 1024 / THREADS : 1)
sort_segments_kernel(const Segment* __restrict__ queue, const uint32_t* __restrict__ queue_count, uint32_t queue_cap,
                     u64* __restrict__ keys, uint32_t* __restrict__ ticket) {
  extern __shared__ u64 sbuf[];
  __shared__ uint32_t s_next;
  const uint32_t nseg = min(*queue_count, queue_cap);
  for (uint32_t sidx = blockIdx.x; sidx < nseg;) {
    const Segment sg = queue[sidx];
    if (threadIdx.x == 0) s_next = gridDim.x + atomicAdd(ticket, 1u);
    const uint32_t start = sg.start, n = sg.count;
    for (uint32_t t = threadIdx.x; t < n; t += THREADS) sbuf[swz(t)] = keys[start + t];
    __syncthreads();
    smem_merge_sort<THREADS, CAP / (16 * THREADS)>(sbuf, n);
    for (uint32_t t = threadIdx.x; t < n; t += THREADS) keys[start + t] = sbuf[swz(t)];
    sidx = s_next;    // written before the barriers inside the sort
    __syncthreads();  // buffer (and s_next) are reused by the next segment of this CTA
  }
}

// Bins larger than kSortSmemKeys: partition the bin by depth into sub-segments that fit shared memory.
// One CTA per huge bin.  256 buckets linear in the depth bits between the bin's min and max; keys are
// moved bucket by bucket through scratch space (the bin's own, not yet written, splat slice), then
// consecutive buckets are grouped greedily into segments <= kSortSmemKeys and appended to the medium
// queue.  A single bucket that is still too large (many equal depths) is sorted in place with the
// global-memory bitonic network.
constexpr int kRefineBuckets = 256;
__global__ void __launch_bounds__(512)
refine_big_bins_kernel(Queues Q, u64* __restrict__ keys, Splat* __restrict__ splats,
                       long long* __restrict__ status) {
  __shared__ uint32_t hist[kRefineBuckets + 1];
  __shared__ uint32_t cursor[kRefineBuckets];
  __shared__ uint32_t red_min[16], red_max[16];
  __shared__ uint32_t s_min, s_max;
  __shared__ uint32_t big_lo[8], big_n[8];
  __shared__ uint32_t n_big;
  const uint32_t nh = min(Q.count[2], Q.cap[2]);
  for (uint32_t h = blockIdx.x; h < nh; h += gridDim.x) {
    const Segment hs = Q.q[2][h];
    const uint32_t start = hs.start, n = hs.count;
    const uint32_t bin = hs.bin;
    u64* K = keys + start;
    u64* S = reinterpret_cast<u64*>(splats + start);  // scratch: 48 B per entry available, 8 used
    // 1. depth range
    uint32_t mn = 0xffffffffu, mx = 0u;
#pragma unroll 4
    for (uint32_t t = threadIdx.x; t < n; t += blockDim.x) {
      const uint32_t d = (uint32_t)(K[t] >> 32);
      mn = min(mn, d);
      mx = max(mx, d);
    }
    for (int o = 16; o > 0; o >>= 1) {
      mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, o));
      mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    if ((threadIdx.x & 31) == 0) {
      red_min[threadIdx.x >> 5] = mn;
      red_max[threadIdx.x >> 5] = mx;
    }
    for (int k = threadIdx.x; k <= kRefineBuckets; k += blockDim.x) hist[k] = 0u;
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t a = 0xffffffffu, b = 0u;
      for (int k = 0; k < (int)(blockDim.x >> 5); ++k) {
        a = min(a, red_min[k]);
        b = max(b, red_max[k]);
      }
      s_min = a;
      s_max = b;
    }
    __syncthreads();
    const uint32_t dmin = s_min;
    const u64 range = (u64)(s_max - dmin) + 1ull;
    auto bucket_of = [&](u64 key) -> uint32_t {
      return (uint32_t)((((key >> 32) - dmin) * (u64)kRefineBuckets) / range);
    };
    // 2. histogram
#pragma unroll 4
    for (uint32_t t = threadIdx.x; t < n; t += blockDim.x) atomicAdd(&hist[bucket_of(K[t])], 1u);
    __syncthreads();
    // 3. exclusive scan of 256 counters by warp 0 (8 per lane)
    if (threadIdx.x < 32) {
      uint32_t v[8], sum = 0;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        v[k] = hist[threadIdx.x * 8 + k];
        sum += v[k];
      }
      uint32_t incl = sum;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
        if ((int)threadIdx.x >= d) incl += t;
      }
      uint32_t run = incl - sum;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        hist[threadIdx.x * 8 + k] = run;
        cursor[threadIdx.x * 8 + k] = run;
        run += v[k];
      }
      if (threadIdx.x == 31) hist[kRefineBuckets] = run;
    }
    __syncthreads();
    // 4. scatter to scratch, copy back
#pragma unroll 4
    for (uint32_t t = threadIdx.x; t < n; t += blockDim.x) {
      const u64 key = K[t];
      S[atomicAdd(&cursor[bucket_of(key)], 1u)] = key;
    }
    __syncthreads();
#pragma unroll 4
    for (uint32_t t = threadIdx.x; t < n; t += blockDim.x) K[t] = S[t];
    __syncthreads();
    // 5. group buckets into segments (thread 0), oversize buckets sorted here
    if (threadIdx.x == 0) {
      n_big = 0;
      uint32_t seg_lo = 0;
      auto emit = [&](uint32_t lo, uint32_t hi) {
        if (hi <= lo) return;
        if (hi - lo > (uint32_t)kSortSmemKeys) {  // one bucket, still too large
          if (n_big < 8) {
            big_lo[n_big] = lo;
            big_n[n_big] = hi - lo;
            ++n_big;
          } else {
            status[2] = 2;
            status[kStatusPoison] = 1;
          }
          return;
        }
        const int k = hi - lo <= (uint32_t)kMidKeys ? 3 : 1;
        const uint32_t q = atomicAdd(&Q.count[k], 1u);
        if (q < Q.cap[k]) {
          Segment sg;
          sg.start = start + lo;
          sg.count = hi - lo;
          sg.bin = bin;
          sg.pad = 0;
          Q.q[k][q] = sg;
        } else {
          status[2] = 3;
          status[kStatusPoison] = 1;
        }
      };
      for (int k = 0; k < kRefineBuckets; ++k) {
        const uint32_t b_hi = hist[k + 1];
        if (b_hi - seg_lo > (uint32_t)kSortSmemKeys) {
          emit(seg_lo, hist[k]);  // close the segment before bucket k
          seg_lo = hist[k];
          if (b_hi - seg_lo > (uint32_t)kSortSmemKeys) {  // bucket k alone is oversize
            emit(seg_lo, b_hi);
            seg_lo = b_hi;
          }
        }
      }
      emit(seg_lo, n);
    }
    __syncthreads();
    for (uint32_t k = 0; k < n_big; ++k) {
      const uint32_t lo = big_lo[k], m = big_n[k];
      bitonic_sort(K + lo, m, pow2ceil(m));
      __syncthreads();
    }
    __syncthreads();
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

// medium and mid queue (half each): whole bins (<= nbins each) plus sub-segments of huge bins, of which every
// one but the last of a bin pairs with a neighbour to more than kSortSmemKeys / 2 entries; small and huge
// queues: one slot per bin each
size_t seg_queue_bytes(int64_t cap, int64_t nbins) {
  return (size_t)(8 * (cap / kSortSmemKeys) + 6 * nbins + 32) * sizeof(Segment);
}

// out_excl must hold n+1 elements (out[n] = total).  tmp: scan_tmp_elems(n) u32.
int launch_scan_u32(const uint32_t* in, uint32_t* out_excl, int64_t n, uint32_t* tmp, cudaStream_t s) {
  if (n <= 0) {
    GSW_CUDA(cudaMemsetAsync(out_excl, 0, sizeof(uint32_t), s));
    return 0;
  }
  static const bool multi = getenv("GSW_SCAN_MULTI") != nullptr;  // tuning aid: always the 3-kernel form
  if (n <= kScanSingleMax && !multi) {
    scan_single_kernel<<<1, kScanBlock, 0, s>>>(in, out_excl, n);
    count_launches(1);
    GSW_CUDA(cudaGetLastError());
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

// Sorts the bins [bin0, bin0 + nb) of the last scan.  `set` selects one of two independent sets of queue storage,
// queue counters and work tickets (status words), so that the static lists and the dynamic lists can be sorted by two
// calls on two streams at the same time; set 0 with the full bin range is the single-call form.  `side`: a stream
// this call may fork onto for the refinement and the medium class (s itself: everything in order on s).
int launch_sort_bins(const gsw_ctx* ctx, int64_t bin0, int64_t nb, int64_t cap, int set, cudaStream_t s,
                     cudaStream_t side, int32_t bins_per_list) {
  if (nb <= 0) return 0;
  auto& w = ctx->ws;
  constexpr size_t smem_m = (size_t)kSortSmemKeys * sizeof(unsigned long long);  // 32 KB
  constexpr size_t smem_s = (size_t)kSmallKeys * sizeof(unsigned long long);     // 8 KB
  // queue storage of a set (half of seg_queue): [small: nb][huge: nb][medium: half of the rest][mid: the other
  // half]; a queued mid/medium segment that is not a whole bin pairs with a neighbour to more than kMidKeys entries
  Queues Q;
  const size_t half = (w.seg_queue.bytes / sizeof(Segment)) / 2;
  Segment* base = w.seg_queue.as<Segment>() + (set ? half : 0);
  if (half < (size_t)(2 * nb + 16)) {
    set_error("sort queues too small for %lld bins", (long long)nb);
    return GSW_E_NOMEM;
  }
  const uint32_t total = (uint32_t)half;
  Q.q[0] = base;
  Q.cap[0] = (uint32_t)nb;
  Q.q[2] = base + nb;
  Q.cap[2] = (uint32_t)nb;
  Q.cap[1] = Q.cap[3] = (total - 2 * (uint32_t)nb) / 2;
  Q.q[1] = base + 2 * nb;
  Q.q[3] = Q.q[1] + Q.cap[1];
  // status words (zeroed every launch): set 0 -> counters 12..15, tickets 16..18; set 1 -> 24..27, 28..30
  Q.count = w.status.as<uint32_t>() + (set ? 24 : 12);
  uint32_t* ticket = w.status.as<uint32_t>() + (set ? 28 : 16);
  // (the kept prefix of the static lists starts at bin 0: only a call whose range starts there may skip it)
  classify_bins_kernel<<<(unsigned)((nb + 255) / 256), 256, 0, s>>>(
      w.bin_off.as<uint32_t>() + bin0, nb, cap, Q, w.status.as<long long>(),
      (ctx->cache_on && bin0 == 0) ? w.cache_m.as<uint32_t>() : nullptr, bins_per_list);
  // the depth partition of the huge bins (one CTA per bin, half the SMs idle) runs on the side stream next to
  // the small-class sort, which does not depend on it; the mid / medium queues it appends to are sorted after
  cudaStream_t ss = side ? side : s;
  if (ss != s) {
    GSW_CUDA(cudaEventRecord(ctx->fork_evt, s));
    GSW_CUDA(cudaStreamWaitEvent(ss, ctx->fork_evt, 0));
  }
  refine_big_bins_kernel<<<148 * 2, 512, 0, ss>>>(Q, w.keys.as<unsigned long long>(), w.splats.as<Splat>(),
                                                  w.status.as<long long>());
  if (ss != s) GSW_CUDA(cudaEventRecord(ctx->join_evt, ss));  // the medium / mid queues are complete
  // side stream: the medium class right behind the refinement; main stream: the small class, then (once the
  // refinement is done) the mid class.  The classes touch disjoint key ranges and each of them alone ends in a
  // tail of half-empty SMs, so they run next to each other.
  sort_segments_kernel<kSortSmemKeys, 256><<<148 * 4, 256, smem_m, ss>>>(Q.q[1], Q.count + 1, Q.cap[1],
                                                                         w.keys.as<unsigned long long>(), ticket);
  sort_segments_kernel<kSmallKeys, 64><<<148 * 16, 64, smem_s, s>>>(Q.q[0], Q.count, Q.cap[0],
                                                                      w.keys.as<unsigned long long>(), ticket + 1);
  if (ss != s) GSW_CUDA(cudaStreamWaitEvent(s, ctx->join_evt, 0));
  sort_segments_kernel<kMidKeys, 128><<<148 * 8, 128, 2 * smem_s, s>>>(Q.q[3], Q.count + 3, Q.cap[3],
                                                                      w.keys.as<unsigned long long>(), ticket + 2);
  if (ss != s) {
    GSW_CUDA(cudaEventRecord(ctx->fork_evt, ss));
    GSW_CUDA(cudaStreamWaitEvent(s, ctx->fork_evt, 0));
  }
  count_launches(5);
  GSW_CUDA(cudaGetLastError());
  return 0;
}

int launch_gather(const gsw_ctx* ctx, int64_t nbins, int32_t V, int64_t Ns, int64_t Nm, const FrameGeom& g, int64_t cap,
                  bool with_labels, cudaStream_t s) {
  if (nbins <= 0) return 0;
  auto& w = ctx->ws;
  const int32_t bpl = g.T * g.NB;
  (void)Nm;
  // enough CTAs per view to fill the machine when a single view is shared (the usual case: one world-fixed camera)
  const unsigned per_view = V >= 16 ? 74u : (unsigned)(148 * 8 / V);
  gather_splats_kernel<<<dim3(per_view, (unsigned)V), 256, 0, s>>>(
      w.bin_off.as<uint32_t>(), nbins, V, Ns, bpl, ctx->last.pos_of, w.keys.as<unsigned long long>(),
      w.rec.as<float4>(), w.splats.as<Splat>(), cap, ctx->labels, with_labels ? w.lab_stream.as<int32_t>() : nullptr,
      w.view_cnt.as<uint32_t>(), ctx->cache_on ? w.cache_m.as<uint32_t>() : nullptr);
  count_launches(1);
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
