// hs_block.cuh — CTA-level building blocks of the row mover: stable digit ranking with
// warp-private counters (no atomics), block scans, min/max reductions.
#pragma once
#include "hs_common.cuh"

// Threads and rows of a tile. -DPT_THREADS=256 builds the 2048-row-tile variant (4 CTAs per SM) for experiments.
#ifndef PT_THREADS
#define PT_THREADS 512
#endif
#define PT_WARPS (PT_THREADS / 32)
#ifndef PT_ITEMS
#define PT_ITEMS 8
#endif
#define PT_TILE (PT_THREADS * PT_ITEMS)  // 4096 rows per tile
#define PT_SCRATCH 32                    // words of scan scratch (>= PT_WARPS + 1)
#ifndef PT_MIN_CTAS
#define PT_MIN_CTAS (1024 / PT_THREADS)  // resident CTAs per SM the kernels are compiled for (64 registers per thread)
#endif

__device__ __forceinline__ uint32_t warp_incl_scan_u32(uint32_t v) {
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    uint32_t t = __shfl_up_sync(0xffffffffu, v, d);
    if (lane_id() >= (uint32_t)d) v += t;
  }
  return v;
}

// exclusive scan of one value per thread over PT_THREADS threads; `scratch` >= PT_WARPS+1 words
// (a one-barrier variant — every warp scanning the warp totals for itself — measured no faster: the leaf sort
// went from 1.68 to 1.72 ms, the scatter passes gained 0.01 ms)
__device__ __forceinline__ uint32_t block_excl_scan_u32(uint32_t v, uint32_t* scratch, uint32_t* total) {
  uint32_t inc = warp_incl_scan_u32(v);
  int w = threadIdx.x >> 5;
  __syncthreads();
  if (lane_id() == 31) scratch[w] = inc;
  __syncthreads();
  if (w == 0) {
    uint32_t s = lane_id() < PT_WARPS ? scratch[lane_id()] : 0;
    uint32_t si = warp_incl_scan_u32(s);
    if (lane_id() < PT_WARPS) scratch[lane_id()] = si - s;
    if (lane_id() == PT_WARPS - 1) scratch[PT_WARPS] = si;
  }
  __syncthreads();
  if (total) *total = scratch[PT_WARPS];
  return inc - v + scratch[w];
}

// the same scan for a CTA of any size up to 1024 threads whose first `nthreads` threads carry a value
// (k_msd_plan: one thread per digit); scratch: 32 words
__device__ __forceinline__ uint32_t block256_excl_scan_u32(uint32_t v, uint32_t* scratch, int nthreads = 256) {
  uint32_t inc = warp_incl_scan_u32(v);
  int w = threadIdx.x >> 5;
  const int nw = (nthreads + 31) >> 5;
  __syncthreads();
  if (lane_id() == 31 && w < 32) scratch[w] = inc;
  __syncthreads();
  if (w == 0) {
    uint32_t s = (int)lane_id() < nw ? scratch[lane_id()] : 0;
    uint32_t si = warp_incl_scan_u32(s);
    if ((int)lane_id() < nw) scratch[lane_id()] = si - s;
  }
  __syncthreads();
  return inc - v + (w < nw ? scratch[w] : 0);
}

// Stable ranking of a tile by digit. Item k of lane l in warp w is tile element
// w*chunk + k*32 + l (warp-striped), so (warp, k, lane) order == element order.
//   digit[k], valid bit k of vmask : inputs
//   cnt  : PT_WARPS * radix_pad words   tot, loc : radix_pad words   scratch : PT_WARPS+1 words
// On return pos[k] = position of the item in the digit-sorted tile, tot[d] = items with digit d,
// loc[d] = first position of digit d.
template <int ITEMS>
__device__ __forceinline__ void tile_rank(const uint32_t (&digit)[ITEMS], uint32_t vmask, int items, int bits,
                                          int radix_pad, uint32_t* cnt, uint32_t* tot, uint32_t* loc,
                                          uint32_t* scratch, uint32_t (&pos)[ITEMS]) {
  const int w = threadIdx.x >> 5;
  const uint32_t lt = (1u << lane_id()) - 1u;
  for (int i = threadIdx.x; i < PT_WARPS * radix_pad; i += PT_THREADS) cnt[i] = 0;
  __syncthreads();
  uint32_t* mycnt = cnt + w * radix_pad;
#pragma unroll
  for (int k = 0; k < ITEMS; k++) {
    if (k < items) {
      bool v = (vmask >> k) & 1u;
      uint32_t d = digit[k];
      uint32_t m = __ballot_sync(0xffffffffu, v);
      for (int b = 0; b < bits; b++) {
        bool bit = (d >> b) & 1u;
        uint32_t bal = __ballot_sync(0xffffffffu, bit);
        m &= bit ? bal : ~bal;
      }
      uint32_t r = __popc(m & lt);
      uint32_t pre = v ? mycnt[d] : 0;
      __syncwarp();
      if (v && r == 0) mycnt[d] = pre + __popc(m);
      __syncwarp();
      pos[k] = pre + r;
    }
  }
  __syncthreads();
  for (int d = threadIdx.x; d < radix_pad; d += PT_THREADS) {
    uint32_t run = 0;
#pragma unroll
    for (int ww = 0; ww < PT_WARPS; ww++) {
      uint32_t c = cnt[ww * radix_pad + d];
      cnt[ww * radix_pad + d] = run;
      run += c;
    }
    tot[d] = run;
  }
  __syncthreads();
  // exclusive scan of tot over digits: each thread owns `per` consecutive digits
  const int per = (radix_pad + PT_THREADS - 1) / PT_THREADS;
  const int d0 = threadIdx.x * per, d1 = d0 + per < radix_pad ? d0 + per : radix_pad;
  uint32_t s = 0;
  for (int j = d0; j < d1; j++) s += tot[j];
  uint32_t ex = block_excl_scan_u32(s, scratch, nullptr);
  for (int j = d0; j < d1; j++) {
    loc[j] = ex;
    ex += tot[j];
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < ITEMS; k++) {
    if (k < items && ((vmask >> k) & 1u)) {
      uint32_t d = digit[k];
      pos[k] += loc[d] + mycnt[d];
    }
  }
}

// Unstable ranking of a tile by digit with shared-memory atomics (B200: ~11 ATOMS lanes/clk/SM on random
// addresses, an order of magnitude cheaper than the ballot match above). Use wherever the order of rows
// with equal digits does not matter: the bucket scatter, MSD levels, the first pass of an LSD sort.
//   cnt : radix_pad words; on return cnt[d] = first position of digit d (exclusive scan, in place),
//   tot : radix_pad words = items per digit (may be nullptr when the caller does not need it).
// On return pos[k] = position of item k in the digit-grouped tile, rnk[k] = its rank inside its digit.
template <int ITEMS>
__device__ __forceinline__ void tile_rank_atomic(const uint32_t (&digit)[ITEMS], uint32_t vmask, int items,
                                                 int radix_pad, uint32_t* cnt, uint32_t* tot, uint32_t* scratch,
                                                 uint32_t (&pos)[ITEMS]) {
  for (int i = threadIdx.x; i < radix_pad; i += PT_THREADS) cnt[i] = 0;
  __syncthreads();
#pragma unroll
  for (int k = 0; k < ITEMS; k++)
    if (k < items && ((vmask >> k) & 1u)) pos[k] = atomicAdd(&cnt[digit[k]], 1u);
  __syncthreads();
  const int per = (radix_pad + PT_THREADS - 1) / PT_THREADS;
  const int d0 = threadIdx.x * per, d1 = d0 + per < radix_pad ? d0 + per : radix_pad;
  uint32_t s = 0;
  for (int j = d0; j < d1; j++) s += cnt[j];
  uint32_t ex = block_excl_scan_u32(s, scratch, nullptr);
  for (int j = d0; j < d1; j++) {
    uint32_t c = cnt[j];
    if (tot) tot[j] = c;
    cnt[j] = ex;
    ex += c;
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < ITEMS; k++)
    if (k < items && ((vmask >> k) & 1u)) pos[k] += cnt[digit[k]];
}

// tile_rank_atomic that also reserves the global output runs: gbase[d] = atomicAdd(&cursor[d], rows of d).
// The global atomics are issued as soon as the counts are final and their results are only consumed after
// the block scan, so the ~1 us round trip to L2 hides behind the scan (one digit per thread; wider radixes
// issue them after the scan).
template <int ITEMS>
__device__ __forceinline__ void tile_rank_atomic_reserve(const uint32_t (&digit)[ITEMS], uint32_t vmask, int items,
                                                         int radix_pad, uint32_t* cnt, uint32_t* tot, uint32_t* gbase,
                                                         uint32_t* cursor, uint32_t* scratch, uint32_t (&pos)[ITEMS]) {
  for (int i = threadIdx.x; i < radix_pad; i += PT_THREADS) cnt[i] = 0;
  __syncthreads();
#pragma unroll
  for (int k = 0; k < ITEMS; k++)
    if (k < items && ((vmask >> k) & 1u)) pos[k] = atomicAdd(&cnt[digit[k]], 1u);
  __syncthreads();
  const bool early = radix_pad <= PT_THREADS;
  uint32_t my_t = 0, my_g = 0;
  if (early && (int)threadIdx.x < radix_pad) {
    my_t = cnt[threadIdx.x];
    if (my_t) my_g = atomicAdd(&cursor[threadIdx.x], my_t);
  }
  const int per = (radix_pad + PT_THREADS - 1) / PT_THREADS;
  const int d0 = threadIdx.x * per, d1 = d0 + per < radix_pad ? d0 + per : radix_pad;
  uint32_t s = 0;
  for (int j = d0; j < d1; j++) s += cnt[j];
  uint32_t ex = block_excl_scan_u32(s, scratch, nullptr);
  for (int j = d0; j < d1; j++) {
    uint32_t c = cnt[j];
    tot[j] = c;
    cnt[j] = ex;
    ex += c;
  }
  if (early && (int)threadIdx.x < radix_pad) gbase[threadIdx.x] = my_g;
  __syncthreads();
  if (!early) {
    for (int d = threadIdx.x; d < radix_pad; d += PT_THREADS) {
      uint32_t t = tot[d];
      gbase[d] = t ? atomicAdd(&cursor[d], t) : 0;
    }
    __syncthreads();
  }
#pragma unroll
  for (int k = 0; k < ITEMS; k++)
    if (k < items && ((vmask >> k) & 1u)) pos[k] += cnt[digit[k]];
}

// Variants that take the digit as a bit field of a 32-bit word (no digit array in registers).
template <int ITEMS>
__device__ __forceinline__ void tile_rank_atomic_w(const uint32_t (&word)[ITEMS], int shift, uint32_t mask,
                                                   uint32_t vmask, int items, int radix_pad, uint32_t* cnt,
                                                   uint32_t* scratch, uint32_t (&pos)[ITEMS]) {
  for (int i = threadIdx.x; i < radix_pad; i += PT_THREADS) cnt[i] = 0;
  __syncthreads();
#pragma unroll
  for (int k = 0; k < ITEMS; k++)
    if (k < items && ((vmask >> k) & 1u)) pos[k] = atomicAdd(&cnt[(word[k] >> shift) & mask], 1u);
  __syncthreads();
  const int per = (radix_pad + PT_THREADS - 1) / PT_THREADS;
  const int d0 = threadIdx.x * per, d1 = d0 + per < radix_pad ? d0 + per : radix_pad;
  uint32_t s = 0;
  for (int j = d0; j < d1; j++) s += cnt[j];
  uint32_t ex = block_excl_scan_u32(s, scratch, nullptr);
  for (int j = d0; j < d1; j++) {
    uint32_t c = cnt[j];
    cnt[j] = ex;
    ex += c;
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < ITEMS; k++)
    if (k < items && ((vmask >> k) & 1u)) pos[k] += cnt[(word[k] >> shift) & mask];
}

template <int ITEMS>
__device__ __forceinline__ void tile_rank_w(const uint32_t (&word)[ITEMS], int shift, uint32_t vmask, int items,
                                            int bits, int radix_pad, uint32_t* cnt, uint32_t* tot, uint32_t* loc,
                                            uint32_t* scratch, uint32_t (&pos)[ITEMS]) {
  const int w = threadIdx.x >> 5;
  const uint32_t lt = (1u << lane_id()) - 1u;
  const uint32_t mask = (1u << bits) - 1u;
  for (int i = threadIdx.x; i < PT_WARPS * radix_pad; i += PT_THREADS) cnt[i] = 0;
  __syncthreads();
  uint32_t* mycnt = cnt + w * radix_pad;
#pragma unroll
  for (int k = 0; k < ITEMS; k++) {
    if (k < items) {
      bool v = (vmask >> k) & 1u;
      uint32_t d = (word[k] >> shift) & mask;
      uint32_t m = __ballot_sync(0xffffffffu, v);
#pragma unroll
      for (int b = 0; b < 8; b++) {   // bits <= 8; unrolled so that every bit test has a constant mask
        if (b < bits) {
          const bool bit = (d & (1u << b)) != 0;
          const uint32_t bal = __ballot_sync(0xffffffffu, bit);
          m &= bal ^ ((uint32_t)bit - 1u);   // bit ? bal : ~bal
        }
      }
      uint32_t r = __popc(m & lt);
      uint32_t pre = v ? mycnt[d] : 0;
      __syncwarp();
      if (v && r == 0) mycnt[d] = pre + __popc(m);
      __syncwarp();
      pos[k] = pre + r;
    }
  }
  __syncthreads();
  for (int d = threadIdx.x; d < radix_pad; d += PT_THREADS) {
    uint32_t run = 0;
#pragma unroll
    for (int ww = 0; ww < PT_WARPS; ww++) {
      uint32_t c = cnt[ww * radix_pad + d];
      cnt[ww * radix_pad + d] = run;
      run += c;
    }
    tot[d] = run;
  }
  __syncthreads();
  const int per = (radix_pad + PT_THREADS - 1) / PT_THREADS;
  const int d0 = threadIdx.x * per, d1 = d0 + per < radix_pad ? d0 + per : radix_pad;
  uint32_t s = 0;
  for (int j = d0; j < d1; j++) s += tot[j];
  uint32_t ex = block_excl_scan_u32(s, scratch, nullptr);
  for (int j = d0; j < d1; j++) {
    loc[j] = ex;
    ex += tot[j];
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < ITEMS; k++) {
    if (k < items && ((vmask >> k) & 1u)) {
      uint32_t d = (word[k] >> shift) & mask;
      pos[k] += loc[d] + mycnt[d];
    }
  }
}

__device__ __forceinline__ uint64_t warp_min_u64(uint64_t v) {
#pragma unroll
  for (int d = 16; d; d >>= 1) {
    uint64_t o = __shfl_xor_sync(0xffffffffu, v, d);
    v = o < v ? o : v;
  }
  return v;
}
__device__ __forceinline__ uint64_t warp_max_u64(uint64_t v) {
#pragma unroll
  for (int d = 16; d; d >>= 1) {
    uint64_t o = __shfl_xor_sync(0xffffffffu, v, d);
    v = o > v ? o : v;
  }
  return v;
}
