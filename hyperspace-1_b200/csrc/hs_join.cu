// hs_join.cu — K5 + K6: bucket-local inner sort-merge join of two indexes, no exchange.
// Executes the plan JoinIndexRule.apply produces (HS/index/rules/JoinIndexRule.scala:57-90) over the
// bucketed relations of RuleUtils.transformPlanToUseIndexOnlyScan (HS/index/rules/RuleUtils.scala:264-292):
// bucket i of the left index meets bucket i of the right index; equal keys produce the cross product;
// NULL keys never match.
// Both sides are sorted per bucket, so neighbouring left rows probe neighbouring right rows: each left
// row binary-searches its bucket of the right run (the probes of a warp share cache lines), a scan turns
// match counts into output offsets, and the pairs are expanded and gathered.
#include <algorithm>
#include <stdlib.h>

#include "hs_keys.cuh"
#include "hs_kernels.h"
#include "hs_tma.cuh"

struct JoinSide {
  KeySpec ks;
  const uint32_t* off;  // device, nb + 1
};

__device__ __forceinline__ bool key_has_null(const KeySpec& ks, int64_t i) {
#pragma unroll
  for (int k = 0; k < HS_MAX_KEYS; k++)
    if (k < ks.nkeys && ks.k[k].has_nulls && !kvalid(ks.k[k].validity, i)) return true;
  return false;
}
// keys of left row i vs right row j (no NULLs on either side)
__device__ __forceinline__ int cmp_keys(const KeySpec& a, int64_t i, const KeySpec& b, int64_t j) {
#pragma unroll
  for (int k = 0; k < HS_MAX_KEYS; k++) {
    if (k < a.nkeys) {
      int c = cmp_cells(a.k[k], i, b.k[k], j);
      if (c) return c;
    }
  }
  return 0;
}
// right rows with NULL keys sort first inside a bucket: skip them by treating NULL as smaller
__device__ __forceinline__ int cmp_left_right(const KeySpec& a, int64_t i, const KeySpec& b, int64_t j) {
#pragma unroll
  for (int k = 0; k < HS_MAX_KEYS; k++) {
    if (k < a.nkeys) {
      if (b.k[k].has_nulls && !kvalid(b.k[k].validity, j)) return 1;  // left (non-null) > right NULL
      int c = cmp_cells(a.k[k], i, b.k[k], j);
      if (c) return c;
    }
  }
  return 0;
}

template <bool FAST_I64>
__global__ void __launch_bounds__(256) k_join_probe(JoinSide L, JoinSide R, int64_t nl, int32_t nb,
                                                    uint32_t* __restrict__ lb_out, uint32_t* __restrict__ cnt_out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nl) return;
  // bucket of left row i: largest b with L.off[b] <= i
  int lo = 0, hi = nb;
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (L.off[mid] <= (uint32_t)i) lo = mid; else hi = mid;
  }
  uint32_t r0 = R.off[lo], r1 = R.off[lo + 1];
  uint32_t lb, ub;
  if (FAST_I64) {
    const int64_t* lk = (const int64_t*)L.ks.k[0].data;
    const int64_t* rk = (const int64_t*)R.ks.k[0].data;
    int64_t key = lk[i];
    uint32_t a = r0, b = r1;
    while (a < b) { uint32_t m = (a + b) >> 1; if (rk[m] < key) a = m + 1; else b = m; }
    lb = a;
    b = r1;
    while (a < b) { uint32_t m = (a + b) >> 1; if (rk[m] <= key) a = m + 1; else b = m; }
    ub = a;
  } else {
    if (key_has_null(L.ks, i)) { lb_out[i] = r0; cnt_out[i] = 0; return; }
    uint32_t a = r0, b = r1;
    while (a < b) { uint32_t m = (a + b) >> 1; if (cmp_left_right(L.ks, i, R.ks, m) > 0) a = m + 1; else b = m; }
    lb = a;
    b = r1;
    while (a < b) { uint32_t m = (a + b) >> 1; if (cmp_left_right(L.ks, i, R.ks, m) >= 0) a = m + 1; else b = m; }
    ub = a;
  }
  lb_out[i] = lb;
  cnt_out[i] = ub - lb;
}

__global__ void __launch_bounds__(256) k_join_expand(const uint32_t* __restrict__ lb, const uint32_t* __restrict__ cnt,
                                                     const uint32_t* __restrict__ off, int64_t nl,
                                                     uint32_t* __restrict__ li, uint32_t* __restrict__ ri) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nl) return;
  uint32_t c = cnt[i], o = off[i], b = lb[i];
  for (uint32_t t = 0; t < c; t++) { li[o + t] = (uint32_t)i; ri[o + t] = b + t; }
}

__global__ void __launch_bounds__(256) k_join_bucket_offsets(const uint32_t* __restrict__ loff, const uint32_t* __restrict__ off,
                                                             const uint64_t* total, int64_t nl, int32_t nb,
                                                             uint64_t* __restrict__ out) {
  int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b > nb) return;
  uint32_t r = b < nb ? loff[b] : (uint32_t)nl;
  out[b] = r < nl ? off[r] : *total;
}

// join types other than inner: how many rows a left row contributes, given its number of matches
__global__ void __launch_bounds__(256) k_join_type_counts(const uint32_t* __restrict__ cnt, int64_t nl, int type,
                                                          uint32_t* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nl) return;
  const uint32_t c = cnt[i];
  out[i] = type == HS_JOIN_LEFT_OUTER ? (c ? c : 1u) : type == HS_JOIN_LEFT_SEMI ? (c ? 1u : 0u) : (c ? 0u : 1u);
}
// expansion for LEFT_OUTER (a row without a match pairs with HS_MISS), LEFT_SEMI and LEFT_ANTI (left rows only)
__global__ void __launch_bounds__(256) k_join_expand_typed(const uint32_t* __restrict__ lb, const uint32_t* __restrict__ cnt,
                                                           const uint32_t* __restrict__ tcnt, const uint32_t* __restrict__ off,
                                                           int64_t nl, uint32_t* __restrict__ li, uint32_t* __restrict__ ri) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nl) return;
  const uint32_t c = cnt[i], t = tcnt[i], o = off[i], b = lb[i];
  for (uint32_t k = 0; k < t; k++) {
    li[o + k] = (uint32_t)i;
    if (ri) ri[o + k] = c ? b + k : HS_MISS;
  }
}
// FULL OUTER = the LEFT OUTER pairs plus the right rows without a match; bucket b of the result holds its
// left-outer pairs first (o1 = their bucket offsets), then its unmatched right rows (o2)
__global__ void __launch_bounds__(256) k_full_outer_merge(const uint32_t* __restrict__ li1, const uint32_t* __restrict__ ri1,
                                                          int64_t m1, const uint64_t* __restrict__ o1,
                                                          const uint32_t* __restrict__ rj, int64_t m2,
                                                          const uint64_t* __restrict__ o2, int32_t nb,
                                                          uint32_t* __restrict__ li, uint32_t* __restrict__ ri) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= m1 + m2) return;
  const bool second = t >= m1;
  const uint64_t q = second ? (uint64_t)(t - m1) : (uint64_t)t;
  const uint64_t* o = second ? o2 : o1;
  int lo = 0, hi = nb;            // largest b with o[b] <= q
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (o[mid] <= q) lo = mid; else hi = mid;
  }
  if (!second) {
    const uint64_t d = q + o2[lo];
    li[d] = li1[q]; ri[d] = ri1[q];
  } else {
    const uint64_t d = o1[lo + 1] + q;
    li[d] = HS_MISS; ri[d] = rj[q];
  }
}

// ------------------------------------------------------------------ fast path: merge-path tiles
// One signed-integer key without NULLs on both sides and fixed-width, NULL-free output columns (the
// lineitem x orders shape). Left rows are cut into tiles of JT rows that never cross a bucket; because both
// runs are sorted, a tile's matches lie in one narrow window of the right run: the window is located with
// two warp-wide searches, its keys are staged in shared memory, every left row finds [lower, upper) there.
// Pass 1 counts matches per tile, a scan turns the counts into output offsets, pass 2 repeats the search,
// expands the pairs in shared memory and writes every output column directly (no index arrays in HBM).
#define JT 2048          // left rows per tile
#ifndef JR
#define JR 2048          // right rows staged per tile (falls back to global probes when the window is larger)
#endif
#define JTHREADS 512
#ifndef JCTAS
#define JCTAS 3          // resident CTAs per SM the tile kernel is compiled for
#endif
#define JPER (JT / JTHREADS)   // consecutive left rows owned by one thread
#define JOMAX 4096       // outputs of one tile that are expanded through shared memory
#define JBUF_BYTES (JR * 8 + 128)

struct JoinFast {
  const void* lkey; const void* rkey;     // key columns
  const uint32_t* roff;                   // right bucket offsets (device)
  const uint32_t* tile_bucket;            // per tile
  const uint32_t* tile_start;             // first left row of the tile
  const uint32_t* tile_len;
  uint32_t ntiles;
  uint32_t* tile_win;                     // [2 * ntiles]: window [lo, hi) of the right run per tile
  uint32_t* tile_cnt;                     // pass 1 out / pass 2 in: matches per tile, then exclusive offsets
  // single-pass mode: chained scan over the tiles (decoupled look-back) instead of a separate count pass
  unsigned long long* tile_state;         // [ntiles]: (flag << 62) | value; flag 1 = tile total, 2 = inclusive prefix
  uint32_t* ticket;                       // dynamic tile ids: a tile only ever waits for tiles that already run
  uint32_t capacity;                      // rows the output buffers can hold
  uint32_t* overflow;                     // set when the output does not fit: the caller falls back to two passes
  int32_t n_lout, n_rout;
  const void* lsrc[16]; const void* rsrc[16];
  void* ldst[16]; void* rdst[16];
  int32_t les[16], res[16];
};

template <typename K>
__device__ __forceinline__ uint32_t warp_lower_bound(const K* a, uint32_t lo, uint32_t hi, K key, bool upper) {
  // first index in [lo, hi) with a[i] >= key (upper: > key); 32-ary search by one warp
  while (hi - lo > 32) {
    uint32_t step = (hi - lo + 31) / 32;
    uint32_t idx = lo + lane_id() * step;
    bool before = idx < hi && (upper ? a[idx] <= key : a[idx] < key);
    uint32_t m = __ballot_sync(0xffffffffu, before);
    uint32_t k = __popc(m);                      // probes 0..k-1 are before the answer
    uint32_t nlo = k ? lo + (k - 1) * step + 1 : lo;
    uint32_t nhi = k < 32 ? lo + k * step : hi;
    lo = nlo;
    hi = nhi < hi ? nhi : hi;
    if (lo > hi) lo = hi;
  }
  uint32_t idx = lo + lane_id();
  bool before = idx < hi && (upper ? a[idx] <= key : a[idx] < key);
  return lo + __popc(__ballot_sync(0xffffffffu, before));
}

// one warp per tile: the window of the right run that can hold matches for the tile's left rows
template <typename K>
__global__ void __launch_bounds__(256) k_join_windows(const __grid_constant__ JoinFast p) {
  const uint32_t t = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (t >= p.ntiles) return;
  const uint32_t b = p.tile_bucket[t], l0 = p.tile_start[t], ln = p.tile_len[t];
  const K* lk = (const K*)p.lkey + l0;
  const K* rk = (const K*)p.rkey;
  const uint32_t r0 = p.roff[b], r1 = p.roff[b + 1];
  uint32_t lo = warp_lower_bound<K>(rk, r0, r1, lk[0], false);
  uint32_t hi = warp_lower_bound<K>(rk, lo, r1, lk[ln - 1], true);
  if (lane_id() == 0) { p.tile_win[2 * t] = lo; p.tile_win[2 * t + 1] = hi; }
}

// first position >= pos in the staged window whose key is >= key (short forward scan, binary search for
// long gaps)
template <typename K>
__device__ __forceinline__ uint32_t window_advance(const K* sR, uint32_t pos, uint32_t wn, K key) {
  if (pos + 8 < wn && sR[pos + 8] < key) {
    uint32_t a = pos + 9, e = wn;
    while (a < e) { uint32_t m = (a + e) >> 1; if (sR[m] < key) a = m + 1; else e = m; }
    return a;
  }
  while (pos < wn && sR[pos] < key) pos++;
  return pos;
}

// One tile of left rows against its window of the right run.
// MODE 0: count matches per tile.  MODE 1: write, offsets from the scanned counts.
// MODE 2: single pass — the tile's output offset comes from a chained scan over the tiles.
// The tile's left keys and the window's right keys arrive in shared memory by TMA bulk copies; every thread
// merges its JPER consecutive left keys against the staged window. For the output, the (left row, right row) of
// every output position is laid out in shared memory once; each output column is then produced from a staged
// slice of its source column (bulk copies, double buffered) with coalesced stores — no per-thread strided
// global loads, no index arrays in HBM. Tiles whose window or output exceeds the staging areas take the
// global-memory paths below (same results).
template <typename K, int MODE>
__global__ void __launch_bounds__(JTHREADS, JCTAS) k_join_tiles(const __grid_constant__ JoinFast p) {
  extern __shared__ __align__(128) unsigned char jsm[];
  unsigned char* buf0 = jsm;
  unsigned char* buf1 = jsm + JBUF_BYTES;
  unsigned char* buf2 = jsm + 2 * JBUF_BYTES;
  uint16_t* omap = (uint16_t*)(jsm + 3 * JBUF_BYTES);   // JOMAX: tile-local left row of every output
  uint16_t* orel = omap + JOMAX;                        // JOMAX: window-local right row of every output
  uint32_t* slb = (uint32_t*)omap;                      // expanding path only (aliases the two maps): JT
  uint32_t* scn = slb + JT;                             // JT + 1
  uint64_t* bar = (uint64_t*)(jsm + 3 * JBUF_BYTES + 4 * JOMAX + 16);   // 3 mbarriers
  __shared__ uint32_t s_red[JTHREADS / 32 + 1];
  __shared__ uint32_t s_tile, s_obase;
  if (MODE == 2) {
    if (threadIdx.x == 0) s_tile = atomicAdd(p.ticket, 1u);
  }
  if (threadIdx.x == 0) {
    mbar_init(&bar[0], 1);
    mbar_init(&bar[1], 1);
    mbar_init(&bar[2], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();
  const uint32_t t = MODE == 2 ? s_tile : blockIdx.x;
  const uint32_t l0 = p.tile_start[t], ln = p.tile_len[t];
  const K* rk = (const K*)p.rkey;
  const uint32_t wlo = p.tile_win[2 * t], whi = p.tile_win[2 * t + 1], wn = whi - wlo;
  const bool staged = wn <= JR;
  uint32_t ph0 = 0, ph1 = 0, ph2 = 0;   // phase parity of the three mbarriers
  if (threadIdx.x == 0) {
    issue_slice(p.lkey, (int)sizeof(K), l0, ln, buf0, &bar[0]);
    if (staged && wn) issue_slice(p.rkey, (int)sizeof(K), wlo, wn, buf1, &bar[1]);
  }
  const K* sLK = (const K*)(buf0 + slice_lead(p.lkey, (int)sizeof(K), l0));
  const K* sR = (const K*)(buf1 + slice_lead(p.rkey, (int)sizeof(K), wlo));
  mbar_wait(&bar[0], ph0); ph0 ^= 1;
  if (staged && wn) { mbar_wait(&bar[1], ph1); ph1 ^= 1; }
  const uint32_t i0 = threadIdx.x * JPER;
  K key[JPER];
#pragma unroll
  for (int j = 0; j < JPER; j++) key[j] = i0 + j < ln ? sLK[i0 + j] : (K)0;
  uint32_t lb[JPER], cnt[JPER];   // first match relative to the window, number of matches
  uint32_t mysum = 0;
  if (staged) {
    uint32_t pos = 0;
    if (i0 < ln) {
      uint32_t a = 0, e = wn;
      while (a < e) { uint32_t m = (a + e) >> 1; if (sR[m] < key[0]) a = m + 1; else e = m; }
      pos = a;
    }
#pragma unroll
    for (int j = 0; j < JPER; j++) {
      lb[j] = 0; cnt[j] = 0;
      if (i0 + j < ln) {
        pos = window_advance<K>(sR, pos, wn, key[j]);
        uint32_t e = pos;
        while (e < wn && sR[e] == key[j]) e++;
        lb[j] = pos;
        cnt[j] = e - pos;
        mysum += cnt[j];
      }
    }
  } else {
#pragma unroll
    for (int j = 0; j < JPER; j++) {
      lb[j] = 0; cnt[j] = 0;
      if (i0 + j < ln) {
        uint32_t a = wlo, e = whi;
        while (a < e) { uint32_t m = (a + e) >> 1; if (rk[m] < key[j]) a = m + 1; else e = m; }
        uint32_t f = a;
        e = whi;
        while (a < e) { uint32_t m = (a + e) >> 1; if (rk[m] <= key[j]) a = m + 1; else e = m; }
        lb[j] = f - wlo;
        cnt[j] = a - f;
        mysum += cnt[j];
      }
    }
  }
  // block exclusive scan of the per-thread sums
  uint32_t inc = mysum;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    uint32_t o = __shfl_up_sync(0xffffffffu, inc, d);
    if (lane_id() >= (uint32_t)d) inc += o;
  }
  if (lane_id() == 31) s_red[threadIdx.x >> 5] = inc;
  __syncthreads();
  // ---- the output columns' source slices are requested now, so that the copies run beside the scan, the
  // look-back and the lay-out of the outputs: every thread follows the same column sequence, thread 0 starts
  // the copies (the staged right keys in buf1 are no longer read: every thread has passed the barrier above)
  const int ncol = p.n_lout + p.n_rout;
  auto col_src = [&](int c, const void*& src, int& es, bool& left) {
    left = c < p.n_lout;
    src = left ? p.lsrc[c] : p.rsrc[c - p.n_lout];
    es = left ? p.les[c] : p.res[c - p.n_lout];
  };
  auto start_load = [&](int c, unsigned char* b, uint64_t* br) {
    const void* src; int es; bool left;
    col_src(c, src, es, left);
    if (left) issue_slice(src, es, l0, ln, b, br);
    else issue_slice(src, es, wlo, wn, b, br);
  };
  int keycol = -1;
  for (int c = 0; c < p.n_lout; c++)
    if (p.lsrc[c] == p.lkey && p.les[c] == (int)sizeof(K)) { keycol = c; break; }
  // order of the staged columns: all but keycol
  int nxt = 0;            // next column to load
  auto next_col = [&]() { while (nxt == keycol) nxt++; return nxt < ncol ? nxt++ : -1; };
  int inb1 = -1, inb2 = -1;
  const bool early = MODE != 0 && staged;
  if (early) {
    inb1 = next_col();
    inb2 = next_col();
    if (threadIdx.x == 0) {
      if (inb1 >= 0) start_load(inb1, buf1, &bar[1]);
      if (inb2 >= 0) start_load(inb2, buf2, &bar[2]);
    }
  }
  // a CTA must not retire with a bulk copy in flight
  auto drain = [&]() {
    if (inb1 >= 0) mbar_wait(&bar[1], ph1);
    if (inb2 >= 0) mbar_wait(&bar[2], ph2);
  };
  if (threadIdx.x == 0) {
    uint32_t run = 0;
    for (int w = 0; w < JTHREADS / 32; w++) { uint32_t v = s_red[w]; s_red[w] = run; run += v; }
    s_red[JTHREADS / 32] = run;
  }
  __syncthreads();
  const uint32_t total = s_red[JTHREADS / 32];
  if (MODE == 0) {
    if (threadIdx.x == 0) p.tile_cnt[t] = total;
    return;
  }
  uint32_t ex = inc - mysum + s_red[threadIdx.x >> 5];
  if (MODE == 2) {
    // chained scan over the tiles (decoupled look-back): warp 0 publishes the tile's total, then looks at 32
    // predecessors at a time until it meets one that has published its inclusive prefix
    if (threadIdx.x < 32) {
      const unsigned long long FA = 1ull << 62, FP = 2ull << 62, VM = (1ull << 62) - 1;
      volatile unsigned long long* st = p.tile_state;
      unsigned long long excl = 0;
      if (t > 0) {
        if (lane_id() == 0) st[t] = FA | total;
        __threadfence();
        for (int64_t q = (int64_t)t - 1; q >= 0; q -= 32) {
          const int64_t idx = q - lane_id();
          unsigned long long v = FP;               // before the first tile: an inclusive prefix of 0
          if (idx >= 0) {
            do { v = st[idx]; } while ((v >> 62) == 0);   // the predecessor has started (ticket order): it will publish
          }
          const uint32_t has_prefix = __ballot_sync(0xffffffffu, (v >> 62) == 2);
          const uint32_t upto = has_prefix ? (uint32_t)__ffs(has_prefix) - 1u : 31u;   // nearest lane with a prefix
          unsigned long long part = lane_id() <= upto ? (v & VM) : 0ull;
#pragma unroll
          for (int d = 16; d; d >>= 1) part += __shfl_xor_sync(0xffffffffu, part, d);
          excl += part;
          if (has_prefix) break;
        }
      }
      if (lane_id() == 0) {
        __threadfence();
        st[t] = FP | (excl + total);
        s_obase = (uint32_t)excl;
        p.tile_cnt[t] = (uint32_t)excl;
        if (excl + total > p.capacity) *p.overflow = 1;
      }
    }
    __syncthreads();
    if ((unsigned long long)s_obase + total > p.capacity) { drain(); return; }   // does not fit: count only
  }
  const uint32_t obase = MODE == 2 ? s_obase : p.tile_cnt[t];   // exclusive offset of this tile in the output
  if (total == 0) { drain(); return; }
  if (staged && total <= JOMAX) {
    // ---- (left row, right row) of every output position of the tile
    {
      uint32_t o = ex;
#pragma unroll
      for (int j = 0; j < JPER; j++) {
        for (uint32_t q = 0; q < cnt[j]; q++) {
          omap[o + q] = (uint16_t)(i0 + j);
          orel[o + q] = (uint16_t)(lb[j] + q);
        }
        o += cnt[j];
      }
    }
    __syncthreads();   // also: every thread is done with the staged right keys (buf1)
    // ---- columns: a left output that is the key column itself is served from buf0; the others are staged in
    // buf1 / buf2 alternately, the copy of column k + 2 starting as soon as column k has been written
    auto write_col = [&](int c, const unsigned char* b) {
      const void* src; int es; bool left;
      col_src(c, src, es, left);
      const unsigned char* s = b + slice_lead(src, es, left ? l0 : wlo);
      void* dst = left ? p.ldst[c] : p.rdst[c - p.n_lout];
      const uint16_t* map = left ? omap : orel;
      if (es == 8) {
        for (uint32_t o = threadIdx.x; o < total; o += JTHREADS) ((uint64_t*)dst)[obase + o] = ((const uint64_t*)s)[map[o]];
      } else {
        for (uint32_t o = threadIdx.x; o < total; o += JTHREADS) ((uint32_t*)dst)[obase + o] = ((const uint32_t*)s)[map[o]];
      }
    };
    if (keycol >= 0) {
      write_col(keycol, buf0);
      __syncthreads();
    }
    while (inb1 >= 0 || inb2 >= 0) {
      if (inb1 >= 0) {
        mbar_wait(&bar[1], ph1); ph1 ^= 1;
        write_col(inb1, buf1);
        __syncthreads();
        inb1 = next_col();
        if (inb1 >= 0 && threadIdx.x == 0) start_load(inb1, buf1, &bar[1]);
      }
      if (inb2 >= 0) {
        mbar_wait(&bar[2], ph2); ph2 ^= 1;
        write_col(inb2, buf2);
        __syncthreads();
        inb2 = next_col();
        if (inb2 >= 0 && threadIdx.x == 0) start_load(inb2, buf2, &bar[2]);
      }
    }
    return;
  }
  // ---- large cross products or windows that were not staged: spread the outputs over the whole CTA and read
  // the sources from global memory
  drain();
  __syncthreads();
#pragma unroll
  for (int j = 0; j < JPER; j++) {
    if (i0 + j < ln) { slb[i0 + j] = wlo + lb[j]; scn[i0 + j] = ex; }
    ex += cnt[j];
  }
  if (threadIdx.x == 0) scn[ln] = total;
  __syncthreads();
  for (uint32_t chunk = 0; chunk < total; chunk += JTHREADS * 8) {
    uint32_t li[8], rj[8];
#pragma unroll
    for (int q = 0; q < 8; q++) {
      uint32_t o = chunk + q * JTHREADS + threadIdx.x;
      li[q] = 0; rj[q] = 0;
      if (o < total) {
        uint32_t a = 0, e = ln;   // last row i with scn[i] <= o
        while (e - a > 1) { uint32_t m = (a + e) >> 1; if (scn[m] <= o) a = m; else e = m; }
        li[q] = a;
        rj[q] = slb[a] + (o - scn[a]);
      }
    }
    for (int c = 0; c < p.n_lout; c++) {
#pragma unroll
      for (int q = 0; q < 8; q++) {
        uint32_t o = chunk + q * JTHREADS + threadIdx.x;
        if (o < total) {
          if (p.les[c] == 8) ((uint64_t*)p.ldst[c])[obase + o] = ((const uint64_t*)p.lsrc[c])[l0 + li[q]];
          else ((uint32_t*)p.ldst[c])[obase + o] = ((const uint32_t*)p.lsrc[c])[l0 + li[q]];
        }
      }
    }
    for (int c = 0; c < p.n_rout; c++) {
#pragma unroll
      for (int q = 0; q < 8; q++) {
        uint32_t o = chunk + q * JTHREADS + threadIdx.x;
        if (o < total) {
          if (p.res[c] == 8) ((uint64_t*)p.rdst[c])[obase + o] = ((const uint64_t*)p.rsrc[c])[rj[q]];
          else ((uint32_t*)p.rdst[c])[obase + o] = ((const uint32_t*)p.rsrc[c])[rj[q]];
        }
      }
    }
  }
}

static size_t join_smem_bytes() { return (size_t)3 * JBUF_BYTES + (size_t)4 * JOMAX + 16 + 3 * 8 + 16; }

static void check_view(const hs_index_view* v) {
  if (!v || !v->cols || v->ncols <= 0 || v->n_indexed <= 0 || v->n_indexed > v->ncols || v->num_buckets <= 0 ||
      !v->bucket_offsets)
    HS_THROW(HS_EINVAL, "malformed index view");
  if (v->n_indexed > HS_MAX_KEYS) HS_THROW(HS_EUNSUPPORTED, "more than %d indexed columns", HS_MAX_KEYS);
}

void hs_offsets_to_device(hs_ctx* c, const int64_t* boff, int32_t nb, int64_t nrows, DBuf& out) {
  std::vector<uint32_t> h((size_t)nb + 1);
  for (int b = 0; b <= nb; b++) {
    if (boff[b] < 0 || boff[b] > nrows || (b && boff[b] < boff[b - 1])) HS_THROW(HS_EINVAL, "bad bucket_offsets");
    h[b] = (uint32_t)boff[b];
  }
  if (boff[0] != 0 || boff[nb] != nrows) HS_THROW(HS_EINVAL, "bucket_offsets do not cover the rows");
  out.alloc(c, ((size_t)nb + 1) * 4);
  // h is pageable memory: the call returns once the bytes sit in the driver's staging buffer, so h may go out of
  // scope without waiting for the stream
  hs_upload(c, out.p, h.data(), ((size_t)nb + 1) * 4);
}

// (left row, right row) pairs of a join driven from side A: for every row of A the matching range of B's run in
// the same bucket, turned into output counts according to the join type (INNER, LEFT_OUTER, LEFT_SEMI, LEFT_ANTI)
struct PairList {
  DBuf li, ri, d_ooff;          // row ids into A and B (ri unset for SEMI/ANTI); bucket offsets of the pairs (u64)
  int64_t m = 0;
  std::vector<int64_t> ooff;
};
static void probe_join(hs_ctx* ctx, const JoinSide& A, const JoinSide& B, int nk, int64_t na, int32_t nb, const DBuf& aoff,
                       int type, PairList& out) {
  DBuf lb(ctx, (size_t)na * 4), cnt(ctx, (size_t)na * 4), tcnt, off(ctx, (size_t)na * 4), total(ctx, 8);
  uint64_t h_total = 0;
  const uint32_t* counts = cnt.as<uint32_t>();
  {
    Stage s(ctx, "join_probe");
    const KCol& ak = A.ks.k[0];
    const KCol& bk = B.ks.k[0];
    bool fast = nk == 1 && ak.type == HS_INT64 && !ak.has_nulls && !bk.has_nulls;
    if (na > 0) {
      unsigned g = (unsigned)((na + 255) / 256);
      if (fast) LAUNCH(ctx, k_join_probe<true>, g, 256, 0, A, B, na, nb, lb.as<uint32_t>(), cnt.as<uint32_t>());
      else LAUNCH(ctx, k_join_probe<false>, g, 256, 0, A, B, na, nb, lb.as<uint32_t>(), cnt.as<uint32_t>());
      if (type != HS_JOIN_INNER) {
        tcnt.alloc(ctx, (size_t)na * 4);
        LAUNCH(ctx, k_join_type_counts, g, 256, 0, cnt.as<uint32_t>(), na, type, tcnt.as<uint32_t>());
        counts = tcnt.as<uint32_t>();
      }
    }
    hs_exclusive_scan_u32(ctx, counts, off.as<uint32_t>(), na, total.as<uint64_t>());
    CK(cudaMemcpyAsync(&h_total, total.p, 8, cudaMemcpyDeviceToHost, ctx->stream));
    hs_sync(ctx);
  }
  if (h_total >= 0xFFFFFFF0ull) HS_THROW(HS_EUNSUPPORTED, "join output exceeds 2^32-16 rows in one call");
  const int64_t m = (int64_t)h_total;
  const bool pairs = type == HS_JOIN_INNER || type == HS_JOIN_LEFT_OUTER;
  out.m = m;
  out.li.alloc(ctx, (size_t)m * 4);
  if (pairs) out.ri.alloc(ctx, (size_t)m * 4);
  out.d_ooff.alloc(ctx, ((size_t)nb + 1) * 8);
  out.ooff.assign((size_t)nb + 1, 0);
  Stage s(ctx, "join_expand");
  if (na > 0) {
    unsigned g = (unsigned)((na + 255) / 256);
    if (type == HS_JOIN_INNER)
      LAUNCH(ctx, k_join_expand, g, 256, 0, lb.as<uint32_t>(), cnt.as<uint32_t>(), off.as<uint32_t>(), na,
             out.li.as<uint32_t>(), out.ri.as<uint32_t>());
    else
      LAUNCH(ctx, k_join_expand_typed, g, 256, 0, lb.as<uint32_t>(), cnt.as<uint32_t>(), tcnt.as<uint32_t>(),
             off.as<uint32_t>(), na, out.li.as<uint32_t>(), pairs ? out.ri.as<uint32_t>() : (uint32_t*)nullptr);
  }
  LAUNCH(ctx, k_join_bucket_offsets, (unsigned)((nb + 256) / 256), 256, 0, aoff.as<uint32_t>(), off.as<uint32_t>(),
         total.as<uint64_t>(), na, nb, out.d_ooff.as<uint64_t>());
  CK(cudaMemcpyAsync(out.ooff.data(), out.d_ooff.p, ((size_t)nb + 1) * 8, cudaMemcpyDeviceToHost, ctx->stream));
  hs_sync(ctx);
}

// per key column: min and max of the order-preserving image of the (NULL-free) cells
__global__ void __launch_bounds__(256) k_keys_range(const __grid_constant__ KeySpec ks, int64_t n, uint64_t* __restrict__ mm) {
  uint64_t mn[HS_MAX_KEYS], mx[HS_MAX_KEYS];
#pragma unroll
  for (int k = 0; k < HS_MAX_KEYS; k++) { mn[k] = ~0ull; mx[k] = 0; }
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
#pragma unroll
    for (int k = 0; k < HS_MAX_KEYS; k++)
      if (k < ks.nkeys) {
        const uint64_t v = omap_cell(ks.k[k], i);
        mn[k] = v < mn[k] ? v : mn[k];
        mx[k] = v > mx[k] ? v : mx[k];
      }
  }
#pragma unroll
  for (int k = 0; k < HS_MAX_KEYS; k++)
    if (k < ks.nkeys) {
      for (int d = 16; d; d >>= 1) {
        const uint64_t a = __shfl_xor_sync(0xffffffffu, mn[k], d), b = __shfl_xor_sync(0xffffffffu, mx[k], d);
        mn[k] = a < mn[k] ? a : mn[k];
        mx[k] = b > mx[k] ? b : mx[k];
      }
      if (lane_id() == 0 && mn[k] <= mx[k]) {
        atomicMin((unsigned long long*)&mm[2 * k], (unsigned long long)mn[k]);
        atomicMax((unsigned long long*)&mm[2 * k + 1], (unsigned long long)mx[k]);
      }
    }
}
struct KeyPack {
  uint64_t min[HS_MAX_KEYS];
  int32_t shift[HS_MAX_KEYS];
};
// key tuple -> one u64: sum of (cell - min) << shift, first column in the highest bits (lexicographic order kept)
__global__ void __launch_bounds__(256) k_pack_keys(const __grid_constant__ KeySpec ks, int64_t n,
                                                   const __grid_constant__ KeyPack kp, uint64_t* __restrict__ out) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    uint64_t v = 0;
#pragma unroll
    for (int k = 0; k < HS_MAX_KEYS; k++)
      if (k < ks.nkeys) v |= (omap_cell(ks.k[k], i) - kp.min[k]) << kp.shift[k];
    out[i] = v;
  }
}

// Both sides' NULL-free fixed-width key tuples as one order-preserving u64 per row, with a common layout
// (per-column minimum over both sides, widths from the joint ranges). false if the ranges need more than 64 bits.
bool hs_pack_key_tuples(hs_ctx* ctx, const KeySpec& a, int64_t na, const KeySpec& b, int64_t nrows_b, int nk, DBuf& pa,
                        DBuf& pb) {
  DBuf mm(ctx, 2 * HS_MAX_KEYS * 8);
  uint64_t init[2 * HS_MAX_KEYS], h_mm[2 * HS_MAX_KEYS];
  for (int k = 0; k < HS_MAX_KEYS; k++) { init[2 * k] = ~0ull; init[2 * k + 1] = 0; }
  hs_upload(ctx, mm.p, init, sizeof init);
  if (na > 0) LAUNCH(ctx, k_keys_range, HS_NUM_SMS * 8, 256, 0, a, na, mm.as<uint64_t>());
  if (nrows_b > 0) LAUNCH(ctx, k_keys_range, HS_NUM_SMS * 8, 256, 0, b, nrows_b, mm.as<uint64_t>());
  CK(cudaMemcpyAsync(h_mm, mm.p, sizeof h_mm, cudaMemcpyDeviceToHost, ctx->stream));
  hs_sync(ctx);
  KeyPack kp;
  memset(&kp, 0, sizeof kp);
  int total_bits = 0;
  for (int k = nk - 1; k >= 0; k--) {
    uint64_t range = h_mm[2 * k] <= h_mm[2 * k + 1] ? h_mm[2 * k + 1] - h_mm[2 * k] : 0;
    int bits = 0;
    while (bits < 64 && (range >> bits)) bits++;
    kp.min[k] = h_mm[2 * k] <= h_mm[2 * k + 1] ? h_mm[2 * k] : 0;
    kp.shift[k] = total_bits;
    total_bits += bits;
  }
  if (total_bits > 64) return false;
  pa.alloc(ctx, (size_t)(na > 0 ? na : 1) * 8);
  pb.alloc(ctx, (size_t)(nrows_b > 0 ? nrows_b : 1) * 8);
  if (na > 0) LAUNCH(ctx, k_pack_keys, HS_NUM_SMS * 16, 256, 0, a, na, kp, pa.as<uint64_t>());
  if (nrows_b > 0) LAUNCH(ctx, k_pack_keys, HS_NUM_SMS * 16, 256, 0, b, nrows_b, kp, pb.as<uint64_t>());
  return true;
}

// (bucket, first left row, rows) of every tile from the left bucket offsets and the tiles before each bucket
__global__ void __launch_bounds__(256) k_join_make_tiles(const uint32_t* __restrict__ loff, const uint32_t* __restrict__ first_tile,
                                                         int nb, uint32_t ntiles, uint32_t* __restrict__ tb,
                                                         uint32_t* __restrict__ ts, uint32_t* __restrict__ tl) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= ntiles) return;
  int lo = 0, hi = nb;   // last bucket whose first tile is <= t (buckets without rows share their successor's)
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (first_tile[mid] <= t) lo = mid; else hi = mid;
  }
  const uint32_t a = loff[lo] + (t - first_tile[lo]) * JT, e = loff[lo + 1];
  tb[t] = (uint32_t)lo;
  ts[t] = a;
  tl[t] = e - a < JT ? e - a : JT;
}
// output offset of every bucket = output offset of its first tile (m for buckets after the last tile)
__global__ void __launch_bounds__(256) k_join_bucket_offsets(const uint32_t* __restrict__ tile_off, const uint32_t* __restrict__ first_tile,
                                                             int nb, uint32_t ntiles, const unsigned long long* last_state,
                                                             const uint32_t* overflow, uint32_t* __restrict__ out) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b > nb) return;
  const unsigned long long last = *last_state;
  const uint32_t m = (uint32_t)(last & ((1ull << 62) - 1));
  out[b] = first_tile[b] < ntiles ? tile_off[first_tile[b]] : m;
  if (b == 0 && overflow) {   // everything the host wants back, behind the offsets: one read
    out[nb + 1] = *overflow;
    out[nb + 2] = (uint32_t)last;
    out[nb + 3] = (uint32_t)(last >> 32);
  }
}

extern "C" int32_t hs_join_bucketed(hs_ctx* ctx, const hs_index_view* left, const hs_index_view* right,
                                    int32_t join_type, const int32_t* left_out, int32_t n_left_out,
                                    const int32_t* right_out, int32_t n_right_out, hs_table** out) {
  API_BEGIN(ctx)
  if (!out) HS_THROW(HS_EINVAL, "null out");
  *out = nullptr;
  check_view(left);
  check_view(right);
  if (join_type < HS_JOIN_INNER || join_type > HS_JOIN_LEFT_ANTI) HS_THROW(HS_EINVAL, "unknown join type %d", join_type);
  if ((join_type == HS_JOIN_LEFT_SEMI || join_type == HS_JOIN_LEFT_ANTI) && n_right_out != 0)
    HS_THROW(HS_EINVAL, "semi and anti joins return left columns only");
  if (left->n_indexed != right->n_indexed) HS_THROW(HS_EINVAL, "indexes differ in number of indexed columns");
  if (n_left_out < 0 || n_right_out < 0 || n_left_out + n_right_out <= 0) HS_THROW(HS_EINVAL, "no output columns");
  for (int i = 0; i < n_left_out; i++)
    if (left_out[i] < 0 || left_out[i] >= left->ncols) HS_THROW(HS_EINVAL, "left output column out of range");
  for (int i = 0; i < n_right_out; i++)
    if (right_out[i] < 0 || right_out[i] >= right->ncols) HS_THROW(HS_EINVAL, "right output column out of range");
  hs_stages_reset(ctx);
  const int nk = left->n_indexed;
  // the bucket count of the join: when the counts differ, that of the side with more rows (see below)
  int nb = left->num_buckets;
  std::vector<DevCol> lc, rc;
  {
    Stage s(ctx, "h2d");
    hs_stage_columns(ctx, left->cols, left->ncols, lc);
    hs_stage_columns(ctx, right->cols, right->ncols, rc);
  }
  for (int k = 0; k < nk; k++)
    if (lc[k].type != rc[k].type) HS_THROW(HS_EUNSUPPORTED, "join key %d differs in type between the indexes", k);
  const int64_t nl = lc[0].length, nr = rc[0].length;
  // Different bucket counts (JoinIndexRule accepts such a pair, JoinIndexRule.scala:515-516; Spark then
  // re-shuffles one side): the side with fewer rows is re-bucketed here to the other side's count — the same
  // hash + scatter + sort as a build, on the fly — and the join runs bucket by bucket as usual.
  std::vector<OutCol> rebuilt;
  std::vector<int64_t> rebuilt_off;
  const int64_t* lboff = left->bucket_offsets;
  const int64_t* rboff = right->bucket_offsets;
  if (left->num_buckets != right->num_buckets) {
    const bool redo_left = nl < nr;
    std::vector<DevCol>& side = redo_left ? lc : rc;
    const int32_t target = redo_left ? right->num_buckets : left->num_buckets;
    hs_build_core(ctx, side, nk, target, 0, target, rebuilt, rebuilt_off, nullptr);
    for (size_t i = 0; i < side.size(); i++) {
      DevCol v;
      v.type = rebuilt[i].type;
      v.length = rebuilt[i].length;
      v.null_count = rebuilt[i].null_count;
      v.validity = rebuilt[i].validity.p && rebuilt[i].null_count ? rebuilt[i].validity.as<uint8_t>() : nullptr;
      v.offsets = rebuilt[i].offsets.as<int32_t>();
      v.data = rebuilt[i].data.p;
      v.data_bytes = rebuilt[i].data_bytes;
      side[i] = std::move(v);
    }
    (redo_left ? lboff : rboff) = rebuilt_off.data();
    nb = target;
  }
  DBuf loff, roff;
  hs_offsets_to_device(ctx, lboff, nb, nl, loff);
  hs_offsets_to_device(ctx, rboff, nb, nr, roff);
  JoinSide L{hs_make_keyspec(lc, nk), loff.as<uint32_t>()};
  JoinSide R{hs_make_keyspec(rc, nk), roff.as<uint32_t>()};

  // ---- fast path: the key is one integer column, or any NULL-free fixed-width key tuple whose value ranges
  // fit 64 bits together (packed below); fixed-width NULL-free outputs
  const bool int_key = nk == 1 && (lc[0].type == HS_INT64 || lc[0].type == HS_INT32);
  bool fast_join = join_type == HS_JOIN_INNER && n_left_out <= 16 && n_right_out <= 16 && nl > 0 && nr > 0;
  for (int k = 0; k < nk && fast_join; k++)
    if (hs_is_varlen(lc[k].type) || lc[k].validity || rc[k].validity) fast_join = false;
  for (int i = 0; i < n_left_out && fast_join; i++) {
    int w = hs_fixed_width(lc[left_out[i]].type);
    if ((w != 4 && w != 8) || lc[left_out[i]].validity) fast_join = false;
  }
  for (int i = 0; i < n_right_out && fast_join; i++) {
    int w = hs_fixed_width(rc[right_out[i]].type);
    if ((w != 4 && w != 8) || rc[right_out[i]].validity) fast_join = false;
  }
  // composite (or non-integer) keys: both sides are rewritten as one order-preserving u64 per row, so that the
  // tile merge below compares plain integers (cfg5's (int32, int64) key: 12 bytes -> 8)
  bool packed = false;
  DBuf lpk, rpk;
  if (fast_join && !int_key) {
    Stage s(ctx, "join_pack_keys");
    packed = hs_pack_key_tuples(ctx, L.ks, nl, R.ks, nr, nk, lpk, rpk);
    if (!packed) fast_join = false;
  }
  if (fast_join) {
    // tiles of JT left rows that never cross a bucket: the host only counts them per bucket, a small kernel
    // writes the (bucket, first row, rows) of every tile
    std::vector<uint32_t> first_tile((size_t)nb + 1);
    uint32_t ntiles = 0;
    for (int b = 0; b < nb; b++) {
      first_tile[b] = ntiles;
      ntiles += (uint32_t)((lboff[b + 1] - lboff[b] + JT - 1) / JT);
    }
    first_tile[nb] = ntiles;
    DBuf d_tb(ctx, (size_t)ntiles * 4 + 4), d_ts(ctx, (size_t)ntiles * 4 + 4), d_tl(ctx, (size_t)ntiles * 4 + 4),
        d_cnt(ctx, (size_t)ntiles * 4 + 4), d_total(ctx, 8), d_win(ctx, (size_t)ntiles * 8 + 8),
        d_first(ctx, ((size_t)nb + 1) * 4), d_boff(ctx, ((size_t)nb + 4) * 4);
    hs_upload(ctx, d_first.p, first_tile.data(), ((size_t)nb + 1) * 4);
    if (ntiles)
      LAUNCH(ctx, k_join_make_tiles, (ntiles + 255) / 256, 256, 0, loff.as<uint32_t>(), d_first.as<uint32_t>(), nb, ntiles,
             d_tb.as<uint32_t>(), d_ts.as<uint32_t>(), d_tl.as<uint32_t>());
    JoinFast jf;
    memset(&jf, 0, sizeof jf);
    jf.lkey = packed ? lpk.p : lc[0].data;
    jf.rkey = packed ? rpk.p : rc[0].data;
    jf.roff = roff.as<uint32_t>();
    jf.tile_bucket = d_tb.as<uint32_t>(); jf.tile_start = d_ts.as<uint32_t>(); jf.tile_len = d_tl.as<uint32_t>();
    jf.ntiles = ntiles;
    jf.tile_cnt = d_cnt.as<uint32_t>();
    jf.tile_win = d_win.as<uint32_t>();
    // key kind: 0 = int32 column, 1 = int64 column, 2 = packed u64 (built above)
    const int kkind = packed ? 2 : (lc[0].type == HS_INT64 ? 1 : 0);
    const size_t smem = join_smem_bytes();
    auto launch_windows = [&]() {
      if (kkind == 2) LAUNCH(ctx, k_join_windows<uint64_t>, (ntiles + 7) / 8, 256, 0, jf);
      else if (kkind == 1) LAUNCH(ctx, k_join_windows<int64_t>, (ntiles + 7) / 8, 256, 0, jf);
      else LAUNCH(ctx, k_join_windows<int32_t>, (ntiles + 7) / 8, 256, 0, jf);
    };
    auto launch_tiles = [&](int mode) {
#define HS_JT(K, M) LAUNCH(ctx, (k_join_tiles<K, M>), ntiles, JTHREADS, smem, jf)
      if (kkind == 2) { if (mode == 0) HS_JT(uint64_t, 0); else if (mode == 1) HS_JT(uint64_t, 1); else HS_JT(uint64_t, 2); }
      else if (kkind == 1) { if (mode == 0) HS_JT(int64_t, 0); else if (mode == 1) HS_JT(int64_t, 1); else HS_JT(int64_t, 2); }
      else { if (mode == 0) HS_JT(int32_t, 0); else if (mode == 1) HS_JT(int32_t, 1); else HS_JT(int32_t, 2); }
#undef HS_JT
    };
    static bool attr_done[64] = {false};
    int dev = 0;
    CK(cudaGetDevice(&dev));
    if (!attr_done[dev & 63]) {
      CK(cudaFuncSetAttribute(k_join_tiles<int64_t, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)join_smem_bytes()));
      CK(cudaFuncSetAttribute(k_join_tiles<int64_t, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)join_smem_bytes()));
      CK(cudaFuncSetAttribute(k_join_tiles<int32_t, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)join_smem_bytes()));
      CK(cudaFuncSetAttribute(k_join_tiles<int32_t, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)join_smem_bytes()));
      CK(cudaFuncSetAttribute(k_join_tiles<int64_t, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)join_smem_bytes()));
      CK(cudaFuncSetAttribute(k_join_tiles<int32_t, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)join_smem_bytes()));
      CK(cudaFuncSetAttribute(k_join_tiles<uint64_t, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)join_smem_bytes()));
      CK(cudaFuncSetAttribute(k_join_tiles<uint64_t, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)join_smem_bytes()));
      CK(cudaFuncSetAttribute(k_join_tiles<uint64_t, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)join_smem_bytes()));
      attr_done[dev & 63] = true;
    }
    uint64_t h_tot = 0;
    std::vector<uint32_t> h_off((size_t)ntiles), h_boff((size_t)nb + 4);
    // ---- single pass first: the output is sized for one match per left row (a key -> unique-key join, the
    // shape JoinIndexRule mostly produces); tiles learn their output offset from a chained scan. If the matches
    // do not fit, the pass has only counted and the two-pass path below runs instead.
    const char* no_single = getenv("HSGPU_JOIN_TWO_PASS");
    if (!(no_single && *no_single && *no_single != '0')) {
      const int64_t cap = nl;
      std::vector<OutCol> ocols((size_t)(n_left_out + n_right_out));
      auto bind = [&](OutCol& o, const DevCol& sc) {
        o.type = sc.type; o.length = 0; o.null_count = 0;
        o.data.alloc(ctx, (size_t)cap * hs_fixed_width(sc.type) + 16);
      };
      for (int i = 0; i < n_left_out; i++) {
        const DevCol& sc = lc[left_out[i]];
        bind(ocols[i], sc);
        jf.lsrc[i] = sc.data; jf.ldst[i] = ocols[i].data.p; jf.les[i] = hs_fixed_width(sc.type);
      }
      for (int i = 0; i < n_right_out; i++) {
        const DevCol& sc = rc[right_out[i]];
        bind(ocols[n_left_out + i], sc);
        jf.rsrc[i] = sc.data; jf.rdst[i] = ocols[n_left_out + i].data.p; jf.res[i] = hs_fixed_width(sc.type);
      }
      jf.n_lout = n_left_out; jf.n_rout = n_right_out;
      DBuf d_state(ctx, (size_t)ntiles * 8 + 16);     // tile states, then the ticket and the overflow flag
      d_state.zero();
      jf.tile_state = d_state.as<unsigned long long>();
      jf.ticket = (uint32_t*)(d_state.as<unsigned long long>() + ntiles);
      jf.overflow = jf.ticket + 1;
      jf.capacity = (uint32_t)cap;
      {
        Stage s(ctx, "join_fused");
        launch_windows();
        launch_tiles(2);
        LAUNCH(ctx, k_join_bucket_offsets, (nb + 256) / 256, 256, 0, d_cnt.as<uint32_t>(), d_first.as<uint32_t>(), nb, ntiles,
               d_state.as<unsigned long long>() + (ntiles - 1), (const uint32_t*)jf.overflow, d_boff.as<uint32_t>());
        CK(cudaMemcpyAsync(h_boff.data(), d_boff.p, ((size_t)nb + 4) * 4, cudaMemcpyDeviceToHost, ctx->stream));
        hs_sync(ctx);
      }
      if (!h_boff[(size_t)nb + 1]) {
        const unsigned long long h_last = (unsigned long long)h_boff[(size_t)nb + 2] | ((unsigned long long)h_boff[(size_t)nb + 3] << 32);
        const int64_t m = (int64_t)(h_last & ((1ull << 62) - 1));
        for (size_t i = 0; i < ocols.size(); i++) {
          ocols[i].length = m;
          ocols[i].data_bytes = m * hs_fixed_width(ocols[i].type);
        }
        std::vector<int64_t> ooff((size_t)nb + 1, 0);
        for (int b = 0; b <= nb; b++) ooff[b] = (int64_t)h_boff[b];
        Stage s(ctx, "d2h");
        *out = hs_finish_table(ctx, ocols, m, nb, &ooff);
        return HS_OK;
      }
      // more matches than left rows: fall through to the exact two-pass path
    }
    {
      Stage s(ctx, "join_count");
      launch_windows();
      launch_tiles(0);
      hs_exclusive_scan_u32(ctx, d_cnt.as<uint32_t>(), d_cnt.as<uint32_t>(), ntiles, d_total.as<uint64_t>());
      CK(cudaMemcpyAsync(&h_tot, d_total.p, 8, cudaMemcpyDeviceToHost, ctx->stream));
      CK(cudaMemcpyAsync(h_off.data(), d_cnt.p, (size_t)ntiles * 4, cudaMemcpyDeviceToHost, ctx->stream));
      hs_sync(ctx);
    }
    if (h_tot >= 0xFFFFFFF0ull) HS_THROW(HS_EUNSUPPORTED, "join output exceeds 2^32-16 rows in one call");
    const int64_t m = (int64_t)h_tot;
    std::vector<OutCol> ocols((size_t)(n_left_out + n_right_out));
    for (int i = 0; i < n_left_out; i++) {
      const DevCol& sc = lc[left_out[i]];
      OutCol& o = ocols[i];
      o.type = sc.type; o.length = m; o.null_count = 0;
      o.data_bytes = m * hs_fixed_width(sc.type);
      o.data.alloc(ctx, (size_t)o.data_bytes);
      jf.lsrc[i] = sc.data; jf.ldst[i] = o.data.p; jf.les[i] = hs_fixed_width(sc.type);
    }
    for (int i = 0; i < n_right_out; i++) {
      const DevCol& sc = rc[right_out[i]];
      OutCol& o = ocols[n_left_out + i];
      o.type = sc.type; o.length = m; o.null_count = 0;
      o.data_bytes = m * hs_fixed_width(sc.type);
      o.data.alloc(ctx, (size_t)o.data_bytes);
      jf.rsrc[i] = sc.data; jf.rdst[i] = o.data.p; jf.res[i] = hs_fixed_width(sc.type);
    }
    jf.n_lout = n_left_out; jf.n_rout = n_right_out;
    {
      Stage s(ctx, "join_write");
      launch_tiles(1);
    }
    std::vector<int64_t> ooff((size_t)nb + 1, 0);
    for (int b = 0; b <= nb; b++) ooff[b] = first_tile[b] < ntiles ? (int64_t)h_off[first_tile[b]] : m;
    Stage s(ctx, "d2h");
    *out = hs_finish_table(ctx, ocols, m, nb, &ooff);
    return HS_OK;
  }

  // ---- general path: per-row search of the other side's run, scan, pair lists, gathers
  PairList pl;
  bool miss_left = false, miss_right = false;
  if (join_type == HS_JOIN_RIGHT_OUTER) {
    probe_join(ctx, R, L, nk, nr, nb, roff, HS_JOIN_LEFT_OUTER, pl);
    std::swap(pl.li, pl.ri);
    miss_left = true;
  } else if (join_type == HS_JOIN_FULL_OUTER) {
    PairList a, b;
    probe_join(ctx, L, R, nk, nl, nb, loff, HS_JOIN_LEFT_OUTER, a);
    probe_join(ctx, R, L, nk, nr, nb, roff, HS_JOIN_LEFT_ANTI, b);
    pl.m = a.m + b.m;
    if ((uint64_t)pl.m >= 0xFFFFFFF0ull) HS_THROW(HS_EUNSUPPORTED, "join output exceeds 2^32-16 rows in one call");
    pl.li.alloc(ctx, (size_t)pl.m * 4);
    pl.ri.alloc(ctx, (size_t)pl.m * 4);
    pl.ooff.resize((size_t)nb + 1);
    for (int i = 0; i <= nb; i++) pl.ooff[i] = a.ooff[i] + b.ooff[i];
    if (pl.m > 0) {
      Stage s(ctx, "join_expand");
      LAUNCH(ctx, k_full_outer_merge, (unsigned)((pl.m + 255) / 256), 256, 0, a.li.as<uint32_t>(), a.ri.as<uint32_t>(), a.m,
             a.d_ooff.as<uint64_t>(), b.li.as<uint32_t>(), b.m, b.d_ooff.as<uint64_t>(), nb, pl.li.as<uint32_t>(),
             pl.ri.as<uint32_t>());
    }
    miss_left = miss_right = true;
  } else {
    probe_join(ctx, L, R, nk, nl, nb, loff, join_type, pl);
    miss_right = join_type == HS_JOIN_LEFT_OUTER;
  }
  const int64_t m = pl.m;
  std::vector<OutCol> ocols((size_t)(n_left_out + n_right_out));
  {
    Stage s(ctx, "join_gather");
    for (int i = 0; i < n_left_out; i++) hs_gather_column(ctx, lc[left_out[i]], pl.li.as<uint32_t>(), m, ocols[i], miss_left);
    for (int i = 0; i < n_right_out; i++)
      hs_gather_column(ctx, rc[right_out[i]], pl.ri.as<uint32_t>(), m, ocols[n_left_out + i], miss_right);
  }
  Stage s(ctx, "d2h");
  *out = hs_finish_table(ctx, ocols, m, nb, &pl.ooff);
  API_END(ctx)
}
