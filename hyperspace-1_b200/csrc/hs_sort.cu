// hs_sort.cu — K2 + K3: the row mover.
//  * bucket scatter: rows -> bucket-contiguous layout (stands in for Spark's ShuffleExchangeExec under
//    df.repartition(numBuckets, cols), HS/actions/CreateActionBase.scala:129-130)
//  * segmented sort: every bucket sorted on the normalised key (stands in for the SortExec that
//    FileFormatWriter plans for BucketSpec(n, cols, cols), HS/index/DataFrameWriterExtensions.scala:57-66)
// One tile kernel (k_partition) does both movements: the column slices of a 4096-row tile arrive in shared
// memory by TMA bulk copies, the rows are ranked by digit (bucket id, or the next key bits) with shared-memory
// atomics, and every column is written out as runs per digit. The sort is an MSD radix partition (up to 8 bits
// per level, only for segments larger than a leaf) finished by an in-shared-memory LSD radix sort of each leaf
// segment over the bits that actually vary inside it (k_leaf_sort32 / k_leaf_sort / k_leaf_small).
#include <algorithm>

#include "hs_block.cuh"
#include "hs_kernels.h"
#include "hs_tma.cuh"

#define LEAF_MAX PT_TILE
#ifndef MSD_BITS
#define MSD_BITS 8
#endif
#define MSD_RADIX (1 << MSD_BITS)
#define MSD_TARGET_ROWS (3300 * PT_TILE / 4096)   // aimed-at child size of an MSD level (a leaf holds at most PT_TILE rows)

// Digit width of the MSD level that splits a segment of `len` rows whose keys still differ in their low
// `r` bits: as many bits (at most 8) as keep the expected child, len / 2^bits, near MSD_TARGET_ROWS. A huge
// segment takes all 8 bits; a segment just above a leaf is only halved, so its children stay full leaves
// instead of 256 crumbs. Every kernel of a level derives the same value from (len, r).
__host__ __device__ __forceinline__ int msd_bits(uint32_t len, uint32_t r) {
  int b = 1;
  while (b < MSD_BITS && ((uint64_t)len >> b) > MSD_TARGET_ROWS) b++;
  return b < (int)r ? b : (int)r;
}

struct SegList {           // device arrays
  uint32_t* start;
  uint32_t* len;
  uint32_t* shift;         // unsorted low bits of (key - base) inside the segment
  uint64_t* base;          // subtracted from the key before digits are taken (0 until a level re-ranges)
};

#define HS_MAX_PEERS 8
#define HS_MAX_PEER_COLS 16

struct MoveArgs {
  int32_t n;
  int32_t esize[HS_MAX_MOVE];
  const void* src[HS_MAX_MOVE];
  void* dst[HS_MAX_MOVE];
};

__device__ __forceinline__ uint64_t load_key(const MoveArgs& mv, const KeySrc& ks, uint32_t row) {
  const void* p = mv.src[ks.col];
  if (ks.mode == KS_U64) return ((const uint64_t*)p)[row];
  if (ks.mode == KS_I64) return (((const uint64_t*)p)[row] ^ 0x8000000000000000ull) - ks.min;
  return ((uint64_t)(int64_t)((const int32_t*)p)[row] ^ 0x8000000000000000ull) - ks.min;
}

__device__ __forceinline__ uint64_t load_key_at(const void* p, const KeySrc& ks, uint32_t row) {
  if (ks.mode == KS_U64) return ((const uint64_t*)p)[row];
  if (ks.mode == KS_I64) return (((const uint64_t*)p)[row] ^ 0x8000000000000000ull) - ks.min;
  return ((uint64_t)(int64_t)((const int32_t*)p)[row] ^ 0x8000000000000000ull) - ks.min;
}

// shared memory carve-up of the partition kernel
#define STAGE_BYTES (PT_TILE * 8 + 128)   // one column slice of a tile + alignment slack
struct TileSmem {
  unsigned char* stage[2];  // TMA landing buffers (double buffered over the columns)
  uint32_t* gdst;           // PT_TILE: global destination row of every tile-sorted position
  uint16_t* inv;            // PT_TILE: tile row that lands at every tile-sorted position
  uint32_t* cnt;            // radix_pad: counters, then first position of every digit
  uint32_t* tot;            // radix_pad: rows per digit
  uint32_t* gbase;          // radix_pad: reserved global run start per digit
  uint32_t* scratch;        // PT_SCRATCH
  uint64_t* bar;            // 3 mbarriers
  unsigned char* bstage;    // TMA landing buffer of the tile's bucket ids (u16)
};
__device__ __forceinline__ TileSmem carve(unsigned char* base, int radix_pad) {
  TileSmem s;
  s.stage[0] = base;
  s.stage[1] = base + STAGE_BYTES;
  s.gdst = (uint32_t*)(base + 2 * STAGE_BYTES);
  s.inv = (uint16_t*)(s.gdst + PT_TILE);
  s.cnt = (uint32_t*)(s.inv + PT_TILE);
  s.tot = s.cnt + radix_pad;
  s.gbase = s.tot + radix_pad;
  s.scratch = s.gbase + radix_pad;
  s.bar = (uint64_t*)(s.scratch + PT_SCRATCH);
  s.bstage = (unsigned char*)(s.bar + 4);
  return s;
}
static size_t tile_smem_bytes(int radix_pad) {
  return (size_t)2 * STAGE_BYTES + (size_t)PT_TILE * 4 + (size_t)PT_TILE * 2 + (size_t)3 * radix_pad * 4 +
         PT_SCRATCH * 4 + 32 + (size_t)PT_TILE * 2 + 32;
}

struct PartParams {
  MoveArgs mv;
  KeySrc key;               // MODE 1
  const uint16_t* bucket;   // MODE 0
  uint32_t n;               // MODE 0: rows
  int32_t radix;            // MODE 0: number of buckets
  int32_t radix_pad;
  int32_t bits;             // MODE 0: bits of a bucket id
  int32_t bucket_shift;     // MODE 0: digit = bucket id >> bucket_shift (coarse pass of the two-level scatter)
  int32_t bucket_base;      // MODE 0: subtracted from the bucket id first (ids of a rank's own bucket range)
  const uint16_t* bin_lut;  // MODE 0: when set, digit = bin_lut[bucket id] (bins that are unions of buckets)
  int32_t by_owner;         // MODE 2: digit = owner of the bucket (nranks bins) instead of the bucket itself
  int32_t nb_total;         // MODE 2, by_owner: number of buckets of the index
  int32_t conv_col1;        // 1 + the column converted on the way (0 = none): MODE 2 writes int64 as u32
                            // ((v ^ sign) - conv_min), MODE 0 reads such u32 and writes the int64 back
  uint64_t conv_min;
  uint32_t* cursor;         // MODE 0: [radix]; MODE 1: [nseg * 256]
  SegList segs;             // MODE 1: the large segments of this level
  const uint32_t* tile_prefix;  // MODE 1: [nseg + 1]
  uint32_t nseg;
  // MODE 2 (multi-GPU): bucket d belongs to rank d * nranks / radix; column c of that rank's receive buffer is
  // pdst[rank][c] (peer-mapped memory); cursor[d] starts at this sender's run inside the owner's bucket
  int32_t nranks;
  void* pdst[HS_MAX_PEERS][HS_MAX_PEER_COLS];
};

__device__ __forceinline__ uint32_t find_segment(const uint32_t* tile_prefix, uint32_t nseg, uint32_t tile) {
  uint32_t lo = 0, hi = nseg;  // largest s with tile_prefix[s] <= tile
  while (hi - lo > 1) {
    uint32_t mid = (lo + hi) >> 1;
    if (tile_prefix[mid] <= tile) lo = mid; else hi = mid;
  }
  return lo;
}

template <typename T>
__device__ __forceinline__ void scatter_column(const unsigned char* stage, uint32_t lead, T* __restrict__ dst,
                                               uint32_t tile_len, const uint16_t* inv, const uint32_t* gdst) {
  const T* col = (const T*)(stage + lead);
#pragma unroll 4
  for (int k = 0; k < PT_ITEMS; k++) {
    uint32_t j = threadIdx.x + k * PT_THREADS;
    if (j < tile_len) dst[gdst[j]] = col[inv[j]];
  }
}

template <typename T>
__device__ __forceinline__ void scatter_column_peers(const unsigned char* stage, uint32_t lead, void* const (*pdst)[HS_MAX_PEER_COLS],
                                                     int c, uint32_t tile_len, const uint16_t* inv, const uint32_t* gdst) {
  const T* col = (const T*)(stage + lead);
#pragma unroll 4
  for (int k = 0; k < PT_ITEMS; k++) {
    uint32_t j = threadIdx.x + k * PT_THREADS;
    if (j < tile_len) {
      const uint32_t iv = inv[j];
      T* dst = (T*)pdst[iv >> 12][c];            // the owner's buffer (local or over NVLink)
      dst[gdst[j]] = col[iv & 0xFFFu];
    }
  }
}

#define HS_SIGN64 0x8000000000000000ull
__device__ __forceinline__ void scatter_column_peers_narrow(const unsigned char* stage, uint32_t lead,
                                                            void* const (*pdst)[HS_MAX_PEER_COLS], int c, uint32_t tile_len,
                                                            const uint16_t* inv, const uint32_t* gdst, uint64_t mn) {
  const uint64_t* col = (const uint64_t*)(stage + lead);
#pragma unroll 4
  for (int k = 0; k < PT_ITEMS; k++) {
    uint32_t j = threadIdx.x + k * PT_THREADS;
    if (j < tile_len) {
      const uint32_t iv = inv[j];
      ((uint32_t*)pdst[iv >> 12][c])[gdst[j]] = (uint32_t)((col[iv & 0xFFFu] ^ HS_SIGN64) - mn);
    }
  }
}
__device__ __forceinline__ void scatter_column_widen(const unsigned char* stage, uint32_t lead, uint64_t* __restrict__ dst,
                                                     uint32_t tile_len, const uint16_t* inv, const uint32_t* gdst,
                                                     uint64_t mn) {
  const uint32_t* col = (const uint32_t*)(stage + lead);
#pragma unroll 4
  for (int k = 0; k < PT_ITEMS; k++) {
    uint32_t j = threadIdx.x + k * PT_THREADS;
    if (j < tile_len) dst[gdst[j]] = ((uint64_t)col[inv[j]] + mn) ^ HS_SIGN64;
  }
}

// MODE 0: scatter by bucket id.  MODE 1: one MSD level inside large segments.
// MODE 2: scatter by bucket id straight into the receive buffers of the GPUs that own the buckets.
// Column slices of the tile arrive in shared memory by TMA bulk copies (double buffered, one mbarrier
// per buffer); rows are ranked by digit with shared-memory atomics; one global atomicAdd per (tile, digit)
// reserves the output run; every column is then read through the tile permutation and written as
// contiguous runs per digit.
// The tile body shared by the three launch modes and by the pipelined sort kernel. dst_bias is added (mod 2^32) to
// every destination row: the pipelined sort writes an MSD level into a small re-used scratch area instead of
// the rows' own positions.
template <int MODE>
__device__ __forceinline__ void partition_tile(const PartParams& p, unsigned char* smem_raw, uint32_t tile_start,
                                               uint32_t tile_len, int bits, uint32_t shift, uint64_t seg_base,
                                               uint32_t* cursor, uint32_t dst_bias, uint64_t* ext_bar = nullptr,
                                               uint32_t* ext_phase = nullptr) {
  // ext_bar / ext_phase: three mbarriers that live on across calls (persistent kernel) and the parity each one
  // completes next (bit b of *ext_phase); without them the barriers of the carve-up are initialised here
  TileSmem sm = carve(smem_raw, p.radix_pad);
  if (ext_bar) sm.bar = ext_bar;
  const uint32_t ph_in = ext_phase ? *ext_phase : 0u;
  const int ncols = p.mv.n;
  // column order: in MODE 1 the key column goes first (its slice also feeds the digits)
  const int first = MODE == 1 ? p.key.col : 0;
  auto col_at = [&](int i) { return i == 0 ? first : (i <= first ? i - 1 : i); };
  if (!ext_bar) {
    if (threadIdx.x == 0) {
      mbar_init(&sm.bar[0], 1);
      mbar_init(&sm.bar[1], 1);
      mbar_init(&sm.bar[2], 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    if (MODE != 1) issue_slice(p.bucket, 2, tile_start, tile_len, sm.bstage, &sm.bar[2]);
    int c0 = col_at(0);
    issue_slice(p.mv.src[c0], p.mv.esize[c0], tile_start, tile_len, sm.stage[0], &sm.bar[0]);
    if (ncols > 1) {
      int c1 = col_at(1);
      issue_slice(p.mv.src[c1], p.mv.esize[c1], tile_start, tile_len, sm.stage[1], &sm.bar[1]);
    }
  }
  uint32_t digit[PT_ITEMS], pos[PT_ITEMS];
  uint32_t vmask = 0;
  if (MODE == 1) mbar_wait(&sm.bar[0], ph_in & 1u);
  else mbar_wait(&sm.bar[2], (ph_in >> 2) & 1u);
  const uint16_t* sbucket = (const uint16_t*)(sm.bstage + slice_lead(p.bucket, 2, tile_start));
#pragma unroll
  for (int k = 0; k < PT_ITEMS; k++) {
    uint32_t idx = threadIdx.x + k * PT_THREADS;
    bool v = idx < tile_len;
    vmask |= (uint32_t)v << k;
    digit[k] = 0;
    if (v) {
      if (MODE == 0) {
        digit[k] = p.bin_lut ? (uint32_t)__ldg(p.bin_lut + sbucket[idx])
                             : ((uint32_t)sbucket[idx] - (uint32_t)p.bucket_base) >> p.bucket_shift;
      } else if (MODE == 2) {
        digit[k] = p.by_owner ? ((uint32_t)sbucket[idx] * (uint32_t)p.nranks) / (uint32_t)p.nb_total : sbucket[idx];
      } else {
        const unsigned char* ks = sm.stage[0] + slice_lead(p.mv.src[first], p.mv.esize[first], tile_start);
        uint64_t key;
        if (p.key.mode == KS_U64) key = ((const uint64_t*)ks)[idx];
        else if (p.key.mode == KS_I64) key = (((const uint64_t*)ks)[idx] ^ 0x8000000000000000ull) - p.key.min;
        else key = ((uint64_t)(int64_t)((const int32_t*)ks)[idx] ^ 0x8000000000000000ull) - p.key.min;
        digit[k] = (uint32_t)((key - seg_base) >> shift) & ((1u << bits) - 1u);
      }
    }
  }
  // the order of rows inside a digit is irrelevant here (the sort below fixes it): rank with atomics
  tile_rank_atomic_reserve<PT_ITEMS>(digit, vmask, PT_ITEMS, p.radix_pad, sm.cnt, sm.tot, sm.gbase, cursor,
                                     sm.scratch, pos);
#pragma unroll
  for (int k = 0; k < PT_ITEMS; k++) {
    if ((vmask >> k) & 1u) {
      sm.gdst[pos[k]] = sm.gbase[digit[k]] + pos[k] - sm.cnt[digit[k]] + dst_bias;
      uint32_t iv = threadIdx.x + k * PT_THREADS;
      if (MODE == 2)
        iv |= (p.by_owner ? digit[k] : (uint32_t)(((uint64_t)digit[k] * (uint32_t)p.nranks) / (uint32_t)p.radix)) << 12;
      sm.inv[pos[k]] = (uint16_t)iv;
    }
  }
  __syncthreads();
  for (int i = 0; i < ncols; i++) {
    const int c = col_at(i), b = i & 1, es = p.mv.esize[c];
    mbar_wait(&sm.bar[b], ((uint32_t)(i >> 1) ^ (ph_in >> b)) & 1u);
    const uint32_t lead = slice_lead(p.mv.src[c], es, tile_start);
    if (MODE == 2 && c + 1 == p.conv_col1) {
      scatter_column_peers_narrow(sm.stage[b], lead, p.pdst, c, tile_len, sm.inv, sm.gdst, p.conv_min);
    } else if (MODE == 0 && c + 1 == p.conv_col1) {
      scatter_column_widen(sm.stage[b], lead, (uint64_t*)p.mv.dst[c], tile_len, sm.inv, sm.gdst, p.conv_min);
    } else if (MODE == 2) {
      switch (es) {
        case 8: scatter_column_peers<uint64_t>(sm.stage[b], lead, p.pdst, c, tile_len, sm.inv, sm.gdst); break;
        case 4: scatter_column_peers<uint32_t>(sm.stage[b], lead, p.pdst, c, tile_len, sm.inv, sm.gdst); break;
        case 2: scatter_column_peers<uint16_t>(sm.stage[b], lead, p.pdst, c, tile_len, sm.inv, sm.gdst); break;
        default: scatter_column_peers<uint8_t>(sm.stage[b], lead, p.pdst, c, tile_len, sm.inv, sm.gdst); break;
      }
    } else
    switch (es) {
      case 8: scatter_column<uint64_t>(sm.stage[b], lead, (uint64_t*)p.mv.dst[c], tile_len, sm.inv, sm.gdst); break;
      case 4: scatter_column<uint32_t>(sm.stage[b], lead, (uint32_t*)p.mv.dst[c], tile_len, sm.inv, sm.gdst); break;
      case 2: scatter_column<uint16_t>(sm.stage[b], lead, (uint16_t*)p.mv.dst[c], tile_len, sm.inv, sm.gdst); break;
      default: scatter_column<uint8_t>(sm.stage[b], lead, (uint8_t*)p.mv.dst[c], tile_len, sm.inv, sm.gdst); break;
    }
    __syncthreads();  // everyone is done with stage[b]: it can be refilled
    if (threadIdx.x == 0 && i + 2 < ncols) {
      int cn = col_at(i + 2);
      issue_slice(p.mv.src[cn], p.mv.esize[cn], tile_start, tile_len, sm.stage[b], &sm.bar[b]);
    }
  }
  if (ext_phase)
    *ext_phase = ph_in ^ (uint32_t)(((ncols + 1) >> 1) & 1) ^ (uint32_t)(((ncols >> 1) & 1) << 1) ^ (MODE != 1 ? 4u : 0u);
}

template <int MODE>
__global__ void __launch_bounds__(PT_THREADS, 2 * PT_MIN_CTAS / 2) k_partition(const __grid_constant__ PartParams p) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  uint32_t tile_start, tile_len, shift = 0;
  uint64_t seg_base = 0;
  int bits = p.bits;
  uint32_t* cursor = p.cursor;
  if (MODE == 0 || MODE == 2) {
    tile_start = blockIdx.x * (uint32_t)PT_TILE;
    uint32_t left = p.n - tile_start;
    tile_len = left < PT_TILE ? left : PT_TILE;
  } else {
    uint32_t s = find_segment(p.tile_prefix, p.nseg, blockIdx.x);
    uint32_t t = blockIdx.x - p.tile_prefix[s];
    uint32_t slen = p.segs.len[s];
    tile_start = p.segs.start[s] + t * PT_TILE;
    uint32_t left = slen - t * PT_TILE;
    tile_len = left < PT_TILE ? left : PT_TILE;
    uint32_t sh = p.segs.shift[s];
    bits = msd_bits(slen, sh);
    shift = sh - bits;
    seg_base = p.segs.base[s];
    cursor = p.cursor + (size_t)s * MSD_RADIX;
  }
  partition_tile<MODE>(p, smem_raw, tile_start, tile_len, bits, shift, seg_base, cursor, 0u);
}

// per-segment digit histogram of one MSD level
__device__ __forceinline__ void msd_hist_tile(const MoveArgs& mv, const KeySrc& key, uint32_t tile_start,
                                              uint32_t tile_len, uint64_t seg_base, uint32_t shift, int bits,
                                              uint32_t* hist /* shared, MSD_RADIX */, uint32_t* cursor) {
  for (int i = threadIdx.x; i < MSD_RADIX; i += PT_THREADS) hist[i] = 0;
  __syncthreads();
  // all keys of the thread first (independent loads), then the shared-memory updates
  uint64_t kk[PT_ITEMS];
#pragma unroll
  for (int k = 0; k < PT_ITEMS; k++) {
    const uint32_t j = threadIdx.x + k * PT_THREADS;
    kk[k] = j < tile_len ? load_key(mv, key, tile_start + j) : 0;
  }
#pragma unroll
  for (int k = 0; k < PT_ITEMS; k++) {
    const uint32_t j = threadIdx.x + k * PT_THREADS;
    if (j < tile_len) atomicAdd(&hist[(uint32_t)((kk[k] - seg_base) >> shift) & ((1u << bits) - 1u)], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < MSD_RADIX; i += PT_THREADS) {
    uint32_t v = hist[i];
    if (v) atomicAdd(&cursor[i], v);
  }
}

__global__ void __launch_bounds__(PT_THREADS) k_msd_hist(const __grid_constant__ PartParams p) {
  __shared__ uint32_t hist[MSD_RADIX];
  uint32_t s = find_segment(p.tile_prefix, p.nseg, blockIdx.x);
  uint32_t t = blockIdx.x - p.tile_prefix[s];
  uint32_t slen = p.segs.len[s];
  uint32_t tile_start = p.segs.start[s] + t * PT_TILE;
  uint32_t left = slen - t * PT_TILE;
  uint32_t tile_len = left < PT_TILE ? left : PT_TILE;
  uint32_t sh = p.segs.shift[s];
  int bits = msd_bits(slen, sh);
  msd_hist_tile(p.mv, p.key, tile_start, tile_len, p.segs.base[s], sh - bits, bits, hist,
                p.cursor + (size_t)s * MSD_RADIX);
}

#define SMALL_MAX 64   // segments up to this many rows are sorted by one warp each (k_leaf_small)
struct PlanOut {
  SegList leaf, large, copy, small;
  uint32_t* counters;  // [0] leaf, [1] large, [2] copy, [3] tiles of large, [4] wide leaves, [5] small
};

__device__ __forceinline__ void classify_push(const PlanOut& o, uint32_t start, uint32_t len, uint32_t shift,
                                              uint64_t base) {
  if (len == 0) return;
  if (len <= SMALL_MAX) {
    uint32_t i = atomicAdd(&o.counters[5], 1u);
    o.small.start[i] = start; o.small.len[i] = len; o.small.shift[i] = shift;
  } else if (len <= LEAF_MAX) {
    uint32_t i = atomicAdd(&o.counters[0], 1u);
    o.leaf.start[i] = start; o.leaf.len[i] = len; o.leaf.shift[i] = shift; o.leaf.base[i] = base;
  } else if (shift == 0) {
    uint32_t i = atomicAdd(&o.counters[2], 1u);
    o.copy.start[i] = start; o.copy.len[i] = len; o.copy.shift[i] = 0;
  } else {
    uint32_t i = atomicAdd(&o.counters[1], 1u);
    o.large.start[i] = start; o.large.len[i] = len; o.large.shift[i] = shift; o.large.base[i] = base;
    atomicAdd(&o.counters[3], (len + PT_TILE - 1) / PT_TILE);
  }
}

// turn the histograms of one level into write cursors and child segments
__global__ void __launch_bounds__(MSD_RADIX) k_msd_plan(SegList segs, uint32_t* cursor, PlanOut out) {
  __shared__ uint32_t scratch[32];
  uint32_t s = blockIdx.x;
  uint32_t h = cursor[(size_t)s * MSD_RADIX + threadIdx.x];
  uint32_t ex = block256_excl_scan_u32(h, scratch, MSD_RADIX);
  uint32_t start = segs.start[s] + ex;
  cursor[(size_t)s * MSD_RADIX + threadIdx.x] = start;
  uint32_t sh = segs.shift[s];
  uint32_t bits = (uint32_t)msd_bits(segs.len[s], sh);
  classify_push(out, start, h, sh - bits, segs.base[s]);
}

// level 0: classify the caller's segments
__global__ void __launch_bounds__(256) k_classify(const uint32_t* seg_off, const uint32_t* lstart, const uint32_t* llen,
                                                  int64_t nseg, uint32_t key_bits, PlanOut out) {
  int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= nseg) return;
  uint32_t start, len;
  if (seg_off) { start = seg_off[s]; len = seg_off[s + 1] - start; }
  else { start = lstart[s]; len = llen[s]; }
  if (len < 2) return;  // nothing to sort; level 0 runs in place (src == dst == buffer A)
  classify_push(out, start, len, key_bits, 0);
}

// Deeper levels only see segments that one level could not break up (skewed or duplicate-heavy keys): measure
// the real key range of each, so that the next digit is taken from bits that actually vary.
__global__ void __launch_bounds__(PT_THREADS) k_seg_range(const __grid_constant__ PartParams p, uint64_t* rmin,
                                                          uint64_t* rmax) {
  uint32_t s = find_segment(p.tile_prefix, p.nseg, blockIdx.x);
  uint32_t t = blockIdx.x - p.tile_prefix[s];
  uint32_t slen = p.segs.len[s];
  uint32_t tile_start = p.segs.start[s] + t * PT_TILE;
  uint32_t left = slen - t * PT_TILE;
  uint32_t tile_len = left < PT_TILE ? left : PT_TILE;
  uint64_t mn = ~0ull, mx = 0;
  for (uint32_t j = threadIdx.x; j < tile_len; j += PT_THREADS) {
    uint64_t k = load_key(p.mv, p.key, tile_start + j);
    mn = k < mn ? k : mn;
    mx = k > mx ? k : mx;
  }
  mn = warp_min_u64(mn);
  mx = warp_max_u64(mx);
  if (lane_id() == 0 && mn <= mx) {
    atomicMin((unsigned long long*)&rmin[s], (unsigned long long)mn);
    atomicMax((unsigned long long*)&rmax[s], (unsigned long long)mx);
  }
}
__global__ void __launch_bounds__(256) k_seg_rerange(SegList segs, uint32_t nseg, const uint64_t* rmin,
                                                     const uint64_t* rmax) {
  uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= nseg) return;
  uint64_t mn = rmin[s], mx = rmax[s];
  uint64_t range = mx >= mn ? mx - mn : 0;
  segs.base[s] = mn;
  segs.shift[s] = range ? 64 - __clzll((long long)range) : 0;
}

// exclusive prefix of tiles over the large list (single CTA; the list is short)
__global__ void __launch_bounds__(1024) k_tile_prefix(SegList segs, const uint32_t* counters, uint32_t* tile_prefix) {
  __shared__ uint32_t carry;
  __shared__ uint32_t wsum[32];
  uint32_t n = counters[1];
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (uint32_t base = 0; base < n; base += 1024) {
    uint32_t i = base + threadIdx.x;
    uint32_t v = i < n ? (segs.len[i] + PT_TILE - 1) / PT_TILE : 0;
    uint32_t inc = warp_incl_scan_u32(v);
    int w = threadIdx.x >> 5;
    if (lane_id() == 31) wsum[w] = inc;
    __syncthreads();
    if (w == 0) {
      uint32_t s = wsum[lane_id()];
      uint32_t si = warp_incl_scan_u32(s);
      wsum[lane_id()] = si - s;
    }
    __syncthreads();
    uint32_t ex = inc - v + wsum[w] + carry;
    if (i < n) tile_prefix[i] = ex;
    __syncthreads();
    if (threadIdx.x == 1023) carry = ex + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) tile_prefix[n] = carry;
}

struct LeafParams {
  MoveArgs mv;        // src = where the leaf segments live now, dst = buffer A
  KeySrc key;
  SegList segs;
  int32_t in_place;   // src == dst
  uint8_t* wide;      // per segment: 1 = key range needs the 64-bit kernel (written by k_leaf_sort32)
  const uint32_t* count;  // device: number of segments in the list (the grid is an upper bound)
  uint32_t* n_wide;       // device: how many segments were flagged wide
  long long* trace;       // diagnostics: 16 clock values per segment, or null
};

// Sort one segment of at most LEAF_MAX rows entirely in shared memory: LSD radix over the bits that
// vary inside the segment (min/max of its keys), 8 bits per pass, stable; then move every column.
__device__ __forceinline__ void leaf64_body(unsigned char* smem_raw, const int32_t* esize, int ncols,
                                            const void* const* src, void* const* dst, const KeySrc& kspec,
                                            uint32_t start, uint32_t dstart, uint32_t len, bool in_place) {
  uint64_t* skey = (uint64_t*)smem_raw;                 // LEAF_MAX
  uint16_t* sidx = (uint16_t*)(skey + LEAF_MAX);        // LEAF_MAX
  uint32_t* cnt = (uint32_t*)(sidx + LEAF_MAX);         // PT_WARPS * 256
  uint32_t* tot = cnt + PT_WARPS * MSD_RADIX;
  uint32_t* loc = tot + MSD_RADIX;
  uint32_t* scratch = loc + MSD_RADIX;                  // PT_SCRATCH
  uint64_t* red = (uint64_t*)(scratch + PT_SCRATCH);    // 2 * PT_WARPS + 2

  const int w = threadIdx.x >> 5;
  const uint32_t chunk = ((len + PT_THREADS - 1) / PT_THREADS) * 32;  // elements per warp, multiple of 32
  const int items = (int)(chunk >> 5);
  const uint32_t base = w * chunk + lane_id();

  uint64_t key[PT_ITEMS];
  uint32_t lidx[PT_ITEMS], digit[PT_ITEMS], pos[PT_ITEMS];
  uint32_t vmask = 0;
  uint64_t mn = ~0ull, mx = 0;
#pragma unroll
  for (int k = 0; k < PT_ITEMS; k++) {
    uint32_t idx = base + k * 32;
    bool v = k < items && idx < len;
    vmask |= (uint32_t)v << k;
    lidx[k] = idx;
    key[k] = 0;
    if (v) {
      key[k] = load_key_at(src[kspec.col], kspec, start + idx);
      mn = key[k] < mn ? key[k] : mn;
      mx = key[k] > mx ? key[k] : mx;
    }
  }
  mn = warp_min_u64(mn);
  mx = warp_max_u64(mx);
  if (lane_id() == 0) { red[w] = mn; red[PT_WARPS + w] = mx; }
  __syncthreads();
  if (threadIdx.x == 0) {
    uint64_t a = red[0], b = red[PT_WARPS];
    for (int i = 1; i < PT_WARPS; i++) {
      a = red[i] < a ? red[i] : a;
      b = red[PT_WARPS + i] > b ? red[PT_WARPS + i] : b;
    }
    red[2 * PT_WARPS] = a;
    red[2 * PT_WARPS + 1] = b;
  }
  __syncthreads();
  const uint64_t kmin = red[2 * PT_WARPS];
  const uint64_t range = red[2 * PT_WARPS + 1] - kmin;
  const int nbits = range ? 64 - __clzll((long long)range) : 0;

  if (nbits == 0) {
    if (in_place) return;
    for (uint32_t j = threadIdx.x; j < len; j += PT_THREADS) sidx[j] = (uint16_t)j;
  }
  for (int lo = 0; lo < nbits; lo += MSD_BITS) {
    int bits = nbits - lo < MSD_BITS ? nbits - lo : MSD_BITS;
#pragma unroll
    for (int k = 0; k < PT_ITEMS; k++) digit[k] = (uint32_t)((key[k] - kmin) >> lo) & ((1u << bits) - 1u);
    tile_rank<PT_ITEMS>(digit, vmask, items, bits, MSD_RADIX, cnt, tot, loc, scratch, pos);
#pragma unroll
    for (int k = 0; k < PT_ITEMS; k++)
      if ((vmask >> k) & 1u) { skey[pos[k]] = key[k]; sidx[pos[k]] = (uint16_t)lidx[k]; }
    __syncthreads();
    if (lo + MSD_BITS < nbits) {
#pragma unroll
      for (int k = 0; k < PT_ITEMS; k++)
        if ((vmask >> k) & 1u) { key[k] = skey[base + k * 32]; lidx[k] = sidx[base + k * 32]; }
      __syncthreads();
    }
  }
  __syncthreads();
  // move every column: read all (registers), then write, so that src == dst is safe
  for (int c = 0; c < ncols; c++) {
    const int es = esize[c];
    uint64_t v[PT_ITEMS];
#pragma unroll
    for (int k = 0; k < PT_ITEMS; k++) {
      uint32_t j = threadIdx.x + k * PT_THREADS;
      if (j < len) {
        uint32_t r = start + sidx[j];
        if (es == 8) v[k] = ((const uint64_t*)src[c])[r];
        else if (es == 4) v[k] = ((const uint32_t*)src[c])[r];
        else if (es == 2) v[k] = ((const uint16_t*)src[c])[r];
        else v[k] = ((const uint8_t*)src[c])[r];
      }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < PT_ITEMS; k++) {
      uint32_t j = threadIdx.x + k * PT_THREADS;
      if (j < len) {
        uint32_t r = dstart + j;
        if (es == 8) ((uint64_t*)dst[c])[r] = v[k];
        else if (es == 4) ((uint32_t*)dst[c])[r] = (uint32_t)v[k];
        else if (es == 2) ((uint16_t*)dst[c])[r] = (uint16_t)v[k];
        else ((uint8_t*)dst[c])[r] = (uint8_t)v[k];
      }
    }
  }
}

__global__ void __launch_bounds__(PT_THREADS, 2 * PT_MIN_CTAS / 2) k_leaf_sort(const __grid_constant__ LeafParams p) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const uint32_t s = blockIdx.x;
  if (s >= *p.count) return;
  if (p.wide && !p.wide[s]) return;  // already sorted by the packed 32-bit kernel
  const uint32_t start = p.segs.start[s], len = p.segs.len[s];
  leaf64_body(smem_raw, p.mv.esize, p.mv.n, p.mv.src, p.mv.dst, p.key, start, start, len, p.in_place != 0);
}

#define L32_IDX_BITS 12
#define L32_MAX_KEY_BITS 20
#define L32_ATOMIC_BITS 11   // the atomic pass takes min(nbits, 11) bits, or 12 when 20 bits vary

template <typename T>
__device__ __forceinline__ void leaf_move_column(const T* __restrict__ src, T* __restrict__ dst, uint32_t start,
                                                 uint32_t len, const uint32_t* sdata) {
  T v[PT_ITEMS];
#pragma unroll
  for (int k = 0; k < PT_ITEMS; k++) {
    uint32_t j = threadIdx.x + k * PT_THREADS;
    if (j < len) v[k] = src[start + (sdata[j] & ((1u << L32_IDX_BITS) - 1u))];
  }
  __syncthreads();  // every read of the segment precedes every write: src == dst is safe
#pragma unroll
  for (int k = 0; k < PT_ITEMS; k++) {
    uint32_t j = threadIdx.x + k * PT_THREADS;
    if (j < len) dst[start + j] = v[k];
  }
}

// The leaf sort for segments whose keys span at most 2^20 values (the common case after one MSD level):
// key - min and the row's position are packed into one 32-bit word, the first LSD pass (up to 11 bits)
// ranks with shared-memory atomics (no stability needed), the remaining <= 9 bits take one stable
// ballot-ranked pass. Segments with a wider key range are flagged for k_leaf_sort.
template <typename T, int ITEMS>
__device__ __forceinline__ void leaf_permute_column(const unsigned char* stage, uint32_t lead, T* __restrict__ dst,
                                                    uint32_t start, uint32_t len, const uint32_t* sdata) {
  const T* col = (const T*)(stage + lead);
#pragma unroll 4
  for (int k = 0; k < ITEMS; k++) {
    uint32_t j = threadIdx.x + k * PT_THREADS;
    if (j < len) dst[start + j] = col[sdata[j] & ((1u << L32_IDX_BITS) - 1u)];
  }
}

// The leaf sort for segments whose keys span at most 2^20 values (the common case after one MSD level).
// Every column slice of the segment arrives in shared memory through a TMA bulk copy (double buffered, the
// key column first); key - min and the row's position are packed into one 32-bit word; the first LSD pass
// (up to 11-12 bits) ranks with shared-memory atomics (no stability needed), the remaining <= 8 bits take
// one stable ballot-ranked pass; each column is then read through the permutation from shared memory and
// written out coalesced. Because a slice is completely in shared memory before the first row of that
// column is written, sorting in place (src == dst) is safe. Segments with a wider key range are flagged
// for k_leaf_sort.
// Returns 0 when the segment has been written sorted to dst, 1 when it holds a run of more than 64 rows with
// equal 20-bit prefixes (nothing written: the 64-bit kernel has to take it).
// ITEMS: rows per thread the code is unrolled for (len <= ITEMS * PT_THREADS): a half-full leaf takes the 4-row
// variant and issues half the instructions.
template <int KMODE, int ITEMS>
__device__ __forceinline__ int leaf32_body(unsigned char* smem_raw, const int32_t* esize, int ncols_,
                                           const void* const* src, void* const* dst, const KeySrc& kspec,
                                           uint32_t start, uint32_t dstart, uint32_t len, uint32_t seg_bits_,
                                           uint64_t seg_base_, bool in_place, uint64_t* ext_bar = nullptr,
                                           uint32_t* ext_phase = nullptr, long long* tr = nullptr) {
  // tr (diagnostics, HSGPU_LEAF_TRACE): SM clock at the phase boundaries of this segment, written by thread 0
  auto mark = [&](int i) { if (tr && threadIdx.x == 0) tr[i] = clock64(); };
  mark(0);
  // landing areas of the column slices: two of a full tile, or — for a leaf of at most half a tile — four of
  // half a tile, so that up to four columns are requested at once, before the sort starts
  constexpr int NS = ITEMS * PT_THREADS * 2 <= PT_TILE ? 4 : 2;
  constexpr uint32_t SLOT_BYTES = 2 * STAGE_BYTES / NS;
  auto stage_of = [&](int b) { return smem_raw + (size_t)b * SLOT_BYTES; };
  unsigned char* stage0 = smem_raw;
  uint32_t* sdata = (uint32_t*)(smem_raw + 2 * STAGE_BYTES);  // LEAF_MAX packed words
  uint32_t* cnt = sdata + LEAF_MAX;                 // 4096 words: atomic counters or PT_WARPS x 256 warp counters
  uint32_t* tot = cnt + 4096;                       // 256
  uint32_t* loc = tot + 256;                        // 256
  uint32_t* scratch = loc + 256;                    // PT_SCRATCH
  uint64_t* red = (uint64_t*)(scratch + PT_SCRATCH);  // 2 * PT_WARPS + 2
  uint64_t* bar = ext_bar ? ext_bar : red + 2 * PT_WARPS + 2;   // 4 mbarriers (see partition_tile for ext_*)
  const uint32_t ph_in = ext_phase ? *ext_phase : 0u;
  // barrier b completes once for each of the columns b, b + NS, ...; the first min(ncols, NS) slices are
  // requested whatever happens next
  auto phase_after = [&](int cols_done) {
    uint32_t ph = ph_in;
    for (int b = 0; b < NS; b++) ph ^= (uint32_t)(((cols_done > b ? (cols_done - b + NS - 1) / NS : 0) & 1) << b);
    return ph;
  };
  if (ext_phase) *ext_phase = phase_after(ncols_ < NS ? ncols_ : NS);

  const int w = threadIdx.x >> 5;
  const int items = (int)((len + PT_THREADS - 1) / PT_THREADS);
  const int ncols = ncols_;
  const int first = kspec.col;
  auto col_at = [&](int i) { return i == 0 ? first : (i <= first ? i - 1 : i); };

  if (!ext_bar) {
    if (threadIdx.x == 0) {
      for (int b = 0; b < 4; b++) mbar_init(&bar[b], 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();
  }
  const int n_first = ncols < NS ? ncols : NS;   // slices requested up front
  if (threadIdx.x == 0) {
    for (int i = 0; i < n_first; i++) {
      const int c = col_at(i);
      issue_slice(src[c], esize[c], start, len, stage_of(i), &bar[i]);
    }
  }
  // a CTA must not retire (or move on) with a bulk copy in flight
  auto drain_first = [&]() {
    for (int b = 1; b < n_first; b++) mbar_wait(&bar[b], (ph_in >> b) & 1u);
  };
  // ---- keys from the staged slice. The segment descriptor says how many low bits of (key - base) can still
  // differ inside the segment; when that fits the packed word no min/max pass is needed at all.
  const uint32_t seg_bits = seg_bits_;
  const uint64_t seg_base = seg_base_;
  const bool known = seg_bits <= L32_MAX_KEY_BITS;
  const uint64_t seg_mask = known ? ((1ull << seg_bits) - 1ull) : ~0ull;
  mbar_wait(&bar[0], ph_in & 1u);
  mark(1);
  const unsigned char* ks = stage0 + slice_lead(src[first], esize[first], start);
  uint64_t key[ITEMS];
  uint32_t vmask = 0;
  uint64_t mn = ~0ull, mx = 0;
#pragma unroll
  for (int k = 0; k < ITEMS; k++) {
    uint32_t idx = threadIdx.x + k * PT_THREADS;
    bool v = idx < len;
    vmask |= (uint32_t)v << k;
    key[k] = 0;
    if (v) {
      if (KMODE == KS_U64) key[k] = ((const uint64_t*)ks)[idx];
      else if (KMODE == KS_I64) key[k] = (((const uint64_t*)ks)[idx] ^ 0x8000000000000000ull) - kspec.min;
      else key[k] = ((uint64_t)(int64_t)((const int32_t*)ks)[idx] ^ 0x8000000000000000ull) - kspec.min;
      key[k] = (key[k] - seg_base) & seg_mask;
      if (!known) {
        mn = key[k] < mn ? key[k] : mn;
        mx = key[k] > mx ? key[k] : mx;
      }
    }
  }
  uint64_t kmin = 0;
  int nbits = (int)seg_bits;
  if (!known) {
    mn = warp_min_u64(mn);
    mx = warp_max_u64(mx);
    if (lane_id() == 0) { red[w] = mn; red[PT_WARPS + w] = mx; }
    __syncthreads();
    if (w == 0) {
      uint64_t a = lane_id() < PT_WARPS ? red[lane_id()] : ~0ull;
      uint64_t b = lane_id() < PT_WARPS ? red[PT_WARPS + lane_id()] : 0;
      a = warp_min_u64(a);
      b = warp_max_u64(b);
      if (lane_id() == 0) { red[2 * PT_WARPS] = a; red[2 * PT_WARPS + 1] = b; }
    }
    __syncthreads();
    kmin = red[2 * PT_WARPS];
    const uint64_t range = red[2 * PT_WARPS + 1] - kmin;
    nbits = range ? 64 - __clzll((long long)range) : 0;
  }
  // A key range wider than the packed field: sort on the top L32_MAX_KEY_BITS bits of (key - kmin) and repair
  // the (rare) runs of equal prefixes afterwards by comparing full keys (see the fix-up below).
  const int drop = nbits > L32_MAX_KEY_BITS ? nbits - L32_MAX_KEY_BITS : 0;
  if (nbits == 0 && in_place) {
    // all keys equal and the rows are in place; a CTA must not retire with a bulk copy in flight
    drain_first();
    return 0;
  }
  if (drop) nbits = L32_MAX_KEY_BITS;

  uint32_t word[ITEMS], pos[ITEMS];
#pragma unroll
  for (int k = 0; k < ITEMS; k++)
    word[k] = ((uint32_t)((key[k] - kmin) >> drop) << L32_IDX_BITS) | (threadIdx.x + k * PT_THREADS);

  if (nbits == 0) {
#pragma unroll
    for (int k = 0; k < ITEMS; k++)
      if ((vmask >> k) & 1u) sdata[threadIdx.x + k * PT_THREADS] = word[k];
    __syncthreads();
  } else {
    // ---- pass A: low bits, atomic ranks
    int ba = nbits < L32_ATOMIC_BITS ? nbits : L32_ATOMIC_BITS;
    if (nbits - ba > 8) ba = nbits - 8;  // the stable pass handles at most 8 bits
    const int radix_a = 1 << ba;
    tile_rank_atomic_w<ITEMS>(word, L32_IDX_BITS, (1u << ba) - 1u, vmask, items, radix_a, cnt, scratch, pos);
#pragma unroll
    for (int k = 0; k < ITEMS; k++)
      if ((vmask >> k) & 1u) sdata[pos[k]] = word[k];
    __syncthreads();
    mark(2);
    // ---- pass B: the remaining high bits, stable
    const int bb = nbits - ba;
    if (bb > 0) {
      const uint32_t chunk = (uint32_t)items * 32;
      const uint32_t base = w * chunk + lane_id();
      uint32_t vm2 = 0;
#pragma unroll
      for (int k = 0; k < ITEMS; k++) {
        uint32_t idx = base + k * 32;
        bool v = k < items && idx < len;
        vm2 |= (uint32_t)v << k;
        word[k] = v ? sdata[idx] : 0;
      }
      tile_rank_w<ITEMS>(word, L32_IDX_BITS + ba, vm2, items, bb, 256, cnt, tot, loc, scratch, pos);
#pragma unroll
      for (int k = 0; k < ITEMS; k++)
        if ((vm2 >> k) & 1u) sdata[pos[k]] = word[k];
      __syncthreads();
    }
  }
  mark(3);
  // ---- prefix ties: rows whose dropped low bits differ may be out of order inside a run of equal prefixes.
  // Every row of such a run ranks itself inside the run by the full key (the key slice is still staged).
  // Runs are short for keys that are not clustered; a run longer than 64 rows sends the segment to the
  // 64-bit kernel instead.
  bool too_long = false;
  if (drop) {
    uint32_t npos[ITEMS];
#pragma unroll
    for (int k = 0; k < ITEMS; k++) {
      const uint32_t j = threadIdx.x + k * PT_THREADS;
      npos[k] = j;
      if (j < len) {
        const uint32_t wj = sdata[j], pj = wj >> L32_IDX_BITS;
        uint32_t a = j, e = j + 1;
        while (a > 0 && (sdata[a - 1] >> L32_IDX_BITS) == pj && j - a < 65) a--;
        while (e < len && (sdata[e] >> L32_IDX_BITS) == pj && e - j < 65) e++;
        if (e - a > 64) too_long = true;
        else if (e - a > 1) {
          auto full = [&](uint32_t w) -> uint64_t {
            const uint32_t r = w & ((1u << L32_IDX_BITS) - 1u);
            if (KMODE == KS_U64) return ((const uint64_t*)ks)[r];
            if (KMODE == KS_I64) return ((const uint64_t*)ks)[r] ^ 0x8000000000000000ull;
            return (uint64_t)(int64_t)((const int32_t*)ks)[r] ^ 0x8000000000000000ull;
          };
          const uint64_t kj = full(wj);
          uint32_t r = 0;
          for (uint32_t i = a; i < e; i++) {
            const uint64_t ki = full(sdata[i]);
            r += (ki < kj) || (ki == kj && i < j);
          }
          npos[k] = a + r;
        }
      }
    }
    too_long = __syncthreads_or(too_long);
    if (!too_long) {
      uint32_t wv[ITEMS];
#pragma unroll
      for (int k = 0; k < ITEMS; k++) {
        const uint32_t j = threadIdx.x + k * PT_THREADS;
        wv[k] = j < len ? sdata[j] : 0;
      }
      __syncthreads();
#pragma unroll
      for (int k = 0; k < ITEMS; k++) {
        const uint32_t j = threadIdx.x + k * PT_THREADS;
        if (j < len) sdata[npos[k]] = wv[k];
      }
      __syncthreads();
    }
  }
  if (too_long) {
    drain_first();
    return 1;
  }
  // ---- every column: staged slice -> permutation -> coalesced store
  for (int i = 0; i < ncols; i++) {
    const int c = col_at(i), b = i % NS, es = esize[c];
    unsigned char* stage = stage_of(b);
    mbar_wait(&bar[b], ((uint32_t)(i / NS) ^ (ph_in >> b)) & 1u);
    const uint32_t lead = slice_lead(src[c], es, start);
    switch (es) {
      case 8: leaf_permute_column<uint64_t, ITEMS>(stage, lead, (uint64_t*)dst[c], dstart, len, sdata); break;
      case 4: leaf_permute_column<uint32_t, ITEMS>(stage, lead, (uint32_t*)dst[c], dstart, len, sdata); break;
      case 2: leaf_permute_column<uint16_t, ITEMS>(stage, lead, (uint16_t*)dst[c], dstart, len, sdata); break;
      default: leaf_permute_column<uint8_t, ITEMS>(stage, lead, (uint8_t*)dst[c], dstart, len, sdata); break;
    }
    mark(4 + (i < 3 ? i : 3));
    if (i + NS < ncols) {
      __syncthreads();  // the slot is free again
      if (threadIdx.x == 0) {
        const int cn = col_at(i + NS);
        issue_slice(src[cn], esize[cn], start, len, stage, &bar[b]);
      }
    }
  }
  __syncthreads();
  mark(8);
  if (ext_phase) *ext_phase = phase_after(ncols);
  return 0;
}

template <int KMODE>
__global__ void __launch_bounds__(PT_THREADS, 2 * PT_MIN_CTAS / 2) k_leaf_sort32(const __grid_constant__ LeafParams p) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const uint32_t s = blockIdx.x;
  if (s >= *p.count) return;
  const uint32_t start = p.segs.start[s], len = p.segs.len[s];
  long long* tr = p.trace ? p.trace + 16 * (size_t)s : nullptr;
  if (tr && threadIdx.x == 0) { tr[9] = len; uint32_t sm; asm volatile("mov.u32 %0, %%smid;" : "=r"(sm)); tr[10] = sm; }
  const int r = len <= 4 * PT_THREADS
      ? leaf32_body<KMODE, 4>(smem_raw, p.mv.esize, p.mv.n, p.mv.src, p.mv.dst, p.key, start, start, len,
                              p.segs.shift[s], p.segs.base[s], p.in_place != 0, nullptr, nullptr, tr)
      : leaf32_body<KMODE, PT_ITEMS>(smem_raw, p.mv.esize, p.mv.n, p.mv.src, p.mv.dst, p.key, start, start, len,
                                     p.segs.shift[s], p.segs.base[s], p.in_place != 0, nullptr, nullptr, tr);
  if (threadIdx.x == 0) {
    p.wide[s] = r ? 1 : 0;
    if (r) atomicAdd(p.n_wide, 1u);
  }
}
static size_t leaf32_smem_bytes() {
  return (size_t)2 * STAGE_BYTES + (size_t)LEAF_MAX * 4 + 4096 * 4 + 512 * 4 + PT_SCRATCH * 4 +
         (2 * PT_WARPS + 2) * 8 + 32;
}

static size_t leaf_smem_bytes() {
  return (size_t)LEAF_MAX * 8 + (size_t)LEAF_MAX * 2 + (size_t)(PT_WARPS + 2) * MSD_RADIX * 4 + PT_SCRATCH * 4 +
         (2 * PT_WARPS + 2) * 8;
}

// Tiny segments (tie groups of a string sort, buckets of a small table): one warp sorts one segment of at
// most SMALL_MAX rows by counting, for every row, the rows that precede it (key, then position).
__device__ __forceinline__ void leaf_small_warp(const int32_t* esize, int ncols, const void* const* src,
                                                void* const* dst, const KeySrc& kspec, uint32_t start,
                                                uint32_t dstart, uint32_t len) {
  const uint32_t lane = lane_id();
  const uint32_t i0 = lane, i1 = lane + 32;
  uint64_t k0 = i0 < len ? load_key_at(src[kspec.col], kspec, start + i0) : ~0ull;
  uint64_t k1 = i1 < len ? load_key_at(src[kspec.col], kspec, start + i1) : ~0ull;
  uint32_t r0 = 0, r1 = 0;
  for (uint32_t j = 0; j < len; j++) {
    uint64_t kj = __shfl_sync(0xffffffffu, j < 32 ? k0 : k1, j & 31);
    r0 += (kj < k0) || (kj == k0 && j < i0);
    r1 += (kj < k1) || (kj == k1 && j < i1);
  }
  // four columns at a time: all loads are issued before the first store (one memory round trip per group of
  // columns instead of one per column; also what makes src == dst safe)
  auto ld = [](const void* base, size_t i, int es) -> uint64_t {
    switch (es) {
      case 8: return ((const uint64_t*)base)[i];
      case 4: return ((const uint32_t*)base)[i];
      case 2: return ((const uint16_t*)base)[i];
      default: return ((const uint8_t*)base)[i];
    }
  };
  auto st = [](void* base, size_t i, int es, uint64_t v) {
    switch (es) {
      case 8: ((uint64_t*)base)[i] = v; break;
      case 4: ((uint32_t*)base)[i] = (uint32_t)v; break;
      case 2: ((uint16_t*)base)[i] = (uint16_t)v; break;
      default: ((uint8_t*)base)[i] = (uint8_t)v; break;
    }
  };
  for (int c0 = 0; c0 < ncols; c0 += 4) {
    uint64_t v0[4], v1[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const int c = c0 + u;
      v0[u] = v1[u] = 0;
      if (c < ncols) {
        if (i0 < len) v0[u] = ld(src[c], (size_t)start + i0, esize[c]);
        if (i1 < len) v1[u] = ld(src[c], (size_t)start + i1, esize[c]);
      }
    }
    __syncwarp();
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const int c = c0 + u;
      if (c < ncols) {
        if (i0 < len) st(dst[c], (size_t)dstart + r0, esize[c], v0[u]);
        if (i1 < len) st(dst[c], (size_t)dstart + r1, esize[c], v1[u]);
      }
    }
    __syncwarp();
  }
}

__global__ void __launch_bounds__(PT_THREADS) k_leaf_small(const __grid_constant__ LeafParams p) {
  const uint32_t s = blockIdx.x * PT_WARPS + (threadIdx.x >> 5);
  if (s >= *p.count) return;
  const uint32_t start = p.segs.start[s], len = p.segs.len[s];
  leaf_small_warp(p.mv.esize, p.mv.n, p.mv.src, p.mv.dst, p.key, start, start, len);
}

// segments whose keys are all equal but that are larger than a leaf: plain copy src -> dst. Such a segment can
// be a whole bucket (a key column with a handful of distinct values), so COPY_SPLIT CTAs share each segment and
// move 16 bytes per thread where the alignment allows (one CTA per segment with byte moves took 143 ms for ten
// segments of 5 M rows).
#define COPY_SPLIT 128
__global__ void __launch_bounds__(256) k_copy_segments(MoveArgs mv, SegList segs, const uint32_t* count) {
  const uint32_t s = blockIdx.x;
  if (s >= *count) return;
  const uint32_t start = segs.start[s], len = segs.len[s];
  // this CTA's rows: a multiple of 16 rows per slice keeps every column's slice 16-byte aligned relative to the
  // segment start
  uint32_t per = (len + COPY_SPLIT - 1) / COPY_SPLIT;
  per = (per + 15u) & ~15u;
  const uint32_t r0 = blockIdx.y * per;
  if (r0 >= len) return;
  const uint32_t rows = len - r0 < per ? len - r0 : per;
  for (int c = 0; c < mv.n; c++) {
    const int es = mv.esize[c];
    const uint8_t* src = (const uint8_t*)mv.src[c] + (size_t)(start + r0) * es;
    uint8_t* dst = (uint8_t*)mv.dst[c] + (size_t)(start + r0) * es;
    const size_t bytes = (size_t)rows * es;
    if ((((uintptr_t)src | (uintptr_t)dst) & 15) == 0) {
      const size_t nv = bytes >> 4;
      for (size_t i = threadIdx.x; i < nv; i += blockDim.x) ((uint4*)dst)[i] = ((const uint4*)src)[i];
      for (size_t i = (nv << 4) + threadIdx.x; i < bytes; i += blockDim.x) dst[i] = src[i];
    } else if ((((uintptr_t)src | (uintptr_t)dst) & 3) == 0 && (bytes & 3) == 0) {
      for (size_t i = threadIdx.x; i < (bytes >> 2); i += blockDim.x) ((uint32_t*)dst)[i] = ((const uint32_t*)src)[i];
    } else {
      for (size_t i = threadIdx.x; i < bytes; i += blockDim.x) dst[i] = src[i];
    }
  }
}

// ---------------------------------------------------------------- host drivers
#define BUF_A 0
#define BUF_B 1
#define BUF_S0 2
static void* const* buf_of(const MoveSet& ms, int code) { return code == BUF_A ? ms.a : code == BUF_B ? ms.b : ms.s0; }
static int other_buf(int code) { return code == BUF_B ? BUF_A : BUF_B; }
static MoveArgs make_move(const MoveSet& ms, int src_code, int dst_code) {
  MoveArgs mv;
  memset(&mv, 0, sizeof mv);
  mv.n = ms.n;
  for (int i = 0; i < ms.n; i++) {
    mv.esize[i] = ms.esize[i];
    mv.src[i] = buf_of(ms, src_code)[i];
    mv.dst[i] = buf_of(ms, dst_code)[i];
  }
  return mv;
}

static int ilog2_ceil(uint32_t v) {
  int b = 0;
  while ((1u << b) < v) b++;
  return b;
}

__global__ void k_coarse_offsets(const uint32_t* __restrict__ bucket_off, int nb, int shift, int ncoarse,
                                 uint32_t* __restrict__ out) {
  int cidx = blockIdx.x * blockDim.x + threadIdx.x;
  if (cidx < ncoarse) {
    int b = cidx << shift;
    out[cidx] = bucket_off[b < nb ? b : nb];
  }
}

static void launch_bucket_scatter(hs_ctx* c, const PartParams& p, int64_t n) {
  size_t smem = tile_smem_bytes(p.radix_pad);
  static size_t attr_set_dev[64] = {0};
  int dev = 0;
  CK(cudaGetDevice(&dev));
  size_t& attr_set = attr_set_dev[dev & 63];
  if (smem > attr_set) {
    CK(cudaFuncSetAttribute(k_partition<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_set = smem;
  }
  unsigned tiles = (unsigned)((n + PT_TILE - 1) / PT_TILE);
  LAUNCH(c, k_partition<0>, tiles, PT_THREADS, smem, p);
}

void hs_bucket_scatter(hs_ctx* c, const void* const* src, const MoveSet& ms, const uint16_t* d_bucket, int64_t n,
                       int32_t nb, const uint32_t* d_bucket_off, int32_t bucket_base, int widen_col, uint64_t widen_min,
                       uint32_t* d_cursor) {
  // d_cursor: the next free row of every bucket, kept by the caller across several calls that fill the same
  // buckets from different sources (else the cursors start at d_bucket_off)
  if (n == 0) return;
  if (d_cursor && nb > 2048) HS_THROW(HS_EINVAL, "a caller-kept cursor cannot pass the two-level scatter");
  if (widen_col >= 0 && nb > 2048) HS_THROW(HS_EINVAL, "a narrowed key cannot pass the two-level scatter");
  if (nb > 8192) HS_THROW(HS_EUNSUPPORTED, "the bucket scatter supports at most 8192 buckets per GPU");
  PartParams p;
  memset(&p, 0, sizeof p);
  p.n = (uint32_t)n;
  p.bucket_base = bucket_base;
  DBuf cursor;
  if (!d_cursor) {
    cursor.alloc(c, (size_t)nb * 4);
    CK(cudaMemcpyAsync(cursor.p, d_bucket_off, (size_t)nb * 4, cudaMemcpyDeviceToDevice, c->stream));
  }
  const char* one = getenv("HSGPU_ONE_LEVEL_SCATTER");
  if (nb > 2048 && n >= ((int64_t)1 << 22) && ms.n + 1 <= HS_MAX_MOVE && !(one && *one && *one != '0')) {
    // Thousands of buckets: a tile of 4096 rows holds less than two rows per bucket, so a direct scatter writes
    // every row on its own (measured 4.7 ms per 100 M rows of 28 bytes at 8192 buckets, against 1.1 ms at 200).
    // Two passes with long runs are cheaper: by the top bits of the bucket id (<= 128 bins, the ids travel with
    // the rows), then by the full id — the rows of a tile then fall into 2^shift neighbouring buckets only.
    const int shift = ilog2_ceil((uint32_t)nb) - 7;
    const int ncoarse = (nb + (1 << shift) - 1) >> shift;
    std::vector<DBuf> tmp((size_t)ms.n + 1);
    for (int i = 0; i < ms.n; i++) tmp[i].alloc(c, (size_t)n * ms.esize[i] + 16);
    tmp[ms.n].alloc(c, (size_t)n * 2 + 16);
    DBuf ccur(c, (size_t)ncoarse * 4);
    LAUNCH(c, k_coarse_offsets, (ncoarse + 127) / 128, 128, 0, d_bucket_off, nb, shift, ncoarse, ccur.as<uint32_t>());
    p.mv.n = ms.n + 1;
    for (int i = 0; i < ms.n; i++) { p.mv.esize[i] = ms.esize[i]; p.mv.src[i] = src[i]; p.mv.dst[i] = tmp[i].p; }
    p.mv.esize[ms.n] = 2; p.mv.src[ms.n] = d_bucket; p.mv.dst[ms.n] = tmp[ms.n].p;
    p.bucket = d_bucket;
    p.bucket_shift = shift;
    p.radix = ncoarse;
    p.radix_pad = ((ncoarse + PT_THREADS - 1) / PT_THREADS) * PT_THREADS;
    p.bits = ilog2_ceil((uint32_t)ncoarse);
    p.cursor = ccur.as<uint32_t>();
    launch_bucket_scatter(c, p, n);
    p.mv.n = ms.n;
    for (int i = 0; i < ms.n; i++) { p.mv.src[i] = tmp[i].p; p.mv.dst[i] = ms.a[i]; }
    p.bucket = tmp[ms.n].as<uint16_t>();
    p.bucket_shift = 0;
  } else {
    p.mv.n = ms.n;
    for (int i = 0; i < ms.n; i++) { p.mv.esize[i] = ms.esize[i]; p.mv.src[i] = src[i]; p.mv.dst[i] = ms.a[i]; }
    if (widen_col >= 0) {
      p.mv.esize[widen_col] = 4;          // what the tile stages; the store widens to 8 bytes
      p.conv_col1 = widen_col + 1;
      p.conv_min = widen_min;
    }
    p.bucket = d_bucket;
  }
  p.radix = nb;
  p.radix_pad = ((nb + PT_THREADS - 1) / PT_THREADS) * PT_THREADS;
  p.bits = ilog2_ceil((uint32_t)nb);
  p.cursor = d_cursor ? d_cursor : cursor.as<uint32_t>();
  launch_bucket_scatter(c, p, n);
}

// Rows grouped by bins that are unions of buckets (d_lut[bucket id] = bin; d_bin_off[bin] = first row of the bin
// in dst): what a sender does before whole spans of rows travel to their owners (hs_dist.cu, grouped spans).
void hs_scatter_by_bins(hs_ctx* c, const void* const* src, void* const* dst, const int32_t* esize, int ncols,
                        const uint16_t* d_bucket, int64_t n, const uint16_t* d_lut, int32_t nbins, const uint32_t* d_bin_off) {
  if (n == 0) return;
  if (ncols > HS_MAX_MOVE) HS_THROW(HS_EUNSUPPORTED, "too many columns for the grouped scatter");
  PartParams p;
  memset(&p, 0, sizeof p);
  p.n = (uint32_t)n;
  p.mv.n = ncols;
  for (int i = 0; i < ncols; i++) { p.mv.esize[i] = esize[i]; p.mv.src[i] = src[i]; p.mv.dst[i] = dst[i]; }
  p.bucket = d_bucket;
  p.bin_lut = d_lut;
  p.radix = nbins;
  p.radix_pad = ((nbins + PT_THREADS - 1) / PT_THREADS) * PT_THREADS;
  p.bits = ilog2_ceil((uint32_t)nbins);
  DBuf cursor(c, (size_t)nbins * 4);
  CK(cudaMemcpyAsync(cursor.p, d_bin_off, (size_t)nbins * 4, cudaMemcpyDeviceToDevice, c->stream));
  p.cursor = cursor.as<uint32_t>();
  launch_bucket_scatter(c, p, n);
}

// Multi-GPU variant of the bucket scatter: rows go straight into the owners' receive buffers.
// d_run_base[b] = first row, inside the owner's buffer, of THIS rank's run for bucket b.
void hs_bucket_scatter_peers(hs_ctx* c, const void* const* src, const int32_t* esize, int ncols, void* const* pdst,
                             int nranks, const uint16_t* d_bucket, int64_t n, int32_t nb, const uint32_t* d_run_base,
                             bool by_owner, int narrow_col, uint64_t narrow_min) {
  if (n == 0) return;
  if (nb > 8192) HS_THROW(HS_EUNSUPPORTED, "the bucket scatter supports at most 8192 buckets");
  if (ncols > HS_MAX_PEER_COLS || nranks > HS_MAX_PEERS) HS_THROW(HS_EUNSUPPORTED, "too many columns or ranks for the peer scatter");
  PartParams p;
  memset(&p, 0, sizeof p);
  p.mv.n = ncols;
  for (int i = 0; i < ncols; i++) { p.mv.esize[i] = esize[i]; p.mv.src[i] = src[i]; p.mv.dst[i] = nullptr; }
  for (int r = 0; r < nranks; r++)
    for (int i = 0; i < ncols; i++) p.pdst[r][i] = pdst[r * ncols + i];
  p.nranks = nranks;
  p.bucket = d_bucket;
  p.n = (uint32_t)n;
  // by_owner: rows are only grouped by the GPU that owns their bucket (nranks bins, d_run_base has nranks entries):
  // a tile then holds runs of hundreds of rows per destination, which is what NVLink stores want
  p.by_owner = by_owner ? 1 : 0;
  p.nb_total = nb;
  if (narrow_col >= 0) {
    p.mv.esize[narrow_col] = 8;           // the source is the int64 column; the store narrows it to 4 bytes
    p.conv_col1 = narrow_col + 1;
    p.conv_min = narrow_min;
  }
  const int32_t radix = by_owner ? nranks : nb;
  p.radix = radix;
  p.radix_pad = ((radix + PT_THREADS - 1) / PT_THREADS) * PT_THREADS;
  p.bits = 0;
  DBuf cursor(c, (size_t)radix * 4);
  CK(cudaMemcpyAsync(cursor.p, d_run_base, (size_t)radix * 4, cudaMemcpyDeviceToDevice, c->stream));
  p.cursor = cursor.as<uint32_t>();
  size_t smem = tile_smem_bytes(p.radix_pad);
  static size_t attr_set_dev[64] = {0};
  int dev = 0;
  CK(cudaGetDevice(&dev));
  size_t& attr_set = attr_set_dev[dev & 63];
  if (smem > attr_set) {
    CK(cudaFuncSetAttribute(k_partition<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_set = smem;
  }
  unsigned tiles = (unsigned)((n + PT_TILE - 1) / PT_TILE);
  LAUNCH(c, k_partition<2>, tiles, PT_THREADS, smem, p);
}

struct ListBufs {
  DBuf start, len, shift, base;
  void alloc(hs_ctx* c, size_t cap, bool with_base = false) {
    start.alloc(c, cap * 4); len.alloc(c, cap * 4); shift.alloc(c, cap * 4);
    if (with_base) { base.alloc(c, cap * 8); base.zero(); }
  }
  SegList view() const {
    return SegList{start.as<uint32_t>(), len.as<uint32_t>(), shift.as<uint32_t>(), base.as<uint64_t>()};
  }
};

static void set_smem_attrs_once() {
  static bool done_dev[64] = {false};
  int dev = 0;
  CK(cudaGetDevice(&dev));
  bool& done = done_dev[dev & 63];
  if (done) return;
  CK(cudaFuncSetAttribute(k_leaf_sort32<KS_U64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)leaf32_smem_bytes()));
  CK(cudaFuncSetAttribute(k_leaf_sort32<KS_I64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)leaf32_smem_bytes()));
  CK(cudaFuncSetAttribute(k_leaf_sort32<KS_I32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)leaf32_smem_bytes()));
  CK(cudaFuncSetAttribute(k_leaf_sort, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)leaf_smem_bytes()));
  CK(cudaFuncSetAttribute(k_partition<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tile_smem_bytes(MSD_RADIX)));
  done = true;
}

// Launch the leaf and copy kernels over device-resident lists. The grids are upper bounds: the kernels
// read the real list lengths from `counters` ([0] leaves, [2] copies, [4] receives the number of segments
// the packed kernel could not take), so the host does not have to wait for them before launching.
struct LeafLaunch {
  DBuf wide;
  LeafParams lp;
  uint32_t max_leaf = 0;
};
static void run_leaf_and_copy(hs_ctx* c, MoveSet& ms, const KeySrc& key, int src_code, const ListBufs& leaf,
                              uint32_t max_leaf, const ListBufs& copy, uint32_t max_copy, uint32_t* counters,
                              LeafLaunch& ll, const ListBufs* small = nullptr, uint32_t max_small = 0) {
  ll.max_leaf = max_leaf;
  if (small && max_small) {
    LeafParams sp;
    memset(&sp, 0, sizeof sp);
    sp.mv = make_move(ms, src_code, BUF_A);
    sp.key = key;
    sp.segs = small->view();
    sp.in_place = src_code == BUF_A ? 1 : 0;
    sp.count = counters + 5;
    Stage st(c, "leaf_small");
    LAUNCH(c, k_leaf_small, (max_small + PT_WARPS - 1) / PT_WARPS, PT_THREADS, 0, sp);
  }
  if (max_leaf) {
    LeafParams& lp = ll.lp;
    memset(&lp, 0, sizeof lp);
    lp.mv = make_move(ms, src_code, BUF_A);
    lp.key = key;
    lp.segs = leaf.view();
    lp.in_place = src_code == BUF_A ? 1 : 0;
    lp.count = counters;
    lp.n_wide = counters + 4;
    ll.wide.alloc(c, max_leaf);
    lp.wide = ll.wide.as<uint8_t>();
    Stage st(c, "leaf_sort");
    const char* trace_path = getenv("HSGPU_LEAF_TRACE");   // diagnostics: per-segment phase clocks of this launch
    DBuf d_trace;
    if (trace_path && *trace_path) {
      d_trace.alloc(c, (size_t)max_leaf * 16 * 8);
      d_trace.zero();
      lp.trace = d_trace.as<long long>();
    }
    if (key.mode == KS_U64) LAUNCH(c, k_leaf_sort32<KS_U64>, max_leaf, PT_THREADS, leaf32_smem_bytes(), lp);
    else if (key.mode == KS_I64) LAUNCH(c, k_leaf_sort32<KS_I64>, max_leaf, PT_THREADS, leaf32_smem_bytes(), lp);
    else LAUNCH(c, k_leaf_sort32<KS_I32>, max_leaf, PT_THREADS, leaf32_smem_bytes(), lp);
    if (lp.trace) {
      std::vector<long long> h((size_t)max_leaf * 16);
      CK(cudaMemcpyAsync(h.data(), lp.trace, h.size() * 8, cudaMemcpyDeviceToHost, c->stream));
      hs_sync(c);
      if (FILE* f = fopen(trace_path, "wb")) { fwrite(h.data(), 8, h.size(), f); fclose(f); }
      lp.trace = nullptr;
    }
  }
  if (max_copy && src_code != BUF_A) {
    MoveArgs mv = make_move(ms, src_code, BUF_A);
    Stage st(c, "copy_segments");
    LAUNCH(c, k_copy_segments, dim3(max_copy, COPY_SPLIT), 256, 0, mv, copy.view(), counters + 2);
  }
}
// after the counters have been read back: the 64-bit kernel, only if some segment needs it
static void run_wide_leaves(hs_ctx* c, LeafLaunch& ll, uint32_t n_wide) {
  if (!n_wide || !ll.max_leaf) return;
  Stage st(c, "leaf_sort_wide");
  LAUNCH(c, k_leaf_sort, ll.max_leaf, PT_THREADS, leaf_smem_bytes(), ll.lp);
}

// The 64-bit kernel runs only for segments the packed kernel flagged. With few leaves it is cheaper to launch it
// blind (unflagged CTAs exit at once) than to ask the device how many were flagged.
static void launch_wide(hs_ctx* c, LeafLaunch& ll, uint32_t* d_counters) {
  if (!ll.max_leaf) return;
  uint32_t nw = 1;
  if (ll.max_leaf > 8192) nw = *(const uint32_t*)hs_read_wait(c, hs_read_begin(c, d_counters + 4, 4));
  run_wide_leaves(c, ll, nw);
}

// ---------------------------------------------------------------- pipelined sort of the large segments
// One launch sorts every level-0 segment that is larger than a leaf: per segment a digit histogram of the MSD
// level (H tiles; the last tile of a segment also turns the histogram into write cursors), the MSD scatter (S
// tiles) and one leaf sort per child (L items). Work items are handed out through an atomic ticket in an order
// in which an item only ever waits for items with smaller tickets (those already run or have finished), so the
// kernel needs no co-residency guarantee:   step k = [ S(group k-1) | H(group k) | L(group k-2) ].
// The scatter does not write the rows to their own positions in buffer B but into one of PIPE_SLOTS slots at the
// start of B, re-used group after group: the leaf sort reads a slot shortly after it was written, while its
// lines are still in L2, and the next group overwrites them there. Memory-bound scatter tiles and issue-bound
// leaf sorts run side by side on every SM.
#define PIPE_SLOTS 3
#define PK_SCATTER 0u
#define PK_HIST 1u
#define PK_LEAF 2u
struct PipePhase { uint32_t first, kind, group, pad; };
struct PipeParams {
  PartParams part;              // mv.src: where the rows are (buffer A, or S0), mv.dst: buffer B (the slots), key
  void* fin[HS_MAX_MOVE];       // buffer A: receives the sorted rows
  uint32_t key_bits;            // unsorted low key bits of every level-0 segment
  uint32_t slot_rows;
  const uint32_t* seg_start;    // [ns]
  const uint32_t* seg_len;      // [ns]
  const uint32_t* seg_tile0;    // [ns + 1] tiles before segment s
  const uint32_t* seg_child0;   // [ns + 1] children before segment s
  const uint32_t* grp_seg0;     // [ng + 1] first segment of group g
  const PipePhase* phases;
  uint32_t nphases, ntickets;
  uint32_t* ticket;             // state below is zeroed before the launch
  uint32_t* hist_done;          // [ns] H tiles finished
  uint32_t* plan_done;          // [ns] cursors ready
  uint32_t* scat_done;          // [ns] S tiles finished
  uint32_t* leaf_done;          // [ng] L items finished
  uint32_t* cursor;             // [ns * MSD_RADIX]
  uint32_t* child_start;        // [ns * MSD_RADIX]
  uint32_t* child_len;          // [ns * MSD_RADIX]
  SegList large;                // children that are still larger than a leaf: left to the level loop
  uint32_t* counters;           // [1] entries of `large`, [3] their tiles
};

__device__ __forceinline__ uint32_t ld_acquire_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// the CTA waits until *p1 >= need1 and *p2 >= need2 (counters bumped by CTAs with smaller tickets); p1 may be null
__device__ __forceinline__ void pipe_wait2(const uint32_t* p1, uint32_t need1, const uint32_t* p2, uint32_t need2) {
  if (threadIdx.x == 0) {
    uint32_t a = p1 ? ld_acquire_u32(p1) : need1, b = ld_acquire_u32(p2);
    while (a < need1 || b < need2) {
      __nanosleep(40);
      if (a < need1) a = ld_acquire_u32(p1);
      if (b < need2) b = ld_acquire_u32(p2);
    }
    __threadfence();   // also drops this SM's L1 lines: a slot may have been read here under its previous contents
    asm volatile("fence.proxy.async.global;" ::: "memory");   // what was acquired is read by bulk copies next
  }
  __syncthreads();
}
// the CTA's global writes are done: publish them and bump a counter; returns the counter's old value to thread 0
__device__ __forceinline__ uint32_t pipe_signal(uint32_t* p) {
  __syncthreads();
  uint32_t old = 0;
  if (threadIdx.x == 0) {
    __threadfence();
    old = atomicAdd(p, 1u);
  }
  return old;
}
__device__ __forceinline__ unsigned long long pipe_clock() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// largest s in [lo, hi) with pre[s] <= x
__device__ __forceinline__ uint32_t pipe_find(const uint32_t* pre, uint32_t lo, uint32_t hi, uint32_t x) {
  while (hi - lo > 1) {
    uint32_t mid = (lo + hi) >> 1;
    if (pre[mid] <= x) lo = mid; else hi = mid;
  }
  return lo;
}

// One descriptor per ticket, written by k_pipe_describe before the pipeline starts, so that a CTA learns all about
// its next work item from one 32-byte load instead of a chain of dependent table look-ups.
struct PipeItem {
  uint32_t kind;     // PK_*
  uint32_t group;
  uint32_t seg;
  uint32_t a;        // H/S: first row of the tile; L: index of the child inside the segment
  uint32_t b;        // H/S: rows of the tile
  uint32_t bias;     // slot row = row + bias (mod 2^32)
  uint32_t need;     // H: tiles of the segment; S: leaf items of the group that used the slot last (0: none);
                     // L: scatter tiles of the segment
  uint32_t pad;
};
__global__ void __launch_bounds__(256) k_pipe_describe(const __grid_constant__ PipeParams p, PipeItem* items,
                                                       uint32_t n_end) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= p.ntickets) {
    if (t < p.ntickets + n_end) {     // end markers: one per CTA of the pipeline kernel
      PipeItem e;
      e.kind = 0xFFFFFFFFu; e.group = e.seg = e.a = e.b = e.bias = e.need = e.pad = 0;
      items[t] = e;
    }
    return;
  }
  uint32_t lo = 0, hi = p.nphases;
  while (hi - lo > 1) {
    uint32_t mid = (lo + hi) >> 1;
    if (p.phases[mid].first <= t) lo = mid; else hi = mid;
  }
  const PipePhase ph = p.phases[lo];
  const uint32_t idx = t - ph.first, g = ph.group;
  const uint32_t s0 = p.grp_seg0[g], s1 = p.grp_seg0[g + 1];
  PipeItem it;
  it.kind = ph.kind;
  it.group = g;
  it.bias = (g % PIPE_SLOTS) * p.slot_rows - p.seg_start[s0];
  it.pad = 0;
  if (ph.kind == PK_LEAF) {
    const uint32_t C = p.seg_child0[s0] + idx;
    const uint32_t s = pipe_find(p.seg_child0, s0, s1, C);
    it.seg = s;
    it.a = C - p.seg_child0[s];
    it.b = 0;
    it.need = p.seg_tile0[s + 1] - p.seg_tile0[s];
  } else {
    const uint32_t T = p.seg_tile0[s0] + idx;
    const uint32_t s = pipe_find(p.seg_tile0, s0, s1, T);
    const uint32_t tt = T - p.seg_tile0[s];
    const uint32_t left = p.seg_len[s] - tt * PT_TILE;
    it.seg = s;
    it.a = p.seg_start[s] + tt * PT_TILE;
    it.b = left < PT_TILE ? left : PT_TILE;
    if (ph.kind == PK_HIST) it.need = p.seg_tile0[s + 1] - p.seg_tile0[s];
    else if (g >= PIPE_SLOTS) {
      const uint32_t q0 = p.grp_seg0[g - PIPE_SLOTS], q1 = p.grp_seg0[g - PIPE_SLOTS + 1];
      it.need = p.seg_child0[q1] - p.seg_child0[q0];
    } else it.need = 0;
  }
  items[t] = it;
}

// Persistent CTAs (two per SM): each takes tickets until none is left. While an item is being worked on, one
// lane of warp 1 already fetches the next ticket and its descriptor.
template <int KMODE>
__global__ void __launch_bounds__(PT_THREADS, 2 * PT_MIN_CTAS / 2) k_sort_pipeline(const __grid_constant__ PipeParams p,
                                                                 const PipeItem* __restrict__ items,
                                                                 uint32_t bar_offset, uint32_t prefetch,
                                                                 unsigned long long* trace) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ __align__(16) PipeItem s_item[2];
  __shared__ uint32_t s_flag;
  uint64_t* bar = (uint64_t*)(smem_raw + bar_offset);   // 4 mbarriers behind every body's shared-memory layout
  uint32_t phase = 0;                                    // parity each of them completes next
  const int ncols = p.part.mv.n;
  const void* const* ring = (const void* const*)p.part.mv.dst;
  // the next item is fetched by lane 0 of warp 1 in three non-blocking steps spread over the current item: take
  // a ticket (top), load its descriptor (after the item's first wait), park it in shared memory (bottom)
  // (tickets past the end read one of the end markers that follow the descriptors)
  uint32_t nt = 0;
  auto load_desc = [&](int slot) {
    const PipeItem* q = items + nt;
    const uint32_t d = smem_u32(&s_item[slot]);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(d), "l"(q) : "memory");
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(d + 16), "l"((const char*)q + 16) : "memory");
  };
  auto park = [&]() { asm volatile("cp.async.wait_all;" ::: "memory"); };
  if (threadIdx.x == 0) {
    for (int b = 0; b < 4; b++) mbar_init(&bar[b], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (threadIdx.x == 32) {
    nt = atomicAdd(p.ticket, 1u);
    load_desc(0);
    park();
  }
  __syncthreads();
  for (int cur = 0;; cur ^= 1) {
    const PipeItem it = s_item[cur];
    if (it.kind == 0xFFFFFFFFu) break;
    if (prefetch && threadIdx.x == 32) nt = atomicAdd(p.ticket, 1u);
    unsigned long long tr0 = 0, tr1 = 0;
    if (trace && threadIdx.x == 0) tr0 = tr1 = pipe_clock();
    const uint32_t s = it.seg, g = it.group, bias = it.bias;
    if (it.kind == PK_HIST || it.kind == PK_SCATTER) {
      const uint32_t slen = p.seg_len[s];
      const int bits = msd_bits(slen, p.key_bits);
      const uint32_t shift = p.key_bits - bits;
      uint32_t* cursor = p.cursor + (size_t)s * MSD_RADIX;
      if (it.kind == PK_HIST) {
        uint32_t* hist = (uint32_t*)smem_raw;
        uint32_t* scratch = hist + MSD_RADIX;
        msd_hist_tile(p.part.mv, p.part.key, it.a, it.b, 0, shift, bits, hist, cursor);
        if (prefetch && threadIdx.x == 32) load_desc(cur ^ 1);
        const uint32_t old = pipe_signal(&p.hist_done[s]);
        if (threadIdx.x == 0) s_flag = old + 1 == it.need;
        __syncthreads();
        if (s_flag) {
          // the last histogram tile of the segment: counts -> write cursors and the children's extents
          __threadfence();
          const uint32_t d = threadIdx.x;
          const uint32_t hcount = d < MSD_RADIX ? __ldcg(&cursor[d]) : 0;
          const uint32_t ex = block256_excl_scan_u32(hcount, scratch);
          if (d < MSD_RADIX) {
            const uint32_t start = p.seg_start[s] + ex;
            cursor[d] = start;
            p.child_start[(size_t)s * MSD_RADIX + d] = start;
            p.child_len[(size_t)s * MSD_RADIX + d] = hcount;
          }
          pipe_signal(&p.plan_done[s]);
        }
      } else {
        // the slot must have been read by every leaf of the group that used it last; the cursors must be ready
        pipe_wait2(it.need ? &p.leaf_done[g - PIPE_SLOTS] : nullptr, it.need, &p.plan_done[s], 1u);
        if (trace && threadIdx.x == 0) tr1 = pipe_clock();
        if (prefetch && threadIdx.x == 32) load_desc(cur ^ 1);
        partition_tile<1>(p.part, smem_raw, it.a, it.b, bits, shift, 0, cursor, bias, bar, &phase);
        pipe_signal(&p.scat_done[s]);
      }
    } else {
      // ---- PK_LEAF: one child of one segment
      pipe_wait2(nullptr, 0, &p.scat_done[s], it.need);
      if (trace && threadIdx.x == 0) tr1 = pipe_clock();
      if (prefetch && threadIdx.x == 32) load_desc(cur ^ 1);
      const uint32_t cs = __ldcg(&p.child_start[(size_t)s * MSD_RADIX + it.a]);
      const uint32_t cl = __ldcg(&p.child_len[(size_t)s * MSD_RADIX + it.a]);
      const uint32_t sh = p.key_bits - (uint32_t)msd_bits(p.seg_len[s], p.key_bits);
      if (cl == 0) {
      } else if (cl <= SMALL_MAX) {
        if (threadIdx.x < 32) leaf_small_warp(p.part.mv.esize, ncols, ring, p.fin, p.part.key, cs + bias, cs, cl);
      } else if (cl <= LEAF_MAX) {
        const int r = cl <= 4 * PT_THREADS
            ? leaf32_body<KMODE, 4>(smem_raw, p.part.mv.esize, ncols, ring, p.fin, p.part.key, cs + bias, cs, cl, sh, 0,
                                    false, bar, &phase)
            : leaf32_body<KMODE, PT_ITEMS>(smem_raw, p.part.mv.esize, ncols, ring, p.fin, p.part.key, cs + bias, cs, cl,
                                           sh, 0, false, bar, &phase);
        if (r) {
          __syncthreads();
          leaf64_body(smem_raw, p.part.mv.esize, ncols, ring, p.fin, p.part.key, cs + bias, cs, cl, false);
        }
      } else {
        // still larger than a leaf (skewed keys, or all keys equal): the rows leave the slot for the place they
        // came from; if key bits are left, the level loop after this kernel splits the child further
        for (int c = 0; c < ncols; c++) {
          const int es = p.part.mv.esize[c];
          const unsigned char* a = (const unsigned char*)ring[c] + (size_t)(uint32_t)(cs + bias) * es;
          unsigned char* b = (sh ? (unsigned char*)p.part.mv.src[c] : (unsigned char*)p.fin[c]) + (size_t)cs * es;
          const size_t bytes = (size_t)cl * es;
          if (((((uintptr_t)a) | ((uintptr_t)b)) & 7) == 0) {
            for (size_t i = threadIdx.x; i < (bytes >> 3); i += PT_THREADS) ((uint64_t*)b)[i] = ((const uint64_t*)a)[i];
            for (size_t i = (bytes & ~(size_t)7) + threadIdx.x; i < bytes; i += PT_THREADS) b[i] = a[i];
          } else {
            for (size_t i = threadIdx.x; i < bytes; i += PT_THREADS) b[i] = a[i];
          }
        }
        if (sh && threadIdx.x == 0) {
          const uint32_t i = atomicAdd(&p.counters[1], 1u);
          p.large.start[i] = cs; p.large.len[i] = cl; p.large.shift[i] = sh; p.large.base[i] = 0;
          atomicAdd(&p.counters[3], (cl + PT_TILE - 1) / PT_TILE);
        }
      }
      pipe_signal(&p.leaf_done[g]);
    }
    if (trace && threadIdx.x == 0) {
      // one record per item: [kind | segment << 8 | CTA << 32 | SM << 44], start, end of the wait, end (ns)
      const unsigned long long i = atomicAdd(&trace[0], 1ull);
      uint32_t smid;
      asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
      unsigned long long* r = trace + 4 + 4 * i;
      r[0] = (unsigned long long)it.kind | ((unsigned long long)(s & 0xFFFFFFu) << 8) |
             ((unsigned long long)(blockIdx.x & 0xFFFu) << 32) | ((unsigned long long)smid << 44);
      r[1] = tr0; r[2] = tr1; r[3] = pipe_clock();
    }
    if (threadIdx.x == 32) {
      if (!prefetch) {
        nt = atomicAdd(p.ticket, 1u);
        load_desc(cur ^ 1);
      }
      park();
    }
    __syncthreads();   // the item's shared memory is free, s_item[cur ^ 1] is in place
  }
}

static bool pipeline_enabled() {
  // off unless asked for: measured on B200 (profiles/README.md) the single launch does not beat the separate
  // kernels — co-resident scatter tiles and leaf sorts take as long as they do alone
  const char* e = getenv("HSGPU_SORT_PIPELINE");
  return MSD_RADIX <= PT_THREADS && MSD_RADIX == 256 && e && *e == '1';
}
static size_t pipeline_smem_bytes() {
  size_t a = leaf32_smem_bytes(), b = leaf_smem_bytes(), t = tile_smem_bytes(MSD_RADIX);
  return std::max(a, std::max(b, t));
}

// The level-0 large segments (start/len, host-known, ascending) are split into those whose rows fit a slot of the
// pipelined sort (ps/pl, in groups) and the rest (left to the level loop).
struct PipePlan {
  std::vector<uint32_t> ps, pl, grp_seg0, rest_s, rest_l;
  uint32_t slot_rows = 0;
};
static PipePlan plan_sort_pipeline(const MoveSet& ms, int64_t n, const std::vector<uint32_t>& lg_s,
                                   const std::vector<uint32_t>& lg_l) {
  PipePlan pl_;
  size_t row_bytes = 0;
  for (int i = 0; i < ms.n; i++) row_bytes += (size_t)ms.esize[i];
  size_t ring_mb = 48;
  if (const char* e = getenv("HSGPU_RING_MB")) { long v = atol(e); if (v > 0) ring_mb = (size_t)v; }
  uint64_t slot_target = (uint64_t)ring_mb * 1024 * 1024 / PIPE_SLOTS / row_bytes;
  if (slot_target > (uint64_t)n / PIPE_SLOTS) slot_target = (uint64_t)n / PIPE_SLOTS;
  // groups: consecutive segments whose address span fits a slot
  std::vector<uint32_t>&ps = pl_.ps, &pl = pl_.pl, &grp_seg0 = pl_.grp_seg0;
  uint64_t g_first = 0;
  for (size_t i = 0; i < lg_s.size(); i++) {
    if (lg_l[i] > slot_target) { pl_.rest_s.push_back(lg_s[i]); pl_.rest_l.push_back(lg_l[i]); continue; }
    const uint64_t end = (uint64_t)lg_s[i] + lg_l[i];
    if (grp_seg0.empty() || end - g_first > slot_target) {
      grp_seg0.push_back((uint32_t)ps.size());
      g_first = lg_s[i];
    }
    if (end - g_first > pl_.slot_rows) pl_.slot_rows = (uint32_t)(end - g_first);
    ps.push_back(lg_s[i]);
    pl.push_back(lg_l[i]);
  }
  if (!ps.empty()) grp_seg0.push_back((uint32_t)ps.size());
  return pl_;
}

// Launches k_sort_pipeline over a plan. Children of pipelined segments that are still larger than a leaf are
// appended on the device to `large` (counters[1] entries, counters[3] tiles).
static void run_sort_pipeline(hs_ctx* c, MoveSet& ms, const KeySrc& key, int key_bits, int src_code,
                              const PipePlan& plan, const ListBufs& large, uint32_t* d_counters) {
  const std::vector<uint32_t>&ps = plan.ps, &pl = plan.pl, &grp_seg0 = plan.grp_seg0;
  const uint32_t slot_rows = plan.slot_rows;
  const size_t ns = ps.size();
  if (ns == 0) return;
  const size_t ng = grp_seg0.size() - 1;
  std::vector<uint32_t> tile0(ns + 1, 0), child0(ns + 1, 0);
  for (size_t s = 0; s < ns; s++) {
    tile0[s + 1] = tile0[s] + (pl[s] + PT_TILE - 1) / PT_TILE;
    child0[s + 1] = child0[s] + (1u << msd_bits(pl[s], (uint32_t)key_bits));
  }
  // ticket order: step k = S(k-1), H(k), L(k-2)
  std::vector<PipePhase> phases;
  uint32_t ticket = 0;
  auto tiles_of = [&](size_t g) { return tile0[grp_seg0[g + 1]] - tile0[grp_seg0[g]]; };
  auto children_of = [&](size_t g) { return child0[grp_seg0[g + 1]] - child0[grp_seg0[g]]; };
  for (size_t k = 0; k < ng + 2; k++) {
    if (k >= 1 && k - 1 < ng) { phases.push_back(PipePhase{ticket, PK_SCATTER, (uint32_t)(k - 1), 0}); ticket += tiles_of(k - 1); }
    if (k < ng) { phases.push_back(PipePhase{ticket, PK_HIST, (uint32_t)k, 0}); ticket += tiles_of(k); }
    if (k >= 2 && k - 2 < ng) { phases.push_back(PipePhase{ticket, PK_LEAF, (uint32_t)(k - 2), 0}); ticket += children_of(k - 2); }
  }
  // one upload: [seg_start | seg_len | tile0 | child0 | grp_seg0 | phases]
  std::vector<uint32_t> up;
  auto put = [&](const std::vector<uint32_t>& v) { size_t o = up.size(); up.insert(up.end(), v.begin(), v.end()); return o; };
  const size_t o_start = put(ps), o_len = put(pl), o_tile = put(tile0), o_child = put(child0), o_grp = put(grp_seg0);
  const size_t o_ph = up.size();
  up.resize(o_ph + phases.size() * 4);
  memcpy(up.data() + o_ph, phases.data(), phases.size() * sizeof(PipePhase));
  DBuf d_up(c, up.size() * 4);
  CK(cudaMemcpyAsync(d_up.p, up.data(), up.size() * 4, cudaMemcpyHostToDevice, c->stream));
  // state: [ticket(4 words) | hist_done ns | plan_done ns | scat_done ns | leaf_done ng | cursor ns*256], zeroed
  const size_t n_state = 4 + 3 * ns + ng + ns * MSD_RADIX;
  DBuf d_state(c, n_state * 4), d_child(c, 2 * ns * MSD_RADIX * 4);
  d_state.zero();
  uint32_t* st = d_state.as<uint32_t>();
  const uint32_t* dup = d_up.as<uint32_t>();

  PipeParams* pp = new PipeParams;
  memset(pp, 0, sizeof *pp);
  pp->part.mv = make_move(ms, src_code, BUF_B);
  pp->part.key = key;
  pp->part.radix = MSD_RADIX;
  pp->part.radix_pad = MSD_RADIX;
  pp->part.bits = MSD_BITS;
  for (int i = 0; i < ms.n; i++) pp->fin[i] = ms.a[i];
  pp->key_bits = (uint32_t)key_bits;
  pp->slot_rows = slot_rows;
  pp->seg_start = dup + o_start; pp->seg_len = dup + o_len; pp->seg_tile0 = dup + o_tile; pp->seg_child0 = dup + o_child;
  pp->grp_seg0 = dup + o_grp;
  pp->phases = (const PipePhase*)(dup + o_ph);
  pp->nphases = (uint32_t)phases.size();
  pp->ntickets = ticket;
  pp->ticket = st;
  pp->hist_done = st + 4; pp->plan_done = pp->hist_done + ns; pp->scat_done = pp->plan_done + ns;
  pp->leaf_done = pp->scat_done + ns; pp->cursor = pp->leaf_done + ng;
  pp->child_start = d_child.as<uint32_t>(); pp->child_len = pp->child_start + ns * MSD_RADIX;
  pp->large = large.view();
  pp->counters = d_counters;
  {
    static bool attr_dev[64] = {false};
    int dev = 0;
    CK(cudaGetDevice(&dev));
    if (!attr_dev[dev & 63]) {
      const int sm = (int)((pipeline_smem_bytes() + 15) & ~(size_t)15) + 48;
      CK(cudaFuncSetAttribute(k_sort_pipeline<KS_U64>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm));
      CK(cudaFuncSetAttribute(k_sort_pipeline<KS_I64>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm));
      CK(cudaFuncSetAttribute(k_sort_pipeline<KS_I32>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm));
      attr_dev[dev & 63] = true;
    }
  }
  {
    Stage stg(c, "sort_pipeline");
    unsigned grid = 2 * (unsigned)hs_sm_count();
    if (grid > ticket) grid = ticket;
    DBuf d_items(c, ((size_t)ticket + grid) * sizeof(PipeItem));
    LAUNCH(c, k_pipe_describe, (ticket + grid + 255) / 256, 256, 0, *pp, d_items.as<PipeItem>(), (uint32_t)grid);
    const size_t bar_off = (pipeline_smem_bytes() + 15) & ~(size_t)15;
    const size_t sm = bar_off + 48;
    const PipeItem* di = d_items.as<PipeItem>();
    const char* pf = getenv("HSGPU_PIPE_PREFETCH");
    const uint32_t prefetch = (pf && *pf == '0') ? 0u : 1u;
    // HSGPU_PIPE_TRACE=<file>: per-item timestamps of this launch (diagnostics; costs a device sync)
    const char* trace_path = getenv("HSGPU_PIPE_TRACE");
    DBuf d_trace;
    unsigned long long* tr = nullptr;
    if (trace_path && *trace_path) {
      d_trace.alloc(c, ((size_t)ticket * 4 + 4) * 8);
      d_trace.zero();
      tr = d_trace.as<unsigned long long>();
    }
    if (key.mode == KS_U64) LAUNCH(c, k_sort_pipeline<KS_U64>, grid, PT_THREADS, sm, *pp, di, (uint32_t)bar_off, prefetch, tr);
    else if (key.mode == KS_I64) LAUNCH(c, k_sort_pipeline<KS_I64>, grid, PT_THREADS, sm, *pp, di, (uint32_t)bar_off, prefetch, tr);
    else LAUNCH(c, k_sort_pipeline<KS_I32>, grid, PT_THREADS, sm, *pp, di, (uint32_t)bar_off, prefetch, tr);
    if (tr) {
      std::vector<unsigned long long> h((size_t)ticket * 4 + 4);
      CK(cudaMemcpyAsync(h.data(), tr, h.size() * 8, cudaMemcpyDeviceToHost, c->stream));
      hs_sync(c);
      if (FILE* f = fopen(trace_path, "wb")) { fwrite(h.data(), 8, h.size(), f); fclose(f); }
    }
  }
  delete pp;
}

// h_seg_off (optional): the segment boundaries on the host. With them level 0 is classified on the host
// and the sort starts without a device round trip.
static void segsort_impl(hs_ctx* c, MoveSet& ms, const KeySrc& key, int key_bits, const uint32_t* d_seg_off,
                         const uint32_t* d_start, const uint32_t* d_len, int64_t nseg, int64_t n,
                         const uint32_t* h_seg_off) {
  if (n == 0 || nseg == 0) return;
  set_smem_attrs_once();
  size_t cap_leaf = (size_t)nseg, cap_large = (size_t)(n / LEAF_MAX + 1);
  ListBufs leaf, large, copy, small;
  DBuf counters(c, 8 * 4), tile_prefix(c, (cap_large + 1) * 4);
  uint32_t h[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  const bool ext = ms.s0[0] != nullptr;   // the rows start outside buffer A: every row has to be moved
  int src = ext ? BUF_S0 : BUF_A;
  LeafLaunch ll0;
  bool pipe_children = false;   // the level loop starts with children of the pipelined level
  if (h_seg_off) {
    // ---- level 0 on the host
    std::vector<uint32_t> tp;
    std::vector<uint32_t> lf_s, lf_l, lg_s, lg_l, sm_s, sm_l, cp_s, cp_l;
    tp.push_back(0);
    for (int64_t sidx = 0; sidx < nseg; sidx++) {
      uint32_t st = h_seg_off[sidx], len = h_seg_off[sidx + 1] - st;
      if (len == 0 || (len < 2 && !ext)) continue;
      if (len <= SMALL_MAX) { sm_s.push_back(st); sm_l.push_back(len); }
      else if (len <= LEAF_MAX) { lf_s.push_back(st); lf_l.push_back(len); }
      else if (key_bits > 0) { lg_s.push_back(st); lg_l.push_back(len); tp.push_back(tp.back() + (len + PT_TILE - 1) / PT_TILE); }
      else if (ext) { cp_s.push_back(st); cp_l.push_back(len); }
      // larger than a leaf with no key bits: all keys equal; in place when the rows already are in A
    }
    PipePlan plan;
    const bool piped = pipeline_enabled() && !lg_s.empty() && ms.b[0] != nullptr;
    if (piped) {
      plan = plan_sort_pipeline(ms, n, lg_s, lg_l);
      lg_s = plan.rest_s;
      lg_l = plan.rest_l;
      tp.assign(1, 0);
      for (uint32_t len : lg_l) tp.push_back(tp.back() + (len + PT_TILE - 1) / PT_TILE);
    }
    h[0] = (uint32_t)lf_s.size();
    h[1] = (uint32_t)lg_s.size();
    h[2] = (uint32_t)cp_s.size();
    h[3] = tp.back();
    h[5] = (uint32_t)sm_s.size();
    auto upload = [&](ListBufs& lb, const std::vector<uint32_t>& st, const std::vector<uint32_t>& ln, size_t cap = 0) {
      size_t m = st.size();
      lb.alloc(c, std::max(cap, m ? m : 1));
      if (!m) return;
      std::vector<uint32_t> sh(m, (uint32_t)key_bits);
      CK(cudaMemcpyAsync(lb.start.p, st.data(), m * 4, cudaMemcpyHostToDevice, c->stream));
      CK(cudaMemcpyAsync(lb.len.p, ln.data(), m * 4, cudaMemcpyHostToDevice, c->stream));
      CK(cudaMemcpyAsync(lb.shift.p, sh.data(), m * 4, cudaMemcpyHostToDevice, c->stream));
    };
    const size_t large_cap = !plan.ps.empty() ? cap_large : 0;   // the pipeline may append children
    upload(leaf, lf_s, lf_l);
    upload(large, lg_s, lg_l, large_cap);
    large.base.alloc(c, std::max(large_cap, lg_s.size() ? lg_s.size() : (size_t)1) * 8);
    large.base.zero();
    leaf.base.alloc(c, (lf_s.size() ? lf_s.size() : 1) * 8);
    leaf.base.zero();
    upload(small, sm_s, sm_l);
    upload(copy, cp_s, cp_l);
    CK(cudaMemcpyAsync(counters.p, h, 32, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(tile_prefix.p, tp.data(), tp.size() * 4, cudaMemcpyHostToDevice, c->stream));
    run_leaf_and_copy(c, ms, key, src, leaf, h[0], copy, h[2], counters.as<uint32_t>(), ll0, &small, h[5]);
    launch_wide(c, ll0, counters.as<uint32_t>());
    if (!plan.ps.empty()) {
      run_sort_pipeline(c, ms, key, key_bits, src, plan, large, counters.as<uint32_t>());
      // children that one level could not bring down to a leaf (skewed keys) were appended to `large`
      uint32_t hh[8];
      memcpy(hh, hs_read_wait(c, hs_read_begin(c, counters.p, 32)), 32);
      if (hh[1] != h[1]) {
        h[1] = hh[1];
        h[3] = hh[3];
        LAUNCH(c, k_tile_prefix, 1, 1024, 0, large.view(), counters.as<uint32_t>(), tile_prefix.as<uint32_t>());
        pipe_children = true;
      }
    }
  } else {
    leaf.alloc(c, cap_leaf, true);
    large.alloc(c, cap_large, true);
    copy.alloc(c, cap_large);
    small.alloc(c, cap_leaf);
    counters.zero();
    PlanOut po{leaf.view(), large.view(), copy.view(), small.view(), counters.as<uint32_t>()};
    hs_stage_begin(c, "sort_plan");
    LAUNCH(c, k_classify, (unsigned)((nseg + 255) / 256), 256, 0, d_seg_off, d_start, d_len, nseg, (uint32_t)key_bits, po);
    LAUNCH(c, k_tile_prefix, 1, 1024, 0, large.view(), counters.as<uint32_t>(), tile_prefix.as<uint32_t>());
    hs_stage_end(c);
    // the list lengths come back first: grids sized by upper bounds (rows / 64 CTAs) cost more in empty CTAs
    // than the round trip does
    if (ext) HS_THROW(HS_EINVAL, "external start buffers need host-known segments");
    memcpy(h, hs_read_wait(c, hs_read_begin(c, counters.p, 32)), 32);
    run_leaf_and_copy(c, ms, key, BUF_A, leaf, h[0], copy, 0, counters.as<uint32_t>(), ll0, &small, h[5]);
    launch_wide(c, ll0, counters.as<uint32_t>());
  }
  uint32_t n_large = h[1], n_tiles = h[3];
  int level = pipe_children ? 1 : 0;   // children of a level are re-ranged before they are split again
  while (n_large > 0) {
    if (++level > 80) HS_THROW(HS_ECUDA, "segmented sort did not converge");
    // histograms -> cursors + child lists
    DBuf cursor(c, (size_t)n_large * MSD_RADIX * 4);
    cursor.zero();
    PartParams p;
    memset(&p, 0, sizeof p);
    p.mv = make_move(ms, src, other_buf(src));
    p.key = key;
    p.radix = MSD_RADIX;
    p.radix_pad = MSD_RADIX;  // digits of an MSD level
    p.bits = MSD_BITS;
    p.cursor = cursor.as<uint32_t>();
    p.segs = large.view();
    p.tile_prefix = tile_prefix.as<uint32_t>();
    p.nseg = n_large;
    if (level >= 2) {
      // what is still large after a level is skewed: take the next digit from its real key range
      DBuf rmin(c, (size_t)n_large * 8), rmax(c, (size_t)n_large * 8);
      CK(cudaMemsetAsync(rmin.p, 0xFF, (size_t)n_large * 8, c->stream));
      rmax.zero();
      hs_stage_begin(c, "msd_range");
      LAUNCH(c, k_seg_range, n_tiles, PT_THREADS, 0, p, rmin.as<uint64_t>(), rmax.as<uint64_t>());
      LAUNCH(c, k_seg_rerange, (n_large + 255) / 256, 256, 0, large.view(), n_large, rmin.as<uint64_t>(),
             rmax.as<uint64_t>());
      hs_stage_end(c);
    }
    hs_stage_begin(c, "msd_hist");
    LAUNCH(c, k_msd_hist, n_tiles, PT_THREADS, 0, p);
    hs_stage_end(c);
    ListBufs nleaf, nlarge, ncopy, nsmall;
    size_t children = (size_t)n_large * MSD_RADIX;
    nleaf.alloc(c, children, true);
    nlarge.alloc(c, cap_large, true);
    ncopy.alloc(c, cap_large);
    nsmall.alloc(c, children);
    DBuf ncounters(c, 8 * 4);
    ncounters.zero();
    PlanOut npo{nleaf.view(), nlarge.view(), ncopy.view(), nsmall.view(), ncounters.as<uint32_t>()};
    hs_stage_begin(c, "msd_plan");
    LAUNCH(c, k_msd_plan, n_large, MSD_RADIX, 0, large.view(), cursor.as<uint32_t>(), npo);
    DBuf ntile_prefix(c, (cap_large + 1) * 4);
    LAUNCH(c, k_tile_prefix, 1, 1024, 0, nlarge.view(), ncounters.as<uint32_t>(), ntile_prefix.as<uint32_t>());
    hs_stage_end(c);
    // the child counts travel to the host while the scatter runs: the leaf grids below are exact
    const int slot = hs_read_begin(c, ncounters.p, 32);
    hs_stage_begin(c, "msd_scatter");
    LAUNCH(c, k_partition<1>, n_tiles, PT_THREADS, tile_smem_bytes(MSD_RADIX), p);
    hs_stage_end(c);
    src = other_buf(src);
    memcpy(h, hs_read_wait(c, slot), 32);
    LeafLaunch ll;
    run_leaf_and_copy(c, ms, key, src, nleaf, h[0], ncopy, h[2], ncounters.as<uint32_t>(), ll, &nsmall, h[5]);
    launch_wide(c, ll, ncounters.as<uint32_t>());
    large = std::move(nlarge);
    leaf = std::move(nleaf);
    copy = std::move(ncopy);
    small = std::move(nsmall);
    tile_prefix = std::move(ntile_prefix);
    n_large = h[1];
    n_tiles = h[3];
  }
}

void hs_segmented_sort(hs_ctx* c, MoveSet& ms, const KeySrc& key, int key_bits, const uint32_t* d_seg_off,
                       int64_t nseg, int64_t n, const uint32_t* h_seg_off) {
  segsort_impl(c, ms, key, key_bits, d_seg_off, nullptr, nullptr, nseg, n, h_seg_off);
}
void hs_segmented_sort_list(hs_ctx* c, MoveSet& ms, const KeySrc& key, int key_bits, const uint32_t* d_start,
                            const uint32_t* d_len, int64_t nseg, int64_t n) {
  segsort_impl(c, ms, key, key_bits, nullptr, d_start, d_len, nseg, n, nullptr);
}
