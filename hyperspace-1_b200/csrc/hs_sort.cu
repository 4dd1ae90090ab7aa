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

#include "hs_sort_shared.cuh"

// shared memory carve-up of the partition kernel
#define STAGE_BYTES (PT_TILE * 8 + 128)   // one column slice of a tile + alignment slack
// Landing buffers for the column slices of a tile. With two, the next column is already in flight while the
// current one is scattered, but a tile then needs 104 KB and two tiles fill an SM; with one (64 KB) THREE tiles are
// resident, each waiting for its next column while the other two work — measured faster (the kernel is a chain
// of dependent steps per tile and runs at 30 % issue utilisation: a third chain pays more than the prefetch).
#ifndef PT_STAGES
#define PT_STAGES 1
#endif
#define PT_PART_CTAS (PT_STAGES == 1 ? 3 : 2)   // resident CTAs per SM k_partition is compiled for
struct TileSmem {
  unsigned char* stage[2];  // TMA landing buffers (stage[1] == stage[0] when PT_STAGES == 1)
  uint32_t* gdst;           // PT_TILE: global destination row of every tile-sorted position
  uint16_t* inv;            // PT_TILE: tile row that lands at every tile-sorted position
  uint32_t* cnt;            // radix_pad: counters, then first position of every digit
  uint32_t* tot;            // radix_pad: rows per digit
  uint32_t* gbase;          // radix_pad: reserved global run start per digit
  uint32_t* scratch;        // PT_SCRATCH
  uint64_t* bar;            // 3 mbarriers
  unsigned char* bstage;    // TMA landing buffer of the tile's bucket ids (u16): the ids are in registers before
                            // the first destination is written, so it shares the bytes of gdst
};
__device__ __forceinline__ TileSmem carve(unsigned char* base, int radix_pad) {
  TileSmem s;
  s.stage[0] = base;
  s.stage[1] = base + (PT_STAGES - 1) * STAGE_BYTES;
  s.gdst = (uint32_t*)(base + PT_STAGES * STAGE_BYTES);
  s.inv = (uint16_t*)(s.gdst + PT_TILE);
  s.cnt = (uint32_t*)(s.inv + PT_TILE);
  s.tot = s.cnt + radix_pad;
  s.gbase = s.tot + radix_pad;
  s.scratch = s.gbase + radix_pad;
  s.bar = (uint64_t*)(s.scratch + PT_SCRATCH);
  s.bstage = (unsigned char*)s.gdst;
  return s;
}
static size_t tile_smem_bytes(int radix_pad) {
  return (size_t)PT_STAGES * STAGE_BYTES + (size_t)PT_TILE * 4 + (size_t)PT_TILE * 2 + (size_t)3 * radix_pad * 4 +
         PT_SCRATCH * 4 + 32;
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
  int32_t conv_narrow;      // MODE 0: narrow like MODE 2 instead of widening (the sender's grouped scatter)
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
__device__ __forceinline__ void scatter_column_narrow(const unsigned char* stage, uint32_t lead, uint32_t* __restrict__ dst,
                                                      uint32_t tile_len, const uint16_t* inv, const uint32_t* gdst,
                                                      uint64_t mn) {
  const uint64_t* col = (const uint64_t*)(stage + lead);
#pragma unroll 4
  for (int k = 0; k < PT_ITEMS; k++) {
    uint32_t j = threadIdx.x + k * PT_THREADS;
    if (j < tile_len) dst[gdst[j]] = (uint32_t)((col[inv[j]] ^ HS_SIGN64) - mn);
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
// The tile body shared by the three launch modes. dst_bias is added (mod 2^32) to every destination row and
// ext_bar / ext_phase let a persistent caller keep the mbarriers across tiles (both used by the one-launch sort
// pipeline of round 2, measured slower and removed again: DESIGN.md section 4).
template <int MODE>
__device__ __forceinline__ void partition_tile(const PartParams& p, unsigned char* smem_raw, uint32_t tile_start,
                                               uint32_t tile_len, int bits, uint32_t shift, uint64_t seg_base,
                                               uint32_t* cursor, uint32_t dst_bias, uint64_t* ext_bar = nullptr,
                                               uint32_t* ext_phase = nullptr, uint32_t seg_mul = 0) {
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
    if (PT_STAGES == 2 && ncols > 1) {
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
        digit[k] = msd_digit(key - seg_base, shift + (uint32_t)bits, bits, seg_mul);
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
    const int c = col_at(i), b = PT_STAGES == 2 ? (i & 1) : 0, es = p.mv.esize[c];
    mbar_wait(&sm.bar[b], ((uint32_t)(PT_STAGES == 2 ? i >> 1 : i) ^ (ph_in >> b)) & 1u);
    const uint32_t lead = slice_lead(p.mv.src[c], es, tile_start);
    if (MODE == 2 && c + 1 == p.conv_col1) {
      scatter_column_peers_narrow(sm.stage[b], lead, p.pdst, c, tile_len, sm.inv, sm.gdst, p.conv_min);
    } else if (MODE == 0 && c + 1 == p.conv_col1) {
      if (p.conv_narrow) scatter_column_narrow(sm.stage[b], lead, (uint32_t*)p.mv.dst[c], tile_len, sm.inv, sm.gdst, p.conv_min);
      else scatter_column_widen(sm.stage[b], lead, (uint64_t*)p.mv.dst[c], tile_len, sm.inv, sm.gdst, p.conv_min);
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
    if (threadIdx.x == 0 && i + PT_STAGES < ncols) {
      int cn = col_at(i + PT_STAGES);
      issue_slice(p.mv.src[cn], p.mv.esize[cn], tile_start, tile_len, sm.stage[b], &sm.bar[b]);
    }
  }
  if (ext_phase)
    *ext_phase = PT_STAGES == 2
        ? ph_in ^ (uint32_t)(((ncols + 1) >> 1) & 1) ^ (uint32_t)(((ncols >> 1) & 1) << 1) ^ (MODE != 1 ? 4u : 0u)
        : ph_in ^ (uint32_t)(ncols & 1) ^ (MODE != 1 ? 4u : 0u);
}

template <int MODE>
__global__ void __launch_bounds__(PT_THREADS, PT_PART_CTAS) k_partition(const __grid_constant__ PartParams p) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  uint32_t tile_start, tile_len, shift = 0, seg_mul = 0;
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
    seg_mul = p.segs.mul[s];
    cursor = p.cursor + (size_t)s * MSD_RADIX;
  }
  partition_tile<MODE>(p, smem_raw, tile_start, tile_len, bits, shift, seg_base, cursor, 0u, nullptr, nullptr, seg_mul);
}

// per-segment digit histogram of one MSD level
__device__ __forceinline__ void msd_hist_tile(const MoveArgs& mv, const KeySrc& key, uint32_t tile_start,
                                              uint32_t tile_len, uint64_t seg_base, uint32_t shift, int bits,
                                              uint32_t* hist /* shared, MSD_RADIX */, uint32_t* cursor, uint32_t seg_mul = 0) {
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
    if (j < tile_len) atomicAdd(&hist[msd_digit(kk[k] - seg_base, shift + (uint32_t)bits, bits, seg_mul)], 1u);
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
                p.cursor + (size_t)s * MSD_RADIX, p.segs.mul[s]);
}

struct PlanOut {
  SegList leaf, large, copy, small;
  uint32_t* counters;  // [0] leaf, [1] large, [2] copy, [3] tiles of large, [4] wide leaves, [5] small
  uint32_t* lcounters; // where [1] and [3] are counted: `counters`, or the carry list shared by several sort calls
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
    uint32_t i = atomicAdd(&o.lcounters[1], 1u);
    o.large.start[i] = start; o.large.len[i] = len; o.large.shift[i] = shift; o.large.base[i] = base;
    o.large.mul[i] = 0;
    atomicAdd(&o.lcounters[3], (len + PT_TILE - 1) / PT_TILE);
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
  // the keys of child d start at the first cell of digit d (see msd_digit): its base, with sh - bits bits left
  const int e = msd_extra(sh, (int)bits);
  const uint32_t mul = segs.mul[s];
  const uint32_t c0 = msd_cell_lo(threadIdx.x, e, mul), c1 = msd_cell_lo(threadIdx.x + 1, e, mul);
  const uint64_t child_base = segs.base[s] + ((uint64_t)c0 << (sh - bits - e));
  // bits still varying inside the child: those below a cell plus what numbers the child's cells
  const uint32_t ncell = c1 > c0 ? c1 - c0 : 1;
  const uint32_t child_bits = (sh - bits - e) + (ncell > 1 ? 32 - __clz(ncell - 1) : 0);
  classify_push(out, start, h, child_bits, child_base);
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
  segs.mul[s] = range ? msd_mul_for(range, segs.shift[s], msd_bits(segs.len[s], segs.shift[s]), segs.len[s]) : 0;
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

// (the leaf kernels and their launchers live in hs_leaf.cu)

// ---------------------------------------------------------------- host drivers

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
                        const uint16_t* d_bucket, int64_t n, const uint16_t* d_lut, int32_t nbins, const uint32_t* d_bin_off,
                        int narrow_col, uint64_t narrow_min) {
  // narrow_col >= 0: that column is an int64 key at the source and is written as u32 (sign-flipped value -
  // narrow_min); esize[narrow_col] is the 4 bytes of the destination
  if (n == 0) return;
  if (ncols > HS_MAX_MOVE) HS_THROW(HS_EUNSUPPORTED, "too many columns for the grouped scatter");
  PartParams p;
  memset(&p, 0, sizeof p);
  p.n = (uint32_t)n;
  p.mv.n = ncols;
  for (int i = 0; i < ncols; i++) { p.mv.esize[i] = esize[i]; p.mv.src[i] = src[i]; p.mv.dst[i] = dst[i]; }
  if (narrow_col >= 0) {
    p.mv.esize[narrow_col] = 8;           // what the tile stages
    p.conv_col1 = narrow_col + 1;
    p.conv_min = narrow_min;
    p.conv_narrow = 1;
  }
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


static void set_smem_attrs_once() {
  static bool done_dev[64] = {false};
  int dev = 0;
  CK(cudaGetDevice(&dev));
  bool& done = done_dev[dev & 63];
  if (done) return;
  CK(cudaFuncSetAttribute(k_partition<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tile_smem_bytes(MSD_RADIX)));
  done = true;
}

struct SortCarry {         // children still larger than a leaf after the first level of several sort calls
  ListBufs large;
  DBuf counters;           // [1] segments, [3] their tiles
  size_t cap_large = 0;
};

// The MSD levels over `large` until nothing is larger than a leaf (with a carry: exactly one level).
static void level_loop(hs_ctx* c, MoveSet& ms, const KeySrc& key, ListBufs& large, DBuf& tile_prefix, uint32_t n_large,
                       uint32_t n_tiles, int src, int level, size_t cap_large, SortCarry* carry) {
  uint32_t h[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  ListBufs leaf, copy, small;
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
    // with a carry, the children that are still large join the list all calls share: one more level for all of
    // them at the end (hs_sort_carry_finish) instead of one per call
    const bool to_carry = carry != nullptr;
    PlanOut npo{nleaf.view(), to_carry ? carry->large.view() : nlarge.view(), ncopy.view(), nsmall.view(),
                ncounters.as<uint32_t>(), to_carry ? carry->counters.as<uint32_t>() : ncounters.as<uint32_t>()};
    hs_stage_begin(c, "msd_plan");
    LAUNCH(c, k_msd_plan, n_large, MSD_RADIX, 0, large.view(), cursor.as<uint32_t>(), npo);
    DBuf ntile_prefix(c, (cap_large + 1) * 4);
    if (!to_carry) LAUNCH(c, k_tile_prefix, 1, 1024, 0, nlarge.view(), ncounters.as<uint32_t>(), ntile_prefix.as<uint32_t>());
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
    if (to_carry) return;
    large = std::move(nlarge);
    leaf = std::move(nleaf);
    copy = std::move(ncopy);
    small = std::move(nsmall);
    tile_prefix = std::move(ntile_prefix);
    n_large = h[1];
    n_tiles = h[3];
  }
}

static void segsort_impl(hs_ctx* c, MoveSet& ms, const KeySrc& key, int key_bits, const uint32_t* d_seg_off,
                         const uint32_t* d_start, const uint32_t* d_len, int64_t nseg, int64_t n,
                         const uint32_t* h_seg_off, SortCarry* carry = nullptr) {
  if (carry && !h_seg_off) HS_THROW(HS_EINVAL, "a sort carry needs host-known segments");
  if (n == 0 || nseg == 0) return;
  set_smem_attrs_once();
  size_t cap_leaf = (size_t)nseg, cap_large = (size_t)(n / LEAF_MAX + 1);
  ListBufs leaf, large, copy, small;
  DBuf counters(c, 8 * 4), tile_prefix(c, (cap_large + 1) * 4);
  uint32_t h[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  const bool ext = ms.s0[0] != nullptr;   // the rows start outside buffer A: every row has to be moved
  int src = ext ? BUF_S0 : BUF_A;
  LeafLaunch ll0;
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
      hs_upload(c, lb.start.p, st.data(), m * 4);
      hs_upload(c, lb.len.p, ln.data(), m * 4);
      hs_upload(c, lb.shift.p, sh.data(), m * 4);
    };
    const size_t large_cap = 0;
    upload(leaf, lf_s, lf_l);
    upload(large, lg_s, lg_l, large_cap);
    large.base.alloc(c, std::max(large_cap, lg_s.size() ? lg_s.size() : (size_t)1) * 8);
    large.base.zero();
    {
      // the key span is known for the whole table: the first level spreads it over all its bins
      std::vector<uint32_t> mu(std::max<size_t>(lg_s.size(), 1), 0);
      if (key.span && key_bits > 0)
        for (size_t i = 0; i < lg_s.size(); i++) mu[i] = msd_mul_for(key.span, (uint32_t)key_bits, msd_bits(lg_l[i], (uint32_t)key_bits), lg_l[i]);
      large.mul.alloc(c, mu.size() * 4);
      hs_upload(c, large.mul.p, mu.data(), mu.size() * 4);
    }
    leaf.base.alloc(c, (lf_s.size() ? lf_s.size() : 1) * 8);
    leaf.base.zero();
    upload(small, sm_s, sm_l);
    upload(copy, cp_s, cp_l);
    hs_upload(c, counters.p, h, 32);
    hs_upload(c, tile_prefix.p, tp.data(), tp.size() * 4);
    run_leaf_and_copy(c, ms, key, src, leaf, h[0], copy, h[2], counters.as<uint32_t>(), ll0, &small, h[5]);
    launch_wide(c, ll0, counters.as<uint32_t>());
  } else {
    leaf.alloc(c, cap_leaf, true);
    large.alloc(c, cap_large, true);
    copy.alloc(c, cap_large);
    small.alloc(c, cap_leaf);
    counters.zero();
    PlanOut po{leaf.view(), large.view(), copy.view(), small.view(), counters.as<uint32_t>(), counters.as<uint32_t>()};
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
  level_loop(c, ms, key, large, tile_prefix, h[1], h[3], src, 0, cap_large, carry);
}

void hs_segmented_sort(hs_ctx* c, MoveSet& ms, const KeySrc& key, int key_bits, const uint32_t* d_seg_off,
                       int64_t nseg, int64_t n, const uint32_t* h_seg_off, SortCarry* carry) {
  segsort_impl(c, ms, key, key_bits, d_seg_off, nullptr, nullptr, nseg, n, h_seg_off, carry);
}
SortCarry* hs_sort_carry_new(hs_ctx* c, int64_t n) {
  SortCarry* k = new SortCarry();
  k->cap_large = (size_t)(n / LEAF_MAX + 1);
  k->large.alloc(c, k->cap_large, true);
  k->counters.alloc(c, 8 * 4);
  k->counters.zero();
  return k;
}
void hs_sort_carry_free(SortCarry* k) { delete k; }
void hs_sort_carry_finish(hs_ctx* c, MoveSet& ms, const KeySrc& key, SortCarry* k) {
  uint32_t h[8];
  memcpy(h, hs_read_wait(c, hs_read_begin(c, k->counters.p, 32)), 32);
  if (!h[1]) return;
  // the children were written into buffer B by the first level of their call
  DBuf tile_prefix(c, (k->cap_large + 1) * 4);
  LAUNCH(c, k_tile_prefix, 1, 1024, 0, k->large.view(), k->counters.as<uint32_t>(), tile_prefix.as<uint32_t>());
  level_loop(c, ms, key, k->large, tile_prefix, h[1], h[3], BUF_B, 1, k->cap_large, nullptr);
}
void hs_segmented_sort_list(hs_ctx* c, MoveSet& ms, const KeySrc& key, int key_bits, const uint32_t* d_start,
                            const uint32_t* d_len, int64_t nseg, int64_t n) {
  segsort_impl(c, ms, key, key_bits, nullptr, d_start, d_len, nseg, n, nullptr);
}
