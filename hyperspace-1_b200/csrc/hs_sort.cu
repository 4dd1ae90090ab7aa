// hs_sort.cu — K2 + K3: the row mover.
//  * bucket scatter: rows -> bucket-contiguous layout (stands in for Spark's ShuffleExchangeExec under
//    df.repartition(numBuckets, cols), HS/actions/CreateActionBase.scala:129-130)
//  * segmented sort: every bucket sorted on the normalised key (stands in for the SortExec that
//    FileFormatWriter plans for BucketSpec(n, cols, cols), HS/index/DataFrameWriterExtensions.scala:57-66)
// One tile kernel does both: rank the tile's rows by an up-to-11-bit digit with warp-private counters,
// stage each column through shared memory in digit order, write runs out coalesced. The sort is an MSD
// radix partition (8 bits per level, only segments larger than a tile) finished by an in-shared-memory
// LSD radix sort of each leaf segment over the bits that actually vary inside it.
#include "hs_block.cuh"
#include "hs_kernels.h"

#define LEAF_MAX PT_TILE
#define MSD_BITS 8
#define MSD_RADIX 256

struct SegList {           // device arrays
  uint32_t* start;
  uint32_t* len;
  uint32_t* shift;         // unsorted low key bits of the segment
};

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

// shared memory carve-up of the tile kernels
struct TileSmem {
  uint32_t* cnt;      // PT_WARPS * radix_pad
  uint32_t* tot;      // radix_pad
  uint32_t* loc;      // radix_pad
  uint32_t* gbase;    // radix_pad
  uint32_t* scratch;  // 16
  uint32_t* gdst;     // PT_TILE
  uint64_t* stage;    // PT_TILE
};
__device__ __forceinline__ TileSmem carve(unsigned char* base, int radix_pad) {
  TileSmem s;
  s.stage = (uint64_t*)base;
  s.gdst = (uint32_t*)(s.stage + PT_TILE);
  s.cnt = s.gdst + PT_TILE;
  s.tot = s.cnt + PT_WARPS * radix_pad;
  s.loc = s.tot + radix_pad;
  s.gbase = s.loc + radix_pad;
  s.scratch = s.gbase + radix_pad;
  return s;
}
static size_t tile_smem_bytes(int radix_pad) {
  return (size_t)PT_TILE * 8 + (size_t)PT_TILE * 4 + (size_t)(PT_WARPS + 3) * radix_pad * 4 + 64;
}

template <typename T>
__device__ __forceinline__ void move_column(const T* __restrict__ src, T* __restrict__ dst, uint32_t tile_start,
                                            uint32_t tile_len, const uint32_t (&pos)[PT_ITEMS], uint32_t vmask,
                                            T* stage, const uint32_t* gdst) {
  const int w = threadIdx.x >> 5;
  const uint32_t base = w * (PT_ITEMS * 32) + lane_id();
  T v[PT_ITEMS];
#pragma unroll
  for (int k = 0; k < PT_ITEMS; k++)
    if ((vmask >> k) & 1u) v[k] = src[tile_start + base + k * 32];
#pragma unroll
  for (int k = 0; k < PT_ITEMS; k++)
    if ((vmask >> k) & 1u) stage[pos[k]] = v[k];
  __syncthreads();
#pragma unroll
  for (int k = 0; k < PT_ITEMS; k++) {
    uint32_t j = threadIdx.x + k * PT_THREADS;
    if (j < tile_len) dst[gdst[j]] = stage[j];
  }
  __syncthreads();
}

struct PartParams {
  MoveArgs mv;
  KeySrc key;               // MODE 1
  const uint16_t* bucket;   // MODE 0
  uint32_t n;               // MODE 0: rows
  int32_t radix;            // MODE 0: number of buckets
  int32_t radix_pad;
  int32_t bits;             // MODE 0: bits of a bucket id
  uint32_t* cursor;         // MODE 0: [radix]; MODE 1: [nseg * 256]
  SegList segs;             // MODE 1: the large segments of this level
  const uint32_t* tile_prefix;  // MODE 1: [nseg + 1]
  uint32_t nseg;
};

__device__ __forceinline__ uint32_t find_segment(const uint32_t* tile_prefix, uint32_t nseg, uint32_t tile) {
  uint32_t lo = 0, hi = nseg;  // largest s with tile_prefix[s] <= tile
  while (hi - lo > 1) {
    uint32_t mid = (lo + hi) >> 1;
    if (tile_prefix[mid] <= tile) lo = mid; else hi = mid;
  }
  return lo;
}

// MODE 0: scatter by bucket id.  MODE 1: one MSD level inside large segments.
template <int MODE>
__global__ void __launch_bounds__(PT_THREADS, 2) k_partition(const __grid_constant__ PartParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  TileSmem sm = carve(smem_raw, p.radix_pad);
  uint32_t tile_start, tile_len, shift = 0;
  int bits = p.bits;
  uint32_t* cursor = p.cursor;
  if (MODE == 0) {
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
    bits = sh < MSD_BITS ? (int)sh : MSD_BITS;
    shift = sh - bits;
    cursor = p.cursor + (size_t)s * MSD_RADIX;
  }
  const int w = threadIdx.x >> 5;
  const uint32_t base = w * (PT_ITEMS * 32) + lane_id();
  uint32_t digit[PT_ITEMS], pos[PT_ITEMS];
  uint32_t vmask = 0;
#pragma unroll
  for (int k = 0; k < PT_ITEMS; k++) {
    uint32_t idx = base + k * 32;
    bool v = idx < tile_len;
    vmask |= (uint32_t)v << k;
    digit[k] = 0;
    if (v) {
      if (MODE == 0) digit[k] = p.bucket[tile_start + idx];
      else digit[k] = (uint32_t)(load_key(p.mv, p.key, tile_start + idx) >> shift) & ((1u << bits) - 1u);
    }
  }
  tile_rank<PT_ITEMS>(digit, vmask, PT_ITEMS, bits, p.radix_pad, sm.cnt, sm.tot, sm.loc, sm.scratch, pos);
  for (int d = threadIdx.x; d < p.radix_pad; d += PT_THREADS) {
    uint32_t t = sm.tot[d];
    sm.gbase[d] = t ? atomicAdd(&cursor[d], t) : 0;
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < PT_ITEMS; k++)
    if ((vmask >> k) & 1u) sm.gdst[pos[k]] = sm.gbase[digit[k]] + pos[k] - sm.loc[digit[k]];
  __syncthreads();
  for (int c = 0; c < p.mv.n; c++) {
    switch (p.mv.esize[c]) {
      case 8: move_column<uint64_t>((const uint64_t*)p.mv.src[c], (uint64_t*)p.mv.dst[c], tile_start, tile_len, pos, vmask, sm.stage, sm.gdst); break;
      case 4: move_column<uint32_t>((const uint32_t*)p.mv.src[c], (uint32_t*)p.mv.dst[c], tile_start, tile_len, pos, vmask, (uint32_t*)sm.stage, sm.gdst); break;
      case 2: move_column<uint16_t>((const uint16_t*)p.mv.src[c], (uint16_t*)p.mv.dst[c], tile_start, tile_len, pos, vmask, (uint16_t*)sm.stage, sm.gdst); break;
      default: move_column<uint8_t>((const uint8_t*)p.mv.src[c], (uint8_t*)p.mv.dst[c], tile_start, tile_len, pos, vmask, (uint8_t*)sm.stage, sm.gdst); break;
    }
  }
}

// per-segment digit histogram of one MSD level
__global__ void __launch_bounds__(PT_THREADS) k_msd_hist(const __grid_constant__ PartParams p) {
  __shared__ uint32_t hist[MSD_RADIX];
  uint32_t s = find_segment(p.tile_prefix, p.nseg, blockIdx.x);
  uint32_t t = blockIdx.x - p.tile_prefix[s];
  uint32_t slen = p.segs.len[s];
  uint32_t tile_start = p.segs.start[s] + t * PT_TILE;
  uint32_t left = slen - t * PT_TILE;
  uint32_t tile_len = left < PT_TILE ? left : PT_TILE;
  uint32_t sh = p.segs.shift[s];
  int bits = sh < MSD_BITS ? (int)sh : MSD_BITS;
  uint32_t shift = sh - bits;
  hist[threadIdx.x] = 0;
  __syncthreads();
  for (uint32_t j = threadIdx.x; j < tile_len; j += PT_THREADS) {
    uint32_t d = (uint32_t)(load_key(p.mv, p.key, tile_start + j) >> shift) & ((1u << bits) - 1u);
    atomicAdd(&hist[d], 1u);
  }
  __syncthreads();
  uint32_t v = hist[threadIdx.x];
  if (v) atomicAdd(&p.cursor[(size_t)s * MSD_RADIX + threadIdx.x], v);
}

struct PlanOut {
  SegList leaf, large, copy;
  uint32_t* counters;  // [0] leaf, [1] large, [2] copy, [3] tiles of large
};

__device__ __forceinline__ void classify_push(const PlanOut& o, uint32_t start, uint32_t len, uint32_t shift) {
  if (len == 0) return;
  if (len <= LEAF_MAX) {
    uint32_t i = atomicAdd(&o.counters[0], 1u);
    o.leaf.start[i] = start; o.leaf.len[i] = len; o.leaf.shift[i] = shift;
  } else if (shift == 0) {
    uint32_t i = atomicAdd(&o.counters[2], 1u);
    o.copy.start[i] = start; o.copy.len[i] = len; o.copy.shift[i] = 0;
  } else {
    uint32_t i = atomicAdd(&o.counters[1], 1u);
    o.large.start[i] = start; o.large.len[i] = len; o.large.shift[i] = shift;
    atomicAdd(&o.counters[3], (len + PT_TILE - 1) / PT_TILE);
  }
}

// turn the histograms of one level into write cursors and child segments
__global__ void __launch_bounds__(MSD_RADIX) k_msd_plan(SegList segs, uint32_t* cursor, PlanOut out) {
  __shared__ uint32_t scratch[PT_WARPS + 1];
  uint32_t s = blockIdx.x;
  uint32_t h = cursor[(size_t)s * MSD_RADIX + threadIdx.x];
  uint32_t ex = block_excl_scan_u32(h, scratch, nullptr);
  uint32_t start = segs.start[s] + ex;
  cursor[(size_t)s * MSD_RADIX + threadIdx.x] = start;
  uint32_t sh = segs.shift[s];
  uint32_t bits = sh < MSD_BITS ? sh : MSD_BITS;
  classify_push(out, start, h, sh - bits);
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
  classify_push(out, start, len, key_bits);
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
};

// Sort one segment of at most LEAF_MAX rows entirely in shared memory: LSD radix over the bits that
// vary inside the segment (min/max of its keys), 8 bits per pass, stable; then move every column.
__global__ void __launch_bounds__(PT_THREADS, 2) k_leaf_sort(const __grid_constant__ LeafParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  uint64_t* skey = (uint64_t*)smem_raw;                 // LEAF_MAX
  uint16_t* sidx = (uint16_t*)(skey + LEAF_MAX);        // LEAF_MAX
  uint32_t* cnt = (uint32_t*)(sidx + LEAF_MAX);         // PT_WARPS * 256
  uint32_t* tot = cnt + PT_WARPS * MSD_RADIX;
  uint32_t* loc = tot + MSD_RADIX;
  uint32_t* scratch = loc + MSD_RADIX;                  // 16
  uint64_t* red = (uint64_t*)(scratch + 16);            // 2 * PT_WARPS + 2

  const uint32_t s = blockIdx.x;
  const uint32_t start = p.segs.start[s], len = p.segs.len[s];
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
      key[k] = load_key(p.mv, p.key, start + idx);
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
    if (p.in_place) return;
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
  for (int c = 0; c < p.mv.n; c++) {
    const int es = p.mv.esize[c];
    uint64_t v[PT_ITEMS];
#pragma unroll
    for (int k = 0; k < PT_ITEMS; k++) {
      uint32_t j = threadIdx.x + k * PT_THREADS;
      if (j < len) {
        uint32_t r = start + sidx[j];
        if (es == 8) v[k] = ((const uint64_t*)p.mv.src[c])[r];
        else if (es == 4) v[k] = ((const uint32_t*)p.mv.src[c])[r];
        else if (es == 2) v[k] = ((const uint16_t*)p.mv.src[c])[r];
        else v[k] = ((const uint8_t*)p.mv.src[c])[r];
      }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < PT_ITEMS; k++) {
      uint32_t j = threadIdx.x + k * PT_THREADS;
      if (j < len) {
        uint32_t r = start + j;
        if (es == 8) ((uint64_t*)p.mv.dst[c])[r] = v[k];
        else if (es == 4) ((uint32_t*)p.mv.dst[c])[r] = (uint32_t)v[k];
        else if (es == 2) ((uint16_t*)p.mv.dst[c])[r] = (uint16_t)v[k];
        else ((uint8_t*)p.mv.dst[c])[r] = (uint8_t)v[k];
      }
    }
  }
}
static size_t leaf_smem_bytes() {
  return (size_t)LEAF_MAX * 8 + (size_t)LEAF_MAX * 2 + (size_t)(PT_WARPS + 2) * MSD_RADIX * 4 + 64 +
         (2 * PT_WARPS + 2) * 8;
}

// segments whose keys are all equal but that are larger than a leaf: plain copy src -> dst
__global__ void __launch_bounds__(256) k_copy_segments(MoveArgs mv, SegList segs) {
  uint32_t s = blockIdx.x;
  uint32_t start = segs.start[s], len = segs.len[s];
  for (int c = 0; c < mv.n; c++) {
    const int es = mv.esize[c];
    const uint8_t* src = (const uint8_t*)mv.src[c] + (size_t)start * es;
    uint8_t* dst = (uint8_t*)mv.dst[c] + (size_t)start * es;
    size_t bytes = (size_t)len * es;
    for (size_t i = threadIdx.x; i < bytes; i += blockDim.x) dst[i] = src[i];
  }
}

// ---------------------------------------------------------------- host drivers
static MoveArgs make_move(const MoveSet& ms, bool src_is_a, bool dst_is_a) {
  MoveArgs mv;
  memset(&mv, 0, sizeof mv);
  mv.n = ms.n;
  for (int i = 0; i < ms.n; i++) {
    mv.esize[i] = ms.esize[i];
    mv.src[i] = src_is_a ? ms.a[i] : ms.b[i];
    mv.dst[i] = dst_is_a ? ms.a[i] : ms.b[i];
  }
  return mv;
}

static int ilog2_ceil(uint32_t v) {
  int b = 0;
  while ((1u << b) < v) b++;
  return b;
}

void hs_bucket_scatter(hs_ctx* c, const void* const* src, const MoveSet& ms, const uint16_t* d_bucket, int64_t n,
                       int32_t nb, const uint32_t* d_bucket_off) {
  if (n == 0) return;
  if (nb > 2048) HS_THROW(HS_EUNSUPPORTED, "single-pass bucket scatter supports at most 2048 buckets per GPU");
  PartParams p;
  memset(&p, 0, sizeof p);
  p.mv.n = ms.n;
  for (int i = 0; i < ms.n; i++) { p.mv.esize[i] = ms.esize[i]; p.mv.src[i] = src[i]; p.mv.dst[i] = ms.a[i]; }
  p.bucket = d_bucket;
  p.n = (uint32_t)n;
  p.radix = nb;
  p.radix_pad = ((nb + PT_THREADS - 1) / PT_THREADS) * PT_THREADS;
  p.bits = ilog2_ceil((uint32_t)nb);
  DBuf cursor(c, (size_t)nb * 4);
  CK(cudaMemcpyAsync(cursor.p, d_bucket_off, (size_t)nb * 4, cudaMemcpyDeviceToDevice, c->stream));
  p.cursor = cursor.as<uint32_t>();
  size_t smem = tile_smem_bytes(p.radix_pad);
  CK(cudaFuncSetAttribute(k_partition<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  unsigned tiles = (unsigned)((n + PT_TILE - 1) / PT_TILE);
  LAUNCH(c, k_partition<0>, tiles, PT_THREADS, smem, p);
}

struct ListBufs {
  DBuf start, len, shift;
  void alloc(hs_ctx* c, size_t cap) {
    start.alloc(c, cap * 4); len.alloc(c, cap * 4); shift.alloc(c, cap * 4);
  }
  SegList view() const { return SegList{start.as<uint32_t>(), len.as<uint32_t>(), shift.as<uint32_t>()}; }
};

static void run_leaf_and_copy(hs_ctx* c, MoveSet& ms, const KeySrc& key, bool src_is_a, const ListBufs& leaf,
                              uint32_t n_leaf, const ListBufs& copy, uint32_t n_copy) {
  if (n_leaf) {
    LeafParams lp;
    memset(&lp, 0, sizeof lp);
    lp.mv = make_move(ms, src_is_a, true);
    lp.key = key;
    lp.segs = leaf.view();
    lp.in_place = src_is_a ? 1 : 0;
    size_t smem = leaf_smem_bytes();
    CK(cudaFuncSetAttribute(k_leaf_sort, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    Stage st(c, "leaf_sort");
    LAUNCH(c, k_leaf_sort, n_leaf, PT_THREADS, smem, lp);
  }
  if (n_copy && !src_is_a) {
    MoveArgs mv = make_move(ms, false, true);
    Stage st(c, "copy_segments");
    LAUNCH(c, k_copy_segments, n_copy, 256, 0, mv, copy.view());
  }
}

static void segsort_impl(hs_ctx* c, MoveSet& ms, const KeySrc& key, int key_bits, const uint32_t* d_seg_off,
                         const uint32_t* d_start, const uint32_t* d_len, int64_t nseg, int64_t n) {
  if (n == 0 || nseg == 0) return;
  // level 0 lists
  size_t cap_leaf = (size_t)nseg, cap_large = (size_t)(n / LEAF_MAX + 1);
  ListBufs leaf, large, copy;
  leaf.alloc(c, cap_leaf);
  large.alloc(c, cap_large);
  copy.alloc(c, cap_large);
  DBuf counters(c, 8 * 4);
  counters.zero();
  PlanOut po{leaf.view(), large.view(), copy.view(), counters.as<uint32_t>()};
  hs_stage_begin(c, "sort_plan");
  LAUNCH(c, k_classify, (unsigned)((nseg + 255) / 256), 256, 0, d_seg_off, d_start, d_len, nseg, (uint32_t)key_bits, po);
  DBuf tile_prefix(c, (cap_large + 1) * 4);
  LAUNCH(c, k_tile_prefix, 1, 1024, 0, large.view(), counters.as<uint32_t>(), tile_prefix.as<uint32_t>());
  uint32_t h[4];
  CK(cudaMemcpyAsync(h, counters.p, 16, cudaMemcpyDeviceToHost, c->stream));
  hs_stage_end(c);
  hs_sync(c);
  bool src_is_a = true;
  run_leaf_and_copy(c, ms, key, src_is_a, leaf, h[0], copy, h[2]);
  uint32_t n_large = h[1], n_tiles = h[3];
  int level = 0;
  while (n_large > 0) {
    if (++level > 10) HS_THROW(HS_ECUDA, "segmented sort did not converge");
    // histograms -> cursors + child lists
    DBuf cursor(c, (size_t)n_large * MSD_RADIX * 4);
    cursor.zero();
    PartParams p;
    memset(&p, 0, sizeof p);
    p.mv = make_move(ms, src_is_a, !src_is_a);
    p.key = key;
    p.radix = MSD_RADIX;
    p.radix_pad = MSD_RADIX;
    p.bits = MSD_BITS;
    p.cursor = cursor.as<uint32_t>();
    p.segs = large.view();
    p.tile_prefix = tile_prefix.as<uint32_t>();
    p.nseg = n_large;
    hs_stage_begin(c, "msd_hist");
    LAUNCH(c, k_msd_hist, n_tiles, PT_THREADS, 0, p);
    hs_stage_end(c);
    ListBufs nleaf, nlarge, ncopy;
    size_t children = (size_t)n_large * MSD_RADIX;
    nleaf.alloc(c, children);
    nlarge.alloc(c, cap_large);
    ncopy.alloc(c, cap_large);
    DBuf ncounters(c, 8 * 4);
    ncounters.zero();
    PlanOut npo{nleaf.view(), nlarge.view(), ncopy.view(), ncounters.as<uint32_t>()};
    hs_stage_begin(c, "msd_plan");
    LAUNCH(c, k_msd_plan, n_large, MSD_RADIX, 0, large.view(), cursor.as<uint32_t>(), npo);
    hs_stage_end(c);
    size_t smem = tile_smem_bytes(MSD_RADIX);
    CK(cudaFuncSetAttribute(k_partition<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    hs_stage_begin(c, "msd_scatter");
    LAUNCH(c, k_partition<1>, n_tiles, PT_THREADS, smem, p);
    hs_stage_end(c);
    DBuf ntile_prefix(c, (cap_large + 1) * 4);
    hs_stage_begin(c, "msd_plan");
    LAUNCH(c, k_tile_prefix, 1, 1024, 0, nlarge.view(), ncounters.as<uint32_t>(), ntile_prefix.as<uint32_t>());
    CK(cudaMemcpyAsync(h, ncounters.p, 16, cudaMemcpyDeviceToHost, c->stream));
    hs_stage_end(c);
    hs_sync(c);
    src_is_a = !src_is_a;
    run_leaf_and_copy(c, ms, key, src_is_a, nleaf, h[0], ncopy, h[2]);
    // the lists of this level stay alive until the kernels above are queued; stream order makes the
    // frees below safe
    large = std::move(nlarge);
    leaf = std::move(nleaf);
    copy = std::move(ncopy);
    tile_prefix = std::move(ntile_prefix);
    n_large = h[1];
    n_tiles = h[3];
  }
}

void hs_segmented_sort(hs_ctx* c, MoveSet& ms, const KeySrc& key, int key_bits, const uint32_t* d_seg_off,
                       int64_t nseg, int64_t n) {
  segsort_impl(c, ms, key, key_bits, d_seg_off, nullptr, nullptr, nseg, n);
}
void hs_segmented_sort_list(hs_ctx* c, MoveSet& ms, const KeySrc& key, int key_bits, const uint32_t* d_start,
                            const uint32_t* d_len, int64_t nseg, int64_t n) {
  segsort_impl(c, ms, key, key_bits, nullptr, d_start, d_len, nseg, n);
}
