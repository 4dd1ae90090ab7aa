// hs_sort_shared.cuh — what the two translation units of the row mover share: hs_sort.cu (bucket scatter, MSD
// levels; 512 threads x 8 rows per tile, two tiles per SM) and hs_leaf.cu (the leaf sorts; 384 threads x 8 rows,
// THREE leaves per SM). The leaf kernels are chains of dependent steps per segment, so one more resident chain
// per SM pays (measured: 1.68 -> 1.3 ms per 100 M rows), while the scatter passes want the longer runs per digit
// of the larger tile; the geometry macros of hs_block.cuh are therefore set per translation unit and nothing in
// this header depends on them.
#pragma once
#include <algorithm>

#include "hs_common.cuh"
#include "hs_kernels.h"

#define LEAF_MAX 3072      // rows a leaf segment holds at most (= the tile of hs_leaf.cu)
#ifndef MSD_BITS
#define MSD_BITS 8
#endif
#define MSD_RADIX (1 << MSD_BITS)
#define MSD_TARGET_ROWS (3400 * LEAF_MAX / 4096)   // aimed-at child size of an MSD level (a leaf holds at most LEAF_MAX rows)

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
  uint64_t* base;          // subtracted from the key before digits are taken
  uint32_t* mul;           // large segments: digit multiplier (msd_digit), 0 = plain top bits
};

// The digit of an MSD level. A key range rarely fills a power of two (cfg2: 100 M keys in 27 bits use 191 of the
// 256 top-8-bit values, so the children are 34 % larger than len / 256 and many miss the leaf size). A level
// therefore looks at MSD_EXTRA more bits than its digit has — `cells` — and, when the segment's key span is known,
// maps the cells that can occur onto all 2^bits digits with one multiplication: digit = cell * mul >> 16. The
// cells of a digit are contiguous, so a child is still described by (base, bits left): msd_cell_lo gives its base.
#define MSD_EXTRA 4
__host__ __device__ __forceinline__ int msd_extra(uint32_t r, int bits) {
  const int e = (int)r - bits;
  return e < MSD_EXTRA ? e : MSD_EXTRA;
}
__host__ __device__ __forceinline__ uint32_t msd_digit(uint64_t rel, uint32_t r, int bits, uint32_t mul) {   // rel = key - base
  const int e = msd_extra(r, bits);
  const uint32_t cell = (uint32_t)(rel >> (r - bits - e)) & ((1u << (bits + e)) - 1u);
  return mul ? (uint32_t)(((uint64_t)cell * mul) >> 16) : cell >> e;
}
__host__ __device__ __forceinline__ uint32_t msd_cell_lo(uint32_t d, int e, uint32_t mul) {   // first cell of digit d
  return mul ? (uint32_t)((((uint64_t)d << 16) + mul - 1) / mul) : d << e;
}
// multiplier for a segment of `len` rows whose keys (minus base) lie in [0, span]: the cells in use are spread over
// just as many digits as bring the expected child to MSD_TARGET_ROWS (a leaf costs about the same whether it is
// full or half full, so children should be as large as a leaf allows); 0 when plain top bits do the same
__host__ __device__ __forceinline__ uint32_t msd_mul_for(uint64_t span, uint32_t r, int bits, uint32_t len) {
  const int e = msd_extra(r, bits);
  const uint64_t ncells = (span >> (r - bits - e)) + 1;
  uint32_t nbins = (len + MSD_TARGET_ROWS - 1) / MSD_TARGET_ROWS;
  if (nbins > (1u << bits)) nbins = 1u << bits;
  if (nbins < 2) nbins = 2;
  if (nbins == (1u << bits) && ncells >= (1ull << (bits + e))) return 0;
  return (uint32_t)(((uint64_t)nbins << 16) / ncells);
}

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

#define SMALL_MAX 64   // segments up to this many rows are sorted by one warp each (k_leaf_small)

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

#define BUF_A 0
#define BUF_B 1
#define BUF_S0 2
static inline void* const* buf_of(const MoveSet& ms, int code) { return code == BUF_A ? ms.a : code == BUF_B ? ms.b : ms.s0; }
static inline int other_buf(int code) { return code == BUF_B ? BUF_A : BUF_B; }
static inline MoveArgs make_move(const MoveSet& ms, int src_code, int dst_code) {
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

struct ListBufs {
  DBuf start, len, shift, base, mul;
  void alloc(hs_ctx* c, size_t cap, bool with_base = false) {
    start.alloc(c, cap * 4); len.alloc(c, cap * 4); shift.alloc(c, cap * 4);
    if (with_base) { base.alloc(c, cap * 8); base.zero(); mul.alloc(c, cap * 4); mul.zero(); }
  }
  SegList view() const {
    return SegList{start.as<uint32_t>(), len.as<uint32_t>(), shift.as<uint32_t>(), base.as<uint64_t>(), mul.as<uint32_t>()};
  }
};

// state of one leaf launch that the 64-bit kernel (segments the packed kernel flagged) needs afterwards
struct LeafLaunch {
  DBuf wide;
  LeafParams lp;
  uint32_t max_leaf = 0;
};

// hs_leaf.cu
void hs_leaf_set_attrs();
void run_leaf_and_copy(hs_ctx* c, MoveSet& ms, const KeySrc& key, int src_code, const ListBufs& leaf, uint32_t max_leaf,
                       const ListBufs& copy, uint32_t max_copy, uint32_t* counters, LeafLaunch& ll,
                       const ListBufs* small = nullptr, uint32_t max_small = 0);
void launch_wide(hs_ctx* c, LeafLaunch& ll, uint32_t* d_counters);
