// hs_parquet.cu — K4: the bodies of Parquet v1 data pages, PLAIN encoding, built on the device.
// Stands where Spark's DynamicPartitionDataWriter + ParquetFileFormat stand behind saveWithBuckets
// (HS/index/DataFrameWriterExtensions.scala:58-65): the rows of a bucket become column chunks of pages. The page
// and file framing (Thrift headers, footer) is a few hundred bytes per file and is written by the host
// (hyperspace-1_b200/parquet_plain.py); what touches every value happens here:
//   * definition levels of an OPTIONAL column: RLE/bit-packed hybrid of bit width 1 — one RLE run for a page
//     without NULLs, otherwise bit-packed groups, which are the validity bitmap re-aligned to the page start;
//   * PLAIN values of the non-NULL rows: fixed-width values compacted, booleans re-packed, strings as
//     4-byte little-endian length + bytes.
// One call encodes all pages of one column: per-row destinations come from two exclusive scans (valid rows,
// valid string bytes), so every value is read once and written once.
#include <algorithm>

#include "hs_common.cuh"

// flag[i] = row i is valid; len[i] = bytes of its PLAIN image (0 for NULL rows)
template <typename TO>
__global__ void __launch_bounds__(256) k_pq_flags(KCol c, int64_t n, int fixed_w, uint8_t* __restrict__ flag,
                                                  uint32_t* __restrict__ len) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    const bool v = !c.has_nulls || kvalid(c.validity, i);
    flag[i] = v;
    if (len) {
      uint32_t l = 0;
      if (v) {
        if (fixed_w) l = (uint32_t)fixed_w;
        else {
          const TO* o = (const TO*)c.offsets;
          l = 4u + (uint32_t)(o[i + 1] - o[i]);
        }
      }
      len[i] = l;
    }
  }
}

// values at the page boundaries of the two scans (page_rows has npages + 1 entries; the entry for row n is the total)
__global__ void k_pq_pick(const uint32_t* __restrict__ vpos, const uint32_t* __restrict__ bpos, const uint64_t* vtotal,
                          const uint64_t* btotal, const int64_t* __restrict__ page_rows, int npages, int64_t n,
                          uint64_t* __restrict__ out) {
  int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p > npages) return;
  const int64_t r = page_rows[p];
  out[2 * p] = r < n ? vpos[r] : *vtotal;
  out[2 * p + 1] = bpos ? (r < n ? bpos[r] : *btotal) : 0;
}

struct PqPages {
  int32_t npages;
  const int64_t* page_rows;     // device, npages + 1
  const uint64_t* val_off;      // device, npages: byte offset of the page's values inside the output
  const uint64_t* def_off;      // device, npages: byte offset of the page's bit-packed def-level bytes (or ~0)
  const uint64_t* first;        // device, 2 * (npages + 1): (valid rows, valid bytes) before the page
};
__device__ __forceinline__ int pq_page_of(const PqPages& pg, int64_t row) {
  int lo = 0, hi = pg.npages;   // largest p with page_rows[p] <= row
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (pg.page_rows[mid] <= row) lo = mid; else hi = mid;
  }
  return lo;
}

// PLAIN values, fixed width w: value of valid row i -> val_off[p] + (vpos[i] - first_valid[p]) * w
__global__ void __launch_bounds__(256) k_pq_fixed(const uint8_t* __restrict__ data, int w, const uint8_t* __restrict__ flag,
                                                  const uint32_t* __restrict__ vpos, int64_t n, PqPages pg,
                                                  uint8_t* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    if (!flag[i]) continue;
    const int p = pq_page_of(pg, i);
    if (pg.val_off[p] == ~0ull) continue;        // dictionary-encoded page
    uint8_t* d = out + pg.val_off[p] + (uint64_t)(vpos[i] - pg.first[2 * p]) * w;
    const uint8_t* s = data + (size_t)i * w;
    // page bodies start at arbitrary byte offsets: copy by bytes unless both ends happen to be aligned
    if (w == 8 && (((uintptr_t)d) & 7) == 0) *(uint64_t*)d = *(const uint64_t*)s;
    else if (w == 4 && (((uintptr_t)d) & 3) == 0) *(uint32_t*)d = *(const uint32_t*)s;
    else for (int b = 0; b < w; b++) d[b] = s[b];
  }
}
// PLAIN booleans: bit (vpos[i] - first) of the page's value bytes; one thread per OUTPUT byte would need a
// search per bit, so set bits with atomics on zeroed output (booleans are rare and small)
__global__ void __launch_bounds__(256) k_pq_bool(const uint8_t* __restrict__ data, const uint8_t* __restrict__ flag,
                                                 const uint32_t* __restrict__ vpos, int64_t n, PqPages pg,
                                                 uint8_t* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    if (!flag[i] || !((data[i >> 3] >> (i & 7)) & 1)) continue;
    const int p = pq_page_of(pg, i);
    const uint64_t bit = vpos[i] - pg.first[2 * p];
    const uint64_t byte = pg.val_off[p] + (bit >> 3);
    atomicOr((unsigned int*)(out + (byte & ~(uint64_t)3)), (1u << (bit & 7)) << (8 * (byte & 3)));
  }
}
// PLAIN byte arrays: 4-byte LE length + bytes; 8 lanes per row
template <typename TO>
__global__ void __launch_bounds__(256) k_pq_bytes(const TO* __restrict__ off, const uint8_t* __restrict__ data,
                                                  const uint8_t* __restrict__ flag, const uint32_t* __restrict__ bpos,
                                                  int64_t n, PqPages pg, uint8_t* __restrict__ out) {
  const int sub = threadIdx.x & 7;
  int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 3;
  const int64_t stride = ((int64_t)gridDim.x * blockDim.x) >> 3;
  for (; i < n; i += stride) {
    if (!flag[i]) continue;
    const int p = pq_page_of(pg, i);
    const int64_t s0 = (int64_t)off[i];
    const uint32_t len = (uint32_t)((int64_t)off[i + 1] - s0);
    uint8_t* d = out + pg.val_off[p] + (uint64_t)(bpos[i] - pg.first[2 * p + 1]);
    if (sub < 4) d[sub] = (uint8_t)(len >> (8 * sub));
    const uint8_t* s = data + s0;
    for (uint32_t b = sub; b < len; b += 8) d[4 + b] = s[b];
  }
}
// definition levels of the pages that hold NULLs: byte g of the page = validity bits of its rows 8g .. 8g+7
__global__ void __launch_bounds__(256) k_pq_deflevels(const uint8_t* __restrict__ flag, int64_t n, PqPages pg,
                                                      uint8_t* __restrict__ out) {
  // one thread per (page, group) pair, pages walked by the grid's y dimension
  const int p = blockIdx.y;
  if (pg.def_off[p] == ~0ull) return;
  const int64_t r0 = pg.page_rows[p], r1 = pg.page_rows[p + 1];
  const int64_t groups = (r1 - r0 + 7) >> 3;
  for (int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; g < groups; g += (int64_t)gridDim.x * blockDim.x) {
    unsigned v = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
      const int64_t r = r0 + g * 8 + k;
      if (r < r1 && flag[r]) v |= 1u << k;
    }
    out[pg.def_off[p] + g] = (uint8_t)v;
  }
}

// ------------------------------------------------------------------ dictionary encoding (PLAIN_DICTIONARY)
// What parquet-mr does first for every column chunk: values go through a dictionary as long as it stays small.
// Per chunk (= bucket file) an open-addressing table in global memory collects the distinct values (atomicCAS on
// the 64-bit slot); a chunk whose table passes max_dict entries falls back to PLAIN. Dictionary-encoded pages hold
// [definition levels][bit width][one bit-packed run of the indices].
#define PQD_EMPTY 0xFFFFFFFFFFFFFFFFull
struct PqDict {
  int32_t nchunks, slots, max_dict, w;       // slots: power of two >= 2 * max_dict
  unsigned long long* keys;                  // [nchunks * slots]
  uint32_t* index;                           // [nchunks * slots]: dictionary index of the slot's value
  uint32_t* count;                           // [nchunks] distinct values seen
  uint32_t* overflow;                        // [nchunks] 1 = too many distinct values: PLAIN
  uint32_t* has_empty;                       // [nchunks] the value whose bits equal PQD_EMPTY occurs (8-byte types)
  const int32_t* page_chunk;                 // [npages]
};
__device__ __forceinline__ uint64_t pqd_load(const uint8_t* data, int w, int64_t i) {
  return w == 8 ? ((const uint64_t*)data)[i] : (uint64_t)((const uint32_t*)data)[i];
}
__device__ __forceinline__ uint32_t pqd_hash(uint64_t v) {
  v ^= v >> 33; v *= 0xff51afd7ed558ccdull; v ^= v >> 29;
  return (uint32_t)v;
}
__global__ void __launch_bounds__(256) k_pqd_insert(const uint8_t* __restrict__ data, const uint8_t* __restrict__ flag, int64_t n,
                                                    PqPages pg, PqDict d) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    if (!flag[i]) continue;
    const int c = d.page_chunk[pq_page_of(pg, i)];
    if (d.overflow[c]) continue;
    const uint64_t v = pqd_load(data, d.w, i);
    if (v == PQD_EMPTY) { d.has_empty[c] = 1; continue; }
    unsigned long long* tab = d.keys + (size_t)c * d.slots;
    uint32_t h = pqd_hash(v) & (uint32_t)(d.slots - 1);
    for (int probe = 0; probe < d.slots; probe++) {
      const unsigned long long old = atomicCAS(&tab[h], (unsigned long long)PQD_EMPTY, (unsigned long long)v);
      if (old == PQD_EMPTY) {
        if (atomicAdd(&d.count[c], 1u) >= (uint32_t)d.max_dict) d.overflow[c] = 1;
        break;
      }
      if (old == v) break;
      h = (h + 1) & (uint32_t)(d.slots - 1);
      if (probe == d.slots - 1) d.overflow[c] = 1;
    }
  }
}
// one CTA per chunk: used slots get consecutive indices (slot order); the dictionary page body is written
__global__ void __launch_bounds__(256) k_pqd_assign(PqDict d, const uint64_t* __restrict__ dict_off, uint8_t* __restrict__ out) {
  __shared__ uint32_t s_base, s_warp[8];
  const int c = blockIdx.x;
  if (dict_off[c] == ~0ull) return;
  const unsigned long long* tab = d.keys + (size_t)c * d.slots;
  uint32_t* idx = d.index + (size_t)c * d.slots;
  uint8_t* body = out + dict_off[c];
  if (threadIdx.x == 0) s_base = 0;
  __syncthreads();
  for (int s0 = 0; s0 < d.slots; s0 += 256) {
    const int s = s0 + threadIdx.x;
    const bool used = s < d.slots && tab[s] != PQD_EMPTY;
    const uint32_t bal = __ballot_sync(0xffffffffu, used);
    if (lane_id() == 0) s_warp[threadIdx.x >> 5] = __popc(bal);
    __syncthreads();
    uint32_t before = s_base;
    for (int w = 0; w < (int)(threadIdx.x >> 5); w++) before += s_warp[w];
    const uint32_t my = before + __popc(bal & ((1u << lane_id()) - 1u));
    if (used) {
      idx[s] = my;
      const uint64_t v = tab[s];
      for (int b = 0; b < d.w; b++) body[(size_t)my * d.w + b] = (uint8_t)(v >> (8 * b));
    }
    __syncthreads();
    if (threadIdx.x == 0) { uint32_t t = 0; for (int w = 0; w < 8; w++) t += s_warp[w]; s_base += t; }
    __syncthreads();
  }
  if (threadIdx.x == 0 && d.has_empty[c]) {   // the all-ones value takes the last index
    for (int b = 0; b < d.w; b++) body[(size_t)s_base * d.w + b] = 0xFF;
  }
}
// index of every non-NULL value of a dictionary-encoded chunk, in row order
__global__ void __launch_bounds__(256) k_pqd_lookup(const uint8_t* __restrict__ data, const uint8_t* __restrict__ flag,
                                                    const uint32_t* __restrict__ vpos, int64_t n, PqPages pg, PqDict d,
                                                    const uint64_t* __restrict__ dict_off, uint32_t* __restrict__ idxv) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    if (!flag[i]) continue;
    const int c = d.page_chunk[pq_page_of(pg, i)];
    if (dict_off[c] == ~0ull) continue;
    const uint64_t v = pqd_load(data, d.w, i);
    uint32_t r;
    if (v == PQD_EMPTY) r = d.count[c];
    else {
      const unsigned long long* tab = d.keys + (size_t)c * d.slots;
      uint32_t h = pqd_hash(v) & (uint32_t)(d.slots - 1);
      while (tab[h] != v) h = (h + 1) & (uint32_t)(d.slots - 1);
      r = d.index[(size_t)c * d.slots + h];
    }
    idxv[vpos[i]] = r;
  }
}
// bit-packed groups of the dictionary-encoded pages: one thread per output byte
__global__ void __launch_bounds__(256) k_pqd_pack(const uint32_t* __restrict__ idxv, const uint64_t* __restrict__ pack_off,
                                                  const uint32_t* __restrict__ pack_bytes, const uint32_t* __restrict__ bw_of,
                                                  PqPages pg, uint8_t* __restrict__ out) {
  const int p = blockIdx.y;
  if (pack_off[p] == ~0ull) return;
  const uint32_t bw = bw_of[p], nbytes = pack_bytes[p];
  const uint64_t v0 = pg.first[2 * p], v1 = pg.first[2 * (p + 1)];   // non-NULL values before / through the page
  for (uint32_t b = blockIdx.x * blockDim.x + threadIdx.x; b < nbytes; b += gridDim.x * blockDim.x) {
    uint32_t byte = 0;
    const uint64_t bit0 = (uint64_t)b * 8;
    for (uint64_t j = bit0 / bw; j * bw < bit0 + 8; j++) {
      const uint64_t x = v0 + j < v1 ? idxv[v0 + j] : 0;        // padding of the last group: zeros
      const int64_t sh = (int64_t)(j * bw) - (int64_t)bit0;     // position of the value's bit 0 inside this byte
      byte |= sh >= 0 ? (uint32_t)(x << sh) : (uint32_t)(x >> (-sh));
    }
    out[pack_off[p] + b] = (uint8_t)byte;
  }
}

// ------------------------------------------------------------------ snappy (codec 1)
// Page bodies compressed on the device, in the raw snappy format parquet-mr's SnappyCodec writes. A body is cut
// into fragments of SN_FRAG bytes, each compressed by one warp on its own (matches never reach back past the
// fragment start, as in the reference compressor, which also resets its table per block — only the block is
// 64 KiB there): the fragment sits in shared memory next to a table of the last position of every 4-byte hash;
// per round the 32 lanes test 32 consecutive positions against the table, the first hit is extended by all lanes
// together, and literals / copies are written by all lanes. Fragments land in worst-case-sized slots; a second
// kernel packs them behind the page's length preamble.
#define SN_FRAG 8192
#define SN_HT_BITS 11
#define SN_WARPS 8
#define SN_SLOT (SN_FRAG + SN_FRAG / 6 + 64)
struct SnFrag { uint64_t src; uint32_t len; uint32_t pad; };
struct SnWarp {
  uint8_t in[SN_FRAG + 32];           // the fragment at in[src & 15] (16-byte loads keep their alignment)
  uint16_t ht[1 << SN_HT_BITS];
};
__device__ __forceinline__ uint32_t sn_ld32(const uint8_t* base, uint32_t a) {   // base is 16-byte aligned
  const uint32_t* w = (const uint32_t*)base;
  return __funnelshift_r(w[a >> 2], w[(a >> 2) + 1], (a & 3) * 8);
}
__device__ __forceinline__ uint32_t sn_literal(uint8_t* out, uint32_t o, const uint8_t* in, uint32_t start, uint32_t len, int lane) {
  if (len == 0) return o;
  const uint32_t n1 = len - 1;
  uint32_t hdr;
  if (n1 < 60) { if (lane == 0) out[o] = (uint8_t)(n1 << 2); hdr = 1; }
  else if (n1 < 256) { if (lane == 0) { out[o] = 60 << 2; out[o + 1] = (uint8_t)n1; } hdr = 2; }
  else { if (lane == 0) { out[o] = 61 << 2; out[o + 1] = (uint8_t)n1; out[o + 2] = (uint8_t)(n1 >> 8); } hdr = 3; }
  for (uint32_t i = lane; i < len; i += 32) out[o + hdr + i] = in[start + i];
  return o + hdr + len;
}
__device__ __forceinline__ uint32_t sn_copy(uint8_t* out, uint32_t o, uint32_t off, uint32_t len, int lane) {
  if (len <= 11 && off < 2048) {      // 2-byte form: length 4..11, 11-bit offset
    if (lane == 0) { out[o] = (uint8_t)(((off >> 8) << 5) | ((len - 4) << 2) | 1); out[o + 1] = (uint8_t)off; }
    return o + 2;
  }
  const uint32_t nel = (len + 63) >> 6;   // 3-byte form: length 1..64, 16-bit offset
  for (uint32_t k = lane; k < nel; k += 32) {
    const uint32_t l = k + 1 < nel ? 64 : len - 64 * (nel - 1);
    out[o + 3 * k] = (uint8_t)(((l - 1) << 2) | 2);
    out[o + 3 * k + 1] = (uint8_t)off;
    out[o + 3 * k + 2] = (uint8_t)(off >> 8);
  }
  return o + 3 * nel;
}
__global__ void __launch_bounds__(SN_WARPS * 32) k_snappy_frag(const uint8_t* __restrict__ body, const SnFrag* __restrict__ frags,
                                                               int nfrag, uint8_t* __restrict__ scratch, uint32_t* __restrict__ sizes) {
  extern __shared__ __align__(16) uint8_t sn_smem[];
  const int lane = lane_id(), f = blockIdx.x * SN_WARPS + (threadIdx.x >> 5);
  if (f >= nfrag) return;
  SnWarp* sw = (SnWarp*)sn_smem + (threadIdx.x >> 5);
  const SnFrag fr = frags[f];
  const uint32_t len = fr.len, skew = (uint32_t)(fr.src & 15);
  const uint4* g = (const uint4*)(body + (fr.src - skew));
  for (uint32_t i = lane; i * 16 < skew + len; i += 32) ((uint4*)sw->in)[i] = g[i];
  for (uint32_t i = lane; i < (1u << SN_HT_BITS); i += 32) sw->ht[i] = 0xFFFF;
  __syncwarp();
  const uint8_t* base = sw->in;          // 16-byte aligned; fragment byte i is base[skew + i]
  const uint8_t* in = base + skew;
  uint8_t* out = scratch + (size_t)f * SN_SLOT;
  uint32_t pos = 0, lit = 0, o = 0;
  while (pos + 4 <= len) {
    const uint32_t p = pos + lane;
    const bool can = p + 4 <= len;
    const uint32_t w = can ? sn_ld32(base, skew + p) : 0;
    const uint32_t h = (w * 0x1e35a7bdu) >> (32 - SN_HT_BITS);
    const uint32_t cand = can ? sw->ht[h] : 0xFFFFu;
    __syncwarp();
    if (can) sw->ht[h] = (uint16_t)p;
    __syncwarp();
    const bool hit = can && cand < p && sn_ld32(base, skew + cand) == w;
    const uint32_t bal = __ballot_sync(0xffffffffu, hit);
    if (!bal) { pos += 32; continue; }
    const int first = __ffs(bal) - 1;
    const uint32_t m = pos + first, c = __shfl_sync(0xffffffffu, cand, first);
    uint32_t mlen = 4;
    for (;;) {
      const uint32_t i = m + mlen + lane;
      const bool same = i < len && in[i] == in[c + mlen + lane];
      const uint32_t diff = __ballot_sync(0xffffffffu, !same);
      if (diff) { mlen += __ffs(diff) - 1; break; }
      mlen += 32;
    }
    o = sn_literal(out, o, in, lit, m - lit, lane);
    o = sn_copy(out, o, m - c, mlen, lane);
    pos = m + mlen;
    lit = pos;
  }
  o = sn_literal(out, o, in, lit, len - lit, lane);
  if (lane == 0) sizes[f] = o;
}
// fragment f of the scratch slots -> its place in the packed output
__global__ void __launch_bounds__(256) k_snappy_pack(const uint8_t* __restrict__ scratch, const uint32_t* __restrict__ sizes,
                                                     const uint64_t* __restrict__ dst, int nfrag, uint8_t* __restrict__ out) {
  const int lane = lane_id(), f = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (f >= nfrag) return;
  const uint8_t* s = scratch + (size_t)f * SN_SLOT;
  uint8_t* d = out + dst[f];
  for (uint32_t i = lane; i < sizes[f]; i += 32) d[i] = s[i];
}

static int varint(uint64_t v, uint8_t* out) {
  int n = 0;
  while (v >= 0x80) { out[n++] = (uint8_t)(v | 0x80); v >>= 7; }
  out[n++] = (uint8_t)v;
  return n;
}

// page_chunk / nchunks / max_dict: dictionary encoding per column chunk (max_dict = 0: PLAIN only)
static void pq_encode(hs_ctx* ctx, const hs_column* col, int32_t optional, const int64_t* page_rows, int32_t npages,
                      const int32_t* page_chunk, int32_t nchunks, int32_t max_dict, int32_t codec, hs_table** out) {
  if (!out) HS_THROW(HS_EINVAL, "null out");
  *out = nullptr;
  if (codec != 0 && codec != 1) HS_THROW(HS_EUNSUPPORTED, "codec must be 0 (UNCOMPRESSED) or 1 (SNAPPY)");
  if (!col || !page_rows || npages <= 0 || npages > 65535) HS_THROW(HS_EINVAL, "bad page list (1..65535 pages per call)");
  if (max_dict < 0 || max_dict > 32768) HS_THROW(HS_EINVAL, "max_dict is at most 32768 entries");
  if (page_chunk) {
    if (nchunks <= 0 || nchunks > npages) HS_THROW(HS_EINVAL, "nchunks must be in 1..npages");
    for (int p = 0; p < npages; p++)
      if (page_chunk[p] < 0 || page_chunk[p] >= nchunks || (p && page_chunk[p] < page_chunk[p - 1]))
        HS_THROW(HS_EINVAL, "page_chunk must be non-decreasing and inside [0, nchunks)");
  } else {
    if (max_dict > 0) HS_THROW(HS_EINVAL, "dictionary encoding needs the chunk of every page");
    nchunks = 0;
  }
  hs_stages_reset(ctx);
  DevCol dc;
  {
    Stage s(ctx, "h2d");
    hs_stage_column(ctx, *col, dc);
  }
  const int64_t n = dc.length;
  if (page_rows[0] != 0 || page_rows[npages] != n) HS_THROW(HS_EINVAL, "pages must cover the column");
  for (int p = 0; p < npages; p++)
    if (page_rows[p + 1] < page_rows[p]) HS_THROW(HS_EINVAL, "page boundaries must not decrease");
  if (!optional && dc.validity) HS_THROW(HS_EINVAL, "a REQUIRED column cannot hold NULLs");
  const bool varlen = hs_is_varlen(dc.type), large = hs_is_large(dc.type), is_bool = dc.type == HS_BOOL;
  const int w = hs_fixed_width(dc.type);
  if (varlen && dc.data_bytes + 4 * n >= 0xFFFFFFF0ll) HS_THROW(HS_EUNSUPPORTED, "more than 4 GiB of page data in one call");
  const KCol kc = hs_kcol(dc);
  const int64_t n1 = n > 0 ? n : 1;
  DBuf flag(ctx, (size_t)n1), vpos(ctx, (size_t)n1 * 4), vtotal(ctx, 8), len, bpos, btotal(ctx, 8);
  DBuf d_rows(ctx, ((size_t)npages + 1) * 8), picks(ctx, ((size_t)npages + 1) * 16);
  std::vector<uint64_t> h_first(2 * ((size_t)npages + 1));
  {
    Stage s(ctx, "parquet_plan");
    CK(cudaMemcpyAsync(d_rows.p, page_rows, ((size_t)npages + 1) * 8, cudaMemcpyHostToDevice, ctx->stream));
    const unsigned g = (unsigned)std::min<int64_t>((n1 + 255) / 256, (int64_t)HS_NUM_SMS * 16);
    if (varlen) { len.alloc(ctx, (size_t)n1 * 4); bpos.alloc(ctx, (size_t)n1 * 4); }
    if (n > 0) {
      if (large) LAUNCH(ctx, k_pq_flags<int64_t>, g, 256, 0, kc, n, 0, flag.as<uint8_t>(), varlen ? len.as<uint32_t>() : nullptr);
      else LAUNCH(ctx, k_pq_flags<int32_t>, g, 256, 0, kc, n, 0, flag.as<uint8_t>(), varlen ? len.as<uint32_t>() : nullptr);
    }
    hs_exclusive_scan_flags(ctx, flag.as<uint8_t>(), vpos.as<uint32_t>(), n, vtotal.as<uint64_t>());
    if (varlen) hs_exclusive_scan_u32(ctx, len.as<uint32_t>(), bpos.as<uint32_t>(), n, btotal.as<uint64_t>());
    else btotal.zero();
    LAUNCH(ctx, k_pq_pick, (npages + 256) / 256, 256, 0, vpos.as<uint32_t>(), varlen ? bpos.as<uint32_t>() : nullptr,
           vtotal.as<uint64_t>(), btotal.as<uint64_t>(), d_rows.as<int64_t>(), npages, n, picks.as<uint64_t>());
    CK(cudaMemcpyAsync(h_first.data(), picks.p, h_first.size() * 8, cudaMemcpyDeviceToHost, ctx->stream));
    hs_sync(ctx);
  }
  // ---- dictionary pass: distinct values per chunk; chunks that stay within max_dict are dictionary-encoded
  const bool try_dict = max_dict > 0 && !varlen && !is_bool && (w == 4 || w == 8) && n > 0;
  std::vector<uint32_t> entries((size_t)nchunks, 0);          // dictionary entries per chunk; 0 = PLAIN
  DBuf d_keys, d_index, d_counts, d_pchunk, d_dictoff, idxv;
  PqDict pd{};
  if (try_dict) {
    Stage s(ctx, "parquet_dict");
    int slots = 64;
    while (slots < 2 * max_dict) slots <<= 1;
    d_keys.alloc(ctx, (size_t)nchunks * slots * 8);
    d_index.alloc(ctx, (size_t)nchunks * slots * 4);
    d_counts.alloc(ctx, (size_t)nchunks * 12);
    d_pchunk.alloc(ctx, (size_t)npages * 4);
    CK(cudaMemsetAsync(d_keys.p, 0xFF, (size_t)nchunks * slots * 8, ctx->stream));
    CK(cudaMemsetAsync(d_counts.p, 0, (size_t)nchunks * 12, ctx->stream));
    CK(cudaMemcpyAsync(d_pchunk.p, page_chunk, (size_t)npages * 4, cudaMemcpyHostToDevice, ctx->stream));
    pd = PqDict{nchunks, slots, max_dict, w, d_keys.as<unsigned long long>(), d_index.as<uint32_t>(), d_counts.as<uint32_t>(),
                d_counts.as<uint32_t>() + nchunks, d_counts.as<uint32_t>() + 2 * (size_t)nchunks, d_pchunk.as<int32_t>()};
    PqPages pg0{npages, d_rows.as<int64_t>(), nullptr, nullptr, picks.as<uint64_t>()};
    const unsigned g = (unsigned)std::min<int64_t>((n1 + 255) / 256, (int64_t)HS_NUM_SMS * 16);
    LAUNCH(ctx, k_pqd_insert, g, 256, 0, (const uint8_t*)dc.data, flag.as<uint8_t>(), n, pg0, pd);
    std::vector<uint32_t> h_counts(3 * (size_t)nchunks);
    CK(cudaMemcpyAsync(h_counts.data(), d_counts.p, h_counts.size() * 4, cudaMemcpyDeviceToHost, ctx->stream));
    hs_sync(ctx);
    for (int c = 0; c < nchunks; c++) {
      const uint32_t cnt = h_counts[c], over = h_counts[nchunks + c], has_e = h_counts[2 * (size_t)nchunks + c];
      if (!over && cnt <= (uint32_t)max_dict) entries[c] = cnt + (has_e ? 1u : 0u);
    }
  }
  auto bit_width = [](uint32_t nentries) { int b = 1; while ((1u << b) < nentries) b++; return b; };
  // ---- page layout (host): [def levels][values] per page, pages back to back; then the dictionary page bodies
  const int nout = npages + nchunks;
  std::vector<int64_t> page_off((size_t)nout + 1, 0), nulls((size_t)nout, 0);
  std::vector<uint64_t> val_off((size_t)npages), def_off((size_t)npages, ~0ull), pack_off((size_t)npages, ~0ull);
  std::vector<uint64_t> dict_off((size_t)nchunks, ~0ull);
  std::vector<uint32_t> pack_bytes((size_t)npages, 0), bw_of((size_t)npages, 0);
  bool any_dict = false;
  std::vector<std::pair<uint64_t, std::vector<uint8_t>>> headers;   // (offset, bytes) written from the host
  uint64_t at = 0;
  for (int p = 0; p < npages; p++) {
    page_off[p] = (int64_t)at;
    const int64_t rows = page_rows[p + 1] - page_rows[p];
    const int64_t valid = (int64_t)(h_first[2 * (p + 1)] - h_first[2 * p]);
    nulls[p] = rows - valid;
    if (optional) {
      std::vector<uint8_t> h(4 + 12);
      int k;
      uint32_t body;
      if (nulls[p] == 0) {                       // one RLE run: varint(rows << 1), value 1
        k = varint((uint64_t)rows << 1, h.data() + 4);
        h[4 + k] = 1;
        k += 1;
        body = (uint32_t)k;
      } else {                                   // bit-packed: varint((groups << 1) | 1), then groups bytes
        const int64_t groups = (rows + 7) >> 3;
        k = varint(((uint64_t)groups << 1) | 1u, h.data() + 4);
        body = (uint32_t)(k + groups);
        def_off[p] = at + 4 + (uint64_t)k;
      }
      memcpy(h.data(), &body, 4);
      h.resize(4 + (size_t)k);
      headers.emplace_back(at, std::move(h));
      at += 4 + body;
    }
    val_off[p] = at;
    if (nchunks && entries[page_chunk[p]]) {       // [bit width][varint((groups << 1) | 1)][groups * bw bytes]
      const int bw = bit_width(entries[page_chunk[p]]);
      std::vector<uint8_t> h(12);
      h[0] = (uint8_t)bw;
      int k = 1;
      if (valid > 0) {
        const int64_t groups = (valid + 7) >> 3;
        k += varint(((uint64_t)groups << 1) | 1u, h.data() + 1);
        pack_off[p] = at + (uint64_t)k;
        pack_bytes[p] = (uint32_t)(groups * bw);
        bw_of[p] = (uint32_t)bw;
      }
      h.resize((size_t)k);
      headers.emplace_back(at, std::move(h));
      at += (uint64_t)k + pack_bytes[p];
      val_off[p] = ~0ull;                          // no PLAIN values in this page
      any_dict = true;
      continue;
    }
    if (varlen) at += h_first[2 * (p + 1) + 1] - h_first[2 * p + 1];
    else if (is_bool) at += (uint64_t)((valid + 7) >> 3);
    else at += (uint64_t)valid * w;
  }
  for (int c = 0; c < nchunks; c++) {
    page_off[(size_t)npages + c] = (int64_t)at;
    nulls[(size_t)npages + c] = entries[c];        // second column, rows past the pages: dictionary entries
    if (entries[c]) { dict_off[c] = at; at += (uint64_t)entries[c] * w; }
  }
  page_off[nout] = (int64_t)at;
  if (at >= 0x7FFFFFF0ull) HS_THROW(HS_EUNSUPPORTED, "more than 2 GiB of page data in one call: encode fewer rows at a time");
  std::vector<OutCol> ocols(codec ? 3 : 2);
  OutCol& body = ocols[0];
  body.type = HS_BINARY;
  body.length = nout;
  body.data_bytes = (int64_t)at;
  body.data.alloc(ctx, (size_t)at + 32);   // slack: the compressor reads whole 16-byte words
  body.offsets.alloc(ctx, ((size_t)nout + 1) * 4);
  {
    Stage s(ctx, "parquet_encode");
    std::vector<int32_t> off32((size_t)nout + 1);
    for (int p = 0; p <= nout; p++) off32[p] = (int32_t)page_off[p];
    CK(cudaMemcpyAsync(body.offsets.p, off32.data(), off32.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
    if (is_bool) CK(cudaMemsetAsync(body.data.p, 0, (size_t)at + 32, ctx->stream));
    for (auto& h : headers)
      CK(cudaMemcpyAsync((uint8_t*)body.data.p + h.first, h.second.data(), h.second.size(), cudaMemcpyHostToDevice, ctx->stream));
    DBuf d_val(ctx, (size_t)npages * 8), d_def(ctx, (size_t)npages * 8);
    CK(cudaMemcpyAsync(d_val.p, val_off.data(), (size_t)npages * 8, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(d_def.p, def_off.data(), (size_t)npages * 8, cudaMemcpyHostToDevice, ctx->stream));
    PqPages pg{npages, d_rows.as<int64_t>(), d_val.as<uint64_t>(), d_def.as<uint64_t>(), picks.as<uint64_t>()};
    const unsigned g = (unsigned)std::min<int64_t>((n1 + 255) / 256, (int64_t)HS_NUM_SMS * 16);
    uint8_t* o = body.data.as<uint8_t>();
    if (n > 0) {
      if (varlen) {
        const unsigned g8 = (unsigned)std::min<int64_t>((n1 * 8 + 255) / 256, (int64_t)HS_NUM_SMS * 32);
        if (large) LAUNCH(ctx, k_pq_bytes<int64_t>, g8, 256, 0, (const int64_t*)dc.offsets, (const uint8_t*)dc.data, flag.as<uint8_t>(), bpos.as<uint32_t>(), n, pg, o);
        else LAUNCH(ctx, k_pq_bytes<int32_t>, g8, 256, 0, dc.offsets, (const uint8_t*)dc.data, flag.as<uint8_t>(), bpos.as<uint32_t>(), n, pg, o);
      } else if (is_bool) {
        LAUNCH(ctx, k_pq_bool, g, 256, 0, (const uint8_t*)dc.data, flag.as<uint8_t>(), vpos.as<uint32_t>(), n, pg, o);
      } else {
        LAUNCH(ctx, k_pq_fixed, g, 256, 0, (const uint8_t*)dc.data, w, flag.as<uint8_t>(), vpos.as<uint32_t>(), n, pg, o);
      }
      if (any_dict) {
        DBuf d_pack(ctx, (size_t)npages * 8), d_pbytes(ctx, (size_t)npages * 4), d_bw(ctx, (size_t)npages * 4);
        d_dictoff.alloc(ctx, (size_t)nchunks * 8);
        idxv.alloc(ctx, (size_t)std::max<uint64_t>(h_first[2 * (size_t)npages], 1) * 4);
        CK(cudaMemcpyAsync(d_pack.p, pack_off.data(), (size_t)npages * 8, cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaMemcpyAsync(d_pbytes.p, pack_bytes.data(), (size_t)npages * 4, cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaMemcpyAsync(d_bw.p, bw_of.data(), (size_t)npages * 4, cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaMemcpyAsync(d_dictoff.p, dict_off.data(), (size_t)nchunks * 8, cudaMemcpyHostToDevice, ctx->stream));
        LAUNCH(ctx, k_pqd_assign, (unsigned)nchunks, 256, 0, pd, d_dictoff.as<uint64_t>(), o);
        LAUNCH(ctx, k_pqd_lookup, g, 256, 0, (const uint8_t*)dc.data, flag.as<uint8_t>(), vpos.as<uint32_t>(), n, pg, pd,
               d_dictoff.as<uint64_t>(), idxv.as<uint32_t>());
        dim3 grid(64, (unsigned)npages);
        k_pqd_pack<<<grid, 256, 0, ctx->stream>>>(idxv.as<uint32_t>(), d_pack.as<uint64_t>(), d_pbytes.as<uint32_t>(),
                                                  d_bw.as<uint32_t>(), pg, o);
        hs_count_launch(ctx);
        CK(cudaGetLastError());
        hs_sync(ctx);   // the small per-page buffers above go out of scope here
      }
      bool any_def = false;
      for (int p = 0; p < npages; p++) any_def |= def_off[p] != ~0ull;
      if (any_def) {
        dim3 grid(64, (unsigned)npages);
        k_pq_deflevels<<<grid, 256, 0, ctx->stream>>>(flag.as<uint8_t>(), n, pg, o);
        hs_count_launch(ctx);
        CK(cudaGetLastError());
      }
    }
    hs_sync(ctx);   // headers, val_off, def_off are host vectors: the copies above must have left them
  }
  if (codec == 1) {
    // ---- snappy: every row of column 0 (pages and dictionary pages) compressed on its own
    Stage s(ctx, "parquet_snappy");
    std::vector<SnFrag> frags;
    std::vector<int32_t> frag_first((size_t)nout + 1, 0);
    for (int r = 0; r < nout; r++) {
      frag_first[r] = (int32_t)frags.size();
      for (int64_t a = page_off[r]; a < page_off[r + 1]; a += SN_FRAG)
        frags.push_back(SnFrag{(uint64_t)a, (uint32_t)std::min<int64_t>(SN_FRAG, page_off[r + 1] - a), 0});
    }
    const int nfrag = (int)frags.size();
    frag_first[nout] = nfrag;
    std::vector<int64_t> usize((size_t)nout), coff((size_t)nout + 1, 0);
    for (int r = 0; r < nout; r++) usize[r] = page_off[r + 1] - page_off[r];
    OutCol& us = ocols[2];
    us.type = HS_INT64;
    us.length = nout;
    us.data_bytes = (int64_t)nout * 8;
    us.data.alloc(ctx, (size_t)nout * 8);
    CK(cudaMemcpyAsync(us.data.p, usize.data(), (size_t)nout * 8, cudaMemcpyHostToDevice, ctx->stream));
    DBuf packed;
    std::vector<std::pair<uint64_t, std::vector<uint8_t>>> pre;
    if (nfrag > 0) {
      DBuf d_frags(ctx, (size_t)nfrag * sizeof(SnFrag)), d_sizes(ctx, (size_t)nfrag * 4), d_dst(ctx, (size_t)nfrag * 8);
      DBuf scratch(ctx, (size_t)nfrag * SN_SLOT);
      CK(cudaMemcpyAsync(d_frags.p, frags.data(), (size_t)nfrag * sizeof(SnFrag), cudaMemcpyHostToDevice, ctx->stream));
      const size_t smem = sizeof(SnWarp) * SN_WARPS;
      CK(cudaFuncSetAttribute(k_snappy_frag, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      LAUNCH(ctx, k_snappy_frag, (unsigned)((nfrag + SN_WARPS - 1) / SN_WARPS), SN_WARPS * 32, smem, body.data.as<uint8_t>(),
             d_frags.as<SnFrag>(), nfrag, scratch.as<uint8_t>(), d_sizes.as<uint32_t>());
      std::vector<uint32_t> sizes((size_t)nfrag);
      CK(cudaMemcpyAsync(sizes.data(), d_sizes.p, (size_t)nfrag * 4, cudaMemcpyDeviceToHost, ctx->stream));
      hs_sync(ctx);
      std::vector<uint64_t> dst((size_t)nfrag);
      uint64_t cat = 0;
      for (int r = 0; r < nout; r++) {
        coff[r] = (int64_t)cat;
        if (usize[r] == 0) continue;               // an empty row stays empty (no dictionary for this chunk)
        std::vector<uint8_t> v(10);
        v.resize((size_t)varint((uint64_t)usize[r], v.data()));
        cat += v.size();
        pre.emplace_back((uint64_t)coff[r], std::move(v));
        for (int f = frag_first[r]; f < frag_first[r + 1]; f++) { dst[f] = cat; cat += sizes[f]; }
      }
      coff[nout] = (int64_t)cat;
      if (cat >= 0x7FFFFFF0ull) HS_THROW(HS_EUNSUPPORTED, "more than 2 GiB of compressed page data in one call");
      packed.alloc(ctx, (size_t)cat + 16);
      CK(cudaMemcpyAsync(d_dst.p, dst.data(), (size_t)nfrag * 8, cudaMemcpyHostToDevice, ctx->stream));
      for (auto& h : pre)
        CK(cudaMemcpyAsync((uint8_t*)packed.p + h.first, h.second.data(), h.second.size(), cudaMemcpyHostToDevice, ctx->stream));
      LAUNCH(ctx, k_snappy_pack, (unsigned)((nfrag + 7) / 8), 256, 0, scratch.as<uint8_t>(), d_sizes.as<uint32_t>(), d_dst.as<uint64_t>(),
             nfrag, packed.as<uint8_t>());
      std::vector<int32_t> off32((size_t)nout + 1);
      for (int r = 0; r <= nout; r++) off32[r] = (int32_t)coff[r];
      CK(cudaMemcpyAsync(body.offsets.p, off32.data(), off32.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
      hs_sync(ctx);
      body.data = std::move(packed);
      body.data_bytes = (int64_t)cat;
    }
  }
  OutCol& nn = ocols[1];
  nn.type = HS_INT64;
  nn.length = nout;
  nn.data_bytes = (int64_t)nout * 8;
  nn.data.alloc(ctx, (size_t)nout * 8);
  CK(cudaMemcpyAsync(nn.data.p, nulls.data(), (size_t)nout * 8, cudaMemcpyHostToDevice, ctx->stream));
  hs_sync(ctx);
  Stage s(ctx, "d2h");
  *out = hs_finish_table(ctx, ocols, nout, 0, nullptr);
}

extern "C" int32_t hs_parquet_plain_pages(hs_ctx* ctx, const hs_column* col, int32_t optional, const int64_t* page_rows,
                                          int32_t npages, hs_table** out) {
  API_BEGIN(ctx)
  pq_encode(ctx, col, optional, page_rows, npages, nullptr, 0, 0, 0, out);
  API_END(ctx)
}

extern "C" int32_t hs_parquet_encode_pages(hs_ctx* ctx, const hs_column* col, int32_t optional, const int64_t* page_rows,
                                           int32_t npages, const int32_t* page_chunk, int32_t nchunks, int32_t max_dict,
                                           int32_t codec, hs_table** out) {
  API_BEGIN(ctx)
  pq_encode(ctx, col, optional, page_rows, npages, page_chunk, nchunks, max_dict, codec, out);
  API_END(ctx)
}
