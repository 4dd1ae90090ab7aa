// hs_build.cu — the index build behind hs_build_index / hs_merge_runs.
// Orchestrates K1 (hash + histogram), K2 (bucket scatter) and K3 (per-bucket sort) for the body of
// CreateActionBase.write (HS/actions/CreateActionBase.scala:122-140):
//   repartition(numBuckets, indexedCols)  ->  saveWithBuckets(BucketSpec(n, cols, cols))
// Data layout in HBM: every fixed-width column is its own array (SoA). Validity bitmaps and boolean
// columns travel as one byte per row and are re-packed at the end; string columns do not travel —
// a 32-bit source-row column does, and the bytes are gathered once at the end.
#include <algorithm>

#include "hs_keys.cuh"
#include "hs_kernels.h"

// ------------------------------------------------------------------ small kernels
__global__ void __launch_bounds__(256) k_bits_to_bytes(const uint8_t* __restrict__ bits, int64_t n,
                                                       uint8_t* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) out[i] = (bits[i >> 3] >> (i & 7)) & 1;
}
__global__ void __launch_bounds__(256) k_bytes_to_bits(const uint8_t* __restrict__ in, int64_t n,
                                                       uint8_t* __restrict__ bits) {
  int64_t nb = (n + 7) >> 3;
  int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; b < nb; b += stride) {
    unsigned v = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
      int64_t i = b * 8 + k;
      if (i < n && in[i]) v |= 1u << k;
    }
    bits[b] = (uint8_t)v;
  }
}
// same, also counting the zero bytes (= NULL rows) into *zeros
__global__ void __launch_bounds__(256) k_bytes_to_bits_count(const uint8_t* __restrict__ in, int64_t n,
                                                             uint8_t* __restrict__ bits, unsigned long long* zeros) {
  int64_t nb = (n + 7) >> 3;
  int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  unsigned z = 0;
  for (; b < nb; b += stride) {
    unsigned v = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
      int64_t i = b * 8 + k;
      if (i < n) { if (in[i]) v |= 1u << k; else z++; }
    }
    bits[b] = (uint8_t)v;
  }
  if (z) atomicAdd(zeros, (unsigned long long)z);
}
__global__ void __launch_bounds__(256) k_iota_u32(uint32_t* out, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) out[i] = (uint32_t)i;
}
// chunk d of the normalised key of the rows rows[i] (or i when rows == NULL)
__global__ void __launch_bounds__(256) k_build_nkey(KeyLayout L, KeySpec ks, int64_t n, int d,
                                                    const uint32_t* __restrict__ rows,
                                                    const uint8_t* __restrict__ flag,
                                                    uint64_t* __restrict__ nkey, uint8_t* __restrict__ more) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    if (flag) {  // refinement: only rows inside a tie group of two or more rows
      bool single = flag[i] && (i + 1 == n || flag[i + 1]);
      if (single) continue;
    }
    bool m = false;
    uint64_t v = nkey_chunk(L, ks, rows ? rows[i] : i, d, &m);
    nkey[i] = v;
    if (more) more[i] = m ? 1 : 0;
  }
}
__global__ void __launch_bounds__(256) k_mark_starts(const uint32_t* __restrict__ seg_off, int64_t nseg,
                                                     uint8_t* __restrict__ flag, uint32_t n) {
  int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s < nseg) {
    uint32_t a = seg_off[s];
    if (a < n) flag[a] = 1;
  }
}
// a row starts a new group when its chunk differs from the previous row's, or the key is complete
__global__ void __launch_bounds__(256) k_tie_flags(const uint64_t* __restrict__ nkey, const uint8_t* __restrict__ more,
                                                   int64_t n, uint8_t* __restrict__ flag) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    uint8_t f = flag[i];
    if (!f) {
      if (i == 0 || nkey[i] != nkey[i - 1] || !more[i]) f = 1;
      flag[i] = f;
    }
  }
}
__global__ void __launch_bounds__(256) k_head_positions(const uint8_t* __restrict__ flag, const uint32_t* __restrict__ rank,
                                                        int64_t n, uint32_t* __restrict__ head_pos) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride)
    if (flag[i]) head_pos[rank[i]] = (uint32_t)i;
}
// groups of two or more rows: up to TIE_SMALL rows -> the small list (finished in one go by k_tie_finish),
// longer ones -> the big list (next chunk, another round). counter[0] = small, counter[1] = big.
#define TIE_SMALL 64
#define TF_PRE 3
__global__ void __launch_bounds__(256) k_group_segments(const uint32_t* __restrict__ head_pos, const uint64_t* ngroups,
                                                        uint32_t n, uint32_t* __restrict__ sstart,
                                                        uint32_t* __restrict__ slen, uint32_t* __restrict__ gstart,
                                                        uint32_t* __restrict__ glen, uint32_t* counter) {
  int64_t G = (int64_t)*ngroups;
  int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; g < G; g += stride) {
    uint32_t a = head_pos[g];
    uint32_t b = g + 1 < G ? head_pos[g + 1] : n;
    if (b - a > TIE_SMALL) {
      uint32_t k = atomicAdd(counter + 1, 1u);
      gstart[k] = a;
      glen[k] = b - a;
    } else if (b - a >= 2) {
      // small groups are many (one per shared prefix): one atomic per warp, not per group
      const uint32_t peers = __activemask();
      const uint32_t lane = threadIdx.x & 31, leader = __ffs(peers) - 1;
      uint32_t base = 0;
      if (lane == leader) base = atomicAdd(counter, (uint32_t)__popc(peers));
      base = __shfl_sync(peers, base, leader);
      const uint32_t k = base + __popc(peers & ((1u << lane) - 1u));
      sstart[k] = a;
      slen[k] = b - a;
    }
  }
}

// ---- rounds after the first work on the rows of the big groups only, through a compact index space:
// compact row ci of group g = real row gstart[g] + (ci - cstart[g])
__global__ void __launch_bounds__(256) k_tie_rowmap(const uint32_t* __restrict__ gstart, const uint32_t* __restrict__ cstart,
                                                    uint32_t ngroups, uint32_t m, uint32_t* __restrict__ rowmap,
                                                    uint8_t* __restrict__ cflag) {
  uint32_t ci = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t stride = gridDim.x * blockDim.x;
  for (; ci < m; ci += stride) {
    uint32_t lo = 0, hi = ngroups;      // largest g with cstart[g] <= ci
    while (hi - lo > 1) {
      const uint32_t mid = (lo + hi) >> 1;
      if (cstart[mid] <= ci) lo = mid; else hi = mid;
    }
    rowmap[ci] = gstart[lo] + (ci - cstart[lo]);
    cflag[ci] = ci == cstart[lo];       // a big group starts here: forced head
  }
}
__global__ void __launch_bounds__(256) k_build_nkey_rows(KeyLayout L, KeySpec ks, const uint32_t* __restrict__ rowmap,
                                                         uint32_t m, int d, const uint32_t* __restrict__ srcrow,
                                                         uint64_t* __restrict__ nkey, uint8_t* __restrict__ more) {
  uint32_t ci = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t stride = gridDim.x * blockDim.x;
  for (; ci < m; ci += stride) {
    const uint32_t i = rowmap[ci];
    bool mr = false;
    nkey[i] = nkey_chunk(L, ks, srcrow[i], d, &mr);
    more[i] = mr ? 1 : 0;
  }
}
__global__ void __launch_bounds__(256) k_tie_flags_rows(const uint64_t* __restrict__ nkey, const uint8_t* __restrict__ more,
                                                        const uint32_t* __restrict__ rowmap, uint32_t m,
                                                        uint8_t* __restrict__ cflag) {
  uint32_t ci = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t stride = gridDim.x * blockDim.x;
  for (; ci < m; ci += stride) {
    if (cflag[ci]) continue;
    const uint32_t i = rowmap[ci];      // not a group start: row i - 1 belongs to the same big group
    if (nkey[i] != nkey[i - 1] || !more[i]) cflag[ci] = 1;
  }
}
// as k_group_segments, heads given in compact positions; the groups go out in real row positions
__global__ void __launch_bounds__(256) k_group_segments_rows(const uint32_t* __restrict__ head_pos, const uint64_t* ngroups,
                                                             uint32_t m, const uint32_t* __restrict__ rowmap,
                                                             uint32_t* __restrict__ sstart, uint32_t* __restrict__ slen,
                                                             uint32_t* __restrict__ gstart, uint32_t* __restrict__ glen,
                                                             uint32_t* counter) {
  int64_t G = (int64_t)*ngroups;
  int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; g < G; g += stride) {
    const uint32_t a = head_pos[g];
    const uint32_t b = g + 1 < G ? head_pos[g + 1] : m;
    if (b - a > TIE_SMALL) {
      uint32_t k = atomicAdd(counter + 1, 1u);
      gstart[k] = rowmap[a];
      glen[k] = b - a;
    } else if (b - a >= 2) {
      uint32_t k = atomicAdd(counter, 1u);
      sstart[k] = rowmap[a];
      slen[k] = b - a;
    }
  }
}

struct TieCols {
  int32_t n, srcrow_col;
  int32_t esize[HS_MAX_MOVE];
  void* a[HS_MAX_MOVE];
};
// -1 / 0 / +1: rows ra, rb of the SOURCE table compared on key chunks d0, d0+1, ... (equal chunks carry equal
// "more" flags: the tag byte of a string chunk says whether the string goes on)
__device__ __forceinline__ int cmp_key_tail(const KeyLayout& L, const KeySpec& ks, uint32_t ra, uint32_t rb, int d0) {
  for (int d = d0;; d++) {
    bool ma = false, mb = false;
    const uint64_t a = nkey_chunk(L, ks, ra, d, &ma);
    const uint64_t b = nkey_chunk(L, ks, rb, d, &mb);
    if (a != b) return a < b ? -1 : 1;
    if (!ma || !mb) return 0;
  }
}
// One warp finishes one small tie group: every row ranks itself inside the group by comparing the rest of its
// key (chunks d0...) with every other row's, straight from the source columns; the moved columns are permuted in
// place and the rows are flagged as settled. Most ties of a string key are a handful of rows sharing a prefix:
// this replaces one full round (flags, scan, group list, next chunk, segmented sort) per 7 more key bytes.
__global__ void __launch_bounds__(256) k_tie_finish(const __grid_constant__ KeyLayout L, const __grid_constant__ KeySpec ks,
                                                    const __grid_constant__ TieCols tc,
                                                    const uint32_t* __restrict__ sstart, const uint32_t* __restrict__ slen,
                                                    uint32_t ngroups, int d0, uint8_t* __restrict__ flag) {
  const uint32_t g = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (g >= ngroups) return;
  const uint32_t start = sstart[g], len = slen[g];
  const uint32_t lane = lane_id();
  const uint32_t i0 = lane, i1 = lane + 32;
  const uint32_t* srcrow = (const uint32_t*)tc.a[tc.srcrow_col];
  const uint32_t s0 = i0 < len ? srcrow[start + i0] : 0, s1 = i1 < len ? srcrow[start + i1] : 0;
  // the next TF_PRE chunks of both rows go to registers (independent loads, one round trip); the ranking loop
  // then runs on shuffles and touches memory only for pairs that still tie after them
  uint64_t a0[TF_PRE], a1[TF_PRE];
  bool m0 = i0 < len, m1 = i1 < len;     // "the key goes on" after the chunks read so far
#pragma unroll
  for (int t = 0; t < TF_PRE; t++) {
    a0[t] = a1[t] = 0;
    if (m0) a0[t] = nkey_chunk(L, ks, s0, d0 + t, &m0);
    if (m1) a1[t] = nkey_chunk(L, ks, s1, d0 + t, &m1);
  }
  uint32_t r0 = 0, r1 = 0;
  for (uint32_t j = 0; j < len; j++) {
    const bool hi = j >= 32;
    const uint32_t sj = __shfl_sync(0xffffffffu, hi ? s1 : s0, j & 31);
    const bool mj = __shfl_sync(0xffffffffu, (int)(hi ? m1 : m0), j & 31) != 0;
    int c0 = 0, c1 = 0;
#pragma unroll
    for (int t = 0; t < TF_PRE; t++) {
      const uint64_t b = __shfl_sync(0xffffffffu, hi ? a1[t] : a0[t], j & 31);
      if (c0 == 0 && b != a0[t]) c0 = b < a0[t] ? -1 : 1;
      if (c1 == 0 && b != a1[t]) c1 = b < a1[t] ? -1 : 1;
    }
    if (i0 < len && j != i0) {
      if (c0 == 0 && mj && m0) c0 = cmp_key_tail(L, ks, sj, s0, d0 + TF_PRE);
      r0 += c0 < 0 || (c0 == 0 && j < i0);
    }
    if (i1 < len && j != i1) {
      if (c1 == 0 && mj && m1) c1 = cmp_key_tail(L, ks, sj, s1, d0 + TF_PRE);
      r1 += c1 < 0 || (c1 == 0 && j < i1);
    }
  }
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
  for (int c0 = 0; c0 < tc.n; c0 += 4) {
    uint64_t v0[4], v1[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const int c = c0 + u;
      v0[u] = v1[u] = 0;
      if (c < tc.n) {
        if (i0 < len) v0[u] = ld(tc.a[c], (size_t)start + i0, tc.esize[c]);
        if (i1 < len) v1[u] = ld(tc.a[c], (size_t)start + i1, tc.esize[c]);
      }
    }
    __syncwarp();
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const int c = c0 + u;
      if (c < tc.n) {
        if (i0 < len) st(tc.a[c], (size_t)start + r0, tc.esize[c], v0[u]);
        if (i1 < len) st(tc.a[c], (size_t)start + r1, tc.esize[c], v1[u]);
      }
    }
    __syncwarp();
  }
  if (i0 < len) flag[start + i0] = 1;
  if (i1 < len) flag[start + i1] = 1;
}

// Integer key columns out of the sorted single-chunk normalised key: the key IS the column values (valid flag,
// value - minimum), so those columns need not travel through the scatter and sort passes at all.
template <typename T>
__global__ void __launch_bounds__(256) k_decode_key(const uint64_t* __restrict__ nkey, int64_t n, NGroup g,
                                                    T* __restrict__ out, uint8_t* __restrict__ valid_bytes) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const uint64_t mask = g.width >= 64 ? ~0ull : ((1ull << g.width) - 1ull);
  for (; i < n; i += stride) {
    const uint64_t k = nkey[i];
    const bool valid = g.flag_shift < 0 || ((k >> g.flag_shift) & 1ull);
    const uint64_t mapped = (g.width ? ((k >> g.shift) & mask) : 0ull) + g.min;
    out[i] = valid ? (T)(int64_t)(mapped ^ 0x8000000000000000ull) : (T)0;
    if (valid_bytes) valid_bytes[i] = valid;
  }
}

static unsigned grid1d(int64_t n) {
  int64_t g = (n + 255) / 256;
  int64_t cap = (int64_t)HS_NUM_SMS * 16;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (unsigned)g;
}

static int bitlen64(uint64_t v) {
  int b = 0;
  while (v) { b++; v >>= 1; }
  return b;
}

// ------------------------------------------------------------------ key layout (host)
// Packs the fixed-width key columns into static chunks (MSB first, right aligned); see hs_keys.cuh.
static KeyLayout make_layout(const std::vector<DevCol>& cols, int nk, const uint64_t* minmax,
                             const std::vector<uint8_t>& has_nulls) {
  KeyLayout L;
  memset(&L, 0, sizeof L);
  L.nkeys = nk;
  int run = -1;
  bool prev_fixed = false;
  bool any_string = false;
  struct Item { int col; int bits; bool is_flag; int chunk; };
  std::vector<std::vector<Item>> runs;
  for (int k = 0; k < nk; k++) {
    NGroup& g = L.g[k];
    g.flag_shift = -1;
    g.flag_chunk = -1;
    g.chunk = -1;
    if (hs_is_varlen(cols[k].type)) {
      g.is_string = 1;
      prev_fixed = false;
      any_string = true;
      continue;
    }
    if (!prev_fixed) { run++; runs.emplace_back(); }
    prev_fixed = true;
    g.run_id = run;
    uint64_t mn = minmax[2 * k], mx = minmax[2 * k + 1];
    int width = mn <= mx ? bitlen64(mx - mn) : 0;
    g.min = mn <= mx ? mn : 0;
    g.width = width;
    if (has_nulls[k]) runs[run].push_back({k, 1, true, 0});
    if (width > 0) runs[run].push_back({k, width, false, 0});
  }
  L.nruns = run + 1;
  int total_static_chunks = 0;
  for (int r = 0; r <= run; r++) {
    auto& items = runs[r];
    int chunk = 0, used = 0;
    for (auto& it : items) {
      if (used + it.bits > 64) { chunk++; used = 0; }
      it.chunk = chunk;
      used += it.bits;
    }
    int nch = items.empty() ? 1 : chunk + 1;
    L.run_chunks[r] = nch;
    total_static_chunks += nch;
    // shifts: right aligned inside each chunk
    for (int ch = 0; ch < nch; ch++) {
      int tot = 0;
      for (auto& it : items) if (it.chunk == ch) tot += it.bits;
      int consumed = 0;
      for (auto& it : items) {
        if (it.chunk != ch) continue;
        consumed += it.bits;
        int shift = tot - consumed;
        if (it.is_flag) { L.g[it.col].flag_shift = shift; L.g[it.col].flag_chunk = ch; }
        else { L.g[it.col].shift = shift; L.g[it.col].chunk = ch; }
      }
      if (r == 0 && ch == 0) L.bits0 = tot;
    }
  }
  L.single_chunk = (!any_string && total_static_chunks <= 1) ? 1 : 0;
  if (!L.single_chunk) L.bits0 = 64;
  return L;
}

struct MovedCol {
  int user_col;     // index of the user column, or -1 for internal columns
  int role;         // 0 data, 1 validity bytes, 2 bool bytes, 3 nkey, 4 more, 5 srcrow
  int esize;
  const void* src;  // device source in input row order
  DBuf a, b, tmp;   // tmp: expanded bits / generated columns (input order)
};

// The build over staged device columns; shared by hs_build_index, hs_merge_runs and the multi-GPU build.
void hs_build_core(hs_ctx* c, std::vector<DevCol>& cols, int n_indexed, int32_t nb, int32_t blo, int32_t nbl,
                   std::vector<OutCol>& out, std::vector<int64_t>& boff, BuildExchange* xchg) {
  const int ncols = (int)cols.size();
  int64_t n = cols[0].length;
  boff.assign((size_t)nb + 1, 0);
  out.resize(ncols);
  if (n == 0 && !xchg) {
    for (int i = 0; i < ncols; i++) hs_copy_column(c, cols[i], out[i]);
    return;
  }
  KeySpec ks = hs_make_keyspec(cols, n_indexed);
  if (xchg) { blo = 0; nbl = nb; }   // the local phase sees every bucket; ownership is applied by the exchange

  // ---- K1: bucket ids, histogram, key ranges
  DBuf bucket(c, (size_t)n * 2 + 16), hist(c, (size_t)nbl * 4), minmax(c, (2 * HS_MAX_KEYS + 1) * 8);
  {
    Stage s(c, "hash_hist");
    hs_hash_bucket_hist(c, ks, cols, n, nb, blo, nbl, bucket.as<uint16_t>(), hist.as<uint32_t>(),
                        minmax.as<uint64_t>());
  }
  std::vector<uint32_t> h_hist((size_t)nbl);
  uint64_t h_minmax[2 * HS_MAX_KEYS + 1];
  CK(cudaMemcpyAsync(h_hist.data(), hist.p, (size_t)nbl * 4, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaMemcpyAsync(h_minmax, minmax.p, sizeof h_minmax, cudaMemcpyDeviceToHost, c->stream));
  hs_sync(c);
  if (h_minmax[2 * HS_MAX_KEYS] != 0)
    HS_THROW(HS_EINVAL, "%llu rows hash outside this rank's bucket range", (unsigned long long)h_minmax[2 * HS_MAX_KEYS]);
  std::vector<uint8_t> has_nulls((size_t)ncols);
  for (int i = 0; i < ncols; i++) has_nulls[i] = cols[i].validity != nullptr;
  // key ranges and NULL presence become global: every rank builds the same nkey and moves the same columns
  if (xchg) xchg->share_stats(c, h_hist, h_minmax, has_nulls);
  // local offsets (device, u32) drive the kernels; the global table has empty ranges outside [blo, blo+nbl)
  std::vector<uint32_t> off32((size_t)nbl + 1, 0);
  for (int b = 0; b < nbl; b++) off32[b + 1] = off32[b] + h_hist[b];
  for (int b = 0; b < nb; b++) {
    int lb = b - blo;
    boff[b + 1] = boff[b] + ((lb >= 0 && lb < nbl) ? (int64_t)h_hist[lb] : 0);
  }
  DBuf d_off(c, ((size_t)nbl + 1) * 4);
  hs_upload(c, d_off.p, off32.data(), ((size_t)nbl + 1) * 4);

  // ---- key layout and the set of travelling columns
  KeyLayout L = make_layout(cols, n_indexed, h_minmax, has_nulls);
  bool fast_key = n_indexed == 1 && !has_nulls[0] && (cols[0].type == HS_INT64 || cols[0].type == HS_INT32);
  bool any_varlen = false;
  for (auto& dc : cols) any_varlen |= hs_is_varlen(dc.type);
  bool need_srcrow = any_varlen || !L.single_chunk;
  // composite / small integer keys that fit one chunk: the key columns stay behind, the 8-byte key stands for them
  bool packed_keys = !fast_key && L.single_chunk;
  for (int k = 0; k < n_indexed && packed_keys; k++)
    if (cols[k].type < HS_INT8 || cols[k].type > HS_INT64) packed_keys = false;
  {
    const char* off = getenv("HSGPU_MOVE_KEY_COLUMNS");
    if (off && *off && *off != '0') packed_keys = false;
  }
  if (xchg && need_srcrow) {
    xchg->rows_first = true;
    return;
  }

  std::vector<MovedCol> mv;
  mv.reserve(HS_MAX_MOVE);
  auto add = [&](int user, int role, int esize, const void* src) -> MovedCol& {
    if ((int)mv.size() >= HS_MAX_MOVE) HS_THROW(HS_EUNSUPPORTED, "too many columns in one index (limit %d moved arrays)", HS_MAX_MOVE);
    mv.emplace_back();
    MovedCol& m = mv.back();
    m.user_col = user; m.role = role; m.esize = esize; m.src = src;
    return m;
  };
  {
    Stage s(c, "prepare");
    for (int i = 0; i < ncols; i++) {
      DevCol& dc = cols[i];
      if (packed_keys && i < n_indexed) continue;
      if (has_nulls[i]) {
        MovedCol& m = add(i, 1, 1, nullptr);
        m.tmp.alloc(c, (size_t)n);
        if (dc.validity) LAUNCH(c, k_bits_to_bytes, grid1d(n), 256, 0, dc.validity, n, m.tmp.as<uint8_t>());
        else CK(cudaMemsetAsync(m.tmp.p, 1, (size_t)(n ? n : 1), c->stream));   // NULLs only on other ranks
        m.src = m.tmp.p;
      }
      if (hs_is_varlen(dc.type)) continue;
      if (dc.type == HS_BOOL) {
        MovedCol& m = add(i, 2, 1, nullptr);
        m.tmp.alloc(c, (size_t)n);
        LAUNCH(c, k_bits_to_bytes, grid1d(n), 256, 0, (const uint8_t*)dc.data, n, m.tmp.as<uint8_t>());
        m.src = m.tmp.p;
      } else {
        add(i, 0, hs_fixed_width(dc.type), dc.data);
      }
    }
    if (!fast_key) {
      MovedCol& m = add(-1, 3, 8, nullptr);
      m.tmp.alloc(c, (size_t)n * 8);
      MovedCol* mm = nullptr;
      if (!L.single_chunk) {
        mm = &add(-1, 4, 1, nullptr);
        mm->tmp.alloc(c, (size_t)n);
        mm->src = mm->tmp.p;
      }
      MovedCol& nk = mv[mv.size() - (mm ? 2 : 1)];
      nk.src = nk.tmp.p;
      LAUNCH(c, k_build_nkey, grid1d(n), 256, 0, L, ks, n, 0, (const uint32_t*)nullptr, (const uint8_t*)nullptr,
             nk.tmp.as<uint64_t>(), mm ? mm->tmp.as<uint8_t>() : (uint8_t*)nullptr);
    }
    if (need_srcrow) {
      MovedCol& m = add(-1, 5, 4, nullptr);
      m.tmp.alloc(c, (size_t)n * 4);
      LAUNCH(c, k_iota_u32, grid1d(n), 256, 0, m.tmp.as<uint32_t>(), n);
      m.src = m.tmp.p;
    }
  }
  MoveSet ms;
  memset(&ms, 0, sizeof ms);
  ms.n = (int)mv.size();
  std::vector<const void*> srcs(mv.size());
  int key_mv = -1, more_mv = -1, srcrow_mv = -1;
  std::vector<int> mv_es;
  for (auto& m : mv) mv_es.push_back(m.esize);
  if (xchg && fast_key && cols[0].type == HS_INT64) {
    for (size_t i = 0; i < mv.size(); i++)
      if (mv[i].role == 0 && mv[i].user_col == 0) xchg->key_col = (int)i;
    xchg->key_min = h_minmax[0];
    xchg->key_span = h_minmax[1] - h_minmax[0];
  }
  const bool fused = xchg && xchg->try_fused(c, mv_es);   // rows go straight to their owners' buffers
  for (size_t i = 0; i < mv.size(); i++) {
    if (!fused) mv[i].a.alloc(c, (size_t)n * mv[i].esize + 16);
    ms.esize[i] = mv[i].esize;
    ms.a[i] = mv[i].a.p;
    ms.b[i] = nullptr;
    srcs[i] = mv[i].src;
    if (mv[i].role == 3) key_mv = (int)i;
    if (mv[i].role == 4) more_mv = (int)i;
    if (mv[i].role == 5) srcrow_mv = (int)i;
    if (fast_key && mv[i].role == 0 && mv[i].user_col == 0) key_mv = (int)i;
  }
  KeySrc key;
  memset(&key, 0, sizeof key);
  key.col = key_mv;
  int key_bits;
  if (fast_key) {
    key.mode = cols[0].type == HS_INT64 ? KS_I64 : KS_I32;
    key.min = h_minmax[0];
    key.span = h_minmax[1] - h_minmax[0];
    key_bits = bitlen64(h_minmax[1] - h_minmax[0]);
  } else {
    key.mode = KS_U64;
    key.min = 0;
    key_bits = L.bits0;
  }

  // ---- K2: rows -> bucket-contiguous layout
  std::vector<void*> recv_ptrs;
  if (fused) {
    // multi-GPU, fused: one kernel scatters by bucket and delivers over NVLink at the same time
    std::vector<uint32_t> seg_off;
    n = xchg->scatter_exchange(c, srcs.data(), mv_es, bucket.as<uint16_t>(), n, recv_ptrs, seg_off);
    blo = xchg->blo;
    nbl = xchg->nbl;
    off32 = seg_off;
    for (size_t i = 0; i < mv.size(); i++) {
      mv[i].a.alloc(c, (size_t)n * mv[i].esize + 16);
      ms.a[i] = mv[i].a.p;
      ms.s0[i] = recv_ptrs[i];
    }
    std::fill(boff.begin(), boff.end(), 0);
    for (int b = 0; b < nb; b++) {
      int lb = b - blo;
      boff[b + 1] = boff[b] + ((lb >= 0 && lb < nbl) ? (int64_t)(off32[lb + 1] - off32[lb]) : 0);
    }
    d_off.alloc(c, ((size_t)nbl + 1) * 4);
    hs_upload(c, d_off.p, off32.data(), ((size_t)nbl + 1) * 4);
  } else {
    Stage s(c, "bucket_scatter");
    hs_bucket_scatter(c, srcs.data(), ms, bucket.as<uint16_t>(), n, nbl, d_off.as<uint32_t>());
  }
  for (auto& m : mv) m.tmp.release();
  bucket.release();

  // ---- multi-GPU: bucket runs travel to the GPU that owns the bucket and land at their final place
  if (xchg && !fused) {
    std::vector<DBuf*> bufs;
    std::vector<int> es;
    for (auto& m : mv) { bufs.push_back(&m.a); es.push_back(m.esize); }
    std::vector<uint32_t> seg_off;
    n = xchg->exchange(c, bufs, es, off32, seg_off);
    blo = xchg->blo;
    nbl = xchg->nbl;
    off32 = seg_off;
    for (size_t i = 0; i < mv.size(); i++) ms.a[i] = mv[i].a.p;
    std::fill(boff.begin(), boff.end(), 0);
    for (int b = 0; b < nb; b++) {
      int lb = b - blo;
      boff[b + 1] = boff[b] + ((lb >= 0 && lb < nbl) ? (int64_t)(off32[lb + 1] - off32[lb]) : 0);
    }
    d_off.alloc(c, ((size_t)nbl + 1) * 4);
    hs_upload(c, d_off.p, off32.data(), ((size_t)nbl + 1) * 4);
  }
  for (size_t i = 0; i < mv.size(); i++) {
    mv[i].b.alloc(c, (size_t)n * mv[i].esize + 16);
    ms.b[i] = mv[i].b.p;
  }

  // ---- K3: every bucket sorted on the key
  if (fused && xchg->ngroups > 0) {
    // grouped spans: the buckets of group g are sorted while the rows of group g + 1 still cross NVLink
    struct CarryGuard {
      SortCarry* k;
      ~CarryGuard() { hs_sort_carry_free(k); }
    } carry{hs_sort_carry_new(c, n)};
    for (int g = 0; g < xchg->ngroups; g++) {
      int lb0 = 0, lb1 = 0;
      xchg->deliver_group(c, g, &lb0, &lb1);
      if (lb1 > lb0) hs_segmented_sort(c, ms, key, key_bits, nullptr, lb1 - lb0, n, off32.data() + lb0, carry.k);
    }
    hs_sort_carry_finish(c, ms, key, carry.k);
  } else {
    hs_segmented_sort(c, ms, key, key_bits, d_off.as<uint32_t>(), nbl, n, off32.data());
  }

  // ---- keys wider than one chunk (strings, wide composites): refine tie groups chunk by chunk
  if (!L.single_chunk) {
    DBuf flag(c, (size_t)n), rank(c, (size_t)n * 4), head_pos(c, (size_t)n * 4), total(c, 8);
    DBuf gstart(c, (size_t)(n / TIE_SMALL + 1) * 4), glen(c, (size_t)(n / TIE_SMALL + 1) * 4), counter(c, 8);
    DBuf sstart(c, (size_t)(n / 2 + 1) * 4), slen(c, (size_t)(n / 2 + 1) * 4);
    flag.zero();
    LAUNCH(c, k_mark_starts, grid1d(nbl + 1), 256, 0, d_off.as<uint32_t>(), (int64_t)nbl, flag.as<uint8_t>(), (uint32_t)n);
    uint64_t* nkey = (uint64_t*)ms.a[key_mv];
    uint8_t* more = (uint8_t*)ms.a[more_mv];
    const uint32_t* srcrow = (const uint32_t*)ms.a[srcrow_mv];
    TieCols tc;
    memset(&tc, 0, sizeof tc);
    tc.n = ms.n;
    tc.srcrow_col = srcrow_mv;
    for (int i = 0; i < ms.n; i++) { tc.esize[i] = ms.esize[i]; tc.a[i] = ms.a[i]; }
    // round 0: ties of chunk 0 among all rows
    uint32_t cnt[2];
    {
      hs_stage_begin(c, "tie_groups");
      LAUNCH(c, k_tie_flags, grid1d(n), 256, 0, nkey, more, n, flag.as<uint8_t>());
      hs_exclusive_scan_flags(c, flag.as<uint8_t>(), rank.as<uint32_t>(), n, total.as<uint64_t>());
      LAUNCH(c, k_head_positions, grid1d(n), 256, 0, flag.as<uint8_t>(), rank.as<uint32_t>(), n, head_pos.as<uint32_t>());
      counter.zero();
      LAUNCH(c, k_group_segments, grid1d(n), 256, 0, head_pos.as<uint32_t>(), total.as<uint64_t>(), (uint32_t)n,
             sstart.as<uint32_t>(), slen.as<uint32_t>(), gstart.as<uint32_t>(), glen.as<uint32_t>(), counter.as<uint32_t>());
      hs_stage_end(c);
      memcpy(cnt, hs_read_wait(c, hs_read_begin(c, counter.p, 8)), 8);
      if (getenv("HSGPU_TRACE")) fprintf(stderr, "[hsgpu] tie round 0: %u small groups, %u big groups\n", cnt[0], cnt[1]);
      if (cnt[0]) {
        // small groups are settled for good, whatever the length of the keys
        Stage st(c, "tie_finish");
        LAUNCH(c, k_tie_finish, (cnt[0] + 7) / 8, 256, 0, L, ks, tc, sstart.as<uint32_t>(), slen.as<uint32_t>(), cnt[0], 1,
               flag.as<uint8_t>());
      }
    }
    rank.release();
    head_pos.release();
    // Long runs of equal prefixes (duplicates, shared path-like prefixes): next chunk, sorted by the segmented
    // sort, and another look — only over the rows of those runs, addressed through a compact index space.
    DBuf gstart2(c, gstart.bytes), glen2(c, glen.bytes);
    DBuf* gs[2] = {&gstart, &gstart2};
    DBuf* gl[2] = {&glen, &glen2};
    int cur = 0;
    for (int d = 1; cnt[1] > 0 && d < 1 << 20; d++) {
      const uint32_t G = cnt[1];
      DBuf cstart(c, (size_t)G * 4), mtot(c, 8);
      hs_stage_begin(c, "tie_groups");
      hs_exclusive_scan_u32(c, gl[cur]->as<uint32_t>(), cstart.as<uint32_t>(), G, mtot.as<uint64_t>());
      const uint32_t m = (uint32_t) * (const uint64_t*)hs_read_wait(c, hs_read_begin(c, mtot.p, 8));
      DBuf rowmap(c, (size_t)m * 4), cflag(c, (size_t)m), crank(c, (size_t)m * 4), chead(c, (size_t)m * 4), ctotal(c, 8);
      LAUNCH(c, k_tie_rowmap, grid1d(m), 256, 0, gs[cur]->as<uint32_t>(), cstart.as<uint32_t>(), G, m, rowmap.as<uint32_t>(),
             cflag.as<uint8_t>());
      hs_stage_end(c);
      {
        Stage st(c, "tie_next_chunk");
        LAUNCH(c, k_build_nkey_rows, grid1d(m), 256, 0, L, ks, rowmap.as<uint32_t>(), m, d, srcrow, nkey, more);
      }
      hs_segmented_sort_list(c, ms, key, 64, gs[cur]->as<uint32_t>(), gl[cur]->as<uint32_t>(), G, n);
      hs_stage_begin(c, "tie_groups");
      LAUNCH(c, k_tie_flags_rows, grid1d(m), 256, 0, nkey, more, rowmap.as<uint32_t>(), m, cflag.as<uint8_t>());
      hs_exclusive_scan_flags(c, cflag.as<uint8_t>(), crank.as<uint32_t>(), m, ctotal.as<uint64_t>());
      LAUNCH(c, k_head_positions, grid1d(m), 256, 0, cflag.as<uint8_t>(), crank.as<uint32_t>(), (int64_t)m, chead.as<uint32_t>());
      counter.zero();
      LAUNCH(c, k_group_segments_rows, grid1d(m), 256, 0, chead.as<uint32_t>(), ctotal.as<uint64_t>(), m, rowmap.as<uint32_t>(),
             sstart.as<uint32_t>(), slen.as<uint32_t>(), gs[cur ^ 1]->as<uint32_t>(), gl[cur ^ 1]->as<uint32_t>(),
             counter.as<uint32_t>());
      hs_stage_end(c);
      memcpy(cnt, hs_read_wait(c, hs_read_begin(c, counter.p, 8)), 8);
      if (getenv("HSGPU_TRACE"))
        fprintf(stderr, "[hsgpu] tie round %d: %u rows in %u big groups -> %u small groups, %u big groups\n", d, m, G, cnt[0], cnt[1]);
      if (cnt[0]) {
        Stage st(c, "tie_finish");
        LAUNCH(c, k_tie_finish, (cnt[0] + 7) / 8, 256, 0, L, ks, tc, sstart.as<uint32_t>(), slen.as<uint32_t>(), cnt[0],
               d + 1, flag.as<uint8_t>());
      }
      cur ^= 1;
    }
  }

  // ---- assemble the result (buffer A of every user column becomes the output buffer)
  {
    Stage s(c, "finalize");
    for (auto& m : mv) m.b.release();
    for (int i = 0; i < ncols; i++) {
      out[i].type = cols[i].type;
      out[i].length = n;
      out[i].null_count = cols[i].validity ? cols[i].null_count : 0;  // recounted below after an exchange
    }
    for (auto& m : mv) {
      if (m.user_col < 0) continue;
      OutCol& o = out[m.user_col];
      if (m.role == 0) {
        o.data_bytes = n * m.esize;
        o.data = std::move(m.a);
      } else if (m.role == 1 && xchg) {
        // the rows of this table came from every rank: count its NULLs afresh
        DBuf bits(c, (size_t)((n + 7) / 8)), zeros(c, 8);
        zeros.zero();
        LAUNCH(c, k_bytes_to_bits_count, grid1d((n + 7) / 8), 256, 0, m.a.as<uint8_t>(), n, bits.as<uint8_t>(),
               zeros.as<unsigned long long>());
        unsigned long long z = 0;
        CK(cudaMemcpyAsync(&z, zeros.p, 8, cudaMemcpyDeviceToHost, c->stream));
        hs_sync(c);
        o.null_count = (int64_t)z;
        if (z) o.validity = std::move(bits);
      } else {
        DBuf bits(c, (size_t)((n + 7) / 8));
        LAUNCH(c, k_bytes_to_bits, grid1d((n + 7) / 8), 256, 0, m.a.as<uint8_t>(), n, bits.as<uint8_t>());
        if (m.role == 1) o.validity = std::move(bits);
        else { o.data_bytes = (n + 7) / 8; o.data = std::move(bits); }
      }
    }
    for (int i = 0; packed_keys && i < n_indexed; i++) {
      OutCol& o = out[i];
      const int w = hs_fixed_width(cols[i].type);
      o.data_bytes = n * w;
      o.data.alloc(c, (size_t)o.data_bytes + 16);
      DBuf vbytes;
      if (has_nulls[i]) vbytes.alloc(c, (size_t)(n > 0 ? n : 1));
      const uint64_t* sorted_key = (const uint64_t*)ms.a[key_mv];
      uint8_t* vb = has_nulls[i] ? vbytes.as<uint8_t>() : nullptr;
      if (n > 0) switch (w) {
        case 1: LAUNCH(c, k_decode_key<int8_t>, grid1d(n), 256, 0, sorted_key, n, L.g[i], o.data.as<int8_t>(), vb); break;
        case 2: LAUNCH(c, k_decode_key<int16_t>, grid1d(n), 256, 0, sorted_key, n, L.g[i], o.data.as<int16_t>(), vb); break;
        case 4: LAUNCH(c, k_decode_key<int32_t>, grid1d(n), 256, 0, sorted_key, n, L.g[i], o.data.as<int32_t>(), vb); break;
        default: LAUNCH(c, k_decode_key<int64_t>, grid1d(n), 256, 0, sorted_key, n, L.g[i], o.data.as<int64_t>(), vb); break;
      }
      if (has_nulls[i]) {
        DBuf bits(c, (size_t)((n + 7) / 8)), zeros(c, 8);
        zeros.zero();
        LAUNCH(c, k_bytes_to_bits_count, grid1d((n + 7) / 8), 256, 0, vbytes.as<uint8_t>(), n, bits.as<uint8_t>(),
               zeros.as<unsigned long long>());
        unsigned long long z = 0;
        CK(cudaMemcpyAsync(&z, zeros.p, 8, cudaMemcpyDeviceToHost, c->stream));
        hs_sync(c);
        o.null_count = (int64_t)z;
        if (z) o.validity = std::move(bits);
      }
    }
    for (int i = 0; i < ncols; i++) {
      if (!hs_is_varlen(cols[i].type)) continue;
      OutCol g;
      DevCol nov;  // values only: validity already travelled as bytes
      nov.type = cols[i].type; nov.length = n; nov.offsets = cols[i].offsets; nov.data = cols[i].data;
      nov.data_bytes = cols[i].data_bytes;
      hs_gather_column(c, nov, (const uint32_t*)ms.a[srcrow_mv], n, g);
      out[i].type = g.type;
      out[i].offsets = std::move(g.offsets);
      out[i].data = std::move(g.data);
      out[i].data_bytes = g.data_bytes;
    }
  }
}

// ------------------------------------------------------------------ concatenation of runs (merge)
__global__ void __launch_bounds__(256) k_concat_validity(const uint8_t* __restrict__ v, int64_t len, int64_t dst0,
                                                         uint8_t* __restrict__ out_bytes) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < len; i += stride) out_bytes[dst0 + i] = v ? ((v[i >> 3] >> (i & 7)) & 1) : 1;
}
template <typename TI, typename TO>
__global__ void __launch_bounds__(256) k_rebase_offsets(const TI* __restrict__ off, int64_t len, int64_t add,
                                                        TO* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i <= len; i += stride) out[i] = (TO)((int64_t)off[i] + add);
}

static void concat_runs(hs_ctx* c, std::vector<std::vector<DevCol>>& runs, std::vector<DevCol>& out) {
  int ncols = (int)runs[0].size();
  int64_t n = 0;
  for (auto& r : runs) n += r[0].length;
  out.resize(ncols);
  for (int i = 0; i < ncols; i++) {
    DevCol& o = out[i];
    o.type = runs[0][i].type;
    o.length = n;
    bool any_null = false;
    int64_t bytes = 0, nulls = 0;
    for (auto& r : runs) {
      if (r[i].type != o.type) HS_THROW(HS_EINVAL, "runs differ in schema");
      any_null |= r[i].validity != nullptr;
      bytes += r[i].data_bytes;
      nulls += r[i].validity ? r[i].null_count : 0;
    }
    // runs with int32 offsets that together exceed 2 GiB come out with int64 offsets (the LARGE twin of the type)
    const bool in_large = hs_is_large(o.type);
    if (hs_is_varlen(o.type) && !in_large && bytes > 0x7fffffffll) o.type = o.type == HS_UTF8 ? HS_LARGE_UTF8 : HS_LARGE_BINARY;
    if (any_null) {
      DBuf vb(c, (size_t)(n > 0 ? n : 1));
      int64_t at = 0;
      for (auto& r : runs) {
        if (r[i].length) LAUNCH(c, k_concat_validity, grid1d(r[i].length), 256, 0, r[i].validity, r[i].length, at, vb.as<uint8_t>());
        at += r[i].length;
      }
      o.owned.emplace_back(c, (size_t)((n + 7) / 8));
      LAUNCH(c, k_bytes_to_bits, grid1d((n + 7) / 8), 256, 0, vb.as<uint8_t>(), n, o.owned.back().as<uint8_t>());
      o.validity = o.owned.back().as<uint8_t>();
      o.null_count = nulls;
    }
    if (hs_is_varlen(o.type)) {
      const bool out_large = hs_is_large(o.type);
      o.owned.emplace_back(c, (size_t)(n + 1) * (out_large ? 8 : 4));
      int32_t* offs = o.owned.back().as<int32_t>();
      o.owned.emplace_back(c, (size_t)bytes);
      uint8_t* data = o.owned.back().as<uint8_t>();
      int64_t at = 0, bat = 0;
      for (auto& r : runs) {
        const unsigned g = grid1d(r[i].length + 1);
        if (in_large)
          LAUNCH(c, (k_rebase_offsets<int64_t, int64_t>), g, 256, 0, (const int64_t*)r[i].offsets, r[i].length, bat, (int64_t*)offs + at);
        else if (out_large)
          LAUNCH(c, (k_rebase_offsets<int32_t, int64_t>), g, 256, 0, r[i].offsets, r[i].length, bat, (int64_t*)offs + at);
        else
          LAUNCH(c, (k_rebase_offsets<int32_t, int32_t>), g, 256, 0, r[i].offsets, r[i].length, bat, offs + at);
        if (r[i].data_bytes)
          CK(cudaMemcpyAsync(data + bat, r[i].data, (size_t)r[i].data_bytes, cudaMemcpyDeviceToDevice, c->stream));
        at += r[i].length;
        bat += r[i].data_bytes;
      }
      o.offsets = offs;
      o.data = data;
      o.data_bytes = bytes;
    } else if (o.type == HS_BOOL) {
      DBuf vb(c, (size_t)(n > 0 ? n : 1));
      int64_t at = 0;
      for (auto& r : runs) {
        if (r[i].length) LAUNCH(c, k_bits_to_bytes, grid1d(r[i].length), 256, 0, (const uint8_t*)r[i].data, r[i].length, vb.as<uint8_t>() + at);
        at += r[i].length;
      }
      o.owned.emplace_back(c, (size_t)((n + 7) / 8));
      LAUNCH(c, k_bytes_to_bits, grid1d((n + 7) / 8), 256, 0, vb.as<uint8_t>(), n, o.owned.back().as<uint8_t>());
      o.data = o.owned.back().p;
      o.data_bytes = (n + 7) / 8;
    } else {
      o.owned.emplace_back(c, (size_t)bytes);
      uint8_t* data = o.owned.back().as<uint8_t>();
      int64_t bat = 0;
      for (auto& r : runs) {
        if (r[i].data_bytes)
          CK(cudaMemcpyAsync(data + bat, r[i].data, (size_t)r[i].data_bytes, cudaMemcpyDeviceToDevice, c->stream));
        bat += r[i].data_bytes;
      }
      o.data = data;
      o.data_bytes = bytes;
    }
  }
}

static void check_build_args(const hs_column* cols, int32_t ncols, int32_t n_indexed, int32_t nb) {
  if (!cols || ncols <= 0) HS_THROW(HS_EINVAL, "no columns");
  if (n_indexed <= 0 || n_indexed > ncols) HS_THROW(HS_EINVAL, "n_indexed must be in [1, ncols]");
  if (n_indexed > HS_MAX_KEYS) HS_THROW(HS_EUNSUPPORTED, "more than %d indexed columns", HS_MAX_KEYS);
  if (nb <= 0) HS_THROW(HS_EINVAL, "num_buckets must be positive");
}

extern "C" {

int32_t hs_build_index(hs_ctx* ctx, const hs_column* cols, int32_t ncols, int32_t n_indexed, int32_t num_buckets,
                       hs_table** out) {
  API_BEGIN(ctx)
  if (!out) HS_THROW(HS_EINVAL, "null out");
  *out = nullptr;
  check_build_args(cols, ncols, n_indexed, num_buckets);
  hs_stages_reset(ctx);
  std::vector<DevCol> dcols;
  {
    Stage s(ctx, "h2d");
    hs_stage_columns(ctx, cols, ncols, dcols);
  }
  std::vector<OutCol> ocols;
  std::vector<int64_t> boff;
  hs_build_core(ctx, dcols, n_indexed, num_buckets, 0, num_buckets, ocols, boff, nullptr);
  Stage s(ctx, "d2h");
  *out = hs_finish_table(ctx, ocols, dcols[0].length, num_buckets, &boff);
  API_END(ctx)
}

int32_t hs_merge_runs(hs_ctx* ctx, const hs_index_view* runs, int32_t nruns, hs_table** out) {
  API_BEGIN(ctx)
  if (!out || !runs || nruns <= 0) HS_THROW(HS_EINVAL, "bad arguments");
  *out = nullptr;
  hs_stages_reset(ctx);
  for (int r = 0; r < nruns; r++) {
    check_build_args(runs[r].cols, runs[r].ncols, runs[r].n_indexed, runs[r].num_buckets);
    if (runs[r].ncols != runs[0].ncols || runs[r].n_indexed != runs[0].n_indexed ||
        runs[r].num_buckets != runs[0].num_buckets)
      HS_THROW(HS_EINVAL, "runs differ in schema or bucket count");
  }
  std::vector<std::vector<DevCol>> staged((size_t)nruns);
  {
    Stage s(ctx, "h2d");
    for (int r = 0; r < nruns; r++) hs_stage_columns(ctx, runs[r].cols, runs[r].ncols, staged[r]);
  }
  // ---- sorted runs merge pairwise along the merge path when the shape allows it (hs_merge.cu)
  const int32_t nbk = runs[0].num_buckets;
  bool offsets_ok = true;
  for (int r = 0; r < nruns && offsets_ok; r++) {
    const int64_t* o = runs[r].bucket_offsets;
    offsets_ok = o && o[0] == 0 && o[nbk] == staged[r][0].length;
    for (int32_t k = 0; k < nbk && offsets_ok; k++) offsets_ok = o[k + 1] >= o[k];
  }
  if (offsets_ok && nruns >= 2) {
    std::vector<OutCol> cur_out;
    std::vector<int64_t> cur_off(runs[0].bucket_offsets, runs[0].bucket_offsets + nbk + 1);
    std::vector<DevCol> cur;
    std::vector<DevCol>* left = &staged[0];
    bool merged = true;
    for (int r = 1; r < nruns && merged; r++) {
      std::vector<OutCol> next_out;
      std::vector<int64_t> next_off;
      merged = hs_merge_two_runs(ctx, *left, cur_off.data(), staged[r], runs[r].bucket_offsets, runs[0].n_indexed, nbk,
                                 next_out, next_off);
      if (!merged) break;
      cur_out = std::move(next_out);     // the previous intermediate (if any) is released here, in stream order
      cur_off = std::move(next_off);
      cur.clear();
      cur.resize(cur_out.size());
      for (size_t i = 0; i < cur_out.size(); i++) {
        cur[i].type = cur_out[i].type;
        cur[i].length = cur_out[i].length;
        cur[i].null_count = cur_out[i].null_count;
        cur[i].validity = cur_out[i].null_count ? cur_out[i].validity.as<uint8_t>() : nullptr;
        cur[i].data = cur_out[i].data.p;
        cur[i].data_bytes = cur_out[i].data_bytes;
      }
      left = &cur;
    }
    if (merged) {
      Stage s(ctx, "d2h");
      *out = hs_finish_table(ctx, cur_out, cur_out[0].length, nbk, &cur_off);
      return HS_OK;
    }
  }
  std::vector<DevCol> all;
  {
    Stage s(ctx, "concat");
    concat_runs(ctx, staged, all);
  }
  std::vector<OutCol> ocols;
  std::vector<int64_t> boff;
  // Re-bucketing the concatenation gives every row the bucket it already had (same hash, same n),
  // and the per-bucket sort merges the runs: exactly what Spark's repartition + sort does here.
  hs_build_core(ctx, all, runs[0].n_indexed, runs[0].num_buckets, 0, runs[0].num_buckets, ocols, boff, nullptr);
  Stage s(ctx, "d2h");
  *out = hs_finish_table(ctx, ocols, all[0].length, runs[0].num_buckets, &boff);
  API_END(ctx)
}

}  // extern "C"
