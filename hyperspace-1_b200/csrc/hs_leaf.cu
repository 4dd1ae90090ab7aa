// hs_leaf.cu — K3, last step: every leaf segment (at most LEAF_MAX rows) sorted inside shared memory and written
// out column by column. Compiled with its own tile geometry — 384 threads x 8 rows, three CTAs per SM (see
// hs_sort_shared.cuh) — which is why it is a translation unit of its own.
#define PT_THREADS 384
#define PT_MIN_CTAS 3
#include "hs_block.cuh"
#include "hs_sort_shared.cuh"
#include "hs_tma.cuh"

static_assert(LEAF_MAX == PT_TILE, "a leaf is one tile of this translation unit");
#define STAGE_BYTES (PT_TILE * 8 + 128)   // one column slice of a leaf + alignment slack


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

// Persistent grid over the leaf list: only the segments the packed kernel flagged are sorted here, and when it
// flagged none every CTA leaves at once — so the launch needs no answer from the device first.
__global__ void __launch_bounds__(PT_THREADS, 2 * PT_MIN_CTAS / 2) k_leaf_sort(const __grid_constant__ LeafParams p) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  if (p.wide && p.n_wide && *p.n_wide == 0) return;
  const uint32_t n = *p.count;
  for (uint32_t s = blockIdx.x; s < n; s += gridDim.x) {
    if (p.wide && !p.wide[s]) continue;  // already sorted by the packed 32-bit kernel
    const uint32_t start = p.segs.start[s], len = p.segs.len[s];
    leaf64_body(smem_raw, p.mv.esize, p.mv.n, p.mv.src, p.mv.dst, p.key, start, start, len, p.in_place != 0);
    __syncthreads();
  }
}

#define L32_IDX_BITS 12
// counters of the two ranking passes: 2^bits atomic counters in pass A, PT_WARPS x 256 warp counters in pass B
#define L32_CNT_WORDS (PT_WARPS * 256 > 2048 ? PT_WARPS * 256 : 2048)
#define L32_MAX_KEY_BITS (L32_CNT_WORDS >= 4096 ? 20 : 19)
#define L32_ATOMIC_BITS 11   // the atomic pass takes min(nbits, 11) bits (12 when 20 bits vary and the counters allow)

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
  uint32_t* cnt = sdata + LEAF_MAX;                 // L32_CNT_WORDS: atomic counters or PT_WARPS x 256 warp counters
  uint32_t* tot = cnt + L32_CNT_WORDS;              // 256
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
  return (size_t)2 * STAGE_BYTES + (size_t)LEAF_MAX * 4 + L32_CNT_WORDS * 4 + 512 * 4 + PT_SCRATCH * 4 +
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
void hs_leaf_set_attrs() {
  static bool done_dev[64] = {false};
  int dev = 0;
  CK(cudaGetDevice(&dev));
  bool& done = done_dev[dev & 63];
  if (done) return;
  CK(cudaFuncSetAttribute(k_leaf_sort32<KS_U64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)leaf32_smem_bytes()));
  CK(cudaFuncSetAttribute(k_leaf_sort32<KS_I64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)leaf32_smem_bytes()));
  CK(cudaFuncSetAttribute(k_leaf_sort32<KS_I32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)leaf32_smem_bytes()));
  CK(cudaFuncSetAttribute(k_leaf_sort, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)leaf_smem_bytes()));
  done = true;
}

// Launch the leaf and copy kernels over device-resident lists. The grids are upper bounds: the kernels
// read the real list lengths from `counters` ([0] leaves, [2] copies, [4] receives the number of segments
// the packed kernel could not take), so the host does not have to wait for them before launching.
void run_leaf_and_copy(hs_ctx* c, MoveSet& ms, const KeySrc& key, int src_code, const ListBufs& leaf, uint32_t max_leaf,
                       const ListBufs& copy, uint32_t max_copy, uint32_t* counters, LeafLaunch& ll, const ListBufs* small,
                       uint32_t max_small) {
  hs_leaf_set_attrs();
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
// The 64-bit kernel for the segments the packed kernel flagged (it reads the flags and their count itself).
void launch_wide(hs_ctx* c, LeafLaunch& ll, uint32_t* d_counters) {
  (void)d_counters;
  if (!ll.max_leaf) return;
  Stage st(c, "leaf_sort_wide");
  const uint32_t grid = std::min<uint32_t>(ll.max_leaf, (uint32_t)HS_NUM_SMS * 4u);
  LAUNCH(c, k_leaf_sort, grid, PT_THREADS, leaf_smem_bytes(), ll.lp);
}
