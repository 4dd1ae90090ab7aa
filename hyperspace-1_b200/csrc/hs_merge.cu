// hs_merge.cu — two sorted runs per bucket become one without sorting again (merge path).
// This is the compaction of OptimizeAction (HS/actions/OptimizeAction.scala:85-99) and the per-side sort Spark
// plans when a bucket has several files after an incremental refresh (RefreshIncrementalAction.scala:129-144):
// both inputs are already sorted inside every bucket, so a row only has to learn its position in the merged
// bucket. Keys of both runs are brought to one u64 per row (the raw integer key, or the packed tuple of
// hs_pack_key_tuples); tiles of MT outputs find their piece of both runs by a binary search along the merge-path
// diagonal, merge it in shared memory and emit, per output row, where it comes from; every column is then
// gathered from the two runs with that list (reads stay nearly sequential: both runs are consumed in order).
// Ties take run A first (stable). Shapes not covered here — NULL keys, string keys or columns, booleans, key
// ranges wider than 64 bits — are merged by the caller the old way (concatenate + sort).
#include <algorithm>

#include "hs_keys.cuh"
#include "hs_kernels.h"

#define MT 2048          // outputs per tile
#define MTHREADS 256
#define MPER (MT / MTHREADS)
#define FROM_B 0x80000000u

struct MergeParams {
  const uint64_t* ka; const uint64_t* kb;   // u64 keys of both runs (order: unsigned compare)
  const uint32_t* aoff; const uint32_t* boff;   // bucket offsets (device, nb + 1)
  const uint32_t* tile_bucket; const uint32_t* tile_diag;   // per tile: bucket and first output of the tile inside it
  uint32_t ntiles;
  uint32_t* src;          // out: per output row, row of A or FROM_B | row of B
};

// number of A elements among the first d outputs of merge(A[0..na), B[0..nb)), ties to A
__device__ __forceinline__ uint32_t merge_split(const uint64_t* a, uint32_t na, const uint64_t* b, uint32_t nb, uint32_t d) {
  uint32_t lo = d > nb ? d - nb : 0, hi = d < na ? d : na;
  while (lo < hi) {
    const uint32_t mid = (lo + hi) >> 1;
    if (a[mid] <= b[d - 1 - mid]) lo = mid + 1; else hi = mid;
  }
  return lo;
}

__global__ void __launch_bounds__(MTHREADS) k_merge_tiles(const __grid_constant__ MergeParams p) {
  __shared__ uint64_t sa[MT], sb[MT];
  __shared__ uint32_t s_src[MT];
  __shared__ uint32_t s_i0, s_i1;
  const uint32_t t = blockIdx.x;
  const uint32_t bk = p.tile_bucket[t], d0 = p.tile_diag[t];
  const uint32_t a0 = p.aoff[bk], na = p.aoff[bk + 1] - a0, b0 = p.boff[bk], nb = p.boff[bk + 1] - b0;
  const uint32_t total = na + nb;
  const uint32_t d1 = d0 + MT < total ? d0 + MT : total;
  const uint64_t* ga = p.ka + a0;
  const uint64_t* gb = p.kb + b0;
  if (threadIdx.x == 0) s_i0 = merge_split(ga, na, gb, nb, d0);
  if (threadIdx.x == 32) s_i1 = merge_split(ga, na, gb, nb, d1);
  __syncthreads();
  const uint32_t i0 = s_i0, i1 = s_i1, j0 = d0 - i0, j1 = d1 - i1;
  const uint32_t wa = i1 - i0, wb = j1 - j0;      // wa + wb = d1 - d0 <= MT
  for (uint32_t k = threadIdx.x; k < wa; k += MTHREADS) sa[k] = ga[i0 + k];
  for (uint32_t k = threadIdx.x; k < wb; k += MTHREADS) sb[k] = gb[j0 + k];
  __syncthreads();
  const uint32_t n = wa + wb;
  uint32_t d = threadIdx.x * MPER;
  if (d < n) {
    uint32_t i = merge_split(sa, wa, sb, wb, d), j = d - i;
#pragma unroll
    for (int e = 0; e < MPER; e++) {
      if (d + e >= n) break;
      const bool take_a = j >= wb || (i < wa && sa[i] <= sb[j]);
      s_src[d + e] = take_a ? (a0 + i0 + i) : (FROM_B | (b0 + j0 + j));
      if (take_a) i++; else j++;
    }
  }
  __syncthreads();
  uint32_t* out = p.src + (size_t)a0 + b0 + d0;
  for (uint32_t k = threadIdx.x; k < n; k += MTHREADS) out[k] = s_src[k];
}

template <typename T>
__global__ void __launch_bounds__(256) k_gather2(const T* __restrict__ a, const T* __restrict__ b,
                                                 const uint32_t* __restrict__ src, int64_t m, T* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < m; i += stride) {
    const uint32_t s = src[i];
    out[i] = (s & FROM_B) ? b[s & ~FROM_B] : a[s];
  }
}
// validity of a column that has a bitmap in at least one run (NULL bitmap = all valid); one thread per output byte
__global__ void __launch_bounds__(256) k_gather2_bits(const uint8_t* __restrict__ va, const uint8_t* __restrict__ vb,
                                                      const uint32_t* __restrict__ src, int64_t m,
                                                      uint8_t* __restrict__ out, unsigned long long* ones) {
  const int64_t nbytes = (m + 7) >> 3;
  int64_t bt = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  unsigned cnt = 0;
  for (; bt < nbytes; bt += stride) {
    unsigned v = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
      const int64_t i = bt * 8 + k;
      if (i < m) {
        const uint32_t s = src[i], r = s & ~FROM_B;
        const uint8_t* bits = (s & FROM_B) ? vb : va;
        v |= (bits ? ((bits[r >> 3] >> (r & 7)) & 1u) : 1u) << k;
      }
    }
    out[bt] = (uint8_t)v;
    cnt += __popc(v);
  }
  for (int d = 16; d; d >>= 1) cnt += __shfl_down_sync(0xffffffffu, cnt, d);
  if (lane_id() == 0 && cnt) atomicAdd(ones, (unsigned long long)cnt);
}
// raw signed integer key column -> order-preserving u64
template <typename T>
__global__ void __launch_bounds__(256) k_int_keys(const T* __restrict__ in, int64_t n, uint64_t* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) out[i] = (uint64_t)(int64_t)in[i] ^ 0x8000000000000000ull;
}

static unsigned mgrid(int64_t n) {
  int64_t g = (n + 255) / 256, cap = (int64_t)HS_NUM_SMS * 16;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (unsigned)g;
}

bool hs_merge_two_runs(hs_ctx* c, std::vector<DevCol>& a, const int64_t* aoff, std::vector<DevCol>& b, const int64_t* boff_b,
                       int nk, int32_t nb, std::vector<OutCol>& out, std::vector<int64_t>& out_off) {
  const char* off = getenv("HSGPU_MERGE_BY_SORT");
  if (off && *off && *off != '0') return false;
  const int ncols = (int)a.size();
  const int64_t na = a[0].length, nbr = b[0].length, m = na + nbr;
  if (m >= (int64_t)0x7FFFFFF0ll) return false;
  for (int i = 0; i < ncols; i++) {
    const int w = hs_fixed_width(a[i].type);
    if (w == 0 || a[i].type != b[i].type) return false;            // strings, booleans: the general path
    if (i < nk && (a[i].validity || b[i].validity)) return false;   // NULL keys
  }
  // ---- u64 keys of both runs
  DBuf ka, kb;
  {
    Stage s(c, "merge_keys");
    if (nk == 1 && (a[0].type == HS_INT64 || a[0].type == HS_INT32)) {
      ka.alloc(c, (size_t)(na > 0 ? na : 1) * 8);
      kb.alloc(c, (size_t)(nbr > 0 ? nbr : 1) * 8);
      if (a[0].type == HS_INT64) {
        if (na) LAUNCH(c, k_int_keys<int64_t>, mgrid(na), 256, 0, (const int64_t*)a[0].data, na, ka.as<uint64_t>());
        if (nbr) LAUNCH(c, k_int_keys<int64_t>, mgrid(nbr), 256, 0, (const int64_t*)b[0].data, nbr, kb.as<uint64_t>());
      } else {
        if (na) LAUNCH(c, k_int_keys<int32_t>, mgrid(na), 256, 0, (const int32_t*)a[0].data, na, ka.as<uint64_t>());
        if (nbr) LAUNCH(c, k_int_keys<int32_t>, mgrid(nbr), 256, 0, (const int32_t*)b[0].data, nbr, kb.as<uint64_t>());
      }
    } else {
      KeySpec ksa = hs_make_keyspec(a, nk), ksb = hs_make_keyspec(b, nk);
      if (!hs_pack_key_tuples(c, ksa, na, ksb, nbr, nk, ka, kb)) return false;
    }
  }
  // ---- tiles along the merge path of every bucket
  std::vector<uint32_t> tb, td, ha((size_t)nb + 1), hb((size_t)nb + 1);
  out_off.assign((size_t)nb + 1, 0);
  for (int32_t k = 0; k <= nb; k++) {
    ha[k] = (uint32_t)aoff[k];
    hb[k] = (uint32_t)boff_b[k];
    out_off[k] = aoff[k] + boff_b[k];
  }
  for (int32_t k = 0; k < nb; k++) {
    const int64_t tot = (aoff[k + 1] - aoff[k]) + (boff_b[k + 1] - boff_b[k]);
    for (int64_t d = 0; d < tot; d += MT) { tb.push_back((uint32_t)k); td.push_back((uint32_t)d); }
  }
  const uint32_t ntiles = (uint32_t)tb.size();
  DBuf d_tb(c, (size_t)ntiles * 4 + 4), d_td(c, (size_t)ntiles * 4 + 4), d_a(c, ((size_t)nb + 1) * 4), d_b(c, ((size_t)nb + 1) * 4);
  DBuf src(c, (size_t)(m > 0 ? m : 1) * 4);
  {
    Stage s(c, "merge_path");
    if (ntiles) {
      CK(cudaMemcpyAsync(d_tb.p, tb.data(), (size_t)ntiles * 4, cudaMemcpyHostToDevice, c->stream));
      CK(cudaMemcpyAsync(d_td.p, td.data(), (size_t)ntiles * 4, cudaMemcpyHostToDevice, c->stream));
    }
    CK(cudaMemcpyAsync(d_a.p, ha.data(), ((size_t)nb + 1) * 4, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(d_b.p, hb.data(), ((size_t)nb + 1) * 4, cudaMemcpyHostToDevice, c->stream));
    MergeParams p;
    p.ka = ka.as<uint64_t>(); p.kb = kb.as<uint64_t>();
    p.aoff = d_a.as<uint32_t>(); p.boff = d_b.as<uint32_t>();
    p.tile_bucket = d_tb.as<uint32_t>(); p.tile_diag = d_td.as<uint32_t>();
    p.ntiles = ntiles;
    p.src = src.as<uint32_t>();
    if (ntiles) LAUNCH(c, k_merge_tiles, ntiles, MTHREADS, 0, p);
    hs_sync(c);   // the host vectors above must outlive their copies
  }
  ka.release();
  kb.release();
  // ---- every column through the source list
  Stage s(c, "merge_gather");
  out.clear();
  out.resize((size_t)ncols);
  for (int i = 0; i < ncols; i++) {
    OutCol& o = out[i];
    const int w = hs_fixed_width(a[i].type);
    o.type = a[i].type;
    o.length = m;
    o.null_count = 0;
    o.data_bytes = m * w;
    o.data.alloc(c, (size_t)o.data_bytes + 16);
    if (m > 0) {
      const unsigned g = mgrid(m);
      switch (w) {
        case 1: LAUNCH(c, k_gather2<uint8_t>, g, 256, 0, (const uint8_t*)a[i].data, (const uint8_t*)b[i].data, src.as<uint32_t>(), m, o.data.as<uint8_t>()); break;
        case 2: LAUNCH(c, k_gather2<uint16_t>, g, 256, 0, (const uint16_t*)a[i].data, (const uint16_t*)b[i].data, src.as<uint32_t>(), m, o.data.as<uint16_t>()); break;
        case 4: LAUNCH(c, k_gather2<uint32_t>, g, 256, 0, (const uint32_t*)a[i].data, (const uint32_t*)b[i].data, src.as<uint32_t>(), m, o.data.as<uint32_t>()); break;
        default: LAUNCH(c, k_gather2<uint64_t>, g, 256, 0, (const uint64_t*)a[i].data, (const uint64_t*)b[i].data, src.as<uint32_t>(), m, o.data.as<uint64_t>()); break;
      }
    }
    if ((a[i].validity || b[i].validity) && m > 0) {
      o.validity.alloc(c, (size_t)((m + 7) / 8));
      DBuf ones(c, 8);
      ones.zero();
      LAUNCH(c, k_gather2_bits, mgrid((m + 7) / 8), 256, 0, a[i].validity, b[i].validity, src.as<uint32_t>(), m,
             o.validity.as<uint8_t>(), ones.as<unsigned long long>());
      unsigned long long h = 0;
      CK(cudaMemcpyAsync(&h, ones.p, 8, cudaMemcpyDeviceToHost, c->stream));
      hs_sync(c);
      o.null_count = m - (int64_t)h;
      if (o.null_count == 0) o.validity.release();
    }
  }
  return true;
}
