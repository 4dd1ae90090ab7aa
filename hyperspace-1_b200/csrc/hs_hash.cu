// hs_hash.cu — K1: Spark-compatible Murmur3 bucket ids, the per-bucket histogram (plain shared-memory
// atomics: measured on B200 they run at ~11 lanes/clk/SM on spread addresses, ~20x cheaper per row than
// aggregating with __match_any_sync, and the hardware already collapses a warp that hits one address) and
// the per-column key ranges the sort uses.
// Stands in for HashPartitioning.partitionIdExpression = Pmod(Murmur3Hash(cols, 42), n)
// reached from df.repartition(numBuckets, indexedCols) at
// HS/actions/CreateActionBase.scala:129-130.
#include "hs_keys.cuh"
#include "hs_kernels.h"

#define HASH_THREADS 256

__device__ __forceinline__ void minmax_flush(uint64_t mn, uint64_t mx, uint64_t* g) {
#pragma unroll
  for (int d = 16; d; d >>= 1) {
    uint64_t a = __shfl_xor_sync(0xffffffffu, mn, d), b = __shfl_xor_sync(0xffffffffu, mx, d);
    mn = a < mn ? a : mn;
    mx = b > mx ? b : mx;
  }
  if (lane_id() == 0 && mn <= mx) {
    atomicMin((unsigned long long*)&g[0], (unsigned long long)mn);
    atomicMax((unsigned long long*)&g[1], (unsigned long long)mx);
  }
}

// MODE 0: out32 = raw hash; 1: out32 = bucket (any number of key columns)
template <int MODE>
__global__ void __launch_bounds__(HASH_THREADS) k_hash_generic(KeySpec ks, int64_t n, int32_t nb, FastPmod fm,
                                                               int32_t* __restrict__ out32) {
  int64_t stride = (int64_t)gridDim.x * HASH_THREADS;
  for (int64_t i = (int64_t)blockIdx.x * HASH_THREADS + threadIdx.x; i < n; i += stride) {
    uint32_t h = mm_hash_row(ks, i);
    out32[i] = MODE == 0 ? (int32_t)h : (int32_t)fast_pmod(h, fm);
  }
}

// The build's K1 for any key shape: out16 = local bucket id, hist[nbl] += 1, min/max of every fixed-width key
// column. NK = number of key columns (compile time: the per-column state stays in registers).
template <int NK>
__global__ void __launch_bounds__(HASH_THREADS) k_hash_build(KeySpec ks, int64_t n, FastPmod fm, int32_t blo,
                                                             int32_t nbl, uint16_t* __restrict__ out16,
                                                             uint32_t* __restrict__ g_hist, int smem_hist,
                                                             uint64_t* __restrict__ g_minmax) {
  extern __shared__ uint32_t s_hist[];
  uint32_t* hist = g_hist;
  if (smem_hist) {
    for (int b = threadIdx.x; b < nbl; b += HASH_THREADS) s_hist[b] = 0;
    __syncthreads();
    hist = s_hist;
  }
  uint64_t mn[NK], mx[NK];
#pragma unroll
  for (int k = 0; k < NK; k++) { mn[k] = ~0ull; mx[k] = 0; }
  int64_t stride = (int64_t)gridDim.x * HASH_THREADS;
  for (int64_t i = (int64_t)blockIdx.x * HASH_THREADS + threadIdx.x; i < n; i += stride) {
    uint32_t h = 42u;
#pragma unroll
    for (int k = 0; k < NK; k++) {
      uint64_t m;
      bool has;
      h = mm_hash_and_map(ks.k[k], i, h, &m, &has);
      if (has) {
        mn[k] = m < mn[k] ? m : mn[k];
        mx[k] = m > mx[k] ? m : mx[k];
      }
    }
    int32_t b = (int32_t)fast_pmod(h, fm) - blo;
    if (b < 0 || b >= nbl) {  // a row that does not belong to this rank's bucket range
      atomicAdd((unsigned long long*)&g_minmax[2 * HS_MAX_KEYS], 1ull);
      b = 0;
    }
    out16[i] = (uint16_t)b;
    atomicAdd(&hist[b], 1u);   // shared atomics: ~11 lanes/clk/SM; a hot bucket serialises within the warp only
  }
#pragma unroll
  for (int k = 0; k < NK; k++) minmax_flush(mn[k], mx[k], g_minmax + 2 * k);
  if (smem_hist) {
    __syncthreads();
    for (int b = threadIdx.x; b < nbl; b += HASH_THREADS) {
      uint32_t v = s_hist[b];
      if (v) atomicAdd(&g_hist[b], v);
    }
  }
}

// Fast path: one non-null INT64 key. Each thread hashes two rows from one 128-bit load and
// stores both 16-bit bucket ids with one 32-bit store.
#define HASH_UNROLL 2
__global__ void __launch_bounds__(HASH_THREADS) k_hash_i64x2(const uint64_t* __restrict__ keys, int64_t n,
                                                             int32_t nb, FastPmod fm, uint16_t* __restrict__ out16,
                                                             uint32_t* __restrict__ g_hist,
                                                             uint64_t* __restrict__ g_minmax) {
  extern __shared__ uint32_t s_hist[];
  for (int b = threadIdx.x; b < nb; b += HASH_THREADS) s_hist[b] = 0;
  __syncthreads();
  uint64_t mn = ~0ull, mx = 0;
  long long smn = 0x7fffffffffffffffll, smx = -0x7fffffffffffffffll - 1;
  int64_t npairs = n >> 1;
  int64_t stride = (int64_t)gridDim.x * HASH_THREADS;
  const ulonglong2* k2 = (const ulonglong2*)keys;
  // HASH_UNROLL 128-bit loads in flight per thread before the first dependent instruction (the loads are the
  // latency: 0.30 ms with one, 0.27 with two)
  for (int64_t base = (int64_t)blockIdx.x * HASH_THREADS; base < npairs; base += HASH_UNROLL * stride) {
    ulonglong2 v[HASH_UNROLL];
#pragma unroll
    for (int h = 0; h < HASH_UNROLL; h++) {
      const int64_t p = base + threadIdx.x + h * stride;
      v[h].x = v[h].y = 0;
      if (p < npairs) asm volatile("ld.global.nc.L1::no_allocate.v2.u64 {%0, %1}, [%2];" : "=l"(v[h].x), "=l"(v[h].y) : "l"(k2 + p));
    }
#pragma unroll
    for (int h = 0; h < HASH_UNROLL; h++) {
      const int64_t p = base + threadIdx.x + h * stride;
      if (p >= npairs) continue;
      const uint32_t b0 = fast_pmod(mm_hash_long(v[h].x, 42u), fm);
      const uint32_t b1 = fast_pmod(mm_hash_long(v[h].y, 42u), fm);
      ((uint32_t*)out16)[p] = b0 | (b1 << 16);
      const long long s0 = (long long)v[h].x, s1 = (long long)v[h].y;
      smn = min(smn, min(s0, s1));
      smx = max(smx, max(s0, s1));
      // plain shared atomics: ~11 lanes/clk/SM on spread buckets; when both rows of a thread share a bucket
      // (hot key) they collapse to one update
      if (b0 == b1) atomicAdd(&s_hist[b0], 2u);
      else { atomicAdd(&s_hist[b0], 1u); atomicAdd(&s_hist[b1], 1u); }
    }
  }
  if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
    uint64_t v = keys[n - 1];
    uint32_t b = fast_pmod(mm_hash_long(v, 42u), fm);
    out16[n - 1] = (uint16_t)b;
    atomicAdd(&s_hist[b], 1u);
    uint64_t m = v ^ 0x8000000000000000ull;
    mn = m < mn ? m : mn;
    mx = m > mx ? m : mx;
  }
  if (smn <= smx) {
    uint64_t a = (uint64_t)smn ^ 0x8000000000000000ull, b = (uint64_t)smx ^ 0x8000000000000000ull;
    mn = a < mn ? a : mn;
    mx = b > mx ? b : mx;
  }
  minmax_flush(mn, mx, g_minmax);
  __syncthreads();
  for (int b = threadIdx.x; b < nb; b += HASH_THREADS) {
    uint32_t v = s_hist[b];
    if (v) atomicAdd(&g_hist[b], v);
  }
}

static unsigned hash_grid(int64_t n) {
  int64_t g = (n + HASH_THREADS - 1) / HASH_THREADS;
  int64_t cap = (int64_t)HS_NUM_SMS * 8;  // 8 resident CTAs of 256 threads per SM
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (unsigned)g;
}

KeySpec hs_make_keyspec(const std::vector<DevCol>& cols, int nkeys) {
  if (nkeys <= 0 || nkeys > HS_MAX_KEYS) HS_THROW(HS_EINVAL, "between 1 and %d key columns are supported", HS_MAX_KEYS);
  KeySpec ks;
  memset(&ks, 0, sizeof ks);
  ks.nkeys = nkeys;
  for (int k = 0; k < nkeys; k++) ks.k[k] = hs_kcol(cols[k]);
  return ks;
}

void hs_hash_to_i32(hs_ctx* c, const KeySpec& ks, int64_t n, int32_t nb, int32_t* d_out) {
  if (n == 0) return;
  if (nb <= 0) LAUNCH(c, k_hash_generic<0>, hash_grid(n), HASH_THREADS, 0, ks, n, 1, make_fastpmod(1), d_out);
  else LAUNCH(c, k_hash_generic<1>, hash_grid(n), HASH_THREADS, 0, ks, n, nb, make_fastpmod(nb), d_out);
}

// Local bucket ids (bucket - blo, u16) + histogram over the nbl buckets [blo, blo + nbl) + min/max of
// every fixed-width key column. d_hist: nbl u32 (zeroed here); d_minmax: 2*HS_MAX_KEYS + 1 u64
// (initialised here; the last word counts rows outside the bucket range).
void hs_hash_bucket_hist(hs_ctx* c, const KeySpec& ks, const std::vector<DevCol>& cols, int64_t n, int32_t nb,
                         int32_t blo, int32_t nbl, uint16_t* d_bucket, uint32_t* d_hist, uint64_t* d_minmax) {
  if (nbl > 65536) HS_THROW(HS_EUNSUPPORTED, "more than 65536 buckets on one GPU");
  CK(cudaMemsetAsync(d_hist, 0, (size_t)nbl * 4, c->stream));
  uint64_t init[2 * HS_MAX_KEYS + 1];
  init[2 * HS_MAX_KEYS] = 0;
  for (int k = 0; k < HS_MAX_KEYS; k++) { init[2 * k] = ~0ull; init[2 * k + 1] = 0; }
  CK(cudaMemcpyAsync(d_minmax, init, sizeof init, cudaMemcpyHostToDevice, c->stream));
  if (n == 0) return;
  size_t smem = (size_t)nbl * 4;
  bool fast = blo == 0 && nbl == nb && ks.nkeys == 1 && ks.k[0].type == HS_INT64 && !ks.k[0].has_nulls &&
              (((uintptr_t)ks.k[0].data) & 15) == 0 && (((uintptr_t)d_bucket) & 3) == 0 && smem <= 48 * 1024;
  if (fast) {
    LAUNCH(c, k_hash_i64x2, hash_grid((n + 1) / 2), HASH_THREADS, smem, (const uint64_t*)ks.k[0].data, n, nb,
           make_fastpmod(nb), d_bucket, d_hist, d_minmax);
    return;
  }
  int smem_hist = smem <= 160 * 1024;
  size_t sm = smem_hist ? smem : 0;
  FastPmod fm = make_fastpmod(nb);
#define HS_LAUNCH_BUILD(NK)                                                                                        \
  do {                                                                                                             \
    if (sm > 48 * 1024) CK(cudaFuncSetAttribute(k_hash_build<NK>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm)); \
    LAUNCH(c, k_hash_build<NK>, hash_grid(n), HASH_THREADS, sm, ks, n, fm, blo, nbl, d_bucket, d_hist, smem_hist,   \
           d_minmax);                                                                                              \
  } while (0)
  switch (ks.nkeys) {
    case 1: HS_LAUNCH_BUILD(1); break;
    case 2: HS_LAUNCH_BUILD(2); break;
    case 3: HS_LAUNCH_BUILD(3); break;
    case 4: HS_LAUNCH_BUILD(4); break;
    case 5: HS_LAUNCH_BUILD(5); break;
    case 6: HS_LAUNCH_BUILD(6); break;
    case 7: HS_LAUNCH_BUILD(7); break;
    default: HS_LAUNCH_BUILD(8); break;
  }
#undef HS_LAUNCH_BUILD
}

extern "C" {

static int32_t hash_api(hs_ctx* ctx, const hs_column* keys, int32_t nkeys, int32_t nb, int32_t* out) {
  API_BEGIN(ctx)
  if (!keys || !out) HS_THROW(HS_EINVAL, "null argument");
  hs_stages_reset(ctx);
  std::vector<DevCol> cols;
  hs_stage_columns(ctx, keys, nkeys, cols);
  KeySpec ks = hs_make_keyspec(cols, nkeys);
  int64_t n = cols[0].length;
  bool dev_out = hs_is_device_ptr(out);
  DBuf tmp;
  int32_t* d_out = out;
  if (!dev_out) { tmp.alloc(ctx, (size_t)n * 4); d_out = tmp.as<int32_t>(); }
  {
    Stage s(ctx, "hash");
    hs_hash_to_i32(ctx, ks, n, nb, d_out);
  }
  if (!dev_out && n) CK(cudaMemcpyAsync(out, d_out, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  hs_sync(ctx);
  API_END(ctx)
}

int32_t hs_hash_bucket(hs_ctx* ctx, const hs_column* keys, int32_t nkeys, int32_t num_buckets, int32_t* out) {
  if (ctx && num_buckets <= 0) { ctx->err = "num_buckets must be positive"; return HS_EINVAL; }
  return hash_api(ctx, keys, nkeys, num_buckets, out);
}
int32_t hs_hash_rows(hs_ctx* ctx, const hs_column* keys, int32_t nkeys, int32_t* out) {
  return hash_api(ctx, keys, nkeys, 0, out);
}

}  // extern "C"
