// hs_scan.cu — device-wide exclusive scan (reduce / scan-of-sums / apply), used for
// string offsets, tie-group discovery, join output offsets and filter compaction.
#include "hs_common.cuh"

#define SCAN_THREADS 256
#define SCAN_ITEMS 16
#define SCAN_TILE (SCAN_THREADS * SCAN_ITEMS)

__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v) {
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    uint32_t t = __shfl_up_sync(0xffffffffu, v, d);
    if (lane_id() >= (uint32_t)d) v += t;
  }
  return v;
}

// exclusive scan of one value per thread across a 256-thread block; returns block total in *total
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t* total) {
  __shared__ uint32_t wsum[SCAN_THREADS / 32];
  __shared__ uint32_t wtot;
  uint32_t inc = warp_incl_scan(v);
  int w = threadIdx.x >> 5;
  if (lane_id() == 31) wsum[w] = inc;
  __syncthreads();
  if (w == 0) {
    uint32_t s = lane_id() < SCAN_THREADS / 32 ? wsum[lane_id()] : 0;
    uint32_t si = warp_incl_scan(s);
    if (lane_id() < SCAN_THREADS / 32) wsum[lane_id()] = si - s;
    if (lane_id() == SCAN_THREADS / 32 - 1) wtot = si;
  }
  __syncthreads();
  uint32_t r = inc - v + wsum[w];
  *total = wtot;
  __syncthreads();
  return r;
}

template <typename T>
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_reduce(const T* __restrict__ in, int64_t n,
                                                               uint64_t* __restrict__ sums) {
  int64_t base = (int64_t)blockIdx.x * SCAN_TILE;
  uint32_t s = 0;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) {
    int64_t i = base + k * SCAN_THREADS + threadIdx.x;
    if (i < n) s += (uint32_t)in[i];
  }
  uint32_t tot;
  block_excl_scan(s, &tot);
  if (threadIdx.x == 0) sums[blockIdx.x] = tot;
}

__global__ void __launch_bounds__(1024) k_scan_sums(uint64_t* sums, int64_t nb, uint64_t* total) {
  __shared__ uint64_t carry;
  __shared__ uint64_t wsum[32];
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int64_t base = 0; base < nb; base += 1024) {
    int64_t i = base + threadIdx.x;
    uint64_t v = i < nb ? sums[i] : 0, inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      uint64_t t = __shfl_up_sync(0xffffffffu, inc, d);
      if (lane_id() >= (uint32_t)d) inc += t;
    }
    int w = threadIdx.x >> 5;
    if (lane_id() == 31) wsum[w] = inc;
    __syncthreads();
    if (w == 0) {
      uint64_t s = wsum[lane_id()], si = s;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        uint64_t t = __shfl_up_sync(0xffffffffu, si, d);
        if (lane_id() >= (uint32_t)d) si += t;
      }
      wsum[lane_id()] = si - s;
    }
    __syncthreads();
    uint64_t excl = inc - v + wsum[w] + carry;
    if (i < nb) sums[i] = excl;
    __syncthreads();
    if (threadIdx.x == 1023) carry = excl + v;
    __syncthreads();
  }
  if (threadIdx.x == 0 && total) *total = carry;
}

template <typename T, typename TO>
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_apply(const T* in, TO* out, int64_t n,
                                                              const uint64_t* __restrict__ sums, int write_total) {
  int64_t base = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
  uint32_t v[SCAN_ITEMS];
  uint32_t s = 0;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) {
    int64_t i = base + k;
    v[k] = i < n ? (uint32_t)in[i] : 0;
    s += v[k];
  }
  uint32_t tot;
  TO ex = (TO)block_excl_scan(s, &tot) + (TO)sums[blockIdx.x];
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) {
    int64_t i = base + k;
    if (i < n) out[i] = ex;
    ex += v[k];
    if (write_total && i == n - 1) out[n] = ex;   // offsets arrays carry the total as their last entry
  }
}

template <typename T>
static void scan_impl(hs_ctx* c, const T* in, uint32_t* out, int64_t n, uint64_t* d_total) {
  if (n <= 0) {
    if (d_total) CK(cudaMemsetAsync(d_total, 0, 8, c->stream));
    return;
  }
  int64_t nb = (n + SCAN_TILE - 1) / SCAN_TILE;
  DBuf sums(c, (size_t)nb * 8);
  LAUNCH(c, k_scan_reduce<T>, (unsigned)nb, SCAN_THREADS, 0, in, n, sums.as<uint64_t>());
  LAUNCH(c, k_scan_sums, 1, 1024, 0, sums.as<uint64_t>(), nb, d_total);
  LAUNCH(c, (k_scan_apply<T, uint32_t>), (unsigned)nb, SCAN_THREADS, 0, in, out, n, sums.as<uint64_t>(), 0);
}

void hs_exclusive_scan_u32(hs_ctx* c, const uint32_t* in, uint32_t* out, int64_t n, uint64_t* d_total) {
  scan_impl<uint32_t>(c, in, out, n, d_total);
}
void hs_exclusive_scan_flags(hs_ctx* c, const uint8_t* flags, uint32_t* out, int64_t n, uint64_t* d_total) {
  scan_impl<uint8_t>(c, flags, out, n, d_total);
}

// lens[0..m) -> Arrow offsets (m+1 entries) in `out`, int32 while the bytes fit 2 GiB, int64 (and the LARGE twin
// of `type`) beyond that or when `type` is LARGE already. Returns the total number of bytes.
int64_t hs_offsets_from_lens(hs_ctx* c, const uint32_t* lens, int64_t m, int32_t type, OutCol& out) {
  bool large = hs_is_large(type);
  if (m <= 0) {
    out.type = type;
    out.offsets.alloc(c, large ? 8 : 4);
    out.offsets.zero();
    return 0;
  }
  int64_t nb = (m + SCAN_TILE - 1) / SCAN_TILE;
  DBuf sums(c, (size_t)nb * 8), total(c, 8);
  LAUNCH(c, k_scan_reduce<uint32_t>, (unsigned)nb, SCAN_THREADS, 0, lens, m, sums.as<uint64_t>());
  LAUNCH(c, k_scan_sums, 1, 1024, 0, sums.as<uint64_t>(), nb, total.as<uint64_t>());
  uint64_t h = 0;
  CK(cudaMemcpyAsync(&h, total.p, 8, cudaMemcpyDeviceToHost, c->stream));
  hs_sync(c);
  if (!large && h > 0x7fffffffull) {
    large = true;
    type = type == HS_UTF8 ? HS_LARGE_UTF8 : HS_LARGE_BINARY;
  }
  out.type = type;
  out.offsets.alloc(c, (size_t)(m + 1) * (large ? 8 : 4));
  if (large)
    LAUNCH(c, (k_scan_apply<uint32_t, int64_t>), (unsigned)nb, SCAN_THREADS, 0, lens, out.offsets.as<int64_t>(), m,
           sums.as<uint64_t>(), 1);
  else
    LAUNCH(c, (k_scan_apply<uint32_t, int32_t>), (unsigned)nb, SCAN_THREADS, 0, lens, out.offsets.as<int32_t>(), m,
           sums.as<uint64_t>(), 1);
  return (int64_t)h;
}
