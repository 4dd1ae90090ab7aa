// hs_noalloc.cu — hs_bucket_histogram_noalloc: bucket ids + per-bucket row counts without taking any
// memory from the CUDA allocator.
//
// Who this is for. Every other entry point of libhsgpu borrows scratch from the stream-ordered
// allocator. Two kinds of caller cannot do that:
//   * a process whose GPU memory already belongs to somebody else's pool — e.g. a Spark executor
//     running next to an RMM/arena allocator that reserved the whole device at start-up, where a
//     further cudaMalloc simply fails;
//   * a latency-critical thread that may not stall the other streams of the device: cudaMalloc and
//     cudaFree synchronise the whole GPU, which a concurrently running query cannot afford.
// For them the histogram counters of the K1 kernel — the only scratch that kernel needs — live in a
// fixed array compiled into the library image instead of in allocated memory.
//
// Consequences the caller must know about:
//   * the array is HS_NOALLOC_MAX_BUCKETS counters (32 KiB); every GPU that loads libhsgpu carries it
//     until the process exits; larger bucket counts are refused, not truncated;
//   * all calls on a device share the one array. It still holds the previous call's counts when a new
//     call starts, so each call zeroes the part it is going to use first. Two calls that are in flight
//     on the same GPU at the same time (two contexts, two threads) would count into the same words and
//     corrupt each other: serialise them;
//   * the first call on each device looks up the device address of the array
//     (cudaGetSymbolAddress); make that first call outside of any stream capture;
//   * inputs and outputs are the caller's: key columns and out_bucket_ids must already be in device
//     memory, because staging host data would need an allocation.
#include <mutex>

#include "hs_keys.cuh"

#define NA_THREADS 256  // HS_NOALLOC_MAX_BUCKETS comes from include/hsgpu.h

// The scratch. In host code this name is only a handle for the runtime, never a device address;
// host code gets the address once per device through cudaGetSymbolAddress (see noalloc_scratch()).
static __device__ __align__(256) uint32_t g_noalloc_hist[HS_NOALLOC_MAX_BUCKETS];

__device__ __forceinline__ void na_hist_add(uint32_t* hist, uint32_t b, bool valid) {
  unsigned active = __ballot_sync(0xffffffffu, valid);
  if (!valid) return;
  unsigned peers = __match_any_sync(active, b);
  if ((int)lane_id() == __ffs(peers) - 1) atomicAdd(&hist[b], (uint32_t)__popc(peers));
}

// The K1 kernel (hash -> bucket id -> warp-aggregated shared-memory histogram), with the global
// counters taken from `counters` = the device address of g_noalloc_hist.
__global__ void __launch_bounds__(NA_THREADS) k_bucket_hist_static(KeySpec ks, int64_t n, int32_t nb,
                                                                   int32_t* __restrict__ out_ids,
                                                                   uint32_t* __restrict__ counters) {
  extern __shared__ uint32_t s_hist[];
  for (int b = threadIdx.x; b < nb; b += NA_THREADS) s_hist[b] = 0;
  __syncthreads();
  int64_t stride = (int64_t)gridDim.x * NA_THREADS;
  for (int64_t base = (int64_t)blockIdx.x * NA_THREADS; base < n; base += stride) {
    int64_t i = base + threadIdx.x;
    bool valid = i < n;
    int32_t b = 0;
    if (valid) {
      b = pmod_i32((int32_t)mm_hash_row(ks, i), nb);
      out_ids[i] = b;
    }
    na_hist_add(s_hist, (uint32_t)b, valid);
  }
  __syncthreads();
  for (int b = threadIdx.x; b < nb; b += NA_THREADS) {
    uint32_t v = s_hist[b];
    if (v) atomicAdd(&counters[b], v);
  }
}

// Set-up, once per device, with that device current: ask the runtime where the array lives.
static uint32_t* noalloc_scratch(int device) {
  static std::mutex mu;
  static uint32_t* addr[64] = {nullptr};
  if (device < 0 || device >= 64) HS_THROW(HS_EINVAL, "device index %d out of range", device);
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[device]) {
    uint32_t* p = nullptr;
    CK(cudaGetSymbolAddress((void**)&p, g_noalloc_hist));
    addr[device] = p;
  }
  return addr[device];
}

static void require_device(const void* p, const char* what) {
  if (p && !hs_is_device_ptr(p))
    HS_THROW(HS_EINVAL, "%s must be in device memory: this entry point cannot stage host data without allocating", what);
}

extern "C" int32_t hs_bucket_histogram_noalloc(hs_ctx* ctx, const hs_column* keys, int32_t nkeys,
                                               int32_t num_buckets, int32_t* out_bucket_ids,
                                               uint32_t* out_counts) {
  API_BEGIN(ctx)
  if (!keys || !out_bucket_ids || !out_counts) HS_THROW(HS_EINVAL, "null argument");
  if (nkeys <= 0 || nkeys > HS_MAX_KEYS) HS_THROW(HS_EINVAL, "between 1 and %d key columns are supported", HS_MAX_KEYS);
  if (num_buckets <= 0) HS_THROW(HS_EINVAL, "num_buckets must be positive");
  if (num_buckets > HS_NOALLOC_MAX_BUCKETS)
    HS_THROW(HS_EINVAL, "num_buckets %d is beyond the %d counters of the built-in scratch of "
             "hs_bucket_histogram_noalloc; use hs_hash_bucket or hs_build_index", num_buckets, HS_NOALLOC_MAX_BUCKETS);
  const int64_t n = keys[0].length;
  // the counters, in shared memory, in the scratch array and in out_counts, are 32 bits wide
  if (n >= 0xFFFFFFF0ll)
    HS_THROW(HS_EUNSUPPORTED, "hs_bucket_histogram_noalloc takes fewer than 2^32-16 rows per call (got %lld)", (long long)n);
  KeySpec ks;
  memset(&ks, 0, sizeof ks);
  ks.nkeys = nkeys;
  for (int k = 0; k < nkeys; k++) {
    const hs_column& in = keys[k];
    if (!hs_type_ok(in.type)) HS_THROW(HS_EUNSUPPORTED, "unsupported column type %d", in.type);
    if (in.length != n || n < 0) HS_THROW(HS_EINVAL, "key columns differ in length");
    bool nulls = in.validity && in.null_count != 0;
    if (nulls) require_device(in.validity, "validity bitmap");
    require_device(in.data, "key data");
    if (hs_is_varlen(in.type)) {
      if (!in.offsets) HS_THROW(HS_EINVAL, "UTF8/BINARY column without offsets");
      require_device(in.offsets, "offsets");
    } else if (n > 0 && !in.data) HS_THROW(HS_EINVAL, "column without data buffer");
    ks.k[k].type = in.type;
    ks.k[k].has_nulls = nulls;
    ks.k[k].validity = nulls ? in.validity : nullptr;
    ks.k[k].offsets = in.offsets;
    ks.k[k].data = in.data;
  }
  require_device(out_bucket_ids, "out_bucket_ids");
  uint32_t* counters = noalloc_scratch(ctx->device);
  // the array still holds what the previous call counted: clear the part this call uses
  CK(cudaMemsetAsync(counters, 0, (size_t)num_buckets * 4, ctx->stream));
  if (n > 0) {
    int64_t g = (n + NA_THREADS - 1) / NA_THREADS, cap = (int64_t)HS_NUM_SMS * 8;
    if (g > cap) g = cap;
    LAUNCH(ctx, k_bucket_hist_static, (unsigned)g, NA_THREADS, (size_t)num_buckets * 4, ks, n, num_buckets,
           out_bucket_ids, counters);
  }
  // out_counts may be host or device memory; the copy allocates nothing
  CK(cudaMemcpyAsync(out_counts, counters, (size_t)num_buckets * 4, cudaMemcpyDefault, ctx->stream));
  hs_sync(ctx);
  API_END(ctx)
}
