// hs_common.cuh — context, error plumbing, device buffers, column staging.
// Internal to libhsgpu; the public surface is include/hsgpu.h.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <map>
#include <string>
#include <vector>

#include "../../include/hsgpu.h"

#define HS_MAX_KEYS 8
#define HS_MAX_MOVE 40
// SMs of the current device (148 on B200): grids are sized in multiples of this. Queried once per device.
int hs_sm_count();
#define HS_NUM_SMS hs_sm_count()

struct HsError {
  int32_t status;
  std::string msg;
};

#define HS_THROW(st, ...)                                   \
  do {                                                      \
    char _b[512];                                           \
    snprintf(_b, sizeof _b, __VA_ARGS__);                   \
    throw HsError{(int32_t)(st), std::string(_b)};          \
  } while (0)

#define CK(call)                                                                          \
  do {                                                                                    \
    cudaError_t _e = (call);                                                              \
    if (_e != cudaSuccess)                                                                \
      HS_THROW(_e == cudaErrorMemoryAllocation ? HS_ENOMEM : HS_ECUDA, "%s:%d %s: %s",    \
               __FILE__, __LINE__, #call, cudaGetErrorString(_e));                        \
  } while (0)

struct StageRec {
  char name[32];
  cudaEvent_t e0, e1;
  int launches;
};

struct hs_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  std::string err;
  int out_loc = HS_MEM_HOST;
  bool deferred = false;             // host results are copied on copy_stream after the call returns
  cudaStream_t copy_stream = nullptr;
  bool timing = false;
  std::vector<StageRec> stages;      // of the last call
  std::vector<cudaEvent_t> ev_free;  // recycled events
  int cur_stage = -1;
  int64_t launches = 0;
  std::multimap<size_t, void*> pinned_free;  // cached pinned host blocks by size
  std::map<void*, size_t> pinned_size;
  // large device blocks are recycled inside the context (all work is ordered on one stream, so a block
  // freed by one call can be handed to the next without touching the allocator: cudaMallocAsync of
  // ~GB blocks was measured to cost up to 3 ms per build)
  std::multimap<size_t, void*> dev_free;
  std::map<void*, size_t> dev_size;
  size_t dev_cached_bytes = 0;
  size_t dev_cache_cap = 0;          // bytes the block cache may hold (hs_ctx_create: half of the device memory,
                                     // HSGPU_CACHE_MB overrides); beyond it the oldest blocks go back to the pool
  size_t pinned_cached_bytes = 0, pinned_cache_cap = 0;   // same for pinned result buffers (HSGPU_PINNED_CACHE_MB)
  std::map<void*, uint64_t> cache_age;                    // when a cached block (device or pinned) was freed
  uint64_t cache_clock = 0;
  cudaMemPool_t pool = nullptr;      // the context's own stream-ordered pool: the device's default pool is not touched
  // multi-GPU (hs_dist.cu)
  void* nccl_lib = nullptr;
  void* comm = nullptr;
  void* p2p = nullptr;     // peer-mapped receive buffers of the fused multi-GPU scatter (hs_dist.cu)
  int rank = 0, nranks = 1;
  int64_t last_xchg_bytes = 0;
  float last_xchg_ms = 0.f;
  // small device->host reads that must not stall the queue behind them (hs_read_begin / hs_read_wait)
  // small host->device uploads (lists, offsets): a pageable source makes cudaMemcpyAsync wait for the stream
  // first, which drains the queue at every upload; they go through this pinned ring instead (hs_upload)
  unsigned char* up_ring = nullptr;
  size_t up_at = 0;
  void* rd_pin = nullptr;            // pinned, HS_READ_SLOTS x HS_READ_BYTES
  cudaEvent_t rd_ev[4] = {nullptr, nullptr, nullptr, nullptr};
  int rd_next = 0;
};
#define HS_READ_SLOTS 4
#define HS_READ_BYTES 256

// counts a kernel launch (for gpu_launches in bench.py and the per-stage table)
static inline void hs_count_launch(hs_ctx* c) {
  c->launches++;
  if (c->cur_stage >= 0) c->stages[c->cur_stage].launches++;
}
#define LAUNCH(ctx, kern, grid, block, smem, ...)                       \
  do {                                                                  \
    kern<<<(grid), (block), (smem), (ctx)->stream>>>(__VA_ARGS__);      \
    hs_count_launch(ctx);                                               \
    CK(cudaGetLastError());                                             \
  } while (0)

void hs_stage_begin(hs_ctx* c, const char* name);
void hs_stage_end(hs_ctx* c);
void hs_stages_reset(hs_ctx* c);
struct Stage {
  hs_ctx* c;
  Stage(hs_ctx* c_, const char* n) : c(c_) { hs_stage_begin(c, n); }
  ~Stage() { hs_stage_end(c); }
};

void* hs_dev_alloc(hs_ctx* c, size_t bytes);
void hs_dev_free(hs_ctx* c, void* p);
void hs_dev_cache_flush(hs_ctx* c);

// stream-ordered device buffer
struct DBuf {
  hs_ctx* c = nullptr;
  void* p = nullptr;
  size_t bytes = 0;
  DBuf() {}
  DBuf(hs_ctx* c_, size_t n) { alloc(c_, n); }
  DBuf(const DBuf&) = delete;
  DBuf& operator=(const DBuf&) = delete;
  DBuf(DBuf&& o) noexcept : c(o.c), p(o.p), bytes(o.bytes) { o.p = nullptr; o.bytes = 0; }
  DBuf& operator=(DBuf&& o) noexcept {
    if (this != &o) { release(); c = o.c; p = o.p; bytes = o.bytes; o.p = nullptr; o.bytes = 0; }
    return *this;
  }
  void alloc(hs_ctx* c_, size_t n) {
    release();
    c = c_;
    bytes = n;
    p = hs_dev_alloc(c, n);
  }
  void zero() { if (p) CK(cudaMemsetAsync(p, 0, bytes ? bytes : 16, c->stream)); }
  void release() {
    if (p) hs_dev_free(c, p);
    p = nullptr;
    bytes = 0;
  }
  void* detach() { void* r = p; p = nullptr; bytes = 0; return r; }
  template <typename T> T* as() const { return (T*)p; }
  ~DBuf() { release(); }
};

// A column whose buffers are on the device (borrowed from the caller or staged).
struct DevCol {
  int32_t type = 0;
  int64_t length = 0;
  int64_t null_count = 0;
  const uint8_t* validity = nullptr;
  const int32_t* offsets = nullptr;
  const void* data = nullptr;
  int64_t data_bytes = 0;
  std::vector<DBuf> owned;  // staging copies of host inputs
  DevCol() {}
  DevCol(DevCol&&) = default;
  DevCol& operator=(DevCol&&) = default;
};

// kernel-side view of a column
struct KCol {
  int32_t type;
  int32_t has_nulls;
  const uint8_t* validity;
  const int32_t* offsets;
  const void* data;
};

// an output column under construction (device buffers)
struct OutCol {
  int32_t type = 0;
  int64_t length = 0;
  int64_t null_count = 0;
  DBuf validity, offsets, data;
  int64_t data_bytes = 0;
  OutCol() {}
  OutCol(OutCol&&) = default;
  OutCol& operator=(OutCol&&) = default;
};

static inline int hs_fixed_width(int32_t t) {
  switch (t) {
    case HS_INT8: return 1;
    case HS_INT16: return 2;
    case HS_INT32: case HS_FLOAT32: return 4;
    case HS_INT64: case HS_FLOAT64: return 8;
    default: return 0;
  }
}
__host__ __device__ __forceinline__ bool hs_is_large(int32_t t) { return t == HS_LARGE_UTF8 || t == HS_LARGE_BINARY; }
__host__ __device__ __forceinline__ bool hs_is_varlen(int32_t t) { return t >= HS_UTF8 && t <= HS_LARGE_BINARY; }
static inline bool hs_type_ok(int32_t t) { return t >= HS_BOOL && t <= HS_LARGE_BINARY; }
static inline size_t hs_offset_width(int32_t t) { return hs_is_large(t) ? 8 : 4; }

bool hs_is_device_ptr(const void* p);
void hs_stage_column(hs_ctx* c, const hs_column& in, DevCol& out);
void hs_stage_columns(hs_ctx* c, const hs_column* in, int n, std::vector<DevCol>& out);
KCol hs_kcol(const DevCol& d);
void* hs_pinned_alloc(hs_ctx* c, size_t bytes);
void hs_pinned_free(hs_ctx* c, void* p);
hs_table* hs_finish_table(hs_ctx* c, std::vector<OutCol>& cols, int64_t nrows, int32_t num_buckets,
                          const std::vector<int64_t>* boff);
void hs_sync(hs_ctx* c);
// Copy up to HS_READ_BYTES from the device into a pinned slot and mark the point with an event; work queued
// afterwards keeps running while the host waits for just that copy. Returns the slot.
int hs_read_begin(hs_ctx* c, const void* d_src, size_t bytes);
// Host -> device copy on the context's stream that never waits for the stream: the bytes are staged in a pinned
// ring (HS_UP_RING bytes; a wrap-around waits for the stream once); h_src may be reused at once. Larger copies
// fall back to cudaMemcpyAsync from h_src.
#define HS_UP_RING (4u << 20)
void hs_upload(hs_ctx* c, void* d_dst, const void* h_src, size_t bytes);
const void* hs_read_wait(hs_ctx* c, int slot);

#define API_BEGIN(ctx)                      \
  if (!(ctx)) return HS_EINVAL;             \
  try {                                     \
    CK(cudaSetDevice((ctx)->device));

#define API_END(ctx)                        \
  }                                         \
  catch (const HsError& e) {                \
    (ctx)->err = e.msg;                     \
    cudaStreamSynchronize((ctx)->stream);   \
    cudaGetLastError();                     \
    return e.status;                        \
  }                                         \
  catch (const std::exception& e) {         \
    (ctx)->err = e.what();                  \
    return HS_ENOMEM;                       \
  }                                         \
  return HS_OK;

// ---- shared device helpers -------------------------------------------------
__device__ __forceinline__ bool kvalid(const uint8_t* v, int64_t i) {
  return v == nullptr || ((v[i >> 3] >> (i & 7)) & 1);
}
__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31; }
// start and length of string cell i (int32 or int64 offsets behind the same pointer)
__device__ __forceinline__ int64_t kstr_start(const KCol& c, int64_t i, int32_t* len) {
  if (hs_is_large(c.type)) {
    const int64_t* o = (const int64_t*)c.offsets;
    int64_t s = o[i];
    *len = (int32_t)(o[i + 1] - s);
    return s;
  }
  int32_t s = c.offsets[i];
  *len = c.offsets[i + 1] - s;
  return s;
}

// primitives implemented in hs_scan.cu / hs_gather.cu
// out[i] = sum_{j<i} in[j] (out may alias in); returns nothing, total written to *d_total (device, u64) if non-null
void hs_exclusive_scan_u32(hs_ctx* c, const uint32_t* in, uint32_t* out, int64_t n, uint64_t* d_total);
void hs_exclusive_scan_flags(hs_ctx* c, const uint8_t* flags, uint32_t* out, int64_t n, uint64_t* d_total);
int64_t hs_offsets_from_lens(hs_ctx* c, const uint32_t* lens, int64_t m, int32_t type, OutCol& out);
// gather rows idx[0..m) of `src` into a new column
// may_miss: idx may hold HS_MISS for "no source row" (outer joins): those rows come out NULL
#define HS_MISS 0xFFFFFFFFu
void hs_gather_column(hs_ctx* c, const DevCol& src, const uint32_t* idx, int64_t m, OutCol& out, bool may_miss = false);
void hs_copy_column(hs_ctx* c, const DevCol& src, OutCol& out);
