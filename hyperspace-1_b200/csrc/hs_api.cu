// hs_api.cu — context lifecycle, column staging, result tables (C ABI glue).
#include "hs_common.cuh"
#include <stdlib.h>

struct TableImpl {
  std::vector<void*> dev;     // device buffers owned by the table
  std::vector<void*> pinned;  // pinned host buffers owned by the table
  std::vector<hs_column> cols;
  std::vector<int64_t> boff;
  // deferred results: the device -> host copies run on the context's copy stream after the call has returned;
  // `ready` marks their end, `staged` are the device buffers they read (released once the copies are done)
  cudaEvent_t ready = nullptr;
  std::vector<void*> staged;
};

static void table_settle(hs_ctx* c, TableImpl* ti) {
  if (!ti || !ti->ready) return;
  cudaError_t e = cudaEventSynchronize(ti->ready);
  cudaEventDestroy(ti->ready);
  ti->ready = nullptr;
  for (void* p : ti->staged) hs_dev_free(c, p);
  ti->staged.clear();
  if (e != cudaSuccess) HS_THROW(HS_ECUDA, "copy of a deferred result failed: %s", cudaGetErrorString(e));
}

void hs_sync(hs_ctx* c) { CK(cudaStreamSynchronize(c->stream)); }

int hs_sm_count() {
  static int n[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  int& v = n[dev & 63];
  if (!v && cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) v = 148;
  return v;
}

void hs_upload(hs_ctx* c, void* d_dst, const void* h_src, size_t bytes) {
  if (!bytes) return;
  if (bytes > HS_UP_RING / 4) {
    CK(cudaMemcpyAsync(d_dst, h_src, bytes, cudaMemcpyHostToDevice, c->stream));
    return;
  }
  if (!c->up_ring) CK(cudaMallocHost((void**)&c->up_ring, HS_UP_RING));
  const size_t need = (bytes + 63) & ~(size_t)63;
  if (c->up_at + need > HS_UP_RING) {
    CK(cudaStreamSynchronize(c->stream));   // everything staged so far has been copied
    c->up_at = 0;
  }
  unsigned char* slot = c->up_ring + c->up_at;
  c->up_at += need;
  memcpy(slot, h_src, bytes);
  CK(cudaMemcpyAsync(d_dst, slot, bytes, cudaMemcpyHostToDevice, c->stream));
}

int hs_read_begin(hs_ctx* c, const void* d_src, size_t bytes) {
  if (bytes > HS_READ_BYTES) HS_THROW(HS_EINVAL, "hs_read_begin: %zu bytes", bytes);
  if (!c->rd_pin) {
    CK(cudaMallocHost(&c->rd_pin, HS_READ_SLOTS * HS_READ_BYTES));
    for (int i = 0; i < HS_READ_SLOTS; i++) CK(cudaEventCreateWithFlags(&c->rd_ev[i], cudaEventDisableTiming));
  }
  int slot = c->rd_next;
  c->rd_next = (slot + 1) % HS_READ_SLOTS;
  CK(cudaMemcpyAsync((char*)c->rd_pin + slot * HS_READ_BYTES, d_src, bytes, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaEventRecord(c->rd_ev[slot], c->stream));
  return slot;
}
const void* hs_read_wait(hs_ctx* c, int slot) {
  CK(cudaEventSynchronize(c->rd_ev[slot]));
  return (const char*)c->rd_pin + slot * HS_READ_BYTES;
}

static cudaEvent_t get_event(hs_ctx* c) {
  if (!c->ev_free.empty()) {
    cudaEvent_t e = c->ev_free.back();
    c->ev_free.pop_back();
    return e;
  }
  cudaEvent_t e;
  CK(cudaEventCreate(&e));
  return e;
}

void hs_stages_reset(hs_ctx* c) {
  for (auto& s : c->stages) {
    c->ev_free.push_back(s.e0);
    c->ev_free.push_back(s.e1);
  }
  c->stages.clear();
  c->cur_stage = -1;
}

#include <chrono>
static double host_now_ms() {
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}
void hs_stage_begin(hs_ctx* c, const char* name) {
  static int trace = getenv("HSGPU_TRACE") ? 1 : 0;
  if (trace) fprintf(stderr, "[hsgpu r%d] %10.3f ms  %s\n", c->rank, host_now_ms(), name);
  if (!c->timing) return;
  StageRec r;
  memset(&r, 0, sizeof r);
  strncpy(r.name, name, sizeof(r.name) - 1);
  r.e0 = get_event(c);
  r.e1 = get_event(c);
  r.launches = 0;
  CK(cudaEventRecord(r.e0, c->stream));
  c->stages.push_back(r);
  c->cur_stage = (int)c->stages.size() - 1;
}

void hs_stage_end(hs_ctx* c) {
  static int trace = getenv("HSGPU_TRACE") ? 1 : 0;
  if (trace) fprintf(stderr, "[hsgpu r%d] %10.3f ms    end\n", c->rank, host_now_ms());
  if (!c->timing || c->cur_stage < 0) return;
  cudaEventRecord(c->stages[c->cur_stage].e1, c->stream);
  c->cur_stage = -1;
}

bool hs_is_device_ptr(const void* p) {
  if (!p) return false;
  cudaPointerAttributes a;
  cudaError_t e = cudaPointerGetAttributes(&a, p);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

static const void* to_device(hs_ctx* c, const void* p, size_t bytes, DevCol& out) {
  if (!p || bytes == 0) return p && hs_is_device_ptr(p) ? p : nullptr;
  if (hs_is_device_ptr(p)) return p;
  out.owned.emplace_back(c, bytes);
  CK(cudaMemcpyAsync(out.owned.back().p, p, bytes, cudaMemcpyHostToDevice, c->stream));
  return out.owned.back().p;
}

void hs_stage_column(hs_ctx* c, const hs_column& in, DevCol& out) {
  if (!hs_type_ok(in.type)) HS_THROW(HS_EUNSUPPORTED, "unsupported column type %d", in.type);
  if (in.length < 0) HS_THROW(HS_EINVAL, "negative column length");
  if (in.length >= (int64_t)0xFFFFFFF0ll) HS_THROW(HS_EUNSUPPORTED, "more than 2^32-16 rows in one call");
  out.type = in.type;
  out.length = in.length;
  out.null_count = in.validity ? in.null_count : 0;
  int64_t n = in.length;
  out.validity = nullptr;
  if (in.validity && in.null_count != 0)
    out.validity = (const uint8_t*)to_device(c, in.validity, (size_t)((n + 7) / 8), out);
  if (hs_is_varlen(in.type)) {
    if (!in.offsets) HS_THROW(HS_EINVAL, "UTF8/BINARY column without offsets");
    const size_t ow = hs_offset_width(in.type);
    int64_t last = 0, first = 0;
    const char* op = (const char*)in.offsets;
    auto rd = [&](const void* p) -> int64_t {
      if (ow == 8) { int64_t v; memcpy(&v, p, 8); return v; }
      int32_t v; memcpy(&v, p, 4); return v;
    };
    if (hs_is_device_ptr(in.offsets)) {
      char tmp[16];
      CK(cudaMemcpyAsync(tmp, op + (size_t)n * ow, ow, cudaMemcpyDeviceToHost, c->stream));
      CK(cudaMemcpyAsync(tmp + 8, op, ow, cudaMemcpyDeviceToHost, c->stream));
      hs_sync(c);
      last = rd(tmp);
      first = rd(tmp + 8);
    } else {
      last = rd(op + (size_t)n * ow);
      first = rd(op);
    }
    if (first != 0) HS_THROW(HS_EINVAL, "offsets[0] must be 0 (pass an unsliced array)");
    if (last < 0) HS_THROW(HS_EINVAL, "negative string offsets");
    out.offsets = (const int32_t*)to_device(c, in.offsets, (size_t)(n + 1) * ow, out);
    out.data_bytes = last;
    out.data = to_device(c, in.data, (size_t)last, out);
  } else if (in.type == HS_BOOL) {
    out.data_bytes = (n + 7) / 8;
    out.data = to_device(c, in.data, (size_t)out.data_bytes, out);
  } else {
    out.data_bytes = n * hs_fixed_width(in.type);
    out.data = to_device(c, in.data, (size_t)out.data_bytes, out);
  }
  if (n > 0 && !hs_is_varlen(in.type) && !out.data) HS_THROW(HS_EINVAL, "column without data buffer");
}

void hs_stage_columns(hs_ctx* c, const hs_column* in, int n, std::vector<DevCol>& out) {
  out.resize(n);
  for (int i = 0; i < n; i++) {
    hs_stage_column(c, in[i], out[i]);
    if (in[i].length != in[0].length) HS_THROW(HS_EINVAL, "columns differ in length");
  }
}

KCol hs_kcol(const DevCol& d) {
  KCol k;
  k.type = d.type;
  k.has_nulls = d.validity != nullptr;
  k.validity = d.validity;
  k.offsets = d.offsets;
  k.data = d.data;
  return k;
}

#define HS_DEV_CACHE_MIN (size_t(1) << 20)

void hs_dev_cache_flush(hs_ctx* c) {
  for (auto& kv : c->dev_free) cudaFreeAsync(kv.second, c->stream);
  for (auto& kv : c->dev_free) { c->dev_size.erase(kv.second); c->cache_age.erase(kv.second); }
  c->dev_free.clear();
  c->dev_cached_bytes = 0;
}

static cudaError_t pool_alloc(hs_ctx* c, void** p, size_t bytes) {
  return c->pool ? cudaMallocFromPoolAsync(p, bytes, c->pool, c->stream) : cudaMallocAsync(p, bytes, c->stream);
}

void* hs_dev_alloc(hs_ctx* c, size_t bytes) {
  if (bytes == 0) bytes = 16;
  if (bytes >= HS_DEV_CACHE_MIN) {
    auto it = c->dev_free.lower_bound(bytes);
    if (it != c->dev_free.end() && it->first <= bytes + bytes / 4) {
      void* p = it->second;
      c->dev_cached_bytes -= it->first;
      c->cache_age.erase(p);
      c->dev_free.erase(it);
      return p;
    }
  }
  void* p = nullptr;
  cudaError_t e = pool_alloc(c, &p, bytes);
  if (e == cudaErrorMemoryAllocation && !c->dev_free.empty()) {
    cudaGetLastError();
    hs_dev_cache_flush(c);
    cudaStreamSynchronize(c->stream);
    e = pool_alloc(c, &p, bytes);
  }
  if (e != cudaSuccess) {
    cudaGetLastError();
    HS_THROW(e == cudaErrorMemoryAllocation ? HS_ENOMEM : HS_ECUDA, "cudaMallocAsync(%zu bytes): %s", bytes,
             cudaGetErrorString(e));
  }
  if (bytes >= HS_DEV_CACHE_MIN) c->dev_size[p] = bytes;
  return p;
}

void hs_dev_free(hs_ctx* c, void* p) {
  if (!p) return;
  auto it = c->dev_size.find(p);
  if (it == c->dev_size.end()) {
    cudaFreeAsync(p, c->stream);
    return;
  }
  c->dev_free.insert({it->second, p});
  c->dev_cached_bytes += it->second;
  c->cache_age[p] = ++c->cache_clock;
  // over the cap: the blocks freed longest ago go back to the pool (the cache holds a few dozen blocks at most)
  while (c->dev_cached_bytes > c->dev_cache_cap && !c->dev_free.empty()) {
    auto old = c->dev_free.begin();
    for (auto j = c->dev_free.begin(); j != c->dev_free.end(); ++j)
      if (c->cache_age[j->second] < c->cache_age[old->second]) old = j;
    cudaFreeAsync(old->second, c->stream);
    c->dev_cached_bytes -= old->first;
    c->dev_size.erase(old->second);
    c->cache_age.erase(old->second);
    c->dev_free.erase(old);
  }
}

void* hs_pinned_alloc(hs_ctx* c, size_t bytes) {
  if (bytes == 0) bytes = 16;
  size_t cap = 256;
  while (cap < bytes) cap <<= 1;
  if (bytes > (size_t(1) << 26)) cap = (bytes + (size_t(1) << 24) - 1) & ~((size_t(1) << 24) - 1);
  auto it = c->pinned_free.lower_bound(cap);
  if (it != c->pinned_free.end() && it->first <= cap * 2) {
    void* p = it->second;
    c->pinned_cached_bytes -= it->first;
    c->cache_age.erase(p);
    c->pinned_free.erase(it);
    return p;
  }
  void* p = nullptr;
  CK(cudaMallocHost(&p, cap));
  c->pinned_size[p] = cap;
  return p;
}

void hs_pinned_free(hs_ctx* c, void* p) {
  if (!p) return;
  auto it = c->pinned_size.find(p);
  if (it == c->pinned_size.end()) return;
  c->pinned_free.insert({it->second, p});
  c->pinned_cached_bytes += it->second;
  c->cache_age[p] = ++c->cache_clock;
  while (c->pinned_cached_bytes > c->pinned_cache_cap && !c->pinned_free.empty()) {
    auto old = c->pinned_free.begin();
    for (auto j = c->pinned_free.begin(); j != c->pinned_free.end(); ++j)
      if (c->cache_age[j->second] < c->cache_age[old->second]) old = j;
    cudaFreeHost(old->second);
    c->pinned_cached_bytes -= old->first;
    c->pinned_size.erase(old->second);
    c->cache_age.erase(old->second);
    c->pinned_free.erase(old);
  }
}

hs_table* hs_finish_table(hs_ctx* c, std::vector<OutCol>& cols, int64_t nrows, int32_t num_buckets,
                          const std::vector<int64_t>* boff) {
  TableImpl* ti = new TableImpl();
  hs_table* t = new hs_table();
  memset(t, 0, sizeof *t);
  t->impl = ti;
  try {
    ti->cols.resize(cols.size());
    bool host = c->out_loc == HS_MEM_HOST;
    const bool deferred = host && c->deferred;
    cudaStream_t cs = c->stream;
    if (deferred) {
      // the copies wait for everything queued on the compute stream so far and then run beside whatever the
      // caller starts next (the upload of the next build's input: PCIe carries both directions at once)
      if (!c->copy_stream) CK(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
      cudaEvent_t done;
      CK(cudaEventCreateWithFlags(&done, cudaEventDisableTiming));
      CK(cudaEventRecord(done, c->stream));
      CK(cudaStreamWaitEvent(c->copy_stream, done, 0));
      CK(cudaEventDestroy(done));
      cs = c->copy_stream;
    }
    auto place = [&](DBuf& b, size_t bytes) -> void* {
      if (!b.p) return nullptr;
      if (!host) {
        void* p = b.detach();
        ti->dev.push_back(p);
        return p;
      }
      void* h = hs_pinned_alloc(c, bytes);
      ti->pinned.push_back(h);
      if (bytes) CK(cudaMemcpyAsync(h, b.p, bytes, cudaMemcpyDeviceToHost, cs));
      if (deferred) ti->staged.push_back(b.detach());
      return h;
    };
    for (size_t i = 0; i < cols.size(); i++) {
      OutCol& o = cols[i];
      hs_column& hc = ti->cols[i];
      memset(&hc, 0, sizeof hc);
      hc.type = o.type;
      hc.length = o.length;
      hc.null_count = o.null_count;
      hc.validity = (const uint8_t*)place(o.validity, (size_t)((o.length + 7) / 8));
      hc.offsets = (const int32_t*)place(o.offsets, (size_t)(o.length + 1) * hs_offset_width(o.type));
      hc.data = place(o.data, (size_t)o.data_bytes);
      if (!hc.validity) hc.null_count = 0;
    }
    if (deferred) {
      CK(cudaEventCreateWithFlags(&ti->ready, cudaEventDisableTiming));
      CK(cudaEventRecord(ti->ready, cs));
    } else {
      hs_sync(c);
    }
    cols.clear();  // frees the device copies in host mode
    t->ncols = (int32_t)ti->cols.size();
    t->location = host ? HS_MEM_HOST : HS_MEM_DEVICE;
    t->nrows = nrows;
    t->cols = ti->cols.data();
    if (boff) {
      ti->boff = *boff;
      t->num_buckets = num_buckets;
      t->bucket_offsets = ti->boff.data();
    }
  } catch (...) {
    if (c->copy_stream) cudaStreamSynchronize(c->copy_stream);
    if (ti->ready) cudaEventDestroy(ti->ready);
    for (void* p : ti->staged) hs_dev_free(c, p);
    for (void* p : ti->dev) hs_dev_free(c, p);
    for (void* p : ti->pinned) hs_pinned_free(c, p);
    delete ti;
    delete t;
    throw;
  }
  return t;
}

extern "C" {

const char* hs_version(void) { return "hsgpu 0.1 (sm_100a)"; }

int32_t hs_ctx_create(int32_t device, hs_ctx** out) {
  if (!out) return HS_EINVAL;
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    return HS_ECUDA;  // no CPU fallback: without a CUDA device there is no context
  }
  if (device < 0 || device >= ndev) return HS_EINVAL;
  hs_ctx* c = new hs_ctx();
  c->device = device;
  if (cudaSetDevice(device) != cudaSuccess ||
      cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) {
    cudaGetLastError();
    delete c;
    return HS_ECUDA;
  }
  // a stream-ordered pool of the context's own, which keeps freed blocks (repeated builds reuse them without
  // cudaMalloc); the device's default pool, which belongs to the host application, is left as it is
  {
    cudaMemPoolProps props;
    memset(&props, 0, sizeof props);
    props.allocType = cudaMemAllocationTypePinned;
    props.handleTypes = cudaMemHandleTypeNone;
    props.location.type = cudaMemLocationTypeDevice;
    props.location.id = device;
    if (cudaMemPoolCreate(&c->pool, &props) == cudaSuccess) {
      uint64_t thr = UINT64_MAX;
      cudaMemPoolSetAttribute(c->pool, cudaMemPoolAttrReleaseThreshold, &thr);
    } else {
      cudaGetLastError();
      c->pool = nullptr;   // allocations fall back to the default pool, untouched
    }
  }
  // bounds of the two block caches (oldest blocks are released beyond them)
  size_t free_b = 0, total_b = 0;
  if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) { cudaGetLastError(); total_b = size_t(16) << 30; }
  c->dev_cache_cap = total_b / 2;
  c->pinned_cache_cap = size_t(16) << 30;
  if (const char* e = getenv("HSGPU_CACHE_MB")) c->dev_cache_cap = (size_t)atoll(e) << 20;
  if (const char* e = getenv("HSGPU_PINNED_CACHE_MB")) c->pinned_cache_cap = (size_t)atoll(e) << 20;
  *out = c;
  return HS_OK;
}

int32_t hs_comm_destroy(hs_ctx* ctx);

int32_t hs_ctx_destroy(hs_ctx* c) {
  if (!c) return HS_EINVAL;
  cudaSetDevice(c->device);
  hs_comm_destroy(c);
  hs_dev_cache_flush(c);
  cudaStreamSynchronize(c->stream);
  for (auto& s : c->stages) { cudaEventDestroy(s.e0); cudaEventDestroy(s.e1); }
  for (auto e : c->ev_free) cudaEventDestroy(e);
  for (auto& kv : c->pinned_size) cudaFreeHost(kv.first);
  if (c->up_ring) cudaFreeHost(c->up_ring);
  if (c->rd_pin) {
    cudaFreeHost(c->rd_pin);
    for (int i = 0; i < HS_READ_SLOTS; i++) cudaEventDestroy(c->rd_ev[i]);
  }
  if (c->copy_stream) { cudaStreamSynchronize(c->copy_stream); cudaStreamDestroy(c->copy_stream); }
  cudaStreamDestroy(c->stream);
  if (c->pool) cudaMemPoolDestroy(c->pool);
  delete c;
  return HS_OK;
}

const char* hs_last_error(hs_ctx* c) { return c ? c->err.c_str() : "null context"; }

int32_t hs_ctx_set_output_location(hs_ctx* c, int32_t loc) {
  if (!c || (loc != HS_MEM_HOST && loc != HS_MEM_DEVICE)) return HS_EINVAL;
  c->out_loc = loc;
  return HS_OK;
}

int32_t hs_ctx_get_stream(hs_ctx* c, void** s) {
  if (!c || !s) return HS_EINVAL;
  *s = (void*)c->stream;
  return HS_OK;
}

int32_t hs_ctx_enable_timing(hs_ctx* c, int32_t on) {
  if (!c) return HS_EINVAL;
  c->timing = on != 0;
  return HS_OK;
}

int32_t hs_ctx_last_timings(hs_ctx* c, hs_stage_timing* out, int32_t cap, int32_t* n) {
  if (!c || !n) return HS_EINVAL;
  API_BEGIN(c)
  hs_sync(c);
  int m = 0;
  for (auto& s : c->stages) {
    if (m >= cap || !out) break;
    memset(&out[m], 0, sizeof out[m]);
    strncpy(out[m].name, s.name, sizeof(out[m].name) - 1);
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, s.e0, s.e1) != cudaSuccess) { cudaGetLastError(); ms = -1.f; }
    out[m].ms = ms;
    out[m].launches = s.launches;
    m++;
  }
  *n = out ? m : (int32_t)c->stages.size();
  API_END(c)
}

int32_t hs_ctx_launch_count(hs_ctx* c, int64_t* out) {
  if (!c || !out) return HS_EINVAL;
  *out = c->launches;
  return HS_OK;
}

int32_t hs_table_free(hs_ctx* c, hs_table* t) {
  if (!c) return HS_EINVAL;
  if (!t) return HS_OK;
  cudaSetDevice(c->device);
  TableImpl* ti = (TableImpl*)t->impl;
  if (ti) {
    try { table_settle(c, ti); } catch (const HsError&) {}
    for (void* p : ti->dev) hs_dev_free(c, p);
    for (void* p : ti->pinned) hs_pinned_free(c, p);
    delete ti;
  }
  delete t;
  return HS_OK;
}

int32_t hs_ctx_trim(hs_ctx* c) {
  API_BEGIN(c)
  hs_sync(c);
  hs_dev_cache_flush(c);
  for (auto& kv : c->pinned_free) { cudaFreeHost(kv.second); c->pinned_size.erase(kv.second); c->cache_age.erase(kv.second); }
  c->pinned_free.clear();
  c->pinned_cached_bytes = 0;
  hs_sync(c);
  if (c->pool) CK(cudaMemPoolTrimTo(c->pool, 0));
  API_END(c)
}

int32_t hs_ctx_set_deferred_results(hs_ctx* c, int32_t on) {
  if (!c) return HS_EINVAL;
  c->deferred = on != 0;
  return HS_OK;
}

int32_t hs_table_wait(hs_ctx* c, hs_table* t) {
  API_BEGIN(c)
  if (!t) HS_THROW(HS_EINVAL, "null table");
  table_settle(c, (TableImpl*)t->impl);
  API_END(c)
}

}  // extern "C"
