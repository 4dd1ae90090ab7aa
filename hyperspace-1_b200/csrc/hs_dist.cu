// hs_dist.cu — multi-GPU index build: one process per GPU, rows travel once to the GPU that owns
// their bucket through an NCCL all-to-all over NVLink/NVSwitch; everything after is bucket-local.
// Replaces Spark's ShuffleExchangeExec under df.repartition(numBuckets, cols)
// (HS/actions/CreateActionBase.scala:129-130) for N > 1 executors.
// NCCL is bound at run time (dlopen of libnccl.so.2) so that libhsgpu loads on hosts without it.
#include <dlfcn.h>

#include "hs_keys.cuh"
#include "hs_kernels.h"

typedef struct { char internal[128]; } nccl_uid;
typedef void* nccl_comm;
typedef int (*fn_get_uid)(nccl_uid*);
typedef int (*fn_comm_init)(nccl_comm*, int, nccl_uid, int);
typedef int (*fn_comm_destroy)(nccl_comm);
typedef int (*fn_group)(void);
typedef int (*fn_sendrecv)(void*, size_t, int, int, nccl_comm, cudaStream_t);
typedef int (*fn_allgather)(const void*, void*, size_t, int, nccl_comm, cudaStream_t);
typedef const char* (*fn_errstr)(int);

struct NcclApi {
  void* lib = nullptr;
  fn_get_uid get_uid = nullptr;
  fn_comm_init comm_init = nullptr;
  fn_comm_destroy comm_destroy = nullptr;
  fn_group group_start = nullptr, group_end = nullptr;
  fn_sendrecv send = nullptr, recv = nullptr;
  fn_allgather allgather = nullptr;
  fn_errstr errstr = nullptr;
};
static NcclApi g_nccl;

static void nccl_load() {
  if (g_nccl.lib) return;
  const char* env = getenv("HSGPU_NCCL_LIB");
  const char* names[] = {env, "libnccl.so.2", "libnccl.so"};
  void* lib = nullptr;
  for (const char* nm : names) {
    if (!nm || !*nm) continue;
    lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
    if (lib) break;
  }
  if (!lib) HS_THROW(HS_ENCCL, "cannot load libnccl.so.2 (set HSGPU_NCCL_LIB): %s", dlerror());
  NcclApi a;
  a.lib = lib;
  a.get_uid = (fn_get_uid)dlsym(lib, "ncclGetUniqueId");
  a.comm_init = (fn_comm_init)dlsym(lib, "ncclCommInitRank");
  a.comm_destroy = (fn_comm_destroy)dlsym(lib, "ncclCommDestroy");
  a.group_start = (fn_group)dlsym(lib, "ncclGroupStart");
  a.group_end = (fn_group)dlsym(lib, "ncclGroupEnd");
  a.send = (fn_sendrecv)dlsym(lib, "ncclSend");
  a.recv = (fn_sendrecv)dlsym(lib, "ncclRecv");
  a.allgather = (fn_allgather)dlsym(lib, "ncclAllGather");
  a.errstr = (fn_errstr)dlsym(lib, "ncclGetErrorString");
  if (!a.get_uid || !a.comm_init || !a.comm_destroy || !a.group_start || !a.group_end || !a.send || !a.recv ||
      !a.allgather)
    HS_THROW(HS_ENCCL, "libnccl lacks a required symbol");
  g_nccl = a;
}
#define NCK(call)                                                                        \
  do {                                                                                   \
    int _r = (call);                                                                     \
    if (_r != 0)                                                                         \
      HS_THROW(HS_ENCCL, "%s: %s", #call, g_nccl.errstr ? g_nccl.errstr(_r) : "nccl error"); \
  } while (0)

#define NCCL_UINT8 1
#define NCCL_UINT64 5

// owner(b) = b * N / nb: contiguous bucket ranges, so every GPU writes a contiguous file-id range
static inline int owner_of(int b, int nranks, int nb) { return (int)((int64_t)b * nranks / nb); }

__global__ void __launch_bounds__(256) k_bucket_to_owner(const uint16_t* __restrict__ bucket, int64_t n, int nranks,
                                                         int nb, uint16_t* __restrict__ owner) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) owner[i] = (uint16_t)(((int64_t)bucket[i] * nranks) / nb);
}
__global__ void __launch_bounds__(256) k_bits_to_bytes_d(const uint8_t* __restrict__ bits, int64_t n,
                                                         uint8_t* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) out[i] = (bits[i >> 3] >> (i & 7)) & 1;
}
__global__ void __launch_bounds__(256) k_bytes_to_bits_d(const uint8_t* __restrict__ in, int64_t n,
                                                         uint8_t* __restrict__ bits, unsigned long long* zeros) {
  int64_t nbytes = (n + 7) >> 3;
  int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  unsigned z = 0;
  for (; b < nbytes; b += stride) {
    unsigned v = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
      int64_t i = b * 8 + k;
      if (i < n) { if (in[i]) v |= 1u << k; else z++; }
    }
    bits[b] = (uint8_t)v;
  }
  if (z && zeros) atomicAdd(zeros, (unsigned long long)z);
}

static unsigned dgrid(int64_t n) {
  int64_t g = (n + 255) / 256, cap = (int64_t)HS_NUM_SMS * 16;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (unsigned)g;
}

extern "C" {

int32_t hs_comm_unique_id(hs_ctx* ctx, void* out_id128) {
  API_BEGIN(ctx)
  if (!out_id128) HS_THROW(HS_EINVAL, "null id buffer");
  nccl_load();
  nccl_uid id;
  NCK(g_nccl.get_uid(&id));
  memcpy(out_id128, &id, sizeof id);
  API_END(ctx)
}

int32_t hs_comm_init(hs_ctx* ctx, int32_t rank, int32_t nranks, const void* id128) {
  API_BEGIN(ctx)
  if (nranks < 1 || rank < 0 || rank >= nranks) HS_THROW(HS_EINVAL, "bad rank/nranks");
  if (nranks > 1 && !id128) HS_THROW(HS_EINVAL, "null unique id");
  if (ctx->comm) HS_THROW(HS_EINVAL, "communicator already initialised");
  ctx->rank = rank;
  ctx->nranks = nranks;
  if (nranks > 1) {
    nccl_load();
    nccl_uid id;
    memcpy(&id, id128, sizeof id);
    nccl_comm comm = nullptr;
    NCK(g_nccl.comm_init(&comm, nranks, id, rank));
    ctx->comm = comm;
  }
  API_END(ctx)
}

int32_t hs_comm_destroy(hs_ctx* ctx) {
  if (!ctx) return HS_EINVAL;
  if (ctx->comm && g_nccl.comm_destroy) {
    cudaStreamSynchronize(ctx->stream);
    g_nccl.comm_destroy((nccl_comm)ctx->comm);
  }
  ctx->comm = nullptr;
  ctx->rank = 0;
  ctx->nranks = 1;
  return HS_OK;
}

int32_t hs_comm_last_exchange(hs_ctx* ctx, int64_t* bytes_sent, float* ms) {
  if (!ctx) return HS_EINVAL;
  if (bytes_sent) *bytes_sent = ctx->last_xchg_bytes;
  if (ms) *ms = ctx->last_xchg_ms;
  return HS_OK;
}

int32_t hs_build_index_dist(hs_ctx* ctx, const hs_column* cols, int32_t ncols, int32_t n_indexed,
                            int32_t num_buckets, hs_table** out) {
  API_BEGIN(ctx)
  if (!out) HS_THROW(HS_EINVAL, "null out");
  *out = nullptr;
  if (!cols || ncols <= 0 || n_indexed <= 0 || n_indexed > ncols || num_buckets <= 0)
    HS_THROW(HS_EINVAL, "bad arguments");
  const int N = ctx->nranks, me = ctx->rank, nb = num_buckets;
  if (N > 1 && !ctx->comm) HS_THROW(HS_EINVAL, "hs_comm_init has not been called");
  if (nb < N) HS_THROW(HS_EINVAL, "num_buckets must be >= number of ranks");
  hs_stages_reset(ctx);
  ctx->last_xchg_bytes = 0;
  ctx->last_xchg_ms = 0.f;
  std::vector<DevCol> dcols;
  {
    Stage s(ctx, "h2d");
    hs_stage_columns(ctx, cols, ncols, dcols);
  }
  // bucket range of this rank
  int blo = 0, bhi = nb;
  if (N > 1) {
    blo = nb; bhi = 0;
    for (int b = 0; b < nb; b++)
      if (owner_of(b, N, nb) == me) { if (b < blo) blo = b; if (b + 1 > bhi) bhi = b + 1; }
  }
  std::vector<OutCol> ocols;
  std::vector<int64_t> boff;
  if (N == 1) {
    hs_build_core(ctx, dcols, n_indexed, nb, 0, nb, ocols, boff);
    Stage s(ctx, "d2h");
    *out = hs_finish_table(ctx, ocols, dcols[0].length, nb, &boff);
    return HS_OK;
  }
  for (auto& dc : dcols)
    if (hs_is_varlen(dc.type)) HS_THROW(HS_EUNSUPPORTED, "string columns in the multi-GPU build");
  const int64_t n = dcols[0].length;

  // ---- local hash -> owner of every row, rows grouped by owner
  KeySpec ks = hs_make_keyspec(dcols, n_indexed);
  DBuf bucket(ctx, (size_t)n * 2 + 16), hist(ctx, (size_t)nb * 4), minmax(ctx, (2 * HS_MAX_KEYS + 1) * 8);
  std::vector<uint32_t> h_hist((size_t)nb);
  {
    Stage s(ctx, "hash_hist");
    hs_hash_bucket_hist(ctx, ks, dcols, n, nb, 0, nb, bucket.as<uint16_t>(), hist.as<uint32_t>(), minmax.as<uint64_t>());
    CK(cudaMemcpyAsync(h_hist.data(), hist.p, (size_t)nb * 4, cudaMemcpyDeviceToHost, ctx->stream));
    hs_sync(ctx);
  }
  std::vector<uint64_t> send_cnt((size_t)N, 0);
  for (int b = 0; b < nb; b++) send_cnt[owner_of(b, N, nb)] += h_hist[b];
  std::vector<uint32_t> send_off((size_t)N + 1, 0);
  for (int r = 0; r < N; r++) send_off[r + 1] = send_off[r] + (uint32_t)send_cnt[r];

  // travelling arrays: values, validity bytes, boolean bytes
  struct Trav { int col; int role; int esize; const void* src; DBuf tmp, sendbuf, recvbuf; };
  std::vector<Trav> tv;
  tv.reserve(2 * (size_t)ncols);
  {
    Stage s(ctx, "prepare");
    for (int i = 0; i < ncols; i++) {
      DevCol& dc = dcols[i];
      if (dc.validity) {
        tv.emplace_back();
        Trav& t = tv.back();
        t.col = i; t.role = 1; t.esize = 1;
        t.tmp.alloc(ctx, (size_t)(n ? n : 1));
        if (n) LAUNCH(ctx, k_bits_to_bytes_d, dgrid(n), 256, 0, dc.validity, n, t.tmp.as<uint8_t>());
        t.src = t.tmp.p;
      }
      tv.emplace_back();
      Trav& t = tv.back();
      t.col = i;
      if (dc.type == HS_BOOL) {
        t.role = 2; t.esize = 1;
        t.tmp.alloc(ctx, (size_t)(n ? n : 1));
        if (n) LAUNCH(ctx, k_bits_to_bytes_d, dgrid(n), 256, 0, (const uint8_t*)dc.data, n, t.tmp.as<uint8_t>());
        t.src = t.tmp.p;
      } else {
        t.role = 0; t.esize = hs_fixed_width(dc.type); t.src = dc.data;
      }
    }
    if ((int)tv.size() > HS_MAX_MOVE) HS_THROW(HS_EUNSUPPORTED, "too many columns");
  }
  MoveSet ms;
  memset(&ms, 0, sizeof ms);
  ms.n = (int)tv.size();
  std::vector<const void*> srcs(tv.size());
  for (size_t i = 0; i < tv.size(); i++) {
    tv[i].sendbuf.alloc(ctx, (size_t)n * tv[i].esize);
    ms.esize[i] = tv[i].esize;
    ms.a[i] = tv[i].sendbuf.p;
    ms.b[i] = nullptr;
    srcs[i] = tv[i].src;
  }
  {
    Stage s(ctx, "owner_scatter");
    DBuf owner(ctx, (size_t)n * 2 + 16), d_soff(ctx, ((size_t)N + 1) * 4);
    if (n) LAUNCH(ctx, k_bucket_to_owner, dgrid(n), 256, 0, bucket.as<uint16_t>(), n, N, nb, owner.as<uint16_t>());
    CK(cudaMemcpyAsync(d_soff.p, send_off.data(), ((size_t)N + 1) * 4, cudaMemcpyHostToDevice, ctx->stream));
    hs_bucket_scatter(ctx, srcs.data(), ms, owner.as<uint16_t>(), n, N, d_soff.as<uint32_t>());
    hs_sync(ctx);
  }
  bucket.release();
  for (auto& t : tv) t.tmp.release();

  // ---- counts, then the all-to-all
  std::vector<uint64_t> all_cnt((size_t)N * N, 0);
  {
    Stage s(ctx, "count_exchange");
    DBuf d_send(ctx, (size_t)N * 8), d_all(ctx, (size_t)N * N * 8);
    CK(cudaMemcpyAsync(d_send.p, send_cnt.data(), (size_t)N * 8, cudaMemcpyHostToDevice, ctx->stream));
    NCK(g_nccl.allgather(d_send.p, d_all.p, (size_t)N, NCCL_UINT64, (nccl_comm)ctx->comm, ctx->stream));
    CK(cudaMemcpyAsync(all_cnt.data(), d_all.p, (size_t)N * N * 8, cudaMemcpyDeviceToHost, ctx->stream));
    hs_sync(ctx);
  }
  std::vector<uint64_t> recv_off((size_t)N + 1, 0);
  for (int r = 0; r < N; r++) recv_off[r + 1] = recv_off[r] + all_cnt[(size_t)r * N + me];
  const int64_t nrecv = (int64_t)recv_off[N];
  if (nrecv >= (int64_t)0xFFFFFFF0ll) HS_THROW(HS_EUNSUPPORTED, "more than 2^32-16 rows on one GPU");
  int64_t bytes_sent = 0;
  {
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    Stage s(ctx, "all_to_all");
    for (auto& t : tv) t.recvbuf.alloc(ctx, (size_t)nrecv * t.esize);
    CK(cudaEventRecord(e0, ctx->stream));
    NCK(g_nccl.group_start());
    for (auto& t : tv) {
      for (int r = 0; r < N; r++) {
        size_t sb = (size_t)send_cnt[r] * t.esize, rb = (size_t)all_cnt[(size_t)r * N + me] * t.esize;
        if (r == me) {
          if (sb) CK(cudaMemcpyAsync((char*)t.recvbuf.p + recv_off[r] * t.esize, (char*)t.sendbuf.p + (size_t)send_off[r] * t.esize,
                                     sb, cudaMemcpyDeviceToDevice, ctx->stream));
          continue;
        }
        if (sb) { NCK(g_nccl.send((char*)t.sendbuf.p + (size_t)send_off[r] * t.esize, sb, NCCL_UINT8, r, (nccl_comm)ctx->comm, ctx->stream)); bytes_sent += (int64_t)sb; }
        if (rb) NCK(g_nccl.recv((char*)t.recvbuf.p + recv_off[r] * t.esize, rb, NCCL_UINT8, r, (nccl_comm)ctx->comm, ctx->stream));
      }
    }
    NCK(g_nccl.group_end());
    CK(cudaEventRecord(e1, ctx->stream));
    hs_sync(ctx);
    float ms_x = 0.f;
    cudaEventElapsedTime(&ms_x, e0, e1);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    ctx->last_xchg_bytes = bytes_sent;
    ctx->last_xchg_ms = ms_x;
    for (auto& t : tv) t.sendbuf.release();
  }

  // ---- bucket-local build on the received rows
  std::vector<DevCol> rcols((size_t)ncols);
  {
    Stage s(ctx, "repack");
    for (int i = 0; i < ncols; i++) {
      rcols[i].type = dcols[i].type;
      rcols[i].length = nrecv;
    }
    for (auto& t : tv) {
      DevCol& rc = rcols[t.col];
      if (t.role == 0) {
        rc.data = t.recvbuf.p;
        rc.data_bytes = nrecv * t.esize;
        rc.owned.push_back(std::move(t.recvbuf));
      } else {
        DBuf bits(ctx, (size_t)((nrecv + 7) / 8)), zeros(ctx, 8);
        zeros.zero();
        LAUNCH(ctx, k_bytes_to_bits_d, dgrid((nrecv + 7) / 8), 256, 0, t.recvbuf.as<uint8_t>(), nrecv, bits.as<uint8_t>(),
               t.role == 1 ? zeros.as<unsigned long long>() : (unsigned long long*)nullptr);
        if (t.role == 1) {
          unsigned long long z = 0;
          CK(cudaMemcpyAsync(&z, zeros.p, 8, cudaMemcpyDeviceToHost, ctx->stream));
          hs_sync(ctx);
          if (z) { rc.validity = bits.as<uint8_t>(); rc.null_count = (int64_t)z; rc.owned.push_back(std::move(bits)); }
        } else {
          rc.data = bits.p;
          rc.data_bytes = (nrecv + 7) / 8;
          rc.owned.push_back(std::move(bits));
        }
      }
    }
  }
  dcols.clear();
  hs_build_core(ctx, rcols, n_indexed, nb, blo, bhi - blo, ocols, boff);
  Stage s(ctx, "d2h");
  *out = hs_finish_table(ctx, ocols, nrecv, nb, &boff);
  API_END(ctx)
}

}  // extern "C"
