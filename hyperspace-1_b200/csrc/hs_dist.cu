// hs_dist.cu — multi-GPU index build: one process per GPU, rows travel once to the GPU that owns their
// bucket — stored by the scatter kernel itself into the owners' peer-mapped receive buffers over
// NVLink/NVSwitch (fused / owner / copy modes, see try_fused), or through an NCCL all-to-all when peers
// cannot be mapped; everything after is bucket-local.
// Replaces Spark's ShuffleExchangeExec under df.repartition(numBuckets, cols)
// (HS/actions/CreateActionBase.scala:129-130) for N > 1 executors.
// NCCL is bound at run time (dlopen of libnccl.so.2) so that libhsgpu loads on hosts without it.
#include <dlfcn.h>

#include <map>
#include <string>

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

#define HS_MAX_PEERS_HOST 8
#define HS_MAX_PEER_COLS_HOST 16
#define NCCL_UINT8 1
#define NCCL_UINT64 5

// owner(b) = b * N / nb: contiguous bucket ranges, so every GPU writes a contiguous file-id range
static inline int owner_of(int b, int nranks, int nb) { return (int)((int64_t)b * nranks / nb); }

// copies runs of rows: run j = (src row, dst row, rows); 16-byte vector moves where alignment allows
__global__ void __launch_bounds__(256) k_copy_runs(const unsigned char* __restrict__ src, unsigned char* __restrict__ dst,
                                                   const uint32_t* __restrict__ runs, int esize) {
  const uint32_t s0 = runs[3 * blockIdx.x], d0 = runs[3 * blockIdx.x + 1], m = runs[3 * blockIdx.x + 2];
  const unsigned char* sp = src + (size_t)s0 * esize;
  unsigned char* dp = dst + (size_t)d0 * esize;
  size_t bytes = (size_t)m * esize;
  if ((((uintptr_t)sp | (uintptr_t)dp) & 15) == 0) {
    size_t nv = bytes >> 4;
    const uint4* s4 = (const uint4*)sp;
    uint4* d4 = (uint4*)dp;
    for (size_t i = threadIdx.x; i < nv; i += blockDim.x) d4[i] = s4[i];
    for (size_t i = (nv << 4) + threadIdx.x; i < bytes; i += blockDim.x) dp[i] = sp[i];
  } else if ((((uintptr_t)sp | (uintptr_t)dp) & 7) == 0 && (esize & 7) == 0) {
    size_t nv = bytes >> 3;
    const uint64_t* s8 = (const uint64_t*)sp;
    uint64_t* d8 = (uint64_t*)dp;
    for (size_t i = threadIdx.x; i < nv; i += blockDim.x) d8[i] = s8[i];
  } else {
    for (size_t i = threadIdx.x; i < bytes; i += blockDim.x) dp[i] = sp[i];
  }
}

// ---- peer-mapped receive buffers (fused path) ---------------------------------------------------------
// Every rank keeps one receive buffer per travelling column, allocated with cudaMalloc (cudaIpc needs that),
// grown when a build needs more, and exported to the other processes with cudaIpcGetMemHandle. The buffers
// are persistent: the sort reads them as its level-0 source and never hands them to the caller.
struct P2PSlot {
  void* ptr = nullptr;
  size_t cap = 0;
  cudaIpcMemHandle_t handle;
};
struct P2PState {
  std::vector<P2PSlot> slots;
  std::vector<void*> retired;                         // old buffers, freed once every peer has let go of them
  std::map<std::string, void*> opened;                // peer handle bytes -> local mapping
  std::map<int, std::string> current;                 // rank * 64 + slot -> handle currently mapped for it
  std::vector<void*> peer_ptr;                        // [rank][slot] for the current build
  bool unusable = false;
  cudaStream_t xs = nullptr;                          // grouped spans: the stream the barriers run on
  std::vector<cudaStream_t> cs;                       // grouped spans: streams the peer copies are spread over
  std::vector<cudaEvent_t> cs_ev;
  std::vector<cudaEvent_t> evs;                       // recycled events (no timing) for the groups
  cudaEvent_t x_e0 = nullptr, x_e1 = nullptr;
};
static P2PState* p2p_of(hs_ctx* c) {
  if (!c->p2p) c->p2p = new P2PState();
  return (P2PState*)c->p2p;
}
static void p2p_destroy(hs_ctx* c) {
  P2PState* st = (P2PState*)c->p2p;
  if (!st) return;
  for (auto& kv : st->opened) cudaIpcCloseMemHandle(kv.second);
  if (st->xs) cudaStreamDestroy(st->xs);
  for (cudaStream_t q : st->cs) cudaStreamDestroy(q);
  for (cudaEvent_t e : st->cs_ev) cudaEventDestroy(e);
  for (cudaEvent_t e : st->evs) cudaEventDestroy(e);
  if (st->x_e0) cudaEventDestroy(st->x_e0);
  if (st->x_e1) cudaEventDestroy(st->x_e1);
  for (auto& sl : st->slots) if (sl.ptr) cudaFree(sl.ptr);
  for (void* p : st->retired) cudaFree(p);
  delete st;
  c->p2p = nullptr;
  cudaGetLastError();
}

// copies runs of rows into the owners' buffers: run j = (src row, dst row, rows, owner); dst[owner] are
// peer-mapped pointers, so the stores of a warp leave as full 128-byte lines over NVLink
struct PeerCols {
  int32_t ncol;
  int32_t esize[HS_MAX_PEER_COLS_HOST];
  const unsigned char* src[HS_MAX_PEER_COLS_HOST];
  unsigned char* dst[HS_MAX_PEERS_HOST][HS_MAX_PEER_COLS_HOST];
};
template <typename V>
__device__ __forceinline__ void copy_vec4(const V* __restrict__ s, V* __restrict__ d, size_t nv) {
  size_t i = threadIdx.x;
  for (; i + 3 * 256 < nv; i += 4 * 256) {   // four loads in flight per thread before the first remote store
    const V a = s[i], b = s[i + 256], c = s[i + 512], e = s[i + 768];
    d[i] = a; d[i + 256] = b; d[i + 512] = c; d[i + 768] = e;
  }
  for (; i < nv; i += 256) d[i] = s[i];
}
// One CTA takes one run chunk and pushes it column after column (all columns in one launch: no tail per column).
__global__ void __launch_bounds__(256) k_copy_runs_peers(const __grid_constant__ PeerCols pc,
                                                         const uint32_t* __restrict__ runs, uint32_t nruns) {
  for (uint32_t j = blockIdx.x; j < nruns; j += gridDim.x) {
    const uint32_t s0 = runs[4 * j], d0 = runs[4 * j + 1], m = runs[4 * j + 2], owner = runs[4 * j + 3];
    for (int col = 0; col < pc.ncol; col++) {
      const int esize = pc.esize[col];
      const unsigned char* sp = pc.src[col] + (size_t)s0 * esize;
      unsigned char* dp = pc.dst[owner][col] + (size_t)d0 * esize;
      const size_t bytes = (size_t)m * esize;
      const uintptr_t al = (uintptr_t)sp | (uintptr_t)dp;
      if ((al & 15) == 0) {
        const size_t nv = bytes >> 4;
        copy_vec4<uint4>((const uint4*)sp, (uint4*)dp, nv);
        for (size_t i = (nv << 4) + threadIdx.x; i < bytes; i += blockDim.x) dp[i] = sp[i];
      } else if ((al & 7) == 0 && (bytes & 7) == 0) {
        copy_vec4<uint64_t>((const uint64_t*)sp, (uint64_t*)dp, bytes >> 3);
      } else if ((al & 3) == 0 && (bytes & 3) == 0) {
        copy_vec4<uint32_t>((const uint32_t*)sp, (uint32_t*)dp, bytes >> 2);
      } else {
        copy_vec4<unsigned char>(sp, dp, bytes);
      }
    }
  }
}

// The exchange step of the build. Each rank has scattered its rows by bucket; the run of bucket b goes to
// owner(b), which places the runs of all senders for that bucket next to each other. The receive therefore
// finishes the bucket partition: nothing is re-hashed or re-scattered after it.
struct NcclExchange : BuildExchange {
  int N, me, nb;
  std::vector<uint32_t> all_hist;   // [rank][bucket]
  NcclExchange(hs_ctx* c, int nb_) : N(c->nranks), me(c->rank), nb(nb_) {
    blo = nb; int bhi = 0;
    for (int b = 0; b < nb; b++)
      if (owner_of(b, N, nb) == me) { if (b < blo) blo = b; if (b + 1 > bhi) bhi = b + 1; }
    nbl = bhi > blo ? bhi - blo : 0;
    if (nbl == 0) blo = 0;
  }
  // Grouped spans queue copies on a second stream that read span_tmp: if the build unwinds early (an error on
  // this rank), those copies must have finished before the buffers go back to the cache.
  cudaStream_t busy_stream = nullptr;
  ~NcclExchange() override {
    if (busy_stream) cudaStreamSynchronize(busy_stream);
  }
  void share_stats(hs_ctx* c, const std::vector<uint32_t>& local_hist, uint64_t* minmax,
                   std::vector<uint8_t>& has_nulls) override {
    Stage s(c, "stats_exchange");
    // one all-gather carries the histogram, the key ranges and the NULL presence of every rank
    const size_t ncol = has_nulls.size();
    const size_t words = (size_t)nb + 4 * HS_MAX_KEYS + ncol;   // hist (u32) + minmax (u64 = 2 words each) + flags
    std::vector<uint32_t> mine(words), all(words * N);
    memcpy(mine.data(), local_hist.data(), (size_t)nb * 4);
    memcpy(mine.data() + nb, minmax, (size_t)2 * HS_MAX_KEYS * 8);
    for (size_t i = 0; i < ncol; i++) mine[nb + 4 * HS_MAX_KEYS + i] = has_nulls[i];
    DBuf d_mine(c, words * 4), d_all(c, words * 4 * N);
    hs_upload(c, d_mine.p, mine.data(), words * 4);
    NCK(g_nccl.allgather(d_mine.p, d_all.p, words * 4, NCCL_UINT8, (nccl_comm)c->comm, c->stream));
    CK(cudaMemcpyAsync(all.data(), d_all.p, words * 4 * N, cudaMemcpyDeviceToHost, c->stream));
    hs_sync(c);
    all_hist.assign((size_t)N * nb, 0);
    for (int k = 0; k < HS_MAX_KEYS; k++) { minmax[2 * k] = ~0ull; minmax[2 * k + 1] = 0; }
    for (int r = 0; r < N; r++) {
      memcpy(&all_hist[(size_t)r * nb], &all[words * r], (size_t)nb * 4);
      uint64_t mm[2 * HS_MAX_KEYS];
      memcpy(mm, &all[words * r + nb], sizeof mm);
      for (int k = 0; k < HS_MAX_KEYS; k++) {
        if (mm[2 * k] < minmax[2 * k]) minmax[2 * k] = mm[2 * k];
        if (mm[2 * k + 1] > minmax[2 * k + 1]) minmax[2 * k + 1] = mm[2 * k + 1];
      }
      for (size_t i = 0; i < ncol; i++) has_nulls[i] |= (uint8_t)(all[words * r + nb + 4 * HS_MAX_KEYS + i] != 0);
    }
  }
  int64_t exchange(hs_ctx* c, std::vector<DBuf*>& bufs, const std::vector<int>& esize,
                   const std::vector<uint32_t>& local_off, std::vector<uint32_t>& seg_off) override {
    // Wire format: one message per (peer, column) — the runs of a peer's buckets are adjacent in the
    // bucket-scattered buffer, so they travel as one span (NCCL is slow on hundreds of small messages).
    // The receiver then holds [sender][bucket] blocks and re-orders them to [bucket][sender] with one
    // run-copy kernel per column: after that every owned bucket is contiguous.
    std::vector<uint64_t> recv_cnt(N, 0), recv_off(N + 1, 0), send_cnt(N, 0), send_off(N, 0);
    for (int r = 0; r < N; r++) {
      for (int lb = 0; lb < nbl; lb++) recv_cnt[r] += all_hist[(size_t)r * nb + blo + lb];
      recv_off[r + 1] = recv_off[r] + recv_cnt[r];
    }
    for (int b = 0; b < nb; b++) send_cnt[owner_of(b, N, nb)] += all_hist[(size_t)me * nb + b];
    for (int r = 0, first = 0; r < N; r++) {
      while (first < nb && owner_of(first, N, nb) < r) first++;
      send_off[r] = first < nb ? local_off[first] : local_off[nb];
    }
    const uint64_t total = recv_off[N];
    // every rank holds every rank's histogram: a rank over the limit fails the build on ALL ranks, so that nobody is
    // left waiting in the exchange that follows
    for (int o = 0; o < N; o++) {
      uint64_t t = 0;
      for (int b = 0; b < nb; b++)
        if (owner_of(b, N, nb) == o)
          for (int r = 0; r < N; r++) t += all_hist[(size_t)r * nb + b];
      if (t >= 0xFFFFFFF0ull) HS_THROW(HS_EUNSUPPORTED, "more than 2^32-16 rows on GPU %d", o);
    }
    const int64_t nrecv = (int64_t)total;
    seg_off.assign((size_t)nbl + 1, 0);
    // run table: (src row, dst row, rows) for every (sender, owned bucket)
    std::vector<uint32_t> runs;
    {
      uint64_t at = 0;
      std::vector<uint64_t> src_at(N);
      for (int r = 0; r < N; r++) src_at[r] = recv_off[r];
      for (int lb = 0; lb < nbl; lb++) {
        seg_off[lb] = (uint32_t)at;
        for (int r = 0; r < N; r++) {
          uint32_t cnt = all_hist[(size_t)r * nb + blo + lb];
          // split long runs so that the copy has enough CTAs
          for (uint32_t done = 0; done < cnt; done += 65536) {
            uint32_t m = cnt - done < 65536 ? cnt - done : 65536;
            runs.push_back((uint32_t)(src_at[r] + done));
            runs.push_back((uint32_t)(at + done));
            runs.push_back(m);
          }
          src_at[r] += cnt;
          at += cnt;
        }
      }
      seg_off[nbl] = (uint32_t)at;
    }
    int64_t bytes_sent = 0;
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    std::vector<DBuf> recv(bufs.size());
    for (size_t i = 0; i < bufs.size(); i++) recv[i].alloc(c, (size_t)nrecv * esize[i] + 16);
    {
      Stage s(c, "all_to_all");
      CK(cudaEventRecord(e0, c->stream));
      for (size_t i = 0; i < bufs.size(); i++)
        if (send_cnt[me])
          CK(cudaMemcpyAsync((char*)recv[i].p + recv_off[me] * esize[i], (char*)bufs[i]->p + send_off[me] * esize[i],
                             send_cnt[me] * esize[i], cudaMemcpyDeviceToDevice, c->stream));
      NCK(g_nccl.group_start());
      for (size_t i = 0; i < bufs.size(); i++) {
        for (int r = 0; r < N; r++) {
          if (r == me) continue;
          if (send_cnt[r]) {
            NCK(g_nccl.send((char*)bufs[i]->p + send_off[r] * esize[i], send_cnt[r] * esize[i], NCCL_UINT8, r,
                            (nccl_comm)c->comm, c->stream));
            bytes_sent += (int64_t)(send_cnt[r] * esize[i]);
          }
          if (recv_cnt[r])
            NCK(g_nccl.recv((char*)recv[i].p + recv_off[r] * esize[i], recv_cnt[r] * esize[i], NCCL_UINT8, r,
                            (nccl_comm)c->comm, c->stream));
        }
      }
      NCK(g_nccl.group_end());
      CK(cudaEventRecord(e1, c->stream));
    }
    {
      // [sender][bucket] -> [bucket][sender]; the send buffers are reused as destination when they are
      // large enough, else fresh buffers are taken
      Stage s(c, "run_transpose");
      DBuf d_runs(c, runs.size() * 4 + 16);
      if (!runs.empty())
        hs_upload(c, d_runs.p, runs.data(), runs.size() * 4);
      for (size_t i = 0; i < bufs.size(); i++) {
        DBuf dst(c, (size_t)nrecv * esize[i] + 16);
        if (!runs.empty())
          LAUNCH(c, k_copy_runs, (unsigned)(runs.size() / 3), 256, 0, (const unsigned char*)recv[i].p,
                 (unsigned char*)dst.p, d_runs.as<uint32_t>(), esize[i]);
        *bufs[i] = std::move(dst);
        recv[i].release();
      }
    }
    hs_sync(c);
    float ms_x = 0.f;
    cudaEventElapsedTime(&ms_x, e0, e1);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    c->last_xchg_bytes = bytes_sent;
    c->last_xchg_ms = ms_x;
    return nrecv;
  }

  // ---- fused path ------------------------------------------------------------------------------------
  std::vector<uint64_t> rank_rows;   // rows every rank receives
  int xmode = 0;                     // 0 fused, 1 owner, 2 copy, 3 grouped spans (see try_fused)
  std::vector<int> xes;              // element sizes of the travelling columns (+ the bucket ids in owner mode)
  std::vector<DBuf> owner_sorted;    // owner mode: the received rows, scattered by bucket (the sort's source)
  bool narrow = false;               // owner mode: the INT64 key travels as u32 (value - minimum)

  bool try_fused(hs_ctx* c, const std::vector<int>& esize) override {
    const char* off = getenv("HSGPU_NO_P2P");
    if (off && *off && *off != '0') return false;
    if (N > HS_MAX_PEERS_HOST || (int)esize.size() > HS_MAX_PEER_COLS_HOST) return false;
    P2PState* st = p2p_of(c);
    if (st->unusable) return false;
    Stage s(c, "p2p_setup");
    // How the rows travel (HSGPU_P2P_MODE=fused|owner|copy overrides):
    //  fused  the bucket scatter stores every row straight at its place in the owner's bucket. Few buckets:
    //         a tile of 4096 rows holds runs of tens of rows per bucket, decent NVLink packets.
    //  owner  thousands of buckets would make every row a packet of its own (measured 3.5 GB/s per GPU at 8192
    //         buckets). The scatter then groups rows by OWNER only (runs of hundreds of rows per destination),
    //         the bucket ids travel as one more column, and the owner scatters what it received by bucket.
    //  copy   scatter by bucket locally, then push whole bucket runs (kept for comparison).
    //  spans  the sender groups its rows by (owner, bucket group) — a few dozen bins, an HBM-speed pass — and
    //         the spans travel as plain peer copies on the copy engines, group after group on a second stream;
    //         the owner scatters a group by bucket and sorts its buckets while the next group is still on the
    //         wire (spans_exchange / deliver_group). The SMs never wait for NVLink except for the first group.
    const char* mode = getenv("HSGPU_P2P_MODE");
    xmode = nb <= 256 ? 0 : 3;
    if (mode && !strcmp(mode, "fused")) xmode = 0;
    if (mode && !strcmp(mode, "owner")) xmode = 1;
    if (mode && !strcmp(mode, "copy")) xmode = 2;
    if (mode && !strcmp(mode, "spans")) xmode = 3;
    if (xmode == 3 && (nb + N - 1) / N > 2048) xmode = 1;   // the owner-side scatter of a group is one-level
    ngroups = 0;
    xes = esize;
    // owner mode: an INT64 key whose values span less than 2^32 crosses NVLink as 4 bytes (HSGPU_WIDE_KEYS=1: no)
    const char* wide = getenv("HSGPU_WIDE_KEYS");
    narrow = (xmode == 1 || xmode == 3) && key_col >= 0 && key_col < (int)esize.size() && esize[key_col] == 8 &&
             key_span < 0xFFFFFFFFull && nbl <= 2048 && !(wide && *wide && *wide != '0');
    if (narrow) xes[key_col] = 4;
    if (xmode == 1 || xmode == 3) xes.push_back(2);       // the bucket-id column
    if ((int)xes.size() > HS_MAX_PEER_COLS_HOST) return false;
    const int ncol = (int)xes.size();
    const std::vector<int>& esize_x = xes;
    rank_rows.assign(N, 0);
    for (int b = 0; b < nb; b++)
      for (int r = 0; r < N; r++) rank_rows[owner_of(b, N, nb)] += all_hist[(size_t)r * nb + b];
    for (int o = 0; o < N; o++)   // decided alike on every rank (see exchange())
      if (rank_rows[o] >= 0xFFFFFFF0ull) HS_THROW(HS_EUNSUPPORTED, "more than 2^32-16 rows on GPU %d", o);
    // peers have finished with buffers retired one build ago (they passed an all-gather since)
    for (void* p : st->retired) cudaFree(p);
    st->retired.clear();
    bool ok = true;
    if ((int)st->slots.size() < ncol) st->slots.resize(ncol);
    for (int i = 0; i < ncol && ok; i++) {
      size_t need = (size_t)rank_rows[me] * esize_x[i] + 256;
      P2PSlot& sl = st->slots[i];
      if (sl.cap < need) {
        if (sl.ptr) st->retired.push_back(sl.ptr);
        sl.ptr = nullptr;
        size_t cap = need + need / 4;
        if (cudaMalloc(&sl.ptr, cap) != cudaSuccess) { cudaGetLastError(); sl.ptr = nullptr; sl.cap = 0; ok = false; break; }
        sl.cap = cap;
        if (cudaIpcGetMemHandle(&sl.handle, sl.ptr) != cudaSuccess) { cudaGetLastError(); ok = false; }
      }
    }
    // all-gather: [ok flag | handles]; doubles as the barrier that says "nobody still reads the buffers"
    const size_t rec = 8 + (size_t)ncol * sizeof(cudaIpcMemHandle_t);
    std::vector<unsigned char> mine(rec, 0), all(rec * N, 0);
    mine[0] = ok ? 1 : 0;
    for (int i = 0; i < ncol && ok; i++) memcpy(&mine[8 + i * sizeof(cudaIpcMemHandle_t)], &st->slots[i].handle, sizeof(cudaIpcMemHandle_t));
    DBuf d_mine(c, rec), d_all(c, rec * N);
    hs_upload(c, d_mine.p, mine.data(), rec);
    NCK(g_nccl.allgather(d_mine.p, d_all.p, rec, NCCL_UINT8, (nccl_comm)c->comm, c->stream));
    CK(cudaMemcpyAsync(all.data(), d_all.p, rec * N, cudaMemcpyDeviceToHost, c->stream));
    hs_sync(c);
    for (int r = 0; r < N; r++) ok = ok && all[rec * r] == 1;
    st->peer_ptr.assign((size_t)N * ncol, nullptr);
    for (int r = 0; r < N && ok; r++) {
      for (int i = 0; i < ncol && ok; i++) {
        if (r == me) { st->peer_ptr[(size_t)r * ncol + i] = st->slots[i].ptr; continue; }
        std::string key((const char*)&all[rec * r + 8 + i * sizeof(cudaIpcMemHandle_t)], sizeof(cudaIpcMemHandle_t));
        // a peer that re-allocated a slot published a new handle: let go of the old mapping first, so that
        // the owner may free the old buffer after the next all-gather
        auto cur = st->current.find(r * 64 + i);
        if (cur != st->current.end() && cur->second != key) {
          auto old = st->opened.find(cur->second);
          if (old != st->opened.end()) { cudaIpcCloseMemHandle(old->second); st->opened.erase(old); }
        }
        st->current[r * 64 + i] = key;
        auto it = st->opened.find(key);
        if (it == st->opened.end()) {
          cudaIpcMemHandle_t h;
          memcpy(&h, key.data(), sizeof h);
          void* p = nullptr;
          if (cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); ok = false; break; }
          it = st->opened.emplace(key, p).first;
        }
        st->peer_ptr[(size_t)r * ncol + i] = it->second;
      }
    }
    // every rank must take the same path
    unsigned char flag = ok ? 1 : 0;
    std::vector<unsigned char> flags(N, 0);
    DBuf d_f(c, 16), d_fa(c, 16 * (size_t)N);
    hs_upload(c, d_f.p, &flag, 1);
    NCK(g_nccl.allgather(d_f.p, d_fa.p, 1, NCCL_UINT8, (nccl_comm)c->comm, c->stream));
    CK(cudaMemcpyAsync(flags.data(), d_fa.p, N, cudaMemcpyDeviceToHost, c->stream));
    hs_sync(c);
    for (int r = 0; r < N; r++) ok = ok && flags[r] == 1;
    if (!ok) st->unusable = true;   // fall back to the NCCL exchange, now and later
    return ok;
  }

  int64_t scatter_exchange(hs_ctx* c, const void* const* srcs, const std::vector<int>& esize, const uint16_t* d_bucket,
                           int64_t n_local, std::vector<void*>& recv, std::vector<uint32_t>& seg_off) override {
    P2PState* st = p2p_of(c);
    const int ncol = (int)esize.size();
    // where this rank's run of every bucket starts inside the owner's buffer
    std::vector<uint32_t> run_base((size_t)nb, 0);
    std::vector<uint64_t> at(N, 0);
    for (int b = 0; b < nb; b++) {
      const int r = owner_of(b, N, nb);
      uint64_t before = 0;
      for (int q = 0; q < me; q++) before += all_hist[(size_t)q * nb + b];
      run_base[b] = (uint32_t)(at[r] + before);
      for (int q = 0; q < N; q++) at[r] += all_hist[(size_t)q * nb + b];
    }
    seg_off.assign((size_t)nbl + 1, 0);
    {
      uint64_t acc = 0;
      for (int lb = 0; lb < nbl; lb++) {
        seg_off[lb] = (uint32_t)acc;
        for (int q = 0; q < N; q++) acc += all_hist[(size_t)q * nb + blo + lb];
      }
      seg_off[nbl] = (uint32_t)acc;
    }
    int64_t bytes_sent = 0;
    for (int b = 0; b < nb; b++)
      if (owner_of(b, N, nb) != me)
        for (size_t i = 0; i < xes.size(); i++) bytes_sent += (int64_t)all_hist[(size_t)me * nb + b] * xes[i];
    if (xmode == 3) {
      c->last_xchg_bytes = bytes_sent;
      return spans_exchange(c, st, srcs, esize, d_bucket, n_local, recv, seg_off);
    }
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    const bool one_kernel = xmode == 0;
    if (xmode == 1) {
      // grouped by owner: my rows for owner o land behind those of the ranks before me
      Stage s(c, "scatter_to_owners");
      std::vector<uint32_t> base((size_t)N, 0);
      for (int o = 0; o < N; o++)
        for (int q = 0; q < me; q++)
          for (int b = 0; b < nb; b++)
            if (owner_of(b, N, nb) == o) base[o] += all_hist[(size_t)q * nb + b];
      DBuf d_base(c, (size_t)N * 4);
      hs_upload(c, d_base.p, base.data(), (size_t)N * 4);
      std::vector<const void*> xsrc(srcs, srcs + ncol);
      xsrc.push_back(d_bucket);
      CK(cudaEventRecord(e0, c->stream));
      hs_bucket_scatter_peers(c, xsrc.data(), xes.data(), ncol + 1, st->peer_ptr.data(), N, d_bucket, n_local, nb,
                              d_base.as<uint32_t>(), true, narrow ? key_col : -1, key_min);
      CK(cudaEventRecord(e1, c->stream));
      hs_sync(c);   // my stores have left this GPU
    } else if (one_kernel) {
      // one kernel: the bucket scatter stores every row at its final place in the owner's memory
      Stage s(c, "scatter_to_owners");
      DBuf d_base(c, (size_t)nb * 4);
      hs_upload(c, d_base.p, run_base.data(), (size_t)nb * 4);
      CK(cudaEventRecord(e0, c->stream));
      hs_bucket_scatter_peers(c, srcs, esize.data(), ncol, st->peer_ptr.data(), N, d_bucket, n_local, nb,
                              d_base.as<uint32_t>());
      CK(cudaEventRecord(e1, c->stream));
      hs_sync(c);   // my stores have left this GPU
    } else {
      // two kernels (HSGPU_P2P_MODE=copy; 3.6 ms vs 3.06 ms fused at N=2): scatter locally, then push whole
      // bucket runs to their final place in the owners' memory with full-line peer stores
      std::vector<uint32_t> local_off((size_t)nb + 1, 0);
      for (int b = 0; b < nb; b++) local_off[b + 1] = local_off[b] + all_hist[(size_t)me * nb + b];
      std::vector<DBuf> tmp((size_t)ncol);
      MoveSet ms;
      memset(&ms, 0, sizeof ms);
      ms.n = ncol;
      for (int i = 0; i < ncol; i++) {
        tmp[i].alloc(c, (size_t)n_local * esize[i] + 16);
        ms.esize[i] = esize[i];
        ms.a[i] = tmp[i].p;
      }
      {
        Stage s(c, "bucket_scatter");
        DBuf d_loff(c, ((size_t)nb + 1) * 4);
        hs_upload(c, d_loff.p, local_off.data(), ((size_t)nb + 1) * 4);
        hs_bucket_scatter(c, srcs, ms, d_bucket, n_local, nb, d_loff.as<uint32_t>());
      }
      // every sender starts with a different owner (me+1, me+2, ..., me last): if all ranks walked the buckets in
      // the same order they would all push into the same GPU at the same time (measured: 32 ms of the 8-GPU
      // cfg5 build waiting at the barrier for the ranks served last)
      std::vector<uint32_t> runs;
      std::vector<int> order;
      for (int k = 1; k <= N; k++)
        for (int b = 0; b < nb; b++)
          if (owner_of(b, N, nb) == (me + k) % N) order.push_back(b);
      for (int b : order) {
        uint32_t cnt = all_hist[(size_t)me * nb + b];
        for (uint32_t done = 0; done < cnt; done += 16384) {
          runs.push_back(local_off[b] + done);
          runs.push_back(run_base[b] + done);
          runs.push_back(cnt - done < 16384 ? cnt - done : 16384);
          runs.push_back((uint32_t)owner_of(b, N, nb));
        }
      }
      Stage s(c, "runs_to_owners");
      DBuf d_runs(c, runs.size() * 4 + 16);
      if (!runs.empty()) hs_upload(c, d_runs.p, runs.data(), runs.size() * 4);
      CK(cudaEventRecord(e0, c->stream));
      if (!runs.empty()) {
        PeerCols pc;
        memset(&pc, 0, sizeof pc);
        pc.ncol = ncol;
        for (int i = 0; i < ncol; i++) {
          pc.esize[i] = esize[i];
          pc.src[i] = (const unsigned char*)tmp[i].p;
          for (int r = 0; r < N; r++) pc.dst[r][i] = (unsigned char*)st->peer_ptr[(size_t)r * ncol + i];
        }
        const uint32_t nruns = (uint32_t)(runs.size() / 4);
        LAUNCH(c, k_copy_runs_peers, std::min<uint32_t>(nruns, HS_NUM_SMS * 8), 256, 0, pc, d_runs.as<uint32_t>(), nruns);
      }
      CK(cudaEventRecord(e1, c->stream));
      hs_sync(c);   // my stores have left this GPU
    }
    {
      // nobody may read its receive buffer before every sender has finished: a tiny all-gather is the barrier
      Stage s(c, "p2p_barrier");
      DBuf d_f(c, 16), d_fa(c, 16 * (size_t)N);
      NCK(g_nccl.allgather(d_f.p, d_fa.p, 1, NCCL_UINT8, (nccl_comm)c->comm, c->stream));
      hs_sync(c);
    }
    float ms_x = 0.f;
    cudaEventElapsedTime(&ms_x, e0, e1);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    c->last_xchg_bytes = bytes_sent;
    c->last_xchg_ms = ms_x;
    recv.resize(ncol);
    if (xmode == 1) {
      // what arrived is grouped by sender only: scatter it by bucket (ids in the last received column, this
      // rank's buckets start at blo) into buffers that then serve as the sort's source
      Stage s(c, "bucket_scatter");
      const int64_t nrecv = (int64_t)rank_rows[me];
      owner_sorted.clear();
      owner_sorted.resize((size_t)ncol);
      MoveSet ms;
      memset(&ms, 0, sizeof ms);
      ms.n = ncol;
      std::vector<const void*> rsrc((size_t)ncol);
      for (int i = 0; i < ncol; i++) {
        owner_sorted[i].alloc(c, (size_t)nrecv * esize[i] + 16);
        ms.esize[i] = esize[i];
        ms.a[i] = owner_sorted[i].p;
        rsrc[i] = st->slots[i].ptr;
        recv[i] = owner_sorted[i].p;
      }
      DBuf d_off(c, ((size_t)nbl + 1) * 4);
      hs_upload(c, d_off.p, seg_off.data(), ((size_t)nbl + 1) * 4);
      hs_bucket_scatter(c, rsrc.data(), ms, (const uint16_t*)st->slots[ncol].ptr, nrecv, nbl, d_off.as<uint32_t>(), blo,
                        narrow ? key_col : -1, key_min);
      hs_sync(c);   // seg_off is the caller's vector, d_off dies here
      return nrecv;
    }
    for (int i = 0; i < ncol; i++) recv[i] = st->slots[i].ptr;
    return (int64_t)rank_rows[me];
  }

  // ---- grouped spans --------------------------------------------------------------------------------
  int G = 0;
  std::vector<uint64_t> grp_off;      // [G + 1] rows of this rank's receive buffers before group g
  std::vector<uint32_t> own_off;      // [G + 1] where this rank's own rows of group g sit in span_tmp (they do not travel)
  std::vector<int> grp_lb;            // [G + 1] first local bucket of group g
  std::vector<cudaEvent_t> grp_ev;    // group g has arrived from every sender
  std::vector<DBuf> span_tmp;         // the sender's rows grouped by (owner, group); read by the copies
  DBuf bar_a, bar_b, d_seg;
  std::vector<int> col_es;            // element sizes of the travelling columns (without the ids)

  int64_t spans_exchange(hs_ctx* c, P2PState* st, const void* const* srcs, const std::vector<int>& esize,
                         const uint16_t* d_bucket, int64_t n_local, std::vector<void*>& recv,
                         const std::vector<uint32_t>& seg_off) {
    const int ncol = (int)esize.size(), ncx = ncol + 1;   // + the bucket ids
    const char* ge = getenv("HSGPU_XCHG_GROUPS");
    G = ge && atoi(ge) > 0 ? std::min(atoi(ge), 16) : 4;   // measured at N=8: 3 groups 8.30 ms, 4 groups 8.02, 5 groups 9.39
    col_es = esize;
    // owner o holds buckets [olo[o], olo[o + 1]); its group g the local buckets lb with lb * G / count == g
    std::vector<int> olo((size_t)N + 1, nb);
    for (int b = nb - 1; b >= 0; b--) olo[owner_of(b, N, nb)] = b;
    for (int o = N - 1; o >= 0; o--) if (olo[o] > olo[o + 1]) olo[o] = olo[o + 1];   // owners without buckets
    // group sizes 1, 2, 3, ..., 3, 2, 1: the first group is small, so the owners start sorting early; the last one
    // is small, so little is left to sort when the last row has arrived; the middle ones are large, so the sort
    // kernels are not cut into too many short launches
    std::vector<int> wcum((size_t)G + 1, 0);
    for (int g = 0; g < G; g++) wcum[g + 1] = wcum[g] + std::min(std::min(g + 1, G - g), 3);
    auto group_of = [&](int b, int o) {
      const int64_t pos = (int64_t)(b - olo[o]) * wcum[G] / (olo[o + 1] - olo[o]);   // in [0, sum of weights)
      int g = 0;
      while (wcum[g + 1] <= pos) g++;
      return g;
    };
    const int nbins = N * G;
    std::vector<uint16_t> lut((size_t)nb);
    std::vector<uint64_t> cnt((size_t)N * nbins, 0);
    for (int b = 0; b < nb; b++) {
      const int o = owner_of(b, N, nb);
      lut[b] = (uint16_t)(o * G + group_of(b, o));
      for (int r = 0; r < N; r++) cnt[(size_t)r * nbins + lut[b]] += all_hist[(size_t)r * nb + b];
    }
    std::vector<uint32_t> my_off((size_t)nbins + 1, 0);
    for (int i = 0; i < nbins; i++) my_off[i + 1] = my_off[i] + (uint32_t)cnt[(size_t)me * nbins + i];
    // owner o's receive buffers are laid out [group][sender]; its own rows stay where its grouping put them
    auto recv_off = [&](int o, int g, int r) {
      uint64_t at = 0;
      for (int gg = 0; gg < g; gg++)
        for (int q = 0; q < N; q++) if (q != o) at += cnt[(size_t)q * nbins + o * G + gg];
      for (int q = 0; q < r; q++) if (q != o) at += cnt[(size_t)q * nbins + o * G + g];
      return at;
    };
    own_off.assign((size_t)G + 1, 0);
    for (int g = 0; g <= G; g++) own_off[g] = my_off[(size_t)me * G + g];
    grp_off.assign((size_t)G + 1, 0);
    grp_lb.assign((size_t)G + 1, nbl);
    for (int g = 0; g <= G; g++) grp_off[g] = recv_off(me, g, 0);
    for (int lb = nbl - 1; lb >= 0; lb--) grp_lb[group_of(blo + lb, me)] = lb;
    for (int g = G - 1; g >= 0; g--) if (grp_lb[g] > grp_lb[g + 1]) grp_lb[g] = grp_lb[g + 1];   // empty groups
    if (!st->xs) {
      int lo = 0, hi = 0;
      CK(cudaDeviceGetStreamPriorityRange(&lo, &hi));
      CK(cudaStreamCreateWithPriority(&st->xs, cudaStreamNonBlocking, hi));
      CK(cudaEventCreate(&st->x_e0));
      CK(cudaEventCreate(&st->x_e1));
    }
    // the copies of a group can be spread over several streams (HSGPU_XCHG_STREAMS); measured at N=8 that does
    // not help — 437 GB/s per GPU with one stream, 440 with four, 428 with seven — so one is the default
    const char* se = getenv("HSGPU_XCHG_STREAMS");
    const int S = std::max(1, std::min(se && atoi(se) > 0 ? atoi(se) : 1, 8));
    while ((int)st->cs.size() < S) {
      int lo = 0, hi = 0;
      cudaStream_t q;
      cudaEvent_t e;
      CK(cudaDeviceGetStreamPriorityRange(&lo, &hi));
      CK(cudaStreamCreateWithPriority(&q, cudaStreamNonBlocking, hi));
      CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
      st->cs.push_back(q);
      st->cs_ev.push_back(e);
    }
    while ((int)st->evs.size() < G + 1) {
      cudaEvent_t e;
      CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
      st->evs.push_back(e);
    }
    grp_ev.assign(st->evs.begin(), st->evs.begin() + G);
    cudaEvent_t ev_grouped = st->evs[G];
    span_tmp.clear();
    span_tmp.resize((size_t)ncx);
    {
      Stage s(c, "group_scatter");
      DBuf d_lut(c, (size_t)nb * 2 + 16), d_boff(c, ((size_t)nbins + 1) * 4);
      hs_upload(c, d_lut.p, lut.data(), (size_t)nb * 2);
      hs_upload(c, d_boff.p, my_off.data(), ((size_t)nbins + 1) * 4);
      std::vector<const void*> xsrc(srcs, srcs + ncol);
      xsrc.push_back(d_bucket);
      std::vector<void*> xdst((size_t)ncx);
      for (int i = 0; i < ncx; i++) {
        span_tmp[i].alloc(c, (size_t)n_local * xes[i] + 16);
        xdst[i] = span_tmp[i].p;
      }
      hs_scatter_by_bins(c, xsrc.data(), xdst.data(), xes.data(), ncx, d_bucket, n_local, (const uint16_t*)d_lut.p, nbins,
                         d_boff.as<uint32_t>(), narrow ? key_col : -1, key_min);
      CK(cudaEventRecord(ev_grouped, c->stream));
    }
    bar_a.alloc(c, 16);
    bar_b.alloc(c, 16 * (size_t)N);
    // everything that crosses NVLink is queued now, on the second stream: per group one copy per (owner, column),
    // then the barrier that tells every owner its group is complete
    busy_stream = st->xs;
    CK(cudaStreamWaitEvent(st->xs, ev_grouped, 0));
    CK(cudaEventRecord(st->x_e0, st->xs));
    for (int q = 0; q < S; q++) CK(cudaStreamWaitEvent(st->cs[q], ev_grouped, 0));
    for (int g = 0; g < G; g++) {
      for (int k = 1; k < N; k++) {        // every sender starts with a different owner
        const int o = (me + k) % N, bin = o * G + g;
        const uint64_t rows = cnt[(size_t)me * nbins + bin];
        if (!rows) continue;
        const uint64_t dst_row = recv_off(o, g, me);
        cudaStream_t q = st->cs[(size_t)(k - 1) % S];
        for (int i = 0; i < ncx; i++)
          CK(cudaMemcpyAsync((char*)st->peer_ptr[(size_t)o * ncx + i] + dst_row * xes[i],
                             (const char*)span_tmp[i].p + (uint64_t)my_off[bin] * xes[i], rows * xes[i], cudaMemcpyDefault, q));
      }
      // the barrier stream joins the copy streams (which go straight on with the next group)
      for (int q = 0; q < S; q++) {
        CK(cudaEventRecord(st->cs_ev[q], st->cs[q]));
        CK(cudaStreamWaitEvent(st->xs, st->cs_ev[q], 0));
      }
      if (g == G - 1) CK(cudaEventRecord(st->x_e1, st->xs));
      NCK(g_nccl.allgather(bar_a.p, bar_b.p, 1, NCCL_UINT8, (nccl_comm)c->comm, st->xs));
      CK(cudaEventRecord(grp_ev[g], st->xs));
    }
    // the owner side: buffers the groups are scattered into by bucket (the sort's source)
    const int64_t nrecv = (int64_t)rank_rows[me];
    owner_sorted.clear();
    owner_sorted.resize((size_t)ncol);
    recv.resize(ncol);
    for (int i = 0; i < ncol; i++) {
      owner_sorted[i].alloc(c, (size_t)nrecv * esize[i] + 16);
      recv[i] = owner_sorted[i].p;
    }
    d_seg.alloc(c, ((size_t)nbl + 1) * 4);
    hs_upload(c, d_seg.p, seg_off.data(), ((size_t)nbl + 1) * 4);
    ngroups = G;
    return nrecv;
  }

  void deliver_group(hs_ctx* c, int g, int* lb0, int* lb1) override {
    P2PState* st = p2p_of(c);
    const int ncol = (int)col_es.size();
    CK(cudaStreamWaitEvent(c->stream, grp_ev[g], 0));
    *lb0 = grp_lb[g];
    *lb1 = grp_lb[g + 1];
    // two sources fill the group's buckets: what arrived from the peers and this rank's own rows; d_seg holds
    // the next free row of every bucket across both
    const int64_t rows = (int64_t)(grp_off[g + 1] - grp_off[g]), own = (int64_t)own_off[g + 1] - (int64_t)own_off[g];
    if (rows > 0 || own > 0) {
      Stage s(c, "bucket_scatter");
      MoveSet ms;
      memset(&ms, 0, sizeof ms);
      ms.n = ncol;
      std::vector<const void*> rsrc((size_t)ncol);
      for (int i = 0; i < ncol; i++) {
        ms.esize[i] = col_es[i];
        ms.a[i] = owner_sorted[i].p;
        rsrc[i] = (const char*)st->slots[i].ptr + grp_off[g] * xes[i];   // xes: a narrowed key is 4 bytes here
      }
      const int wcol = narrow ? key_col : -1;
      hs_bucket_scatter(c, rsrc.data(), ms, (const uint16_t*)st->slots[ncol].ptr + grp_off[g], rows, nbl, nullptr, blo, wcol,
                        key_min, d_seg.as<uint32_t>());
      for (int i = 0; i < ncol; i++) rsrc[i] = (const char*)span_tmp[i].p + (size_t)own_off[g] * xes[i];
      hs_bucket_scatter(c, rsrc.data(), ms, (const uint16_t*)span_tmp[ncol].p + own_off[g], own, nbl, nullptr, blo, wcol,
                        key_min, d_seg.as<uint32_t>());
    }
    if (g == G - 1) {
      // the stream has passed the last barrier: the copies are done with the grouped rows
      busy_stream = nullptr;
      span_tmp.clear();
      bar_a.release();
      bar_b.release();
      float ms_x = 0.f;
      CK(cudaEventSynchronize(st->x_e1));
      cudaEventElapsedTime(&ms_x, st->x_e0, st->x_e1);
      c->last_xchg_ms = ms_x;
    }
  }
};

// ------------------------------------------------------------------ rows first, then a local build
// The general multi-GPU build, for shapes the fused scatter cannot move (string columns travel as a row id and
// are gathered from the SOURCE column at the end, which lives on the sender; so do keys wider than one 64-bit
// chunk): every rank groups its rows by the GPU that owns their bucket, the groups change hands in one grouped
// ncclSend/ncclRecv round (fixed-width values, one byte per row where a column holds NULLs anywhere, string
// lengths and string bytes as separate arrays), and each GPU then runs the ordinary single-GPU build over the
// rows it received, restricted to its own bucket range. One more pass over the rows than the fused path.
static __global__ void __launch_bounds__(256) k_rf_iota(uint32_t* out, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = (uint32_t)i;
}
template <typename OFF>
static __global__ void __launch_bounds__(256) k_rf_lens(const OFF* __restrict__ off, int64_t n, uint32_t* __restrict__ lens) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    lens[i] = (uint32_t)(off[i + 1] - off[i]);
}
static __global__ void __launch_bounds__(256) k_rf_bits_to_bytes(const uint8_t* __restrict__ bits, int64_t n, uint8_t* __restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = bits ? (bits[i >> 3] >> (i & 7)) & 1 : 1;
}
static __global__ void __launch_bounds__(256) k_rf_bytes_to_bits(const uint8_t* __restrict__ in, int64_t n, uint8_t* __restrict__ bits,
                                                                 unsigned long long* zeros) {
  const int64_t nb = (n + 7) >> 3;
  unsigned z = 0;
  for (int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; b < nb; b += (int64_t)gridDim.x * blockDim.x) {
    unsigned v = 0;
    for (int k = 0; k < 8; k++) {
      const int64_t i = b * 8 + k;
      if (i < n) { if (in[i]) v |= 1u << k; else z++; }
    }
    bits[b] = (uint8_t)v;
  }
  if (zeros && z) atomicAdd(zeros, (unsigned long long)z);
}
static unsigned rf_grid(int64_t n) {
  int64_t g = (n + 255) / 256;
  return (unsigned)std::max<int64_t>(1, std::min<int64_t>(g, HS_NUM_SMS * 16));
}

struct RfPart {           // one array that changes hands
  const uint8_t* src = nullptr;            // sender side, grouped by owner
  std::vector<uint64_t> cut;               // N + 1 byte offsets of the owners' pieces inside src
  DBuf recv;                               // receiver side: the pieces of rank 0, 1, ... one after the other
  uint64_t recv_bytes = 0;
};

static void dist_build_rows_first(hs_ctx* c, std::vector<DevCol>& cols, int n_indexed, int nb,
                                  std::vector<OutCol>& out, std::vector<int64_t>& boff) {
  const int N = c->nranks, me = c->rank, ncols = (int)cols.size();
  nccl_comm comm = (nccl_comm)c->comm;
  const int64_t n = cols[0].length;
  if (n >= 0xFFFFFFF0ll) HS_THROW(HS_EUNSUPPORTED, "more than 2^32-16 rows on one GPU");
  if (nb > 8192 * N) HS_THROW(HS_EUNSUPPORTED, "more than 8192 buckets per GPU");
  // ---- bucket of every row, rows per bucket; rows grouped by bucket (hence by owner: owners hold bucket ranges)
  std::vector<uint32_t> h_hist((size_t)nb, 0);
  DBuf perm(c, (size_t)std::max<int64_t>(n, 1) * 4);
  {
    Stage s(c, "group_by_owner");
    KeySpec ks = hs_make_keyspec(cols, n_indexed);
    DBuf bucket(c, (size_t)std::max<int64_t>(n, 1) * 2 + 16), hist(c, (size_t)nb * 4), minmax(c, (2 * HS_MAX_KEYS + 1) * 8);
    hs_hash_bucket_hist(c, ks, cols, n, nb, 0, nb, bucket.as<uint16_t>(), hist.as<uint32_t>(), minmax.as<uint64_t>());
    CK(cudaMemcpyAsync(h_hist.data(), hist.p, (size_t)nb * 4, cudaMemcpyDeviceToHost, c->stream));
    hs_sync(c);
    std::vector<uint32_t> off((size_t)nb + 1, 0);
    for (int b = 0; b < nb; b++) off[b + 1] = off[b] + h_hist[b];
    DBuf d_off(c, ((size_t)nb + 1) * 4), iota(c, (size_t)std::max<int64_t>(n, 1) * 4);
    hs_upload(c, d_off.p, off.data(), ((size_t)nb + 1) * 4);
    if (n > 0) {
      LAUNCH(c, k_rf_iota, rf_grid(n), 256, 0, iota.as<uint32_t>(), n);
      MoveSet ms;
      memset(&ms, 0, sizeof ms);
      ms.n = 1;
      ms.esize[0] = 4;
      ms.a[0] = perm.p;
      const void* src0 = iota.p;
      if (nb <= 8192) {
        hs_bucket_scatter(c, &src0, ms, bucket.as<uint16_t>(), n, nb, d_off.as<uint32_t>());
      } else {
        HS_THROW(HS_EUNSUPPORTED, "the rows-first multi-GPU build takes at most 8192 buckets");
      }
    }
    hs_sync(c);   // off and the scatter's temporaries go out of scope
  }
  std::vector<uint64_t> row_cut((size_t)N + 1, 0);    // rows for owner o: [row_cut[o], row_cut[o+1])
  {
    uint64_t run = 0;
    int o = 0;
    for (int b = 0; b < nb; b++) {
      while (o < owner_of(b, N, nb)) row_cut[++o] = run;
      run += h_hist[b];
    }
    while (o < N) row_cut[++o] = run;
  }
  // ---- which columns hold NULLs on some rank (everyone must send the same arrays)
  std::vector<uint8_t> nulls_here((size_t)ncols, 0), nulls_any((size_t)ncols * N, 0);
  for (int i = 0; i < ncols; i++) nulls_here[i] = cols[i].validity && cols[i].null_count != 0;
  {
    DBuf d_in(c, (size_t)ncols), d_all(c, (size_t)ncols * N);
    hs_upload(c, d_in.p, nulls_here.data(), (size_t)ncols);
    NCK(g_nccl.allgather(d_in.p, d_all.p, (size_t)ncols, NCCL_UINT8, comm, c->stream));
    CK(cudaMemcpyAsync(nulls_any.data(), d_all.p, (size_t)ncols * N, cudaMemcpyDeviceToHost, c->stream));
    hs_sync(c);
  }
  auto has_nulls = [&](int i) { for (int r = 0; r < N; r++) if (nulls_any[(size_t)r * ncols + i]) return true; return false; };

  // ---- every column gathered into owner order and taken apart into the arrays that travel
  std::vector<OutCol> grouped((size_t)ncols);
  std::vector<DBuf> keep;                  // expanded forms (bytes per row, lengths)
  keep.reserve((size_t)ncols * 3);
  std::vector<RfPart> parts;
  struct ColParts { int data = -1, valid = -1, lens = -1; };
  std::vector<ColParts> cp((size_t)ncols);
  auto rows_part = [&](const uint8_t* p, int w) {
    RfPart part;
    part.src = p;
    part.cut.resize((size_t)N + 1);
    for (int o = 0; o <= N; o++) part.cut[o] = row_cut[o] * (uint64_t)w;
    parts.push_back(std::move(part));
    return (int)parts.size() - 1;
  };
  {
    Stage s(c, "gather_by_owner");
    for (int i = 0; i < ncols; i++) {
      hs_gather_column(c, cols[i], perm.as<uint32_t>(), n, grouped[i]);
      OutCol& g = grouped[i];
      if (has_nulls(i)) {
        keep.emplace_back(c, (size_t)std::max<int64_t>(n, 1));
        if (n > 0) LAUNCH(c, k_rf_bits_to_bytes, rf_grid(n), 256, 0, g.validity.as<uint8_t>(), n, keep.back().as<uint8_t>());
        cp[i].valid = rows_part(keep.back().as<uint8_t>(), 1);
      }
      if (hs_is_varlen(g.type)) {
        keep.emplace_back(c, (size_t)std::max<int64_t>(n, 1) * 4);
        uint32_t* lens = keep.back().as<uint32_t>();
        std::vector<uint64_t> bcut((size_t)N + 1, 0);
        const bool large = hs_is_large(g.type);
        if (n > 0) {
          if (large) LAUNCH(c, k_rf_lens<int64_t>, rf_grid(n), 256, 0, g.offsets.as<int64_t>(), n, lens);
          else LAUNCH(c, k_rf_lens<int32_t>, rf_grid(n), 256, 0, g.offsets.as<int32_t>(), n, lens);
          for (int o = 0; o <= N; o++) {   // byte offset of every owner's first row
            if (large) {
              int64_t v = 0;
              CK(cudaMemcpyAsync(&v, g.offsets.as<int64_t>() + row_cut[o], 8, cudaMemcpyDeviceToHost, c->stream));
              hs_sync(c);
              bcut[o] = (uint64_t)v;
            } else {
              int32_t v = 0;
              CK(cudaMemcpyAsync(&v, g.offsets.as<int32_t>() + row_cut[o], 4, cudaMemcpyDeviceToHost, c->stream));
              hs_sync(c);
              bcut[o] = (uint64_t)(uint32_t)v;
            }
          }
        }
        cp[i].lens = rows_part((const uint8_t*)lens, 4);
        RfPart part;
        part.src = g.data.as<uint8_t>();
        part.cut = bcut;
        parts.push_back(std::move(part));
        cp[i].data = (int)parts.size() - 1;
      } else if (g.type == HS_BOOL) {
        keep.emplace_back(c, (size_t)std::max<int64_t>(n, 1));
        if (n > 0) LAUNCH(c, k_rf_bits_to_bytes, rf_grid(n), 256, 0, g.data.as<uint8_t>(), n, keep.back().as<uint8_t>());
        cp[i].data = rows_part(keep.back().as<uint8_t>(), 1);
      } else {
        cp[i].data = rows_part(g.data.as<uint8_t>(), hs_fixed_width(g.type));
      }
    }
  }
  // ---- who sends how much to whom: row [r][o * (P + 1) + p] = bytes of part p from rank r to owner o (p = P: rows)
  const int P = (int)parts.size();
  std::vector<uint64_t> mine((size_t)N * (P + 1)), all((size_t)N * N * (P + 1));
  for (int o = 0; o < N; o++) {
    for (int p = 0; p < P; p++) mine[(size_t)o * (P + 1) + p] = parts[p].cut[o + 1] - parts[p].cut[o];
    mine[(size_t)o * (P + 1) + P] = row_cut[o + 1] - row_cut[o];
  }
  {
    DBuf d_in(c, mine.size() * 8), d_all(c, all.size() * 8);
    hs_upload(c, d_in.p, mine.data(), mine.size() * 8);
    NCK(g_nccl.allgather(d_in.p, d_all.p, mine.size(), NCCL_UINT64, comm, c->stream));
    CK(cudaMemcpyAsync(all.data(), d_all.p, all.size() * 8, cudaMemcpyDeviceToHost, c->stream));
    hs_sync(c);
  }
  auto from = [&](int r, int p) { return all[((size_t)r * N + me) * (P + 1) + p]; };   // bytes (rows for p == P) rank r sends here
  uint64_t m_rows = 0;
  for (int r = 0; r < N; r++) m_rows += from(r, P);
  for (int o = 0; o < N; o++) {   // every rank sees the whole matrix: a rank over the limit fails the build on all ranks
    uint64_t t = 0;
    for (int r = 0; r < N; r++) t += all[((size_t)r * N + o) * (P + 1) + P];
    if (t >= 0xFFFFFFF0ull) HS_THROW(HS_EUNSUPPORTED, "more than 2^32-16 rows on GPU %d", o);
  }
  // ---- the exchange: one grouped round of sends and receives; this rank's own piece is a local copy
  int64_t sent = 0;
  cudaEvent_t x0, x1;
  CK(cudaEventCreate(&x0));
  CK(cudaEventCreate(&x1));
  {
    Stage s(c, "rows_all_to_all");
    for (int p = 0; p < P; p++) {
      uint64_t tot = 0;
      for (int r = 0; r < N; r++) tot += from(r, p);
      parts[p].recv_bytes = tot;
      parts[p].recv.alloc(c, (size_t)std::max<uint64_t>(tot, 16));
    }
    CK(cudaEventRecord(x0, c->stream));
    NCK(g_nccl.group_start());
    for (int p = 0; p < P; p++) {
      uint64_t at = 0;
      for (int r = 0; r < N; r++) {
        const uint64_t in_bytes = from(r, p);
        uint8_t* dst = parts[p].recv.as<uint8_t>() + at;
        at += in_bytes;
        const uint64_t out_bytes = parts[p].cut[r + 1] - parts[p].cut[r];
        const uint8_t* src = parts[p].src + parts[p].cut[r];
        if (r == me) {
          if (in_bytes) CK(cudaMemcpyAsync(dst, src, (size_t)in_bytes, cudaMemcpyDeviceToDevice, c->stream));
          continue;
        }
        if (out_bytes) { NCK(g_nccl.send((void*)src, (size_t)out_bytes, NCCL_UINT8, r, comm, c->stream)); sent += (int64_t)out_bytes; }
        if (in_bytes) NCK(g_nccl.recv(dst, (size_t)in_bytes, NCCL_UINT8, r, comm, c->stream));
      }
    }
    NCK(g_nccl.group_end());
    CK(cudaEventRecord(x1, c->stream));
  }
  // ---- the received arrays as columns again
  const int64_t m = (int64_t)m_rows;
  std::vector<DevCol> rcv((size_t)ncols);
  {
    Stage s(c, "rows_to_columns");
    for (int i = 0; i < ncols; i++) {
      DevCol& d = rcv[i];
      d.type = cols[i].type;
      d.length = m;
      if (cp[i].valid >= 0) {
        DBuf bits(c, (size_t)(m + 7) / 8 + 16), zeros(c, 8);
        zeros.zero();
        if (m > 0) LAUNCH(c, k_rf_bytes_to_bits, rf_grid((m + 7) / 8), 256, 0, parts[cp[i].valid].recv.as<uint8_t>(), m,
                          bits.as<uint8_t>(), zeros.as<unsigned long long>());
        unsigned long long z = 0;
        CK(cudaMemcpyAsync(&z, zeros.p, 8, cudaMemcpyDeviceToHost, c->stream));
        hs_sync(c);
        d.null_count = (int64_t)z;
        if (z) d.validity = bits.as<uint8_t>();
        d.owned.push_back(std::move(bits));
      }
      RfPart& dp = parts[cp[i].data];
      if (hs_is_varlen(d.type)) {
        OutCol tmp;
        tmp.type = d.type;
        const int64_t bytes = hs_offsets_from_lens(c, parts[cp[i].lens].recv.as<uint32_t>(), m, d.type, tmp);
        if ((uint64_t)bytes != dp.recv_bytes) HS_THROW(HS_ENCCL, "string bytes received do not match their lengths");
        d.type = tmp.type;                          // the large type when the received bytes pass 2 GiB
        d.offsets = tmp.offsets.as<int32_t>();
        d.owned.push_back(std::move(tmp.offsets));
        d.data = dp.recv.p;
        d.data_bytes = bytes;
      } else if (d.type == HS_BOOL) {
        DBuf bits(c, (size_t)(m + 7) / 8 + 16);
        if (m > 0) LAUNCH(c, k_rf_bytes_to_bits, rf_grid((m + 7) / 8), 256, 0, dp.recv.as<uint8_t>(), m, bits.as<uint8_t>(),
                          (unsigned long long*)nullptr);
        d.data = bits.p;
        d.data_bytes = (m + 7) / 8;
        d.owned.push_back(std::move(bits));
      } else {
        d.data = dp.recv.p;
        d.data_bytes = m * hs_fixed_width(d.type);
      }
    }
  }
  // the source-side copies are no longer needed; the received pieces stay alive inside `parts` until the build is done
  grouped.clear();
  keep.clear();
  perm.release();
  int blo = 0, bhi = 0;
  for (int b = 0; b < nb; b++) if (owner_of(b, N, nb) == me) { blo = b; break; }
  for (bhi = blo; bhi < nb && owner_of(bhi, N, nb) == me; bhi++) {}
  hs_build_core(c, rcv, n_indexed, nb, blo, bhi - blo, out, boff, nullptr);
  CK(cudaEventSynchronize(x1));
  float ms = 0.f;
  cudaEventElapsedTime(&ms, x0, x1);
  cudaEventDestroy(x0);
  cudaEventDestroy(x1);
  c->last_xchg_bytes = sent;
  c->last_xchg_ms = ms;
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
  cudaStreamSynchronize(ctx->stream);
  p2p_destroy(ctx);
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
  const int N = ctx->nranks, nb = num_buckets;
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
  std::vector<OutCol> ocols;
  std::vector<int64_t> boff;
  if (N == 1) {
    hs_build_core(ctx, dcols, n_indexed, nb, 0, nb, ocols, boff, nullptr);
    Stage s(ctx, "d2h");
    *out = hs_finish_table(ctx, ocols, dcols[0].length, nb, &boff);
    return HS_OK;
  }
  NcclExchange xchg(ctx, nb);
  hs_build_core(ctx, dcols, n_indexed, nb, 0, nb, ocols, boff, &xchg);
  if (xchg.rows_first) {
    // string columns or a key wider than one 64-bit chunk (every rank decides the same from the shared statistics):
    // rows change hands first, then every GPU builds its own buckets
    ocols.clear();
    boff.clear();
    dist_build_rows_first(ctx, dcols, n_indexed, nb, ocols, boff);
  }
  Stage s(ctx, "d2h");
  *out = hs_finish_table(ctx, ocols, boff[nb], nb, &boff);
  API_END(ctx)
}

}  // extern "C"
