// hs_kernels.h — internal entry points between the translation units of libhsgpu.
#pragma once
#include "hs_common.cuh"

struct KeySpec;
struct KeyLayout;

KeySpec hs_make_keyspec(const std::vector<DevCol>& cols, int nkeys);
void hs_hash_to_i32(hs_ctx* c, const KeySpec& ks, int64_t n, int32_t nb, int32_t* d_out);
void hs_hash_bucket_hist(hs_ctx* c, const KeySpec& ks, const std::vector<DevCol>& cols, int64_t n, int32_t nb,
                         int32_t blo, int32_t nbl, uint16_t* d_bucket, uint32_t* d_hist, uint64_t* d_minmax);
// the build over staged device columns; rows must hash into buckets [blo, blo + nbl)
// Multi-GPU hook of the build (implemented over NCCL in hs_dist.cu).
struct BuildExchange {
  int blo = 0, nbl = 0;   // the bucket range this rank owns
  virtual ~BuildExchange() {}
  // in: this rank's histogram over all buckets, its key min/max and which columns hold NULLs here;
  // out: minmax and has_nulls replaced by their global values
  virtual void share_stats(hs_ctx* c, const std::vector<uint32_t>& local_hist, uint64_t* minmax,
                           std::vector<uint8_t>& has_nulls) = 0;
  // in: bufs[i] holds the local rows bucket-contiguously (local_off, nb+1 entries). Every bucket run goes to the
  // bucket's owner; bufs[i] is replaced by a buffer holding this rank's buckets, each contiguous. Returns the
  // rows received; seg_off (nbl+1 entries) are the local offsets of the owned buckets.
  virtual int64_t exchange(hs_ctx* c, std::vector<DBuf*>& bufs, const std::vector<int>& esize,
                           const std::vector<uint32_t>& local_off, std::vector<uint32_t>& seg_off) = 0;
  // Fused path: the bucket scatter kernel itself writes every row into the owner's receive buffer over
  // peer-mapped memory, so there is no separate all-to-all. try_fused sets up the receive buffers on all ranks
  // (collective; false on every rank if any rank cannot map its peers). scatter_exchange then takes the rows
  // in input order; recv[i] is this rank's receive buffer of column i (persistent, owned by the exchange).
  // set by the build before try_fused when the sort key is one non-null INT64 column: that moved column, the
  // mapped (sign-flipped) minimum over all ranks and the span max - min. A span below 2^32 lets the key travel as
  // 4 bytes (value - minimum) and be widened again by the receiver's bucket scatter.
  int key_col = -1;
  uint64_t key_min = 0, key_span = ~0ull;
  // set by the build (on every rank alike, from the shared statistics) when the rows cannot take the fused path:
  // the build returns without having moved anything and the caller exchanges whole rows first
  bool rows_first = false;
  virtual bool try_fused(hs_ctx* c, const std::vector<int>& esize) { return false; }
  virtual int64_t scatter_exchange(hs_ctx* c, const void* const* srcs, const std::vector<int>& esize,
                                   const uint16_t* d_bucket, int64_t n_local, std::vector<void*>& recv,
                                   std::vector<uint32_t>& seg_off) { return -1; }
  // Pipelined delivery (grouped spans): after scatter_exchange the owned buckets become ready group by group
  // while later groups are still travelling. deliver_group(g) makes the context's stream wait for group g and
  // puts its rows bucket-contiguously into recv[]; the build then sorts local buckets [lb0, lb1) and asks for
  // the next group. ngroups = 0: everything was ready when scatter_exchange returned.
  int ngroups = 0;
  virtual void deliver_group(hs_ctx* c, int g, int* lb0, int* lb1) {}
};
void hs_build_core(hs_ctx* c, std::vector<DevCol>& cols, int n_indexed, int32_t nb, int32_t blo, int32_t nbl,
                   std::vector<OutCol>& out, std::vector<int64_t>& boff, BuildExchange* xchg);
void hs_offsets_to_device(hs_ctx* c, const int64_t* boff, int32_t nb, int64_t nrows, DBuf& out);
bool hs_pack_key_tuples(hs_ctx* ctx, const KeySpec& a, int64_t na, const KeySpec& b, int64_t nrows_b, int nk, DBuf& pa,
                        DBuf& pb);
// two sorted runs per bucket -> one (merge path); false if the shape is not covered (the caller re-sorts instead)
bool hs_merge_two_runs(hs_ctx* c, std::vector<DevCol>& a, const int64_t* aoff, std::vector<DevCol>& b, const int64_t* boff_b,
                       int nk, int32_t nb, std::vector<OutCol>& out, std::vector<int64_t>& out_off);

// ---- the row mover / segmented sort (hs_sort.cu) ----
struct MoveSet {           // the columns that travel together; all fixed width
  int32_t n;
  int32_t esize[HS_MAX_MOVE];
  void* a[HS_MAX_MOVE];    // buffer A (holds the result at the end)
  void* b[HS_MAX_MOVE];    // buffer B (ping-pong)
  void* s0[HS_MAX_MOVE];   // optional: where the rows are when the sort starts (else they are in A);
                           // used by the multi-GPU build, whose receive buffers are persistent peer-mapped memory
};
// how kernels read the 64-bit sort key of a row out of the moved columns
struct KeySrc {
  int32_t col;             // index into MoveSet
  int32_t mode;            // 0: u64 as is, 1: int64 (sign flip, minus min), 2: int32 (sign extend, flip, minus min)
  uint64_t min;
  uint64_t span;           // max - min of the key over all rows when known, else 0 (lets the first MSD level use all its bins)
};
#define KS_U64 0
#define KS_I64 1
#define KS_I32 2

// Scatter rows from the caller's column buffers (src[i], element sizes as in ms) into bucket-contiguous
// order in ms.a, given u16 bucket ids and host-known bucket offsets.
// widen_col >= 0: that column arrives as u32 (value - widen_min of the sign-flipped key) and is written as the
// original int64 again
void hs_bucket_scatter(hs_ctx* c, const void* const* src, const MoveSet& ms, const uint16_t* d_bucket, int64_t n,
                       int32_t nb, const uint32_t* d_bucket_off, int32_t bucket_base = 0, int widen_col = -1,
                       uint64_t widen_min = 0, uint32_t* d_cursor = nullptr);
void hs_scatter_by_bins(hs_ctx* c, const void* const* src, void* const* dst, const int32_t* esize, int ncols,
                        const uint16_t* d_bucket, int64_t n, const uint16_t* d_lut, int32_t nbins, const uint32_t* d_bin_off,
                        int narrow_col = -1, uint64_t narrow_min = 0);
void hs_bucket_scatter_peers(hs_ctx* c, const void* const* src, const int32_t* esize, int ncols, void* const* pdst,
                             int nranks, const uint16_t* d_bucket, int64_t n, int32_t nb, const uint32_t* d_run_base,
                             bool by_owner = false, int narrow_col = -1, uint64_t narrow_min = 0);
// Sort every segment [seg_off[s], seg_off[s+1]) of ms.a by the key; result in ms.a. key_bits = significant
// low bits of the key. d_seg_off: device u32[nseg+1].
// carry: several calls over disjoint segments of the same buffers (the grouped multi-GPU build) run only the
// first MSD level each; children still larger than a leaf collect in the carry and hs_sort_carry_finish takes
// them through the remaining levels together (a handful of leftovers would otherwise cost every call a level).
struct SortCarry;
SortCarry* hs_sort_carry_new(hs_ctx* c, int64_t n);
void hs_sort_carry_finish(hs_ctx* c, MoveSet& ms, const KeySrc& key, SortCarry* k);
void hs_sort_carry_free(SortCarry* k);
void hs_segmented_sort(hs_ctx* c, MoveSet& ms, const KeySrc& key, int key_bits, const uint32_t* d_seg_off,
                       int64_t nseg, int64_t n, const uint32_t* h_seg_off, SortCarry* carry = nullptr);
// Same, for an explicit device list of segments (start,len), all with the same key_bits.
void hs_segmented_sort_list(hs_ctx* c, MoveSet& ms, const KeySrc& key, int key_bits, const uint32_t* d_start,
                            const uint32_t* d_len, int64_t nseg, int64_t n);
