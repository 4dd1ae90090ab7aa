// hs_join.cu — K5 + K6: bucket-local inner sort-merge join of two indexes, no exchange.
// Executes the plan JoinIndexRule.apply produces (HS/index/rules/JoinIndexRule.scala:57-90) over the
// bucketed relations of RuleUtils.transformPlanToUseIndexOnlyScan (HS/index/rules/RuleUtils.scala:264-292):
// bucket i of the left index meets bucket i of the right index; equal keys produce the cross product;
// NULL keys never match.
// Both sides are sorted per bucket, so neighbouring left rows probe neighbouring right rows: each left
// row binary-searches its bucket of the right run (the probes of a warp share cache lines), a scan turns
// match counts into output offsets, and the pairs are expanded and gathered.
#include "hs_keys.cuh"
#include "hs_kernels.h"

struct JoinSide {
  KeySpec ks;
  const uint32_t* off;  // device, nb + 1
};

__device__ __forceinline__ bool key_has_null(const KeySpec& ks, int64_t i) {
#pragma unroll
  for (int k = 0; k < HS_MAX_KEYS; k++)
    if (k < ks.nkeys && ks.k[k].has_nulls && !kvalid(ks.k[k].validity, i)) return true;
  return false;
}
// keys of left row i vs right row j (no NULLs on either side)
__device__ __forceinline__ int cmp_keys(const KeySpec& a, int64_t i, const KeySpec& b, int64_t j) {
#pragma unroll
  for (int k = 0; k < HS_MAX_KEYS; k++) {
    if (k < a.nkeys) {
      int c = cmp_cells(a.k[k], i, b.k[k], j);
      if (c) return c;
    }
  }
  return 0;
}
// right rows with NULL keys sort first inside a bucket: skip them by treating NULL as smaller
__device__ __forceinline__ int cmp_left_right(const KeySpec& a, int64_t i, const KeySpec& b, int64_t j) {
#pragma unroll
  for (int k = 0; k < HS_MAX_KEYS; k++) {
    if (k < a.nkeys) {
      if (b.k[k].has_nulls && !kvalid(b.k[k].validity, j)) return 1;  // left (non-null) > right NULL
      int c = cmp_cells(a.k[k], i, b.k[k], j);
      if (c) return c;
    }
  }
  return 0;
}

template <bool FAST_I64>
__global__ void __launch_bounds__(256) k_join_probe(JoinSide L, JoinSide R, int64_t nl, int32_t nb,
                                                    uint32_t* __restrict__ lb_out, uint32_t* __restrict__ cnt_out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nl) return;
  // bucket of left row i: largest b with L.off[b] <= i
  int lo = 0, hi = nb;
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (L.off[mid] <= (uint32_t)i) lo = mid; else hi = mid;
  }
  uint32_t r0 = R.off[lo], r1 = R.off[lo + 1];
  uint32_t lb, ub;
  if (FAST_I64) {
    const int64_t* lk = (const int64_t*)L.ks.k[0].data;
    const int64_t* rk = (const int64_t*)R.ks.k[0].data;
    int64_t key = lk[i];
    uint32_t a = r0, b = r1;
    while (a < b) { uint32_t m = (a + b) >> 1; if (rk[m] < key) a = m + 1; else b = m; }
    lb = a;
    b = r1;
    while (a < b) { uint32_t m = (a + b) >> 1; if (rk[m] <= key) a = m + 1; else b = m; }
    ub = a;
  } else {
    if (key_has_null(L.ks, i)) { lb_out[i] = r0; cnt_out[i] = 0; return; }
    uint32_t a = r0, b = r1;
    while (a < b) { uint32_t m = (a + b) >> 1; if (cmp_left_right(L.ks, i, R.ks, m) > 0) a = m + 1; else b = m; }
    lb = a;
    b = r1;
    while (a < b) { uint32_t m = (a + b) >> 1; if (cmp_left_right(L.ks, i, R.ks, m) >= 0) a = m + 1; else b = m; }
    ub = a;
  }
  lb_out[i] = lb;
  cnt_out[i] = ub - lb;
}

__global__ void __launch_bounds__(256) k_join_expand(const uint32_t* __restrict__ lb, const uint32_t* __restrict__ cnt,
                                                     const uint32_t* __restrict__ off, int64_t nl,
                                                     uint32_t* __restrict__ li, uint32_t* __restrict__ ri) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nl) return;
  uint32_t c = cnt[i], o = off[i], b = lb[i];
  for (uint32_t t = 0; t < c; t++) { li[o + t] = (uint32_t)i; ri[o + t] = b + t; }
}

__global__ void __launch_bounds__(256) k_join_bucket_offsets(const uint32_t* __restrict__ loff, const uint32_t* __restrict__ off,
                                                             const uint64_t* total, int64_t nl, int32_t nb,
                                                             uint64_t* __restrict__ out) {
  int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b > nb) return;
  uint32_t r = b < nb ? loff[b] : (uint32_t)nl;
  out[b] = r < nl ? off[r] : *total;
}

static void check_view(const hs_index_view* v) {
  if (!v || !v->cols || v->ncols <= 0 || v->n_indexed <= 0 || v->n_indexed > v->ncols || v->num_buckets <= 0 ||
      !v->bucket_offsets)
    HS_THROW(HS_EINVAL, "malformed index view");
  if (v->n_indexed > HS_MAX_KEYS) HS_THROW(HS_EUNSUPPORTED, "more than %d indexed columns", HS_MAX_KEYS);
}

void hs_offsets_to_device(hs_ctx* c, const int64_t* boff, int32_t nb, int64_t nrows, DBuf& out) {
  std::vector<uint32_t> h((size_t)nb + 1);
  for (int b = 0; b <= nb; b++) {
    if (boff[b] < 0 || boff[b] > nrows || (b && boff[b] < boff[b - 1])) HS_THROW(HS_EINVAL, "bad bucket_offsets");
    h[b] = (uint32_t)boff[b];
  }
  if (boff[0] != 0 || boff[nb] != nrows) HS_THROW(HS_EINVAL, "bucket_offsets do not cover the rows");
  out.alloc(c, ((size_t)nb + 1) * 4);
  CK(cudaMemcpyAsync(out.p, h.data(), ((size_t)nb + 1) * 4, cudaMemcpyHostToDevice, c->stream));
  hs_sync(c);  // h goes out of scope
}

extern "C" int32_t hs_join_bucketed(hs_ctx* ctx, const hs_index_view* left, const hs_index_view* right,
                                    int32_t join_type, const int32_t* left_out, int32_t n_left_out,
                                    const int32_t* right_out, int32_t n_right_out, hs_table** out) {
  API_BEGIN(ctx)
  if (!out) HS_THROW(HS_EINVAL, "null out");
  *out = nullptr;
  check_view(left);
  check_view(right);
  if (join_type != HS_JOIN_INNER) HS_THROW(HS_EUNSUPPORTED, "only inner joins are implemented");
  if (left->num_buckets != right->num_buckets)
    HS_THROW(HS_EUNSUPPORTED, "indexes with different bucket counts (Spark would re-shuffle one side)");
  if (left->n_indexed != right->n_indexed) HS_THROW(HS_EINVAL, "indexes differ in number of indexed columns");
  if (n_left_out < 0 || n_right_out < 0 || n_left_out + n_right_out <= 0) HS_THROW(HS_EINVAL, "no output columns");
  for (int i = 0; i < n_left_out; i++)
    if (left_out[i] < 0 || left_out[i] >= left->ncols) HS_THROW(HS_EINVAL, "left output column out of range");
  for (int i = 0; i < n_right_out; i++)
    if (right_out[i] < 0 || right_out[i] >= right->ncols) HS_THROW(HS_EINVAL, "right output column out of range");
  hs_stages_reset(ctx);
  const int nk = left->n_indexed, nb = left->num_buckets;
  std::vector<DevCol> lc, rc;
  {
    Stage s(ctx, "h2d");
    hs_stage_columns(ctx, left->cols, left->ncols, lc);
    hs_stage_columns(ctx, right->cols, right->ncols, rc);
  }
  for (int k = 0; k < nk; k++)
    if (lc[k].type != rc[k].type) HS_THROW(HS_EUNSUPPORTED, "join key %d differs in type between the indexes", k);
  const int64_t nl = lc[0].length, nr = rc[0].length;
  DBuf loff, roff;
  hs_offsets_to_device(ctx, left->bucket_offsets, nb, nl, loff);
  hs_offsets_to_device(ctx, right->bucket_offsets, nb, nr, roff);
  JoinSide L{hs_make_keyspec(lc, nk), loff.as<uint32_t>()};
  JoinSide R{hs_make_keyspec(rc, nk), roff.as<uint32_t>()};

  DBuf lb(ctx, (size_t)nl * 4), cnt(ctx, (size_t)nl * 4), off(ctx, (size_t)nl * 4), total(ctx, 8);
  uint64_t h_total = 0;
  {
    Stage s(ctx, "join_probe");
    bool fast = nk == 1 && lc[0].type == HS_INT64 && !lc[0].validity && !rc[0].validity;
    if (nl > 0) {
      unsigned g = (unsigned)((nl + 255) / 256);
      if (fast) LAUNCH(ctx, k_join_probe<true>, g, 256, 0, L, R, nl, nb, lb.as<uint32_t>(), cnt.as<uint32_t>());
      else LAUNCH(ctx, k_join_probe<false>, g, 256, 0, L, R, nl, nb, lb.as<uint32_t>(), cnt.as<uint32_t>());
    }
    hs_exclusive_scan_u32(ctx, cnt.as<uint32_t>(), off.as<uint32_t>(), nl, total.as<uint64_t>());
    CK(cudaMemcpyAsync(&h_total, total.p, 8, cudaMemcpyDeviceToHost, ctx->stream));
    hs_sync(ctx);
  }
  if (h_total >= 0xFFFFFFF0ull) HS_THROW(HS_EUNSUPPORTED, "join output exceeds 2^32-16 rows in one call");
  const int64_t m = (int64_t)h_total;
  DBuf li(ctx, (size_t)m * 4), ri(ctx, (size_t)m * 4), d_ooff(ctx, ((size_t)nb + 1) * 8);
  std::vector<int64_t> ooff((size_t)nb + 1, 0);
  {
    Stage s(ctx, "join_expand");
    if (nl > 0)
      LAUNCH(ctx, k_join_expand, (unsigned)((nl + 255) / 256), 256, 0, lb.as<uint32_t>(), cnt.as<uint32_t>(),
             off.as<uint32_t>(), nl, li.as<uint32_t>(), ri.as<uint32_t>());
    LAUNCH(ctx, k_join_bucket_offsets, (unsigned)((nb + 256) / 256), 256, 0, loff.as<uint32_t>(), off.as<uint32_t>(),
           total.as<uint64_t>(), nl, nb, d_ooff.as<uint64_t>());
    CK(cudaMemcpyAsync(ooff.data(), d_ooff.p, ((size_t)nb + 1) * 8, cudaMemcpyDeviceToHost, ctx->stream));
    hs_sync(ctx);
  }
  lb.release(); cnt.release(); off.release();
  std::vector<OutCol> ocols((size_t)(n_left_out + n_right_out));
  {
    Stage s(ctx, "join_gather");
    for (int i = 0; i < n_left_out; i++) hs_gather_column(ctx, lc[left_out[i]], li.as<uint32_t>(), m, ocols[i]);
    for (int i = 0; i < n_right_out; i++)
      hs_gather_column(ctx, rc[right_out[i]], ri.as<uint32_t>(), m, ocols[n_left_out + i]);
  }
  Stage s(ctx, "d2h");
  *out = hs_finish_table(ctx, ocols, m, nb, &ooff);
  API_END(ctx)
}
