/*
 * hs_oracle.c — CPU restatement of the covering-index hot path. TEST INFRASTRUCTURE ONLY.
 *
 * This file is the parity oracle for libhsgpu. Nothing on the product path may
 * call it: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs load it, and only as the checker or the timed CPU arm.
 *
 * What it restates. The reference (pirz/hyperspace-1) does no row arithmetic of
 * its own; the arithmetic lives in its `provided` dependency Apache Spark 2.4.2
 * (build.sbt:19,29-33 — org.apache.spark:spark-sql/core/catalyst 2.4.2), which is
 * NOT vendored under /root/reference. The functions below restate Spark 2.4.2's
 * published behaviour at the reference's own call sites:
 *   - hso_hash_rows / hso_bucket_ids : Murmur3Hash(cols, 42) and
 *     Pmod(hash, n) = HashPartitioning.partitionIdExpression, reached from
 *     df.repartition(numBuckets, cols) at
 *     src/main/scala/com/microsoft/hyperspace/actions/CreateActionBase.scala:129-130
 *   - hso_build_index : that repartition followed by saveWithBuckets with
 *     BucketSpec(n, cols, cols) at
 *     src/main/scala/com/microsoft/hyperspace/index/DataFrameWriterExtensions.scala:49-67
 *     (per-bucket sort on the indexed columns, ASC NULLS FIRST)
 *   - hso_sort_segments : the SortExec Spark plans per side when a bucket has
 *     more than one file (E2EHyperspaceRulesTest.scala:474-478) and the
 *     re-bucketing of OptimizeAction.op (actions/OptimizeAction.scala:85-99)
 *   - hso_join : the SortMergeJoinExec under the plan of
 *     index/rules/JoinIndexRule.scala:57-90 (inner, NULL keys never match)
 *   - hso_filter : the FilterExec under index/rules/FilterIndexRule.scala:43-85
 * Semantics: SURVEY.md Appendix A.1-A.5.
 *
 * Parity pin: tests/test_oracle_golden.py checks this file against the literal
 * known answer in the reference (src/test/scala/com/microsoft/hyperspace/index/
 * BucketUnionTest.scala:116-122: int ids 2,3 -> partitions 4,1 of 10), Spark's
 * documented answers hash(1) = -559580957 and hash('Spark', array(123), 2) =
 * -1321691492 (the array element is an int column here), and the quick-start
 * fixture of examples/scala/src/main/scala/App.scala:33-98. For int64, string
 * and multi-column hashing the reference itself holds no literal vectors.
 *
 * Build: oracle/build.sh (gcc -O2 -fopenmp -shared).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "../include/hsgpu.h"

/* ---- A.1 Murmur3_x86_32 as Spark uses it ------------------------------- */
static inline uint32_t rotl32(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }
static inline uint32_t mix_k1(uint32_t k) {
  k *= 0xcc9e2d51u; k = rotl32(k, 15); k *= 0x1b873593u; return k;
}
static inline uint32_t mix_h1(uint32_t h, uint32_t k) {
  h ^= k; h = rotl32(h, 13); h = h * 5u + 0xe6546b64u; return h;
}
static inline uint32_t fmix(uint32_t h, uint32_t len) {
  h ^= len; h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16;
  return h;
}
static inline uint32_t hash_int(uint32_t v, uint32_t seed) {
  return fmix(mix_h1(seed, mix_k1(v)), 4);
}
static inline uint32_t hash_long(uint64_t v, uint32_t seed) {
  uint32_t lo = (uint32_t)v, hi = (uint32_t)(v >> 32);
  uint32_t h = mix_h1(seed, mix_k1(lo));
  h = mix_h1(h, mix_k1(hi));
  return fmix(h, 8);
}
/* Spark's hashUnsafeBytes: whole little-endian words, then ONE FULL ROUND per
 * trailing byte, the byte sign-extended (not the standard murmur3 tail). */
static inline uint32_t hash_bytes(const uint8_t* p, int32_t n, uint32_t seed) {
  uint32_t h = seed;
  int32_t nw = n & ~3;
  for (int32_t i = 0; i < nw; i += 4) {
    uint32_t w; memcpy(&w, p + i, 4);
    h = mix_h1(h, mix_k1(w));
  }
  for (int32_t i = nw; i < n; i++) {
    int32_t b = (int8_t)p[i];
    h = mix_h1(h, mix_k1((uint32_t)b));
  }
  return fmix(h, (uint32_t)n);
}

static inline int is_valid(const hs_column* c, int64_t i) {
  return c->validity == NULL || ((c->validity[i >> 3] >> (i & 7)) & 1);
}

/* A.2: one column's contribution; NULL leaves the running hash unchanged. */
static inline uint32_t hash_value(const hs_column* c, int64_t i, uint32_t h) {
  if (!is_valid(c, i)) return h;
  switch (c->type) {
    case HS_BOOL: {
      const uint8_t* d = (const uint8_t*)c->data;
      return hash_int((d[i >> 3] >> (i & 7)) & 1, h);
    }
    case HS_INT8:  return hash_int((uint32_t)(int32_t)((const int8_t*)c->data)[i], h);
    case HS_INT16: return hash_int((uint32_t)(int32_t)((const int16_t*)c->data)[i], h);
    case HS_INT32: return hash_int(((const uint32_t*)c->data)[i], h);
    case HS_INT64: return hash_long(((const uint64_t*)c->data)[i], h);
    case HS_FLOAT32: {  /* floatToIntBits: every NaN -> 0x7fc00000; -0.0f keeps its sign bit */
      float f = ((const float*)c->data)[i];
      uint32_t b; memcpy(&b, &f, 4);
      if (f != f) b = 0x7fc00000u;
      return hash_int(b, h);
    }
    case HS_FLOAT64: {  /* doubleToLongBits: every NaN -> 0x7ff8000000000000 */
      double f = ((const double*)c->data)[i];
      uint64_t b; memcpy(&b, &f, 8);
      if (f != f) b = 0x7ff8000000000000ull;
      return hash_long(b, h);
    }
    case HS_UTF8:
    case HS_BINARY: {
      int32_t s = c->offsets[i], e = c->offsets[i + 1];
      return hash_bytes((const uint8_t*)c->data + s, e - s, h);
    }
    default: return h;
  }
}

static int type_ok(int32_t t) { return t >= HS_BOOL && t <= HS_BINARY; }

int32_t hso_hash_rows(const hs_column* keys, int32_t nkeys, int32_t* out) {
  if (nkeys <= 0) return HS_EINVAL;
  for (int k = 0; k < nkeys; k++) if (!type_ok(keys[k].type)) return HS_EUNSUPPORTED;
  int64_t n = keys[0].length;
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; i++) {
    uint32_t h = 42;
    for (int k = 0; k < nkeys; k++) h = hash_value(&keys[k], i, h);
    out[i] = (int32_t)h;
  }
  return HS_OK;
}

static inline int32_t pmod(int32_t h, int32_t n) {
  int32_t r = h % n;               /* Java remainder: sign of the dividend */
  if (r < 0) r = (r + n) % n;
  return r;
}

int32_t hso_bucket_ids(const hs_column* keys, int32_t nkeys, int32_t nb, int32_t* out) {
  if (nb <= 0) return HS_EINVAL;
  int32_t st = hso_hash_rows(keys, nkeys, out);
  if (st) return st;
  int64_t n = keys[0].length;
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; i++) out[i] = pmod(out[i], nb);
  return HS_OK;
}

/* ---- A.3 ordering: ASC NULLS FIRST; NaN greatest and equal to itself,
 * -0.0 == 0.0; strings unsigned byte-wise, shorter first on a common prefix. */
static inline int cmp_f64(double a, double b) {
  int an = a != a, bn = b != b;
  if (an || bn) return an - bn;
  return (a > b) - (a < b);
}
static inline int cmp_bytes(const uint8_t* a, int32_t la, const uint8_t* b, int32_t lb) {
  int32_t m = la < lb ? la : lb;
  int c = memcmp(a, b, (size_t)m);
  if (c) return c < 0 ? -1 : 1;
  return (la > lb) - (la < lb);
}
static inline int cmp_value(const hs_column* a, int64_t i, const hs_column* b, int64_t j) {
  int va = is_valid(a, i), vb = is_valid(b, j);
  if (!va || !vb) return va - vb;  /* null < non-null, null == null for ordering */
  switch (a->type) {
    case HS_BOOL: {
      int x = (((const uint8_t*)a->data)[i >> 3] >> (i & 7)) & 1;
      int y = (((const uint8_t*)b->data)[j >> 3] >> (j & 7)) & 1;
      return x - y;
    }
    case HS_INT8:  { int x = ((const int8_t*)a->data)[i],  y = ((const int8_t*)b->data)[j];  return (x > y) - (x < y); }
    case HS_INT16: { int x = ((const int16_t*)a->data)[i], y = ((const int16_t*)b->data)[j]; return (x > y) - (x < y); }
    case HS_INT32: { int32_t x = ((const int32_t*)a->data)[i], y = ((const int32_t*)b->data)[j]; return (x > y) - (x < y); }
    case HS_INT64: { int64_t x = ((const int64_t*)a->data)[i], y = ((const int64_t*)b->data)[j]; return (x > y) - (x < y); }
    case HS_FLOAT32: return cmp_f64(((const float*)a->data)[i], ((const float*)b->data)[j]);
    case HS_FLOAT64: return cmp_f64(((const double*)a->data)[i], ((const double*)b->data)[j]);
    case HS_UTF8:
    case HS_BINARY:
      return cmp_bytes((const uint8_t*)a->data + a->offsets[i], a->offsets[i + 1] - a->offsets[i],
                       (const uint8_t*)b->data + b->offsets[j], b->offsets[j + 1] - b->offsets[j]);
    default: return 0;
  }
}
static inline int cmp_rows(const hs_column* a, int64_t i, const hs_column* b, int64_t j, int nk) {
  for (int k = 0; k < nk; k++) {
    int c = cmp_value(&a[k], i, &b[k], j);
    if (c) return c;
  }
  return 0;
}

/* stable bottom-up merge sort of row ids under cmp_rows */
static void merge_sort_ids(int64_t* ids, int64_t* tmp, int64_t n, const hs_column* keys, int nk) {
  /* insertion sort runs of 16, then merge */
  const int64_t RUN = 16;
  for (int64_t s = 0; s < n; s += RUN) {
    int64_t e = s + RUN < n ? s + RUN : n;
    for (int64_t i = s + 1; i < e; i++) {
      int64_t v = ids[i], j = i;
      while (j > s && cmp_rows(keys, ids[j - 1], keys, v, nk) > 0) { ids[j] = ids[j - 1]; j--; }
      ids[j] = v;
    }
  }
  int64_t* src = ids; int64_t* dst = tmp;
  for (int64_t w = RUN; w < n; w *= 2) {
    for (int64_t lo = 0; lo < n; lo += 2 * w) {
      int64_t mid = lo + w < n ? lo + w : n, hi = lo + 2 * w < n ? lo + 2 * w : n;
      int64_t a = lo, b = mid, o = lo;
      while (a < mid && b < hi) {
        if (cmp_rows(keys, src[b], keys, src[a], nk) < 0) dst[o++] = src[b++];
        else dst[o++] = src[a++];
      }
      while (a < mid) dst[o++] = src[a++];
      while (b < hi) dst[o++] = src[b++];
    }
    int64_t* t = src; src = dst; dst = t;
  }
  if (src != ids) memcpy(ids, src, (size_t)n * sizeof(int64_t));
}

/* fast path: a single non-null INT32/INT64 key -> stable LSD radix on (key, id) pairs
 * (Spark's UnsafeExternalRowSorter also radix-sorts an 8-byte prefix). Same result as
 * the merge sort because both are stable. */
static void radix_sort_ids(int64_t* ids, int64_t* tmp, int64_t n, const hs_column* key) {
  uint64_t* k0 = (uint64_t*)malloc((size_t)n * 8);
  uint64_t* k1 = (uint64_t*)malloc((size_t)n * 8);
  uint64_t lo = ~0ull, hi = 0;
  for (int64_t i = 0; i < n; i++) {
    uint64_t v = key->type == HS_INT32
      ? (uint64_t)(((const uint32_t*)key->data)[ids[i]] ^ 0x80000000u)
      : ((const uint64_t*)key->data)[ids[i]] ^ 0x8000000000000000ull;
    k0[i] = v; if (v < lo) lo = v; if (v > hi) hi = v;
  }
  uint64_t range = hi - lo;
  int64_t* is = ids; int64_t* id = tmp; uint64_t* ks = k0; uint64_t* kd = k1;
  for (int sh = 0; sh < 64 && (range >> sh) != 0; sh += 8) {
    int64_t cnt[257]; memset(cnt, 0, sizeof cnt);
    for (int64_t i = 0; i < n; i++) cnt[(((ks[i] - lo) >> sh) & 255) + 1]++;
    for (int d = 0; d < 256; d++) cnt[d + 1] += cnt[d];
    for (int64_t i = 0; i < n; i++) {
      int64_t p = cnt[((ks[i] - lo) >> sh) & 255]++;
      kd[p] = ks[i]; id[p] = is[i];
    }
    uint64_t* tk = ks; ks = kd; kd = tk; int64_t* ti = is; is = id; id = ti;
  }
  if (is != ids) memcpy(ids, is, (size_t)n * sizeof(int64_t));
  free(k0); free(k1);
}

/* Sort the row ids of every segment [seg[s], seg[s+1]) of perm (which holds row ids
 * into cols) by the first nk columns. use_radix != 0 allows the fast path. */
int32_t hso_sort_segments(const hs_column* cols, int32_t nk, const int64_t* seg, int64_t nseg,
                          int64_t* perm, int32_t use_radix) {
  for (int k = 0; k < nk; k++) if (!type_ok(cols[k].type)) return HS_EUNSUPPORTED;
  int64_t n = seg[nseg];
  int64_t* tmp = (int64_t*)malloc((size_t)(n > 0 ? n : 1) * sizeof(int64_t));
  if (!tmp) return HS_ENOMEM;
  int fast = use_radix && nk == 1 && cols[0].null_count == 0 &&
             (cols[0].type == HS_INT32 || cols[0].type == HS_INT64);
#pragma omp parallel for schedule(dynamic, 1)
  for (int64_t s = 0; s < nseg; s++) {
    int64_t a = seg[s], len = seg[s + 1] - seg[s];
    if (len < 2) continue;
    if (fast && len > 64) radix_sort_ids(perm + a, tmp + a, len, &cols[0]);
    else merge_sort_ids(perm + a, tmp + a, len, cols, nk);
  }
  free(tmp);
  return HS_OK;
}

/* The build: perm[i] = input row of output row i; rows grouped by bucket (stable),
 * every bucket sorted stably on the first n_indexed columns. */
int32_t hso_build_index(const hs_column* cols, int32_t ncols, int32_t n_indexed, int32_t nb,
                        int64_t* out_perm, int64_t* out_bucket_offsets, int32_t use_radix) {
  if (ncols <= 0 || n_indexed <= 0 || n_indexed > ncols || nb <= 0) return HS_EINVAL;
  int64_t n = cols[0].length;
  int32_t* b = (int32_t*)malloc((size_t)(n > 0 ? n : 1) * sizeof(int32_t));
  if (!b) return HS_ENOMEM;
  int32_t st = hso_bucket_ids(cols, n_indexed, nb, b);
  if (st) { free(b); return st; }
  /* stable partition by bucket, parallel over contiguous row chunks: chunk t counts its rows per bucket,
   * the (bucket, chunk) counts are scanned bucket-major, and every chunk then writes its rows in input
   * order behind the rows of the chunks before it — the same output as the serial loop. */
  int nt = omp_get_max_threads();
  if (nt < 1) nt = 1;
  if (n < (int64_t)nt * 4096) nt = 1;
  int64_t* cur = (int64_t*)calloc((size_t)nt * (size_t)nb, sizeof(int64_t));
  if (!cur) { free(b); return HS_ENOMEM; }
  int64_t chunk = (n + nt - 1) / nt;
#pragma omp parallel for schedule(static, 1) num_threads(nt)
  for (int t = 0; t < nt; t++) {
    int64_t i0 = (int64_t)t * chunk, i1 = i0 + chunk < n ? i0 + chunk : n;
    int64_t* c = cur + (size_t)t * nb;
    for (int64_t i = i0; i < i1; i++) c[b[i]]++;
  }
  int64_t run = 0;
  for (int32_t k = 0; k < nb; k++) {
    out_bucket_offsets[k] = run;
    for (int t = 0; t < nt; t++) { int64_t c = cur[(size_t)t * nb + k]; cur[(size_t)t * nb + k] = run; run += c; }
  }
  out_bucket_offsets[nb] = run;
#pragma omp parallel for schedule(static, 1) num_threads(nt)
  for (int t = 0; t < nt; t++) {
    int64_t i0 = (int64_t)t * chunk, i1 = i0 + chunk < n ? i0 + chunk : n;
    int64_t* c = cur + (size_t)t * nb;
    for (int64_t i = i0; i < i1; i++) out_perm[c[b[i]]++] = i;
  }
  free(cur); free(b);
  return hso_sort_segments(cols, n_indexed, out_bucket_offsets, nb, out_perm, use_radix);
}

/* ---- A.5 inner sort-merge join, bucket by bucket ---------------------------
 * left/right: first nk columns are the join keys, each bucket one sorted run.
 * out_off[nb+1] receives the output bucket offsets. With li == NULL only out_off
 * is filled (count pass). Output order: left row order, then right row order. */
static int key_has_null(const hs_column* c, int64_t i, int nk) {
  for (int k = 0; k < nk; k++) if (!is_valid(&c[k], i)) return 1;
  return 0;
}
static int64_t join_bucket(const hs_column* L, int64_t l0, int64_t l1, const hs_column* R,
                           int64_t r0, int64_t r1, int nk, int64_t* li, int64_t* ri) {
  int64_t out = 0, a = l0, b = r0;
  while (a < l1 && b < r1) {
    if (key_has_null(L, a, nk)) { a++; continue; }
    if (key_has_null(R, b, nk)) { b++; continue; }
    int c = cmp_rows(L, a, R, b, nk);
    if (c < 0) a++;
    else if (c > 0) b++;
    else {
      int64_t ae = a + 1, be = b + 1;
      while (ae < l1 && cmp_rows(L, ae, L, a, nk) == 0) ae++;
      while (be < r1 && cmp_rows(R, be, R, b, nk) == 0) be++;
      if (li) for (int64_t x = a; x < ae; x++) for (int64_t y = b; y < be; y++) { li[out] = x; ri[out] = y; out++; }
      else out += (ae - a) * (be - b);
      a = ae; b = be;
    }
  }
  return out;
}
int32_t hso_join(const hs_column* left, const hs_column* right, int32_t nk, int32_t nb,
                 const int64_t* loff, const int64_t* roff, int64_t* out_off,
                 int64_t* li, int64_t* ri) {
  if (nk <= 0 || nb <= 0) return HS_EINVAL;
  for (int k = 0; k < nk; k++) {
    if (!type_ok(left[k].type)) return HS_EUNSUPPORTED;
    if (left[k].type != right[k].type) return HS_EUNSUPPORTED;
  }
  if (li == NULL) {
    out_off[0] = 0;
#pragma omp parallel for schedule(dynamic, 1)
    for (int32_t b = 0; b < nb; b++)
      out_off[b + 1] = join_bucket(left, loff[b], loff[b + 1], right, roff[b], roff[b + 1], nk, NULL, NULL);
    for (int32_t b = 0; b < nb; b++) out_off[b + 1] += out_off[b];
    return HS_OK;
  }
#pragma omp parallel for schedule(dynamic, 1)
  for (int32_t b = 0; b < nb; b++)
    join_bucket(left, loff[b], loff[b + 1], right, roff[b], roff[b + 1], nk, li + out_off[b], ri + out_off[b]);
  return HS_OK;
}

/* ---- filter: postfix predicate, SQL three-valued logic; keeps rows that are TRUE.
 * A value is (known, truth): comparisons with NULL are unknown. */
static inline void eval_leaf(const hs_column* c, int64_t i, const hs_pred_term* t, int* known, int* val) {
  if (t->op == HS_PRED_IS_NULL) { *known = 1; *val = !is_valid(c, i); return; }
  if (t->op == HS_PRED_IS_NOT_NULL) { *known = 1; *val = is_valid(c, i); return; }
  if (!is_valid(c, i)) { *known = 0; *val = 0; return; }
  if (t->op == HS_PRED_IN) {   /* SQL IN over a literal list without NULLs: TRUE if any element equals the cell */
    int64_t x = 0;
    switch (c->type) {
      case HS_BOOL: x = (((const uint8_t*)c->data)[i >> 3] >> (i & 7)) & 1; break;
      case HS_INT8: x = ((const int8_t*)c->data)[i]; break;
      case HS_INT16: x = ((const int16_t*)c->data)[i]; break;
      case HS_INT32: x = ((const int32_t*)c->data)[i]; break;
      default: x = ((const int64_t*)c->data)[i]; break;
    }
    const int64_t* list = (const int64_t*)t->lit_bytes;
    int hit = 0;
    for (int32_t k = 0; k < t->lit_len && !hit; k++) hit = list[k] == x;
    *known = 1; *val = hit;
    return;
  }
  int cmp = 0;
  switch (c->type) {
    case HS_BOOL: { int64_t x = (((const uint8_t*)c->data)[i >> 3] >> (i & 7)) & 1; cmp = (x > t->lit_i64) - (x < t->lit_i64); break; }
    case HS_INT8: { int64_t x = ((const int8_t*)c->data)[i]; cmp = (x > t->lit_i64) - (x < t->lit_i64); break; }
    case HS_INT16: { int64_t x = ((const int16_t*)c->data)[i]; cmp = (x > t->lit_i64) - (x < t->lit_i64); break; }
    case HS_INT32: { int64_t x = ((const int32_t*)c->data)[i]; cmp = (x > t->lit_i64) - (x < t->lit_i64); break; }
    case HS_INT64: { int64_t x = ((const int64_t*)c->data)[i]; cmp = (x > t->lit_i64) - (x < t->lit_i64); break; }
    case HS_FLOAT32: cmp = cmp_f64(((const float*)c->data)[i], t->lit_f64); break;
    case HS_FLOAT64: cmp = cmp_f64(((const double*)c->data)[i], t->lit_f64); break;
    case HS_UTF8: case HS_BINARY:
      cmp = cmp_bytes((const uint8_t*)c->data + c->offsets[i], c->offsets[i + 1] - c->offsets[i],
                      (const uint8_t*)t->lit_bytes, t->lit_len);
      break;
    default: break;
  }
  *known = 1;
  switch (t->op) {
    case HS_PRED_EQ: *val = cmp == 0; break;
    case HS_PRED_NE: *val = cmp != 0; break;
    case HS_PRED_LT: *val = cmp < 0; break;
    case HS_PRED_LE: *val = cmp <= 0; break;
    case HS_PRED_GT: *val = cmp > 0; break;
    case HS_PRED_GE: *val = cmp >= 0; break;
    default: *val = 0; break;
  }
}
static int eval_row(const hs_column* cols, int64_t i, const hs_predicate* p) {
  int kn[16], vl[16], sp = 0;
  for (int t = 0; t < p->nterms; t++) {
    const hs_pred_term* tm = &p->terms[t];
    if (tm->op == HS_PRED_AND || tm->op == HS_PRED_OR) {
      int kb = kn[--sp], vb = vl[sp], ka = kn[--sp], va = vl[sp], k, v;
      if (tm->op == HS_PRED_AND) {
        if ((ka && !va) || (kb && !vb)) { k = 1; v = 0; }
        else if (ka && kb) { k = 1; v = 1; } else { k = 0; v = 0; }
      } else {
        if ((ka && va) || (kb && vb)) { k = 1; v = 1; }
        else if (ka && kb) { k = 1; v = 0; } else { k = 0; v = 0; }
      }
      kn[sp] = k; vl[sp] = v; sp++;
    } else if (tm->op == HS_PRED_NOT) {
      if (kn[sp - 1]) vl[sp - 1] = !vl[sp - 1];
    } else {
      eval_leaf(&cols[tm->col], i, tm, &kn[sp], &vl[sp]); sp++;
    }
  }
  return sp == 1 && kn[0] && vl[0];
}
/* out_idx may be NULL (count only). Returns the number of matching rows in *out_n. */
int32_t hso_filter(const hs_column* cols, int32_t ncols, const hs_predicate* pred,
                   int64_t* out_idx, int64_t* out_n) {
  if (ncols <= 0 || pred == NULL || pred->nterms <= 0 || pred->nterms > 32) return HS_EINVAL;
  int depth = 0;
  for (int t = 0; t < pred->nterms; t++) {
    int op = pred->terms[t].op;
    if (op == HS_PRED_AND || op == HS_PRED_OR) { if (depth < 2) return HS_EINVAL; depth--; }
    else if (op == HS_PRED_NOT) { if (depth < 1) return HS_EINVAL; }
    else { if (pred->terms[t].col < 0 || pred->terms[t].col >= ncols) return HS_EINVAL; depth++; if (depth > 16) return HS_EINVAL; }
  }
  if (depth != 1) return HS_EINVAL;
  int64_t n = cols[0].length, m = 0;
  for (int64_t i = 0; i < n; i++)
    if (eval_row(cols, i, pred)) { if (out_idx) out_idx[m] = i; m++; }
  *out_n = m;
  return HS_OK;
}

int32_t hso_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/* libgomp reads OMP_NUM_THREADS once, when it is loaded; launchers such as torch.distributed.run export
 * OMP_NUM_THREADS=1. The timed CPU legs of bench.py call this to use every host core regardless. */
int32_t hso_set_threads(int32_t n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
  return omp_get_max_threads();
#else
  (void)n;
  return 1;
#endif
}

/* dst[i] = src[perm[i]] for a fixed-width column (esize 1, 2, 4 or 8): the materialisation of an index column
 * in its new order, parallel over the output rows (the timed CPU legs of bench.py use it instead of a
 * single-threaded numpy take). */
int32_t hso_take_fixed(const void* src, int32_t esize, const int64_t* perm, int64_t n, void* dst) {
  if (esize != 1 && esize != 2 && esize != 4 && esize != 8) return HS_EINVAL;
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; i++) {
    int64_t r = perm[i];
    switch (esize) {
      case 8: ((uint64_t*)dst)[i] = ((const uint64_t*)src)[r]; break;
      case 4: ((uint32_t*)dst)[i] = ((const uint32_t*)src)[r]; break;
      case 2: ((uint16_t*)dst)[i] = ((const uint16_t*)src)[r]; break;
      default: ((uint8_t*)dst)[i] = ((const uint8_t*)src)[r]; break;
    }
  }
  return HS_OK;
}
