// hs_filter.cu — K7: predicate-evaluating scan over an index.
// Executes the plan FilterIndexRule.apply produces (HS/index/rules/FilterIndexRule.scala:43-85): rows of the
// index for which the predicate is TRUE (SQL three-valued logic), projected to the requested columns.
// `first indexed column = literal` on a single-column index prunes the scan to the literal's bucket.
#include <algorithm>

#include "hs_keys.cuh"
#include "hs_kernels.h"

#define HS_MAX_TERMS 32

struct KTerm {
  int32_t op, col;
  int64_t lit_i64;
  double lit_f64;
  int32_t lit_off, lit_len;  // into the literal byte pool
};
struct KPred {
  int32_t nterms;
  KTerm t[HS_MAX_TERMS];
  const uint8_t* pool;
};
struct KCols {
  KCol c[HS_MAX_MOVE];
};

__device__ __forceinline__ int cmp_f64_total(double a, double b) {
  bool an = a != a, bn = b != b;
  if (an || bn) return (int)an - (int)bn;
  return (a > b) - (a < b);
}

__device__ __forceinline__ void eval_leaf(const KCol& c, int64_t i, const KTerm& t, const uint8_t* pool, bool* known,
                                          bool* val) {
  bool valid = !c.has_nulls || kvalid(c.validity, i);
  if (t.op == HS_PRED_IS_NULL) { *known = true; *val = !valid; return; }
  if (t.op == HS_PRED_IS_NOT_NULL) { *known = true; *val = valid; return; }
  if (!valid) { *known = false; *val = false; return; }
  if (t.op == HS_PRED_IN) {
    // integer cell against a sorted list of int64 in the literal pool (8-byte aligned there)
    int64_t x = 0;
    switch (c.type) {
      case HS_BOOL: x = (((const uint8_t*)c.data)[i >> 3] >> (i & 7)) & 1; break;
      case HS_INT8: x = ((const int8_t*)c.data)[i]; break;
      case HS_INT16: x = ((const int16_t*)c.data)[i]; break;
      case HS_INT32: x = ((const int32_t*)c.data)[i]; break;
      default: x = ((const int64_t*)c.data)[i]; break;
    }
    const int64_t* list = (const int64_t*)(pool + t.lit_off);
    int lo = 0, hi = t.lit_len;
    while (lo < hi) { int mid = (lo + hi) >> 1; if (list[mid] < x) lo = mid + 1; else hi = mid; }
    *known = true;
    *val = lo < t.lit_len && list[lo] == x;
    return;
  }
  int cmp = 0;
  switch (c.type) {
    case HS_BOOL: { int64_t x = (((const uint8_t*)c.data)[i >> 3] >> (i & 7)) & 1; cmp = (x > t.lit_i64) - (x < t.lit_i64); break; }
    case HS_INT8: { int64_t x = ((const int8_t*)c.data)[i]; cmp = (x > t.lit_i64) - (x < t.lit_i64); break; }
    case HS_INT16: { int64_t x = ((const int16_t*)c.data)[i]; cmp = (x > t.lit_i64) - (x < t.lit_i64); break; }
    case HS_INT32: { int64_t x = ((const int32_t*)c.data)[i]; cmp = (x > t.lit_i64) - (x < t.lit_i64); break; }
    case HS_INT64: { int64_t x = ((const int64_t*)c.data)[i]; cmp = (x > t.lit_i64) - (x < t.lit_i64); break; }
    case HS_FLOAT32: cmp = cmp_f64_total((double)((const float*)c.data)[i], t.lit_f64); break;
    case HS_FLOAT64: cmp = cmp_f64_total(((const double*)c.data)[i], t.lit_f64); break;
    case HS_UTF8: case HS_BINARY: case HS_LARGE_UTF8: case HS_LARGE_BINARY: {
      int32_t len;
      int64_t s = kstr_start(c, i, &len);
      cmp = cmp_u8str((const uint8_t*)c.data + s, len, pool + t.lit_off, t.lit_len);
      break;
    }
  }
  *known = true;
  switch (t.op) {
    case HS_PRED_EQ: *val = cmp == 0; break;
    case HS_PRED_NE: *val = cmp != 0; break;
    case HS_PRED_LT: *val = cmp < 0; break;
    case HS_PRED_LE: *val = cmp <= 0; break;
    case HS_PRED_GT: *val = cmp > 0; break;
    case HS_PRED_GE: *val = cmp >= 0; break;
    default: *val = false; break;
  }
}

// the postfix program on one row -> TRUE or not (3-valued logic: kn = "known" bits, vl = value bits; the stack
// is two 32-bit masks)
__device__ __forceinline__ bool eval_pred(const KCols& cols, const KPred& pred, int64_t row) {
  uint32_t kn = 0, vl = 0;
  int sp = 0;
  for (int t = 0; t < pred.nterms; t++) {
    const KTerm& tm = pred.t[t];
    if (tm.op == HS_PRED_AND || tm.op == HS_PRED_OR) {
      bool kb = (kn >> (sp - 1)) & 1, vb = (vl >> (sp - 1)) & 1;
      bool ka = (kn >> (sp - 2)) & 1, va = (vl >> (sp - 2)) & 1;
      bool k, v;
      if (tm.op == HS_PRED_AND) {
        if ((ka && !va) || (kb && !vb)) { k = true; v = false; }
        else if (ka && kb) { k = true; v = true; }
        else { k = false; v = false; }
      } else {
        if ((ka && va) || (kb && vb)) { k = true; v = true; }
        else if (ka && kb) { k = true; v = false; }
        else { k = false; v = false; }
      }
      sp -= 2;
      kn = (kn & ~(3u << sp)) | ((uint32_t)k << sp);
      vl = (vl & ~(3u << sp)) | ((uint32_t)v << sp);
      sp++;
    } else if (tm.op == HS_PRED_NOT) {
      if ((kn >> (sp - 1)) & 1) vl ^= 1u << (sp - 1);
    } else {
      bool k, v;
      eval_leaf(cols.c[tm.col], row, tm, pred.pool, &k, &v);
      kn = (kn & ~(1u << sp)) | ((uint32_t)k << sp);
      vl = (vl & ~(1u << sp)) | ((uint32_t)v << sp);
      sp++;
    }
  }
  return (kn & 1u) && (vl & 1u);
}

// rows [row0, row0 + n): flag[i - row0] = predicate is TRUE
__global__ void __launch_bounds__(256) k_filter_eval(const __grid_constant__ KCols cols,
                                                     const __grid_constant__ KPred pred, int64_t row0, int64_t n,
                                                     uint8_t* __restrict__ flag) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) flag[i] = eval_pred(cols, pred, row0 + i);
}

// ---- fixed-width, NULL-free output columns: two streaming passes, no row-id list
// Pass 1 evaluates the predicate once and keeps one bit per row (a warp's ballot per 32 rows) plus the number of
// matches per tile of FT_TILE rows. After a scan of the tile counts, pass 2 reads the bits, ranks the matches
// of a tile with popcounts and copies every output column straight from its source to its packed place — rows
// keep their order, so a scan of a sorted index stays sorted. Per row: predicate columns once, output columns
// once, 1/8 byte of bits twice.
#define FT_TILE 2048
__global__ void __launch_bounds__(256) k_filter_mask(const __grid_constant__ KCols cols, const __grid_constant__ KPred pred,
                                                     int64_t row0, int64_t n, uint32_t* __restrict__ mask,
                                                     uint32_t* __restrict__ tile_cnt) {
  __shared__ uint32_t s_cnt[8];
  const int64_t base = (int64_t)blockIdx.x * FT_TILE;
  uint32_t c = 0;
#pragma unroll
  for (int k = 0; k < FT_TILE / 256; k++) {
    const int64_t i = base + k * 256 + threadIdx.x;
    const bool f = i < n && eval_pred(cols, pred, row0 + i);
    const uint32_t b = __ballot_sync(0xffffffffu, f);
    if (lane_id() == 0) {
      if (base + k * 256 + (threadIdx.x & ~31) < n) mask[(base + k * 256 + threadIdx.x) >> 5] = b;
      c += __popc(b);
    }
  }
  if (lane_id() == 0) s_cnt[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
    for (int w = 0; w < 8; w++) t += s_cnt[w];
    tile_cnt[blockIdx.x] = t;
  }
}
// The common shape of FilterIndexRule's predicate — one comparison, or a range (two comparisons joined by AND), on
// one NULL-free integer column — needs no interpreter: the tile's values are loaded up front (independent loads)
// and compared against the literals. Same bits, same counts as k_filter_mask.
__device__ __forceinline__ bool cmp_op(int op, int64_t x, int64_t lit) {
  switch (op) {
    case HS_PRED_EQ: return x == lit;
    case HS_PRED_NE: return x != lit;
    case HS_PRED_LT: return x < lit;
    case HS_PRED_LE: return x <= lit;
    case HS_PRED_GT: return x > lit;
    default: return x >= lit;
  }
}
template <typename T>
__global__ void __launch_bounds__(256) k_filter_mask_int(const T* __restrict__ col, int op1, int64_t lit1, int op2,
                                                         int64_t lit2, int64_t row0, int64_t n,
                                                         uint32_t* __restrict__ mask, uint32_t* __restrict__ tile_cnt) {
  __shared__ uint32_t s_cnt[8];
  const int64_t base = (int64_t)blockIdx.x * FT_TILE;
  T v[FT_TILE / 256];
#pragma unroll
  for (int k = 0; k < FT_TILE / 256; k++) {
    const int64_t i = base + k * 256 + threadIdx.x;
    v[k] = i < n ? col[row0 + i] : (T)0;
  }
  uint32_t c = 0;
#pragma unroll
  for (int k = 0; k < FT_TILE / 256; k++) {
    const int64_t i = base + k * 256 + threadIdx.x;
    const bool f = i < n && cmp_op(op1, (int64_t)v[k], lit1) && (op2 == 0 || cmp_op(op2, (int64_t)v[k], lit2));
    const uint32_t b = __ballot_sync(0xffffffffu, f);
    if (lane_id() == 0) {
      if (base + k * 256 + (threadIdx.x & ~31) < n) mask[(base + k * 256 + threadIdx.x) >> 5] = b;
      c += __popc(b);
    }
  }
  if (lane_id() == 0) s_cnt[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
    for (int w = 0; w < 8; w++) t += s_cnt[w];
    tile_cnt[blockIdx.x] = t;
  }
}
struct FilterOut {
  int32_t n;
  int32_t es[HS_MAX_MOVE];
  const void* src[HS_MAX_MOVE];
  void* dst[HS_MAX_MOVE];
};
__global__ void __launch_bounds__(256) k_filter_write(const __grid_constant__ FilterOut fo, int64_t row0, int64_t n,
                                                      const uint32_t* __restrict__ mask,
                                                      const uint32_t* __restrict__ tile_off) {
  __shared__ uint32_t s_word[FT_TILE / 32], s_pre[FT_TILE / 32];
  const int64_t base = (int64_t)blockIdx.x * FT_TILE;
  if (threadIdx.x < FT_TILE / 32) {
    const int64_t w = (base >> 5) + threadIdx.x;
    const uint32_t v = (w << 5) < n ? mask[w] : 0u;
    s_word[threadIdx.x] = v;
    // exclusive prefix of the popcounts of the tile's 64 words: per warp here, the second warp's words get the
    // first warp's total (s_half) added where they are used
    uint32_t inc = __popc(v);
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d);
      if (lane_id() >= (uint32_t)d) inc += o;
    }
    s_pre[threadIdx.x] = inc - __popc(v);
  }
  __syncthreads();
  __shared__ uint32_t s_half;
  if (threadIdx.x == 31) s_half = s_pre[31] + __popc(s_word[31]);
  __syncthreads();
  const uint32_t obase = tile_off[blockIdx.x];
  uint32_t rank[FT_TILE / 256];
  uint32_t fl = 0;
#pragma unroll
  for (int k = 0; k < FT_TILE / 256; k++) {
    const uint32_t r = k * 256 + threadIdx.x, w = r >> 5;
    const uint32_t word = s_word[w];
    if ((word >> (r & 31)) & 1u) fl |= 1u << k;
    rank[k] = obase + s_pre[w] + (w >= 32 ? s_half : 0u) + __popc(word & ((1u << (r & 31)) - 1u));
  }
  for (int c = 0; c < fo.n; c++) {
    const int es = fo.es[c];
#pragma unroll
    for (int k = 0; k < FT_TILE / 256; k++) {
      if (!((fl >> k) & 1u)) continue;
      const int64_t i = row0 + base + k * 256 + threadIdx.x;
      switch (es) {
        case 8: ((uint64_t*)fo.dst[c])[rank[k]] = ((const uint64_t*)fo.src[c])[i]; break;
        case 4: ((uint32_t*)fo.dst[c])[rank[k]] = ((const uint32_t*)fo.src[c])[i]; break;
        case 2: ((uint16_t*)fo.dst[c])[rank[k]] = ((const uint16_t*)fo.src[c])[i]; break;
        default: ((uint8_t*)fo.dst[c])[rank[k]] = ((const uint8_t*)fo.src[c])[i]; break;
      }
    }
  }
}

__global__ void __launch_bounds__(256) k_filter_compact(const uint8_t* __restrict__ flag, const uint32_t* __restrict__ pos,
                                                        int64_t row0, int64_t n, uint32_t* __restrict__ idx) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride)
    if (flag[i]) idx[pos[i]] = (uint32_t)(row0 + i);
}

// host copy of the Murmur3 of hs_keys.cuh for one literal (bucket pruning only)
static uint32_t h_rotl(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }
static uint32_t h_k1(uint32_t k) { k *= 0xcc9e2d51u; k = h_rotl(k, 15); k *= 0x1b873593u; return k; }
static uint32_t h_h1(uint32_t h, uint32_t k) { h ^= k; h = h_rotl(h, 13); return h * 5u + 0xe6546b64u; }
static uint32_t h_fmix(uint32_t h, uint32_t len) {
  h ^= len; h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16; return h;
}
static bool literal_bucket(int32_t type, const hs_pred_term& t, int32_t nb, int32_t* bucket) {
  uint32_t h;
  switch (type) {
    case HS_INT8: case HS_INT16: case HS_INT32: {
      int64_t lo = type == HS_INT8 ? -128 : type == HS_INT16 ? -32768 : -2147483648ll;
      int64_t hi = type == HS_INT8 ? 127 : type == HS_INT16 ? 32767 : 2147483647ll;
      if (t.lit_i64 < lo || t.lit_i64 > hi) return false;
      h = h_fmix(h_h1(42u, h_k1((uint32_t)(int32_t)t.lit_i64)), 4);
      break;
    }
    case HS_INT64: {
      uint64_t v = (uint64_t)t.lit_i64;
      h = h_h1(42u, h_k1((uint32_t)v));
      h = h_fmix(h_h1(h, h_k1((uint32_t)(v >> 32))), 8);
      break;
    }
    case HS_UTF8: case HS_BINARY: case HS_LARGE_UTF8: case HS_LARGE_BINARY: {
      if (!t.lit_bytes && t.lit_len) return false;
      h = 42u;
      int n = t.lit_len, nw = n & ~3;
      for (int i = 0; i < nw; i += 4) { uint32_t w; memcpy(&w, t.lit_bytes + i, 4); h = h_h1(h, h_k1(w)); }
      for (int i = nw; i < n; i++) h = h_h1(h, h_k1((uint32_t)(int32_t)(int8_t)t.lit_bytes[i]));
      h = h_fmix(h, (uint32_t)n);
      break;
    }
    default: return false;
  }
  int32_t r = (int32_t)h % nb;
  *bucket = r < 0 ? r + nb : r;
  return true;
}

static unsigned fgrid(int64_t n) {
  int64_t g = (n + 255) / 256, cap = (int64_t)HS_NUM_SMS * 16;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (unsigned)g;
}

extern "C" int32_t hs_filter_scan(hs_ctx* ctx, const hs_index_view* index, const hs_predicate* pred,
                                  const int32_t* out_cols, int32_t n_out, hs_table** out) {
  API_BEGIN(ctx)
  if (!out) HS_THROW(HS_EINVAL, "null out");
  *out = nullptr;
  if (!index || !index->cols || index->ncols <= 0) HS_THROW(HS_EINVAL, "malformed index view");
  if (index->ncols > HS_MAX_MOVE) HS_THROW(HS_EUNSUPPORTED, "more than %d columns", HS_MAX_MOVE);
  if (!pred || !pred->terms || pred->nterms <= 0 || pred->nterms > HS_MAX_TERMS)
    HS_THROW(HS_EINVAL, "predicate must have 1..%d terms", HS_MAX_TERMS);
  if (n_out <= 0 || !out_cols) HS_THROW(HS_EINVAL, "no output columns");
  for (int i = 0; i < n_out; i++)
    if (out_cols[i] < 0 || out_cols[i] >= index->ncols) HS_THROW(HS_EINVAL, "output column out of range");
  // validate the postfix program
  int depth = 0;
  size_t pool_bytes = 0;
  for (int t = 0; t < pred->nterms; t++) {
    const hs_pred_term& tm = pred->terms[t];
    if (tm.op == HS_PRED_AND || tm.op == HS_PRED_OR) { if (depth < 2) HS_THROW(HS_EINVAL, "malformed predicate"); depth--; }
    else if (tm.op == HS_PRED_NOT) { if (depth < 1) HS_THROW(HS_EINVAL, "malformed predicate"); }
    else if (tm.op >= HS_PRED_EQ && tm.op <= HS_PRED_IN) {
      if (tm.col < 0 || tm.col >= index->ncols) HS_THROW(HS_EINVAL, "predicate column out of range");
      if (++depth > 16) HS_THROW(HS_EINVAL, "predicate too deep");
      if (tm.op == HS_PRED_IN) {
        const int32_t ct = index->cols[tm.col].type;
        if (ct < HS_BOOL || ct > HS_INT64) HS_THROW(HS_EUNSUPPORTED, "IN is implemented for integer columns");
        if (tm.lit_len < 0 || (tm.lit_len > 0 && !tm.lit_bytes)) HS_THROW(HS_EINVAL, "IN list missing");
        pool_bytes = ((pool_bytes + 7) & ~(size_t)7) + (size_t)tm.lit_len * 8;
      }
      if (hs_is_varlen(index->cols[tm.col].type) && tm.op <= HS_PRED_GE) {
        if (tm.lit_len < 0 || (tm.lit_len > 0 && !tm.lit_bytes)) HS_THROW(HS_EINVAL, "string literal missing");
        pool_bytes += (size_t)tm.lit_len;
      }
    } else HS_THROW(HS_EINVAL, "unknown predicate op %d", tm.op);
  }
  if (depth != 1) HS_THROW(HS_EINVAL, "malformed predicate");
  hs_stages_reset(ctx);
  std::vector<DevCol> cols;
  {
    Stage s(ctx, "h2d");
    hs_stage_columns(ctx, index->cols, index->ncols, cols);
  }
  const int64_t nrows = cols[0].length;
  KCols kc;
  memset(&kc, 0, sizeof kc);
  for (int i = 0; i < index->ncols; i++) kc.c[i] = hs_kcol(cols[i]);
  KPred kp;
  memset(&kp, 0, sizeof kp);
  kp.nterms = pred->nterms;
  std::vector<uint8_t> pool(pool_bytes + 16);
  size_t at = 0;
  for (int t = 0; t < pred->nterms; t++) {
    const hs_pred_term& tm = pred->terms[t];
    kp.t[t].op = tm.op;
    kp.t[t].col = tm.col;
    kp.t[t].lit_i64 = tm.lit_i64;
    kp.t[t].lit_f64 = tm.lit_f64;
    if (tm.op == HS_PRED_IN) {
      at = (at + 7) & ~(size_t)7;
      kp.t[t].lit_off = (int32_t)at;
      kp.t[t].lit_len = tm.lit_len;
      if (tm.lit_len) {
        int64_t* dst = (int64_t*)(pool.data() + at);
        memcpy(dst, tm.lit_bytes, (size_t)tm.lit_len * 8);
        std::sort(dst, dst + tm.lit_len);
      }
      at += (size_t)tm.lit_len * 8;
    } else if (tm.op <= HS_PRED_GE && hs_is_varlen(index->cols[tm.col].type)) {
      kp.t[t].lit_off = (int32_t)at;
      kp.t[t].lit_len = tm.lit_len;
      if (tm.lit_len) memcpy(pool.data() + at, tm.lit_bytes, (size_t)tm.lit_len);
      at += (size_t)tm.lit_len;
    }
  }
  DBuf dpool(ctx, pool.size());
  CK(cudaMemcpyAsync(dpool.p, pool.data(), pool.size(), cudaMemcpyHostToDevice, ctx->stream));
  kp.pool = dpool.as<uint8_t>();

  // bucket pruning: key = literal on a single-column index scans one bucket only
  int64_t row0 = 0, n = nrows;
  if (pred->nterms == 1 && pred->terms[0].op == HS_PRED_EQ && pred->terms[0].col == 0 && index->n_indexed == 1 &&
      index->num_buckets > 0 && index->bucket_offsets) {
    int32_t b;
    if (literal_bucket(index->cols[0].type, pred->terms[0], index->num_buckets, &b)) {
      row0 = index->bucket_offsets[b];
      n = index->bucket_offsets[b + 1] - row0;
      if (row0 < 0 || n < 0 || row0 + n > nrows) HS_THROW(HS_EINVAL, "bad bucket_offsets");
    }
  }
  bool packed_out = n > 0;
  for (int i = 0; i < n_out && packed_out; i++) {
    const DevCol& oc = cols[out_cols[i]];
    if (hs_fixed_width(oc.type) == 0 || oc.validity) packed_out = false;
  }
  if (packed_out) {
    const uint32_t ntiles = (uint32_t)((n + FT_TILE - 1) / FT_TILE);
    DBuf mask(ctx, (size_t)ntiles * (FT_TILE / 32) * 4), tcnt(ctx, (size_t)ntiles * 4), total(ctx, 8);
    uint64_t h_total = 0;
    {
      Stage s(ctx, "filter_eval");
      // one comparison or a range on one NULL-free INT32/INT64 column: the specialised kernel
      auto is_cmp = [&](const KTerm& t) { return t.op >= HS_PRED_EQ && t.op <= HS_PRED_GE; };
      int fast_col = -1, op2 = 0;
      int64_t lit2 = 0;
      if (kp.nterms == 1 && is_cmp(kp.t[0])) fast_col = kp.t[0].col;
      else if (kp.nterms == 3 && is_cmp(kp.t[0]) && is_cmp(kp.t[1]) && kp.t[2].op == HS_PRED_AND && kp.t[0].col == kp.t[1].col) {
        fast_col = kp.t[0].col;
        op2 = kp.t[1].op;
        lit2 = kp.t[1].lit_i64;
      }
      if (fast_col >= 0 && (kc.c[fast_col].has_nulls || (kc.c[fast_col].type != HS_INT64 && kc.c[fast_col].type != HS_INT32)))
        fast_col = -1;
      if (fast_col >= 0 && kc.c[fast_col].type == HS_INT64)
        LAUNCH(ctx, k_filter_mask_int<int64_t>, ntiles, 256, 0, (const int64_t*)kc.c[fast_col].data, kp.t[0].op,
               kp.t[0].lit_i64, op2, lit2, row0, n, mask.as<uint32_t>(), tcnt.as<uint32_t>());
      else if (fast_col >= 0)
        LAUNCH(ctx, k_filter_mask_int<int32_t>, ntiles, 256, 0, (const int32_t*)kc.c[fast_col].data, kp.t[0].op,
               kp.t[0].lit_i64, op2, lit2, row0, n, mask.as<uint32_t>(), tcnt.as<uint32_t>());
      else
        LAUNCH(ctx, k_filter_mask, ntiles, 256, 0, kc, kp, row0, n, mask.as<uint32_t>(), tcnt.as<uint32_t>());
      hs_exclusive_scan_u32(ctx, tcnt.as<uint32_t>(), tcnt.as<uint32_t>(), ntiles, total.as<uint64_t>());
      h_total = *(const uint64_t*)hs_read_wait(ctx, hs_read_begin(ctx, total.p, 8));
    }
    const int64_t m = (int64_t)h_total;
    std::vector<OutCol> ocols((size_t)n_out);
    FilterOut fo;
    memset(&fo, 0, sizeof fo);
    fo.n = n_out;
    for (int i = 0; i < n_out; i++) {
      const DevCol& oc = cols[out_cols[i]];
      OutCol& o = ocols[i];
      o.type = oc.type; o.length = m; o.null_count = 0;
      o.data_bytes = m * hs_fixed_width(oc.type);
      o.data.alloc(ctx, (size_t)o.data_bytes);
      fo.es[i] = hs_fixed_width(oc.type); fo.src[i] = oc.data; fo.dst[i] = o.data.p;
    }
    if (m > 0) {
      Stage s(ctx, "filter_write");
      LAUNCH(ctx, k_filter_write, ntiles, 256, 0, fo, row0, n, mask.as<uint32_t>(), tcnt.as<uint32_t>());
    }
    Stage s(ctx, "d2h");
    *out = hs_finish_table(ctx, ocols, m, 0, nullptr);
    return HS_OK;
  }
  DBuf flag(ctx, (size_t)n), pos(ctx, (size_t)n * 4), total(ctx, 8);
  uint64_t h_total = 0;
  {
    Stage s(ctx, "filter_eval");
    if (n > 0) LAUNCH(ctx, k_filter_eval, fgrid(n), 256, 0, kc, kp, row0, n, flag.as<uint8_t>());
    hs_exclusive_scan_flags(ctx, flag.as<uint8_t>(), pos.as<uint32_t>(), n, total.as<uint64_t>());
    CK(cudaMemcpyAsync(&h_total, total.p, 8, cudaMemcpyDeviceToHost, ctx->stream));
    hs_sync(ctx);
  }
  const int64_t m = (int64_t)h_total;
  DBuf idx(ctx, (size_t)m * 4);
  std::vector<OutCol> ocols((size_t)n_out);
  {
    Stage s(ctx, "filter_gather");
    if (n > 0 && m > 0) LAUNCH(ctx, k_filter_compact, fgrid(n), 256, 0, flag.as<uint8_t>(), pos.as<uint32_t>(), row0, n, idx.as<uint32_t>());
    for (int i = 0; i < n_out; i++) hs_gather_column(ctx, cols[out_cols[i]], idx.as<uint32_t>(), m, ocols[i]);
  }
  Stage s(ctx, "d2h");
  *out = hs_finish_table(ctx, ocols, m, 0, nullptr);
  API_END(ctx)
}
