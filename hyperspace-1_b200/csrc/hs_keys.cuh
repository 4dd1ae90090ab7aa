// hs_keys.cuh — device-side key arithmetic shared by the kernels:
//  * Spark 2.4.2 Murmur3_x86_32 / Murmur3Hash(cols, 42) / pmod   (stands in for
//    HashPartitioning.partitionIdExpression reached from HS/actions/CreateActionBase.scala:129-130)
//  * the order-preserving normalised key ("nkey") under Spark's ASC NULLS FIRST ordering
//    (stands in for the SortExec under HS/index/DataFrameWriterExtensions.scala:49-67)
//  * typed key comparison for the join and the filter.
#pragma once
#include "hs_common.cuh"

struct KeySpec {
  int32_t nkeys;
  KCol k[HS_MAX_KEYS];
};

// ---------------------------------------------------------------- Murmur3
__device__ __forceinline__ uint32_t mm_rotl(uint32_t x, int r) { return __funnelshift_l(x, x, r); }
__device__ __forceinline__ uint32_t mm_k1(uint32_t k) {
  k *= 0xcc9e2d51u; k = mm_rotl(k, 15); k *= 0x1b873593u; return k;
}
__device__ __forceinline__ uint32_t mm_h1(uint32_t h, uint32_t k) {
  h ^= k; h = mm_rotl(h, 13); return h * 5u + 0xe6546b64u;
}
__device__ __forceinline__ uint32_t mm_fmix(uint32_t h, uint32_t len) {
  h ^= len; h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16; return h;
}
__device__ __forceinline__ uint32_t mm_hash_int(uint32_t v, uint32_t seed) {
  return mm_fmix(mm_h1(seed, mm_k1(v)), 4);
}
__device__ __forceinline__ uint32_t mm_hash_long(uint64_t v, uint32_t seed) {
  uint32_t h = mm_h1(seed, mm_k1((uint32_t)v));
  h = mm_h1(h, mm_k1((uint32_t)(v >> 32)));
  return mm_fmix(h, 8);
}
// Spark's hashUnsafeBytes: whole LE words, then one FULL round per trailing SIGNED byte.
// Reads aligned 32-bit words and realigns with a funnel shift (no byte loads).
__device__ __forceinline__ uint32_t mm_hash_bytes(const uint8_t* p, int len, uint32_t seed) {
  uint32_t h = seed;
  uintptr_t a = (uintptr_t)p;
  int mis = (int)(a & 3);
  const uint32_t* wp = (const uint32_t*)(a - mis);
  int sh = mis * 8;
  int nw = len >> 2, tail = len & 3;
  int total = (mis + len + 3) >> 2;
  uint32_t cur = total > 0 ? __ldg(wp) : 0;
  int wi = 1;
  for (int k = 0; k < nw; k++) {
    uint32_t nxt = wi < total ? __ldg(wp + wi) : 0;
    wi++;
    uint32_t w = __funnelshift_r(cur, nxt, sh);
    cur = nxt;
    h = mm_h1(h, mm_k1(w));
  }
  if (tail) {
    uint32_t nxt = wi < total ? __ldg(wp + wi) : 0;
    uint32_t w = __funnelshift_r(cur, nxt, sh);
    for (int t = 0; t < tail; t++) {
      int32_t b = (int32_t)(int8_t)((w >> (8 * t)) & 0xff);
      h = mm_h1(h, mm_k1((uint32_t)b));
    }
  }
  return mm_fmix(h, (uint32_t)len);
}

__device__ __forceinline__ uint32_t canon_f32_bits(float f) {   // floatToIntBits
  return f != f ? 0x7fc00000u : __float_as_uint(f);
}
__device__ __forceinline__ uint64_t canon_f64_bits(double f) {  // doubleToLongBits
  return f != f ? 0x7ff8000000000000ull : (uint64_t)__double_as_longlong(f);
}

// one column's contribution to the running hash; NULL leaves it unchanged
__device__ __forceinline__ uint32_t mm_hash_value(const KCol& c, int64_t i, uint32_t h) {
  if (c.has_nulls && !kvalid(c.validity, i)) return h;
  switch (c.type) {
    case HS_BOOL: return mm_hash_int((((const uint8_t*)c.data)[i >> 3] >> (i & 7)) & 1u, h);
    case HS_INT8: return mm_hash_int((uint32_t)(int32_t)((const int8_t*)c.data)[i], h);
    case HS_INT16: return mm_hash_int((uint32_t)(int32_t)((const int16_t*)c.data)[i], h);
    case HS_INT32: return mm_hash_int(((const uint32_t*)c.data)[i], h);
    case HS_INT64: return mm_hash_long(((const uint64_t*)c.data)[i], h);
    case HS_FLOAT32: return mm_hash_int(canon_f32_bits(((const float*)c.data)[i]), h);
    case HS_FLOAT64: return mm_hash_long(canon_f64_bits(((const double*)c.data)[i]), h);
    case HS_UTF8: case HS_BINARY: case HS_LARGE_UTF8: case HS_LARGE_BINARY: {
      int32_t len;
      int64_t s = kstr_start(c, i, &len);
      return mm_hash_bytes((const uint8_t*)c.data + s, len, h);
    }
  }
  return h;
}
__device__ __forceinline__ uint32_t mm_hash_row(const KeySpec& ks, int64_t i) {
  uint32_t h = 42u;
#pragma unroll
  for (int k = 0; k < HS_MAX_KEYS; k++)
    if (k < ks.nkeys) h = mm_hash_value(ks.k[k], i, h);
  return h;
}
__device__ __forceinline__ int32_t pmod_i32(int32_t h, int32_t n) {
  int32_t r = h % n;
  return r < 0 ? r + n : r;
}

// pmod with a precomputed reciprocal (Lemire's fastmod): exact for every 32-bit input, ~10 instructions
// instead of the ~35 of a runtime-divisor `%`. c = 2^32 mod n folds the sign of the Java remainder in.
struct FastPmod {
  uint64_t M;
  uint32_t n, c;
};
static inline FastPmod make_fastpmod(int32_t n) {
  FastPmod f;
  f.n = (uint32_t)n;
  f.M = 0xFFFFFFFFFFFFFFFFull / (uint32_t)n + 1;
  f.c = (uint32_t)((1ull << 32) % (uint32_t)n);
  return f;
}
__device__ __forceinline__ uint32_t fast_pmod(uint32_t h, const FastPmod& f) {
  uint32_t a = (uint32_t)__umul64hi(f.M * h, (uint64_t)f.n);  // h mod n, h read as unsigned
  if ((int32_t)h < 0) a = a >= f.c ? a - f.c : a + f.n - f.c;  // h = u - 2^32
  return a;
}

// ---------------------------------------------------------------- order-preserving maps
__device__ __forceinline__ uint64_t omap_i64(int64_t v) { return (uint64_t)v ^ 0x8000000000000000ull; }
__device__ __forceinline__ uint64_t omap_f64(double f) {
  // NaN greatest and all NaNs equal, -0.0 == 0.0
  if (f != f) return 0xffffffffffffffffull;
  if (f == 0.0) f = 0.0;
  uint64_t b = (uint64_t)__double_as_longlong(f);
  return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
// one cell: contribution to the running hash and, for fixed-width types, the order-preserving image of the
// value (one load, one type switch for both)
__device__ __forceinline__ uint32_t mm_hash_and_map(const KCol& c, int64_t i, uint32_t h, uint64_t* mapped, bool* has) {
  *has = false;
  if (c.has_nulls && !kvalid(c.validity, i)) return h;
  switch (c.type) {
    case HS_BOOL: { uint32_t v = (((const uint8_t*)c.data)[i >> 3] >> (i & 7)) & 1u; *mapped = omap_i64(v); *has = true; return mm_hash_int(v, h); }
    case HS_INT8: { int32_t v = ((const int8_t*)c.data)[i]; *mapped = omap_i64(v); *has = true; return mm_hash_int((uint32_t)v, h); }
    case HS_INT16: { int32_t v = ((const int16_t*)c.data)[i]; *mapped = omap_i64(v); *has = true; return mm_hash_int((uint32_t)v, h); }
    case HS_INT32: { int32_t v = ((const int32_t*)c.data)[i]; *mapped = omap_i64(v); *has = true; return mm_hash_int((uint32_t)v, h); }
    case HS_INT64: { int64_t v = ((const int64_t*)c.data)[i]; *mapped = omap_i64(v); *has = true; return mm_hash_long((uint64_t)v, h); }
    case HS_FLOAT32: { float f = ((const float*)c.data)[i]; *mapped = omap_f64((double)f); *has = true; return mm_hash_int(canon_f32_bits(f), h); }
    case HS_FLOAT64: { double f = ((const double*)c.data)[i]; *mapped = omap_f64(f); *has = true; return mm_hash_long(canon_f64_bits(f), h); }
    case HS_UTF8: case HS_BINARY: case HS_LARGE_UTF8: case HS_LARGE_BINARY: {
      int32_t len;
      int64_t s = kstr_start(c, i, &len);
      return mm_hash_bytes((const uint8_t*)c.data + s, len, h);
    }
  }
  return h;
}
// mapped value of a valid fixed-width cell (callers check validity first)
__device__ __forceinline__ uint64_t omap_cell(const KCol& c, int64_t i) {
  switch (c.type) {
    case HS_BOOL: return omap_i64((((const uint8_t*)c.data)[i >> 3] >> (i & 7)) & 1);
    case HS_INT8: return omap_i64(((const int8_t*)c.data)[i]);
    case HS_INT16: return omap_i64(((const int16_t*)c.data)[i]);
    case HS_INT32: return omap_i64(((const int32_t*)c.data)[i]);
    case HS_INT64: return omap_i64(((const int64_t*)c.data)[i]);
    case HS_FLOAT32: return omap_f64((double)((const float*)c.data)[i]);
    case HS_FLOAT64: return omap_f64(((const double*)c.data)[i]);
  }
  return 0;
}

// ---------------------------------------------------------------- normalised key layout
// The sort key of a row is a sequence of 64-bit chunks compared as unsigned integers, chunk by
// chunk. Fixed-width columns become bit fields [valid flag][mapped value - column minimum] packed
// MSB-first and right-aligned inside static chunks; a string column becomes max(1, ceil(len/7))
// chunks of 7 bytes (big-endian, zero padded) + a tag byte: 0 NULL, 1+k = ends with k bytes in
// this chunk (k = 0..7), 9 = continues. Equal chunk values imply equal "more chunks follow".
struct NGroup {            // one key column
  int32_t is_string;
  int32_t chunk;           // fixed: chunk index inside its run of fixed columns
  int32_t shift;           // fixed: left shift of the value field inside the chunk
  int32_t flag_shift;      // fixed: position of the valid flag, or -1
  int32_t flag_chunk;      // fixed: chunk that holds the flag (a 65-bit field splits)
  int32_t run_id;          // fixed: which run of consecutive fixed columns
  int32_t width;           // fixed: bits of the value field (0: all valid cells hold the minimum)
  uint64_t min;            // fixed: mapped minimum over valid cells
};
struct KeyLayout {
  int32_t nkeys;
  int32_t nruns;
  int32_t run_chunks[HS_MAX_KEYS];  // static chunks per fixed run
  int32_t single_chunk;             // whole key fits chunk 0 (no refinement needed)
  int32_t bits0;                    // significant bits of chunk 0 when single_chunk
  NGroup g[HS_MAX_KEYS];
};

__device__ __forceinline__ int nk_string_chunks(int len) { return len <= 7 ? 1 : (len + 6) / 7; }

// chunk d of row i; *more = further chunks can still distinguish rows that tie on chunks 0..d
// bytes p[0..take) (take <= 7) as the top bytes of a big-endian u64, the rest zero. Two aligned 8-byte loads and
// a funnel shift instead of a byte loop: the loop's loads each wait for the previous one (the value is used at
// once), which made every caller latency-bound. An aligned word that holds a valid byte lies in that byte's
// page, so reading it whole is safe.
__device__ __forceinline__ uint64_t nk_load_be(const uint8_t* p, int take) {
  if (take <= 0) return 0;
  const uintptr_t a = (uintptr_t)p & ~(uintptr_t)7;
  const int mis = (int)((uintptr_t)p & 7);
  const uint64_t w0 = *(const uint64_t*)a;
  const uint64_t w1 = mis + take > 8 ? *(const uint64_t*)(a + 8) : 0;
  const uint64_t le = mis ? (w0 >> (8 * mis)) | (w1 << (64 - 8 * mis)) : w0;   // p[0] in the low byte
  const uint32_t lo = (uint32_t)le, hi = (uint32_t)(le >> 32);
  const uint64_t be = ((uint64_t)__byte_perm(lo, 0, 0x0123) << 32) | __byte_perm(hi, 0, 0x0123);
  return be & ~((1ull << (64 - 8 * take)) - 1ull);
}
__device__ __forceinline__ uint64_t nkey_chunk(const KeyLayout& L, const KeySpec& ks, int64_t i, int d,
                                               bool* more) {
  int base = 0;
  int k = 0;
  while (k < L.nkeys) {
    const NGroup& g = L.g[k];
    if (!g.is_string) {
      int run = g.run_id;
      int nch = L.run_chunks[run];
      int kend = k;
      while (kend < L.nkeys && !L.g[kend].is_string && L.g[kend].run_id == run) kend++;
      if (d < base + nch) {
        int want = d - base;
        uint64_t v = 0;
        for (int j = k; j < kend; j++) {
          const NGroup& f = L.g[j];
          const KCol& c = ks.k[j];
          bool valid = !c.has_nulls || kvalid(c.validity, i);
          if (f.flag_shift >= 0 && f.flag_chunk == want && valid) v |= 1ull << f.flag_shift;
          if (f.chunk == want && valid) v |= (omap_cell(c, i) - f.min) << f.shift;
        }
        *more = (want + 1 < nch) || kend < L.nkeys;
        return v;
      }
      base += nch;
      k = kend;
    } else {
      const KCol& c = ks.k[k];
      bool valid = !c.has_nulls || kvalid(c.validity, i);
      int64_t s = 0;
      int len = 0;
      if (valid) s = kstr_start(c, i, &len);
      int nch = valid ? nk_string_chunks(len) : 1;
      if (d < base + nch) {
        bool later = k + 1 < L.nkeys;
        if (!valid) { *more = later; return 0; }
        int j = d - base;
        int rem = len - 7 * j;
        int take = rem > 7 ? 7 : rem;
        const uint64_t v = nk_load_be(( const uint8_t*)c.data + s + 7 * j, take);
        uint64_t tag = rem > 7 ? 9 : (uint64_t)(1 + rem);
        *more = rem > 7 || later;
        return v | tag;
      }
      base += nch;
      k++;
    }
  }
  *more = false;
  return 0;
}

// ---------------------------------------------------------------- typed comparison (join, filter)
__device__ __forceinline__ int cmp_u8str(const uint8_t* a, int la, const uint8_t* b, int lb) {
  int m = la < lb ? la : lb;
  for (int i = 0; i < m; i++) {
    int d = (int)a[i] - (int)b[i];
    if (d) return d < 0 ? -1 : 1;
  }
  return (la > lb) - (la < lb);
}
// both cells valid
__device__ __forceinline__ int cmp_cells(const KCol& a, int64_t i, const KCol& b, int64_t j) {
  if (hs_is_varlen(a.type)) {
    int32_t la, lb;
    int64_t sa = kstr_start(a, i, &la), sb = kstr_start(b, j, &lb);
    return cmp_u8str((const uint8_t*)a.data + sa, la, (const uint8_t*)b.data + sb, lb);
  }
  uint64_t x = omap_cell(a, i), y = omap_cell(b, j);
  return (x > y) - (x < y);
}
