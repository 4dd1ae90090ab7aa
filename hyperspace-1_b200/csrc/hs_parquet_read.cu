// hs_parquet_read.cu — the read side of K4/K5: Parquet data pages decoded on the device.
// Stands where Spark's ParquetFileFormat reader stands under the bucketed FileSourceScanExec that the rules plan
// (HS/index/rules/RuleUtils.scala:280-286: HadoopFsRelation(..., ParquetFileFormat, ...) over the index files).
// The host parses the few hundred bytes of framing per file (footer, Thrift page headers: parquet_read.py) and
// hands over the raw bytes of the column chunks with one descriptor per page; everything that touches a value
// happens here, for the encodings Spark 2.4 / parquet-mr 1.10 and pyarrow write into index files:
//   * codecs UNCOMPRESSED and SNAPPY (one warp decompresses one page; pages run side by side);
//   * data pages v1 with definition levels as RLE/bit-packed hybrid of bit width 1 (flat OPTIONAL columns);
//   * values PLAIN or dictionary-encoded (PLAIN_DICTIONARY / RLE_DICTIONARY: a PLAIN dictionary page per column
//     chunk, then hybrid runs of indices), for BOOLEAN, INT32, INT64, FLOAT, DOUBLE and BYTE_ARRAY.
// One call decodes ONE column over all pages of all files handed in, into one Arrow-layout column.
#include <algorithm>

#include "hs_common.cuh"

#define PQ_CODEC_UNCOMPRESSED 0
#define PQ_CODEC_SNAPPY 1
#define PQ_ENC_PLAIN 0
#define PQ_ENC_PLAIN_DICT 2
#define PQ_ENC_RLE_DICT 8
#define PQ_KIND_DATA 0
#define PQ_KIND_DICT 2

struct PqPageD {              // device view of one page after decompression
  uint64_t body;              // byte offset of the (uncompressed) body
  uint32_t size;              // body bytes
  uint32_t nvals;             // rows of a data page / entries of a dictionary page
  int32_t kind, enc, chunk, pad;
  uint64_t row0;              // data pages: first row of the page in the output column
};

// ---- snappy: varint length, then literals (tag & 3 == 0) and copies with 1-, 2- or 4-byte offsets.
// One warp per page: every lane walks the element stream (the same bytes, uniform control flow), the bytes of each
// element are copied by the 32 lanes together. A copy that overlaps its own output (offset < length: a repeated
// pattern) reads c[i % offset], which is final before the element starts.
__global__ void __launch_bounds__(32) k_pq_snappy(const uint8_t* __restrict__ in, const uint64_t* __restrict__ in_off,
                                                  const uint32_t* __restrict__ in_size, uint8_t* __restrict__ out,
                                                  const uint64_t* __restrict__ out_off, const uint32_t* __restrict__ out_size,
                                                  int npages, uint32_t* __restrict__ bad) {
  const int p = blockIdx.x;
  if (p >= npages) return;
  const uint32_t lane = threadIdx.x;
  const uint8_t* s = in + in_off[p];
  const uint8_t* const se = s + in_size[p];
  uint8_t* d = out + out_off[p];
  uint8_t* const d0 = d;
  uint8_t* const de = d + out_size[p];
  auto fail = [&]() { if (lane == 0) atomicAdd(bad, 1u); };
  uint64_t ulen = 0;
  int sh = 0;
  while (s < se) {
    const uint8_t b = *s++;
    ulen |= (uint64_t)(b & 0x7F) << sh;
    sh += 7;
    if (!(b & 0x80)) break;
  }
  if (ulen != out_size[p]) { fail(); return; }
  while (s < se && d < de) {
    const uint8_t tag = *s++;
    uint32_t len, off = 0;
    switch (tag & 3) {
      case 0: {
        len = (tag >> 2) + 1;
        if (len > 60) {
          const int nb = (int)len - 60;
          len = 0;
          for (int i = 0; i < nb && s < se; i++) len |= (uint32_t)(*s++) << (8 * i);
          len += 1;
        }
        if (s + len > se || d + len > de) { fail(); return; }
        for (uint32_t i = lane; i < len; i += 32) d[i] = s[i];
        s += len;
        d += len;
        __syncwarp();
        continue;
      }
      case 1:
        len = ((tag >> 2) & 7) + 4;
        off = ((uint32_t)(tag >> 5) << 8) | *s++;
        break;
      case 2:
        len = (tag >> 2) + 1;
        off = (uint32_t)s[0] | ((uint32_t)s[1] << 8);
        s += 2;
        break;
      default:
        len = (tag >> 2) + 1;
        off = (uint32_t)s[0] | ((uint32_t)s[1] << 8) | ((uint32_t)s[2] << 16) | ((uint32_t)s[3] << 24);
        s += 4;
        break;
    }
    if (off == 0 || (uint64_t)(d - d0) < off || d + len > de) { fail(); return; }
    const uint8_t* c = d - off;
    if (off >= len) { for (uint32_t i = lane; i < len; i += 32) d[i] = c[i]; }
    else { for (uint32_t i = lane; i < len; i += 32) d[i] = c[i % off]; }
    d += len;
    __syncwarp();
  }
  if (d != de) fail();
}

// ---- RLE / bit-packed hybrid, read by one warp: calls emit(first index, count, is_rle, value, packed bytes, bit
// width) for every run until `want` values have been produced; returns the bytes consumed
__device__ __forceinline__ uint32_t pq_varint(const uint8_t*& s, const uint8_t* se) {
  uint32_t v = 0;
  int sh = 0;
  while (s < se) {
    const uint8_t b = *s++;
    v |= (uint32_t)(b & 0x7F) << sh;
    sh += 7;
    if (!(b & 0x80)) break;
  }
  return v;
}
// value j of a bit-packed group run (LSB first, bit width bw <= 32)
__device__ __forceinline__ uint32_t pq_unpack(const uint8_t* g, uint32_t j, int bw) {
  const uint64_t bit = (uint64_t)j * bw;
  const uint8_t* q = g + (bit >> 3);
  uint64_t w = 0;
  for (int b = 0; b < 5; b++) w |= (uint64_t)q[b] << (8 * b);   // bw <= 32: at most 5 bytes hold the value
  return (uint32_t)((w >> (bit & 7)) & ((bw == 32) ? 0xFFFFFFFFull : ((1ull << bw) - 1ull)));
}
template <typename F>
__device__ __forceinline__ void pq_hybrid_warp(const uint8_t* s, const uint8_t* se, int bw, uint32_t want, F emit) {
  uint32_t done = 0;
  const int vbytes = (bw + 7) >> 3;
  while (done < want && s < se) {
    const uint32_t h = pq_varint(s, se);
    if (h & 1u) {
      const uint32_t cnt = (h >> 1) * 8u;
      emit(done, cnt < want - done ? cnt : want - done, false, 0u, s);
      s += (size_t)(h >> 1) * bw;
      done += cnt;
    } else {
      const uint32_t cnt = h >> 1;
      uint32_t v = 0;
      for (int b = 0; b < vbytes && s + b < se; b++) v |= (uint32_t)s[b] << (8 * b);
      s += vbytes;
      emit(done, cnt < want - done ? cnt : want - done, true, v, (const uint8_t*)nullptr);
      done += cnt;
    }
  }
}

// definition levels of OPTIONAL flat columns (max level 1): one warp per data page -> one byte per row,
// the number of non-NULL rows of the page and where its values start
__global__ void __launch_bounds__(256) k_pq_levels(const uint8_t* __restrict__ bytes, const PqPageD* __restrict__ pages,
                                                   int npages, uint8_t* __restrict__ valid, uint32_t* __restrict__ nn,
                                                   uint32_t* __restrict__ val_off) {
  const int p = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (p >= npages) return;
  const PqPageD pg = pages[p];
  if (pg.kind != PQ_KIND_DATA) return;
  const uint8_t* b = bytes + pg.body;
  const uint32_t L = (uint32_t)b[0] | ((uint32_t)b[1] << 8) | ((uint32_t)b[2] << 16) | ((uint32_t)b[3] << 24);
  uint32_t cnt = 0;
  uint8_t* v = valid + pg.row0;
  pq_hybrid_warp(b + 4, b + 4 + L, 1, pg.nvals, [&](uint32_t j0, uint32_t n, bool rle, uint32_t val, const uint8_t* g) {
    for (uint32_t j = lane_id(); j < n; j += 32) {
      const uint32_t x = rle ? (val & 1u) : pq_unpack(g, j, 1);
      v[j0 + j] = (uint8_t)x;
      cnt += x;
    }
  });
#pragma unroll
  for (int d = 16; d; d >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, d);
  if (lane_id() == 0) { nn[p] = cnt; val_off[p] = 4 + L; }
}

// dictionary indices of a data page: [bit width][hybrid runs], one warp per page -> idx[first value of page + j]
__global__ void __launch_bounds__(256) k_pq_dict_indices(const uint8_t* __restrict__ bytes, const PqPageD* __restrict__ pages,
                                                         int npages, const uint32_t* __restrict__ nn,
                                                         const uint32_t* __restrict__ val_off, const uint32_t* __restrict__ vfirst,
                                                         uint32_t* __restrict__ idx) {
  const int p = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (p >= npages) return;
  const PqPageD pg = pages[p];
  if (pg.kind != PQ_KIND_DATA || pg.enc == PQ_ENC_PLAIN) return;
  const uint8_t* b = bytes + pg.body + val_off[p];
  const uint8_t* be = bytes + pg.body + pg.size;
  const int bw = b[0];
  uint32_t* out = idx + vfirst[p];
  pq_hybrid_warp(b + 1, be, bw, nn[p], [&](uint32_t j0, uint32_t n, bool rle, uint32_t val, const uint8_t* g) {
    for (uint32_t j = lane_id(); j < n; j += 32) out[j0 + j] = rle ? val : pq_unpack(g, j, bw);
  });
}

// lengths and start offsets of PLAIN byte arrays ([u32 length][bytes] ...): a serial walk, one thread per page.
// Data pages: entry vfirst[p] + j of (start, len); dictionary pages: entry dfirst[chunk] + j
__global__ void __launch_bounds__(32) k_pq_walk_bytes(const uint8_t* __restrict__ bytes, const PqPageD* __restrict__ pages, int npages,
                                                      const uint32_t* __restrict__ nn, const uint32_t* __restrict__ val_off,
                                                      const uint32_t* __restrict__ vfirst, const uint32_t* __restrict__ dfirst,
                                                      uint64_t* __restrict__ v_start, uint32_t* __restrict__ v_len,
                                                      uint64_t* __restrict__ d_start, uint32_t* __restrict__ d_len) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= npages) return;
  const PqPageD pg = pages[p];
  const bool dict = pg.kind == PQ_KIND_DICT;
  if (!dict && pg.enc != PQ_ENC_PLAIN) return;
  uint64_t at = pg.body + (dict ? 0u : val_off[p]);
  const uint64_t end = pg.body + pg.size;
  const uint32_t n = dict ? pg.nvals : nn[p];
  uint64_t* st = dict ? d_start + dfirst[pg.chunk] : v_start + vfirst[p];
  uint32_t* ln = dict ? d_len + dfirst[pg.chunk] : v_len + vfirst[p];
  for (uint32_t j = 0; j < n && at + 4 <= end; j++) {
    const uint8_t* q = bytes + at;
    const uint32_t l = (uint32_t)q[0] | ((uint32_t)q[1] << 8) | ((uint32_t)q[2] << 16) | ((uint32_t)q[3] << 24);
    st[j] = at + 4;
    ln[j] = l;
    at += 4 + (uint64_t)l;
  }
}

struct PqRowCtx {
  const uint8_t* bytes;
  const PqPageD* pages;          // data pages first (nd of them, ascending row0), dictionary pages after
  int nd;
  const uint8_t* valid;          // one byte per row
  const uint32_t* vpos;          // non-NULL rows before row i
  const uint32_t* vfirst;        // per data page: non-NULL rows before the page
  const uint32_t* val_off;       // per data page: where its values start inside the body
  const uint32_t* idx;           // dictionary index of every non-NULL value of a dictionary-encoded page
  const uint64_t* dict_body;     // per chunk: byte offset of the dictionary page's body
};
__device__ __forceinline__ int pq_row_page(const PqRowCtx& c, int64_t row) {
  int lo = 0, hi = c.nd;   // largest p with row0[p] <= row
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (c.pages[mid].row0 <= (uint64_t)row) lo = mid; else hi = mid;
  }
  return lo;
}
// fixed-width values (w_in bytes in the file, the low w_out of them kept: INT8/INT16 ride as INT32)
__global__ void __launch_bounds__(256) k_pq_rows_fixed(PqRowCtx c, int w_in, int w_out, int64_t n, uint8_t* __restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    uint8_t* d = out + (size_t)i * w_out;
    if (!c.valid[i]) { for (int b = 0; b < w_out; b++) d[b] = 0; continue; }
    const int p = pq_row_page(c, i);
    const PqPageD pg = c.pages[p];
    const uint32_t v = c.vpos[i];
    const uint8_t* s = pg.enc == PQ_ENC_PLAIN
                           ? c.bytes + pg.body + c.val_off[p] + (uint64_t)(v - c.vfirst[p]) * w_in
                           : c.bytes + c.dict_body[pg.chunk] + (uint64_t)c.idx[v] * w_in;
    for (int b = 0; b < w_out; b++) d[b] = s[b];
  }
}
// PLAIN booleans: bit j of the page's value bytes -> one byte per row
__global__ void __launch_bounds__(256) k_pq_rows_bool(PqRowCtx c, int64_t n, uint8_t* __restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    uint8_t x = 0;
    if (c.valid[i]) {
      const int p = pq_row_page(c, i);
      const PqPageD pg = c.pages[p];
      const uint32_t j = c.vpos[i] - c.vfirst[p];
      x = (c.bytes[pg.body + c.val_off[p] + (j >> 3)] >> (j & 7)) & 1;
    }
    out[i] = x;
  }
}
// byte arrays: where each row's bytes are and how many
__global__ void __launch_bounds__(256) k_pq_rows_strloc(PqRowCtx c, const uint64_t* __restrict__ v_start, const uint32_t* __restrict__ v_len,
                                                        const uint64_t* __restrict__ d_start, const uint32_t* __restrict__ d_len,
                                                        const uint32_t* __restrict__ dfirst, int64_t n,
                                                        uint64_t* __restrict__ r_start, uint32_t* __restrict__ r_len) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    uint64_t st = 0;
    uint32_t ln = 0;
    if (c.valid[i]) {
      const int p = pq_row_page(c, i);
      const PqPageD pg = c.pages[p];
      const uint32_t v = c.vpos[i];
      if (pg.enc == PQ_ENC_PLAIN) { st = v_start[v]; ln = v_len[v]; }
      else { const uint32_t e = dfirst[pg.chunk] + c.idx[v]; st = d_start[e]; ln = d_len[e]; }
    }
    r_start[i] = st;
    r_len[i] = ln;
  }
}
template <typename OFF>
__global__ void __launch_bounds__(256) k_pq_rows_strcopy(const uint8_t* __restrict__ bytes, const uint64_t* __restrict__ r_start,
                                                         const uint32_t* __restrict__ r_len, const OFF* __restrict__ off,
                                                         int64_t n, uint8_t* __restrict__ out) {
  const int sub = threadIdx.x & 7;
  for (int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 3; i < n; i += ((int64_t)gridDim.x * blockDim.x) >> 3) {
    const uint8_t* s = bytes + r_start[i];
    uint8_t* d = out + (uint64_t)off[i];
    for (uint32_t b = sub; b < r_len[i]; b += 8) d[b] = s[b];
  }
}
__global__ void k_pq_pick_first(const uint32_t* __restrict__ vpos, const uint64_t* vtotal, const PqPageD* __restrict__ pages, int nd,
                                int64_t n, uint32_t* __restrict__ vfirst) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= nd) return;
  const uint64_t r = pages[p].row0;
  vfirst[p] = r < (uint64_t)n ? vpos[r] : (uint32_t)*vtotal;
}
static __global__ void __launch_bounds__(256) k_pq_bytes_to_bits(const uint8_t* __restrict__ in, int64_t n, uint8_t* __restrict__ bits,
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
static unsigned pq_grid(int64_t n) {
  const int64_t g = (n + 255) / 256;
  return (unsigned)std::max<int64_t>(1, std::min<int64_t>(g, HS_NUM_SMS * 32));
}

extern "C" int32_t hs_parquet_decode_pages(hs_ctx* ctx, const uint8_t* bytes, int64_t nbytes, const hs_parquet_page* pages,
                                           int32_t npages, int32_t type, int32_t optional, int32_t codec, hs_table** out) {
  API_BEGIN(ctx)
  if (!out) HS_THROW(HS_EINVAL, "null out");
  *out = nullptr;
  if (!pages || npages < 0 || nbytes < 0 || (nbytes > 0 && !bytes)) HS_THROW(HS_EINVAL, "bad arguments");
  if (!hs_type_ok(type) || hs_is_large(type)) HS_THROW(HS_EUNSUPPORTED, "column type %d cannot be decoded", type);
  if (codec != PQ_CODEC_UNCOMPRESSED && codec != PQ_CODEC_SNAPPY)
    HS_THROW(HS_EUNSUPPORTED, "Parquet codec %d (only UNCOMPRESSED and SNAPPY are decoded on the device)", codec);
  hs_stages_reset(ctx);
  // ---- pages: data pages first (in row order), dictionary pages after; bodies laid out uncompressed
  std::vector<PqPageD> hp;
  std::vector<uint64_t> in_off, out_off;
  std::vector<uint32_t> in_size, out_size;
  int nchunks = 0;
  for (int i = 0; i < npages; i++) nchunks = std::max(nchunks, pages[i].chunk + 1);
  std::vector<uint64_t> dict_body((size_t)std::max(nchunks, 1), 0);
  std::vector<uint32_t> dfirst((size_t)std::max(nchunks, 1) + 1, 0);
  uint64_t rows = 0, ubytes = 0;
  bool any_dict = false;
  for (int pass = 0; pass < 2; pass++) {
    for (int i = 0; i < npages; i++) {
      const hs_parquet_page& pg = pages[i];
      const bool is_dict = pg.kind == PQ_KIND_DICT;
      if ((pass == 1) != is_dict) continue;
      if (pg.kind != PQ_KIND_DATA && pg.kind != PQ_KIND_DICT) HS_THROW(HS_EUNSUPPORTED, "page kind %d (data pages v1 and dictionary pages are decoded)", pg.kind);
      if (pg.encoding != PQ_ENC_PLAIN && pg.encoding != PQ_ENC_PLAIN_DICT && pg.encoding != PQ_ENC_RLE_DICT)
        HS_THROW(HS_EUNSUPPORTED, "page encoding %d", pg.encoding);
      if (pg.offset < 0 || pg.size < 0 || pg.offset + pg.size > nbytes || pg.chunk < 0 || pg.num_values < 0 ||
          pg.uncompressed_size < 0 || pg.uncompressed_size >= (int64_t(1) << 31))
        HS_THROW(HS_EINVAL, "page %d lies outside the bytes handed in", i);
      PqPageD d;
      memset(&d, 0, sizeof d);
      d.size = (uint32_t)(codec ? pg.uncompressed_size : pg.size);
      d.nvals = (uint32_t)pg.num_values;
      d.kind = pg.kind;
      d.enc = is_dict ? PQ_ENC_PLAIN : (pg.encoding == PQ_ENC_PLAIN ? PQ_ENC_PLAIN : PQ_ENC_RLE_DICT);
      d.chunk = pg.chunk;
      d.row0 = rows;
      if (codec) { d.body = ubytes; ubytes += ((uint64_t)d.size + 15) & ~15ull; }
      else d.body = (uint64_t)pg.offset;
      in_off.push_back((uint64_t)pg.offset); in_size.push_back((uint32_t)pg.size);
      out_off.push_back(d.body); out_size.push_back(d.size);
      if (!is_dict) { rows += (uint64_t)pg.num_values; any_dict |= d.enc != PQ_ENC_PLAIN; }
      else { dict_body[pg.chunk] = d.body; dfirst[pg.chunk + 1] = (uint32_t)pg.num_values; }
      hp.push_back(d);
    }
  }
  for (int k = 0; k < nchunks; k++) dfirst[k + 1] += dfirst[k];
  int nd = 0;
  for (auto& d : hp) nd += d.kind == PQ_KIND_DATA;
  if (rows >= 0xFFFFFFF0ull) HS_THROW(HS_EUNSUPPORTED, "more than 2^32-16 rows in one call");
  const int64_t n = (int64_t)rows;
  // ---- bytes on the device (+ padding: the bit unpacker reads up to 4 bytes past a value)
  DBuf d_in;
  const uint8_t* d_bytes = bytes;
  {
    Stage s(ctx, "h2d");
    if (nbytes > 0 && !hs_is_device_ptr(bytes)) {
      d_in.alloc(ctx, (size_t)nbytes + 16);
      CK(cudaMemcpyAsync(d_in.p, bytes, (size_t)nbytes, cudaMemcpyHostToDevice, ctx->stream));
      d_bytes = d_in.as<uint8_t>();
    }
  }
  const int np = (int)hp.size();
  DBuf d_pages(ctx, (size_t)std::max(np, 1) * sizeof(PqPageD));
  if (np) CK(cudaMemcpyAsync(d_pages.p, hp.data(), (size_t)np * sizeof(PqPageD), cudaMemcpyHostToDevice, ctx->stream));
  DBuf d_un;
  if (codec && np) {
    Stage s(ctx, "pq_snappy");
    d_un.alloc(ctx, (size_t)ubytes + 16);
    DBuf io(ctx, (size_t)np * 16), bad(ctx, 4);
    bad.zero();
    uint64_t* a = io.as<uint64_t>();
    CK(cudaMemcpyAsync(a, in_off.data(), (size_t)np * 8, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(a + np, out_off.data(), (size_t)np * 8, cudaMemcpyHostToDevice, ctx->stream));
    DBuf sizes(ctx, (size_t)np * 8);
    CK(cudaMemcpyAsync(sizes.p, in_size.data(), (size_t)np * 4, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(sizes.as<uint32_t>() + np, out_size.data(), (size_t)np * 4, cudaMemcpyHostToDevice, ctx->stream));
    // one warp per page
    LAUNCH(ctx, k_pq_snappy, np, 32, 0, d_bytes, a, sizes.as<uint32_t>(), d_un.as<uint8_t>(), a + np, sizes.as<uint32_t>() + np, np,
           bad.as<uint32_t>());
    uint32_t h_bad = 0;
    CK(cudaMemcpyAsync(&h_bad, bad.p, 4, cudaMemcpyDeviceToHost, ctx->stream));
    hs_sync(ctx);
    if (h_bad) HS_THROW(HS_EINVAL, "%u page(s) are not valid snappy streams of the stated size", h_bad);
    d_bytes = d_un.as<uint8_t>();
  }
  // ---- definition levels -> one byte per row; non-NULL rows before every row and before every page
  DBuf valid(ctx, (size_t)std::max<int64_t>(n, 1)), nn(ctx, (size_t)std::max(np, 1) * 4), val_off(ctx, (size_t)std::max(np, 1) * 4),
      vpos(ctx, (size_t)std::max<int64_t>(n, 1) * 4), vtotal(ctx, 8), vfirst(ctx, (size_t)std::max(nd, 1) * 4);
  const PqPageD* dp = d_pages.as<PqPageD>();
  {
    Stage s(ctx, "pq_levels");
    if (optional && np) {
      LAUNCH(ctx, k_pq_levels, (np * 32 + 255) / 256, 256, 0, d_bytes, dp, np, valid.as<uint8_t>(), nn.as<uint32_t>(), val_off.as<uint32_t>());
    } else {
      CK(cudaMemsetAsync(valid.p, 1, (size_t)std::max<int64_t>(n, 1), ctx->stream));
      val_off.zero();
      std::vector<uint32_t> h_nn((size_t)std::max(np, 1), 0);
      for (int p = 0; p < np; p++) h_nn[p] = hp[p].kind == PQ_KIND_DATA ? hp[p].nvals : 0;
      CK(cudaMemcpyAsync(nn.p, h_nn.data(), (size_t)std::max(np, 1) * 4, cudaMemcpyHostToDevice, ctx->stream));
      hs_sync(ctx);
    }
    hs_exclusive_scan_flags(ctx, valid.as<uint8_t>(), vpos.as<uint32_t>(), n, vtotal.as<uint64_t>());
    if (nd) LAUNCH(ctx, k_pq_pick_first, (nd + 255) / 256, 256, 0, vpos.as<uint32_t>(), vtotal.as<uint64_t>(), dp, nd, n, vfirst.as<uint32_t>());
  }
  uint64_t h_vtotal = 0;
  CK(cudaMemcpyAsync(&h_vtotal, vtotal.p, 8, cudaMemcpyDeviceToHost, ctx->stream));
  hs_sync(ctx);
  const int64_t nv = (int64_t)h_vtotal;
  // ---- dictionary indices
  DBuf idx(ctx, (size_t)std::max<int64_t>(nv, 1) * 4), d_dict(ctx, (size_t)std::max(nchunks, 1) * 8),
      d_dfirst(ctx, ((size_t)std::max(nchunks, 1) + 1) * 4);
  CK(cudaMemcpyAsync(d_dict.p, dict_body.data(), (size_t)std::max(nchunks, 1) * 8, cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemcpyAsync(d_dfirst.p, dfirst.data(), ((size_t)std::max(nchunks, 1) + 1) * 4, cudaMemcpyHostToDevice, ctx->stream));
  if (any_dict) {
    Stage s(ctx, "pq_dict_indices");
    LAUNCH(ctx, k_pq_dict_indices, (np * 32 + 255) / 256, 256, 0, d_bytes, dp, np, nn.as<uint32_t>(), val_off.as<uint32_t>(),
           vfirst.as<uint32_t>(), idx.as<uint32_t>());
  }
  PqRowCtx rc;
  rc.bytes = d_bytes; rc.pages = dp; rc.nd = nd; rc.valid = valid.as<uint8_t>(); rc.vpos = vpos.as<uint32_t>();
  rc.vfirst = vfirst.as<uint32_t>(); rc.val_off = val_off.as<uint32_t>(); rc.idx = idx.as<uint32_t>();
  rc.dict_body = d_dict.as<uint64_t>();
  // ---- values
  std::vector<OutCol> ocols(1);
  OutCol& o = ocols[0];
  o.type = type;
  o.length = n;
  o.null_count = n - nv;
  {
    Stage s(ctx, "pq_values");
    if (hs_is_varlen(type)) {
      const uint32_t ndict = dfirst[(size_t)std::max(nchunks, 1)];
      DBuf v_start(ctx, (size_t)std::max<int64_t>(nv, 1) * 8), v_len(ctx, (size_t)std::max<int64_t>(nv, 1) * 4),
          k_start(ctx, (size_t)std::max<uint32_t>(ndict, 1) * 8), k_len(ctx, (size_t)std::max<uint32_t>(ndict, 1) * 4),
          r_start(ctx, (size_t)std::max<int64_t>(n, 1) * 8), r_len(ctx, (size_t)std::max<int64_t>(n, 1) * 4);
      if (np)
        LAUNCH(ctx, k_pq_walk_bytes, np, 1, 0, d_bytes, dp, np, nn.as<uint32_t>(), val_off.as<uint32_t>(), vfirst.as<uint32_t>(),
               d_dfirst.as<uint32_t>(), v_start.as<uint64_t>(), v_len.as<uint32_t>(), k_start.as<uint64_t>(), k_len.as<uint32_t>());
      if (n > 0)
        LAUNCH(ctx, k_pq_rows_strloc, pq_grid(n), 256, 0, rc, v_start.as<uint64_t>(), v_len.as<uint32_t>(), k_start.as<uint64_t>(),
               k_len.as<uint32_t>(), d_dfirst.as<uint32_t>(), n, r_start.as<uint64_t>(), r_len.as<uint32_t>());
      const int64_t total = hs_offsets_from_lens(ctx, r_len.as<uint32_t>(), n, type, o);
      o.data_bytes = total;
      o.data.alloc(ctx, (size_t)std::max<int64_t>(total, 16));
      if (n > 0 && total > 0) {
        if (hs_is_large(o.type))
          LAUNCH(ctx, k_pq_rows_strcopy<int64_t>, pq_grid(n * 8), 256, 0, d_bytes, r_start.as<uint64_t>(), r_len.as<uint32_t>(),
                 o.offsets.as<int64_t>(), n, o.data.as<uint8_t>());
        else
          LAUNCH(ctx, k_pq_rows_strcopy<int32_t>, pq_grid(n * 8), 256, 0, d_bytes, r_start.as<uint64_t>(), r_len.as<uint32_t>(),
                 o.offsets.as<int32_t>(), n, o.data.as<uint8_t>());
      }
      hs_sync(ctx);   // the location arrays go out of scope
    } else if (type == HS_BOOL) {
      DBuf by(ctx, (size_t)std::max<int64_t>(n, 1));
      if (n > 0) LAUNCH(ctx, k_pq_rows_bool, pq_grid(n), 256, 0, rc, n, by.as<uint8_t>());
      o.data_bytes = (n + 7) / 8;
      o.data.alloc(ctx, (size_t)o.data_bytes + 16);
      if (n > 0) LAUNCH(ctx, k_pq_bytes_to_bits, pq_grid((n + 7) / 8), 256, 0, by.as<uint8_t>(), n, o.data.as<uint8_t>(),
                        (unsigned long long*)nullptr);
      hs_sync(ctx);
    } else {
      const int w_out = hs_fixed_width(type);
      const int w_in = (type == HS_INT8 || type == HS_INT16) ? 4 : w_out;   // Parquet has no 1- and 2-byte integers
      o.data_bytes = n * w_out;
      o.data.alloc(ctx, (size_t)std::max<int64_t>(o.data_bytes, 16));
      if (n > 0) LAUNCH(ctx, k_pq_rows_fixed, pq_grid(n), 256, 0, rc, w_in, w_out, n, o.data.as<uint8_t>());
    }
    if (o.null_count > 0) {
      o.validity.alloc(ctx, (size_t)(n + 7) / 8 + 16);
      LAUNCH(ctx, k_pq_bytes_to_bits, pq_grid((n + 7) / 8), 256, 0, valid.as<uint8_t>(), n, o.validity.as<uint8_t>(),
             (unsigned long long*)nullptr);
    }
  }
  Stage s(ctx, "d2h");
  *out = hs_finish_table(ctx, ocols, n, 0, nullptr);
  API_END(ctx)
}
