// hs_parquet.cu — K4: the bodies of Parquet v1 data pages, PLAIN encoding, built on the device.
// Stands where Spark's DynamicPartitionDataWriter + ParquetFileFormat stand behind saveWithBuckets
// (HS/index/DataFrameWriterExtensions.scala:58-65): the rows of a bucket become column chunks of pages. The page
// and file framing (Thrift headers, footer) is a few hundred bytes per file and is written by the host
// (hyperspace-1_b200/parquet_plain.py); what touches every value happens here:
//   * definition levels of an OPTIONAL column: RLE/bit-packed hybrid of bit width 1 — one RLE run for a page
//     without NULLs, otherwise bit-packed groups, which are the validity bitmap re-aligned to the page start;
//   * PLAIN values of the non-NULL rows: fixed-width values compacted, booleans re-packed, strings as
//     4-byte little-endian length + bytes.
// One call encodes all pages of one column: per-row destinations come from two exclusive scans (valid rows,
// valid string bytes), so every value is read once and written once.
#include <algorithm>

#include "hs_common.cuh"

// flag[i] = row i is valid; len[i] = bytes of its PLAIN image (0 for NULL rows)
template <typename TO>
__global__ void __launch_bounds__(256) k_pq_flags(KCol c, int64_t n, int fixed_w, uint8_t* __restrict__ flag,
                                                  uint32_t* __restrict__ len) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    const bool v = !c.has_nulls || kvalid(c.validity, i);
    flag[i] = v;
    if (len) {
      uint32_t l = 0;
      if (v) {
        if (fixed_w) l = (uint32_t)fixed_w;
        else {
          const TO* o = (const TO*)c.offsets;
          l = 4u + (uint32_t)(o[i + 1] - o[i]);
        }
      }
      len[i] = l;
    }
  }
}

// values at the page boundaries of the two scans (page_rows has npages + 1 entries; the entry for row n is the total)
__global__ void k_pq_pick(const uint32_t* __restrict__ vpos, const uint32_t* __restrict__ bpos, const uint64_t* vtotal,
                          const uint64_t* btotal, const int64_t* __restrict__ page_rows, int npages, int64_t n,
                          uint64_t* __restrict__ out) {
  int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p > npages) return;
  const int64_t r = page_rows[p];
  out[2 * p] = r < n ? vpos[r] : *vtotal;
  out[2 * p + 1] = bpos ? (r < n ? bpos[r] : *btotal) : 0;
}

struct PqPages {
  int32_t npages;
  const int64_t* page_rows;     // device, npages + 1
  const uint64_t* val_off;      // device, npages: byte offset of the page's values inside the output
  const uint64_t* def_off;      // device, npages: byte offset of the page's bit-packed def-level bytes (or ~0)
  const uint64_t* first;        // device, 2 * (npages + 1): (valid rows, valid bytes) before the page
};
__device__ __forceinline__ int pq_page_of(const PqPages& pg, int64_t row) {
  int lo = 0, hi = pg.npages;   // largest p with page_rows[p] <= row
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (pg.page_rows[mid] <= row) lo = mid; else hi = mid;
  }
  return lo;
}

// PLAIN values, fixed width w: value of valid row i -> val_off[p] + (vpos[i] - first_valid[p]) * w
__global__ void __launch_bounds__(256) k_pq_fixed(const uint8_t* __restrict__ data, int w, const uint8_t* __restrict__ flag,
                                                  const uint32_t* __restrict__ vpos, int64_t n, PqPages pg,
                                                  uint8_t* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    if (!flag[i]) continue;
    const int p = pq_page_of(pg, i);
    uint8_t* d = out + pg.val_off[p] + (uint64_t)(vpos[i] - pg.first[2 * p]) * w;
    const uint8_t* s = data + (size_t)i * w;
    // page bodies start at arbitrary byte offsets: copy by bytes unless both ends happen to be aligned
    if (w == 8 && (((uintptr_t)d) & 7) == 0) *(uint64_t*)d = *(const uint64_t*)s;
    else if (w == 4 && (((uintptr_t)d) & 3) == 0) *(uint32_t*)d = *(const uint32_t*)s;
    else for (int b = 0; b < w; b++) d[b] = s[b];
  }
}
// PLAIN booleans: bit (vpos[i] - first) of the page's value bytes; one thread per OUTPUT byte would need a
// search per bit, so set bits with atomics on zeroed output (booleans are rare and small)
__global__ void __launch_bounds__(256) k_pq_bool(const uint8_t* __restrict__ data, const uint8_t* __restrict__ flag,
                                                 const uint32_t* __restrict__ vpos, int64_t n, PqPages pg,
                                                 uint8_t* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    if (!flag[i] || !((data[i >> 3] >> (i & 7)) & 1)) continue;
    const int p = pq_page_of(pg, i);
    const uint64_t bit = vpos[i] - pg.first[2 * p];
    const uint64_t byte = pg.val_off[p] + (bit >> 3);
    atomicOr((unsigned int*)(out + (byte & ~(uint64_t)3)), (1u << (bit & 7)) << (8 * (byte & 3)));
  }
}
// PLAIN byte arrays: 4-byte LE length + bytes; 8 lanes per row
template <typename TO>
__global__ void __launch_bounds__(256) k_pq_bytes(const TO* __restrict__ off, const uint8_t* __restrict__ data,
                                                  const uint8_t* __restrict__ flag, const uint32_t* __restrict__ bpos,
                                                  int64_t n, PqPages pg, uint8_t* __restrict__ out) {
  const int sub = threadIdx.x & 7;
  int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 3;
  const int64_t stride = ((int64_t)gridDim.x * blockDim.x) >> 3;
  for (; i < n; i += stride) {
    if (!flag[i]) continue;
    const int p = pq_page_of(pg, i);
    const int64_t s0 = (int64_t)off[i];
    const uint32_t len = (uint32_t)((int64_t)off[i + 1] - s0);
    uint8_t* d = out + pg.val_off[p] + (uint64_t)(bpos[i] - pg.first[2 * p + 1]);
    if (sub < 4) d[sub] = (uint8_t)(len >> (8 * sub));
    const uint8_t* s = data + s0;
    for (uint32_t b = sub; b < len; b += 8) d[4 + b] = s[b];
  }
}
// definition levels of the pages that hold NULLs: byte g of the page = validity bits of its rows 8g .. 8g+7
__global__ void __launch_bounds__(256) k_pq_deflevels(const uint8_t* __restrict__ flag, int64_t n, PqPages pg,
                                                      uint8_t* __restrict__ out) {
  // one thread per (page, group) pair, pages walked by the grid's y dimension
  const int p = blockIdx.y;
  if (pg.def_off[p] == ~0ull) return;
  const int64_t r0 = pg.page_rows[p], r1 = pg.page_rows[p + 1];
  const int64_t groups = (r1 - r0 + 7) >> 3;
  for (int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; g < groups; g += (int64_t)gridDim.x * blockDim.x) {
    unsigned v = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
      const int64_t r = r0 + g * 8 + k;
      if (r < r1 && flag[r]) v |= 1u << k;
    }
    out[pg.def_off[p] + g] = (uint8_t)v;
  }
}

static int varint(uint64_t v, uint8_t* out) {
  int n = 0;
  while (v >= 0x80) { out[n++] = (uint8_t)(v | 0x80); v >>= 7; }
  out[n++] = (uint8_t)v;
  return n;
}

extern "C" int32_t hs_parquet_plain_pages(hs_ctx* ctx, const hs_column* col, int32_t optional, const int64_t* page_rows,
                                          int32_t npages, hs_table** out) {
  API_BEGIN(ctx)
  if (!out) HS_THROW(HS_EINVAL, "null out");
  *out = nullptr;
  if (!col || !page_rows || npages <= 0 || npages > 65535) HS_THROW(HS_EINVAL, "bad page list (1..65535 pages per call)");
  hs_stages_reset(ctx);
  DevCol dc;
  {
    Stage s(ctx, "h2d");
    hs_stage_column(ctx, *col, dc);
  }
  const int64_t n = dc.length;
  if (page_rows[0] != 0 || page_rows[npages] != n) HS_THROW(HS_EINVAL, "pages must cover the column");
  for (int p = 0; p < npages; p++)
    if (page_rows[p + 1] < page_rows[p]) HS_THROW(HS_EINVAL, "page boundaries must not decrease");
  if (!optional && dc.validity) HS_THROW(HS_EINVAL, "a REQUIRED column cannot hold NULLs");
  const bool varlen = hs_is_varlen(dc.type), large = hs_is_large(dc.type), is_bool = dc.type == HS_BOOL;
  const int w = hs_fixed_width(dc.type);
  if (varlen && dc.data_bytes + 4 * n >= 0xFFFFFFF0ll) HS_THROW(HS_EUNSUPPORTED, "more than 4 GiB of page data in one call");
  const KCol kc = hs_kcol(dc);
  const int64_t n1 = n > 0 ? n : 1;
  DBuf flag(ctx, (size_t)n1), vpos(ctx, (size_t)n1 * 4), vtotal(ctx, 8), len, bpos, btotal(ctx, 8);
  DBuf d_rows(ctx, ((size_t)npages + 1) * 8), picks(ctx, ((size_t)npages + 1) * 16);
  std::vector<uint64_t> h_first(2 * ((size_t)npages + 1));
  {
    Stage s(ctx, "parquet_plan");
    CK(cudaMemcpyAsync(d_rows.p, page_rows, ((size_t)npages + 1) * 8, cudaMemcpyHostToDevice, ctx->stream));
    const unsigned g = (unsigned)std::min<int64_t>((n1 + 255) / 256, (int64_t)HS_NUM_SMS * 16);
    if (varlen) { len.alloc(ctx, (size_t)n1 * 4); bpos.alloc(ctx, (size_t)n1 * 4); }
    if (n > 0) {
      if (large) LAUNCH(ctx, k_pq_flags<int64_t>, g, 256, 0, kc, n, 0, flag.as<uint8_t>(), varlen ? len.as<uint32_t>() : nullptr);
      else LAUNCH(ctx, k_pq_flags<int32_t>, g, 256, 0, kc, n, 0, flag.as<uint8_t>(), varlen ? len.as<uint32_t>() : nullptr);
    }
    hs_exclusive_scan_flags(ctx, flag.as<uint8_t>(), vpos.as<uint32_t>(), n, vtotal.as<uint64_t>());
    if (varlen) hs_exclusive_scan_u32(ctx, len.as<uint32_t>(), bpos.as<uint32_t>(), n, btotal.as<uint64_t>());
    else btotal.zero();
    LAUNCH(ctx, k_pq_pick, (npages + 256) / 256, 256, 0, vpos.as<uint32_t>(), varlen ? bpos.as<uint32_t>() : nullptr,
           vtotal.as<uint64_t>(), btotal.as<uint64_t>(), d_rows.as<int64_t>(), npages, n, picks.as<uint64_t>());
    CK(cudaMemcpyAsync(h_first.data(), picks.p, h_first.size() * 8, cudaMemcpyDeviceToHost, ctx->stream));
    hs_sync(ctx);
  }
  // ---- page layout (host): [def levels][values] per page, pages back to back
  std::vector<int64_t> page_off((size_t)npages + 1, 0), nulls((size_t)npages, 0);
  std::vector<uint64_t> val_off((size_t)npages), def_off((size_t)npages, ~0ull);
  std::vector<std::pair<uint64_t, std::vector<uint8_t>>> headers;   // (offset, bytes) written from the host
  uint64_t at = 0;
  for (int p = 0; p < npages; p++) {
    page_off[p] = (int64_t)at;
    const int64_t rows = page_rows[p + 1] - page_rows[p];
    const int64_t valid = (int64_t)(h_first[2 * (p + 1)] - h_first[2 * p]);
    nulls[p] = rows - valid;
    if (optional) {
      std::vector<uint8_t> h(4 + 12);
      int k;
      uint32_t body;
      if (nulls[p] == 0) {                       // one RLE run: varint(rows << 1), value 1
        k = varint((uint64_t)rows << 1, h.data() + 4);
        h[4 + k] = 1;
        k += 1;
        body = (uint32_t)k;
      } else {                                   // bit-packed: varint((groups << 1) | 1), then groups bytes
        const int64_t groups = (rows + 7) >> 3;
        k = varint(((uint64_t)groups << 1) | 1u, h.data() + 4);
        body = (uint32_t)(k + groups);
        def_off[p] = at + 4 + (uint64_t)k;
      }
      memcpy(h.data(), &body, 4);
      h.resize(4 + (size_t)k);
      headers.emplace_back(at, std::move(h));
      at += 4 + body;
    }
    val_off[p] = at;
    if (varlen) at += h_first[2 * (p + 1) + 1] - h_first[2 * p + 1];
    else if (is_bool) at += (uint64_t)((valid + 7) >> 3);
    else at += (uint64_t)valid * w;
  }
  page_off[npages] = (int64_t)at;
  if (at >= 0x7FFFFFF0ull) HS_THROW(HS_EUNSUPPORTED, "more than 2 GiB of page data in one call: encode fewer rows at a time");
  std::vector<OutCol> ocols(2);
  OutCol& body = ocols[0];
  body.type = HS_BINARY;
  body.length = npages;
  body.data_bytes = (int64_t)at;
  body.data.alloc(ctx, (size_t)at + 8);
  body.offsets.alloc(ctx, ((size_t)npages + 1) * 4);
  {
    Stage s(ctx, "parquet_encode");
    std::vector<int32_t> off32((size_t)npages + 1);
    for (int p = 0; p <= npages; p++) off32[p] = (int32_t)page_off[p];
    CK(cudaMemcpyAsync(body.offsets.p, off32.data(), off32.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
    if (is_bool) CK(cudaMemsetAsync(body.data.p, 0, (size_t)at + 8, ctx->stream));
    for (auto& h : headers)
      CK(cudaMemcpyAsync((uint8_t*)body.data.p + h.first, h.second.data(), h.second.size(), cudaMemcpyHostToDevice, ctx->stream));
    DBuf d_val(ctx, (size_t)npages * 8), d_def(ctx, (size_t)npages * 8);
    CK(cudaMemcpyAsync(d_val.p, val_off.data(), (size_t)npages * 8, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(d_def.p, def_off.data(), (size_t)npages * 8, cudaMemcpyHostToDevice, ctx->stream));
    PqPages pg{npages, d_rows.as<int64_t>(), d_val.as<uint64_t>(), d_def.as<uint64_t>(), picks.as<uint64_t>()};
    const unsigned g = (unsigned)std::min<int64_t>((n1 + 255) / 256, (int64_t)HS_NUM_SMS * 16);
    uint8_t* o = body.data.as<uint8_t>();
    if (n > 0) {
      if (varlen) {
        const unsigned g8 = (unsigned)std::min<int64_t>((n1 * 8 + 255) / 256, (int64_t)HS_NUM_SMS * 32);
        if (large) LAUNCH(ctx, k_pq_bytes<int64_t>, g8, 256, 0, (const int64_t*)dc.offsets, (const uint8_t*)dc.data, flag.as<uint8_t>(), bpos.as<uint32_t>(), n, pg, o);
        else LAUNCH(ctx, k_pq_bytes<int32_t>, g8, 256, 0, dc.offsets, (const uint8_t*)dc.data, flag.as<uint8_t>(), bpos.as<uint32_t>(), n, pg, o);
      } else if (is_bool) {
        LAUNCH(ctx, k_pq_bool, g, 256, 0, (const uint8_t*)dc.data, flag.as<uint8_t>(), vpos.as<uint32_t>(), n, pg, o);
      } else {
        LAUNCH(ctx, k_pq_fixed, g, 256, 0, (const uint8_t*)dc.data, w, flag.as<uint8_t>(), vpos.as<uint32_t>(), n, pg, o);
      }
      bool any_def = false;
      for (int p = 0; p < npages; p++) any_def |= def_off[p] != ~0ull;
      if (any_def) {
        dim3 grid(64, (unsigned)npages);
        k_pq_deflevels<<<grid, 256, 0, ctx->stream>>>(flag.as<uint8_t>(), n, pg, o);
        hs_count_launch(ctx);
        CK(cudaGetLastError());
      }
    }
    hs_sync(ctx);   // headers, val_off, def_off are host vectors: the copies above must have left them
  }
  OutCol& nn = ocols[1];
  nn.type = HS_INT64;
  nn.length = npages;
  nn.data_bytes = (int64_t)npages * 8;
  nn.data.alloc(ctx, (size_t)npages * 8);
  CK(cudaMemcpyAsync(nn.data.p, nulls.data(), (size_t)npages * 8, cudaMemcpyHostToDevice, ctx->stream));
  hs_sync(ctx);
  Stage s(ctx, "d2h");
  *out = hs_finish_table(ctx, ocols, npages, 0, nullptr);
  API_END(ctx)
}
