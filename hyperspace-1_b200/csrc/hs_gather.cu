// hs_gather.cu — gather rows of a column by a u32 row-id list (join/filter output,
// string columns of the build). Fixed-width values, validity/boolean bits, UTF8 bytes.
#include "hs_common.cuh"

template <typename T>
__global__ void __launch_bounds__(256) k_gather_fixed(const T* __restrict__ in, const uint32_t* __restrict__ idx,
                                                      int64_t m, T* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < m; i += stride) {
    const uint32_t s = idx[i];
    out[i] = s == HS_MISS ? T(0) : in[s];
  }
}

// one thread builds one output byte (8 rows); counts set bits into *ones
__global__ void __launch_bounds__(256) k_gather_bits(const uint8_t* __restrict__ in, const uint32_t* __restrict__ idx,
                                                     int64_t m, uint8_t* __restrict__ out,
                                                     unsigned long long* ones) {
  int64_t nbytes = (m + 7) >> 3;
  int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  unsigned cnt = 0;
  for (; b < nbytes; b += stride) {
    unsigned v = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
      int64_t i = b * 8 + k;
      if (i < m) {
        uint32_t s = idx[i];
        if (s != HS_MISS) v |= (in ? ((in[s >> 3] >> (s & 7)) & 1u) : 1u) << k;   // in == NULL: all rows valid
      }
    }
    out[b] = (uint8_t)v;
    cnt += __popc(v);
  }
  for (int d = 16; d; d >>= 1) cnt += __shfl_down_sync(0xffffffffu, cnt, d);
  if (lane_id() == 0 && cnt && ones) atomicAdd(ones, (unsigned long long)cnt);
}

// length and source start of every output row: the one random read of the source offsets; the byte mover below
// then reads its row table sequentially
template <typename TO>
__global__ void __launch_bounds__(256) k_gather_lens(const TO* __restrict__ off, const uint32_t* __restrict__ idx,
                                                     int64_t m, uint32_t* __restrict__ lens, TO* __restrict__ starts) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < m; i += stride) {
    uint32_t s = idx[i];
    TO a = 0, b = 0;
    if (s != HS_MISS) { a = off[s]; b = off[s + 1]; }
    lens[i] = (uint32_t)(b - a);
    starts[i] = a;
  }
}

// String bytes of `rows` consecutive OUTPUT rows are contiguous in the output: a CTA collects them in shared
// memory and then writes the whole span with 16-byte stores (writing string pieces straight to global memory
// costs a partial-sector write per piece). Per tile:
//   1. the row table (source start from k_gather_lens, destination start), one row per thread;
//   2. 16 lanes fetch one string as aligned 4-byte words (64 bytes per step, four strings in flight per
//      group) and drop the bytes at their place in the staging buffer (measured: faster than cp.async words
//      into a raw area + a repack, 6.4 vs 8.4 ms for the whole finalize step of 80 M rows);
//   3. packed span -> global memory, 16 bytes per store.
// Tiles whose bytes do not fit (long strings) write the bytes to global memory directly.
#define GB_MAX_ROWS 256
#define GB_CAP (24 * 1024)
#define GB_UNROLL 4
template <typename TO, typename TN>
__global__ void __launch_bounds__(256) k_gather_bytes(const TO* __restrict__ starts, const uint8_t* __restrict__ in,
                                                      int64_t m, int rows,
                                                      const TN* __restrict__ new_off, uint8_t* __restrict__ out) {
  __shared__ __align__(16) uint8_t s_buf[GB_CAP + 16];
  __shared__ int64_t s_src[GB_MAX_ROWS];
  __shared__ int64_t s_dst[GB_MAX_ROWS + 1];
  const int sub = threadIdx.x & 15, grp = threadIdx.x >> 4;
  const int64_t ntiles = (m + rows - 1) / rows;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t r0 = tile * rows;
    const int nrows = (int)((r0 + rows < m) ? rows : m - r0);
    if ((int)threadIdx.x < nrows) {
      s_src[threadIdx.x] = (int64_t)starts[r0 + threadIdx.x];
      s_dst[threadIdx.x] = (int64_t)new_off[r0 + threadIdx.x];
    }
    if (threadIdx.x == 0) s_dst[nrows] = (int64_t)new_off[r0 + nrows];
    __syncthreads();
    const int64_t b0 = s_dst[0], bytes = s_dst[nrows] - b0;
    const bool staged = bytes <= GB_CAP;
    const uint32_t pad = (uint32_t)((uintptr_t)(out + b0) & 15);
    uint8_t* const dbase = staged ? s_buf + pad : out + b0;
    for (int rb = grp; rb < nrows; rb += 16 * GB_UNROLL) {
      uint32_t w[GB_UNROLL];
      int len[GB_UNROLL], k0[GB_UNROLL];   // k0 = row byte index of this lane's first fetched byte (may be < 0)
      const uint8_t* p[GB_UNROLL];
#pragma unroll
      for (int u = 0; u < GB_UNROLL; u++) {
        const int r = rb + u * 16;
        len[u] = 0;
        k0[u] = 0;
        p[u] = in;
        w[u] = 0;
        if (r < nrows) {
          p[u] = in + s_src[r];
          len[u] = (int)(s_dst[r + 1] - s_dst[r]);
          k0[u] = 4 * sub - (int)((uintptr_t)p[u] & 3);
          if (k0[u] < len[u]) w[u] = *(const uint32_t*)(p[u] + k0[u]);
        }
      }
#pragma unroll
      for (int u = 0; u < GB_UNROLL; u++) {
        const int r = rb + u * 16;
        if (r >= nrows) continue;
        uint8_t* d = dbase + (s_dst[r] - b0);
        int k = k0[u];
        uint32_t word = w[u];
        while (k < len[u]) {
#pragma unroll
          for (int j = 0; j < 4; j++)
            if (k + j >= 0 && k + j < len[u]) d[k + j] = (uint8_t)(word >> (8 * j));
          k += 64;
          if (k < len[u]) word = *(const uint32_t*)(p[u] + k);
        }
      }
    }
    if (staged) {
      __syncthreads();
      uint8_t* g = out + b0 - pad;                       // 16-byte aligned
      const uint32_t total = pad + (uint32_t)bytes;
      for (uint32_t lo = threadIdx.x * 16; lo < total; lo += 256 * 16) {
        const uint32_t hi = lo + 16;
        if (lo >= pad && hi <= total) *(uint4*)(g + lo) = *(const uint4*)(s_buf + lo);
        else
          for (uint32_t b = lo < pad ? pad : lo; b < (hi < total ? hi : total); b++) g[b] = s_buf[b];
      }
    }
    __syncthreads();
  }
}

static unsigned grid_for(int64_t n, int threads = 256) {
  int64_t g = (n + threads - 1) / threads;
  int64_t cap = (int64_t)HS_NUM_SMS * 32;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (unsigned)g;
}

static void gather_validity(hs_ctx* c, const DevCol& src, const uint32_t* idx, int64_t m, OutCol& out, bool may_miss) {
  if (!src.validity && !may_miss) { out.null_count = 0; return; }
  out.validity.alloc(c, (size_t)((m + 7) / 8));
  DBuf ones(c, 8);
  ones.zero();
  LAUNCH(c, k_gather_bits, grid_for((m + 7) / 8), 256, 0, src.validity, idx, m, out.validity.as<uint8_t>(),
         ones.as<unsigned long long>());
  unsigned long long h = 0;
  CK(cudaMemcpyAsync(&h, ones.p, 8, cudaMemcpyDeviceToHost, c->stream));
  hs_sync(c);
  out.null_count = m - (int64_t)h;
  if (out.null_count == 0) out.validity.release();
}

void hs_gather_column(hs_ctx* c, const DevCol& src, const uint32_t* idx, int64_t m, OutCol& out, bool may_miss) {
  out.type = src.type;
  out.length = m;
  gather_validity(c, src, idx, m, out, may_miss);
  if (hs_is_varlen(src.type)) {
    const bool in_large = hs_is_large(src.type);
    DBuf lens(c, (size_t)(m > 0 ? m : 1) * 4), starts(c, (size_t)(m > 0 ? m : 1) * (in_large ? 8 : 4));
    if (m > 0) {
      if (in_large)
        LAUNCH(c, k_gather_lens<int64_t>, grid_for(m), 256, 0, (const int64_t*)src.offsets, idx, m, lens.as<uint32_t>(), starts.as<int64_t>());
      else
        LAUNCH(c, k_gather_lens<int32_t>, grid_for(m), 256, 0, src.offsets, idx, m, lens.as<uint32_t>(), starts.as<int32_t>());
    }
    // rows can repeat (join output): the gathered bytes may pass 2 GiB even when the source did not
    const int64_t h = hs_offsets_from_lens(c, lens.as<uint32_t>(), m, src.type, out);
    const bool out_large = hs_is_large(out.type);
    out.data_bytes = h;
    out.data.alloc(c, (size_t)h);
    if (m > 0 && h > 0) {
      // rows per tile: as many as keep the expected tile inside the staging buffer (3/4 full on average)
      int rows = GB_MAX_ROWS;
      while (rows > 8 && (h / m + 1) * rows > GB_CAP * 3 / 4) rows >>= 1;
      const unsigned g = (unsigned)std::min<int64_t>((m + rows - 1) / rows, (int64_t)HS_NUM_SMS * 32);
      const uint8_t* in = (const uint8_t*)src.data;
      uint8_t* o = out.data.as<uint8_t>();
      if (in_large)   // a LARGE source always gives a LARGE result
        LAUNCH(c, (k_gather_bytes<int64_t, int64_t>), g, 256, 0, starts.as<int64_t>(), in, m, rows, out.offsets.as<int64_t>(), o);
      else if (out_large)
        LAUNCH(c, (k_gather_bytes<int32_t, int64_t>), g, 256, 0, starts.as<int32_t>(), in, m, rows, out.offsets.as<int64_t>(), o);
      else
        LAUNCH(c, (k_gather_bytes<int32_t, int32_t>), g, 256, 0, starts.as<int32_t>(), in, m, rows, out.offsets.as<int32_t>(), o);
    }
  } else if (src.type == HS_BOOL) {
    out.data_bytes = (m + 7) / 8;
    out.data.alloc(c, (size_t)out.data_bytes);
    if (m > 0)
      LAUNCH(c, k_gather_bits, grid_for((m + 7) / 8), 256, 0, (const uint8_t*)src.data, idx, m,
             out.data.as<uint8_t>(), (unsigned long long*)nullptr);
  } else {
    int w = hs_fixed_width(src.type);
    out.data_bytes = m * w;
    out.data.alloc(c, (size_t)out.data_bytes);
    if (m > 0) {
      unsigned g = grid_for(m);
      switch (w) {
        case 1: LAUNCH(c, k_gather_fixed<uint8_t>, g, 256, 0, (const uint8_t*)src.data, idx, m, out.data.as<uint8_t>()); break;
        case 2: LAUNCH(c, k_gather_fixed<uint16_t>, g, 256, 0, (const uint16_t*)src.data, idx, m, out.data.as<uint16_t>()); break;
        case 4: LAUNCH(c, k_gather_fixed<uint32_t>, g, 256, 0, (const uint32_t*)src.data, idx, m, out.data.as<uint32_t>()); break;
        default: LAUNCH(c, k_gather_fixed<uint64_t>, g, 256, 0, (const uint64_t*)src.data, idx, m, out.data.as<uint64_t>()); break;
      }
    }
  }
}

void hs_copy_column(hs_ctx* c, const DevCol& src, OutCol& out) {
  out.type = src.type;
  out.length = src.length;
  out.null_count = src.null_count;
  int64_t n = src.length;
  if (src.validity) {
    out.validity.alloc(c, (size_t)((n + 7) / 8));
    CK(cudaMemcpyAsync(out.validity.p, src.validity, (size_t)((n + 7) / 8), cudaMemcpyDeviceToDevice, c->stream));
  }
  if (hs_is_varlen(src.type)) {
    out.offsets.alloc(c, (size_t)(n + 1) * hs_offset_width(src.type));
    CK(cudaMemcpyAsync(out.offsets.p, src.offsets, (size_t)(n + 1) * hs_offset_width(src.type), cudaMemcpyDeviceToDevice, c->stream));
  }
  out.data_bytes = src.data_bytes;
  out.data.alloc(c, (size_t)src.data_bytes);
  if (src.data_bytes)
    CK(cudaMemcpyAsync(out.data.p, src.data, (size_t)src.data_bytes, cudaMemcpyDeviceToDevice, c->stream));
}
