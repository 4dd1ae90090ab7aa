// hs_gather.cu — gather rows of a column by a u32 row-id list (join/filter output,
// string columns of the build). Fixed-width values, validity/boolean bits, UTF8 bytes.
#include "hs_common.cuh"

template <typename T>
__global__ void __launch_bounds__(256) k_gather_fixed(const T* __restrict__ in, const uint32_t* __restrict__ idx,
                                                      int64_t m, T* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < m; i += stride) out[i] = in[idx[i]];
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
        v |= ((in[s >> 3] >> (s & 7)) & 1u) << k;
      }
    }
    out[b] = (uint8_t)v;
    cnt += __popc(v);
  }
  for (int d = 16; d; d >>= 1) cnt += __shfl_down_sync(0xffffffffu, cnt, d);
  if (lane_id() == 0 && cnt && ones) atomicAdd(ones, (unsigned long long)cnt);
}

__global__ void __launch_bounds__(256) k_gather_lens(const int32_t* __restrict__ off, const uint32_t* __restrict__ idx,
                                                     int64_t m, uint32_t* __restrict__ lens) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < m; i += stride) {
    uint32_t s = idx[i];
    lens[i] = (uint32_t)(off[s + 1] - off[s]);
  }
}

// 8 lanes cooperate on one string: short strings (mean ~24 B) finish in one round of 4-byte moves
__global__ void __launch_bounds__(256) k_gather_bytes(const int32_t* __restrict__ off, const uint8_t* __restrict__ in,
                                                      const uint32_t* __restrict__ idx, int64_t m,
                                                      const int32_t* __restrict__ new_off, uint8_t* __restrict__ out) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = ((int64_t)gridDim.x * blockDim.x) >> 3;
  int sub = threadIdx.x & 7;
  for (int64_t i = t >> 3; i < m; i += stride) {
    uint32_t s = idx[i];
    const uint8_t* src = in + off[s];
    int len = off[s + 1] - off[s];
    uint8_t* dst = out + new_off[i];
    for (int b = sub; b < len; b += 8) dst[b] = src[b];
  }
}

__global__ void k_set_last_offset(int32_t* off, int64_t m, const uint64_t* total) { off[m] = (int32_t)*total; }

static unsigned grid_for(int64_t n, int threads = 256) {
  int64_t g = (n + threads - 1) / threads;
  int64_t cap = (int64_t)HS_NUM_SMS * 32;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (unsigned)g;
}

static void gather_validity(hs_ctx* c, const DevCol& src, const uint32_t* idx, int64_t m, OutCol& out) {
  if (!src.validity) { out.null_count = 0; return; }
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

void hs_gather_column(hs_ctx* c, const DevCol& src, const uint32_t* idx, int64_t m, OutCol& out) {
  out.type = src.type;
  out.length = m;
  gather_validity(c, src, idx, m, out);
  if (hs_is_varlen(src.type)) {
    out.offsets.alloc(c, (size_t)(m + 1) * 4);
    DBuf total(c, 8);
    if (m > 0) LAUNCH(c, k_gather_lens, grid_for(m), 256, 0, src.offsets, idx, m, out.offsets.as<uint32_t>());
    hs_exclusive_scan_u32(c, out.offsets.as<uint32_t>(), out.offsets.as<uint32_t>(), m, total.as<uint64_t>());
    LAUNCH(c, k_set_last_offset, 1, 1, 0, out.offsets.as<int32_t>(), m, total.as<uint64_t>());
    uint64_t h = 0;
    CK(cudaMemcpyAsync(&h, total.p, 8, cudaMemcpyDeviceToHost, c->stream));
    hs_sync(c);
    if (h > 0x7fffffffull) HS_THROW(HS_EUNSUPPORTED, "string column exceeds 2 GiB (int32 offsets)");
    out.data_bytes = (int64_t)h;
    out.data.alloc(c, (size_t)h);
    if (m > 0 && h > 0)
      LAUNCH(c, k_gather_bytes, grid_for(m * 8), 256, 0, src.offsets, (const uint8_t*)src.data, idx, m,
             out.offsets.as<int32_t>(), out.data.as<uint8_t>());
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
    out.offsets.alloc(c, (size_t)(n + 1) * 4);
    CK(cudaMemcpyAsync(out.offsets.p, src.offsets, (size_t)(n + 1) * 4, cudaMemcpyDeviceToDevice, c->stream));
  }
  out.data_bytes = src.data_bytes;
  out.data.alloc(c, (size_t)src.data_bytes);
  if (src.data_bytes)
    CK(cudaMemcpyAsync(out.data.p, src.data, (size_t)src.data_bytes, cudaMemcpyDeviceToDevice, c->stream));
}
