// microbenchmark: one-pass scatter into (bucket, digit) bins with a global atomic per row and direct 8-byte stores
#include <cstdio>
#include <cstdint>
#include <vector>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e)); return 1; } } while (0)
__device__ __forceinline__ uint64_t mix(uint64_t x) { x ^= x >> 33; x *= 0xff51afd7ed558ccdull; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ull; x ^= x >> 33; return x; }
__global__ void k_gen(uint64_t* keys, uint16_t* bucket, uint64_t* v1, uint64_t* v2, int64_t n, uint64_t range, int nb, int skew) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint64_t k = mix(i) % range;
  if (skew) k = k % skew;
  keys[i] = k; bucket[i] = (uint16_t)(mix(k + 77) % nb); v1[i] = i; v2[i] = ~i;
}
__device__ __forceinline__ uint32_t digit(uint64_t k, uint32_t mult, uint32_t D) { uint32_t d = (uint32_t)((k * mult) >> 32); return d >= D ? D - 1 : d; }
__global__ void k_hist(const uint64_t* keys, const uint16_t* bucket, int64_t n, uint32_t mult, uint32_t D, uint32_t* hist) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  atomicAdd(&hist[bucket[i] * D + digit(keys[i], mult, D)], 1u);
}
template <int R>
__global__ void __launch_bounds__(256) k_scatter(const uint64_t* __restrict__ keys, const uint16_t* __restrict__ bucket,
                                                 const uint64_t* __restrict__ v1, const uint64_t* __restrict__ v2, int64_t n,
                                                 uint32_t mult, uint32_t D, uint32_t* cursor, uint64_t* ok, uint64_t* o1, uint64_t* o2) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < n; i0 += stride * R) {
    uint64_t k[R], a[R], b[R]; uint32_t pos[R];
#pragma unroll
    for (int r = 0; r < R; r++) { int64_t i = i0 + r * stride; if (i < n) { k[r] = keys[i]; pos[r] = bucket[i]; } }
#pragma unroll
    for (int r = 0; r < R; r++) { int64_t i = i0 + r * stride; if (i < n) pos[r] = atomicAdd(&cursor[pos[r] * D + digit(k[r], mult, D)], 1u); }
#pragma unroll
    for (int r = 0; r < R; r++) { int64_t i = i0 + r * stride; if (i < n) { a[r] = v1[i]; b[r] = v2[i]; } }
#pragma unroll
    for (int r = 0; r < R; r++) { int64_t i = i0 + r * stride; if (i < n) { ok[pos[r]] = k[r]; o1[pos[r]] = a[r]; o2[pos[r]] = b[r]; } }
  }
}
int main(int argc, char** argv) {
  int64_t n = 100000000; int nb = 200; uint64_t range = 600000000ull;
  uint64_t *keys, *v1, *v2, *ok, *o1, *o2; uint16_t* bucket; uint32_t *hist, *cursor;
  CK(cudaMalloc(&keys, n * 8)); CK(cudaMalloc(&v1, n * 8)); CK(cudaMalloc(&v2, n * 8));
  CK(cudaMalloc(&ok, n * 8)); CK(cudaMalloc(&o1, n * 8)); CK(cudaMalloc(&o2, n * 8)); CK(cudaMalloc(&bucket, n * 2));
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int skew : {0, 1000000, 1000, 1}) {
    k_gen<<<(unsigned)((n + 255) / 256), 256>>>(keys, bucket, v1, v2, n, range, nb, skew);
    for (uint32_t D : {64u, 179u, 256u}) {
      uint32_t nbins = nb * D;
      uint32_t mult = (uint32_t)(((uint64_t)D << 32) / range);
      CK(cudaMalloc(&hist, nbins * 4)); CK(cudaMalloc(&cursor, nbins * 4));
      CK(cudaMemset(hist, 0, nbins * 4));
      k_hist<<<(unsigned)((n + 255) / 256), 256>>>(keys, bucket, n, mult, D, hist);
      std::vector<uint32_t> h(nbins), c(nbins);
      CK(cudaMemcpy(h.data(), hist, nbins * 4, cudaMemcpyDeviceToHost));
      uint32_t acc = 0, mxb = 0; for (uint32_t i = 0; i < nbins; i++) { c[i] = acc; acc += h[i]; if (h[i] > mxb) mxb = h[i]; }
      for (int variant = 0; variant < 6; variant++) {
        int R = variant % 3 == 0 ? 1 : variant % 3 == 1 ? 2 : 4;
        int grid = variant < 3 ? 148 * 8 : 148 * 32;
        float best = 1e9;
        for (int rep = 0; rep < 3; rep++) {
          CK(cudaMemcpy(cursor, c.data(), nbins * 4, cudaMemcpyHostToDevice));
          cudaEventRecord(e0);
          if (R == 1) k_scatter<1><<<grid, 256>>>(keys, bucket, v1, v2, n, mult, D, cursor, ok, o1, o2);
          if (R == 2) k_scatter<2><<<grid, 256>>>(keys, bucket, v1, v2, n, mult, D, cursor, ok, o1, o2);
          if (R == 4) k_scatter<4><<<grid, 256>>>(keys, bucket, v1, v2, n, mult, D, cursor, ok, o1, o2);
          cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
          float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
        }
        printf("skew=%d D=%u maxbin=%u R=%d grid=%d  %.3f ms\n", skew, D, mxb, R, grid, best);
      }
      cudaFree(hist); cudaFree(cursor);
      if (skew == 1) break;
    }
  }
  return 0;
}
