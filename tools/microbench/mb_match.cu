#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
// microbenchmark: per-warp-row cost of (a) ballot loop match on d bits, (b) __match_any_sync, (c) smem atomicAdd
template <int MODE, int BITS>
__global__ void k(const uint32_t* in, uint32_t* out, int iters) {
  __shared__ uint32_t cnt[8 * 1024];
  for (int i = threadIdx.x; i < 8 * 1024; i += blockDim.x) cnt[i] = 0;
  __syncthreads();
  uint32_t x = in[blockIdx.x * blockDim.x + threadIdx.x];
  uint32_t acc = 0;
  const uint32_t lt = (1u << (threadIdx.x & 31)) - 1;
  uint32_t* my = cnt + (threadIdx.x >> 5) * 1024;
  for (int it = 0; it < iters; it++) {
    x = x * 1664525u + 1013904223u;
    uint32_t d = (x >> 7) & ((1u << BITS) - 1);
    if (MODE == 0) {
      uint32_t m = 0xffffffffu;
#pragma unroll
      for (int b = 0; b < BITS; b++) {
        bool bit = (d >> b) & 1;
        uint32_t bal = __ballot_sync(0xffffffffu, bit);
        m &= bit ? bal : ~bal;
      }
      uint32_t r = __popc(m & lt);
      uint32_t pre = my[d];
      __syncwarp();
      if (r == 0) my[d] = pre + __popc(m);
      __syncwarp();
      acc += pre + r;
    } else if (MODE == 1) {
      uint32_t m = __match_any_sync(0xffffffffu, d);
      uint32_t r = __popc(m & lt);
      uint32_t pre = my[d];
      __syncwarp();
      if (r == 0) my[d] = pre + __popc(m);
      __syncwarp();
      acc += pre + r;
    } else if (MODE == 2) {
      acc += atomicAdd(&cnt[d], 1u);
    } else if (MODE == 3) {
      atomicAdd(&cnt[d], 1u);          // result unused: a reduction, no value comes back
    } else {
      // MODE 4: two 4-bit ballot passes with the per-digit counts kept in registers (no shared-memory counters):
      // what a tile ranking by ballots alone would issue per row for an 8-bit digit
      uint32_t r = 0;
#pragma unroll
      for (int half = 0; half < 2; half++) {
        const uint32_t dd = (d >> (4 * half)) & 15u;
        uint32_t m = 0xffffffffu;
#pragma unroll
        for (int b = 0; b < 4; b++) {
          const bool bit = (dd >> b) & 1;
          const uint32_t bal = __ballot_sync(0xffffffffu, bit);
          m &= bit ? bal : ~bal;
        }
        r += __popc(m & lt);
      }
      acc += r;
    }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
template <int MODE, int BITS> void run(const char* name, uint32_t* in, uint32_t* out) {
  int iters = 2000, grid = 148 * 4, block = 256;
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  k<MODE, BITS><<<grid, block>>>(in, out, 10);
  cudaEventRecord(a);
  k<MODE, BITS><<<grid, block>>>(in, out, iters);
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  double items = (double)grid * block * iters;
  printf("%-28s %8.3f ms  %7.1f Gitems/s  (100M rows -> %.3f ms)\n", name, ms, items / ms / 1e6, 1e8 / (items / ms));
}
int main() {
  uint32_t *in, *out; int n = 148 * 4 * 256;
  cudaMalloc(&in, n * 4); cudaMalloc(&out, n * 4);
  uint32_t* h = new uint32_t[n]; for (int i = 0; i < n; i++) h[i] = i * 2654435761u;
  cudaMemcpy(in, h, n * 4, cudaMemcpyHostToDevice);
  run<0, 8>("ballot loop 8 bits", in, out);
  run<0, 10>("ballot loop 10 bits", in, out);
  run<0, 5>("ballot loop 5 bits", in, out);
  run<1, 8>("match_any 8 bits", in, out);
  run<1, 10>("match_any 10 bits", in, out);
  run<2, 8>("smem atomicAdd 256 bins", in, out);
  run<2, 10>("smem atomicAdd 1024 bins", in, out);
  run<3, 8>("smem atomicAdd, no return, 256", in, out);
  run<3, 10>("smem atomicAdd, no return, 1024", in, out);
  run<4, 8>("2 x 4-bit ballots, registers", in, out);
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
