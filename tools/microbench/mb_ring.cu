// mb_ring.cu — does a scratch ring that is re-used stay in B200's L2 while two streams pass through it?
// Emulates the "MSD level writes a ring, the leaf sort reads it back" plan of DESIGN.md: every CTA copies
// chunks src -> its private ring slot -> dst. With a ring the slot is overwritten by the next chunk (the lines
// can die in L2); without (ring = 0) the intermediate is a full-size buffer (every line is written back).
//   mb_ring <total_MB> <chunk_KB> <slots_per_cta> <hint: 0 none, 1 evict_first on the streams + evict_last on the ring>
// Prints ms and GB/s of algorithmic traffic (src read + dst write) for both variants.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ uint64_t pol_first() { uint64_t p; asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p)); return p; }
__device__ __forceinline__ uint64_t pol_last() { uint64_t p; asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p)); return p; }
__device__ __forceinline__ uint4 ld_h(const uint4* a, uint64_t p) {
  uint4 v;
  asm volatile("ld.global.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(a), "l"(p));
  return v;
}
__device__ __forceinline__ void st_h(uint4* a, uint4 v, uint64_t p) {
  asm volatile("st.global.L2::cache_hint.v4.u32 [%0], {%1,%2,%3,%4}, %5;" ::"l"(a), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "l"(p) : "memory");
}

template <int HINT>
__global__ void __launch_bounds__(512) k_ring(const uint4* __restrict__ src, uint4* __restrict__ dst, uint4* mid,
                                              size_t chunk_v, size_t nchunks, int slots, int use_ring) {
  const uint64_t pf = pol_first(), pl = pol_last();
  size_t it = 0;
  for (size_t c = blockIdx.x; c < nchunks; c += gridDim.x, it++) {
    const uint4* s = src + c * chunk_v;
    uint4* d = dst + c * chunk_v;
    uint4* m = use_ring ? mid + ((size_t)blockIdx.x * slots + (it % slots)) * chunk_v : mid + c * chunk_v;
    for (size_t i = threadIdx.x; i < chunk_v; i += blockDim.x) {
      uint4 v = HINT ? ld_h(s + i, pf) : s[i];
      v.x += 1;
      if (HINT) st_h(m + i, v, pl); else m[i] = v;
    }
    __syncthreads();
    __threadfence();
    // read back in a permuted order (another thread's data), as a sort would
    for (size_t i = threadIdx.x; i < chunk_v; i += blockDim.x) {
      size_t j = (i * 33 + 7) % chunk_v;
      uint4 v = HINT ? ld_h(m + j, pl) : m[j];
      if (HINT) st_h(d + i, v, pf); else d[i] = v;
    }
    __syncthreads();
  }
}

int main(int argc, char** argv) {
  size_t total_mb = argc > 1 ? atol(argv[1]) : 2400;
  size_t chunk_kb = argc > 2 ? atol(argv[2]) : 96;
  int slots = argc > 3 ? atoi(argv[3]) : 1;
  int hint = argc > 4 ? atoi(argv[4]) : 0;
  int grid = 148 * 2;
  size_t chunk_v = chunk_kb * 1024 / 16, nchunks = total_mb * 1024 * 1024 / (chunk_kb * 1024);
  size_t bytes = nchunks * chunk_v * 16;
  uint4 *src, *dst, *mid;
  CK(cudaMalloc(&src, bytes)); CK(cudaMalloc(&dst, bytes)); CK(cudaMalloc(&mid, bytes));
  CK(cudaMemset(src, 1, bytes)); CK(cudaMemset(mid, 0, bytes));
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int ring = 0; ring < 2; ring++) {
    float best = 1e9;
    for (int rep = 0; rep < 4; rep++) {
      cudaEventRecord(e0);
      if (hint) k_ring<1><<<grid, 512>>>(src, dst, mid, chunk_v, nchunks, slots, ring);
      else k_ring<0><<<grid, 512>>>(src, dst, mid, chunk_v, nchunks, slots, ring);
      cudaEventRecord(e1);
      CK(cudaDeviceSynchronize());
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      if (rep && ms < best) best = ms;
    }
    printf("total %zu MB chunk %zu KB slots %d hint %d ring %d (ring footprint %.1f MB): %.3f ms, %.0f GB/s algorithmic (src read + dst write)\n",
           total_mb, chunk_kb, slots, hint, ring, ring ? grid * slots * chunk_kb / 1024.0 : (double)total_mb, best,
           2.0 * bytes / best / 1e6);
  }
  return 0;
}
