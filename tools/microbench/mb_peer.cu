// mb_peer.cu — how fast can SMs of GPU 0 write into GPU 1's memory over NVLink, by store pattern?
//   (a) one 8-byte store per thread, warps writing 256 contiguous bytes (what k_partition<2> does per run)
//   (b) one 16-byte store per thread
//   (c) TMA bulk stores shared -> peer global (cp.async.bulk.global.shared::cta) of `chunk` bytes per CTA step
// Each CTA stages its data in shared memory first (as the scatter kernel does), runs of `run` bytes go to
// pseudo-random places of the destination (run-sized slots), like bucket/owner runs.
//   mb_peer <total_MB> <run_bytes>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int MODE>
__global__ void __launch_bounds__(512) k_push(const uint4* __restrict__ src, unsigned char* dst, size_t nruns, uint32_t run) {
  extern __shared__ __align__(128) unsigned char sm[];
  const uint32_t tile = 32768;                       // bytes staged per step
  const uint32_t runs_per_tile = tile / run;
  for (size_t t = blockIdx.x; t * runs_per_tile < nruns; t += gridDim.x) {
    // stage (local HBM -> shared)
    for (uint32_t i = threadIdx.x; i < tile / 16; i += blockDim.x) ((uint4*)sm)[i] = src[t * (tile / 16) + i];
    __syncthreads();
    if (MODE == 2) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      if (threadIdx.x < runs_per_tile && threadIdx.x < 32 * 16) {
        for (uint32_t r = threadIdx.x; r < runs_per_tile; r += blockDim.x) {
          size_t slot = ((t * runs_per_tile + r) * 2654435761ull) % nruns;
          asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst + slot * run),
                       "r"(smem_u32(sm + r * run)), "r"(run) : "memory");
        }
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      }
    } else {
      const uint32_t w = MODE == 0 ? 8 : 16;
      for (uint32_t i = threadIdx.x; i < tile / w; i += blockDim.x) {
        uint32_t byte = i * w, r = byte / run, o = byte % run;
        size_t slot = ((t * runs_per_tile + r) * 2654435761ull) % nruns;
        if (MODE == 0) *(uint64_t*)(dst + slot * run + o) = *(const uint64_t*)(sm + byte);
        else *(uint4*)(dst + slot * run + o) = *(const uint4*)(sm + byte);
      }
    }
    __syncthreads();
  }
}

int main(int argc, char** argv) {
  size_t total_mb = argc > 1 ? atol(argv[1]) : 1024;
  uint32_t run = argc > 2 ? atoi(argv[2]) : 4096;
  int nd = 0;
  CK(cudaGetDeviceCount(&nd));
  size_t bytes = total_mb << 20, nruns = bytes / run;
  uint4* src; unsigned char *dloc, *dpeer = nullptr;
  CK(cudaSetDevice(0));
  CK(cudaMalloc(&src, bytes)); CK(cudaMalloc(&dloc, bytes)); CK(cudaMemset(src, 1, bytes));
  if (nd > 1) {
    CK(cudaSetDevice(1)); CK(cudaMalloc(&dpeer, bytes)); CK(cudaSetDevice(0));
    CK(cudaDeviceEnablePeerAccess(1, 0));
  }
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  CK(cudaFuncSetAttribute(k_push<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768));
  CK(cudaFuncSetAttribute(k_push<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768));
  CK(cudaFuncSetAttribute(k_push<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768));
  for (int peer = 0; peer < (dpeer ? 2 : 1); peer++) {
    unsigned char* d = peer ? dpeer : dloc;
    for (int mode = 0; mode < 3; mode++) {
      for (int grid = 148; grid <= 148 * 4; grid *= 2) {
        float best = 1e9;
        for (int rep = 0; rep < 4; rep++) {
          cudaEventRecord(e0);
          if (mode == 0) k_push<0><<<grid, 512, 32768>>>(src, d, nruns, run);
          else if (mode == 1) k_push<1><<<grid, 512, 32768>>>(src, d, nruns, run);
          else k_push<2><<<grid, 512, 32768>>>(src, d, nruns, run);
          cudaEventRecord(e1);
          CK(cudaDeviceSynchronize());
          float ms; cudaEventElapsedTime(&ms, e0, e1);
          if (rep && ms < best) best = ms;
        }
        printf("%s run %u B mode %s grid %d: %.3f ms = %.0f GB/s\n", peer ? "PEER " : "LOCAL", run,
               mode == 0 ? "st.8" : mode == 1 ? "st.16" : "bulk", grid, best, bytes / best / 1e6);
      }
    }
  }
  if (dpeer) {
    float best = 1e9;
    for (int rep = 0; rep < 4; rep++) {
      cudaEventRecord(e0); CK(cudaMemcpyPeerAsync(dpeer, 1, src, 0, bytes, 0)); cudaEventRecord(e1); CK(cudaDeviceSynchronize());
      float ms; cudaEventElapsedTime(&ms, e0, e1); if (rep && ms < best) best = ms;
    }
    printf("cudaMemcpyPeerAsync: %.3f ms = %.0f GB/s\n", best, bytes / best / 1e6);
  }
  if (dpeer) {
    // both directions at once (what the exchange does: every GPU stores into its peers while they store into it)
    uint4* src1; unsigned char* d0;
    CK(cudaSetDevice(1)); CK(cudaMalloc(&src1, bytes)); CK(cudaMemset(src1, 2, bytes));
    CK(cudaDeviceEnablePeerAccess(0, 0));
    CK(cudaFuncSetAttribute(k_push<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768));
    cudaStream_t s1; CK(cudaStreamCreate(&s1));
    cudaEvent_t f0, f1; cudaEventCreate(&f0); cudaEventCreate(&f1);
    CK(cudaDeviceSynchronize());
    CK(cudaSetDevice(0)); CK(cudaMalloc(&d0, bytes));
    cudaStream_t s0; CK(cudaStreamCreate(&s0));
    for (int kind = 0; kind < 2; kind++) {
      float b0 = 1e9, b1 = 1e9;
      for (int rep = 0; rep < 5; rep++) {
        CK(cudaSetDevice(0)); cudaEventRecord(e0, s0);
        if (kind == 0) k_push<1><<<296, 512, 32768, s0>>>(src, dpeer, nruns, run);
        else CK(cudaMemcpyPeerAsync(dpeer, 1, src, 0, bytes, s0));
        cudaEventRecord(e1, s0);
        CK(cudaSetDevice(1)); cudaEventRecord(f0, s1);
        if (kind == 0) k_push<1><<<296, 512, 32768, s1>>>(src1, d0, nruns, run);
        else CK(cudaMemcpyPeerAsync(d0, 0, src1, 1, bytes, s1));
        cudaEventRecord(f1, s1);
        CK(cudaSetDevice(0)); CK(cudaDeviceSynchronize());
        CK(cudaSetDevice(1)); CK(cudaDeviceSynchronize());
        float m0, m1; cudaEventElapsedTime(&m0, e0, e1); cudaEventElapsedTime(&m1, f0, f1);
        if (rep && m0 < b0) b0 = m0;
        if (rep && m1 < b1) b1 = m1;
      }
      printf("BOTH DIRECTIONS %s run %u B: 0->1 %.3f ms = %.0f GB/s, 1->0 %.3f ms = %.0f GB/s\n",
             kind == 0 ? "st.16 grid 296" : "cudaMemcpyPeerAsync", run, b0, bytes / b0 / 1e6, b1, bytes / b1 / 1e6);
    }
  }
  return 0;
}
