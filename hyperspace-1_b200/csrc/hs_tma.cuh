// hs_tma.cuh — bulk asynchronous copies (TMA, cp.async.bulk) of contiguous column slices into shared memory and
// the mbarrier plumbing around them. Shared by the row mover (hs_sort.cu) and the join (hs_join.cu).
#pragma once
#include <stdint.h>

// ---- TMA (bulk async copy) + mbarrier helpers: contiguous column slices global -> shared memory
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  do {
    asm volatile(
        "{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
  } while (!ok);
}

// one elected thread starts the bulk copy of the tile's slice of column c into a stage buffer.
// The copy covers whole 16-byte granules around the slice: every granule holds at least one byte of the
// slice, so it never touches a page the slice does not touch. Returns nothing; *lead = offset of the
// slice inside the stage buffer.
__device__ __forceinline__ uint32_t slice_lead(const void* col, int esize, uint32_t tile_start) {
  return (uint32_t)(((uintptr_t)col + (size_t)tile_start * esize) & 15);
}
__device__ __forceinline__ void issue_slice(const void* col, int esize, uint32_t tile_start, uint32_t tile_len,
                                            unsigned char* stage, uint64_t* bar) {
  uintptr_t addr = (uintptr_t)col + (size_t)tile_start * esize;
  uintptr_t a0 = addr & ~(uintptr_t)15;
  uint32_t bytes = (uint32_t)(((addr + (size_t)tile_len * esize + 15) & ~(uintptr_t)15) - a0);
  mbar_expect_tx(bar, bytes);
  tma_load_1d(stage, (const void*)a0, bytes, bar);
}

