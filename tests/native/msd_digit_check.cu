// Host-only check of the MSD digit arithmetic of hs_sort_shared.cuh (msd_bits, msd_digit, msd_cell_lo, msd_mul_for):
// built with nvcc and run on the CPU by tests/test_msd_digit_cpu.py. For random segments (len, span) it walks the
// key space and asserts what the sort relies on:
//   1. digits are non-decreasing in the key and stay below 2^bits;
//   2. msd_cell_lo(d) is exactly the first cell whose digit is >= d (the child's base);
//   3. every key of child d, minus the child's base, fits the bits k_msd_plan gives the child;
//   4. with a known span the expected child size is near MSD_TARGET_ROWS (the point of the multiplier).
#include <cstdint>
#include <cstdio>
#include <cstdlib>

#include "hs_sort_shared.cuh"

static uint64_t rng_state = 88172645463325252ull;
static uint64_t rnd() { rng_state ^= rng_state << 13; rng_state ^= rng_state >> 7; rng_state ^= rng_state << 17; return rng_state; }

static int bitlen(uint64_t v) { int b = 0; while (v) { b++; v >>= 1; } return b; }

int main() {
  long checked = 0;
  for (int it = 0; it < 4000; it++) {
    const int r = 1 + (int)(rnd() % 64);                                  // bits that vary
    uint64_t span = r == 64 ? (rnd() | (1ull << 63)) : ((1ull << (r - 1)) | (rnd() & ((1ull << (r - 1)) - 1)));
    if (bitlen(span) != r) { printf("bad span\n"); return 1; }
    const uint32_t len = 3073 + (uint32_t)(rnd() % 2000000);
    const int bits = msd_bits(len, (uint32_t)r);
    const int e = msd_extra((uint32_t)r, bits);
    for (int scaled = 0; scaled < 2; scaled++) {
      const uint32_t mul = scaled ? msd_mul_for(span, (uint32_t)r, bits, len) : 0;
      const uint32_t ncells_all = 1u << (bits + e);
      const uint64_t cell_w = r - bits - e >= 64 ? 0 : (1ull << (r - bits - e));
      const uint64_t last_cell = span >> (r - bits - e);
      uint32_t prev = 0;
      // every cell that can occur: digit monotone, below 2^bits, consistent with cell_lo
      for (uint64_t cell = 0; cell <= last_cell && cell < ncells_all; cell++) {
        const uint64_t rel = cell << (r - bits - e);
        const uint32_t d = msd_digit(rel, (uint32_t)r, bits, mul);
        const uint32_t d_hi = msd_digit(rel + (cell_w - 1), (uint32_t)r, bits, mul);   // last key of the cell
        if (d != d_hi) { printf("digit changes inside a cell r=%d bits=%d mul=%u\n", r, bits, mul); return 1; }
        if (d >= (1u << bits)) { printf("digit %u out of range r=%d bits=%d mul=%u\n", d, r, bits, mul); return 1; }
        if (d < prev) { printf("digit not monotone\n"); return 1; }
        if (msd_cell_lo(d, e, mul) > cell || (d + 1 < (1u << bits) && msd_cell_lo(d + 1, e, mul) <= cell)) {
          printf("cell_lo inconsistent: cell %llu digit %u lo %u next %u (r=%d bits=%d e=%d mul=%u)\n",
                 (unsigned long long)cell, d, msd_cell_lo(d, e, mul), msd_cell_lo(d + 1, e, mul), r, bits, e, mul);
          return 1;
        }
        // the child's bits (as k_msd_plan computes them) hold every key of the child relative to its base
        const uint32_t c0 = msd_cell_lo(d, e, mul), c1 = msd_cell_lo(d + 1, e, mul);
        const uint32_t ncell = c1 > c0 ? c1 - c0 : 1;
        const uint32_t child_bits = (uint32_t)(r - bits - e) + (ncell > 1 ? 32 - __builtin_clz(ncell - 1) : 0);
        const uint64_t rel_child_max = ((cell - c0) << (r - bits - e)) + (cell_w - 1);
        if (child_bits < 64 && rel_child_max >> child_bits) { printf("child bits too few\n"); return 1; }
        prev = d;
        checked++;
      }
      if (scaled && mul) {
        // expected child = len / digits in use: within [target / 2, target * 1.07] unless the level is out of
        // digits (2^bits) or the span has fewer cells than digits wanted
        const uint32_t used = msd_digit(span, (uint32_t)r, bits, mul) + 1;
        const double child = (double)len / used;
        const uint64_t wanted = (len + MSD_TARGET_ROWS - 1) / MSD_TARGET_ROWS;
        if (used < (1u << bits) && wanted <= last_cell + 1 && (child > MSD_TARGET_ROWS * 1.07 || child < MSD_TARGET_ROWS * 0.5)) {
          printf("expected child %.0f rows with %u digits (len %u, r=%d, bits=%d)\n", child, used, len, r, bits);
          return 1;
        }
      }
    }
  }
  // the cfg2 shape: 500 k rows per bucket, keys below 100 M -> 197 digits, children of ~2540 rows in 19 bits
  {
    const uint32_t len = 500000; const uint64_t span = 99999999; const int r = 27;
    const int bits = msd_bits(len, r); const uint32_t mul = msd_mul_for(span, r, bits, len);
    const uint32_t used = msd_digit(span, r, bits, mul) + 1;
    if (bits != 8 || used != 197) { printf("cfg2: bits %d digits %u\n", bits, used); return 1; }
  }
  printf("ok %ld cells\n", checked);
  return 0;
}
