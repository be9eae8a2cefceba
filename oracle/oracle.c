/* TEST INFRASTRUCTURE ONLY -- see oracle.h.  Plain C11, no dependencies beyond libc. */
#include "oracle.h"
#include <stdlib.h>
#include <string.h>

#define T double
#define SUF f64
#include "oracle_impl.inc"
#undef T
#undef SUF

#define T float
#define SUF f32
#include "oracle_impl.inc"
#undef T
#undef SUF

int oracle_string2order(const char* s) /* linalg.hpp:21-37 */
{
  static const char* names[6] = {"ijk", "ikj", "jik", "jki", "kij", "kji"};
  for (int i = 0; i < 6; ++i)
    if (strcmp(s, names[i]) == 0)
      return i;
  return -1;
}

void oracle_srand(unsigned seed) { srand(seed); }

static inline uint64_t splitmix64(uint64_t x)
{
  uint64_t z = x + 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

void oracle_fill_splitmix_f64(double* x, size_t n, uint64_t seed, uint64_t stream, uint64_t offset, int centred)
{
  const uint64_t base = seed + stream * 0x9E3779B97F4A7C15ull + offset;
  for (size_t i = 0; i < n; ++i)
  {
    const double v = (double)(splitmix64(base + i) >> 11) * 0x1.0p-53;
    x[i] = centred ? v - 0.5 : v;
  }
}

void oracle_fill_splitmix_f32(float* x, size_t n, uint64_t seed, uint64_t stream, uint64_t offset, int centred)
{
  const uint64_t base = seed + stream * 0x9E3779B97F4A7C15ull + offset;
  for (size_t i = 0; i < n; ++i)
  {
    const float v = (float)(splitmix64(base + i) >> 40) * 0x1.0p-24f;
    x[i] = centred ? v - 0.5f : v;
  }
}
