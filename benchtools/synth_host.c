// synth_host.c -- the table generator of synth_kernels.cu on the host cores (OpenMP), bit-identical:
// the reference arm of bench.py builds its table with this and never touches the GPU.
//   gcc -O3 -fopenmp -fPIC -shared -o libsynth_host.so synth_host.c
#include <math.h>
#include <stdint.h>
#include <string.h>

static inline uint64_t mix64(uint64_t z) {
  z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull;
  z ^= z >> 27; z *= 0x94D049BB133111EBull;
  z ^= z >> 31;
  return z;
}
static inline float unit24(uint64_t seed, uint64_t row, uint64_t col) {
  const uint64_t z = mix64(row * 0x9E3779B97F4A7C15ull + col * 0xBF58476D1CE4E5B9ull + seed * 0x94D049BB133111EBull);
  return (float)(z >> 40) * (1.0f / 16777216.0f);
}
static inline float table_value(uint64_t seed, uint64_t row, int col, int cvm_cols) {
  const float u = unit24(seed, row, (uint64_t)col);
  if (cvm_cols) {
    if (col == 0) return floorf(u * 16.0f);
    if (col == 1) return floorf(u * (floorf(unit24(seed, row, 0) * 16.0f) + 1.0f));
  }
  return u * 2.0f - 1.0f;
}
static inline uint16_t f32_to_bf16(float v) {  // round to nearest even, as __float2bfloat16_rn
  uint32_t x;
  memcpy(&x, &v, 4);
  return (uint16_t)((x + (0x7FFFu + ((x >> 16) & 1u))) >> 16);
}

// dtype: 0 f32, 1 f16, 2 bf16; rows row0, row0 + row_stride, ... written to t[0 .. rows*dim)
int synth_fill_table_host(void* t, long long rows, int dim, int dtype, unsigned long long seed, long long row0,
                          long long row_stride, int cvm_cols, int threads) {
  if (dtype < 0 || dtype > 2) return 1;
#pragma omp parallel for schedule(static) num_threads(threads > 0 ? threads : 1)
  for (long long r = 0; r < rows; ++r) {
    const uint64_t lr = (uint64_t)(row0 + r * row_stride);
    for (int c = 0; c < dim; ++c) {
      const float v = table_value(seed, lr, c, cvm_cols);
      if (dtype == 0) ((float*)t)[r * dim + c] = v;
      else if (dtype == 1) ((_Float16*)t)[r * dim + c] = (_Float16)v;
      else ((uint16_t*)t)[r * dim + c] = f32_to_bf16(v);
    }
  }
  return 0;
}
