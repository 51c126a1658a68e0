// synth_kernels.cu -- fills an embedding table on the device from a counter-based hash of
// (seed, row, col), so a 100 GB table never crosses PCIe and the CPU side can regenerate
// any row bit-identically (benchtools/synth.py: table_rows_numpy).  Bench/test tooling
// only; not part of libctrfe.so.
//
// Value law: column 0 "show" = integer in [0,16); column 1
// "click" = integer in [0, show]; other columns uniform in [-1, 1) on a 2^-23 grid.
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace {

__host__ __device__ inline uint64_t mix64(uint64_t z) {
  z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull;
  z ^= z >> 27; z *= 0x94D049BB133111EBull;
  z ^= z >> 31;
  return z;
}

__device__ inline float unit24(uint64_t seed, uint64_t row, uint64_t col) {
  const uint64_t z = mix64(row * 0x9E3779B97F4A7C15ull + col * 0xBF58476D1CE4E5B9ull + seed * 0x94D049BB133111EBull);
  return static_cast<float>(z >> 40) * (1.0f / 16777216.0f);  // exact: 24-bit integer * 2^-24
}

__device__ inline float table_value(uint64_t seed, uint64_t row, int col, int cvm_cols) {
  const float u = unit24(seed, row, static_cast<uint64_t>(col));
  if (cvm_cols) {
    if (col == 0) return floorf(u * 16.0f);
    if (col == 1) {
      const float show = floorf(unit24(seed, row, 0) * 16.0f);
      return floorf(u * (show + 1.0f));
    }
  }
  return u * 2.0f - 1.0f;
}

template <typename T> __device__ inline T cast_to(float v);
template <> __device__ inline float cast_to<float>(float v) { return v; }
template <> __device__ inline __half cast_to<__half>(float v) { return __float2half_rn(v); }
template <> __device__ inline __nv_bfloat16 cast_to<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

template <typename T>
__global__ void fill_table_kernel(T* __restrict__ t, long long rows, int dim, uint64_t seed, long long row0,
                                  long long row_stride, int cvm_cols) {
  const long long n = rows * dim;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    const long long r = i / dim;
    const int c = static_cast<int>(i - r * dim);
    t[i] = cast_to<T>(table_value(seed, static_cast<uint64_t>(row0 + r * row_stride), c, cvm_cols));
  }
}

}  // namespace

// dtype: 0 f32, 1 f16, 2 bf16.  Logical rows row0, row0 + row_stride, ... (rows of them) are
// written to t[0 .. rows*dim) -- row_stride > 1 fills one shard of a row-sharded table.
// Returns a cudaError_t.
extern "C" int synth_fill_table(void* t, long long rows, int dim, int dtype, unsigned long long seed,
                                long long row0, long long row_stride, int cvm_cols, void* stream) {
  auto st = static_cast<cudaStream_t>(stream);
  const int threads = 256;
  const int blocks = 148 * 16;
  switch (dtype) {
    case 0: fill_table_kernel<float><<<blocks, threads, 0, st>>>(static_cast<float*>(t), rows, dim, seed, row0, row_stride, cvm_cols); break;
    case 1: fill_table_kernel<__half><<<blocks, threads, 0, st>>>(static_cast<__half*>(t), rows, dim, seed, row0, row_stride, cvm_cols); break;
    case 2: fill_table_kernel<__nv_bfloat16><<<blocks, threads, 0, st>>>(static_cast<__nv_bfloat16*>(t), rows, dim, seed, row0, row_stride, cvm_cols); break;
    default: return 1;
  }
  return static_cast<int>(cudaGetLastError());
}
