// gather_ceiling.cu -- what can a B200 sustain on RANDOM row reads?  A stripped gather: every
// group of G lanes reads rows at hashed (uniform, non-repeating) positions of a large
// buffer with 128-bit loads, U loads in flight per lane, and folds them into a register sum.
// No ids in memory, no LoD, no stores: an upper bound for the fused kernel's row traffic.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o gather_ceiling gather_ceiling.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>

__device__ __forceinline__ uint64_t mix(uint64_t z) {
  z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull; z ^= z >> 27; z *= 0x94D049BB133111EBull; z ^= z >> 31; return z;
}

template <int G, int U>
__global__ void __launch_bounds__(256) gather(const uint4* __restrict__ buf, uint64_t n_rows, int rows_per_group,
                                              float* sink, uint64_t salt) {
  const uint64_t gid = (static_cast<uint64_t>(blockIdx.x) * 256 + threadIdx.x) / G;
  const int gl = threadIdx.x % G;
  float acc = 0.f;
  for (int r0 = 0; r0 < rows_per_group; r0 += U) {
    uint4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const uint64_t row = mix(gid * 0x9E3779B97F4A7C15ull + (r0 + u) + salt) % n_rows;
      v[u] = __ldg(buf + row * G + gl);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) acc += __uint_as_float(v[u].x ^ v[u].y ^ v[u].z ^ v[u].w);
  }
  if (acc == 123.456f) sink[0] = acc;
}

template <int G, int U>
void run(const uint4* buf, size_t bytes, float* sink) {
  const int row_bytes = 16 * G;
  const uint64_t n_rows = bytes / row_bytes;
  const int rows_per_group = 16;
  const uint64_t groups = (1ull << 30) / row_bytes / rows_per_group;  // 1 GiB of rows per launch
  const unsigned blocks = static_cast<unsigned>(groups * G / 256);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int i = 0; i < 3; ++i) gather<G, U><<<blocks, 256>>>(buf, n_rows, rows_per_group, sink, i);
  cudaEventRecord(e0);
  const int iters = 10;
  for (int i = 0; i < iters; ++i) gather<G, U><<<blocks, 256>>>(buf, n_rows, rows_per_group, sink, 100 + i);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
  const double gb = static_cast<double>(groups) * rows_per_group * row_bytes * iters / 1e9;
  printf("row %4d B, %d loads in flight/lane, table %.1f GB: %.0f GB/s (%s)\n", row_bytes, U, bytes / 1e9, gb / (ms * 1e-3),
         cudaGetErrorString(cudaGetLastError()));
}

int main(int argc, char** argv) {
  const size_t gbytes = argc > 1 ? atoll(argv[1]) : 32;
  const size_t bytes = gbytes << 30;
  uint4* buf; float* sink;
  if (cudaMalloc(&buf, bytes) != cudaSuccess) { printf("alloc failed\n"); return 1; }
  cudaMalloc(&sink, 4);
  cudaMemset(buf, 1, bytes);
  run<4, 4>(buf, bytes, sink);  run<4, 8>(buf, bytes, sink);
  run<8, 4>(buf, bytes, sink);  run<8, 8>(buf, bytes, sink);
  run<16, 4>(buf, bytes, sink); run<16, 8>(buf, bytes, sink);
  run<2, 8>(buf, bytes, sink);
  // sequential reference: the same loads at consecutive rows
  return 0;
}
