// log_check.cu -- how far is ctrfe::cvm_log1 (the short fp64 log used by the CVM epilogue)
// from the CUDA library log?  Sweeps every float x in a strided walk over all non-negative
// bit patterns plus dense ranges around the counts CTR data holds, on the GPU.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -ftz=true -fmad=false -I.. -o log_check log_check.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "../paddle_infer_b200/csrc/ctrfe_kernels.cuh"

__global__ void sweep(unsigned first, unsigned stride, unsigned n, double* max_rel, unsigned long long* flips) {
  const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float x = __uint_as_float(first + i * stride);
  if (!(x >= 0.0f) || isinf(x)) return;
  const double ref = log(static_cast<double>(x + 1.0f));
  const double got = ctrfe::cvm_log1(x);
  double rel = ref == 0.0 ? fabs(got) : fabs((got - ref) / ref);
  if (isinf(ref)) rel = (got == ref) ? 0.0 : 1.0;
  unsigned long long* mr = reinterpret_cast<unsigned long long*>(max_rel);
  atomicMax(mr, static_cast<unsigned long long>(__double_as_longlong(rel)));  // rel >= 0: bit order == value order
  if (static_cast<float>(got) != static_cast<float>(ref)) atomicAdd(flips, 1ull);
}

int main() {
  double* d_max; unsigned long long* d_flips;
  cudaMalloc(&d_max, 8); cudaMalloc(&d_flips, 8);
  cudaMemset(d_max, 0, 8); cudaMemset(d_flips, 0, 8);
  const unsigned n = 1u << 24;
  sweep<<<n / 256, 256>>>(0u, 127u, n, d_max, d_flips);            // all magnitudes, stride 127 bit patterns
  sweep<<<n / 256, 256>>>(0x3f800000u, 1u, n, d_max, d_flips);     // every float in [1, 4)
  sweep<<<n / 256, 256>>>(0x41000000u, 1u, n, d_max, d_flips);     // every float in [8, 32)
  sweep<<<n / 256, 256>>>(0x33000000u, 64u, n, d_max, d_flips);    // tiny arguments: x + 1 rounds to 1 or 1 + ulp
  double h_max; unsigned long long h_flips;
  cudaMemcpy(&h_max, d_max, 8, cudaMemcpyDeviceToHost);
  cudaMemcpy(&h_flips, d_flips, 8, cudaMemcpyDeviceToHost);
  printf("cvm_log1 vs library log over %u arguments: max relative difference %.3e, float results that differ: %llu (%s)\n",
         4 * n, h_max, h_flips, cudaGetErrorString(cudaGetLastError()));
  return 0;
}
