// gather_vmm.cu -- does the page size of the mapping change what random rows of a > 100 GB table
// sustain?  The same LDG gather as gather_sweep.cu over (a) cudaMalloc memory and (b) memory from
// cuMemCreate / cuMemMap at the RECOMMENDED granularity, with the reserved address aligned to 512 MB.
// Prints both granularities the driver reports.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o gather_vmm gather_vmm.cu -lcuda
//   ./gather_vmm [table GiB, default 150]
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>

__device__ __forceinline__ uint64_t mix(uint64_t z) {
  z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull; z ^= z >> 27; z *= 0x94D049BB133111EBull; z ^= z >> 31; return z;
}

template <int G, int U>
__global__ void __launch_bounds__(256) k_ldg(const uint4* __restrict__ buf, uint64_t n_rows, int rows_per_group, float* sink,
                                             uint64_t salt) {
  const uint64_t gid = (static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x) / G;
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

template <int G>
static void run(const char* what, const uint4* buf, size_t bytes, float* sink) {
  const int rowb = 16 * G, rpg = 16;
  const uint64_t n_rows = bytes / rowb;
  const uint64_t groups = (1ull << 30) / rowb / rpg;
  const unsigned blocks = static_cast<unsigned>(groups * G / 256);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int i = 0; i < 2; ++i) k_ldg<G, 4><<<blocks, 256>>>(buf, n_rows, rpg, sink, i);
  cudaEventRecord(e0);
  for (int i = 0; i < 6; ++i) k_ldg<G, 4><<<blocks, 256>>>(buf, n_rows, rpg, sink, 100 + i);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  printf("%-28s row %4d B  table %.1f GB  %7.0f GB/s  %s\n", what, rowb, bytes / 1e9,
         static_cast<double>(groups) * rpg * rowb * 6 / 1e9 / (ms * 1e-3), cudaGetErrorString(cudaGetLastError()));
  fflush(stdout);
}

int main(int argc, char** argv) {
  const size_t gib = argc > 1 ? atoll(argv[1]) : 150;
  const size_t bytes = gib << 30;
  float* sink;
  cudaFree(nullptr);
  cudaMalloc(&sink, 4);
  {
    uint4* buf = nullptr;
    if (cudaMalloc(&buf, bytes) == cudaSuccess) {
      cudaMemset(buf, 1, bytes);
      run<4>("cudaMalloc", buf, bytes, sink);
      run<8>("cudaMalloc", buf, bytes, sink);
      run<16>("cudaMalloc", buf, bytes, sink);
      cudaFree(buf);
    } else {
      printf("cudaMalloc failed\n");
    }
  }
  CUmemAllocationProp prop = {};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = 0;
  size_t gmin = 0, grec = 0;
  cuMemGetAllocationGranularity(&gmin, &prop, CU_MEM_ALLOC_GRANULARITY_MINIMUM);
  cuMemGetAllocationGranularity(&grec, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED);
  printf("allocation granularity: minimum %zu, recommended %zu bytes\n", gmin, grec);
  const size_t align = 512ull << 20;
  const size_t sz = (bytes + align - 1) / align * align;
  CUmemGenericAllocationHandle h;
  CUresult r = cuMemCreate(&h, sz, &prop, 0);
  if (r != CUDA_SUCCESS) { printf("cuMemCreate failed: %d\n", (int)r); return 0; }
  CUdeviceptr va = 0;
  r = cuMemAddressReserve(&va, sz, align, 0, 0);
  if (r != CUDA_SUCCESS) { printf("cuMemAddressReserve failed: %d\n", (int)r); return 0; }
  r = cuMemMap(va, sz, 0, h, 0);
  if (r != CUDA_SUCCESS) { printf("cuMemMap failed: %d\n", (int)r); return 0; }
  CUmemAccessDesc acc = {};
  acc.location = prop.location;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = cuMemSetAccess(va, sz, &acc, 1);
  if (r != CUDA_SUCCESS) { printf("cuMemSetAccess failed: %d\n", (int)r); return 0; }
  printf("VMM mapping at %p (512 MB aligned), %zu bytes\n", (void*)va, sz);
  cudaMemset(reinterpret_cast<void*>(va), 1, bytes);
  run<4>("cuMemCreate 512MB-aligned", reinterpret_cast<const uint4*>(va), bytes, sink);
  run<8>("cuMemCreate 512MB-aligned", reinterpret_cast<const uint4*>(va), bytes, sink);
  run<16>("cuMemCreate 512MB-aligned", reinterpret_cast<const uint4*>(va), bytes, sink);
  return 0;
}
