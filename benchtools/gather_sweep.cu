// gather_sweep.cu -- what does a B200 sustain on RANDOM row reads, and by which instruction?
// A stripped gather over a large buffer, rows at hashed (uniform, non-repeating) positions, swept
// over: row width (64 / 128 / 256 B), loads in flight per lane, resident warps per SM, and the
// data-movement instruction:
//   ldg    128-bit LDG per lane into registers (what the pooling kernels do)
//   ldgsts cp.async 16 B per lane into shared memory (LDGSTS), read back with LDS
//   bulk   one cp.async.bulk (UBLKCP, the 1-D TMA copy) per ROW into shared memory, completion on
//          an mbarrier, issued by every lane (bulkL) or by one elected lane (bulk1)
// No ids in memory, no LoD, no stores: an upper bound for the pooling kernels' row traffic.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o gather_sweep gather_sweep.cu
//   ./gather_sweep [table GiB, default 32]
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>

__device__ __forceinline__ uint64_t mix(uint64_t z) {
  z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull; z ^= z >> 27; z *= 0x94D049BB133111EBull; z ^= z >> 31; return z;
}
__device__ __forceinline__ unsigned smem_u32(const void* p) { return static_cast<unsigned>(__cvta_generic_to_shared(p)); }

// ---- ldg: G lanes per row, U rows in flight per lane ---------------------------------------
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

// ---- ldgsts: the same mapping, rows land in shared memory ------------------------------------
template <int G, int U>
__global__ void __launch_bounds__(256) k_ldgsts(const uint4* __restrict__ buf, uint64_t n_rows, int rows_per_group, float* sink,
                                                uint64_t salt) {
  extern __shared__ uint4 sm[];
  uint4* mine = sm + threadIdx.x;  // [u][thread]
  const uint64_t gid = (static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x) / G;
  const int gl = threadIdx.x % G;
  float acc = 0.f;
  for (int r0 = 0; r0 < rows_per_group; r0 += U) {
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const uint64_t row = mix(gid * 0x9E3779B97F4A7C15ull + (r0 + u) + salt) % n_rows;
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(mine + u * blockDim.x)), "l"(buf + row * G + gl) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const uint4 v = mine[u * blockDim.x];
      acc += __uint_as_float(v.x ^ v.y ^ v.z ^ v.w);
    }
  }
  if (acc == 123.456f) sink[0] = acc;
}

// ---- bulk: per warp a double-buffered ring of R rows, one cp.async.bulk per row --------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "W: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@!p bra W;\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// ROWB bytes per row, R rows per batch per warp (a multiple of 32), ONE: one elected lane issues
template <int ROWB, int R, bool ONE>
__global__ void __launch_bounds__(256) k_bulk(const unsigned char* __restrict__ buf, uint64_t n_rows, int batches_per_warp,
                                              float* sink, uint64_t salt) {
  extern __shared__ __align__(128) unsigned char smraw[];
  constexpr int kStage = R * ROWB;
  const int wib = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  unsigned char* ring = smraw + static_cast<size_t>(wib) * 2 * kStage;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smraw + static_cast<size_t>(nw) * 2 * kStage) + wib * 2;
  if (lane == 0) { mbar_init(bars, 1); mbar_init(bars + 1, 1); }
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncwarp();
  const uint64_t wid = static_cast<uint64_t>(blockIdx.x) * nw + wib;
  auto issue = [&](int b) {
    const int s = b & 1;
    if (lane == 0) mbar_expect_tx(bars + s, kStage);
    __syncwarp();
    if (ONE) {
      if (lane == 0)
        for (int r = 0; r < R; ++r) {
          const uint64_t row = mix(wid * 0x9E3779B97F4A7C15ull + static_cast<uint64_t>(b) * R + r + salt) % n_rows;
          bulk_g2s(ring + s * kStage + r * ROWB, buf + row * ROWB, ROWB, bars + s);
        }
    } else {
#pragma unroll
      for (int r = lane; r < R; r += 32) {
        const uint64_t row = mix(wid * 0x9E3779B97F4A7C15ull + static_cast<uint64_t>(b) * R + r + salt) % n_rows;
        bulk_g2s(ring + s * kStage + r * ROWB, buf + row * ROWB, ROWB, bars + s);
      }
    }
  };
  float acc = 0.f;
  issue(0);
  for (int b = 0; b < batches_per_warp; ++b) {
    if (b + 1 < batches_per_warp) issue(b + 1);
    mbar_wait(bars + (b & 1), (b >> 1) & 1);
    const uint4* st = reinterpret_cast<const uint4*>(ring + (b & 1) * kStage);
#pragma unroll 4
    for (int c = lane; c < kStage / 16; c += 32) {
      const uint4 v = st[c];
      acc += __uint_as_float(v.x ^ v.y ^ v.z ^ v.w);
    }
    __syncwarp();
  }
  if (acc == 123.456f) sink[0] = acc;
}

// ---- hybrid: NT of a CTA's 8 warps fetch rows with bulk copies, the others with LDG -----------
// Are the two paths' limits independent (LSU / L1 miss tracking vs the TMA request rate)?
// LDG warps read `rpg` rows per lane group, bulk warps `bpw` batches of 32 rows.
template <int ROWB, int NT>
__global__ void __launch_bounds__(256) k_hybrid(const unsigned char* __restrict__ buf, uint64_t n_rows, int rpg, int bpw,
                                                float* sink, uint64_t salt) {
  extern __shared__ __align__(128) unsigned char smraw[];
  constexpr int R = 32, G = ROWB / 16, U = 4;
  constexpr int kStage = R * ROWB;
  const int wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float acc = 0.f;
  if (wib < NT) {
    unsigned char* ring = smraw + static_cast<size_t>(wib) * 2 * kStage;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smraw + static_cast<size_t>(NT) * 2 * kStage) + wib * 2;
    if (lane == 0) { mbar_init(bars, 1); mbar_init(bars + 1, 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncwarp();
    const uint64_t wid = static_cast<uint64_t>(blockIdx.x) * 8 + wib;
    auto issue = [&](int b) {
      const int s = b & 1;
      if (lane == 0) mbar_expect_tx(bars + s, kStage);
      __syncwarp();
      const uint64_t row = mix(wid * 0x9E3779B97F4A7C15ull + static_cast<uint64_t>(b) * R + lane + salt) % n_rows;
      bulk_g2s(ring + s * kStage + lane * ROWB, buf + row * ROWB, ROWB, bars + s);
    };
    issue(0);
    for (int b = 0; b < bpw; ++b) {
      if (b + 1 < bpw) issue(b + 1);
      mbar_wait(bars + (b & 1), (b >> 1) & 1);
      const uint4* st = reinterpret_cast<const uint4*>(ring + (b & 1) * kStage);
#pragma unroll 4
      for (int c = lane; c < kStage / 16; c += 32) {
        const uint4 v = st[c];
        acc += __uint_as_float(v.x ^ v.y ^ v.z ^ v.w);
      }
      __syncwarp();
    }
  } else {
    const uint64_t gid = (static_cast<uint64_t>(blockIdx.x) * 256 + threadIdx.x) / G + (1ull << 40);
    const int gl = threadIdx.x % G;
    const uint4* b4 = reinterpret_cast<const uint4*>(buf);
    for (int r0 = 0; r0 < rpg; r0 += U) {
      uint4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const uint64_t row = mix(gid * 0x9E3779B97F4A7C15ull + (r0 + u) + salt) % n_rows;
        v[u] = __ldg(b4 + row * G + gl);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) acc += __uint_as_float(v[u].x ^ v[u].y ^ v[u].z ^ v[u].w);
    }
  }
  if (acc == 123.456f) sink[0] = acc;
}

// ---- ldg256: 32-byte loads, G2 lanes per row -------------------------------------------------
template <int G2, int U>
__global__ void __launch_bounds__(256) k_ldg256(const unsigned char* __restrict__ buf, uint64_t n_rows, int rows_per_group,
                                                float* sink, uint64_t salt) {
  const uint64_t gid = (static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x) / G2;
  const int gl = threadIdx.x % G2;
  float acc = 0.f;
  for (int r0 = 0; r0 < rows_per_group; r0 += U) {
    unsigned w[U][8];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const uint64_t row = mix(gid * 0x9E3779B97F4A7C15ull + (r0 + u) + salt) % n_rows;
      const unsigned char* p = buf + (row * G2 + gl) * 32;
      asm volatile("ld.global.nc.v8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                   : "=r"(w[u][0]), "=r"(w[u][1]), "=r"(w[u][2]), "=r"(w[u][3]), "=r"(w[u][4]), "=r"(w[u][5]), "=r"(w[u][6]),
                     "=r"(w[u][7])
                   : "l"(p));
    }
#pragma unroll
    for (int u = 0; u < U; ++u) acc += __uint_as_float(w[u][0] ^ w[u][1] ^ w[u][2] ^ w[u][3] ^ w[u][4] ^ w[u][5] ^ w[u][6] ^ w[u][7]);
  }
  if (acc == 123.456f) sink[0] = acc;
}

struct Timer {
  cudaEvent_t e0, e1;
  Timer() { cudaEventCreate(&e0); cudaEventCreate(&e1); }
  void start() { cudaEventRecord(e0); }
  float stop() { cudaEventRecord(e1); cudaEventSynchronize(e1); float ms = 0; cudaEventElapsedTime(&ms, e0, e1); return ms; }
};

static const uint4* g_buf; static size_t g_bytes; static float* g_sink; static int g_sms = 148;
constexpr uint64_t kTraffic = 1ull << 30;  // bytes of rows per launch
constexpr int kIters = 6;

static void report(const char* mode, int rowb, int inflight, int warps_per_sm, double gb, float ms) {
  cudaError_t e = cudaGetLastError();
  printf("%-7s row %4d B  in-flight/lane %3d  warps/SM %2d  %7.0f GB/s  %s\n", mode, rowb, inflight, warps_per_sm, gb / (ms * 1e-3),
         e == cudaSuccess ? "" : cudaGetErrorString(e));
  fflush(stdout);
}

// occ = resident CTAs of 256 threads per SM (limited with dynamic shared memory)
template <int G, int U>
void run_ldg(int occ) {
  const int rowb = 16 * G;
  const uint64_t n_rows = g_bytes / rowb;
  const int rpg = 16 * (U > 16 ? 2 : 1);
  const uint64_t groups = kTraffic / rowb / rpg;
  const unsigned blocks = static_cast<unsigned>(groups * G / 256);
  const int smem = occ >= 8 ? 0 : (227 * 1024 / occ) - 2048;
  cudaFuncSetAttribute(k_ldg<G, U>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  Timer t;
  for (int i = 0; i < 2; ++i) k_ldg<G, U><<<blocks, 256, smem>>>(g_buf, n_rows, rpg, g_sink, i);
  t.start();
  for (int i = 0; i < kIters; ++i) k_ldg<G, U><<<blocks, 256, smem>>>(g_buf, n_rows, rpg, g_sink, 100 + i);
  const float ms = t.stop();
  report("ldg", rowb, U, occ * 8, static_cast<double>(groups) * rpg * rowb * kIters / 1e9, ms);
}

template <int G, int U>
void run_ldgsts(int occ) {
  const int rowb = 16 * G;
  const uint64_t n_rows = g_bytes / rowb;
  const int rpg = 16 * (U > 16 ? 2 : 1);
  const uint64_t groups = kTraffic / rowb / rpg;
  const unsigned blocks = static_cast<unsigned>(groups * G / 256);
  int smem = U * 256 * 16;
  const int cap = (227 * 1024 / occ) - 2048;
  if (smem > cap) return;
  if (occ < 8) smem = cap;
  cudaFuncSetAttribute(k_ldgsts<G, U>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  Timer t;
  for (int i = 0; i < 2; ++i) k_ldgsts<G, U><<<blocks, 256, smem>>>(g_buf, n_rows, rpg, g_sink, i);
  t.start();
  for (int i = 0; i < kIters; ++i) k_ldgsts<G, U><<<blocks, 256, smem>>>(g_buf, n_rows, rpg, g_sink, 100 + i);
  const float ms = t.stop();
  report("ldgsts", rowb, U, occ * 8, static_cast<double>(groups) * rpg * rowb * kIters / 1e9, ms);
}

template <int ROWB, int R, bool ONE>
void run_bulk(int threads, int occ) {
  const uint64_t n_rows = g_bytes / ROWB;
  const int nw = threads / 32;
  int smem = nw * 2 * R * ROWB + nw * 16;
  const int cap = (227 * 1024 / occ) - 1024;
  if (smem > cap) return;
  const int bpw = 8;  // batches per warp
  const uint64_t warps = kTraffic / ROWB / R / bpw;
  const unsigned blocks = static_cast<unsigned>(warps / nw);
  cudaFuncSetAttribute(k_bulk<ROWB, R, ONE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  Timer t;
  for (int i = 0; i < 2; ++i)
    k_bulk<ROWB, R, ONE><<<blocks, threads, smem>>>(reinterpret_cast<const unsigned char*>(g_buf), n_rows, bpw, g_sink, i);
  t.start();
  for (int i = 0; i < kIters; ++i)
    k_bulk<ROWB, R, ONE><<<blocks, threads, smem>>>(reinterpret_cast<const unsigned char*>(g_buf), n_rows, bpw, g_sink, 100 + i);
  const float ms = t.stop();
  // rows in flight per lane-equivalent: 2 stages * R rows per warp
  report(ONE ? "bulk1" : "bulkL", ROWB, 2 * R / 32 * (ROWB / 16) > 0 ? 2 * R * (ROWB / 16) / 32 : 1, nw * occ,
         static_cast<double>(blocks) * nw * bpw * R * ROWB * kIters / 1e9, ms);
}

template <int ROWB, int NT>
void run_hybrid(int rpg, int bpw) {
  const uint64_t n_rows = g_bytes / ROWB;
  constexpr int G = ROWB / 16;
  const int smem = NT * 2 * 32 * ROWB + NT * 16;
  const double rows_per_cta = static_cast<double>(8 - NT) * (32 / G) * rpg + static_cast<double>(NT) * bpw * 32;
  const unsigned blocks = static_cast<unsigned>(static_cast<double>(kTraffic) / ROWB / rows_per_cta);
  cudaFuncSetAttribute(k_hybrid<ROWB, NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  Timer t;
  for (int i = 0; i < 2; ++i)
    k_hybrid<ROWB, NT><<<blocks, 256, smem>>>(reinterpret_cast<const unsigned char*>(g_buf), n_rows, rpg, bpw, g_sink, i);
  t.start();
  for (int i = 0; i < kIters; ++i)
    k_hybrid<ROWB, NT><<<blocks, 256, smem>>>(reinterpret_cast<const unsigned char*>(g_buf), n_rows, rpg, bpw, g_sink, 100 + i);
  const float ms = t.stop();
  const double gb = blocks * rows_per_cta * ROWB * kIters / 1e9;
  const double tma_share = static_cast<double>(NT) * bpw * 32 / rows_per_cta;
  cudaError_t e = cudaGetLastError();
  printf("hybrid  row %4d B  bulk warps %d of 8  rows/group(ldg) %3d  batches(bulk) %3d  bulk share %.2f  %7.0f GB/s  %s\n", ROWB, NT, rpg,
         bpw, tma_share, gb / (ms * 1e-3), e == cudaSuccess ? "" : cudaGetErrorString(e));
  fflush(stdout);
}

template <int G2, int U>
void run_ldg256() {
  const int rowb = 32 * G2;
  const uint64_t n_rows = g_bytes / rowb;
  const int rpg = 16;
  const uint64_t groups = kTraffic / rowb / rpg;
  const unsigned blocks = static_cast<unsigned>(groups * G2 / 256);
  Timer t;
  for (int i = 0; i < 2; ++i) k_ldg256<G2, U><<<blocks, 256>>>(reinterpret_cast<const unsigned char*>(g_buf), n_rows, rpg, g_sink, i);
  t.start();
  for (int i = 0; i < kIters; ++i)
    k_ldg256<G2, U><<<blocks, 256>>>(reinterpret_cast<const unsigned char*>(g_buf), n_rows, rpg, g_sink, 100 + i);
  const float ms = t.stop();
  report("ldg256", rowb, U, 64, static_cast<double>(groups) * rpg * rowb * kIters / 1e9, ms);
}

template <int ROWB>
void hybrid_grid() {
  // LDG warps: 32/G groups x rpg rows; bulk warps: bpw x 32 rows.  Sweep the share of rows on the bulk path.
  constexpr int G = ROWB / 16;
  for (int rpg : {16, 32}) {
    const int ldg_rows = (32 / G) * rpg;
    for (int mult : {1, 2, 4, 8}) {
      const int bpw = ldg_rows * mult / 4 / 32 > 0 ? ldg_rows * mult / 4 / 32 : 1;
      run_hybrid<ROWB, 1>(rpg, bpw);
      run_hybrid<ROWB, 2>(rpg, bpw);
      run_hybrid<ROWB, 4>(rpg, bpw);
    }
  }
}

int main(int argc, char** argv) {
  const size_t gbytes = argc > 1 ? atoll(argv[1]) : 32;
  g_bytes = gbytes << 30;
  uint4* buf;
  if (cudaMalloc(&buf, g_bytes) != cudaSuccess) { printf("alloc failed\n"); return 1; }
  cudaMalloc(&g_sink, 4);
  cudaMemset(buf, 1, g_bytes);
  g_buf = buf;
  cudaDeviceGetAttribute(&g_sms, cudaDevAttrMultiProcessorCount, 0);
  if (argc > 2 && argv[2][0] == 'w') {  // table-size check: the LDG ceilings only (does a > 100 GB table change them?)
    printf("ldg / ldg256 pass, table %.1f GB\n", g_bytes / 1e9);
    run_ldg<4, 4>(8); run_ldg<8, 4>(8); run_ldg<16, 4>(8);
    run_ldg256<2, 2>(); run_ldg256<4, 2>(); run_ldg256<4, 4>(); run_ldg256<8, 2>(); run_ldg256<8, 4>();
    return 0;
  }
  if (argc > 2 && argv[2][0] == 'h') {  // second pass: hybrid LDG + bulk, and 256-bit loads
    printf("hybrid / ldg256 pass, table %.1f GB\n", g_bytes / 1e9);
    run_ldg256<2, 2>(); run_ldg256<2, 4>(); run_ldg256<4, 2>(); run_ldg256<4, 4>(); run_ldg256<8, 2>(); run_ldg256<8, 4>();
    hybrid_grid<64>();
    hybrid_grid<128>();
    hybrid_grid<256>();
    return 0;
  }
  printf("table %.1f GB, %d SMs, %.2f GB of rows per launch, %d launches timed\n", g_bytes / 1e9, g_sms, kTraffic / 1e9, kIters);
  // --- ldg: row width x loads in flight x occupancy
  for (int occ : {8, 4, 2}) {
    run_ldg<4, 2>(occ);  run_ldg<4, 4>(occ);  run_ldg<4, 8>(occ);  run_ldg<4, 16>(occ);
    run_ldg<8, 2>(occ);  run_ldg<8, 4>(occ);  run_ldg<8, 8>(occ);  run_ldg<8, 16>(occ);
    run_ldg<16, 2>(occ); run_ldg<16, 4>(occ); run_ldg<16, 8>(occ); run_ldg<16, 16>(occ);
  }
  run_ldg<2, 4>(8); run_ldg<2, 8>(8); run_ldg<2, 16>(8);
  // --- ldgsts
  for (int occ : {8, 4}) {
    run_ldgsts<4, 4>(occ); run_ldgsts<4, 8>(occ); run_ldgsts<4, 16>(occ);
    run_ldgsts<8, 4>(occ); run_ldgsts<8, 8>(occ); run_ldgsts<8, 16>(occ);
    run_ldgsts<16, 4>(occ); run_ldgsts<16, 8>(occ);
  }
  // --- bulk copies: rows per batch per warp 32 / 64 / 128, issued by all lanes or by one
  for (int occ : {4, 2, 1}) {
    run_bulk<64, 32, false>(256, occ);  run_bulk<64, 64, false>(256, occ);  run_bulk<64, 128, false>(256, occ);
    run_bulk<128, 32, false>(256, occ); run_bulk<128, 64, false>(256, occ);
    run_bulk<256, 32, false>(256, occ); run_bulk<256, 64, false>(256, occ);
    run_bulk<64, 32, true>(256, occ);   run_bulk<128, 32, true>(256, occ);  run_bulk<256, 32, true>(256, occ);
  }
  return 0;
}
