// ctrfe_fast_kernel.cuh -- the specialised instantiation of the fused gather + pool + CVM
// kernel for the case every BASELINE config is in:
//   * rows are exactly 16 * G bytes (G = 1..32 lanes, a power of two) and 16-byte aligned,
//     outputs 16-byte aligned: every lane always holds a valid 128-bit chunk;
//   * SUM / AVERAGE / SQRT slots, accumulator starts at zero, no padding_idx;
//   * ids present, int64; CVM on or absent.
// ncu showed the general kernel bound by warp-instruction issue (profiles/README.md); here the
// row width is a compile-time constant and every per-row decision that the general kernel
// takes at run time (zero rows, first-row copy, column masks, tiles) is gone:
// a row slot is SHFL + compare + address + LDG.128 + two FADD2.
// Same arithmetic, same order, same results as ctrfe::fused_pool_kernel (the GPU tests run
// both paths on the same inputs).  An out-of-range id is flagged and its row skipped.
#pragma once

#include "ctrfe_kernels.cuh"

namespace ctrfe {

// Programmatic dependent launch: lets a kernel that was launched with the overlap flag
// (ctrfe_plan_launch_ex, CTRFE_LAUNCH_OVERLAP_PREVIOUS) start filling SMs as soon as every CTA
// of this grid has been scheduled, i.e. during this grid's tail.  A no-op otherwise.
__device__ __forceinline__ void launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// Epilogue shared by the specialised kernels: pad / scale / CVM / one 128-bit store per lane
// (same arithmetic as the general kernel).  Every lane of the warp must call it (shuffles).
template <int DT, int GS, bool CVM>
__device__ __forceinline__ void finish_segment(const KParams& kp, const SlotK* slp, float (&acc)[Lane<DT, true, 16>::VEC],
                                               int len, unsigned seg, unsigned gl, bool seg_ok) {
  using L = Lane<DT, true, 16>;
  constexpr int ES = L::ES, VEC = L::VEC, RW = L::RW;
  constexpr unsigned kFull = 0xffffffffu;
  constexpr int G = 1 << GS;
  const unsigned lane_off = gl * 16u;
  const int pooltype = slp->pooltype;
  if (len == 0) {
#pragma unroll
    for (int e = 0; e < VEC; ++e) acc[e] = kp.pad_value;
  } else if (pooltype != CTRFE_POOL_SUM) {
    const float h = static_cast<float>(len);
    if (kp.scale_mode == CTRFE_SCALE_DIVIDE) {
      const float d = pooltype == CTRFE_POOL_AVERAGE ? h : sqrtf(h);
#pragma unroll
      for (int e = 0; e < VEC; ++e) acc[e] = acc[e] / d;
    } else {
      float s = 1.0f;
      s = pooltype == CTRFE_POOL_AVERAGE ? s / h : s / sqrtf(h);
#pragma unroll
      for (int e = 0; e < VEC; ++e) acc[e] = s * acc[e];
    }
  }

  if constexpr (CVM) {
    if constexpr (G >= 2) {
      const float x1 = __shfl_sync(kFull, acc[1], 0, G);
      double lg = 0.0;
      if (gl < 2) lg = cvm_log1(gl == 0 ? acc[0] : x1);
      const float y0 = __shfl_sync(kFull, static_cast<float>(lg), 0, G);
      float y1 = 0.0f;
      if (gl == 1) y1 = static_cast<float>(lg - static_cast<double>(y0));
      y1 = __shfl_sync(kFull, y1, 1, G);
      if (gl == 0) { acc[0] = y0; acc[1] = y1; }
    } else {
      const float y0 = static_cast<float>(cvm_log1(acc[0]));
      acc[1] = static_cast<float>(cvm_log1(acc[1]) - static_cast<double>(y0));
      acc[0] = y0;
    }
  }

  if (!seg_ok) return;
  uint32_t packed[RW];
#pragma unroll
  for (int r = 0; r < RW; ++r) packed[r] = 0u;
#pragma unroll
  for (int e = 0; e < VEC; ++e) {
    const uint32_t b = float_to_elem<DT>(acc[e]);
    if constexpr (ES == 4) packed[e] = b;
    else packed[e >> 1] |= b << ((e & 1) * 16);
  }
  char* dst = slp->out + static_cast<long long>(seg) * slp->out_ld_bytes + lane_off;
  *reinterpret_cast<uint4*>(dst) = make_uint4(packed[0], packed[1], packed[2], packed[3]);
}

// Where rows come from and how an id becomes a row index: the slot's table, the shards of a
// row-sharded table, or the hot-row replicas.
template <int DT, int GS, bool SHARDED, bool HOT>
struct RowSource {
  static constexpr unsigned kRowShift = GS + 4;  // row_bytes = 16 << GS
  const KParams& kp;
  const SlotK* slp;
  const unsigned long long* s_shards;  // shared-memory copy of kp.shard_tables (SHARDED)
  unsigned lane_off;
  unsigned long long lane_base;
  unsigned long long hot_base;

  __device__ __forceinline__ RowSource(const KParams& kp_, const SlotK* slp_, const unsigned long long* sh, unsigned gl)
      : kp(kp_), slp(slp_), s_shards(sh), lane_off(gl * 16u),
        lane_base(reinterpret_cast<unsigned long long>(slp_->table) + gl * 16u),
        hot_base(HOT ? reinterpret_cast<unsigned long long>(kp_.hot_rows) + gl * 16u : 0ull) {
    // keep the lane's base address in registers: left alone, ptxas re-derives it from the
    // parameter block in front of every row load (9 instructions per row)
    asm volatile("" : "+l"(lane_base));
  }

  __device__ __forceinline__ const char* row_ptr(uint32_t rid) const {
    if constexpr (HOT) {
      if (rid & kHotBit)  // a local replica of a hot row
        return reinterpret_cast<const char*>(hot_base + (static_cast<unsigned long long>(rid & ~kHotBit) << kRowShift));
    }
    if constexpr (SHARDED) {
      const unsigned shard_mask = static_cast<unsigned>(kp.n_shards - 1);
      return reinterpret_cast<const char*>(s_shards[rid & shard_mask] +
                                           (static_cast<unsigned long long>(rid >> kp.shard_shift) << kRowShift) + lane_off);
    } else {
      return reinterpret_cast<const char*>(lane_base + (static_cast<unsigned long long>(rid) << kRowShift));
    }
  }

  // row index of the id at position `pos` of the slot's id array; kSkip (and the error flag)
  // for an id outside the table
  __device__ __forceinline__ uint32_t resolve(unsigned pos) const {
    // int64 ids only: slots with int32 ids take the general kernel (the per-fetch test of the
    // id width cost 3 % on cfg2)
    const uint2 v = __ldg(reinterpret_cast<const uint2*>(slp->ids) + pos);
    const uint32_t lo = v.x, hi = v.y;
    if (hi == 0u && lo < slp->table_rows) {
      if constexpr (HOT) {
        // is there a local replica?  open addressing, linear probing, bounded probe count
        unsigned h = (lo * 2654435761u) >> kp.hot_shift;
#pragma unroll 1
        for (int pr = 0; pr < kHotProbes; ++pr) {
          const unsigned k = __ldg(kp.hot_keys + h);
          if (k == lo) return kHotBit | h;  // the replica lives at the hash slot itself
          if (k == 0xffffffffu) break;
          h = (h + 1u) & kp.hot_mask;
        }
      }
      return lo;
    }
    atomicMin(kp.err, (static_cast<unsigned long long>(slp->slot_index) << 44) | pos);
    return kSkip;
  }
};

// Adds U loaded rows into the accumulator, in order.
template <int DT, int U>
__device__ __forceinline__ void add_rows(float (&acc)[Lane<DT, true, 16>::VEC],
                                         const uint32_t (&raw)[U][Lane<DT, true, 16>::RW], const bool (&ok)[U]) {
  using L = Lane<DT, true, 16>;
#pragma unroll
  for (int u = 0; u < U; ++u) {
    if (!ok[u]) continue;
    if constexpr (L::ES == 4) {
#pragma unroll
      for (int e = 0; e < L::VEC; e += 2)
        add_f32x2(acc[e], acc[e + 1], __uint_as_float(raw[u][e]), __uint_as_float(raw[u][e + 1]));
    } else {
#pragma unroll
      for (int r = 0; r < L::RW; ++r) {
        const float2 x = L::unpack2(raw[u][r]);
        add_f32x2(acc[2 * r], acc[2 * r + 1], x.x, x.y);
      }
    }
  }
}

// The warp-level row loop: lanes of a group fetch G consecutive ids with one request and
// broadcast them with shuffles; the whole warp walks to its longest sequence.
template <int DT, int GS, bool SHARDED, bool HOT>
__device__ __forceinline__ void pool_rows_by_warp(const RowSource<DT, GS, SHARDED, HOT>& src, uint32_t start, int len,
                                                  unsigned gl, float (&acc)[Lane<DT, true, 16>::VEC]) {
  using L = Lane<DT, true, 16>;
  constexpr int RW = L::RW;
  constexpr unsigned kFull = 0xffffffffu;
  constexpr int G = 1 << GS;
  constexpr int U = kUnroll;
  constexpr int NI = G >= U ? 1 : U / G;  // ids held per lane
  constexpr int R = G * NI;               // rows consumed per step of the sequence loop
  const int maxlen = __reduce_max_sync(kFull, len);
  auto fetch_rid = [&](int off) -> uint32_t {
    return off < len ? src.resolve(start + static_cast<unsigned>(off)) : kSkip;
  };
  uint32_t cur[NI];
#pragma unroll
  for (int i = 0; i < NI; ++i) cur[i] = fetch_rid(i * G + static_cast<int>(gl));

  for (int base = 0; base < maxlen; base += R) {
    uint32_t nxt[NI];
#pragma unroll
    for (int i = 0; i < NI; ++i) nxt[i] = kSkip;
    if (base + R < maxlen) {  // warp-uniform: only long sequences come back for more ids
#pragma unroll
      for (int i = 0; i < NI; ++i) nxt[i] = fetch_rid(base + R + i * G + static_cast<int>(gl));
    }
#pragma unroll
    for (int c = 0; c < R / U; ++c) {
      if (c > 0 && base + c * U >= maxlen) break;  // warp-uniform
      uint32_t raw[U][RW];
      bool ok[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int j = c * U + u;
        const uint32_t rid = __shfl_sync(kFull, cur[j / G], j % G, G);
        ok[u] = rid != kSkip;
        if (ok[u]) L::load(src.row_ptr(rid), raw[u]);
      }
      add_rows<DT, U>(acc, raw, ok);
    }
#pragma unroll
    for (int i = 0; i < NI; ++i) cur[i] = nxt[i];
  }
}

template <int DT, int GS, bool CVM, bool SHARDED, bool HOT = false>
__global__ void __launch_bounds__(kThreads, (DT == CTRFE_F32 ? CTRFE_MINB32 : CTRFE_MINB16))
fused_pool_fast_kernel(const __grid_constant__ KParams kp) {
  launch_dependents();
  using L = Lane<DT, true, 16>;
  constexpr int VEC = L::VEC;
  constexpr int G = 1 << GS;

  __shared__ unsigned long long s_shards[SHARDED ? kMaxShards : 1];
  if constexpr (SHARDED) {
    if (threadIdx.x < kMaxShards) s_shards[threadIdx.x] = reinterpret_cast<unsigned long long>(kp.shard_tables[threadIdx.x]);
    __syncthreads();
  }

  const SlotK* const slp = &kp.slots[blockIdx.y];
  const unsigned n_seg = slp->n_seg;
  if (((blockIdx.x * static_cast<unsigned>(kThreads)) >> GS) >= n_seg) return;  // CTA past this slot's work
  const unsigned seg = (blockIdx.x * static_cast<unsigned>(kThreads) + threadIdx.x) >> GS;
  const unsigned gl = threadIdx.x & (G - 1);
  const bool seg_ok = seg < n_seg;

  uint32_t start = 0u, endo = 0u;
  if (seg_ok) {
    const unsigned* w = reinterpret_cast<const unsigned*>(slp->lod) + 2u * seg;  // low words: offsets < 2^32
    start = __ldg(w);
    endo = __ldg(w + 2);
  }
  const int len = static_cast<int>(endo - start);

  const RowSource<DT, GS, SHARDED, HOT> src(kp, slp, s_shards, gl);
  float acc[VEC];
#pragma unroll
  for (int e = 0; e < VEC; ++e) acc[e] = 0.0f;
  pool_rows_by_warp<DT, GS, SHARDED, HOT>(src, start, len, gl, acc);
  finish_segment<DT, GS, CVM>(kp, slp, acc, len, seg, gl, seg_ok);
}

__device__ __forceinline__ void cp_async_16(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(static_cast<unsigned>(__cvta_generic_to_shared(smem_dst))),
               "l"(gsrc)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// ---- balanced loads -------------------------------------------------------------------------
// In the kernel above a lane group walks ITS segment four rows at a time, so a warp runs to its
// longest segment, about half of the row-load slots are empty with short ragged sequences
// (mean 4, Poisson), and -- what ncu shows to matter most -- a warp's life is a chain of
// dependent memory latencies (LoD -> ids -> rows -> rows ...) paid once per 32 / G segments.
// Here a warp takes M * 32 / G adjacent segments.  Their ids are one contiguous run, loaded one
// per lane with coalesced requests; the rows of the run are fetched in batches of up to
// 4 * M * 32 / G, every lane group taking exactly its share whatever segment a row belongs to,
// and they go straight from global to shared memory with cp.async (no registers tied up, all
// of a batch in flight at once).  Each group then adds the rows of ITS M segments from shared
// memory, in sequence order -- so the arithmetic and its order are unchanged and results stay
// bit-identical -- and runs the epilogue M times.  Per warp: one LoD wait, one id wait and
// ceil(rows / batch) row waits for M times the segments.
// Measured on B200 (profiles/README.md): cfg3 (128-byte bf16 rows) 328 -> 231 us with M = 2 and
// eight rows in flight per lane; four in flight: 257 us; M = 1: 247 us.  64-byte rows (cfg2) lose
// 10-18 % in every variant -- the staging buffers take the L1 capacity that serves the Zipf
// head -- and rows behind NVLink lose more, so the host uses this kernel for unsharded
// 128-byte rows only.
#ifndef CTRFE_BAL_M
#define CTRFE_BAL_M 2
#endif
#ifndef CTRFE_BAL_M64
#define CTRFE_BAL_M64 CTRFE_BAL_M
#endif
#ifndef CTRFE_BAL_MINB
#define CTRFE_BAL_MINB 6
#endif
constexpr int kBalM128 = CTRFE_BAL_M;   // segments per lane group, 128-byte rows
constexpr int kBalM64 = CTRFE_BAL_M64;  // 64-byte rows (wider rows: 1)

template <int DT, int GS, bool CVM, bool SHARDED, bool HOT, int M, int MODE = kModeAdd>
__global__ void __launch_bounds__(kThreads, CTRFE_BAL_MINB)
fused_pool_bal_kernel(const __grid_constant__ KParams kp) {
  launch_dependents();
  using L = Lane<DT, true, 16>;
  static_assert(GS >= 2 && GS <= 5, "64- to 512-byte rows");
  static_assert(MODE == kModeAdd || (!CVM && M == 1), "the selecting (MAX) form has no CVM and one segment per group");
  constexpr int VEC = L::VEC;
  constexpr unsigned kFull = 0xffffffffu;
  constexpr int G = 1 << GS;
  constexpr int NG = 32 >> GS;        // lane groups per warp
  constexpr int WS = NG * M;          // segments per warp
#ifdef CTRFE_BAL_U
  constexpr int U = CTRFE_BAL_U;
#else
  constexpr int U = GS == 2 ? kUnroll * M : 8;  // rows in flight per lane
#endif
  constexpr int RB = U * NG;          // rows per batch
  constexpr int IPL = RB / 32 > 0 ? RB / 32 : 1;  // ids per lane per batch
  constexpr int kWarps = kThreads / 32;
  static_assert(WS + 1 <= 32 && (RB % 32 == 0 || RB < 32), "one LoD offset per lane");

  __shared__ uint4 s_rows_all[kWarps][U * 32];
  __shared__ unsigned long long s_shards[SHARDED ? kMaxShards : 1];
  if constexpr (SHARDED) {
    if (threadIdx.x < kMaxShards) s_shards[threadIdx.x] = reinterpret_cast<unsigned long long>(kp.shard_tables[threadIdx.x]);
    __syncthreads();
  }

  const SlotK* const slp = &kp.slots[blockIdx.y];
  const unsigned n_seg = slp->n_seg;
  const unsigned wib = threadIdx.x >> 5;
  const unsigned seg0 = (blockIdx.x * kWarps + wib) * WS;  // first segment of this warp
  if (blockIdx.x * kWarps * WS >= n_seg) return;           // CTA past this slot's work
  const unsigned lane = threadIdx.x & 31u;
  uint4* const s_rows = s_rows_all[wib];
  const unsigned g = lane >> GS;
  const unsigned gl = lane & (G - 1);

  // the WS + 1 offsets of the warp's segments, one per lane (low words: offsets < 2^32)
  uint32_t my_off = 0u;
  if (lane <= static_cast<unsigned>(WS)) {
    const unsigned i = min(seg0 + lane, n_seg);
    my_off = __ldg(reinterpret_cast<const unsigned*>(slp->lod) + 2u * i);
  }
  const uint32_t run0 = __shfl_sync(kFull, my_off, 0);
  const uint32_t run1 = __shfl_sync(kFull, my_off, WS);

  const RowSource<DT, GS, SHARDED, HOT> src(kp, slp, s_shards, gl);
  float acc[M][VEC];
  uint32_t sel[M][4];  // selecting form on 16-bit types: raw bits of the selected elements
#pragma unroll
  for (int m = 0; m < M; ++m) {
#pragma unroll
    for (int e = 0; e < VEC; ++e) acc[m][e] = 0.0f;
#pragma unroll
    for (int r = 0; r < 4; ++r) sel[m][r] = 0u;
  }

  // one loaded row (this lane's 16 bytes) into segment m's state; `first`: the segment's row 0
  auto take_row = [&](int m, const uint4& v, bool first) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    if constexpr (MODE == kModeAdd) {
      if constexpr (L::ES == 4) {
        add_f32x2(acc[m][0], acc[m][1], __uint_as_float(w[0]), __uint_as_float(w[1]));
        add_f32x2(acc[m][2], acc[m][3], __uint_as_float(w[2]), __uint_as_float(w[3]));
      } else {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
          const float2 x = L::unpack2(w[r]);
          add_f32x2(acc[m][2 * r], acc[m][2 * r + 1], x.x, x.y);
        }
      }
    } else {
      // MAX: out = row 0, then strictly-greater replaces (math/sequence_pooling.cc:166-174);
      // NaN never replaces and a NaN in row 0 sticks, as on the CPU
      if constexpr (L::ES == 4) {
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float x = __uint_as_float(w[e]);
          if (first || x > acc[m][e]) acc[m][e] = x;
        }
      } else {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
          const float2 x = L::unpack2(w[r]);
          uint32_t mask = 0u;
          if (first || x.x > acc[m][2 * r]) { acc[m][2 * r] = x.x; mask |= 0x0000ffffu; }
          if (first || x.y > acc[m][2 * r + 1]) { acc[m][2 * r + 1] = x.y; mask |= 0xffff0000u; }
          sel[m][r] = (sel[m][r] & ~mask) | (w[r] & mask);
        }
      }
    }
  };

  // lane L holds the row indices of batch positions L, L + 32, ...
  auto stage_ids = [&](uint32_t p0, uint32_t (&ids)[IPL]) {
#pragma unroll
    for (int i = 0; i < IPL; ++i) {
      const uint32_t pos = p0 + i * 32u + lane;
      ids[i] = (i * 32u + lane < static_cast<unsigned>(RB) && pos < run1) ? src.resolve(pos) : kSkip;
    }
  };
  uint32_t cur[IPL];
  stage_ids(run0, cur);
#pragma unroll 1
  for (uint32_t p0 = run0; p0 < run1; p0 += RB) {
    uint32_t nxt[IPL];
#pragma unroll
    for (int i = 0; i < IPL; ++i) nxt[i] = kSkip;
    if (p0 + RB < run1) stage_ids(p0 + RB, nxt);  // warp-uniform: more than one batch
    __syncwarp();  // every lane is done reading the previous batch out of shared memory
    // batch row r = u * NG + g is fetched by group g as its u-th row; a skipped row reads as zeros
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int r = u * NG + static_cast<int>(g);
      const uint32_t rid = __shfl_sync(kFull, cur[(u * NG) / 32], r & 31);
      uint4* const dst = s_rows + u * 32 + lane;
      if (rid != kSkip) cp_async_16(dst, src.row_ptr(rid));
      else *dst = make_uint4(0u, 0u, 0u, 0u);
    }
    cp_async_commit();
    cp_async_wait_all();
    __syncwarp();
#pragma unroll
    for (int m = 0; m < M; ++m) {
      const uint32_t start = __shfl_sync(kFull, my_off, static_cast<int>(g) * M + m);
      const uint32_t endo = __shfl_sync(kFull, my_off, static_cast<int>(g) * M + m + 1);
      // rows of this segment that sit in the batch: batch positions [j0, j1)
      const int j0 = static_cast<int>(max(start, p0) - p0);
      const int j1 = static_cast<int>(min(endo, p0 + RB)) - static_cast<int>(p0);
#pragma unroll 1
      for (int j = j0; j < j1; j += 2) {
        const uint4 a = s_rows[j * G + gl];
        uint4 b = make_uint4(0u, 0u, 0u, 0u);
        const bool two = j + 1 < j1;
        if (two) b = s_rows[(j + 1) * G + gl];
        take_row(m, a, p0 + static_cast<uint32_t>(j) == start);
        if (two) take_row(m, b, false);
      }
    }
#pragma unroll
    for (int i = 0; i < IPL; ++i) cur[i] = nxt[i];
  }
#pragma unroll
  for (int m = 0; m < M; ++m) {
    const unsigned seg = seg0 + g * M + m;
    const uint32_t start = __shfl_sync(kFull, my_off, static_cast<int>(g) * M + m);
    const uint32_t endo = __shfl_sync(kFull, my_off, static_cast<int>(g) * M + m + 1);
    if constexpr (MODE == kModeAdd) {
      finish_segment<DT, GS, CVM>(kp, slp, acc[m], static_cast<int>(endo - start), seg, gl, seg < n_seg);
    } else {
      if (seg < n_seg) {  // the selected bits as they are; an empty sequence takes pad_value
        uint32_t packed[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) {
          if constexpr (L::ES == 4) packed[r] = __float_as_uint(endo == start ? kp.pad_value : acc[m][r]);
          else packed[r] = endo == start ? float_to_elem<DT>(kp.pad_value) * 0x00010001u : sel[m][r];
        }
        char* dst = slp->out + static_cast<long long>(seg) * slp->out_ld_bytes + gl * 16u;
        *reinterpret_cast<uint4*>(dst) = make_uint4(packed[0], packed[1], packed[2], packed[3]);
      }
    }
  }
}

// Fills a hot-row cache: one warp per hot id claims a hash slot (open addressing, linear
// probing) and copies the id's row from src_rows[i] to rows[slot].  An id that finds no free
// slot within kHotProbes is left out.
__global__ void hot_insert_kernel(const long long* __restrict__ hot_ids, long long n, const char* __restrict__ src_rows,
                                  unsigned row_bytes, unsigned* keys, char* rows, unsigned mask, int shift) {
  const long long i = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  if (i >= n) return;
  const unsigned lane = threadIdx.x & 31u;
  unsigned slot = 0xffffffffu;
  if (lane == 0) {
    const long long id = hot_ids[i];
    if (id >= 0 && id < 0x7fffffffLL) {  // the hot bit needs row ids below 2^31
      const unsigned lo = static_cast<unsigned>(id);
      unsigned h = (lo * 2654435761u) >> shift;
      for (int pr = 0; pr < kHotProbes; ++pr) {
        const unsigned old = atomicCAS(keys + h, 0xffffffffu, lo);
        if (old == 0xffffffffu || old == lo) { slot = h; break; }
        h = (h + 1u) & mask;
      }
    }
  }
  slot = __shfl_sync(0xffffffffu, slot, 0);
  if (slot == 0xffffffffu) return;
  const uint4* src = reinterpret_cast<const uint4*>(src_rows + static_cast<unsigned long long>(i) * row_bytes);
  uint4* dst = reinterpret_cast<uint4*>(rows + static_cast<unsigned long long>(slot) * row_bytes);
  for (unsigned c = lane; c < row_bytes / 16u; c += 32u) dst[c] = src[c];
}

}  // namespace ctrfe
