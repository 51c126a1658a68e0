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

#ifndef CTRFE_WIDE_U
#define CTRFE_WIDE_U 2   // rows in flight per lane with 256-bit lanes (2 = the same bytes as four 128-bit loads)
#endif

// Programmatic dependent launch: lets a kernel that was launched with the overlap flag
// (ctrfe_plan_launch_ex, CTRFE_LAUNCH_OVERLAP_PREVIOUS) start filling SMs as soon as every CTA
// of this grid has been scheduled, i.e. during this grid's tail.  A no-op otherwise.
__device__ __forceinline__ void launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// Epilogue shared by the specialised kernels: pad / scale / CVM / one 128-bit store per lane
// (same arithmetic as the general kernel).  Every lane of the warp must call it (shuffles).
template <int DT, int GS, bool CVM, int CB = 16>
__device__ __forceinline__ void finish_segment(const KParams& kp, const SlotK* slp, float (&acc)[Lane<DT, true, CB>::VEC],
                                               int len, unsigned seg, unsigned gl, bool seg_ok) {
  using L = Lane<DT, true, CB>;
  constexpr int ES = L::ES, VEC = L::VEC, RW = L::RW;
  constexpr unsigned kFull = 0xffffffffu;
  constexpr int G = 1 << GS;
  const unsigned lane_off = gl * static_cast<unsigned>(CB);
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
      if (gl < 2) lg = cvm_log1_tab(gl == 0 ? acc[0] : x1);
      const float y0 = __shfl_sync(kFull, static_cast<float>(lg), 0, G);
      float y1 = 0.0f;
      if (gl == 1) y1 = static_cast<float>(lg - static_cast<double>(y0));
      y1 = __shfl_sync(kFull, y1, 1, G);
      if (gl == 0) { acc[0] = y0; acc[1] = y1; }
    } else {
      const float y0 = static_cast<float>(cvm_log1_tab(acc[0]));
      acc[1] = static_cast<float>(cvm_log1_tab(acc[1]) - static_cast<double>(y0));
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
  char* dst = slp->out + static_cast<unsigned long long>(seg) * slp->out_ld_bytes + lane_off;
#pragma unroll
  for (int q = 0; q < RW / 4; ++q)
    reinterpret_cast<uint4*>(dst)[q] = make_uint4(packed[4 * q], packed[4 * q + 1], packed[4 * q + 2], packed[4 * q + 3]);
}

// Where rows come from and how an id becomes a row index: the slot's table, the shards of a
// row-sharded table, or the hot-row replicas.
template <int DT, int GS, bool SHARDED, bool HOT, int CB = 16>
struct RowSource {
  static constexpr unsigned kRowShift = GS + (CB == 32 ? 5 : 4);  // row_bytes = CB << GS
  const KParams& kp;
  const SlotK* slp;
  const unsigned long long* s_shards;  // shared-memory copy of kp.shard_tables (SHARDED)
  unsigned lane_off;
  unsigned long long lane_base;
  unsigned long long hot_base;

  __device__ __forceinline__ RowSource(const KParams& kp_, const SlotK* slp_, const unsigned long long* sh, unsigned gl)
      : kp(kp_), slp(slp_), s_shards(sh), lane_off(gl * static_cast<unsigned>(CB)),
        lane_base(reinterpret_cast<unsigned long long>(slp_->table) + gl * static_cast<unsigned>(CB)),
        hot_base(HOT ? reinterpret_cast<unsigned long long>(kp_.hot_rows) + gl * static_cast<unsigned>(CB) : 0ull) {
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
    atomicMin(kp.err, err_key(kp, slp, pos));
    return kSkip;
  }
};

// Adds U loaded rows into the accumulator, in order.
template <int DT, int U, int CB = 16>
__device__ __forceinline__ void add_rows(float (&acc)[Lane<DT, true, CB>::VEC],
                                         const uint32_t (&raw)[U][Lane<DT, true, CB>::RW], const bool (&ok)[U]) {
  using L = Lane<DT, true, CB>;
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
template <int DT, int GS, bool SHARDED, bool HOT, int CB = 16>
__device__ __forceinline__ void pool_rows_by_warp(const RowSource<DT, GS, SHARDED, HOT, CB>& src, uint32_t start, int len,
                                                  unsigned gl, float (&acc)[Lane<DT, true, CB>::VEC]) {
  using L = Lane<DT, true, CB>;
  constexpr int RW = L::RW;
  constexpr unsigned kFull = 0xffffffffu;
  constexpr int G = 1 << GS;
  constexpr int U = CB == 32 ? CTRFE_WIDE_U : kUnroll;  // row loads in flight per lane
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
      add_rows<DT, U, CB>(acc, raw, ok);
    }
#pragma unroll
    for (int i = 0; i < NI; ++i) cur[i] = nxt[i];
  }
}

// CB: bytes of a row one lane holds.  16: one LDG.E.128 per lane per row.  32: one 256-bit load
// (LDG.E.ENL2.256).  benchtools/gather_sweep.cu measured what uniform-random rows sustain on B200
// (profiles/r2_gather_sweep.txt): 128-byte rows 3.75 TB/s through 128-bit lanes against 5.26 TB/s
// through 256-bit lanes, 256-byte rows 3.79 against 6.52 TB/s (the streaming-copy peak); 64-byte rows
// 2.82 TB/s either way; one cp.async.bulk per row 2.26 / 4.40 / 5.64 TB/s.  Rows of 128 bytes and
// wider therefore take CB = 32.
#ifndef CTRFE_MINB_WIDE32
#define CTRFE_MINB_WIDE32 6
#endif
#ifndef CTRFE_MINB_WIDE16
#define CTRFE_MINB_WIDE16 5
#endif
template <int DT, int CB>
constexpr int fast_min_blocks() {
  if (CB == 16) return DT == CTRFE_F32 ? CTRFE_MINB32 : CTRFE_MINB16;
  return DT == CTRFE_F32 ? CTRFE_MINB_WIDE32 : CTRFE_MINB_WIDE16;
}

template <int DT, int GS, bool CVM, bool SHARDED, bool HOT = false, int CB = 16>
__global__ void __launch_bounds__(kThreads, (fast_min_blocks<DT, CB>()))
fused_pool_fast_kernel(const __grid_constant__ KParams kp) {
  launch_dependents();
  using L = Lane<DT, true, CB>;
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

  uint32_t start = 0u;
  int len = 0;
  if (seg_ok) seg_bounds(kp, slp, seg, start, len);

  const RowSource<DT, GS, SHARDED, HOT, CB> src(kp, slp, s_shards, gl);
  float acc[VEC];
#pragma unroll
  for (int e = 0; e < VEC; ++e) acc[e] = 0.0f;
  pool_rows_by_warp<DT, GS, SHARDED, HOT, CB>(src, start, len, gl, acc);
  finish_segment<DT, GS, CVM, CB>(kp, slp, acc, len, seg, gl, seg_ok);
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
