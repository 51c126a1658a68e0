// ctrfe_kernels.cuh -- sm_100a device code for the CTR sparse front end.
//
// One kernel family does the whole hot path: for every LoD segment of every slot,
// gather the rows its ids name, reduce them IN ROW ORDER with one fp32 accumulator
// per column, apply the CVM transform, and store the pooled row once.  Nothing is
// materialised in between (the reference writes and re-reads [K,D] and [B,D]).
//
// Mapping: a "group" of G = 2^g_shift adjacent lanes owns one (segment, column-tile)
// work item; each lane owns one 16-byte chunk of the row (VECT) or one element
// (scalar fallback for widths/alignments that do not allow 128-bit access).  Lanes of
// a group load G consecutive ids with one coalesced request and broadcast them with
// warp shuffles; the U row loads of a batch are all issued before the first dependent
// add, so each lane keeps U independent 128-bit loads in flight while the adds stay
// strictly sequential per column (bit-exact with the reference's order:
// jit/refer/refer.h:372-380, math/sequence_pooling.cu:93-97).
//
// Compiled with -ftz=true -fmad=false: the reference CPU executor runs under FTZ+DAZ
// (framework/naive_executor.cc:66) and never contracts a*b+c on this path.
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>

#include <type_traits>

#include "../../include/ctrfe.h"

namespace ctrfe {

// Resident CTAs per SM the register allocator must leave room for (tuned on B200 with ncu;
// overridable at build time for experiments).
#ifndef CTRFE_MINB32
#define CTRFE_MINB32 8
#endif
#ifndef CTRFE_MINB16
#define CTRFE_MINB16 8
#endif

#ifndef CTRFE_THREADS
#define CTRFE_THREADS 256
#endif
constexpr int kThreads = CTRFE_THREADS; // threads per CTA of the pooling kernel
#ifndef CTRFE_UNROLL
#define CTRFE_UNROLL 4
#endif
constexpr int kUnroll = CTRFE_UNROLL;   // row loads in flight per lane
constexpr int kMaxSlotsPerLaunch = 64; // slot descriptors carried in the kernel parameters (block <= 4 KB)
constexpr int kMaxShards = 8;           // GPUs of one NVSwitch box
constexpr uint32_t kSkip = 0xffffffffu;  // row slot past the end of the sequence: no load, no add
constexpr uint32_t kZero = 0xfffffffeu;  // padding_idx / out-of-range id: contributes a zero row
constexpr uint32_t kHotBit = 0x80000000u; // row index refers to the hot-row replica cache (sharded mode)
constexpr int kHotProbes = 8;             // linear probes of the hot-row hash
constexpr int kKeyProbes = 32;            // linear probes of the feasign key map
constexpr unsigned long long kKeyHashMul = 0x9E3779B97F4A7C15ull;

// Slot descriptor, passed BY VALUE in the kernel parameter block (constant bank): a CTA
// reads its slot's pointers without a dependent global load.
struct SlotK {
  const char* table;               // [table_rows, dim] or dense X when id_kind == 2
  const void* ids;                 // int64 / int32
  const unsigned long long* lod;   // device offsets [n_seg + 1]
  char* out;
  unsigned int out_ld_bytes;       // bytes between consecutive output rows (< 2^32)
  unsigned int table_rows;         // < 2^32 - 2
  unsigned int n_seg;
  unsigned int n_ids;              // entries in ids (dense rows: table_rows); every LoD offset must be <= it (< 2^31)
  signed char pooltype;
  unsigned char lod32;             // the LoD image holds 32-bit offsets (the library's own upload of a host LoD)
  short id_kind;                   // 0 int64, 1 int32, 2 dense (row index = position), 3 uint64 feasign keys (key map)
  int slot_index;                  // position in the caller's slot array (bad-id report)
};
static_assert(sizeof(SlotK) == 56, "slot descriptor layout");

// Error records: err[0] = min over keys of out-of-range ids, err[1] = min over keys of malformed
// LoD pairs met by a kernel; ~0 = none.  key = tag << 56 | slot << 44 | position (or sequence),
// tag = the plan / call the launch belongs to (0 for the static entry point).
constexpr int kErrSlotShift = 44;
constexpr int kErrTagShift = 56;

struct KParams {
  unsigned long long* err;   // [2], see above
  unsigned long long err_tag;  // tag << kErrTagShift, OR-ed into every key this launch records
  long long padding_idx;
  int dim;
  int row_bytes;
  int tiles;                 // column tiles per row (1 unless a row is wider than one group covers)
  int cvm_mode;
  int cvm_offset;
  int scale_mode;
  int init_mode;
  int reserved;
  float pad_value;
  int n_slots;
  // row-sharded table mode: row r of the logical table lives on shard r % n_shards at local
  // row r / n_shards; shard_tables[] are device pointers valid on THIS GPU (local memory or
  // NVLink peer mappings obtained through CUDA IPC).  n_shards is a power of two, 0 = off.
  int shard_shift;
  int n_shards;
  const char* shard_tables[kMaxShards];
  // hot-row replica cache of the sharded mode (ctrfe_hot_cache): null = none
  const char* hot_rows;
  const unsigned* hot_keys;
  unsigned hot_mask;   // slots - 1
  int hot_shift;       // 32 - log2(slots)
  // feasign key map (ctrfe_keymap): ids are uint64 keys, resolved to table rows in the kernel
  const unsigned long long* km_keys;
  const unsigned* km_rows;
  unsigned km_mask;    // slots - 1
  int km_shift;        // 64 - log2(slots)
  SlotK slots[kMaxSlotsPerLaunch];
};
static_assert(sizeof(KParams) <= 4096, "kernel parameter block must stay within 4 KB");

__device__ __forceinline__ uint4 ldg128(const void* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }

__device__ __forceinline__ unsigned long long err_key(const KParams& kp, const SlotK* slp, unsigned long long pos) {
  return kp.err_tag | (static_cast<unsigned long long>(slp->slot_index) << kErrSlotShift) | pos;
}

// The [start, end) id range of sequence `seg`, with the CheckLoD rules (framework/lod_tensor.cc:
// 116-142) applied to the pair that is read: non-decreasing and inside the id array (n_ids < 2^31,
// so one 64-bit compare also rejects offsets with a non-zero high word).  An offending sequence
// is pooled as empty and recorded in err[1]; the kernel never reads past the ids it was given.
__device__ __forceinline__ void seg_bounds(const KParams& kp, const SlotK* slp, unsigned seg, uint32_t& start, int& len) {
  unsigned long long a, b;
  if (slp->lod32) {  // offsets the library uploaded itself: validated on the host, 4 bytes each
    const unsigned* w = reinterpret_cast<const unsigned*>(slp->lod);
    a = __ldg(w + seg);
    b = __ldg(w + seg + 1);
  } else {
    a = __ldg(slp->lod + seg);
    b = __ldg(slp->lod + seg + 1);
  }
  const bool ok = a <= b && b <= static_cast<unsigned long long>(slp->n_ids);
  start = static_cast<uint32_t>(a);
  len = ok ? static_cast<int>(static_cast<uint32_t>(b) - static_cast<uint32_t>(a)) : 0;
  if (!ok) atomicMin(kp.err + 1, err_key(kp, slp, seg));
}

template <int DT>
__device__ __forceinline__ float elem_to_float(uint32_t bits16_or_32) {
  if constexpr (DT == CTRFE_F32) {
    return __uint_as_float(bits16_or_32);
  } else if constexpr (DT == CTRFE_F16) {
    return __half2float(__ushort_as_half(static_cast<unsigned short>(bits16_or_32)));
  } else {
    return __uint_as_float(bits16_or_32 << 16);  // bf16 -> f32 is a shift
  }
}

template <int DT>
__device__ __forceinline__ uint32_t float_to_elem(float v) {
  if constexpr (DT == CTRFE_F32) {
    return __float_as_uint(v);
  } else if constexpr (DT == CTRFE_F16) {
    return __half_as_ushort(__float2half_rn(v));
  } else {
    return __bfloat16_as_ushort(__float2bfloat16_rn(v));
  }
}

// A lane's slice of one row: CB bytes = RW raw 32-bit words holding VEC elements.  CB is 16
// (one 128-bit load) or, for fp32 rows that are a multiple of 32 bytes, 32 (one 256-bit
// load, LDG.E.256 on sm_100): half as many lanes per row, so half as many instructions per
// byte on a path that ncu shows to be issue-bound.
template <int DT, bool VECT, int CB = 16>
struct Lane {
  static constexpr int ES = (DT == CTRFE_F32) ? 4 : 2;
  static constexpr int VEC = VECT ? CB / ES : 1;
  static constexpr int RW = VECT ? CB / 4 : 1;

  __device__ __forceinline__ static void load(const char* p, uint32_t (&w)[RW]) {
    if constexpr (VECT && CB == 32) {
      asm volatile("ld.global.nc.v8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                   : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7])
                   : "l"(p));
    } else if constexpr (VECT) {
      uint4 v = ldg128(p);
      w[0] = v.x; w[1] = v.y; w[2] = v.z; w[3] = v.w;
    } else if constexpr (ES == 4) {
      w[0] = __ldg(reinterpret_cast<const uint32_t*>(p));
    } else {
      w[0] = __ldg(reinterpret_cast<const unsigned short*>(p));
    }
  }
  // element e of the raw words, as bits (16-bit types: low 16 bits)
  __device__ __forceinline__ static uint32_t bits(const uint32_t (&w)[RW], int e) {
    if constexpr (ES == 4) return w[e];
    else return (w[e >> 1] >> ((e & 1) * 16)) & 0xffffu;
  }
  // both halves of word r as floats (16-bit types)
  __device__ __forceinline__ static float2 unpack2(uint32_t w) {
    if constexpr (DT == CTRFE_F16) {
      return __half22float2(*reinterpret_cast<const __half2*>(&w));
    } else {
      return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u));
    }
  }
};

enum { kModeAdd = 0, kModeSelect = 1 };

// (a0, a1) += (x0, x1): sm_100's packed fp32 add -- two independent IEEE round-to-nearest,
// flush-to-zero adds in one instruction, so the per-column order and rounding are unchanged.
__device__ __forceinline__ void add_f32x2(float& a0, float& a1, float x0, float x1) {
  unsigned long long a, x;
  asm("mov.b64 %0, {%1, %2};" : "=l"(a) : "f"(a0), "f"(a1));
  asm("mov.b64 %0, {%1, %2};" : "=l"(x) : "f"(x0), "f"(x1));
  asm("add.rn.ftz.f32x2 %0, %0, %1;" : "+l"(a) : "l"(x));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(a0), "=f"(a1) : "l"(a));
}

// log(x + 1) as the reference CPU kernel evaluates it: the add in fp32, the log in fp64.
//
// The fp64 log is a fifth of the fused kernel's instruction stream when evaluated with the
// CUDA math library, and the kernel is issue-bound.  Its argument here is always a float
// (24 significant bits) and its result is rounded back to float, so a shorter evaluation
// suffices: x = 2^e * m with m in [sqrt(1/2), sqrt(2)], s = (m-1)/(m+1),
// log x = e*ln2 + 2*atanh(s) with the odd series through s^13 (truncation < 5e-13 relative),
// the division by a float reciprocal refined by two Newton steps in fp64.  Measured against
// the library log on B200 over 6.7e7 arguments (benchtools/log_check.cu): max relative
// difference 1.3e-12, 7 results differ after rounding to float (by one ulp).  Anything that
// is not positive and finite goes to the library routine.
__device__ __forceinline__ double log_pos_normal(double x) {
  long long bits = __double_as_longlong(x);
  int e = static_cast<int>(bits >> 52) - 1023;
  bits = (bits & 0x000fffffffffffffLL) | 0x3ff0000000000000LL;
  double m = __longlong_as_double(bits);  // [1, 2)
  if (m > 1.4142135623730951) { m *= 0.5; e += 1; }
  const double f = m - 1.0;
  const double d = m + 1.0;
  double r = static_cast<double>(__frcp_rn(static_cast<float>(d)));
  r = fma(fma(-d, r, 1.0), r, r);  // Newton: ~2^-46
  r = fma(fma(-d, r, 1.0), r, r);  // ~2^-53
  const double s = f * r;
  const double z = s * s;
  double p = 2.0 / 13.0;
  p = fma(p, z, 2.0 / 11.0);
  p = fma(p, z, 2.0 / 9.0);
  p = fma(p, z, 2.0 / 7.0);
  p = fma(p, z, 2.0 / 5.0);
  p = fma(p, z, 2.0 / 3.0);
  const double lm = fma(s * z, p, 2.0 * s);  // 2*atanh(s)
  return fma(static_cast<double>(e), 0.6931471805599453094, lm);
}

__device__ __forceinline__ double cvm_log1(float x) {
  const double a = static_cast<double>(x + 1.0f);
  // floats are never fp64-subnormal: "positive and finite" is all that needs checking
  if (a > 0.0 && a < 1.0e300) return log_pos_normal(a);
  return log(a);  // zero, negative, inf, NaN: the library's special-case results
}

// log(x + 1) for the show / click sums CTR data actually carries: small non-negative integers.
// g_cvm_log_tab[n] holds cvm_log1(float(n)) -- filled on the device by cvm_table_fill_kernel with
// the very routine above, so a table hit returns the same bits as the evaluation it replaces
// (tests/test_gpu_parity.py compares both paths); anything else falls through to cvm_log1.
// The fp64 evaluation was 37 % of the specialised kernel's instruction stream.
constexpr int kCvmTab = 4096;
__device__ double g_cvm_log_tab[kCvmTab];

__global__ void cvm_table_fill_kernel() {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n < kCvmTab) g_cvm_log_tab[n] = cvm_log1(static_cast<float>(n));
}

__device__ __forceinline__ double cvm_log1_tab(float x) {
#ifndef CTRFE_NO_CVM_TABLE
  if (x >= 0.0f && x < static_cast<float>(kCvmTab)) {
    const int n = static_cast<int>(x);
    if (static_cast<float>(n) == x) return __ldg(&g_cvm_log_tab[n]);
  }
#endif
  return cvm_log1(x);
}

// Pin a value in a register: stops ptxas from re-deriving it from the (dynamically indexed)
// parameter bank inside the row loop.
#define CTRFE_KEEP64(x) asm volatile("" : "+l"(x))
#define CTRFE_KEEP32(x) asm volatile("" : "+r"(x))

// ---------------------------------------------------------------------------------------
// fused gather + segmented pool + CVM.
// grid = (ceil(max_items * G / kThreads), slots in this launch); G = 1 << GS lanes per item.
// MODE: kModeAdd for SUM / AVERAGE / SQRT slots, kModeSelect for MAX / LAST / FIRST slots
// (the host puts slots of different modes into different launches).
// ---------------------------------------------------------------------------------------
template <int DT, bool VECT, int GS, int MODE, int CB, bool SHARDED = false>
__global__ void __launch_bounds__(kThreads, (DT == CTRFE_F32 ? CTRFE_MINB32 : CTRFE_MINB16))
fused_pool_kernel(const __grid_constant__ KParams kp) {
  // lets a launch made with CTRFE_LAUNCH_OVERLAP_PREVIOUS start during this grid's tail
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  using L = Lane<DT, VECT, CB>;
  // sharded mode: the shard base pointers sit in shared memory (the owner differs per lane,
  // and a lane-varying index into the parameter bank would serialise)
  __shared__ unsigned long long s_shards[SHARDED ? kMaxShards : 1];
  if constexpr (SHARDED) {
    if (threadIdx.x < kMaxShards) s_shards[threadIdx.x] = reinterpret_cast<unsigned long long>(kp.shard_tables[threadIdx.x]);
    __syncthreads();
  }
  constexpr int ES = L::ES, VEC = L::VEC, RW = L::RW;
  constexpr unsigned kFull = 0xffffffffu;
  constexpr int G = 1 << GS;
  constexpr int U = kUnroll * 16 / CB;               // row loads in flight per lane (same bytes)
  constexpr int NI = G >= kUnroll ? 1 : kUnroll / G;  // ids held per lane
  constexpr int R = G * NI;                           // rows consumed per step of the sequence loop
  constexpr bool selecting = MODE == kModeSelect;

  const SlotK* const slp = &kp.slots[blockIdx.y];
  const unsigned n_seg = slp->n_seg;
  const int tiles = kp.tiles;
  const unsigned long long first_item = (static_cast<unsigned long long>(blockIdx.x) * kThreads) >> GS;
  if (first_item >= static_cast<unsigned long long>(n_seg) * tiles) return;  // CTA past this slot's work

  const int pooltype = slp->pooltype;
  const int lane = threadIdx.x & 31;
  const int gl = lane & (G - 1);
  const unsigned long long item = (static_cast<unsigned long long>(blockIdx.x) * kThreads + threadIdx.x) >> GS;
  unsigned seg = static_cast<unsigned>(item);
  int tile = 0;
  if (tiles > 1) { seg = static_cast<unsigned>(item / tiles); tile = static_cast<int>(item - static_cast<unsigned long long>(seg) * tiles); }
  const bool seg_ok = seg < n_seg;
  const int col0 = (tile * G + gl) * VEC;
  const bool col_ok = col0 < kp.dim;

  long long start = 0, end = 0;
  if (seg_ok) {
    uint32_t s0;
    int l0;
    seg_bounds(kp, slp, seg, s0, l0);
    start = static_cast<long long>(s0);
    end = start + l0;
  }
  const int seg_len = static_cast<int>(end - start);  // rows of the sequence (for the scale)
  if (selecting && seg_len > 0) {  // LAST / FIRST read exactly one row
    if (pooltype == CTRFE_POOL_FIRST) end = start + 1;
    else if (pooltype == CTRFE_POOL_LAST) start = end - 1;
  }
  const int len = static_cast<int>(end - start);
  const int maxlen = __reduce_max_sync(kFull, len);

  // this CTA's slot, copied out of the parameter bank once
  unsigned long long lane_base = reinterpret_cast<unsigned long long>(slp->table) + static_cast<unsigned>(col0 * ES);
  unsigned table_rows = slp->table_rows;
  unsigned row_bytes = kp.row_bytes;
  int id_kind = slp->id_kind;
  // address of this lane's first id (dense rows: the position itself)
  unsigned long long idp = id_kind == 2 ? static_cast<unsigned long long>(start + gl)
                                        : reinterpret_cast<unsigned long long>(slp->ids) +
                                              static_cast<unsigned long long>(start + gl) * (id_kind == 1 ? 4u : 8u);
  const unsigned id_step = id_kind == 2 ? 1u : (id_kind == 1 ? 4u : 8u);
  const bool pad_on = kp.padding_idx != -1;
  const uint32_t pad_lo = static_cast<uint32_t>(kp.padding_idx);
  const uint32_t pad_hi = static_cast<uint32_t>(static_cast<unsigned long long>(kp.padding_idx) >> 32);
  CTRFE_KEEP64(lane_base);
  const unsigned lane_off = static_cast<unsigned>(col0 * ES);
  const unsigned shard_mask = SHARDED ? static_cast<unsigned>(kp.n_shards - 1) : 0u;
  const int shard_shift = SHARDED ? kp.shard_shift : 0;
  // address of this lane's chunk of table row `rid`: local table, or -- sharded -- the owner
  // shard's memory (rid % n_shards), reached with ordinary loads through the NVLink mapping
  auto row_ptr = [&](uint32_t rid) -> const char* {
    if constexpr (SHARDED) {
      const unsigned long long base = s_shards[rid & shard_mask];
      return reinterpret_cast<const char*>(base + static_cast<unsigned long long>(rid >> shard_shift) * row_bytes + lane_off);
    } else {
      return reinterpret_cast<const char*>(lane_base + static_cast<unsigned long long>(rid) * row_bytes);
    }
  };
  CTRFE_KEEP32(table_rows);
  CTRFE_KEEP32(row_bytes);
  CTRFE_KEEP32(id_kind);

  // Row index of the id at address `a` (sequence position `off`): kSkip past the end; kZero
  // for a row that reads as zeros -- padding_idx (tested before the range check, as
  // embedding_kernel.cc:56-57 does) or an out-of-range id, which is also flagged for the
  // host to raise so that the kernel itself never faults.
  auto fetch_rid = [&](unsigned long long a, int off) -> uint32_t {
    if (off >= len) return kSkip;
    if (id_kind == 3) {
      // a feasign key: key 0 reads as zeros ("0 key fill zero", heter_ps/feature_value.cu:76-80,
      // feature_value.h:498-514); so does a key the map does not hold (the reference leaves the
      // value untouched, hashtable_kernel.cu:79-82)
      const unsigned long long key = __ldg(reinterpret_cast<const unsigned long long*>(a));
      if (key == 0ull) return kZero;
      unsigned h = static_cast<unsigned>((key * kKeyHashMul) >> kp.km_shift);
#pragma unroll 1
      for (int pr = 0; pr < kKeyProbes; ++pr) {
        const unsigned long long k = __ldg(kp.km_keys + h);
        if (k == key) {
          const unsigned r = __ldg(kp.km_rows + h);
          return r < table_rows ? r : kZero;
        }
        if (k == 0ull) break;
        h = (h + 1u) & kp.km_mask;
      }
      return kZero;
    }
    long long id;
    if (id_kind == 0) id = __ldg(reinterpret_cast<const long long*>(a));
    else if (id_kind == 1) id = static_cast<long long>(__ldg(reinterpret_cast<const int*>(a)));
    else id = static_cast<long long>(a);
    const uint32_t lo = static_cast<uint32_t>(id), hi = static_cast<uint32_t>(static_cast<unsigned long long>(id) >> 32);
    const bool in_range = hi == 0u && lo < table_rows;  // table_rows < 2^32 - 2
    const bool is_pad = pad_on && lo == pad_lo && hi == pad_hi;
    if (!in_range && !is_pad && tile == 0) {
      atomicMin(kp.err, err_key(kp, slp, static_cast<unsigned long long>(start + off)));
    }
    return (in_range && !is_pad) ? lo : kZero;
  };

  float acc[VEC];
  uint32_t sel[RW];  // selection pools on 16-bit types: raw bits of the selected elements
  {
    const float init = kp.init_mode == CTRFE_INIT_PAD ? kp.pad_value : 0.0f;
#pragma unroll
    for (int e = 0; e < VEC; ++e) acc[e] = init;
#pragma unroll
    for (int r = 0; r < RW; ++r) sel[r] = 0u;
  }
  // SELECT copies the first row; ADD does so only in EmbSeqPool order (CTRFE_INIT_FIRST_ROW)
  const bool copy_first = selecting || kp.init_mode == CTRFE_INIT_FIRST_ROW;

  uint32_t cur[NI];
#pragma unroll
  for (int i = 0; i < NI; ++i) cur[i] = fetch_rid(idp + static_cast<unsigned long long>(i * G) * id_step, i * G + gl);
  for (int base = 0; base < maxlen; base += R) {
    idp += static_cast<unsigned long long>(R) * id_step;
    uint32_t nxt[NI];  // next step's ids are requested before this step's rows
#pragma unroll
    for (int i = 0; i < NI; ++i) nxt[i] = kSkip;
    if (base + R < maxlen) {  // warp-uniform: only long sequences come back for more ids
#pragma unroll
      for (int i = 0; i < NI; ++i)
        nxt[i] = fetch_rid(idp + static_cast<unsigned long long>(i * G) * id_step, base + R + i * G + gl);
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
        ok[u] = (rid != kSkip) && col_ok;
        // a kZero row (padding_idx / bad id) is added as zeros without touching memory
#pragma unroll
        for (int r = 0; r < RW; ++r) raw[u][r] = 0u;
        if (rid < kZero && col_ok)
          L::load(row_ptr(rid), raw[u]);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (!ok[u]) continue;
        const bool first = copy_first && (c == 0 && u == 0) && (base == 0);
        if constexpr (MODE == kModeSelect) {
          // out = row0; then strictly-greater replaces (math/sequence_pooling.cc:166-174)
          if constexpr (DT == CTRFE_F16) {
            // packed half compare: same order as the fp32 compare (the conversion is exact and
            // monotone, half subnormals are normal floats); NaN never replaces; row 0 sticks
#pragma unroll
            for (int r = 0; r < RW; ++r) {
              const __half2 x2 = *reinterpret_cast<const __half2*>(&raw[u][r]);
              const __half2 a2 = *reinterpret_cast<const __half2*>(&sel[r]);
              const unsigned m = first ? 0xffffffffu : __hgt2_mask(x2, a2);
              sel[r] = (sel[r] & ~m) | (raw[u][r] & m);
            }
          } else if constexpr (ES == 4) {
#pragma unroll
            for (int e = 0; e < VEC; ++e) {
              const float x = __uint_as_float(raw[u][e]);
              if (first || x > acc[e]) acc[e] = x;
            }
          } else {
            uint32_t take[RW];
#pragma unroll
            for (int r = 0; r < RW; ++r) take[r] = 0u;
#pragma unroll
            for (int e = 0; e < VEC; ++e) {
              const float x = elem_to_float<DT>(L::bits(raw[u], e));
              if (first || x > acc[e]) {
                acc[e] = x;
                take[e >> 1] |= 0xffffu << ((e & 1) * 16);
              }
            }
#pragma unroll
            for (int r = 0; r < RW; ++r) sel[r] = (sel[r] & ~take[r]) | (raw[u][r] & take[r]);
          }
        } else {
          if constexpr (ES == 4 && VECT) {
            if (first) {
#pragma unroll
              for (int e = 0; e < VEC; ++e) acc[e] = __uint_as_float(raw[u][e]);
            } else {
#pragma unroll
              for (int e = 0; e < VEC; e += 2)  // two independent fp32 adds per instruction (FADD2)
                add_f32x2(acc[e], acc[e + 1], __uint_as_float(raw[u][e]), __uint_as_float(raw[u][e + 1]));
            }
          } else if constexpr (ES == 4) {
            const float x = __uint_as_float(raw[u][0]);
            acc[0] = first ? x : acc[0] + x;
          } else if constexpr (VECT) {
#pragma unroll
            for (int r = 0; r < RW; ++r) {
              const float2 x = L::unpack2(raw[u][r]);
              if (first) { acc[2 * r] = x.x; acc[2 * r + 1] = x.y; }
              else add_f32x2(acc[2 * r], acc[2 * r + 1], x.x, x.y);
            }
          } else {
            const float x = elem_to_float<DT>(raw[u][0]);
            acc[0] = first ? x : acc[0] + x;
          }
        }
      }
    }
#pragma unroll
    for (int i = 0; i < NI; ++i) cur[i] = nxt[i];
  }

  // ---- epilogue: pad / scale / CVM / store -------------------------------------------
  bool use_sel = selecting && ES == 2;  // 16-bit selection pools store the selected bits
  if (selecting && DT == CTRFE_F16 && kp.cvm_mode == CTRFE_CVM_USE && tile == 0 && gl == 0) {
    // the half-compare path kept no float copy: the CVM columns need one
    if constexpr (VECT && DT == CTRFE_F16) {
      const float2 x = L::unpack2(sel[0]);
      acc[0] = x.x;
      acc[1] = x.y;
    }
  }
  if constexpr (!VECT && DT == CTRFE_F16) {
    if (selecting) acc[0] = elem_to_float<DT>(sel[0]);
  }
  if (seg_len == 0 && kp.init_mode != CTRFE_INIT_PAD) {
    // empty sequence -> pad_value for every pooltype (sequence_pooling.cc:93-98,160-165,395-398)
#pragma unroll
    for (int e = 0; e < VEC; ++e) acc[e] = kp.pad_value;
    use_sel = false;
  } else if (seg_len > 0 && (pooltype == CTRFE_POOL_AVERAGE || pooltype == CTRFE_POOL_SQRT)) {
    const float h = static_cast<float>(seg_len);
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

  bool cvm_cols[2] = {false, false};  // which of this lane's leading elements were CVM-transformed
  if (kp.cvm_mode == CTRFE_CVM_USE) {
    // y0 = log(x0 + 1); y1 = log(x1 + 1) - y0   (cvm_op.h:34-37): the add in fp32, the log in
    // fp64 narrowed to fp32, the subtraction against the ROUNDED y0.  The two logs run on two
    // lanes at once (lane 1 borrows x1), halving the fp64 instruction stream.
    const bool head = tile == 0;
    if constexpr (VEC >= 2 && G >= 2) {
      const float x1 = __shfl_sync(kFull, acc[1], 0, G);
      double lg = 0.0;
      if (head && gl < 2) lg = cvm_log1(gl == 0 ? acc[0] : x1);
      const float y0 = __shfl_sync(kFull, static_cast<float>(lg), 0, G);
      float y1 = 0.0f;
      if (head && gl == 1) y1 = static_cast<float>(lg - static_cast<double>(y0));
      y1 = __shfl_sync(kFull, y1, 1, G);
      if (head && gl == 0) { acc[0] = y0; acc[1] = y1; cvm_cols[0] = cvm_cols[1] = true; }
    } else if constexpr (VEC >= 2) {
      if (head) {
        const float y0 = static_cast<float>(cvm_log1(acc[0]));
        acc[1] = static_cast<float>(cvm_log1(acc[1]) - static_cast<double>(y0));
        acc[0] = y0;
        cvm_cols[0] = cvm_cols[1] = true;
      }
    } else {
      double lg = 0.0;
      if (head && gl < 2) lg = cvm_log1(acc[0]);
      const float y0 = __shfl_sync(kFull, static_cast<float>(lg), 0, G);  // column 1 lives on lane 1
      if (head && gl == 0) { acc[0] = y0; cvm_cols[0] = true; }
      if (head && gl == 1) { acc[0] = static_cast<float>(lg - static_cast<double>(y0)); cvm_cols[0] = true; }
    }
  }

  if (!(seg_ok && col_ok)) return;
  const int skip = kp.cvm_mode == CTRFE_CVM_STRIP ? kp.cvm_offset : 0;
  char* orow = slp->out + static_cast<unsigned long long>(seg) * slp->out_ld_bytes;

  uint32_t packed[RW];
#pragma unroll
  for (int r = 0; r < RW; ++r) packed[r] = 0u;
#pragma unroll
  for (int e = 0; e < VEC; ++e) {
    uint32_t b;
    if (use_sel && !(e < 2 && cvm_cols[e & 1])) b = L::bits(sel, e);
    else b = float_to_elem<DT>(acc[e]);
    if constexpr (ES == 4) packed[e] = b;
    else packed[e >> 1] |= b << ((e & 1) * 16);
  }

  if constexpr (VECT) {
    char* dst = orow + static_cast<long long>(col0 - skip) * ES;
    if (skip == 0 && (reinterpret_cast<uintptr_t>(dst) & 15u) == 0) {
#pragma unroll
      for (int q = 0; q < RW / 4; ++q)
        reinterpret_cast<uint4*>(dst)[q] = make_uint4(packed[4 * q], packed[4 * q + 1], packed[4 * q + 2], packed[4 * q + 3]);
    } else {
#pragma unroll
      for (int e = 0; e < VEC; ++e) {
        if (col0 + e < skip) continue;
        const uint32_t b = L::bits(packed, e);
        if constexpr (ES == 4) reinterpret_cast<uint32_t*>(dst)[e] = b;
        else reinterpret_cast<unsigned short*>(dst)[e] = static_cast<unsigned short>(b);
      }
    }
  } else {
    if (col0 >= skip) {
      char* dst = orow + static_cast<long long>(col0 - skip) * ES;
      if constexpr (ES == 4) *reinterpret_cast<uint32_t*>(dst) = packed[0];
      else *reinterpret_cast<unsigned short*>(dst) = static_cast<unsigned short>(packed[0]);
    }
  }
}

// ---------------------------------------------------------------------------------------
// MaxIndex for sequence_pool(MAX, is_test=false): the row each column's maximum came from
// (MaxSeqPoolFunctor<T,false>, math/sequence_pooling.cc:93-111).  Training-side artefact,
// kept for interface completeness: one thread per output element.
// ---------------------------------------------------------------------------------------
template <int DT>
__global__ void max_index_kernel(const char* __restrict__ x, const unsigned* __restrict__ lod,
                                 long long n_seg, int width, int* __restrict__ index) {
  constexpr int ES = (DT == CTRFE_F32) ? 4 : 2;
  const long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (t >= n_seg * width) return;
  const long long seg = t / width;
  const int k = static_cast<int>(t - seg * width);
  const long long s = static_cast<long long>(lod[seg]), e = static_cast<long long>(lod[seg + 1]);
  if (s == e) { index[t] = -1; return; }
  auto at = [&](long long r) -> float {
    const char* p = x + (r * width + k) * ES;
    if constexpr (ES == 4) return __uint_as_float(__ldg(reinterpret_cast<const uint32_t*>(p)));
    else return elem_to_float<DT>(__ldg(reinterpret_cast<const unsigned short*>(p)));
  };
  float best = at(s);
  int bi = static_cast<int>(s);
  for (long long r = s + 1; r < e; ++r) {
    const float v = at(r);
    if (v > best) { best = v; bi = static_cast<int>(r); }
  }
  index[t] = bi;
}

// ---------------------------------------------------------------------------------------
// lookup_table_v2 on its own: out[k,:] = table[ids[k],:] (zero row for padding_idx).
// Byte copy in units of UB bytes, grid-stride, kUnroll units in flight per thread.
// ---------------------------------------------------------------------------------------
template <int UB> struct UnitT;
template <> struct UnitT<16> { using type = uint4; };
template <> struct UnitT<8> { using type = uint2; };
template <> struct UnitT<4> { using type = uint32_t; };
template <> struct UnitT<2> { using type = unsigned short; };
template <> struct UnitT<1> { using type = unsigned char; };

template <int UB>
__global__ void __launch_bounds__(kThreads) gather_rows_kernel(
    const char* __restrict__ table, long long table_rows, long long row_bytes, const void* __restrict__ ids,
    int ids_i32, long long n_ids, long long padding_idx, char* __restrict__ out, unsigned long long* err) {
  using T = typename UnitT<UB>::type;
  const long long upr = row_bytes / UB;  // units per row
  const long long total = n_ids * upr;
  const long long stride = static_cast<long long>(gridDim.x) * kThreads;
  for (long long i0 = static_cast<long long>(blockIdx.x) * kThreads + threadIdx.x; i0 < total;
       i0 += stride * kUnroll) {
    T v[kUnroll];
    long long dst[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const long long i = i0 + u * stride;
      dst[u] = -1;
      if (i >= total) continue;
      const long long k = i / upr, c = i - k * upr;
      // ids_i32: the CTRFE_I64 / I32 / I16 code of the id width
      const long long id = ids_i32 == 1 ? static_cast<long long>(__ldg(reinterpret_cast<const int*>(ids) + k))
                           : ids_i32 == 2 ? static_cast<long long>(__ldg(reinterpret_cast<const short*>(ids) + k))
                                          : __ldg(reinterpret_cast<const long long*>(ids) + k);
      dst[u] = i;
      v[u] = T{};
      if (padding_idx != -1 && id == padding_idx) continue;
      if (static_cast<unsigned long long>(id) >= static_cast<unsigned long long>(table_rows)) {
        if (c == 0) atomicMin(err, static_cast<unsigned long long>(k));
        continue;
      }
      v[u] = __ldg(reinterpret_cast<const T*>(table + id * row_bytes) + c);
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u)
      if (dst[u] >= 0) reinterpret_cast<T*>(out)[dst[u]] = v[u];
  }
}

// ids of any supported width -> int64 rows (the rows vector of a SelectedRows W@GRAD,
// phi/kernels/gpu/embedding_grad_kernel.cu:168-186 InputTypeConvert / Copy).
__global__ void ids_widen_kernel(const void* __restrict__ ids, int kind, long long n, long long* __restrict__ out) {
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x)
    out[i] = kind == 1 ? static_cast<long long>(reinterpret_cast<const int*>(ids)[i])
             : kind == 2 ? static_cast<long long>(reinterpret_cast<const short*>(ids)[i])
                         : reinterpret_cast<const long long*>(ids)[i];
}

// ---------------------------------------------------------------------------------------
// cvm on its own (cvm_op.h:25-41 semantics; one thread per output element).
// ---------------------------------------------------------------------------------------
template <int DT>
__global__ void cvm_kernel(const char* __restrict__ x, long long rows, int width, int use_cvm,
                           char* __restrict__ y) {
  constexpr int ES = (DT == CTRFE_F32) ? 4 : 2;
  const int wo = use_cvm ? width : width - 2;
  const long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (t >= rows * wo) return;
  const long long r = t / wo;
  const int c = static_cast<int>(t - r * wo);
  auto ld = [&](int col) -> uint32_t {
    const char* p = x + (r * width + col) * ES;
    if constexpr (ES == 4) return __ldg(reinterpret_cast<const uint32_t*>(p));
    else return __ldg(reinterpret_cast<const unsigned short*>(p));
  };
  uint32_t b;
  if (!use_cvm) {
    b = ld(c + 2);
  } else if (c >= 2) {
    b = ld(c);
  } else {
    const float x0 = elem_to_float<DT>(ld(0));
    const float y0 = static_cast<float>(cvm_log1(x0));
    if (c == 0) {
      b = float_to_elem<DT>(y0);
    } else {
      const float x1 = elem_to_float<DT>(ld(1));
      // the reference subtracts the STORED y0 (cvm_op.h:36): round it to T first
      const float y0s = elem_to_float<DT>(float_to_elem<DT>(y0));
      b = float_to_elem<DT>(static_cast<float>(cvm_log1(x1) - static_cast<double>(y0s)));
    }
  }
  char* q = y + t * ES;
  if constexpr (ES == 4) *reinterpret_cast<uint32_t*>(q) = b;
  else *reinterpret_cast<unsigned short*>(q) = static_cast<unsigned short>(b);
}

// cvm on double rows (the reference registers CVMCUDAKernel<double>, cvm_op.cu:192-194): everything
// in fp64 with the library log, as CvmComputeKernel<double> does (cvm_op.cu:28-47).
__global__ void cvm_f64_kernel(const double* __restrict__ x, long long rows, int width, int use_cvm, double* __restrict__ y) {
  const int wo = use_cvm ? width : width - 2;
  const long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (t >= rows * wo) return;
  const long long r = t / wo;
  const int c = static_cast<int>(t - r * wo);
  const double* xr = x + r * width;
  double v;
  if (!use_cvm) v = xr[c + 2];
  else if (c >= 2) v = xr[c];
  else if (c == 0) v = log(xr[0] + 1.0);
  else v = log(xr[1] + 1.0) - log(xr[0] + 1.0);
  y[t] = v;
}

// ---------------------------------------------------------------------------------------
// LoD offsets of a plan travel to the device as 16-bit sequence LENGTHS when every sequence is shorter
// than 65536 (a quarter of the 64-bit offsets Paddle holds, half of a 32-bit image): per slot the
// 32-bit offset at which every chunk of kExpandChunk sequences starts (the host has them anyway), then
// n_seg lengths.  This kernel turns them back into the 32-bit offsets the pooling kernels read: one CTA
// per chunk, eight lengths per thread, one block-wide scan, no dependence between CTAs.
// ---------------------------------------------------------------------------------------
constexpr unsigned kExpandChunk = 2048;  // sequences per CTA of lod_expand_kernel (256 threads x 8)
struct ExpandSlot {
  unsigned long long dst;  // device uint32 [n_seg + 1]
  unsigned base_off;       // byte offset of the slot's chunk bases (uint32 per chunk) in the compact image
  unsigned len_off;        // byte offset of the slot's lengths (uint16 per sequence, 16-byte aligned)
  unsigned n_seg;
  unsigned first_block;    // blockIdx.x of the slot's first chunk
};
struct ExpandParams {
  const char* src;
  int n;           // slots
  int n_blocks;    // chunks of all slots
  ExpandSlot s[kMaxSlotsPerLaunch];
};

__global__ void __launch_bounds__(256) lod_expand_kernel(const __grid_constant__ ExpandParams p) {
  int j = 0;
  while (j + 1 < p.n && p.s[j + 1].first_block <= blockIdx.x) ++j;
  const ExpandSlot& sl = p.s[j];
  const unsigned n_seg = sl.n_seg;
  const unsigned chunk = blockIdx.x - sl.first_block;
  const unsigned c0 = chunk * kExpandChunk;
  const unsigned base = __ldg(reinterpret_cast<const unsigned*>(p.src + sl.base_off) + chunk);
  const uint4* len8 = reinterpret_cast<const uint4*>(p.src + sl.len_off);
  unsigned* dst = reinterpret_cast<unsigned*>(sl.dst);
  __shared__ unsigned s_warp[8];
  const unsigned tid = threadIdx.x, lane = tid & 31u, w = tid >> 5;
  if (chunk == 0 && tid == 0) dst[0] = base;
  const unsigned i0 = c0 + tid * 8u;
  uint4 v = make_uint4(0u, 0u, 0u, 0u);
  if (i0 < n_seg) v = __ldg(len8 + (i0 >> 3));
  unsigned l[8] = {v.x & 0xffffu, v.x >> 16, v.y & 0xffffu, v.y >> 16, v.z & 0xffffu, v.z >> 16, v.w & 0xffffu, v.w >> 16};
  unsigned run = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    if (i0 + k >= n_seg) l[k] = 0u;
    run += l[k];
    l[k] = run;  // inclusive prefix inside the thread
  }
  unsigned incl = run;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const unsigned t = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= static_cast<unsigned>(d)) incl += t;
  }
  if (lane == 31u) s_warp[w] = incl;
  __syncthreads();
  unsigned before = base;
  for (unsigned q = 0; q < w; ++q) before += s_warp[q];
  const unsigned off = before + incl - run;  // offset of this thread's first sequence
#pragma unroll
  for (int k = 0; k < 8; ++k)
    if (i0 + k < n_seg) dst[i0 + k + 1] = off + l[k];
}

// ids [K, W] row-major -> W contiguous arrays of K ids (column w at out + w * K): lets an Ids tensor
// whose last dimension is wider than 1 (fused_embedding_seq_pool, refer.h:477-528 index_width) feed
// the pooling kernels, which want one contiguous id run per slot.
__global__ void ids_transpose_kernel(const long long* __restrict__ ids, long long k, int w, long long* __restrict__ out) {
  const long long t = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (t >= k * w) return;
  const long long r = t / w;
  const int c = static_cast<int>(t - r * w);
  out[static_cast<long long>(c) * k + r] = ids[t];
}

// Fills a feasign key map: keys[i] -> row i.  Open addressing, linear probing, 0 = empty slot
// (key 0 is the padding feasign and is never stored).  A key that finds no slot within
// kKeyProbes is counted in *n_unplaced (the caller rebuilds with more slots).
__global__ void keymap_insert_kernel(const unsigned long long* __restrict__ keys, long long n, unsigned long long* map_keys,
                                     unsigned* map_rows, unsigned mask, int shift, unsigned* n_unplaced) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const unsigned long long key = keys[i];
  if (key == 0ull) return;
  unsigned h = static_cast<unsigned>((key * kKeyHashMul) >> shift);
  for (int pr = 0; pr < kKeyProbes; ++pr) {
    const unsigned long long old = atomicCAS(map_keys + h, 0ull, key);
    if (old == 0ull || old == key) {
      atomicMin(map_rows + h, static_cast<unsigned>(i));  // duplicate keys: the first occurrence wins, whatever the schedule
      return;
    }
    h = (h + 1u) & mask;
  }
  if (n_unplaced) atomicAdd(n_unplaced, 1u);
}

}  // namespace ctrfe
