// ctrfe_api.cu -- the C ABI declared in include/ctrfe.h: argument checking (mirroring the
// reference's enforce checks), one-copy LoD upload, kernel launches.
// No CPU compute path exists in this file: every entry point launches sm_100a kernels.
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <string>
#include <utility>
#include <vector>

#include "ctrfe_fast_kernel.cuh"
#include "ctrfe_grad_kernels.cuh"
#include "ctrfe_kernels.cuh"

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

#define CU(call)                                                                              \
  do {                                                                                        \
    cudaError_t e__ = (call);                                                                 \
    if (e__ != cudaSuccess)                                                                   \
      return fail(CTRFE_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__),     \
                  __FILE__, __LINE__);                                                        \
  } while (0)

constexpr int kRing = 4;  // pinned staging buffers in flight

struct Staging {
  char* host = nullptr;
  size_t cap = 0;
  cudaEvent_t done = nullptr;
  bool pending = false;
};

struct IdSource {  // what ctrfe_check_ids needs to fetch the offending id value
  const void* ids;
  int ids_i32;  // the id width as CTRFE_I64 (0) / CTRFE_I32 (1) / CTRFE_I16 (2)
};

}  // namespace

struct ctrfe_ctx {
  int device = 0;
  unsigned long long* err_dev = nullptr;   // [1]
  unsigned long long* err_host = nullptr;  // pinned [2]
  Staging ring[kRing];
  int ring_next = 0;
  // LoD image of the one-shot ctrfe_fused_forward / ctrfe_sequence_pool paths
  char* scratch = nullptr;
  size_t scratch_cap = 0;
  char* id_scratch = nullptr;  // transposed ids of ctrfe_fused_embedding_seq_pool (index_width > 1)
  size_t id_scratch_cap = 0;
  // Every plan and every one-shot call gets a tag (1..255, reused cyclically) that its launches OR
  // into the error keys they record, so ctrfe_check_ids can fetch the offending id from the batch
  // that met it even with several plans in flight.
  std::vector<IdSource> tag_ids[256];
  unsigned next_tag = 1;
  cudaStream_t copy_stream = nullptr;  // LoD re-uploads of CTRFE_LAUNCH_UPLOAD_LOD
  int sm_count = 148;
};

struct ctrfe_plan {
  ctrfe_ctx* ctx = nullptr;
  char* dev = nullptr;  // LoD image when the offsets came from the host
  bool owns_dev = false;
  std::vector<ctrfe::KParams> chunks;  // <= kMaxSlotsPerLaunch slots each: one launch per chunk
  std::vector<dim3> grids;
  std::vector<int> modes;                      // kernel MODE of each chunk
  std::vector<std::pair<int, int>> where;      // caller slot -> (chunk, index in chunk)
  int dtype = 0;
  bool vect = false;
  int g_shift = 0;
  int chunk_bytes = 16;  // bytes of a row one lane holds (32: 256-bit loads, fp32 only)
  bool fast = false;     // accumulating chunks may use the specialised kernel (ctrfe_fast_kernel.cuh)
  bool cvm_use = false;
  bool hot = false;      // sharded mode with a hot-row replica cache
  std::vector<char> all_max;  // per launch: every slot pools with MAX (selecting launches only)
  std::vector<IdSource> id_src;
  unsigned tag = 0;
  // host-LoD plans: the pinned image the offsets were uploaded from (re-filled by ctrfe_plan_update_lod,
  // re-uploaded by CTRFE_LAUNCH_UPLOAD_LOD) and what is needed to validate new offsets
  char* host_img = nullptr;
  size_t img_bytes = 0;      // bytes of the 32-bit offsets image (device `dev`; also the pinned buffer's capacity)
  // the compact form of the image: 16-bit sequence lengths (+ a 16-byte header per slot), expanded on the
  // device by lod_expand_kernel.  `compact` says which form the pinned image currently holds.
  char* dev_c = nullptr;
  size_t c_bytes = 0;
  bool compact = false;
  std::vector<ctrfe::ExpandParams> expand;
  struct SlotShape { int64_t n_seg; int64_t bound; bool has_ids; };
  std::vector<SlotShape> shapes;
  cudaEvent_t lod_ready = nullptr, done = nullptr;
  std::vector<std::string> kernel_names;  // per launch, for ctrfe_plan_kernel_name
};

namespace {

struct DeviceGuard {
  int prev = -1;
  bool ok = true;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
    if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
  }
  ~DeviceGuard() {
    int cur = -1;
    if (prev >= 0 && cudaGetDevice(&cur) == cudaSuccess && cur != prev) cudaSetDevice(prev);
  }
};

int elem_size(int dtype) {
  switch (dtype) {
    case CTRFE_F32: return 4;
    case CTRFE_F16: case CTRFE_BF16: return 2;
    case CTRFE_F64: return 8;
    default: return 0;
  }
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

unsigned take_tag(ctrfe_ctx* ctx) {
  const unsigned t = ctx->next_tag;
  ctx->next_tag = t % 255u + 1u;
  return t;
}

int stage_acquire(ctrfe_ctx* ctx, size_t bytes, Staging** out) {
  Staging& s = ctx->ring[ctx->ring_next];
  ctx->ring_next = (ctx->ring_next + 1) % kRing;
  if (s.pending) {
    CU(cudaEventSynchronize(s.done));  // normally long complete
    s.pending = false;
  }
  if (s.cap < bytes) {
    if (s.host) CU(cudaFreeHost(s.host));
    s.host = nullptr;
    size_t cap = bytes < (1u << 16) ? (1u << 16) : bytes + bytes / 4;
    CU(cudaHostAlloc(reinterpret_cast<void**>(&s.host), cap, cudaHostAllocDefault));
    s.cap = cap;
  }
  if (!s.done) CU(cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming));
  *out = &s;
  return CTRFE_OK;
}

int scratch_reserve(ctrfe_ctx* ctx, size_t bytes, cudaStream_t stream) {
  if (ctx->scratch_cap >= bytes) return CTRFE_OK;
  if (ctx->scratch) {
    CU(cudaStreamSynchronize(stream));  // work already enqueued may still read the old buffer
    CU(cudaFree(ctx->scratch));
    ctx->scratch = nullptr;
  }
  size_t cap = bytes < (1u << 20) ? (1u << 20) : bytes + bytes / 4;
  CU(cudaMalloc(reinterpret_cast<void**>(&ctx->scratch), cap));
  ctx->scratch_cap = cap;
  return CTRFE_OK;
}

// Validates a descriptor the way the reference ops' InferShape/Compute do, derives the launch
// geometry and fills everything of the plan that does not depend on where the LoD lives.
// *lod_bytes = device bytes needed for the host LoDs (0 when they are already on the device).
int plan_geometry(ctrfe_ctx* ctx, const ctrfe_fused_desc* d, ctrfe_plan* plan, size_t* lod_bytes) {
  if (!d || !d->slots || d->n_slots <= 0) return fail(CTRFE_ERR_ARG, "fused: no slots");
  if (d->n_slots > 4096) return fail(CTRFE_ERR_ARG, "fused: %d slots exceed the limit of 4096 per call", d->n_slots);
  if (d->dtype != CTRFE_F32 && d->dtype != CTRFE_F16 && d->dtype != CTRFE_BF16)
    return fail(CTRFE_ERR_UNSUPPORTED, "fused: dtype %d not supported (f32, f16, bf16)", d->dtype);
  if (d->dim <= 0 || d->dim > (1 << 20)) return fail(CTRFE_ERR_ARG, "fused: bad dim %lld", (long long)d->dim);
  if (d->cvm_mode < CTRFE_CVM_NONE || d->cvm_mode > CTRFE_CVM_STRIP)
    return fail(CTRFE_ERR_ARG, "fused: bad cvm_mode %d", d->cvm_mode);
  if (d->cvm_mode != CTRFE_CVM_NONE && d->dim < 2)
    return fail(CTRFE_ERR_ARG, "cvm needs at least the show and click columns (dim=%lld)", (long long)d->dim);
  if (d->cvm_mode == CTRFE_CVM_STRIP && (d->cvm_offset < 0 || d->cvm_offset >= d->dim))
    return fail(CTRFE_ERR_ARG, "fused: cvm_offset %d outside [0, dim)", d->cvm_offset);
  if (d->scale_mode != CTRFE_SCALE_DIVIDE && d->scale_mode != CTRFE_SCALE_RECIPROCAL)
    return fail(CTRFE_ERR_ARG, "fused: bad scale_mode %d", d->scale_mode);
  if (d->init_mode < CTRFE_INIT_ZERO || d->init_mode > CTRFE_INIT_FIRST_ROW)
    return fail(CTRFE_ERR_ARG, "fused: bad init_mode %d", d->init_mode);
  if (d->lod_location != CTRFE_LOD_HOST && d->lod_location != CTRFE_LOD_DEVICE)
    return fail(CTRFE_ERR_ARG, "fused: bad lod_location %d", d->lod_location);

  const int es = elem_size(d->dtype);
  const long long row_bytes = d->dim * es;
  const int d_out = static_cast<int>(d->dim) - (d->cvm_mode == CTRFE_CVM_STRIP ? d->cvm_offset : 0);
  bool v = (row_bytes % 16) == 0;
  const bool sharded = d->n_shards != 0;
  if (sharded) {
    if (d->n_shards != 2 && d->n_shards != 4 && d->n_shards != 8)
      return fail(CTRFE_ERR_ARG, "fused: n_shards must be 2, 4 or 8 (got %d)", d->n_shards);
    if (!d->shard_tables) return fail(CTRFE_ERR_ARG, "fused: n_shards set but shard_tables is null");
    if (!v) return fail(CTRFE_ERR_UNSUPPORTED, "fused: the sharded table mode needs rows that are a multiple of 16 bytes");
    for (int g = 0; g < d->n_shards; ++g)
      if (!d->shard_tables[g] || reinterpret_cast<uintptr_t>(d->shard_tables[g]) % 16 != 0)
        return fail(CTRFE_ERR_ARG, "fused: shard_tables[%d] is null or not 16-byte aligned", g);
  }
  if (d->hot_cache && d->hot_cache->rows) {
    // replicas of the hottest rows: beside a sharded table they save NVLink trips, beside ONE big
    // unsharded table they are a compact region the TLB and L2 can hold
    const ctrfe_hot_cache* hc = d->hot_cache;
    if (!hc->keys || hc->log2_slots < 4 || hc->log2_slots > 30 || hc->n_rows <= 0 ||
        reinterpret_cast<uintptr_t>(hc->rows) % 16 != 0)
      return fail(CTRFE_ERR_ARG, "fused: malformed hot_cache");
    for (int s = 0; s < d->n_slots; ++s) {
      if (d->slots[s].table_rows >= 0x7fffffffLL)
        return fail(CTRFE_ERR_UNSUPPORTED, "fused: a hot-row cache needs table_rows < 2^31");
      if (!sharded && (d->slots[s].table != d->slots[0].table || d->slots[s].table_rows != d->slots[0].table_rows))
        return fail(CTRFE_ERR_ARG, "fused: a hot-row cache beside unsharded tables needs every slot to read the same table");
    }
  }
  const ctrfe_keymap* km = d->keymap;
  if (km) {
    if (sharded) return fail(CTRFE_ERR_UNSUPPORTED, "fused: a key map cannot be combined with a sharded table");
    if (!km->keys || !km->rows || km->log2_slots < 4 || km->log2_slots > 31)
      return fail(CTRFE_ERR_ARG, "fused: malformed keymap");
    for (int s = 0; s < d->n_slots; ++s)
      if (!d->slots[s].ids || d->slots[s].ids_dtype != CTRFE_I64)
        return fail(CTRFE_ERR_ARG, "slot %d: with a key map the ids are uint64 feasign keys (ids_dtype CTRFE_I64)", s);
  }
  size_t lb = 0;
  for (int s = 0; s < d->n_slots; ++s) {
    const ctrfe_slot& sl = d->slots[s];
    if (sharded && !sl.ids) return fail(CTRFE_ERR_ARG, "slot %d: the sharded table mode needs ids", s);
    if (sl.n_seg < 0 || sl.n_seg > 0x7fffffffLL) return fail(CTRFE_ERR_ARG, "slot %d: n_seg %lld outside [0, 2^31)", s, (long long)sl.n_seg);
    if (sl.n_seg > 0 && ((!sl.table && !sharded) || !sl.lod || !sl.out))
      return fail(CTRFE_ERR_ARG, "slot %d: null table / lod / out", s);
    if (sl.pooltype < CTRFE_POOL_AVERAGE || sl.pooltype > CTRFE_POOL_MAX)
      return fail(CTRFE_ERR_ARG,
                  "slot %d: unsupported pooling pooltype %d (AVERAGE, SUM, SQRT, LAST, FIRST, MAX)", s,
                  sl.pooltype);
    if (sl.ids_dtype != CTRFE_I64 && sl.ids_dtype != CTRFE_I32)
      return fail(CTRFE_ERR_UNSUPPORTED, "slot %d: ids must be int32 or int64", s);
    if (sl.out_ld < d_out) return fail(CTRFE_ERR_ARG, "slot %d: out_ld %lld < output width %d", s,
                                       (long long)sl.out_ld, d_out);
    if (sl.table_rows < 0 || sl.table_rows >= 0xfffffffeLL)
      return fail(CTRFE_ERR_UNSUPPORTED, "slot %d: table_rows %lld outside [0, 2^32 - 2)", s, (long long)sl.table_rows);
    if (sl.ids && (sl.n_ids < 0 || sl.n_ids > 0x7fffffffLL))
      return fail(CTRFE_ERR_ARG, "slot %d: n_ids %lld outside [0, 2^31) (a slot holds fewer than 2^31 ids)", s, (long long)sl.n_ids);
    if (sl.out_ld * es >= (1LL << 32)) return fail(CTRFE_ERR_ARG, "slot %d: out_ld %lld too large", s, (long long)sl.out_ld);
    if (!sharded && reinterpret_cast<uintptr_t>(sl.table) % 16 != 0) v = false;
    lb += align_up(static_cast<size_t>(sl.n_seg + 1) * 4, 16);  // the library's own image holds 32-bit offsets
  }
  // the specialised kernel: rows of exactly 16 << k bytes, zero-initialised accumulating pools, no
  // padding_idx, aligned outputs (CTRFE_NO_FAST=1 forces the general kernel, for A/B runs)
  bool has_select = false;
  {
    const char* nofast = getenv("CTRFE_NO_FAST");
    bool f = v && row_bytes <= 1024 && (row_bytes & (row_bytes - 1)) == 0 && d->padding_idx == -1 &&
             d->init_mode == CTRFE_INIT_ZERO && d->cvm_mode != CTRFE_CVM_STRIP && !km && !(nofast && nofast[0] == '1');
    for (int s = 0; s < d->n_slots; ++s) {
      const ctrfe_slot& sl = d->slots[s];
      if (sl.pooltype >= CTRFE_POOL_LAST) has_select = true;
      if (!sl.ids || sl.ids_dtype != CTRFE_I64 || reinterpret_cast<uintptr_t>(sl.out) % 16 != 0 || (sl.out_ld * es) % 16 != 0) f = false;
    }
    plan->fast = f;
    plan->cvm_use = d->cvm_mode == CTRFE_CVM_USE;
  }
  // Bytes of a row one lane holds: 16 (one 128-bit load) or 32 (one 256-bit load).  A stripped gather of
  // uniform-random rows sustains far more through 256-bit loads once rows are 128 bytes wide (5.3 against
  // 3.8 TB/s, 6.5 against 3.8 at 256 bytes: profiles/r2_gather_sweep.txt) -- but none of the five BASELINE
  // configs is bound there: cfg3 / cfg4 on one GPU hold > 100 GB of tables and run at the TLB's page-walk
  // rate (profiles/r2_gather_table_size.txt), cfg5 at the latency of its long sequences, and the wide form
  // costs a third of the resident warps (40 - 48 registers).  Measured (profiles/r2_wide_lane_sweep.txt):
  // cfg3 203 -> 206 us at best, cfg4 611 -> 617, cfg5 174 -> 195.  So 128-bit lanes stay the default;
  // CTRFE_LDG256=1 selects 256-bit lanes for rows of 128 bytes and wider, =2 for every row of 32 bytes
  // and up.  In the sharded mode only the specialised kernel has the wide form.
  int cb = 16;
  {
    const char* use256 = getenv("CTRFE_LDG256");
    const int wide = use256 && use256[0] ? atoi(use256) : 0;
    const long long min_row = wide >= 2 ? 32 : 128;
    const bool hot_unsharded = !sharded && d->hot_cache && d->hot_cache->rows;
    if (wide >= 1 && !hot_unsharded && v && row_bytes % 32 == 0 && row_bytes >= min_row &&
        (!sharded || (plan->fast && !has_select))) {
      cb = 32;
      if (sharded) {
        for (int g = 0; g < d->n_shards; ++g)
          if (reinterpret_cast<uintptr_t>(d->shard_tables[g]) % 32 != 0) cb = 16;
        if (d->hot_cache && d->hot_cache->rows && reinterpret_cast<uintptr_t>(d->hot_cache->rows) % 32 != 0) cb = 16;
      } else {
        for (int s = 0; s < d->n_slots; ++s)
          if (reinterpret_cast<uintptr_t>(d->slots[s].table) % 32 != 0) cb = 16;
        if (d->hot_cache && d->hot_cache->rows && reinterpret_cast<uintptr_t>(d->hot_cache->rows) % 32 != 0) cb = 16;
      }
    }
    if (cb == 16 && row_bytes > 512) plan->fast = false;  // 128-bit lanes cover at most 512 bytes with one warp
  }
  plan->chunk_bytes = cb;
  const long long units = v ? row_bytes / cb : d->dim;  // lane-sized pieces per row
  int g_shift = 5;                                      // the scalar fallback always uses whole warps
  if (v) { g_shift = 0; while ((1 << g_shift) < units && g_shift < 5) ++g_shift; }
  const int tiles = static_cast<int>((units + (1 << g_shift) - 1) >> g_shift);
  plan->ctx = ctx;
  plan->dtype = d->dtype;
  plan->vect = v;
  plan->g_shift = g_shift;
  plan->chunks.clear();
  plan->grids.clear();
  plan->modes.clear();
  plan->all_max.clear();
  plan->where.assign(d->n_slots, std::make_pair(-1, -1));
  // Slots are grouped by kernel mode (accumulating vs selecting pools) into launches of at
  // most kMaxSlotsPerLaunch descriptors; a CTR batch with one pooltype family is ONE launch.
  int open_chunk[2] = {-1, -1};
  std::vector<long long> max_seg;
  for (int s = 0; s < d->n_slots; ++s) {
    const ctrfe_slot& sl = d->slots[s];
    const int mode = sl.pooltype >= CTRFE_POOL_LAST ? ctrfe::kModeSelect : ctrfe::kModeAdd;
    int c = open_chunk[mode];
    if (c < 0 || plan->chunks[c].n_slots == ctrfe::kMaxSlotsPerLaunch) {
      plan->chunks.emplace_back();
      c = static_cast<int>(plan->chunks.size()) - 1;
      open_chunk[mode] = c;
      plan->modes.push_back(mode);
      plan->all_max.push_back(1);
      max_seg.push_back(0);
      ctrfe::KParams& k = plan->chunks[c];
      memset(&k, 0, sizeof(k));
      k.err = ctx->err_dev;
      k.err_tag = static_cast<unsigned long long>(plan->tag) << ctrfe::kErrTagShift;
      k.padding_idx = d->padding_idx;
      k.dim = static_cast<int>(d->dim);
      k.row_bytes = static_cast<int>(row_bytes);
      k.tiles = tiles;
      k.cvm_mode = d->cvm_mode;
      k.cvm_offset = d->cvm_offset;
      k.scale_mode = d->scale_mode;
      k.init_mode = d->init_mode;
      k.pad_value = d->pad_value;
      k.n_slots = 0;
      if (plan->fast && d->hot_cache && d->hot_cache->rows && d->hot_cache->log2_slots > 0) {
        k.hot_rows = static_cast<const char*>(d->hot_cache->rows);
        k.hot_keys = d->hot_cache->keys;
        k.hot_mask = (1u << d->hot_cache->log2_slots) - 1u;
        k.hot_shift = 32 - d->hot_cache->log2_slots;
      }
      if (km) {
        k.km_keys = reinterpret_cast<const unsigned long long*>(km->keys);
        k.km_rows = km->rows;
        k.km_mask = (1u << km->log2_slots) - 1u;
        k.km_shift = 64 - km->log2_slots;
      }
      if (sharded) {
        k.n_shards = d->n_shards;
        k.shard_shift = d->n_shards == 2 ? 1 : (d->n_shards == 4 ? 2 : 3);
        for (int g = 0; g < ctrfe::kMaxShards; ++g)
          k.shard_tables[g] = static_cast<const char*>(d->shard_tables[g % d->n_shards]);
      }
    }
    ctrfe::KParams& k = plan->chunks[c];
    ctrfe::SlotK& o = k.slots[k.n_slots];
    plan->where[s] = std::make_pair(c, k.n_slots);
    ++k.n_slots;
    o.table = static_cast<const char*>(sl.table);
    o.ids = sl.ids;
    o.lod = reinterpret_cast<const unsigned long long*>(sl.lod);  // re-pointed for host LoDs
    o.out = static_cast<char*>(sl.out);
    o.out_ld_bytes = static_cast<unsigned>(sl.out_ld * es);
    o.table_rows = static_cast<unsigned>(sl.table_rows);
    o.n_seg = static_cast<unsigned>(sl.n_seg);
    // what every LoD offset is checked against: the id count, or -- dense rows -- the rows of X
    o.n_ids = static_cast<unsigned>(sl.ids ? sl.n_ids : (sl.table_rows > 0x7fffffffLL ? 0x7fffffffLL : sl.table_rows));
    o.pooltype = static_cast<signed char>(sl.pooltype);
    o.lod32 = d->lod_location == CTRFE_LOD_HOST ? 1 : 0;
    o.id_kind = static_cast<short>(km ? 3 : (sl.ids ? (sl.ids_dtype == CTRFE_I32 ? 1 : 0) : 2));
    o.slot_index = s;
    if (sl.n_seg > max_seg[c]) max_seg[c] = sl.n_seg;
    if (sl.pooltype != CTRFE_POOL_MAX) plan->all_max[c] = 0;
  }
  for (size_t c = 0; c < plan->chunks.size(); ++c) {
    const long long items = max_seg[c] * tiles;
    const long long gx = ((items << g_shift) + ctrfe::kThreads - 1) / ctrfe::kThreads;
    if (gx > 0x7fffffffLL) return fail(CTRFE_ERR_ARG, "fused: batch too large for one launch");
    plan->grids.push_back(dim3(static_cast<unsigned>(gx > 0 ? gx : 1), static_cast<unsigned>(plan->chunks[c].n_slots), 1));
  }
  plan->shapes.resize(d->n_slots);
  for (int s = 0; s < d->n_slots; ++s)
    plan->shapes[s] = ctrfe_plan::SlotShape{d->slots[s].n_seg, d->slots[s].ids ? d->slots[s].n_ids : d->slots[s].table_rows,
                                            d->slots[s].ids != nullptr};
  plan->id_src.resize(d->n_slots);
  for (int s = 0; s < d->n_slots; ++s)
    plan->id_src[s] = IdSource{d->slots[s].ids, d->slots[s].ids_dtype == CTRFE_I32};
  // the kernel each launch of this plan runs (mirrors launch_plan), for profiles and bench records
  plan->kernel_names.clear();
  for (size_t c = 0; c < plan->chunks.size(); ++c) {
    static const char* const dtn[] = {"f32", "f16", "bf16"};
    char buf[160];
    const bool sh = plan->chunks[c].n_shards > 0;
    if (plan->fast && plan->modes[c] == ctrfe::kModeAdd)
      snprintf(buf, sizeof(buf), "ctrfe::fused_pool_fast_kernel<%s, %d-byte rows, cvm=%d, sharded=%d, hot=%d, %d-byte lanes>",
               dtn[plan->dtype], plan->chunk_bytes << plan->g_shift, plan->cvm_use ? 1 : 0, sh ? 1 : 0,
               plan->chunks[c].hot_rows ? 1 : 0, plan->chunk_bytes);
    else
      snprintf(buf, sizeof(buf), "ctrfe::fused_pool_kernel<%s, vect=%d, lanes=%d, %s, %d-byte lanes, sharded=%d>", dtn[plan->dtype],
               plan->vect ? 1 : 0, 1 << plan->g_shift, plan->modes[c] == ctrfe::kModeSelect ? "select" : "add",
               plan->chunk_bytes, sh ? 1 : 0);
    plan->kernel_names.emplace_back(buf);
  }
  *lod_bytes = d->lod_location == CTRFE_LOD_HOST ? lb : 0;
  return CTRFE_OK;
}

// One slot's host offsets -> 32-bit image, validated in the same pass: non-decreasing (framework
// CheckLoD, lod_tensor.h:74-104), below 2^31, last offset inside the ids (or, dense rows, inside X:
// sequence_pool_op.h:52-60).
// (the narrowing loop has no loop-carried dependency besides an OR, so the AVX2 clone runs it eight
// offsets at a time: a cfg2 batch has 320 K offsets and this pass sits on the host-buffer path)
__attribute__((target_clones("avx2", "default")))
uint64_t narrow_and_check(const uint64_t* lod, int64_t n, uint32_t* dst) {
  uint64_t bad = 0;
  dst[0] = static_cast<uint32_t>(lod[0]);
  for (int64_t i = 1; i < n; ++i) {
    bad |= static_cast<uint64_t>(lod[i] < lod[i - 1]);
    dst[i] = static_cast<uint32_t>(lod[i]);
  }
  return bad;
}

int stage_one_lod(int s, const uint64_t* lod, int64_t n_seg, int64_t bound, bool has_ids, uint32_t* dst) {
  if (n_seg == 0) {
    dst[0] = lod ? static_cast<uint32_t>(lod[0]) : 0u;
    return CTRFE_OK;
  }
  const bool mono = narrow_and_check(lod, n_seg + 1, dst) == 0;
  const uint64_t prev = lod[n_seg];
  if (!mono) return fail(CTRFE_ERR_ARG, "slot %d: LoD offsets must be non-decreasing", s);
  if (prev >= (1ull << 31))
    return fail(CTRFE_ERR_ARG, "slot %d: LoD offset %llu too large (a slot holds fewer than 2^31 ids)", s,
                (unsigned long long)prev);
  if (has_ids && static_cast<int64_t>(prev) > bound)
    return fail(CTRFE_ERR_ARG, "slot %d: the last LoD offset (%llu) exceeds the number of ids (%lld)", s,
                (unsigned long long)prev, (long long)bound);
  if (!has_ids && static_cast<int64_t>(prev) > bound)
    return fail(CTRFE_ERR_ARG, "slot %d: the first dimension of Input(X) (%lld) must be >= the last LoD offset (%llu)", s,
                (long long)bound, (unsigned long long)prev);
  return CTRFE_OK;
}

constexpr int kNeedWide = -1;  // a sequence of 65536 ids or more: the slot cannot travel as 16-bit lengths

__attribute__((target_clones("avx2", "default")))
uint64_t lengths_and_check(const uint64_t* lod, int64_t n_seg, uint16_t* dst) {
  uint64_t any = 0;  // OR of all differences: a bit above 15 means "too long" or "decreasing" (wrapped)
  for (int64_t i = 0; i < n_seg; ++i) {
    const uint64_t d = lod[i + 1] - lod[i];
    any |= d;
    dst[i] = static_cast<uint16_t>(d);
  }
  return any;
}

size_t compact_chunks(int64_t n_seg) { return n_seg > 0 ? (static_cast<size_t>(n_seg) + ctrfe::kExpandChunk - 1) / ctrfe::kExpandChunk : 1; }
size_t compact_base_bytes(int64_t n_seg) { return align_up(compact_chunks(n_seg) * 4, 16); }

// One slot's host offsets -> chunk bases (the offset every kExpandChunk-th sequence starts at) + 16-bit
// lengths, validated like stage_one_lod.
int stage_one_lod_compact(int s, const uint64_t* lod, int64_t n_seg, int64_t bound, bool has_ids, char* dst) {
  auto* bases = reinterpret_cast<uint32_t*>(dst);
  const size_t bb = compact_base_bytes(n_seg);
  memset(dst, 0, bb);
  if (n_seg == 0) {
    bases[0] = lod ? static_cast<uint32_t>(lod[0]) : 0u;
    return CTRFE_OK;
  }
  auto* len = reinterpret_cast<uint16_t*>(dst + bb);
  const uint64_t any = lengths_and_check(lod, n_seg, len);
  for (int64_t i = n_seg; i < static_cast<int64_t>(align_up(static_cast<size_t>(n_seg) * 2, 16) / 2); ++i) len[i] = 0;
  if (any >> 16) {
    for (int64_t i = 0; i < n_seg; ++i)
      if (lod[i + 1] < lod[i]) return fail(CTRFE_ERR_ARG, "slot %d: LoD offsets must be non-decreasing", s);
    return kNeedWide;
  }
  const uint64_t last = lod[n_seg];
  if (last >= (1ull << 31))
    return fail(CTRFE_ERR_ARG, "slot %d: LoD offset %llu too large (a slot holds fewer than 2^31 ids)", s, (unsigned long long)last);
  if (has_ids && static_cast<int64_t>(last) > bound)
    return fail(CTRFE_ERR_ARG, "slot %d: the last LoD offset (%llu) exceeds the number of ids (%lld)", s,
                (unsigned long long)last, (long long)bound);
  if (!has_ids && static_cast<int64_t>(last) > bound)
    return fail(CTRFE_ERR_ARG, "slot %d: the first dimension of Input(X) (%lld) must be >= the last LoD offset (%llu)", s,
                (long long)bound, (unsigned long long)last);
  for (size_t c = 0; c < compact_chunks(n_seg); ++c) bases[c] = static_cast<uint32_t>(lod[c * ctrfe::kExpandChunk]);
  return CTRFE_OK;
}

size_t compact_slot_bytes(int64_t n_seg) { return compact_base_bytes(n_seg) + align_up(static_cast<size_t>(n_seg) * 2, 16); }

// Stages all slots of a plan into its pinned image: the compact form when every sequence allows it,
// else 32-bit offsets.  lods[s] = host offsets of slot s.
int plan_stage_image(ctrfe_plan* plan, const uint64_t* const* lods) {
  const size_t S = plan->shapes.size();
  bool compact = plan->dev_c != nullptr;
  if (compact) {
    size_t off = 0;
    for (size_t s = 0; s < S && compact; ++s) {
      const auto& sh = plan->shapes[s];
      const int rc = stage_one_lod_compact(static_cast<int>(s), lods[s], sh.n_seg, sh.bound, sh.has_ids, plan->host_img + off);
      if (rc == kNeedWide) compact = false;
      else if (rc) return rc;
      off += compact_slot_bytes(sh.n_seg);
    }
  }
  if (!compact) {
    size_t off = 0;
    for (size_t s = 0; s < S; ++s) {
      const auto& sh = plan->shapes[s];
      const int rc = stage_one_lod(static_cast<int>(s), lods[s], sh.n_seg, sh.bound, sh.has_ids,
                                   reinterpret_cast<uint32_t*>(plan->host_img + off));
      if (rc) return rc;
      off += align_up(static_cast<size_t>(sh.n_seg + 1) * 4, 16);
    }
  }
  plan->compact = compact;
  return CTRFE_OK;
}

// The plan's pinned image -> its device offsets, on `st`: one copy, plus the expansion when compact.
int plan_upload_image(ctrfe_plan* plan, cudaStream_t st) {
  if (!plan->compact) {
    CU(cudaMemcpyAsync(plan->dev, plan->host_img, plan->img_bytes, cudaMemcpyHostToDevice, st));
    return CTRFE_OK;
  }
  CU(cudaMemcpyAsync(plan->dev_c, plan->host_img, plan->c_bytes, cudaMemcpyHostToDevice, st));
  for (const auto& ep : plan->expand) ctrfe::lod_expand_kernel<<<static_cast<unsigned>(ep.n_blocks), 256, 0, st>>>(ep);
  CU(cudaGetLastError());
  return CTRFE_OK;
}

// Copies the host LoDs into one pinned image of 32-bit offsets, validating them in the same pass,
// and re-points the descriptors at the device image.
int plan_stage_lods(const ctrfe_fused_desc* d, char* host_img, const char* dev_img, ctrfe_plan* plan) {
  size_t off = 0;
  for (int s = 0; s < d->n_slots; ++s) {
    const ctrfe_slot& sl = d->slots[s];
    ctrfe::SlotK& o = plan->chunks[plan->where[s].first].slots[plan->where[s].second];
    o.lod = reinterpret_cast<const unsigned long long*>(dev_img + off);
    const int rc = stage_one_lod(s, sl.lod, sl.n_seg, sl.ids ? sl.n_ids : sl.table_rows, sl.ids != nullptr,
                                 reinterpret_cast<uint32_t*>(host_img + off));
    if (rc) return rc;
    off += align_up(static_cast<size_t>(sl.n_seg + 1) * 4, 16);
  }
  return CTRFE_OK;
}

// One launch of a pooling kernel; `overlap` adds the programmatic-stream-serialization
// attribute, so the grid may start while the kernel before it in the stream is draining.
template <class K>
void launch_k(K kernel, dim3 grid, cudaStream_t st, const ctrfe::KParams& p, bool overlap) {
  if (!overlap) {
    kernel<<<grid, ctrfe::kThreads, 0, st>>>(p);
    return;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = dim3(ctrfe::kThreads, 1, 1);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaLaunchKernelEx(&cfg, kernel, p);
}

template <int DT, int MODE, int CB, bool SH>
void launch_fused_tm(bool vect, int gs, dim3 grid, cudaStream_t st, const ctrfe::KParams& p, bool overlap) {
  using namespace ctrfe;
  if (!vect) { launch_k(fused_pool_kernel<DT, false, 5, MODE, 16, false>, grid, st, p, overlap); return; }
  switch (gs) {
    case 0: launch_k(fused_pool_kernel<DT, true, 0, MODE, CB, SH>, grid, st, p, overlap); break;
    case 1: launch_k(fused_pool_kernel<DT, true, 1, MODE, CB, SH>, grid, st, p, overlap); break;
    case 2: launch_k(fused_pool_kernel<DT, true, 2, MODE, CB, SH>, grid, st, p, overlap); break;
    case 3: launch_k(fused_pool_kernel<DT, true, 3, MODE, CB, SH>, grid, st, p, overlap); break;
    case 4: launch_k(fused_pool_kernel<DT, true, 4, MODE, CB, SH>, grid, st, p, overlap); break;
    default: launch_k(fused_pool_kernel<DT, true, 5, MODE, CB, SH>, grid, st, p, overlap); break;
  }
}

template <int DT>
void launch_fused_t(int mode, bool vect, int gs, int cb, dim3 grid, cudaStream_t st, const ctrfe::KParams& p,
                    bool overlap = false) {
  const bool sel = mode == ctrfe::kModeSelect;
  if (p.n_shards > 0) {  // row-sharded table: 128-bit lanes, rows fetched through peer mappings
    if (sel) launch_fused_tm<DT, ctrfe::kModeSelect, 16, true>(vect, gs, grid, st, p, overlap);
    else launch_fused_tm<DT, ctrfe::kModeAdd, 16, true>(vect, gs, grid, st, p, overlap);
    return;
  }
  if (cb == 32) {  // 256-bit lanes
    if (sel) launch_fused_tm<DT, ctrfe::kModeSelect, 32, false>(vect, gs, grid, st, p, overlap);
    else launch_fused_tm<DT, ctrfe::kModeAdd, 32, false>(vect, gs, grid, st, p, overlap);
    return;
  }
  if (sel) launch_fused_tm<DT, ctrfe::kModeSelect, 16, false>(vect, gs, grid, st, p, overlap);
  else launch_fused_tm<DT, ctrfe::kModeAdd, 16, false>(vect, gs, grid, st, p, overlap);
}

template <int DT, bool CVM, bool SH, bool HOT, int CB>
void launch_fast_g(int gs, dim3 grid, cudaStream_t st, const ctrfe::KParams& p, bool overlap) {
  using namespace ctrfe;
  switch (gs) {
    case 0: launch_k(fused_pool_fast_kernel<DT, 0, CVM, SH, HOT, CB>, grid, st, p, overlap); break;
    case 1: launch_k(fused_pool_fast_kernel<DT, 1, CVM, SH, HOT, CB>, grid, st, p, overlap); break;
    case 2: launch_k(fused_pool_fast_kernel<DT, 2, CVM, SH, HOT, CB>, grid, st, p, overlap); break;
    case 3: launch_k(fused_pool_fast_kernel<DT, 3, CVM, SH, HOT, CB>, grid, st, p, overlap); break;
    case 4: launch_k(fused_pool_fast_kernel<DT, 4, CVM, SH, HOT, CB>, grid, st, p, overlap); break;
    default: launch_k(fused_pool_fast_kernel<DT, 5, CVM, SH, HOT, CB>, grid, st, p, overlap); break;
  }
}

template <int DT, int CB>
void launch_fast_c(int gs, bool cvm, dim3 grid, cudaStream_t st, const ctrfe::KParams& p, bool overlap) {
  const bool sh = p.n_shards > 0;
  const bool hot = p.hot_rows != nullptr;
  if (hot && sh) {
    if (cvm) launch_fast_g<DT, true, true, true, CB>(gs, grid, st, p, overlap); else launch_fast_g<DT, false, true, true, CB>(gs, grid, st, p, overlap);
  } else if (hot && CB == 16) {  // one unsharded table with a compact region of hot-row replicas (128-bit lanes only)
    if constexpr (CB == 16) {
      if (cvm) launch_fast_g<DT, true, false, true, 16>(gs, grid, st, p, overlap); else launch_fast_g<DT, false, false, true, 16>(gs, grid, st, p, overlap);
    }
  } else if (cvm) {
    if (sh) launch_fast_g<DT, true, true, false, CB>(gs, grid, st, p, overlap); else launch_fast_g<DT, true, false, false, CB>(gs, grid, st, p, overlap);
  } else {
    if (sh) launch_fast_g<DT, false, true, false, CB>(gs, grid, st, p, overlap); else launch_fast_g<DT, false, false, false, CB>(gs, grid, st, p, overlap);
  }
}

template <int DT>
void launch_fast_t(int gs, int cb, bool cvm, dim3 grid, cudaStream_t st, const ctrfe::KParams& p, bool overlap) {
  if (cb == 32) launch_fast_c<DT, 32>(gs, cvm, grid, st, p, overlap);
  else launch_fast_c<DT, 16>(gs, cvm, grid, st, p, overlap);
}

int launch_plan(const ctrfe_plan& plan, cudaStream_t st, bool overlap_first = false) {
  for (size_t c = 0; c < plan.chunks.size(); ++c) {
    const int m = plan.modes[c];
    // the launches of one plan work on different slots: the second may always start while the
    // first drains; the first does so only if the caller said it may (overlap_first)
    const bool overlap = overlap_first || c > 0;
    if (plan.fast && m == ctrfe::kModeAdd) {
      switch (plan.dtype) {
        case CTRFE_F32: launch_fast_t<CTRFE_F32>(plan.g_shift, plan.chunk_bytes, plan.cvm_use, plan.grids[c], st, plan.chunks[c], overlap); break;
        case CTRFE_F16: launch_fast_t<CTRFE_F16>(plan.g_shift, plan.chunk_bytes, plan.cvm_use, plan.grids[c], st, plan.chunks[c], overlap); break;
        default: launch_fast_t<CTRFE_BF16>(plan.g_shift, plan.chunk_bytes, plan.cvm_use, plan.grids[c], st, plan.chunks[c], overlap); break;
      }
      continue;
    }
    switch (plan.dtype) {
      case CTRFE_F32: launch_fused_t<CTRFE_F32>(m, plan.vect, plan.g_shift, plan.chunk_bytes, plan.grids[c], st, plan.chunks[c], overlap); break;
      case CTRFE_F16: launch_fused_t<CTRFE_F16>(m, plan.vect, plan.g_shift, plan.chunk_bytes, plan.grids[c], st, plan.chunks[c], overlap); break;
      default: launch_fused_t<CTRFE_BF16>(m, plan.vect, plan.g_shift, plan.chunk_bytes, plan.grids[c], st, plan.chunks[c], overlap); break;
    }
  }
  CU(cudaGetLastError());
  return CTRFE_OK;
}

// Uploads the host LoDs (if any) into `dev` through a pinned staging slot.
int plan_upload(ctrfe_ctx* ctx, const ctrfe_fused_desc* d, cudaStream_t st, char* dev, size_t bytes,
                ctrfe_plan* plan) {
  if (bytes == 0) return CTRFE_OK;
  Staging* sg = nullptr;
  int rc = stage_acquire(ctx, bytes, &sg);
  if (rc) return rc;
  rc = plan_stage_lods(d, sg->host, dev, plan);
  if (rc) return rc;
  CU(cudaMemcpyAsync(dev, sg->host, bytes, cudaMemcpyHostToDevice, st));
  CU(cudaEventRecord(sg->done, st));
  sg->pending = true;
  plan->dev = dev;
  return CTRFE_OK;
}

}  // namespace

extern "C" {

int ctrfe_version(void) { return CTRFE_VERSION; }
const char* ctrfe_last_error(void) { return g_err; }

int ctrfe_ctx_create(int device, ctrfe_ctx** out) {
  if (!out) return fail(CTRFE_ERR_ARG, "ctx_create: null out");
  int n = 0;
  CU(cudaGetDeviceCount(&n));
  if (device < 0 || device >= n) return fail(CTRFE_ERR_ARG, "ctx_create: device %d of %d", device, n);
  DeviceGuard g(device);
  if (!g.ok) return fail(CTRFE_ERR_CUDA, "ctx_create: cannot select device %d", device);
  cudaDeviceProp prop;
  CU(cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10)
    return fail(CTRFE_ERR_UNSUPPORTED, "this library is built for sm_100a only; device %d is sm_%d%d", device,
                prop.major, prop.minor);
  auto* c = new ctrfe_ctx();
  c->device = device;
  c->sm_count = prop.multiProcessorCount;
  cudaError_t e = cudaMalloc(reinterpret_cast<void**>(&c->err_dev), 16);
  if (e == cudaSuccess) e = cudaMemset(c->err_dev, 0xff, 16);
  if (e == cudaSuccess) e = cudaHostAlloc(reinterpret_cast<void**>(&c->err_host), 32, cudaHostAllocDefault);
  if (e == cudaSuccess) {  // the CVM log table of this device (idempotent: every context refills the same values)
    ctrfe::cvm_table_fill_kernel<<<(ctrfe::kCvmTab + 255) / 256, 256>>>();
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
  }
  if (e != cudaSuccess) {
    ctrfe_ctx_destroy(c);
    return fail(CTRFE_ERR_CUDA, "ctx_create: %s", cudaGetErrorString(e));
  }
  *out = c;
  return CTRFE_OK;
}

int ctrfe_ctx_destroy(ctrfe_ctx* c) {
  if (!c) return CTRFE_OK;
  DeviceGuard g(c->device);
  cudaDeviceSynchronize();
  for (auto& s : c->ring) {
    if (s.done) cudaEventDestroy(s.done);
    if (s.host) cudaFreeHost(s.host);
  }
  if (c->scratch) cudaFree(c->scratch);
  if (c->id_scratch) cudaFree(c->id_scratch);
  if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
  if (c->err_dev) cudaFree(c->err_dev);
  if (c->err_host) cudaFreeHost(c->err_host);
  delete c;
  return CTRFE_OK;
}

// ---- plain device / pinned memory, for host languages that bring no allocator of their own ----
int ctrfe_malloc(int device, uint64_t bytes, void** ptr) {
  if (!ptr) return fail(CTRFE_ERR_ARG, "malloc: null ptr");
  *ptr = nullptr;
  if (bytes == 0) return CTRFE_OK;
  DeviceGuard g(device);
  if (!g.ok) return fail(CTRFE_ERR_CUDA, "malloc: cannot select device %d", device);
  CU(cudaMalloc(ptr, bytes));
  return CTRFE_OK;
}

int ctrfe_free(int device, void* ptr) {
  if (!ptr) return CTRFE_OK;
  DeviceGuard g(device);
  CU(cudaFree(ptr));
  return CTRFE_OK;
}

int ctrfe_host_alloc(uint64_t bytes, void** ptr) {
  if (!ptr) return fail(CTRFE_ERR_ARG, "host_alloc: null ptr");
  *ptr = nullptr;
  if (bytes == 0) return CTRFE_OK;
  CU(cudaHostAlloc(ptr, bytes, cudaHostAllocDefault));
  return CTRFE_OK;
}

int ctrfe_host_free(void* ptr) {
  if (!ptr) return CTRFE_OK;
  CU(cudaFreeHost(ptr));
  return CTRFE_OK;
}

int ctrfe_memcpy(int device, void* dst, const void* src, uint64_t bytes, int kind, void* stream, int wait) {
  if (bytes == 0) return CTRFE_OK;
  if (!dst || !src) return fail(CTRFE_ERR_ARG, "memcpy: null pointer");
  if (kind < CTRFE_COPY_H2D || kind > CTRFE_COPY_D2D) return fail(CTRFE_ERR_ARG, "memcpy: bad kind %d", kind);
  DeviceGuard g(device);
  auto st = static_cast<cudaStream_t>(stream);
  const cudaMemcpyKind k = kind == CTRFE_COPY_H2D ? cudaMemcpyHostToDevice : (kind == CTRFE_COPY_D2H ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice);
  CU(cudaMemcpyAsync(dst, src, bytes, k, st));
  if (wait) CU(cudaStreamSynchronize(st));
  return CTRFE_OK;
}

int ctrfe_stream_synchronize(int device, void* stream) {
  DeviceGuard g(device);
  CU(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
  return CTRFE_OK;
}

int ctrfe_check_ids(ctrfe_ctx* ctx, void* stream, int64_t* slot, int64_t* position, int64_t* id) {
  if (!ctx) return fail(CTRFE_ERR_ARG, "check_ids: null ctx");
  DeviceGuard g(ctx->device);
  auto st = static_cast<cudaStream_t>(stream);
  CU(cudaMemcpyAsync(ctx->err_host, ctx->err_dev, 16, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  const unsigned long long key = ctx->err_host[0], lod_key = ctx->err_host[1];
  if (key == ~0ull && lod_key == ~0ull) return CTRFE_OK;
  CU(cudaMemsetAsync(ctx->err_dev, 0xff, 16, st));
  constexpr unsigned long long kPosMask = (1ull << ctrfe::kErrSlotShift) - 1;
  constexpr unsigned long long kSlotMask = (1ull << (ctrfe::kErrTagShift - ctrfe::kErrSlotShift)) - 1;
  if (lod_key != ~0ull) {  // a kernel met a malformed CTRFE_LOD_DEVICE LoD (the CheckLoD rules, lod_tensor.cc:116-142)
    const int64_t s = static_cast<int64_t>((lod_key >> ctrfe::kErrSlotShift) & kSlotMask);
    const int64_t seq = static_cast<int64_t>(lod_key & kPosMask);
    if (slot) *slot = s;
    if (position) *position = seq;
    if (id) *id = -1;
    return fail(CTRFE_ERR_ARG,
                "slot %lld: the LoD offsets of sequence %lld are decreasing or exceed the number of ids "
                "(the sequence was pooled as empty)", (long long)s, (long long)seq);
  }
  const unsigned tag = static_cast<unsigned>(key >> ctrfe::kErrTagShift);
  const int64_t s = static_cast<int64_t>((key >> ctrfe::kErrSlotShift) & kSlotMask);
  const int64_t pos = static_cast<int64_t>(key & kPosMask);
  long long bad = -1;
  const std::vector<IdSource>& srcs = ctx->tag_ids[tag & 255u];
  if (s < static_cast<int64_t>(srcs.size()) && srcs[s].ids) {
    const IdSource& src = srcs[s];
    if (src.ids_i32 == CTRFE_I16) {
      short v = 0;
      CU(cudaMemcpyAsync(&v, static_cast<const short*>(src.ids) + pos, 2, cudaMemcpyDefault, st));
      CU(cudaStreamSynchronize(st));
      bad = v;
    } else if (src.ids_i32) {
      int v = 0;
      CU(cudaMemcpyAsync(&v, static_cast<const int*>(src.ids) + pos, 4, cudaMemcpyDefault, st));  // (the ids may sit in pinned host memory)
      CU(cudaStreamSynchronize(st));
      bad = v;
    } else {
      CU(cudaMemcpyAsync(&bad, static_cast<const long long*>(src.ids) + pos, 8, cudaMemcpyDefault, st));
      CU(cudaStreamSynchronize(st));
    }
  }
  if (slot) *slot = s;
  if (position) *position = pos;
  if (id) *id = bad;
  return fail(CTRFE_ERR_BAD_ID,
              "Variable value (input) of OP(fluid.layers.embedding) expected >= 0 and < table rows, but got "
              "%lld (slot %lld, position %lld). Please check input value.",
              bad, (long long)s, (long long)pos);
}

int ctrfe_plan_create(ctrfe_ctx* ctx, const ctrfe_fused_desc* desc, void* stream, ctrfe_plan** out) {
  if (!ctx || !out) return fail(CTRFE_ERR_ARG, "plan_create: null ctx / out");
  DeviceGuard g(ctx->device);
  auto* plan = new ctrfe_plan();
  plan->tag = take_tag(ctx);
  size_t bytes = 0;
  int rc = plan_geometry(ctx, desc, plan, &bytes);
  if (rc) { delete plan; return rc; }
  ctx->tag_ids[plan->tag] = plan->id_src;
  if (bytes) {
    // device: the 32-bit offsets the kernels read (+ the compact image they are expanded from);
    // pinned host: one image, in whichever of the two forms the current offsets allow
    size_t cb = 0;
    for (const auto& sh : plan->shapes) cb += compact_slot_bytes(sh.n_seg);
    const char* nocompact = getenv("CTRFE_LOD_WIDE");
    // (below 256 KB the extra launch of the expansion kernel costs more than the bytes it saves: cfg1)
    const char* cmin = getenv("CTRFE_LOD_COMPACT_MIN");  // bytes of the 32-bit image from which the compact form is used
    const size_t compact_min = cmin && cmin[0] ? static_cast<size_t>(atoll(cmin)) : (256u << 10);
    const bool want_compact = !(nocompact && nocompact[0] == '1') && bytes >= compact_min;
    char* dev = nullptr;
    cudaError_t e = cudaMalloc(reinterpret_cast<void**>(&dev), bytes);
    if (e == cudaSuccess && want_compact) e = cudaMalloc(reinterpret_cast<void**>(&plan->dev_c), cb);
    if (e == cudaSuccess) e = cudaHostAlloc(reinterpret_cast<void**>(&plan->host_img), bytes > cb ? bytes : cb, cudaHostAllocDefault);
    if (e != cudaSuccess) {
      if (dev) cudaFree(dev);
      if (plan->dev_c) cudaFree(plan->dev_c);
      delete plan;
      return fail(CTRFE_ERR_CUDA, "plan_create: allocating %zu bytes for the LoD image: %s", bytes, cudaGetErrorString(e));
    }
    plan->owns_dev = true;
    plan->dev = dev;
    plan->img_bytes = bytes;
    plan->c_bytes = cb;
    // where each slot's offsets live on the device (as before), and how to expand them from the compact image
    std::vector<const uint64_t*> lods(desc->n_slots);
    size_t off = 0, coff = 0;
    for (int s = 0; s < desc->n_slots; ++s) {
      ctrfe::SlotK& o = plan->chunks[plan->where[s].first].slots[plan->where[s].second];
      o.lod = reinterpret_cast<const unsigned long long*>(dev + off);
      lods[s] = desc->slots[s].lod;
      if (s % ctrfe::kMaxSlotsPerLaunch == 0) {
        plan->expand.emplace_back();
        memset(&plan->expand.back(), 0, sizeof(ctrfe::ExpandParams));
        plan->expand.back().src = plan->dev_c;
      }
      ctrfe::ExpandParams& ep = plan->expand.back();
      ep.s[ep.n].dst = reinterpret_cast<unsigned long long>(dev + off);
      ep.s[ep.n].base_off = static_cast<unsigned>(coff);
      ep.s[ep.n].len_off = static_cast<unsigned>(coff + compact_base_bytes(desc->slots[s].n_seg));
      ep.s[ep.n].n_seg = static_cast<unsigned>(desc->slots[s].n_seg);
      ep.s[ep.n].first_block = static_cast<unsigned>(ep.n_blocks);
      ep.n_blocks += static_cast<int>(compact_chunks(desc->slots[s].n_seg));
      ++ep.n;
      off += align_up(static_cast<size_t>(desc->slots[s].n_seg + 1) * 4, 16);
      coff += compact_slot_bytes(desc->slots[s].n_seg);
    }
    rc = plan_stage_image(plan, lods.data());
    if (rc == CTRFE_OK) rc = plan_upload_image(plan, static_cast<cudaStream_t>(stream));
    if (rc) { cudaFree(dev); if (plan->dev_c) cudaFree(plan->dev_c); cudaFreeHost(plan->host_img); delete plan; return rc; }
  }
  *out = plan;
  return CTRFE_OK;
}

int ctrfe_plan_launch(ctrfe_plan* plan, void* stream) {
  if (!plan) return fail(CTRFE_ERR_ARG, "plan_launch: null plan");
  DeviceGuard g(plan->ctx->device);
  return launch_plan(*plan, static_cast<cudaStream_t>(stream));
}

int ctrfe_plan_launch_ex(ctrfe_plan* plan, void* stream, uint32_t flags) {
  if (!plan) return fail(CTRFE_ERR_ARG, "plan_launch_ex: null plan");
  if (flags & ~static_cast<uint32_t>(CTRFE_LAUNCH_OVERLAP_PREVIOUS)) return fail(CTRFE_ERR_ARG, "plan_launch_ex: unknown flag");
  DeviceGuard g(plan->ctx->device);
  return launch_plan(*plan, static_cast<cudaStream_t>(stream), (flags & CTRFE_LAUNCH_OVERLAP_PREVIOUS) != 0);
}

int ctrfe_plan_update_lod(ctrfe_plan* plan, const uint64_t* const* lods) {
  if (!plan || !lods) return fail(CTRFE_ERR_ARG, "plan_update_lod: null argument");
  if (!plan->host_img) return fail(CTRFE_ERR_ARG, "plan_update_lod: the plan was not created from host LoD offsets");
  for (size_t s = 0; s < plan->shapes.size(); ++s)
    if (plan->shapes[s].n_seg > 0 && !lods[s]) return fail(CTRFE_ERR_ARG, "plan_update_lod: slot %zu: null LoD", s);
  return plan_stage_image(plan, lods);
}

int ctrfe_plan_launch_many(ctrfe_plan* const* plans, int n, void* stream, uint32_t flags) {
  if (!plans || n < 0) return fail(CTRFE_ERR_ARG, "plan_launch_many: bad argument");
  if (flags & ~static_cast<uint32_t>(CTRFE_LAUNCH_OVERLAP_PREVIOUS | CTRFE_LAUNCH_UPLOAD_LOD))
    return fail(CTRFE_ERR_ARG, "plan_launch_many: unknown flag");
  if (n == 0) return CTRFE_OK;
  for (int i = 0; i < n; ++i)
    if (!plans[i] || plans[i]->ctx != plans[0]->ctx) return fail(CTRFE_ERR_ARG, "plan_launch_many: plan %d is null or of another context", i);
  ctrfe_ctx* ctx = plans[0]->ctx;
  DeviceGuard g(ctx->device);
  const bool overlap = (flags & CTRFE_LAUNCH_OVERLAP_PREVIOUS) != 0;
  const bool upload = (flags & CTRFE_LAUNCH_UPLOAD_LOD) != 0;
  auto st = static_cast<cudaStream_t>(stream);
  if (upload && !ctx->copy_stream) {
    // highest priority: the small expansion kernels of the next block must get SM slots while a chain
    // of overlapped pooling kernels keeps every SM full
    int lo = 0, hi = 0;
    CU(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    CU(cudaStreamCreateWithPriority(&ctx->copy_stream, cudaStreamNonBlocking, hi));
  }
  if (!upload) {
    for (int i = 0; i < n; ++i) {
      const int rc = launch_plan(*plans[i], st, overlap);
      if (rc) return rc;
    }
    return CTRFE_OK;
  }
  // With uploads the list is walked in blocks of up to kUploadBlock distinct plans: the block's LoD
  // images go up on the context's copy stream (beside the previous block's kernels), `stream` waits
  // ONCE for the block's last upload, then the block's kernels follow each other without anything in
  // between -- an event wait in front of a kernel would cost it the overlap with its predecessor.
  constexpr int kUploadBlock = 8;
  int i = 0;
  while (i < n) {
    int j = i;
    while (j < n && j - i < kUploadBlock) {
      bool repeat = false;
      for (int q = i; q < j; ++q) repeat |= plans[q] == plans[j];
      if (repeat) break;
      ++j;
    }
    cudaEvent_t last_ready = nullptr;
    for (int q = i; q < j; ++q) {
      ctrfe_plan* p = plans[q];
      if (!p->host_img) continue;
      // the image is not overwritten before the plan's previous launch is through
      if (!p->lod_ready) CU(cudaEventCreateWithFlags(&p->lod_ready, cudaEventDisableTiming));
      if (p->done) CU(cudaStreamWaitEvent(ctx->copy_stream, p->done, 0));
      const int urc = plan_upload_image(p, ctx->copy_stream);
      if (urc) return urc;
      CU(cudaEventRecord(p->lod_ready, ctx->copy_stream));
      last_ready = p->lod_ready;
    }
    if (last_ready) CU(cudaStreamWaitEvent(st, last_ready, 0));  // the copy stream is in order: the last upload covers the block
    for (int q = i; q < j; ++q) {
      ctrfe_plan* p = plans[q];
      const int rc = launch_plan(*p, st, overlap);
      if (rc) return rc;
      if (p->host_img) {
        if (!p->done) CU(cudaEventCreateWithFlags(&p->done, cudaEventDisableTiming));
        CU(cudaEventRecord(p->done, st));
      }
    }
    i = j;
  }
  return CTRFE_OK;
}

int ctrfe_plan_destroy(ctrfe_plan* plan) {
  if (!plan) return CTRFE_OK;
  DeviceGuard g(plan->ctx->device);
  if (plan->owns_dev && plan->dev) {
    cudaDeviceSynchronize();
    cudaFree(plan->dev);
  }
  if (plan->host_img) cudaFreeHost(plan->host_img);
  if (plan->dev_c) cudaFree(plan->dev_c);
  if (plan->lod_ready) cudaEventDestroy(plan->lod_ready);
  if (plan->done) cudaEventDestroy(plan->done);
  delete plan;
  return CTRFE_OK;
}

int ctrfe_plan_num_kernels(const ctrfe_plan* plan) { return plan ? static_cast<int>(plan->chunks.size()) : 0; }

const char* ctrfe_plan_kernel_name(const ctrfe_plan* plan, int i) {
  if (!plan || i < 0 || i >= static_cast<int>(plan->kernel_names.size())) return "";
  return plan->kernel_names[i].c_str();
}

int ctrfe_fused_forward(ctrfe_ctx* ctx, const ctrfe_fused_desc* desc, void* stream) {
  if (!ctx) return fail(CTRFE_ERR_ARG, "fused_forward: null ctx");
  DeviceGuard g(ctx->device);
  auto st = static_cast<cudaStream_t>(stream);
  ctrfe_plan plan;
  plan.tag = take_tag(ctx);
  size_t bytes = 0;
  int rc = plan_geometry(ctx, desc, &plan, &bytes);
  if (rc) return rc;
  rc = scratch_reserve(ctx, bytes, st);
  if (rc) return rc;
  rc = plan_upload(ctx, desc, st, ctx->scratch, bytes, &plan);
  if (rc) return rc;
  ctx->tag_ids[plan.tag].swap(plan.id_src);
  return launch_plan(plan, st);
}

int ctrfe_fused_embedding_seq_pool(ctrfe_ctx* ctx, const void* table, int64_t table_rows, int64_t dim, int32_t dtype,
                                   const int64_t* ids, int64_t n_rows, int64_t index_width, const uint64_t* lod_host,
                                   int64_t n_seg, int64_t padding_idx, void* out, void* stream) {
  if (!ctx) return fail(CTRFE_ERR_ARG, "fused_embedding_seq_pool: null ctx");
  if (!lod_host) return fail(CTRFE_ERR_ARG, "fused_embedding_seq_pool: The LoD level of Input(Ids) should be 1.");
  if (index_width < 1 || index_width > 4096 || n_rows < 0 || n_seg < 0 || dim <= 0)
    return fail(CTRFE_ERR_ARG, "fused_embedding_seq_pool: bad shape (index_width %lld)", (long long)index_width);
  if (n_seg == 0) return CTRFE_OK;
  if (!table || !out || (n_rows > 0 && !ids)) return fail(CTRFE_ERR_ARG, "fused_embedding_seq_pool: null pointer");
  DeviceGuard g(ctx->device);
  auto st = static_cast<cudaStream_t>(stream);
  const long long* cols = reinterpret_cast<const long long*>(ids);
  if (index_width > 1 && n_rows > 0) {
    const size_t need = static_cast<size_t>(n_rows) * index_width * 8;
    if (ctx->id_scratch_cap < need) {
      if (ctx->id_scratch) { CU(cudaStreamSynchronize(st)); CU(cudaFree(ctx->id_scratch)); ctx->id_scratch = nullptr; }
      CU(cudaMalloc(reinterpret_cast<void**>(&ctx->id_scratch), need + need / 4));
      ctx->id_scratch_cap = need + need / 4;
    }
    const long long total = n_rows * index_width;
    ctrfe::ids_transpose_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, st>>>(
        reinterpret_cast<const long long*>(ids), n_rows, static_cast<int>(index_width), reinterpret_cast<long long*>(ctx->id_scratch));
    CU(cudaGetLastError());
    cols = reinterpret_cast<const long long*>(ctx->id_scratch);
  }
  std::vector<ctrfe_slot> slots(static_cast<size_t>(index_width));
  const int es = elem_size(dtype);
  for (int64_t w = 0; w < index_width; ++w) {
    ctrfe_slot& sl = slots[w];
    sl.table = table;
    sl.table_rows = table_rows;
    sl.ids = cols + w * n_rows;
    sl.n_ids = n_rows;
    sl.lod = lod_host;
    sl.n_seg = n_seg;
    sl.out = static_cast<char*>(out) + w * dim * es;  // column w of Ids pools into columns [w*D, (w+1)*D) of Out
    sl.out_ld = dim * index_width;
    sl.pooltype = CTRFE_POOL_SUM;
    sl.ids_dtype = CTRFE_I64;
  }
  ctrfe_fused_desc d{};
  d.n_slots = static_cast<int32_t>(index_width);
  d.dtype = dtype;
  d.slots = slots.data();
  d.dim = dim;
  d.padding_idx = padding_idx;
  d.pad_value = 0.0f;
  d.cvm_mode = CTRFE_CVM_NONE;
  d.cvm_offset = 2;
  d.scale_mode = CTRFE_SCALE_DIVIDE;
  d.init_mode = CTRFE_INIT_FIRST_ROW;  // refer.h:513-517: the first row is copied, the rest added in id order
  d.lod_location = CTRFE_LOD_HOST;
  return ctrfe_fused_forward(ctx, &d, stream);
}

int ctrfe_embedding(ctrfe_ctx* ctx, const void* table, int64_t table_rows, int64_t row_bytes, const void* ids,
                    int32_t ids_dtype, int64_t n_ids, int64_t padding_idx, void* out, void* stream) {
  if (!ctx) return fail(CTRFE_ERR_ARG, "embedding: null ctx");
  if (ids_dtype != CTRFE_I64 && ids_dtype != CTRFE_I32 && ids_dtype != CTRFE_I16)
    return fail(CTRFE_ERR_UNSUPPORTED, "emebdding input only support int16, int32 and int64");
  if (row_bytes <= 0 || table_rows < 0 || n_ids < 0) return fail(CTRFE_ERR_ARG, "embedding: bad sizes");
  if (n_ids == 0) return CTRFE_OK;
  if (!table || !ids || !out) return fail(CTRFE_ERR_ARG, "embedding: null pointer");
  DeviceGuard g(ctx->device);
  auto st = static_cast<cudaStream_t>(stream);
  const uintptr_t al = reinterpret_cast<uintptr_t>(table) | reinterpret_cast<uintptr_t>(out) |
                       static_cast<uintptr_t>(row_bytes);
  int ub = 16;
  while (ub > 1 && (al % ub) != 0) ub >>= 1;
  const long long total = n_ids * (row_bytes / ub);
  long long blocks = (total + static_cast<long long>(ctrfe::kThreads) * ctrfe::kUnroll - 1) /
                     (static_cast<long long>(ctrfe::kThreads) * ctrfe::kUnroll);
  const long long cap = static_cast<long long>(ctx->sm_count) * 8 * 4;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  ctx->tag_ids[0].assign(1, IdSource{ids, ids_dtype});  // the unfused ops record untagged keys
  const char* t = static_cast<const char*>(table);
  char* o = static_cast<char*>(out);
  const int i32 = ids_dtype;  // CTRFE_I64 / I32 / I16
  const unsigned gb = static_cast<unsigned>(blocks);
#define CTRFE_GATHER(UB) \
  ctrfe::gather_rows_kernel<UB><<<gb, ctrfe::kThreads, 0, st>>>(t, table_rows, row_bytes, ids, i32, n_ids, padding_idx, o, ctx->err_dev)
  switch (ub) {
    case 16: CTRFE_GATHER(16); break;
    case 8: CTRFE_GATHER(8); break;
    case 4: CTRFE_GATHER(4); break;
    case 2: CTRFE_GATHER(2); break;
    default: CTRFE_GATHER(1); break;
  }
#undef CTRFE_GATHER
  CU(cudaGetLastError());
  return CTRFE_OK;
}

int ctrfe_sequence_pool(ctrfe_ctx* ctx, const void* x, int32_t dtype, int64_t width, const uint64_t* lod_host,
                        int64_t n_seg, int32_t pooltype, float pad_value, void* out, int32_t* max_index,
                        void* stream) {
  if (!ctx) return fail(CTRFE_ERR_ARG, "sequence_pool: null ctx");
  if (!lod_host) return fail(CTRFE_ERR_ARG, "sequence_pool: Input(X) must carry LoD (1 or 2 levels)");
  if (n_seg < 0) return fail(CTRFE_ERR_ARG, "sequence_pool: negative n_seg");
  ctrfe_slot sl{};
  sl.table = x;
  sl.table_rows = static_cast<int64_t>(lod_host[n_seg]);
  sl.ids = nullptr;
  sl.lod = lod_host;
  sl.n_seg = n_seg;
  sl.out = out;
  sl.out_ld = width;
  sl.pooltype = pooltype;
  sl.ids_dtype = CTRFE_I64;
  ctrfe_fused_desc d{};
  d.n_slots = 1;
  d.dtype = dtype;
  d.slots = &sl;
  d.dim = width;
  d.padding_idx = -1;
  d.pad_value = pad_value;
  d.cvm_mode = CTRFE_CVM_NONE;
  d.cvm_offset = 2;
  d.scale_mode = CTRFE_SCALE_DIVIDE;
  d.init_mode = CTRFE_INIT_ZERO;
  d.lod_location = CTRFE_LOD_HOST;
  if (n_seg == 0) return CTRFE_OK;
  int rc = ctrfe_fused_forward(ctx, &d, stream);
  if (rc) return rc;
  if (max_index && pooltype == CTRFE_POOL_MAX) {
    DeviceGuard g(ctx->device);
    auto st = static_cast<cudaStream_t>(stream);
    // the single slot's LoD image sits at the start of the context scratch
    const auto* lod_dev = reinterpret_cast<const unsigned*>(ctx->scratch);  // the upload holds 32-bit offsets
    const long long total = n_seg * width;
    const unsigned blocks = static_cast<unsigned>((total + 255) / 256);
    const char* xc = static_cast<const char*>(x);
    switch (dtype) {
      case CTRFE_F32: ctrfe::max_index_kernel<CTRFE_F32><<<blocks, 256, 0, st>>>(xc, lod_dev, n_seg, (int)width, max_index); break;
      case CTRFE_F16: ctrfe::max_index_kernel<CTRFE_F16><<<blocks, 256, 0, st>>>(xc, lod_dev, n_seg, (int)width, max_index); break;
      default: ctrfe::max_index_kernel<CTRFE_BF16><<<blocks, 256, 0, st>>>(xc, lod_dev, n_seg, (int)width, max_index); break;
    }
    CU(cudaGetLastError());
  }
  return CTRFE_OK;
}

int ctrfe_cvm(ctrfe_ctx* ctx, const void* x, int32_t dtype, int64_t rows, int64_t width, int32_t use_cvm,
              void* y, void* stream) {
  if (!ctx) return fail(CTRFE_ERR_ARG, "cvm: null ctx");
  if (dtype != CTRFE_F32 && dtype != CTRFE_F16 && dtype != CTRFE_BF16 && dtype != CTRFE_F64)
    return fail(CTRFE_ERR_UNSUPPORTED, "cvm: dtype %d not supported (f32, f64, f16, bf16)", dtype);
  if (width < 2) return fail(CTRFE_ERR_ARG, "cvm: Input(X) needs at least the show and click columns (width=%lld)", (long long)width);
  if (rows < 0) return fail(CTRFE_ERR_ARG, "cvm: negative rows");
  const long long wo = use_cvm ? width : width - 2;
  const long long total = rows * wo;
  if (total == 0) return CTRFE_OK;
  if (!x || !y) return fail(CTRFE_ERR_ARG, "cvm: null pointer");
  DeviceGuard g(ctx->device);
  auto st = static_cast<cudaStream_t>(stream);
  const unsigned blocks = static_cast<unsigned>((total + 255) / 256);
  const char* xc = static_cast<const char*>(x);
  char* yc = static_cast<char*>(y);
  switch (dtype) {
    case CTRFE_F64:
      ctrfe::cvm_f64_kernel<<<blocks, 256, 0, st>>>(static_cast<const double*>(x), rows, (int)width, use_cvm, static_cast<double*>(y));
      break;
    case CTRFE_F32: ctrfe::cvm_kernel<CTRFE_F32><<<blocks, 256, 0, st>>>(xc, rows, (int)width, use_cvm, yc); break;
    case CTRFE_F16: ctrfe::cvm_kernel<CTRFE_F16><<<blocks, 256, 0, st>>>(xc, rows, (int)width, use_cvm, yc); break;
    default: ctrfe::cvm_kernel<CTRFE_BF16><<<blocks, 256, 0, st>>>(xc, rows, (int)width, use_cvm, yc); break;
  }
  CU(cudaGetLastError());
  return CTRFE_OK;
}


int ctrfe_keymap_build(ctrfe_ctx* ctx, const uint64_t* keys, int64_t n, uint64_t* map_keys, uint32_t* map_rows,
                       int32_t log2_slots, uint32_t* n_unplaced, void* stream) {
  if (!ctx || !keys || !map_keys || !map_rows || n < 0 || n >= 0xfffffffeLL || log2_slots < 4 || log2_slots > 31)
    return fail(CTRFE_ERR_ARG, "keymap_build: bad argument");
  if (n == 0) return CTRFE_OK;
  DeviceGuard g(ctx->device);
  const unsigned blocks = static_cast<unsigned>((n + 255) / 256);
  // rows start at all ones so that duplicate keys resolve to their FIRST occurrence (atomicMin)
  CU(cudaMemsetAsync(map_rows, 0xff, sizeof(uint32_t) << log2_slots, static_cast<cudaStream_t>(stream)));
  ctrfe::keymap_insert_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const unsigned long long*>(keys), n, reinterpret_cast<unsigned long long*>(map_keys), map_rows,
      (1u << log2_slots) - 1u, 64 - log2_slots, n_unplaced);
  CU(cudaGetLastError());
  return CTRFE_OK;
}

// ---- backward of the three ops (fp32) -----------------------------------------------------
namespace {
// host LoD -> the context scratch (through a pinned staging slot), validated: starts anywhere,
// non-decreasing, last offset <= limit
int upload_lod(ctrfe_ctx* ctx, const uint64_t* lod_host, int64_t n, int64_t limit, cudaStream_t st, const char* who) {
  for (int64_t i = 1; i < n; ++i)
    if (lod_host[i] < lod_host[i - 1]) return fail(CTRFE_ERR_ARG, "%s: LoD offsets must not decrease", who);
  if (n > 0 && limit >= 0 && static_cast<int64_t>(lod_host[n - 1]) > limit)
    return fail(CTRFE_ERR_ARG, "%s: the last LoD offset (%llu) exceeds the row count (%lld)", who,
                (unsigned long long)lod_host[n - 1], (long long)limit);
  const size_t bytes = static_cast<size_t>(n) * 8;
  int rc = scratch_reserve(ctx, bytes, st);
  if (rc) return rc;
  Staging* sg = nullptr;
  rc = stage_acquire(ctx, bytes, &sg);
  if (rc) return rc;
  memcpy(sg->host, lod_host, bytes);
  CU(cudaMemcpyAsync(ctx->scratch, sg->host, bytes, cudaMemcpyHostToDevice, st));
  CU(cudaEventRecord(sg->done, st));
  sg->pending = true;
  return CTRFE_OK;
}
}  // namespace

int ctrfe_sequence_pool_grad(ctrfe_ctx* ctx, const float* dout, int64_t width, const uint64_t* lod_host, int64_t n_seg,
                             int64_t x_rows, int32_t pooltype, const int32_t* max_index, float* dx, void* stream) {
  if (!ctx || !lod_host) return fail(CTRFE_ERR_ARG, "sequence_pool_grad: null ctx / LoD");
  if (n_seg < 0 || width <= 0 || x_rows < 0) return fail(CTRFE_ERR_ARG, "sequence_pool_grad: bad shape");
  if (pooltype < CTRFE_POOL_AVERAGE || pooltype > CTRFE_POOL_MAX)
    return fail(CTRFE_ERR_ARG, "sequence_pool_grad: unsupported pooling pooltype %d", pooltype);
  if (pooltype == CTRFE_POOL_MAX && !max_index)
    return fail(CTRFE_ERR_ARG, "sequence_pool_grad: MAX needs the MaxIndex of the forward op");
  if (x_rows == 0) return CTRFE_OK;
  if (!dx || (n_seg > 0 && !dout)) return fail(CTRFE_ERR_ARG, "sequence_pool_grad: null pointer");
  DeviceGuard g(ctx->device);
  auto st = static_cast<cudaStream_t>(stream);
  int rc = upload_lod(ctx, lod_host, n_seg + 1, x_rows, st, "sequence_pool_grad");
  if (rc) return rc;
  const bool scattered = pooltype >= CTRFE_POOL_LAST;
  if (scattered || lod_host[0] != 0 || static_cast<int64_t>(lod_host[n_seg]) != x_rows)
    CU(cudaMemsetAsync(dx, 0, static_cast<size_t>(x_rows) * width * sizeof(float), st));
  if (n_seg > 0) {
    const unsigned blocks = static_cast<unsigned>((n_seg * 32 + 255) / 256);
    ctrfe::seqpool_grad_kernel<<<blocks, 256, 0, st>>>(dout, reinterpret_cast<const unsigned long long*>(ctx->scratch), n_seg,
                                                       static_cast<int>(width), pooltype, max_index, dx);
    CU(cudaGetLastError());
  }
  return CTRFE_OK;
}

int ctrfe_cvm_grad(ctrfe_ctx* ctx, const float* dy, const float* cvm, const uint64_t* lod_host, int64_t n_seq,
                   int64_t rows, int64_t width, int32_t use_cvm, float* dx, void* stream) {
  if (!ctx) return fail(CTRFE_ERR_ARG, "cvm_grad: null ctx");
  if (width < 2 || rows < 0 || n_seq < 0) return fail(CTRFE_ERR_ARG, "cvm_grad: bad shape");
  if (rows == 0) return CTRFE_OK;
  if (!dy && (use_cvm || width > 2)) return fail(CTRFE_ERR_ARG, "cvm_grad: null dy");
  if (!cvm || !dx) return fail(CTRFE_ERR_ARG, "cvm_grad: null pointer");
  if (!lod_host && n_seq < rows) return fail(CTRFE_ERR_ARG, "cvm_grad: CVM has %lld rows, X has %lld", (long long)n_seq, (long long)rows);
  DeviceGuard g(ctx->device);
  auto st = static_cast<cudaStream_t>(stream);
  const unsigned long long* lod_dev = nullptr;
  if (lod_host) {
    if (lod_host[0] != 0 || static_cast<int64_t>(lod_host[n_seq]) != rows)
      return fail(CTRFE_ERR_ARG, "cvm_grad: the LoD of X must cover its %lld rows", (long long)rows);
    int rc = upload_lod(ctx, lod_host, n_seq + 1, rows, st, "cvm_grad");
    if (rc) return rc;
    lod_dev = reinterpret_cast<const unsigned long long*>(ctx->scratch);
  }
  const long long total = rows * width;
  ctrfe::cvm_grad_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, st>>>(dy, cvm, lod_dev, n_seq, rows,
                                                                                      static_cast<int>(width), use_cvm, dx);
  CU(cudaGetLastError());
  return CTRFE_OK;
}

int ctrfe_embedding_grad(ctrfe_ctx* ctx, const float* dout, const void* ids, int32_t ids_dtype, int64_t n_ids,
                         int64_t table_rows, int64_t dim, int64_t padding_idx, float* dw, void* stream) {
  if (!ctx) return fail(CTRFE_ERR_ARG, "embedding_grad: null ctx");
  if (ids_dtype != CTRFE_I64 && ids_dtype != CTRFE_I32) return fail(CTRFE_ERR_UNSUPPORTED, "embedding_grad: ids must be int32 or int64");
  if (n_ids < 0 || table_rows < 0 || dim <= 0) return fail(CTRFE_ERR_ARG, "embedding_grad: bad shape");
  if (table_rows == 0) return CTRFE_OK;
  if (!dw || (n_ids > 0 && (!dout || !ids))) return fail(CTRFE_ERR_ARG, "embedding_grad: null pointer");
  DeviceGuard g(ctx->device);
  auto st = static_cast<cudaStream_t>(stream);
  CU(cudaMemsetAsync(dw, 0, static_cast<size_t>(table_rows) * dim * sizeof(float), st));
  if (n_ids > 0) {
    ctx->tag_ids[0].assign(1, IdSource{ids, ids_dtype == CTRFE_I32});
    const long long total = n_ids * dim;
    ctrfe::embedding_grad_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, st>>>(
        dout, ids, ids_dtype == CTRFE_I32, n_ids, table_rows, static_cast<int>(dim), padding_idx, dw, ctx->err_dev);
    CU(cudaGetLastError());
  }
  return CTRFE_OK;
}

int ctrfe_embedding_grad_sparse(ctrfe_ctx* ctx, const void* dout, int64_t row_bytes, const void* ids, int32_t ids_dtype,
                                int64_t n_ids, int64_t* rows, int64_t* rows_host, void* value, void* stream) {
  if (!ctx) return fail(CTRFE_ERR_ARG, "embedding_grad_sparse: null ctx");
  if (ids_dtype != CTRFE_I64 && ids_dtype != CTRFE_I32 && ids_dtype != CTRFE_I16)
    return fail(CTRFE_ERR_UNSUPPORTED, "emebdding input only support int16, int32 and int64");
  if (n_ids < 0 || row_bytes <= 0) return fail(CTRFE_ERR_ARG, "embedding_grad_sparse: bad shape");
  if (n_ids == 0) return CTRFE_OK;
  if (!dout || !ids || !rows || !value) return fail(CTRFE_ERR_ARG, "embedding_grad_sparse: null pointer");
  DeviceGuard g(ctx->device);
  auto st = static_cast<cudaStream_t>(stream);
  long long blocks = (n_ids + 255) / 256;
  if (blocks > 148LL * 16) blocks = 148LL * 16;
  ctrfe::ids_widen_kernel<<<static_cast<unsigned>(blocks), 256, 0, st>>>(ids, ids_dtype, n_ids, reinterpret_cast<long long*>(rows));
  CU(cudaGetLastError());
  if (rows_host) CU(cudaMemcpyAsync(rows_host, rows, static_cast<size_t>(n_ids) * 8, cudaMemcpyDeviceToHost, st));
  if (value != dout)
    CU(cudaMemcpyAsync(value, dout, static_cast<size_t>(n_ids) * static_cast<size_t>(row_bytes), cudaMemcpyDeviceToDevice, st));
  return CTRFE_OK;
}

int ctrfe_hot_cache_build(ctrfe_ctx* ctx, const int64_t* hot_ids, int64_t n, const void* src_rows, int64_t row_bytes,
                          uint32_t* keys, void* rows, int32_t log2_slots, void* stream) {
  if (!ctx || !hot_ids || !src_rows || !keys || !rows || n < 0 || log2_slots < 4 || log2_slots > 30 || row_bytes <= 0 ||
      row_bytes % 16 != 0 || reinterpret_cast<uintptr_t>(src_rows) % 16 != 0 || reinterpret_cast<uintptr_t>(rows) % 16 != 0)
    return fail(CTRFE_ERR_ARG, "hot_cache_build: bad argument");
  if (n == 0) return CTRFE_OK;
  DeviceGuard g(ctx->device);
  const unsigned blocks = static_cast<unsigned>((n * 32 + 255) / 256);
  ctrfe::hot_insert_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const long long*>(hot_ids), n, static_cast<const char*>(src_rows), static_cast<unsigned>(row_bytes),
      keys, static_cast<char*>(rows), (1u << log2_slots) - 1u, 32 - log2_slots);
  CU(cudaGetLastError());
  return CTRFE_OK;
}

// ---- row-sharded tables: exportable shard memory and peer mappings ----------------------
int ctrfe_shard_alloc(int device, uint64_t bytes, void** ptr) {
  if (!ptr || bytes == 0) return fail(CTRFE_ERR_ARG, "shard_alloc: null ptr / zero bytes");
  DeviceGuard g(device);
  if (!g.ok) return fail(CTRFE_ERR_CUDA, "shard_alloc: cannot select device %d", device);
  CU(cudaMalloc(ptr, bytes));  // plain cudaMalloc: cudaIpcGetMemHandle needs it (no pooled memory)
  return CTRFE_OK;
}

int ctrfe_shard_free(int device, void* ptr) {
  if (!ptr) return CTRFE_OK;
  DeviceGuard g(device);
  CU(cudaFree(ptr));
  return CTRFE_OK;
}

int ctrfe_shard_export(void* ptr, unsigned char handle[CTRFE_IPC_HANDLE_BYTES]) {
  static_assert(sizeof(cudaIpcMemHandle_t) == CTRFE_IPC_HANDLE_BYTES, "IPC handle size");
  if (!ptr || !handle) return fail(CTRFE_ERR_ARG, "shard_export: null argument");
  cudaIpcMemHandle_t h;
  CU(cudaIpcGetMemHandle(&h, ptr));
  memcpy(handle, &h, sizeof(h));
  return CTRFE_OK;
}

int ctrfe_shard_open(int device, const unsigned char handle[CTRFE_IPC_HANDLE_BYTES], void** ptr) {
  if (!ptr || !handle) return fail(CTRFE_ERR_ARG, "shard_open: null argument");
  DeviceGuard g(device);
  if (!g.ok) return fail(CTRFE_ERR_CUDA, "shard_open: cannot select device %d", device);
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  // maps the peer's allocation into this process and enables peer access over NVLink
  CU(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return CTRFE_OK;
}

int ctrfe_shard_close(int device, void* ptr) {
  if (!ptr) return CTRFE_OK;
  DeviceGuard g(device);
  CU(cudaIpcCloseMemHandle(ptr));
  return CTRFE_OK;
}

// ---- shard memory through the virtual-memory-management API -------------------------------
// Legacy CUDA IPC mappings are translated with small pages on the importing GPU: random row
// reads over a multi-GB peer shard then run at the page-walker's pace (measured on B200:
// 20 GB/s at 12.8 GB per shard against full speed at 0.64 GB).  Shards are therefore
// allocated with cuMemCreate at the recommended (2 MiB) granularity and shared as POSIX file
// descriptors; the importer maps them with the same granularity.  The driver entry points
// are looked up at run time, so the library has no link-time dependency on libcuda.
namespace {
struct VmmApi {
  CUresult (*GetGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags) = nullptr;
  CUresult (*Create)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long) = nullptr;
  CUresult (*AddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
  CUresult (*Map)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
  CUresult (*SetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
  CUresult (*Export)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long) = nullptr;
  CUresult (*Import)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType) = nullptr;
  CUresult (*Unmap)(CUdeviceptr, size_t) = nullptr;
  CUresult (*Release)(CUmemGenericAllocationHandle) = nullptr;
  CUresult (*AddressFree)(CUdeviceptr, size_t) = nullptr;
  bool ok = false;
};

int vmm_api(VmmApi** out) {
  static VmmApi api;
  if (!api.ok) {
    auto get = [](const char* name, void** fn) -> bool {
      cudaDriverEntryPointQueryResult q;
      return cudaGetDriverEntryPoint(name, fn, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess && *fn;
    };
    bool ok = get("cuMemGetAllocationGranularity", reinterpret_cast<void**>(&api.GetGranularity)) &&
              get("cuMemCreate", reinterpret_cast<void**>(&api.Create)) &&
              get("cuMemAddressReserve", reinterpret_cast<void**>(&api.AddressReserve)) &&
              get("cuMemMap", reinterpret_cast<void**>(&api.Map)) &&
              get("cuMemSetAccess", reinterpret_cast<void**>(&api.SetAccess)) &&
              get("cuMemExportToShareableHandle", reinterpret_cast<void**>(&api.Export)) &&
              get("cuMemImportFromShareableHandle", reinterpret_cast<void**>(&api.Import)) &&
              get("cuMemUnmap", reinterpret_cast<void**>(&api.Unmap)) &&
              get("cuMemRelease", reinterpret_cast<void**>(&api.Release)) &&
              get("cuMemAddressFree", reinterpret_cast<void**>(&api.AddressFree));
    if (!ok) return fail(CTRFE_ERR_CUDA, "shard: the CUDA virtual-memory-management driver API is unavailable");
    api.ok = true;
  }
  *out = &api;
  return CTRFE_OK;
}

#define CUD(call)                                                                            \
  do {                                                                                       \
    CUresult r__ = (call);                                                                   \
    if (r__ != CUDA_SUCCESS) return fail(CTRFE_ERR_CUDA, "%s failed: CUresult %d (%s:%d)", #call, (int)r__, __FILE__, __LINE__); \
  } while (0)

CUmemAllocationProp shard_prop(int device) {
  CUmemAllocationProp prop;
  memset(&prop, 0, sizeof(prop));
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}

int vmm_map(VmmApi* a, int device, CUmemGenericAllocationHandle h, size_t bytes, size_t gran, void** ptr) {
  CUdeviceptr va = 0;
  CUD(a->AddressReserve(&va, bytes, gran, 0, 0));
  CUD(a->Map(va, bytes, 0, h, 0));
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  CUD(a->SetAccess(va, bytes, &acc, 1));
  *ptr = reinterpret_cast<void*>(va);
  return CTRFE_OK;
}
}  // namespace

// Allocates `bytes` (rounded up to the mapping granularity, returned in *mapped_bytes) on
// `device`, maps it for that device, and exports a POSIX file descriptor for the peers.
int ctrfe_shard_alloc_fd(int device, uint64_t bytes, void** ptr, int* fd, uint64_t* mapped_bytes) {
  if (!ptr || !fd || !mapped_bytes || bytes == 0) return fail(CTRFE_ERR_ARG, "shard_alloc_fd: bad argument");
  DeviceGuard g(device);
  if (!g.ok) return fail(CTRFE_ERR_CUDA, "shard_alloc_fd: cannot select device %d", device);
  CU(cudaFree(nullptr));  // make sure the primary context exists
  VmmApi* a = nullptr;
  int rc = vmm_api(&a);
  if (rc) return rc;
  CUmemAllocationProp prop = shard_prop(device);
  size_t gran = 0;
  CUD(a->GetGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  const size_t sz = align_up(static_cast<size_t>(bytes), gran);
  CUmemGenericAllocationHandle h;
  CUD(a->Create(&h, sz, &prop, 0));
  rc = vmm_map(a, device, h, sz, gran, ptr);
  if (rc) return rc;
  int f = -1;
  CUD(a->Export(&f, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  CUD(a->Release(h));  // the mapping (and the exported descriptor) keep the memory alive
  *fd = f;
  *mapped_bytes = sz;
  return CTRFE_OK;
}

// Maps a peer's shard (received as a file descriptor) for `device`.
int ctrfe_shard_open_fd(int device, int fd, uint64_t mapped_bytes, void** ptr) {
  if (!ptr || fd < 0 || mapped_bytes == 0) return fail(CTRFE_ERR_ARG, "shard_open_fd: bad argument");
  DeviceGuard g(device);
  if (!g.ok) return fail(CTRFE_ERR_CUDA, "shard_open_fd: cannot select device %d", device);
  CU(cudaFree(nullptr));
  VmmApi* a = nullptr;
  int rc = vmm_api(&a);
  if (rc) return rc;
  CUmemAllocationProp prop = shard_prop(device);
  size_t gran = 0;
  CUD(a->GetGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  CUmemGenericAllocationHandle h;
  CUD(a->Import(&h, reinterpret_cast<void*>(static_cast<intptr_t>(fd)), CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  rc = vmm_map(a, device, h, static_cast<size_t>(mapped_bytes), gran, ptr);
  if (rc) return rc;
  CUD(a->Release(h));
  return CTRFE_OK;
}

// Unmaps a shard obtained from ctrfe_shard_alloc_fd or ctrfe_shard_open_fd (the memory is
// freed when its last mapping and descriptor are gone).
int ctrfe_shard_unmap(int device, void* ptr, uint64_t mapped_bytes) {
  if (!ptr) return CTRFE_OK;
  DeviceGuard g(device);
  VmmApi* a = nullptr;
  int rc = vmm_api(&a);
  if (rc) return rc;
  const CUdeviceptr va = reinterpret_cast<CUdeviceptr>(ptr);
  CUD(a->Unmap(va, static_cast<size_t>(mapped_bytes)));
  CUD(a->AddressFree(va, static_cast<size_t>(mapped_bytes)));
  return CTRFE_OK;
}

// =====================================================================================
// ctrfe_fused_forward_static: the fused front end for callers that must not allocate.
//
// The ordinary entry points keep their scratch (the device copy of the host LoD offsets and
// the bad-id flag) in memory obtained with cudaMalloc / cudaHostAlloc.  Some callers of this
// library cannot afford that: an op running inside Paddle's executor shares the device with
// the framework's own allocator, which has usually reserved nearly all of HBM up front, so a
// raw cudaMalloc from a plugin can fail outright or fragment what is left; cudaMalloc and
// cudaFree also synchronise the whole device, stalling every other predictor stream in a
// serving process; and allocation is off limits on latency-critical request threads in
// general.  For them the same scratch lives in the fixed array below, which every GPU that
// loads this library holds from load to exit.
//
// Limits and rules (also in include/ctrfe.h):
//   * all LoD offsets of a call, each slot padded to 16 bytes, must fit CTRFE_STATIC_LOD_BYTES
//     (4 MiB, i.e. 524,288 offsets); at most 64 slots of one pool family (all accumulating
//     or all selecting pooltypes).  Larger calls are rejected with CTRFE_ERR_ARG.
//   * the array is shared by every call on the device and keeps the previous call's
//     contents.  This call rewrites the LoD image it uses and resets the bad-id flag itself;
//     two calls whose GPU work overlaps (different streams) race on it -- serialise them.
//   * the first call on each device resolves the array's device address (set-up below);
//     make that first call before any stream capture begins.
// =====================================================================================
#define CTRFE_STATIC_LOD_BYTES (4u << 20)
#define CTRFE_STATIC_FLAG_BYTES 256u
static __device__ __align__(256) unsigned char g_static_scratch[CTRFE_STATIC_FLAG_BYTES + CTRFE_STATIC_LOD_BYTES];

namespace {
constexpr int kMaxStaticDevices = 64;
unsigned char* g_static_ptr[kMaxStaticDevices] = {nullptr};  // device address of the array, per device

// In host code the name g_static_scratch is only a handle for the runtime, not a device
// address; the address a kernel can use comes from cudaGetSymbolAddress, asked once per
// device with that device current.
int static_scratch_for_current_device(unsigned char** out) {
  int dev = -1;
  CU(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxStaticDevices) return fail(CTRFE_ERR_ARG, "static: device %d out of range", dev);
  if (!g_static_ptr[dev]) {
    unsigned char* p = nullptr;
    CU(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_static_scratch));
    if (!p) return fail(CTRFE_ERR_CUDA, "static: cudaGetSymbolAddress returned a null address");
    g_static_ptr[dev] = p;
  }
  *out = g_static_ptr[dev];
  return CTRFE_OK;
}
}  // namespace

int ctrfe_fused_forward_static(const ctrfe_fused_desc* d, void* stream, uint64_t* bad_id_key_host) {
  auto st = static_cast<cudaStream_t>(stream);
  // ---- argument checks (same rules as ctrfe_fused_forward), no heap, no device memory ----
  if (!d || !d->slots || d->n_slots <= 0) return fail(CTRFE_ERR_ARG, "static: no slots");
  if (d->n_slots > ctrfe::kMaxSlotsPerLaunch)
    return fail(CTRFE_ERR_ARG, "static: %d slots exceed the limit of %d per call", d->n_slots, ctrfe::kMaxSlotsPerLaunch);
  if (d->dtype != CTRFE_F32 && d->dtype != CTRFE_F16 && d->dtype != CTRFE_BF16)
    return fail(CTRFE_ERR_UNSUPPORTED, "static: dtype %d not supported (f32, f16, bf16)", d->dtype);
  if (d->dim <= 0 || d->dim > (1 << 20)) return fail(CTRFE_ERR_ARG, "static: bad dim %lld", (long long)d->dim);
  if (d->cvm_mode < CTRFE_CVM_NONE || d->cvm_mode > CTRFE_CVM_STRIP) return fail(CTRFE_ERR_ARG, "static: bad cvm_mode %d", d->cvm_mode);
  if (d->cvm_mode != CTRFE_CVM_NONE && d->dim < 2) return fail(CTRFE_ERR_ARG, "static: cvm needs the show and click columns");
  if (d->cvm_mode == CTRFE_CVM_STRIP && (d->cvm_offset < 0 || d->cvm_offset >= d->dim))
    return fail(CTRFE_ERR_ARG, "static: cvm_offset %d outside [0, dim)", d->cvm_offset);
  if (d->scale_mode != CTRFE_SCALE_DIVIDE && d->scale_mode != CTRFE_SCALE_RECIPROCAL) return fail(CTRFE_ERR_ARG, "static: bad scale_mode");
  if (d->init_mode < CTRFE_INIT_ZERO || d->init_mode > CTRFE_INIT_FIRST_ROW) return fail(CTRFE_ERR_ARG, "static: bad init_mode");
  if (d->lod_location != CTRFE_LOD_HOST) return fail(CTRFE_ERR_ARG, "static: LoD offsets must be passed from host memory");
  if (d->n_shards != 0) return fail(CTRFE_ERR_UNSUPPORTED, "static: the sharded table mode is not available here");
  if (d->keymap) return fail(CTRFE_ERR_UNSUPPORTED, "static: feasign key maps are not available here");

  const int es = elem_size(d->dtype);
  const long long row_bytes = d->dim * es;
  const int d_out = static_cast<int>(d->dim) - (d->cvm_mode == CTRFE_CVM_STRIP ? d->cvm_offset : 0);
  bool v = (row_bytes % 16) == 0;
  size_t lod_bytes = 0;
  long long max_seg = 0;
  const int mode = d->slots[0].pooltype >= CTRFE_POOL_LAST ? ctrfe::kModeSelect : ctrfe::kModeAdd;
  for (int s = 0; s < d->n_slots; ++s) {
    const ctrfe_slot& sl = d->slots[s];
    if (sl.n_seg < 0 || sl.n_seg > 0x7fffffffLL) return fail(CTRFE_ERR_ARG, "static: slot %d: n_seg outside [0, 2^31)", s);
    if (sl.n_seg > 0 && (!sl.table || !sl.lod || !sl.out)) return fail(CTRFE_ERR_ARG, "static: slot %d: null table / lod / out", s);
    if (sl.pooltype < CTRFE_POOL_AVERAGE || sl.pooltype > CTRFE_POOL_MAX) return fail(CTRFE_ERR_ARG, "static: slot %d: bad pooltype %d", s, sl.pooltype);
    if ((sl.pooltype >= CTRFE_POOL_LAST ? ctrfe::kModeSelect : ctrfe::kModeAdd) != mode)
      return fail(CTRFE_ERR_ARG, "static: slot %d mixes accumulating and selecting pooltypes in one call", s);
    if (sl.ids_dtype != CTRFE_I64 && sl.ids_dtype != CTRFE_I32) return fail(CTRFE_ERR_UNSUPPORTED, "static: slot %d: ids must be int32 or int64", s);
    if (sl.out_ld < d_out) return fail(CTRFE_ERR_ARG, "static: slot %d: out_ld %lld < output width %d", s, (long long)sl.out_ld, d_out);
    if (sl.table_rows < 0 || sl.table_rows >= 0xfffffffeLL) return fail(CTRFE_ERR_UNSUPPORTED, "static: slot %d: table_rows outside [0, 2^32 - 2)", s);
    if (sl.ids && (sl.n_ids < 0 || sl.n_ids > 0x7fffffffLL)) return fail(CTRFE_ERR_ARG, "static: slot %d: n_ids outside [0, 2^31)", s);
    if (sl.out_ld * es >= (1LL << 32)) return fail(CTRFE_ERR_ARG, "static: slot %d: out_ld too large", s);
    if (reinterpret_cast<uintptr_t>(sl.table) % 16 != 0) v = false;
    if (sl.n_seg > max_seg) max_seg = sl.n_seg;
    lod_bytes += align_up(static_cast<size_t>(sl.n_seg + 1) * 8, 16);
    // LoD rules (CheckLoD, lod_tensor.h:74-104), read in place from the caller's host memory
    if (sl.n_seg > 0) {
      uint64_t prev = sl.lod[0];
      bool mono = true;
      for (int64_t i = 1; i <= sl.n_seg; ++i) { mono &= sl.lod[i] >= prev; prev = sl.lod[i]; }
      if (!mono) return fail(CTRFE_ERR_ARG, "static: slot %d: LoD offsets must be non-decreasing", s);
      if (prev >= (1ull << 31)) return fail(CTRFE_ERR_ARG, "static: slot %d: a slot holds fewer than 2^31 ids", s);
      if (sl.ids && static_cast<int64_t>(prev) > sl.n_ids)
        return fail(CTRFE_ERR_ARG, "static: slot %d: the last LoD offset (%llu) exceeds the number of ids (%lld)", s,
                    (unsigned long long)prev, (long long)sl.n_ids);
      if (!sl.ids && static_cast<int64_t>(prev) > sl.table_rows) return fail(CTRFE_ERR_ARG, "static: slot %d: last LoD offset exceeds the rows of X", s);
    }
  }
  if (lod_bytes > CTRFE_STATIC_LOD_BYTES)
    return fail(CTRFE_ERR_ARG, "static: %zu bytes of LoD offsets exceed the %u-byte static scratch (use ctrfe_fused_forward)",
                lod_bytes, CTRFE_STATIC_LOD_BYTES);

  unsigned char* scratch = nullptr;
  int rc = static_scratch_for_current_device(&scratch);
  if (rc) return rc;
  auto* flag = reinterpret_cast<unsigned long long*>(scratch);
  unsigned char* lod_img = scratch + CTRFE_STATIC_FLAG_BYTES;

  // the flag starts at "no bad id" (all ones) on every call; the LoD image is overwritten
  CU(cudaMemsetAsync(flag, 0xff, 16, st));  // both error words (bad id, bad LoD pair)

  ctrfe::KParams k;  // 3.7 KB, on the stack
  memset(&k, 0, sizeof(k));
  const long long units = v ? row_bytes / 16 : d->dim;
  int g_shift = 5;
  if (v) { g_shift = 0; while ((1 << g_shift) < units && g_shift < 5) ++g_shift; }
  const int tiles = static_cast<int>((units + (1 << g_shift) - 1) >> g_shift);
  k.err = flag;
  k.padding_idx = d->padding_idx;
  k.dim = static_cast<int>(d->dim);
  k.row_bytes = static_cast<int>(row_bytes);
  k.tiles = tiles;
  k.cvm_mode = d->cvm_mode;
  k.cvm_offset = d->cvm_offset;
  k.scale_mode = d->scale_mode;
  k.init_mode = d->init_mode;
  k.pad_value = d->pad_value;
  k.n_slots = d->n_slots;
  size_t off = 0;
  for (int s = 0; s < d->n_slots; ++s) {
    const ctrfe_slot& sl = d->slots[s];
    ctrfe::SlotK& o = k.slots[s];
    o.table = static_cast<const char*>(sl.table);
    o.ids = sl.ids;
    o.lod = reinterpret_cast<const unsigned long long*>(lod_img + off);
    o.out = static_cast<char*>(sl.out);
    o.out_ld_bytes = static_cast<unsigned>(sl.out_ld * es);
    o.table_rows = static_cast<unsigned>(sl.table_rows);
    o.n_seg = static_cast<unsigned>(sl.n_seg);
    o.n_ids = static_cast<unsigned>(sl.ids ? sl.n_ids : (sl.table_rows > 0x7fffffffLL ? 0x7fffffffLL : sl.table_rows));
    o.pooltype = static_cast<short>(sl.pooltype);
    o.id_kind = static_cast<short>(sl.ids ? (sl.ids_dtype == CTRFE_I32 ? 1 : 0) : 2);
    o.slot_index = s;
    // straight from the caller's host array into the static image (pinned memory keeps this
    // asynchronous; pageable memory is staged by the driver)
    if (sl.lod) CU(cudaMemcpyAsync(lod_img + off, sl.lod, static_cast<size_t>(sl.n_seg + 1) * 8, cudaMemcpyHostToDevice, st));
    off += align_up(static_cast<size_t>(sl.n_seg + 1) * 8, 16);
  }
  const long long items = max_seg * tiles;
  const long long gx = ((items << g_shift) + ctrfe::kThreads - 1) / ctrfe::kThreads;
  if (gx > 0x7fffffffLL) return fail(CTRFE_ERR_ARG, "static: batch too large for one launch");
  const dim3 grid(static_cast<unsigned>(gx > 0 ? gx : 1), static_cast<unsigned>(d->n_slots), 1);
  switch (d->dtype) {
    case CTRFE_F32: launch_fused_t<CTRFE_F32>(mode, v, g_shift, 16, grid, st, k); break;
    case CTRFE_F16: launch_fused_t<CTRFE_F16>(mode, v, g_shift, 16, grid, st, k); break;
    default: launch_fused_t<CTRFE_BF16>(mode, v, g_shift, 16, grid, st, k); break;
  }
  CU(cudaGetLastError());
  if (bad_id_key_host) CU(cudaMemcpyAsync(bad_id_key_host, flag, 8, cudaMemcpyDeviceToHost, st));
  return CTRFE_OK;
}

}  // extern "C"

#include "ctrfe_host_pipe.inc"
