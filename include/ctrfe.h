/*
 * ctrfe.h -- C ABI of the B200 (sm_100a) sparse-feature front end for Paddle Inference
 * CTR models:  lookup_table_v2 (embedding gather over LoD-packed id slots)
 *              -> sequence_pool (SUM / AVERAGE / SQRT / MAX / LAST / FIRST)
 *              -> cvm (show/click transform).
 *
 * This is the drop-in boundary.  Everything is extern "C", plain pointers and sizes;
 * no C++ types, no exceptions, no torch/paddle types.  Streams are passed as void*
 * (a cudaStream_t).  Every function returns CTRFE_OK (0) or a CTRFE_ERR_* code and
 * leaves a human-readable message for ctrfe_last_error() (thread-local).
 *
 * Each entry point names the reference interface it replaces (paths relative to the
 * reference tree, chao9527/Paddle_infer).  INTEGRATION.md shows the Paddle-side binding
 * (PD_BUILD_OP shim, ctypes) a maintainer would add.
 *
 * There is no CPU fallback anywhere behind this header: every compute entry point
 * launches sm_100a kernels or fails.
 */
#ifndef CTRFE_H_
#define CTRFE_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CTRFE_VERSION 122 /* 0.2.0: ctrfe_slot.n_ids, LoD checks on the device, plan_launch_many, fp64 cvm, router */

/* ---- status codes ------------------------------------------------------------------ */
enum {
  CTRFE_OK = 0,
  CTRFE_ERR_BAD_ID = 1,   /* an id outside [0, table_rows): what EmbeddingCPUFunctor throws,
                             paddle/phi/kernels/cpu/embedding_kernel.cc:59-76 */
  CTRFE_ERR_ARG = 2,      /* malformed argument / LoD (framework CheckLoD, lod_tensor.h:74-104;
                             sequence_pool_op.h:42-60) */
  CTRFE_ERR_CUDA = 3,     /* a CUDA runtime call failed */
  CTRFE_ERR_UNSUPPORTED = 4,
  CTRFE_ERR_NCCL = 5
};

/* ---- element types (phi::DataType subset this path is registered for) --------------- */
enum { CTRFE_F32 = 0, CTRFE_F16 = 1, CTRFE_BF16 = 2, CTRFE_F64 = 3 };
/* ---- id types: embedding_kernel.cu:108-117 dispatches int32 / int64 ----------------- */
enum { CTRFE_I64 = 0, CTRFE_I32 = 1, CTRFE_I16 = 2 /* ctrfe_embedding and ctrfe_embedding_grad_sparse only */ };
/* ---- pooltype, in the order SequencePoolOpMaker lists them
 *      (sequence_ops/sequence_pool_op.cc:64-70) ---------------------------------------- */
enum {
  CTRFE_POOL_AVERAGE = 0, CTRFE_POOL_SUM = 1, CTRFE_POOL_SQRT = 2,
  CTRFE_POOL_LAST = 3, CTRFE_POOL_FIRST = 4, CTRFE_POOL_MAX = 5
};
/* ---- what happens after pooling ------------------------------------------------------ */
enum {
  CTRFE_CVM_NONE = 0,  /* no cvm op downstream: write the pooled row */
  CTRFE_CVM_USE = 1,   /* cvm, use_cvm=true: y0=log(x0+1), y1=log(x1+1)-y0 (cvm_op.h:34-37) */
  CTRFE_CVM_STRIP = 2  /* cvm, use_cvm=false: drop the leading cvm_offset columns (cvm_op.h:30-32) */
};
/* ---- how AVERAGE / SQRT scale the sum ------------------------------------------------ */
enum {
  CTRFE_SCALE_DIVIDE = 0,     /* sum/h, sum/sqrt(h): sequence_pool CPU, math/sequence_pooling.cc:422-426 */
  CTRFE_SCALE_RECIPROCAL = 1  /* sum*(1/h), sum*(1/sqrt(h)): the jit kernel the fused CPU op calls,
                                 jit/refer/refer.h:381-389 */
};
/* ---- what the accumulator starts from ------------------------------------------------ */
enum {
  CTRFE_INIT_ZERO = 0,      /* T(0): jit/refer/refer.h:375, sequence_pooling.cu:93 */
  CTRFE_INIT_PAD = 1,       /* pad_value, also for empty segments: fused_seqpool_cvm_op.cu:49 */
  CTRFE_INIT_FIRST_ROW = 2  /* copy of the first row: refer.h:513-517, jit/gen/seqpool.h:61-64 */
};
/* ---- where a slot's LoD offsets live ------------------------------------------------- */
enum {
  CTRFE_LOD_HOST = 0,   /* as Paddle holds them (tensor_meta.h:43); validated on the host, then uploaded
                           once per call, all slots in ONE async copy, no stream sync (contrast
                           mixed_vector.cc:57-76, which syncs per sequence_pool call).  The image that
                           travels holds 32-bit offsets, or -- for a plan whose sequences are all shorter
                           than 65536 -- 16-bit lengths that a small kernel expands on the device */
  CTRFE_LOD_DEVICE = 1  /* already resident in HBM (or in pinned host memory the GPU can read); cannot be
                           validated on the host, so the kernels check every pair of offsets they read
                           (non-decreasing, below 2^31, last <= n_ids: the CheckLoD rules,
                           framework/lod_tensor.cc:116-142), pool an offending sequence as empty and
                           record it for ctrfe_check_ids */
};

/* Opaque per-device context: pinned staging ring, device scratch, the bad-id flag. */
typedef struct ctrfe_ctx ctrfe_ctx;
/* Opaque prepared batch: slot descriptors + LoD resident on the device. */
typedef struct ctrfe_plan ctrfe_plan;

/* One id slot of a CTR batch.  Replaces the (W, Ids+LoD) -> Out triple of ops for one
 * sparse feature: LookupTableV2Op (lookup_table_v2_op.cc:26-104), SequencePoolOp
 * (sequence_ops/sequence_pool_op.cc:23-112), CVMOp (cvm_op.cc:26-59). */
typedef struct ctrfe_slot {
  const void* table;     /* device [table_rows, dim] row-major ("W").  When ids == NULL this is
                            the pre-gathered X [lod[n_seg], dim] (sequence_pool's input). */
  int64_t table_rows;
  const void* ids;       /* device [lod[n_seg]] int64/int32 ("Ids"), or NULL for dense rows */
  const uint64_t* lod;   /* [n_seg + 1] offsets of the LAST LoD level (size_t, lod_tensor.h:59);
                            host or device per ctrfe_fused_desc.lod_location */
  int64_t n_seg;         /* number of sequences (batch size for this slot) */
  void* out;             /* device; pooled row i is written at out + i*out_ld elements.
                            out_ld > D_out gives the concat(axis=1) layout of
                            fusion_seqpool_cvm_concat_op.cc:132-157 */
  int64_t out_ld;
  int32_t pooltype;      /* CTRFE_POOL_* ("pooltype" attr) */
  int32_t ids_dtype;     /* CTRFE_I64 / CTRFE_I32 */
  int64_t n_ids;         /* entries in `ids` (Ids.numel()); the last LoD offset must not exceed it, so a
                            LoD longer than the id tensor is an error instead of a read past the end.
                            Ignored when ids == NULL (table_rows then bounds the LoD, sequence_pool_op.h:52-60).
                            A slot holds fewer than 2^31 ids. */
} ctrfe_slot;

/* Hot-row replica cache for the row-sharded mode.  CTR ids are Zipf-distributed: a few
 * hundred MB of replicated hot rows per GPU absorb most look-ups, which would otherwise all
 * cross NVLink (peer memory is not cached by the local L2).  `keys` is an open-addressing
 * hash of row ids (2^log2_slots entries, linear probing, at most 8 probes) and `rows` holds, at
 * the same slot index, an exact copy of that table row: a hit costs one 4-byte probe.  The table is read-only at inference
 * time, so there is nothing to keep coherent; results stay bit-identical.  The reference has no
 * counterpart on this path (HeterComm always walks to the owner GPU,
 * framework/fleet/heter_ps/heter_comm_inl.h:1477-1607). */
typedef struct ctrfe_hot_cache {
  const void* rows;       /* device [2^log2_slots, dim], same dtype as the table */
  const uint32_t* keys;   /* device [2^log2_slots]; 0xffffffff = empty */
  int32_t log2_slots;
  int32_t n_rows;         /* replicas held (informational) */
} ctrfe_hot_cache;

/* Feasign key map: the `pull_box_sparse`-shaped front.  With a key map in the descriptor the
 * slots' `ids` are uint64 feasign keys (stored as int64, operators/pull_box_sparse_op.h:36-40)
 * and the kernel resolves each key to a table row itself, so keys -> values -> pooled rows is
 * still one launch and the [K, D] values of pull_box_sparse are never written.  Key 0 reads as
 * a zero row (framework/fleet/heter_ps/feature_value.cu:76-80, feature_value.h:498-514); so does
 * a key the map does not hold (the reference's search leaves the value untouched,
 * heter_ps/hashtable_kernel.cu:79-82).  Open addressing, linear probing, at most 32 probes,
 * 0 = empty slot.  The parameter server that fills the reference's table is out of scope: the
 * table here is the dense [n_keys, dim] image of its pull values (show, click, embed_w, mf). */
typedef struct ctrfe_keymap {
  const uint64_t* keys;   /* device [2^log2_slots]; 0 = empty */
  const uint32_t* rows;   /* device [2^log2_slots]; table row of the key in the same slot */
  int32_t log2_slots;
  int32_t reserved;
  int64_t n_keys;         /* keys held (informational) */
} ctrfe_keymap;

typedef struct ctrfe_fused_desc {
  int32_t n_slots;
  int32_t dtype;          /* CTRFE_F32 / F16 / BF16: storage type of tables and outputs;
                             accumulation is always fp32 */
  const ctrfe_slot* slots;/* host array [n_slots] */
  int64_t dim;            /* row width D in elements */
  int64_t padding_idx;    /* "padding_idx" attr, -1 = none (embedding_util.h:19) */
  float pad_value;        /* "pad_value" attr: output of an empty sequence
                             (sequence_pooling.cc:395-398) */
  int32_t cvm_mode;       /* CTRFE_CVM_* ("use_cvm" attr, or no cvm op) */
  int32_t cvm_offset;     /* columns dropped by CTRFE_CVM_STRIP; 2 for cvm_op, the
                             "cvm_offset" attr for fused_seqpool_cvm (fused_seqpool_cvm_op.cc:141) */
  int32_t scale_mode;     /* CTRFE_SCALE_* */
  int32_t init_mode;      /* CTRFE_INIT_* */
  int32_t lod_location;   /* CTRFE_LOD_* */
  int32_t n_shards;       /* 0: each slot reads its own `table`.  2 / 4 / 8: ROW-SHARDED table mode --
                             every slot reads the one logical table whose row r lives on shard
                             r % n_shards at local row r / n_shards (the HeterComm rule,
                             framework/fleet/heter_ps/heter_comm_kernel.cu:85-93); slot.table is
                             ignored and slot.table_rows is the LOGICAL row count */
  const void* const* shard_tables; /* host array [n_shards] of device pointers valid on the calling
                             GPU: its own shard and the peers' shards mapped over NVLink
                             (ctrfe_shard_open).  Replaces c_embedding + c_allreduce_sum
                             (operators/collective/c_embedding_op.cu:33-62,
                             distributed/fleet/layers/mpu/mp_ops.py:547-575) and HeterComm's
                             bucket / walk_to_dest / gather / walk_to_src pull
                             (heter_comm_inl.h:1477-1607): rows are fetched by the pooling kernel
                             itself with loads through the peer mapping, in sequence order */
  const struct ctrfe_hot_cache* hot_cache; /* may be NULL: replicas of the most frequently looked-up rows
                             (see ctrfe_hot_cache_build).  Beside a sharded table they are local copies
                             that save NVLink trips; beside ONE unsharded table that every slot reads they
                             are a compact region the TLB and L2 can hold when the table itself is beyond
                             the TLB's reach (> ~64 GB on B200: cold rows then cost a page walk each) */
  const struct ctrfe_keymap* keymap; /* may be NULL: ids are uint64 feasign keys looked up in this
                             map (not with sharded tables; ids_dtype must be CTRFE_I64) */
} ctrfe_fused_desc;

/* ---- library ------------------------------------------------------------------------ */
int ctrfe_version(void);
const char* ctrfe_last_error(void);

/* Context bound to one CUDA device.  Re-entrant across contexts; one context must not be
 * used from two host threads at once (one predictor = one thread in Paddle). */
int ctrfe_ctx_create(int device, ctrfe_ctx** out);
int ctrfe_ctx_destroy(ctrfe_ctx* ctx);

/* Plain device / pinned host memory and copies, so that a host language that brings no CUDA
 * allocator of its own (Python with numpy only, Go, Java ...) can use the library: thin wrappers
 * over cudaMalloc / cudaFree / cudaHostAlloc / cudaFreeHost / cudaMemcpyAsync.  The reference's
 * counterpart is its allocator facade (memory/malloc.h) and ZeroCopyTensor::CopyFromCpu / CopyToCpu. */
enum { CTRFE_COPY_H2D = 1, CTRFE_COPY_D2H = 2, CTRFE_COPY_D2D = 3 };
int ctrfe_malloc(int device, uint64_t bytes, void** ptr);
int ctrfe_free(int device, void* ptr);
int ctrfe_host_alloc(uint64_t bytes, void** ptr);
int ctrfe_host_free(void* ptr);
int ctrfe_memcpy(int device, void* dst, const void* src, uint64_t bytes, int kind, void* stream, int wait);
int ctrfe_stream_synchronize(int device, void* stream);

/* Bad-id report.  The kernels never fault on an out-of-range id: they treat the row as
 * zero and record the first offender.  This call synchronises `stream`, and returns
 * CTRFE_ERR_BAD_ID (filling slot / position / the id value when non-NULL) if any launch
 * since the last check saw one; the flag is then cleared.  The record carries a tag of the
 * plan (or one-shot call) that met the id, so with several plans in flight the id value is
 * read from the right batch (tags are reused after 255 plans / calls).  A malformed
 * CTRFE_LOD_DEVICE LoD met by a kernel is reported here too, as CTRFE_ERR_ARG with the slot and
 * the sequence index (the sequence was pooled as empty).
 * Replaces the PADDLE_ENFORCE_LT/GE pair of embedding_kernel.cc:59-76 (the reference GPU
 * kernel has no check at all, embedding_kernel.cu:36-51). */
int ctrfe_check_ids(ctrfe_ctx* ctx, void* stream, int64_t* slot, int64_t* position, int64_t* id);

/* ---- the fused hot path: gather + segmented pool + CVM, all slots in ONE launch ------
 * Replaces, per slot, EmbeddingKernel<GPU> (phi/kernels/gpu/embedding_kernel.cu:99-118)
 * + SequencePoolFunctor<GPUContext> (math/sequence_pooling.cu:192-280)
 * + CVMCUDAKernel (cvm_op.cu:86-125), and as a whole FusedSeqpoolCVMCUDAKernel
 * (fused/fused_seqpool_cvm_op.cu:423-506) and the CPU FusionSeqPoolCVMConcatKernel
 * (fused/fusion_seqpool_cvm_concat_op.cc:97-159).
 * Asynchronous on `stream`.  Summation order per sequence is the reference's (row order,
 * one accumulator per column). */
int ctrfe_fused_forward(ctrfe_ctx* ctx, const ctrfe_fused_desc* desc, void* stream);

/* Same, split in two so a serving loop can keep the launch minimal: create validates the
 * descriptor (and host LoD), uploads descriptors + LoD in one async copy on `stream`;
 * launch enqueues exactly one kernel.  A plan may be launched any number of times (the
 * buffers it names must stay valid) and on any stream ordered after the create stream. */
int ctrfe_plan_create(ctrfe_ctx* ctx, const ctrfe_fused_desc* desc, void* stream, ctrfe_plan** out);
int ctrfe_plan_launch(ctrfe_plan* plan, void* stream);
/* ctrfe_plan_launch with flags.  CTRFE_LAUNCH_OVERLAP_PREVIOUS: the caller states that this
 * launch does not consume what the kernel immediately before it in `stream` produces (typical:
 * back-to-back batches of a serving loop); the grid may then start while that kernel is
 * draining (programmatic dependent launch), which hides the tail of a launch -- 15-20 % of a
 * cfg2 step.  Everything later in the stream still waits for both kernels.  Without the flag
 * the launch is an ordinary stream-ordered one.  No reference counterpart. */
enum {
  CTRFE_LAUNCH_OVERLAP_PREVIOUS = 1,
  CTRFE_LAUNCH_UPLOAD_LOD = 2  /* ctrfe_plan_launch_many only: before each launch, upload the plan's host LoD image
                                  again (one async copy on the context's copy stream, so it runs beside the previous
                                  batch's kernel) -- the serving-loop step in which the offsets change every batch
                                  (ctrfe_plan_update_lod); plans created from device LoDs just launch */
};
int ctrfe_plan_launch_ex(ctrfe_plan* plan, void* stream, uint32_t flags);
/* n plans of one context launched back to back on `stream` in one call, with the same flags for
 * each (CTRFE_LAUNCH_OVERLAP_PREVIOUS: every launch may start while the one before it drains) --
 * the serving loop's inner step without a host-language call per batch.  No reference
 * counterpart (the reference launches per op, per slot). */
int ctrfe_plan_launch_many(ctrfe_plan* const* plans, int n, void* stream, uint32_t flags);
/* New offsets for a plan created from host LoDs: lods[s] = the [n_seg + 1] offsets of slot s (same
 * n_seg as at creation).  Validated like ctrfe_plan_create does (CheckLoD, lod_tensor.h:74-104) and
 * written into the plan's pinned image; they reach the device with the next CTRFE_LAUNCH_UPLOAD_LOD
 * launch.  Host-only, no CUDA call.  Stands where the reference re-copies the LoD per op
 * (framework/mixed_vector.cc:57-76). */
int ctrfe_plan_update_lod(ctrfe_plan* plan, const uint64_t* const* lods);
int ctrfe_plan_destroy(ctrfe_plan* plan);
/* Number of kernels one ctrfe_plan_launch enqueues (for launch accounting). */
int ctrfe_plan_num_kernels(const ctrfe_plan* plan);
/* Name and template arguments of the kernel launch i of the plan runs (for profiles and bench
 * records); "" when i is out of range.  Valid until the plan is destroyed. */
const char* ctrfe_plan_kernel_name(const ctrfe_plan* plan, int i);

/* ---- the fused front end without any allocation ------------------------------------
 * No counterpart in the reference.  Same computation and argument rules as
 * ctrfe_fused_forward, for callers that must not allocate memory: an op running inside
 * Paddle's executor shares the GPU with the framework's allocator, which has normally
 * reserved almost all of HBM already, so a cudaMalloc from a plugin may fail or fragment the
 * remainder; cudaMalloc / cudaFree also synchronise the whole device, which stalls the
 * other predictor streams of a serving process.  This entry point therefore takes no
 * context, and keeps its scratch (the device copy of the host LoD offsets, the bad-id flag)
 * in a fixed array inside the library that each GPU holds from load to exit.
 *   - desc->lod_location must be CTRFE_LOD_HOST (pinned host memory keeps the copy async).
 *   - Limits: the LoD offsets of one call, each slot padded to 16 bytes, must fit 4 MiB
 *     (524,288 offsets); at most 64 slots, all accumulating (SUM/AVERAGE/SQRT) or all
 *     selecting (MAX/LAST/FIRST).  Beyond that the call returns CTRFE_ERR_ARG.
 *   - The array is shared by all calls on a device and keeps the last call's contents; each
 *     call rewrites what it uses and resets the flag.  Calls whose GPU work overlaps (for
 *     example on two streams) race on it: serialise them.
 *   - The first call on a device looks up the array's device address; make that call before
 *     any stream capture starts.
 *   - The LoD offsets are copied asynchronously straight from the caller's arrays: when those
 *     live in pinned memory the copy really is asynchronous, so leave them untouched until
 *     `stream` has reached the launch (with pageable memory the driver stages them first).
 *   - bad_id_key_host (may be NULL): host address that receives, once `stream` has reached
 *     this point, ~0 when every id was in range, else (slot << 44 | position) of the first
 *     offender (the row was treated as zeros, as ctrfe_check_ids documents). */
int ctrfe_fused_forward_static(const ctrfe_fused_desc* desc, void* stream, uint64_t* bad_id_key_host);

/* ---- host buffers in, host buffers out ------------------------------------------------
 * The fused front end for a predictor fed from CPU memory: per slot the ids and LoD offsets in
 * host memory, the pooled rows back in host memory.  Replaces, around the fused launch, the
 * reference's feed / fetch path: ZeroCopyTensor::CopyFromCpu + SetLoD per slot and CopyToCpu on
 * the result (inference/api/details/zero_copy_tensor.cc).  Such a step is PCIe-bound, so the
 * slots are cut into n_groups groups and pipelined over both copy engines: group g goes up while
 * group g-1 is pooled and group g-2's rows come down.  The pipe owns its device staging
 * (ids [n_slots, max_ids_per_slot], outputs, LoD images); a run allocates nothing.
 *   create: desc gives the slots' tables, n_seg, pooltypes, id width (one for all slots) and the
 *           attributes; its ids / lod / out pointers are ignored.
 *   run:    ids_host[s] (n_ids[s] entries) and lods_host[s] ([n_seg + 1] offsets) are read,
 *           out_host[s] receives slot s's [n_seg, D_out] rows with row stride out_ld elements.
 *           The offsets go up as they are: only the last one is compared with n_ids on the host;
 *           the kernels check every pair they read (as for CTRFE_LOD_DEVICE), pool an offending
 *           sequence as empty and leave the record for ctrfe_check_ids.  Pinned host memory keeps
 *           the copies asynchronous; arrays that follow each other in host memory (one feed buffer)
 *           go up in one copy per group; ids, offsets and the output must stay untouched until the
 *           run is through.  wait != 0:
 *           returns once the rows are on the host; wait == 0: `stream` is made to wait for them.
 *           Pooling kernels run on `stream`.  One run at a time per pipe.
 *   n_groups < 0 at create: |n_groups| groups with DIRECT output -- the kernels store the pooled rows
 *           into out_host themselves (out_host must then be pinned, device-mapped memory such as
 *           cudaHostAlloc returns, 16-byte aligned) and only the inputs use a copy engine.
 *   create_ex: the same with flags.  CTRFE_PIPE_DIRECT_OUT as above.  CTRFE_PIPE_DIRECT_IN: the
 *           kernels read ids and offsets from the caller's host memory themselves (PCIe reads issued
 *           by the SMs; nothing is uploaded, only the rows use a copy engine).  ids_host[s] and
 *           lods_host[s] must then be pinned, device-mapped memory (cudaHostAlloc / cudaHostRegister);
 *           run() looks each array up (once per run of arrays that follow each other in memory) and
 *           returns CTRFE_ERR_ARG for pageable memory.  CTRFE_PIPE_RAMP: group sizes grow
 *           1 : 2 : 4 : ... instead of being equal (the download engine idles until the first
 *           group is pooled, so that group is kept small).
 *   run with wait < 0: returns at once and leaves `stream` alone; ctrfe_host_pipe_wait(pipe) then
 *           blocks until that run's rows are on the host.  Two pipes used in turn on one stream
 *           keep two requests in flight: request i + 1 goes up and is pooled while request i's
 *           rows come down. */
typedef struct ctrfe_host_pipe ctrfe_host_pipe;
enum { CTRFE_PIPE_DIRECT_OUT = 1, CTRFE_PIPE_DIRECT_IN = 2, CTRFE_PIPE_RAMP = 4 };
int ctrfe_host_pipe_create(ctrfe_ctx* ctx, const ctrfe_fused_desc* desc, int64_t max_ids_per_slot, int n_groups,
                           ctrfe_host_pipe** out);
int ctrfe_host_pipe_create_ex(ctrfe_ctx* ctx, const ctrfe_fused_desc* desc, int64_t max_ids_per_slot, int n_groups,
                              uint32_t flags, ctrfe_host_pipe** out);
int ctrfe_host_pipe_run(ctrfe_host_pipe* pipe, const void* const* ids_host, const int64_t* n_ids,
                        const uint64_t* const* lods_host, void* const* out_host, int64_t out_ld, void* stream, int wait);
int ctrfe_host_pipe_wait(ctrfe_host_pipe* pipe);
int ctrfe_host_pipe_destroy(ctrfe_host_pipe* pipe);

/* ---- row-sharded tables across the GPUs of one box ------------------------------------
 * One process per GPU.  Each process allocates its shard with ctrfe_shard_alloc (plain
 * cudaMalloc memory, so that it can be exported), exports a 64-byte handle, exchanges the
 * handles with its peers by any means (torch.distributed all_gather in this repo), and maps
 * every peer's shard with ctrfe_shard_open.  The mapped pointers go into
 * ctrfe_fused_desc.shard_tables.  Stands where the reference's HeterComm keeps one
 * HashTable per GPU and walks keys/values between them with cudaMemcpyPeerAsync
 * (framework/fleet/heter_ps/heter_comm_inl.h:193,273-428). */
#define CTRFE_IPC_HANDLE_BYTES 64
int ctrfe_shard_alloc(int device, uint64_t bytes, void** ptr);
int ctrfe_shard_free(int device, void* ptr);
int ctrfe_shard_export(void* ptr, unsigned char handle[CTRFE_IPC_HANDLE_BYTES]);
int ctrfe_shard_open(int device, const unsigned char handle[CTRFE_IPC_HANDLE_BYTES], void** ptr);
int ctrfe_shard_close(int device, void* ptr);
/* Preferred variant of the above: the shard is allocated through the CUDA virtual-memory-
 * management API at 2 MiB granularity and shared as a POSIX file descriptor (pass it to the
 * peers over a Unix socket); the importer maps it with large pages, which random row reads
 * over a multi-GB peer shard need (legacy IPC mappings measured 50x slower there). */
int ctrfe_shard_alloc_fd(int device, uint64_t bytes, void** ptr, int* fd, uint64_t* mapped_bytes);
int ctrfe_shard_open_fd(int device, int fd, uint64_t mapped_bytes, void** ptr);
int ctrfe_shard_unmap(int device, void* ptr, uint64_t mapped_bytes);
/* Fills a ctrfe_hot_cache: for each of the n DISTINCT row ids hot_ids[i], claims a hash slot in
 * `keys` and copies src_rows[i] (row_bytes each, a multiple of 16) to rows[slot].  `keys` must
 * have been set to 0xff bytes.  An id that finds no free slot within 8 probes is left out (it
 * is then always fetched from its shard).  Asynchronous on `stream`. */
int ctrfe_hot_cache_build(ctrfe_ctx* ctx, const int64_t* hot_ids, int64_t n, const void* src_rows, int64_t row_bytes,
                          uint32_t* keys, void* rows, int32_t log2_slots, void* stream);

/* Fills a ctrfe_keymap: keys[i] -> row i for i < n (distinct keys; key 0 is skipped).  map_keys
 * must have been zeroed.  *n_unplaced (device, may be NULL) counts keys that found no slot
 * within 32 probes: rebuild with a larger log2_slots if it is not 0.  Asynchronous on `stream`.
 * Stands where the reference builds its per-GPU HashTable (heter_ps/hashtable_kernel.cu:216-235,
 * insert). */
int ctrfe_keymap_build(ctrfe_ctx* ctx, const uint64_t* keys, int64_t n, uint64_t* map_keys, uint32_t* map_rows,
                       int32_t log2_slots, uint32_t* n_unplaced, void* stream);

/* fused_embedding_seq_pool: gather + SUM per sequence in the EmbSeqPool order (first row copied, the
 * rest added in id order).  ids: device int64 [n_rows, index_width] row-major with ONE LoD level over
 * its rows (lod_host: [n_seg + 1] offsets); out: device [n_seg, dim * index_width], column j of ids
 * pooled into columns [j*dim, (j+1)*dim).  Replaces FusedEmbeddingSeqPoolKernel / EmbeddingVSumFunctor
 * (fused/fused_embedding_seq_pool_op.h:84-127) and jit::refer::EmbSeqPool incl. index_width > 1
 * (jit/refer/refer.h:477-528).  Ids outside the table are reported by ctrfe_check_ids (slot = the
 * ids column); the reference kernel has no padding_idx on this path, here it pools as a zero row. */
int ctrfe_fused_embedding_seq_pool(ctrfe_ctx* ctx, const void* table, int64_t table_rows, int64_t dim, int32_t dtype,
                                   const int64_t* ids, int64_t n_rows, int64_t index_width, const uint64_t* lod_host,
                                   int64_t n_seg, int64_t padding_idx, void* out, void* stream);

/* ---- the three ops on their own (1:1 clones of the reference ops) -------------------- */

/* lookup_table_v2 / phi embedding: out[k,:] = (ids[k]==padding_idx) ? 0 : table[ids[k],:].
 * Pure byte copy, any element size (fp64 tables included); ids int64, int32 or int16 as the reference
 * GPU kernel takes them (embedding_kernel.cu:107-117).  Replaces EmbeddingCUDAFunctor / EmbeddingFW
 * (phi/kernels/gpu/embedding_kernel.cu:25-88). */
int ctrfe_embedding(ctrfe_ctx* ctx, const void* table, int64_t table_rows, int64_t row_bytes,
                    const void* ids, int32_t ids_dtype, int64_t n_ids, int64_t padding_idx,
                    void* out, void* stream);

/* sequence_pool: x device [lod[n_seg], width], lod HOST offsets of the last level,
 * out device [n_seg, width]; max_index device int32 [n_seg, width] or NULL (is_test).
 * Replaces SequencePoolFunctor<GPUContext,T>::operator() (math/sequence_pooling.cu:192-280),
 * with the CPU functor's semantics for MAX (sequence_pooling.cc:159-175). */
int ctrfe_sequence_pool(ctrfe_ctx* ctx, const void* x, int32_t dtype, int64_t width,
                        const uint64_t* lod_host, int64_t n_seg, int32_t pooltype, float pad_value,
                        void* out, int32_t* max_index, void* stream);

/* cvm: x device [rows, width] -> y [rows, width] (use_cvm) or [rows, width-2].
 * Replaces CVMCUDAKernel<T>::Compute (cvm_op.cu:86-125); the op's CVM input is not read
 * in forward (cvm_op.h:60-99) and is not a parameter. */
int ctrfe_cvm(ctrfe_ctx* ctx, const void* x, int32_t dtype, int64_t rows, int64_t width,
              int32_t use_cvm, void* y, void* stream);

/* ---- their backward ops, fp32 (inference never calls these) -------- */

/* sequence_pool_grad: dout device [n_seg, width] -> dx device [x_rows, width]; lod HOST offsets of
 * the last level; max_index: the forward op's MaxIndex (MAX only).  SUM copies, AVERAGE divides
 * by the length, SQRT by its square root, LAST / FIRST / MAX scatter into zeros.  Replaces
 * SequencePoolGradFunctor (math/sequence_pooling.cu:282-494; CPU semantics
 * sequence_pooling.cc:179-239,309-343,438-496).  Bit-exact with the CPU functor. */
int ctrfe_sequence_pool_grad(ctrfe_ctx* ctx, const float* dout, int64_t width, const uint64_t* lod_host,
                             int64_t n_seg, int64_t x_rows, int32_t pooltype, const int32_t* max_index,
                             float* dx, void* stream);

/* cvm_grad: dy device [rows, width] (use_cvm) or [rows, width-2]; cvm device [n_seq, 2] (the op's
 * CVM input: show, click); dx device [rows, width] with dx[r, 0:2] = cvm[seq(r)] and the rest
 * copied from dy.  lod_host NULL: X has no LoD, seq(r) = r; else the HOST offsets of X's level 0.
 * Replaces CVMGradCUDAKernel (cvm_op.cu:49-84,127-187; cvm_op.h:43-58,100-140).  Bit-exact. */
int ctrfe_cvm_grad(ctrfe_ctx* ctx, const float* dy, const float* cvm, const uint64_t* lod_host, int64_t n_seq,
                   int64_t rows, int64_t width, int32_t use_cvm, float* dx, void* stream);

/* lookup_table_v2_grad, dense W@GRAD: dw device [table_rows, dim] = sum over k of dout[k, :] into
 * row ids[k]; ids equal to padding_idx are skipped (cpu/embedding_grad_kernel.cc:63-66); ids
 * outside the table are reported by ctrfe_check_ids (slot 0, position k).  Replaces
 * EmbeddingGradCUDAFunctor / EmbeddingGrad (phi/kernels/gpu/embedding_grad_kernel.cu:38-60,
 * 62-130).  Atomic adds as in the reference GPU kernel: the order of additions into a row is
 * not fixed, so results match the CPU kernel to rounding (<= 1e-6 relative), not bit for bit.
 * The SelectedRows (is_sparse) variant is ctrfe_embedding_grad_sparse below. */
int ctrfe_embedding_grad(ctrfe_ctx* ctx, const float* dout, const void* ids, int32_t ids_dtype, int64_t n_ids,
                         int64_t table_rows, int64_t dim, int64_t padding_idx, float* dw, void* stream);

/* lookup_table_v2_grad with is_sparse: W@GRAD as a SelectedRows -- rows (device, int64 [n_ids]) = the ids
 * widened to int64, one entry per id and not merged; value (device, [n_ids, row_bytes]) = dout (any
 * dtype; value == dout: no copy); the SelectedRows height is the table's row count, which the caller
 * knows.  rows_host (may be NULL) receives the same rows once `stream` has reached this point -- the
 * reference keeps the rows vector on the host (MixVector::CopyToCPU).  padding_idx is not applied and
 * ids are not range-checked, as in the reference.  Replaces EmbeddingSparseGradCUDAFunctor
 * (phi/kernels/gpu/embedding_grad_kernel.cu:141-214). */
int ctrfe_embedding_grad_sparse(ctrfe_ctx* ctx, const void* dout, int64_t row_bytes, const void* ids, int32_t ids_dtype,
                                int64_t n_ids, int64_t* rows, int64_t* rows_host, void* value, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CTRFE_H_ */
