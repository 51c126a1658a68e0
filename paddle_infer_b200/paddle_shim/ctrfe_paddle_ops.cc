// ctrfe_paddle_ops.cc -- Paddle custom-operator shim over the ctrfe C ABI.
//
// Built by a Paddle maintainer against an installed Paddle (paddle/extension.h) and linked
// with libctrfe.so; loaded with paddle.utils.cpp_extension.load / LoadCustomOperatorLib.
// It cannot be loaded on the authoring box (Paddle is not installed there); it is
// syntax-checked against the reference headers by paddle_shim/check_syntax.sh.
//
// Why the op names differ from the built-ins: the custom-op loader skips a name that is
// already registered (framework/custom_operator.cc:722-725), so `lookup_table_v2`,
// `sequence_pool`, `cvm` cannot be re-registered through the public API.  This file
// registers
//     ctr_fused_emb_seqpool_cvm   W, Ids@VECTOR -> Out@VECTOR      (the fused front end)
//     b200_lookup_table_v2        W, Ids -> Out                    (1:1 clone)
//     b200_sequence_pool          X -> Out, MaxIndex               (1:1 clone)
//     b200_cvm                    X, CVM -> Y                      (1:1 clone)
//     b200_fused_seqpool_cvm      X@VECTOR, CVM -> Out@VECTOR      (PaddleBox's multi-slot op, 1:1 clone)
//     b200_fused_embedding_seq_pool  W, Ids -> Out                 (1:1 clone)
//     b200_pull_box_sparse        W, MapKeys, MapRows, Ids@VECTOR -> Out@VECTOR  (resident table + key map)
//     b200_fusion_seqpool_cvm_concat X@VECTOR, CVM -> Out          (the CPU predictor's fused op)
// with the reference ops' input/output/attribute names, and paddle_infer_b200/rewrite.py
// retargets a loaded program's op types (the saved model on disk is untouched).
//
// LoD: paddle::Tensor has no LoD accessor; the DenseTensor behind it carries it
// (custom_operator.cc:159,173 copy meta incl. lod; phi/core/tensor_meta.h:72).
#include <cstdint>
#include <map>
#include <mutex>
#include <string>
#include <utility>
#include <vector>

#include "paddle/extension.h"
#include "paddle/phi/core/dense_tensor.h"

#include "ctrfe.h"

namespace {

const phi::LoD& LodOf(const paddle::Tensor& t) {
  auto* dense = static_cast<phi::DenseTensor*>(t.impl().get());
  PD_CHECK(dense != nullptr, "ctrfe: input is not a DenseTensor");
  return dense->lod();
}

// Untyped device pointers: paddle::Tensor::data<T>() is typed; the DenseTensor behind it
// hands out void*.
const void* Raw(const paddle::Tensor& t) {
  return static_cast<const phi::DenseTensor*>(t.impl().get())->data();
}
void* RawMut(paddle::Tensor& t) {  // NOLINT
  return static_cast<phi::DenseTensor*>(t.impl().get())->data();
}

int DtypeCode(paddle::DataType dt) {
  switch (dt) {
    case paddle::DataType::FLOAT32: return CTRFE_F32;
    case paddle::DataType::FLOAT16: return CTRFE_F16;
    case paddle::DataType::BFLOAT16: return CTRFE_BF16;
    default: PD_THROW("ctrfe: only float32 / float16 / bfloat16 tables have sm_100a kernels");
  }
  return -1;
}

int IdCode(paddle::DataType dt) {
  if (dt == paddle::DataType::INT64) return CTRFE_I64;
  if (dt == paddle::DataType::INT32) return CTRFE_I32;
  PD_THROW("emebdding input only support int32 and int64");
  return -1;
}

int PoolCode(const std::string& p) {
  if (p == "AVERAGE") return CTRFE_POOL_AVERAGE;
  if (p == "SUM") return CTRFE_POOL_SUM;
  if (p == "SQRT") return CTRFE_POOL_SQRT;
  if (p == "LAST") return CTRFE_POOL_LAST;
  if (p == "FIRST") return CTRFE_POOL_FIRST;
  if (p == "MAX") return CTRFE_POOL_MAX;
  PD_THROW("unsupported pooling pooltype: ", p);
  return -1;
}

// One context per (device, stream).  A ctrfe_ctx owns one LoD scratch image, one staging ring and
// one error record, all ordered by the stream it is used with, so two predictor clones -- one
// thread and one stream each on the same GPU (analysis_predictor.cc Clone) -- must not share one.
// The map itself is guarded; a context is only ever used from the thread that runs its stream.
ctrfe_ctx* CtxFor(const paddle::Tensor& t) {
  static std::mutex mu;
  static std::map<std::pair<int, const void*>, ctrfe_ctx*> ctxs;
  const int dev = t.place().GetDeviceId();
  PD_CHECK(dev >= 0, "ctrfe: bad device id");
  const void* stream = reinterpret_cast<const void*>(t.stream());
  std::lock_guard<std::mutex> lock(mu);
  auto it = ctxs.find({dev, stream});
  if (it == ctxs.end()) {
    ctrfe_ctx* c = nullptr;
    PD_CHECK(ctrfe_ctx_create(dev, &c) == CTRFE_OK, ctrfe_last_error());
    it = ctxs.emplace(std::make_pair(dev, stream), c).first;
  }
  return it->second;
}

void Check(int rc) { PD_CHECK(rc == CTRFE_OK, ctrfe_last_error()); }

void NeedGpu(const paddle::Tensor& t, const char* what) {
  PD_CHECK(t.is_gpu(), what, " must be on a CUDAPlace: the ctrfe ops have no CPU kernel");
}

}  // namespace

// ---------------------------------------------------------------------------------------
// ctr_fused_emb_seqpool_cvm
// ---------------------------------------------------------------------------------------
std::vector<paddle::Tensor> CtrFusedForward(const paddle::Tensor& w, const std::vector<paddle::Tensor>& ids,
                                            const std::string& pooltype, float pad_value, int64_t padding_idx,
                                            bool use_cvm, int cvm_offset, bool check_ids) {
  NeedGpu(w, "W");
  const auto wshape = w.shape();
  PD_CHECK(wshape.size() == 2, "ShapeError: The dimensions of the 'lookup table' must be 2.");
  const int64_t dim = wshape[1];
  const int64_t d_out = use_cvm ? dim : dim - cvm_offset;
  const int pool = PoolCode(pooltype);
  std::vector<ctrfe_slot> slots(ids.size());
  std::vector<paddle::Tensor> outs;
  outs.reserve(ids.size());
  for (size_t s = 0; s < ids.size(); ++s) {
    NeedGpu(ids[s], "Ids");
    const auto& lod = LodOf(ids[s]);
    PD_CHECK(!lod.empty() && lod.size() <= 2, "Input(Ids) must carry 1 or 2 LoD levels");
    const auto& last = lod.back();
    const int64_t n_seg = static_cast<int64_t>(last.size()) - 1;
    outs.emplace_back(paddle::empty({n_seg, d_out}, w.dtype(), w.place()));
    ctrfe_slot& sl = slots[s];
    sl.table = Raw(w);
    sl.table_rows = wshape[0];
    sl.ids = Raw(ids[s]);
    sl.lod = reinterpret_cast<const uint64_t*>(last.data());  // size_t offsets, host
    sl.n_seg = n_seg;
    sl.out = RawMut(outs.back());
    sl.out_ld = d_out;
    sl.pooltype = pool;
    sl.ids_dtype = IdCode(ids[s].dtype());
    sl.n_ids = ids[s].numel();  // a LoD that names more ids than the tensor holds is refused, not read
  }
  ctrfe_fused_desc d{};
  d.n_slots = static_cast<int32_t>(slots.size());
  d.dtype = DtypeCode(w.dtype());
  d.slots = slots.data();
  d.dim = dim;
  d.padding_idx = padding_idx;
  d.pad_value = pad_value;
  d.cvm_mode = use_cvm ? CTRFE_CVM_USE : CTRFE_CVM_STRIP;
  d.cvm_offset = cvm_offset;
  d.scale_mode = CTRFE_SCALE_DIVIDE;
  d.init_mode = CTRFE_INIT_ZERO;
  d.lod_location = CTRFE_LOD_HOST;
  ctrfe_ctx* ctx = CtxFor(w);
  Check(ctrfe_fused_forward(ctx, &d, w.stream()));
  // check_ids: synchronise and raise on an id outside the table, as the reference's CPU
  // lookup_table_v2 does (embedding_kernel.cc:59-76).  Without it an offending row pools as zeros
  // and the record stays with THIS launch's tag until some later check reports it.
  if (check_ids) Check(ctrfe_check_ids(ctx, w.stream(), nullptr, nullptr, nullptr));
  return outs;
}

std::vector<std::vector<int64_t>> CtrFusedInferShape(const std::vector<int64_t>& w_shape,
                                                     const std::vector<std::vector<int64_t>>& ids_shapes,
                                                     const std::string& pooltype, float pad_value,
                                                     int64_t padding_idx, bool use_cvm, int cvm_offset, bool check_ids) {
  // batch is only known from the LoD at run time, as in fused_seqpool_cvm_op.cc:82-91
  std::vector<std::vector<int64_t>> out;
  for (size_t i = 0; i < ids_shapes.size(); ++i) out.push_back({-1, use_cvm ? w_shape[1] : w_shape[1] - cvm_offset});
  return out;
}

std::vector<paddle::DataType> CtrFusedInferDtype(const paddle::DataType& w_dtype,
                                                 const std::vector<paddle::DataType>& ids_dtypes) {
  return std::vector<paddle::DataType>(ids_dtypes.size(), w_dtype);
}

PD_BUILD_OP(ctr_fused_emb_seqpool_cvm)
    .Inputs({"W", paddle::Vec("Ids")})
    .Outputs({paddle::Vec("Out")})
    .Attrs({"pooltype: std::string", "pad_value: float", "padding_idx: int64_t", "use_cvm: bool",
            "cvm_offset: int", "check_ids: bool"})
    .SetKernelFn(PD_KERNEL(CtrFusedForward))
    .SetInferShapeFn(PD_INFER_SHAPE(CtrFusedInferShape))
    .SetInferDtypeFn(PD_INFER_DTYPE(CtrFusedInferDtype));

// ---------------------------------------------------------------------------------------
// b200_fusion_seqpool_cvm_concat  (FusionSeqPoolCVMConcatOpMaker,
// fused/fusion_seqpool_cvm_concat_op.cc:76-95): X (duplicable, pre-embedded LoD rows), CVM ->
// Out [B, S*w]; what rewrite.fuse_seqpool_cvm_concat emits.  The jit kernel's reciprocal
// scaling (refer.h:381-389), use_cvm must be true and axis 1 as the reference enforces.
// ---------------------------------------------------------------------------------------
std::vector<paddle::Tensor> B200FusionSeqpoolCvmConcat(const std::vector<paddle::Tensor>& xs, const paddle::Tensor& cvm,
                                                       const std::string& pooltype, bool use_cvm, int axis) {
  PD_CHECK(axis == 1, "fusion_seqpool_cvm_concat only supports concat axis=1 yet.");
  PD_CHECK(use_cvm, "fusion_seqpool_cvm_concat only supports use_cvm is true yet.");
  PD_CHECK(!xs.empty(), "Input(X) of fusion_seqpool_cvm_concat is empty");
  NeedGpu(xs[0], "X");
  const int64_t w = xs[0].numel() / xs[0].shape()[0];
  const int64_t n = static_cast<int64_t>(xs.size());
  const auto& lod0 = LodOf(xs[0]);
  PD_CHECK(!lod0.empty(), "Input(X) must carry LoD");
  const int64_t bs = static_cast<int64_t>(lod0[0].size()) - 1;
  auto out = paddle::empty({bs, n * w}, xs[0].dtype(), xs[0].place());
  const int es = xs[0].dtype() == paddle::DataType::FLOAT32 ? 4 : 2;
  std::vector<ctrfe_slot> slots(xs.size());
  for (size_t s = 0; s < xs.size(); ++s) {
    const auto& lod = LodOf(xs[s]);
    PD_CHECK(!lod.empty() && static_cast<int64_t>(lod[0].size()) - 1 == bs, "Batchsize of all inputs should be equal.");
    PD_CHECK(xs[s].numel() / xs[s].shape()[0] == w, "Width of all inputs should be equal.");
    ctrfe_slot& sl = slots[s];
    sl.table = Raw(xs[s]);
    sl.table_rows = xs[s].shape()[0];
    sl.ids = nullptr;  // dense rows: row index = position
    sl.lod = reinterpret_cast<const uint64_t*>(lod[0].data());
    sl.n_seg = bs;
    sl.out = static_cast<char*>(RawMut(out)) + static_cast<int64_t>(s) * w * es;
    sl.out_ld = n * w;
    sl.pooltype = PoolCode(pooltype);
    sl.ids_dtype = CTRFE_I64;
  }
  ctrfe_fused_desc d{};
  d.n_slots = static_cast<int32_t>(slots.size());
  d.dtype = DtypeCode(xs[0].dtype());
  d.slots = slots.data();
  d.dim = w;
  d.padding_idx = -1;
  d.pad_value = 0.0f;
  d.cvm_mode = CTRFE_CVM_USE;
  d.cvm_offset = 2;
  d.scale_mode = CTRFE_SCALE_RECIPROCAL;
  d.init_mode = CTRFE_INIT_ZERO;
  d.lod_location = CTRFE_LOD_HOST;
  Check(ctrfe_fused_forward(CtxFor(xs[0]), &d, xs[0].stream()));
  return {out};
}

std::vector<std::vector<int64_t>> B200FusionSeqpoolCvmConcatInferShape(const std::vector<std::vector<int64_t>>& x_shapes,
                                                                       const std::vector<int64_t>& cvm_shape,
                                                                       const std::string& pooltype, bool use_cvm,
                                                                       int axis) {
  int64_t w = 1;
  for (size_t i = 1; i < x_shapes[0].size(); ++i) w *= x_shapes[0][i];
  return {{-1, static_cast<int64_t>(x_shapes.size()) * w}};  // batch comes from the LoD at run time
}

std::vector<paddle::DataType> B200FusionSeqpoolCvmConcatInferDtype(const std::vector<paddle::DataType>& x_dtypes,
                                                                   const paddle::DataType& cvm_dtype) {
  return {x_dtypes[0]};
}

PD_BUILD_OP(b200_fusion_seqpool_cvm_concat)
    .Inputs({paddle::Vec("X"), "CVM"})
    .Outputs({"Out"})
    .Attrs({"pooltype: std::string", "use_cvm: bool", "axis: int"})
    .SetKernelFn(PD_KERNEL(B200FusionSeqpoolCvmConcat))
    .SetInferShapeFn(PD_INFER_SHAPE(B200FusionSeqpoolCvmConcatInferShape))
    .SetInferDtypeFn(PD_INFER_DTYPE(B200FusionSeqpoolCvmConcatInferDtype));

// ---------------------------------------------------------------------------------------
// b200_fused_seqpool_cvm  (FusedSeqpoolCVMOpMaker, fused/fused_seqpool_cvm_op.cc:120-150): PaddleBox's
// multi-slot op over pre-pulled values.  X (duplicable, LoD rows), CVM -> Out (duplicable);
// pooltype in {SUM}; the accumulator STARTS at pad_value (fused_seqpool_cvm_op.cu:49);
// use_cvm=false drops cvm_offset leading columns.  One launch for all slots.
// ---------------------------------------------------------------------------------------
std::vector<paddle::Tensor> B200FusedSeqpoolCvm(const std::vector<paddle::Tensor>& xs, const paddle::Tensor& cvm,
                                                const std::string& pooltype, float pad_value, bool use_cvm,
                                                int cvm_offset) {
  PD_CHECK(pooltype == "SUM", "fused_seqpool_cvm only supports pooltype SUM");
  PD_CHECK(!xs.empty(), "Input(X) of fused_seqpool_cvm is empty");
  NeedGpu(xs[0], "X");
  const int64_t w = xs[0].numel() / xs[0].shape()[0];
  const int64_t d_out = use_cvm ? w : w - cvm_offset;
  std::vector<ctrfe_slot> slots(xs.size());
  std::vector<paddle::Tensor> outs;
  outs.reserve(xs.size());
  for (size_t s = 0; s < xs.size(); ++s) {
    const auto& lod = LodOf(xs[s]);
    PD_CHECK(!lod.empty(), "Input(X) of fused_seqpool_cvm must carry LoD");
    PD_CHECK(xs[s].numel() / xs[s].shape()[0] == w, "Width of all inputs should be equal.");
    const auto& last = lod.back();
    const int64_t n_seg = static_cast<int64_t>(last.size()) - 1;
    outs.emplace_back(paddle::empty({n_seg, d_out}, xs[s].dtype(), xs[s].place()));
    ctrfe_slot& sl = slots[s];
    sl.table = Raw(xs[s]);
    sl.table_rows = xs[s].shape()[0];
    sl.ids = nullptr;  // dense rows: row index = position
    sl.n_ids = 0;
    sl.lod = reinterpret_cast<const uint64_t*>(last.data());
    sl.n_seg = n_seg;
    sl.out = RawMut(outs.back());
    sl.out_ld = d_out;
    sl.pooltype = CTRFE_POOL_SUM;
    sl.ids_dtype = CTRFE_I64;
  }
  ctrfe_fused_desc d{};
  d.n_slots = static_cast<int32_t>(slots.size());
  d.dtype = DtypeCode(xs[0].dtype());
  d.slots = slots.data();
  d.dim = w;
  d.padding_idx = -1;
  d.pad_value = pad_value;
  d.cvm_mode = use_cvm ? CTRFE_CVM_USE : CTRFE_CVM_STRIP;
  d.cvm_offset = cvm_offset;
  d.scale_mode = CTRFE_SCALE_DIVIDE;
  d.init_mode = CTRFE_INIT_PAD;
  d.lod_location = CTRFE_LOD_HOST;
  Check(ctrfe_fused_forward(CtxFor(xs[0]), &d, xs[0].stream()));
  return outs;
}

std::vector<std::vector<int64_t>> B200FusedSeqpoolCvmInferShape(const std::vector<std::vector<int64_t>>& x_shapes,
                                                                const std::vector<int64_t>& cvm_shape,
                                                                const std::string& pooltype, float pad_value,
                                                                bool use_cvm, int cvm_offset) {
  std::vector<std::vector<int64_t>> out;  // batch comes from the LoD at run time (fused_seqpool_cvm_op.cc:82-91)
  for (const auto& xs : x_shapes) out.push_back({-1, use_cvm ? xs[1] : xs[1] - cvm_offset});
  return out;
}

std::vector<paddle::DataType> B200FusedSeqpoolCvmInferDtype(const std::vector<paddle::DataType>& x_dtypes,
                                                            const paddle::DataType& cvm_dtype) {
  return x_dtypes;
}

PD_BUILD_OP(b200_fused_seqpool_cvm)
    .Inputs({paddle::Vec("X"), "CVM"})
    .Outputs({paddle::Vec("Out")})
    .Attrs({"pooltype: std::string", "pad_value: float", "use_cvm: bool", "cvm_offset: int"})
    .SetKernelFn(PD_KERNEL(B200FusedSeqpoolCvm))
    .SetInferShapeFn(PD_INFER_SHAPE(B200FusedSeqpoolCvmInferShape))
    .SetInferDtypeFn(PD_INFER_DTYPE(B200FusedSeqpoolCvmInferDtype));

// ---------------------------------------------------------------------------------------
// b200_fused_embedding_seq_pool  (FusedEmbeddingSeqPoolOpMaker, fused/fused_embedding_seq_pool_op.cc:
// 82-130): W, Ids [K, idx_width] int64 with one LoD level -> Out [num_seq, D * idx_width]; combiner
// "sum"; the EmbSeqPool order (first row copied, the rest added in id order,
// jit/refer/refer.h:477-528).  Column j of Ids is pooled into columns [j*D, (j+1)*D) of Out.
// ---------------------------------------------------------------------------------------
std::vector<paddle::Tensor> B200FusedEmbeddingSeqPool(const paddle::Tensor& w, const paddle::Tensor& ids,
                                                      const std::string& combiner, int64_t padding_idx,
                                                      bool grad_inplace, bool is_sparse) {
  PD_CHECK(combiner == "sum", "fused_embedding_seq_pool only supports combiner sum");
  NeedGpu(w, "W");
  NeedGpu(ids, "Ids");
  PD_CHECK(ids.dtype() == paddle::DataType::INT64, "fused_embedding_seq_pool: Ids must be int64");
  const auto wshape = w.shape();
  const auto ishape = ids.shape();
  PD_CHECK(wshape.size() == 2 && ishape.size() == 2, "The dimension of W and Ids should be 2.");
  const auto& lod = LodOf(ids);
  PD_CHECK(lod.size() == 1, "The LoD level of Input(Ids) should be 1.");
  const int64_t n_seg = static_cast<int64_t>(lod[0].size()) - 1;
  const int64_t width = ishape[1], dim = wshape[1];
  auto out = paddle::empty({n_seg, dim * width}, w.dtype(), w.place());
  Check(ctrfe_fused_embedding_seq_pool(CtxFor(w), Raw(w), wshape[0], dim, DtypeCode(w.dtype()), static_cast<const int64_t*>(Raw(ids)), ishape[0], width,
                                       reinterpret_cast<const uint64_t*>(lod[0].data()), n_seg, padding_idx,
                                       RawMut(out), w.stream()));
  Check(ctrfe_check_ids(CtxFor(w), w.stream(), nullptr, nullptr, nullptr));
  return {out};
}

std::vector<std::vector<int64_t>> B200FusedEmbeddingSeqPoolInferShape(const std::vector<int64_t>& w_shape,
                                                                      const std::vector<int64_t>& ids_shape,
                                                                      const std::string& combiner, int64_t padding_idx,
                                                                      bool grad_inplace, bool is_sparse) {
  return {{-1, w_shape[1] * ids_shape[1]}};  // fused_embedding_seq_pool_op.cc:62-71
}

std::vector<paddle::DataType> B200FusedEmbeddingSeqPoolInferDtype(const paddle::DataType& w_dtype,
                                                                  const paddle::DataType& ids_dtype) {
  return {w_dtype};
}

PD_BUILD_OP(b200_fused_embedding_seq_pool)
    .Inputs({"W", "Ids"})
    .Outputs({"Out"})
    .Attrs({"combiner: std::string", "padding_idx: int64_t", "grad_inplace: bool", "is_sparse: bool"})
    .SetKernelFn(PD_KERNEL(B200FusedEmbeddingSeqPool))
    .SetInferShapeFn(PD_INFER_SHAPE(B200FusedEmbeddingSeqPoolInferShape))
    .SetInferDtypeFn(PD_INFER_DTYPE(B200FusedEmbeddingSeqPoolInferDtype));

// ---------------------------------------------------------------------------------------
// b200_pull_box_sparse  (PullBoxSparseOpMaker, pull_box_sparse_op.cc:62-100): Ids (duplicable,
// uint64 feasign keys held as int64) -> Out (duplicable, [K_s, size]).  The reference asks BoxPS for
// the values; here the table is resident: W [n_keys, size] is the dense image of the pull values and
// MapKeys / MapRows the key map built by ctrfe_keymap_build (two extra inputs the reference op does
// not have -- its table lives behind the BoxPS handle).  Key 0 and unknown keys give zero rows.
// ---------------------------------------------------------------------------------------
std::vector<paddle::Tensor> B200PullBoxSparse(const paddle::Tensor& w, const paddle::Tensor& map_keys,
                                              const paddle::Tensor& map_rows, const std::vector<paddle::Tensor>& ids,
                                              int size, bool is_sparse, bool is_distributed) {
  NeedGpu(w, "W");
  const auto wshape = w.shape();
  PD_CHECK(wshape.size() == 2 && wshape[1] == size, "pull_box_sparse: W must be [n_keys, size]");
  const int64_t n_slots_map = map_keys.numel();
  PD_CHECK(n_slots_map == map_rows.numel() && n_slots_map >= 16 && (n_slots_map & (n_slots_map - 1)) == 0,
           "pull_box_sparse: MapKeys / MapRows must have the same power-of-two length");
  int log2 = 0;
  while ((int64_t{1} << log2) < n_slots_map) ++log2;
  ctrfe_keymap km{};
  km.keys = static_cast<const uint64_t*>(Raw(map_keys));
  km.rows = static_cast<const uint32_t*>(Raw(map_rows));
  km.log2_slots = log2;
  km.n_keys = wshape[0];
  std::vector<ctrfe_slot> slots(ids.size());
  std::vector<std::vector<uint64_t>> unit(ids.size());  // one key per "sequence": FIRST pool = plain pull
  std::vector<paddle::Tensor> outs;
  outs.reserve(ids.size());
  for (size_t s = 0; s < ids.size(); ++s) {
    NeedGpu(ids[s], "Ids");
    PD_CHECK(ids[s].dtype() == paddle::DataType::INT64, "pull_box_sparse: Ids must be int64 (uint64 feasigns)");
    const int64_t k = ids[s].numel();
    unit[s].resize(static_cast<size_t>(k) + 1);
    for (int64_t i = 0; i <= k; ++i) unit[s][i] = static_cast<uint64_t>(i);
    outs.emplace_back(paddle::empty({k, static_cast<int64_t>(size)}, w.dtype(), w.place()));
    ctrfe_slot& sl = slots[s];
    sl.table = Raw(w);
    sl.table_rows = wshape[0];
    sl.ids = Raw(ids[s]);
    sl.n_ids = k;
    sl.lod = unit[s].data();
    sl.n_seg = k;
    sl.out = RawMut(outs.back());
    sl.out_ld = size;
    sl.pooltype = CTRFE_POOL_FIRST;
    sl.ids_dtype = CTRFE_I64;
  }
  ctrfe_fused_desc d{};
  d.n_slots = static_cast<int32_t>(slots.size());
  d.dtype = DtypeCode(w.dtype());
  d.slots = slots.data();
  d.dim = size;
  d.padding_idx = -1;
  d.cvm_mode = CTRFE_CVM_NONE;
  d.cvm_offset = 2;
  d.scale_mode = CTRFE_SCALE_DIVIDE;
  d.init_mode = CTRFE_INIT_ZERO;
  d.lod_location = CTRFE_LOD_HOST;
  d.keymap = &km;
  Check(ctrfe_fused_forward(CtxFor(w), &d, w.stream()));
  return outs;
}

std::vector<std::vector<int64_t>> B200PullBoxSparseInferShape(const std::vector<int64_t>& w_shape,
                                                              const std::vector<int64_t>& mk_shape,
                                                              const std::vector<int64_t>& mr_shape,
                                                              const std::vector<std::vector<int64_t>>& ids_shapes, int size,
                                                              bool is_sparse, bool is_distributed) {
  std::vector<std::vector<int64_t>> out;  // Ids dims with the trailing 1 replaced by size (pull_box_sparse_op.cc:40-52)
  for (const auto& is : ids_shapes) {
    auto o = is;
    if (!o.empty() && o.back() == 1) o.back() = size; else o.push_back(size);
    out.push_back(o);
  }
  return out;
}

std::vector<paddle::DataType> B200PullBoxSparseInferDtype(const paddle::DataType& w_dtype, const paddle::DataType& mk,
                                                          const paddle::DataType& mr,
                                                          const std::vector<paddle::DataType>& ids_dtypes) {
  return std::vector<paddle::DataType>(ids_dtypes.size(), w_dtype);
}

PD_BUILD_OP(b200_pull_box_sparse)
    .Inputs({"W", "MapKeys", "MapRows", paddle::Vec("Ids")})
    .Outputs({paddle::Vec("Out")})
    .Attrs({"size: int", "is_sparse: bool", "is_distributed: bool"})
    .SetKernelFn(PD_KERNEL(B200PullBoxSparse))
    .SetInferShapeFn(PD_INFER_SHAPE(B200PullBoxSparseInferShape))
    .SetInferDtypeFn(PD_INFER_DTYPE(B200PullBoxSparseInferDtype));

// ---------------------------------------------------------------------------------------
// b200_lookup_table_v2  (LookupTableV2OpMaker, lookup_table_v2_op.cc:69-104)
// ---------------------------------------------------------------------------------------
std::vector<paddle::Tensor> B200LookupTableV2(const paddle::Tensor& w, const paddle::Tensor& ids,
                                              int64_t padding_idx) {
  NeedGpu(w, "W");
  NeedGpu(ids, "Ids");
  const auto wshape = w.shape();
  PD_CHECK(wshape.size() == 2, "ShapeError: The dimensions of the 'lookup table' must be 2.");
  auto oshape = ids.shape();
  oshape.push_back(wshape[1]);
  auto out = paddle::empty(oshape, w.dtype(), w.place());
  const int64_t es = static_cast<int64_t>(paddle::experimental::SizeOf(w.dtype()));
  ctrfe_ctx* ctx = CtxFor(w);
  Check(ctrfe_embedding(ctx, Raw(w), wshape[0], wshape[1] * es, Raw(ids), IdCode(ids.dtype()), ids.numel(),
                        padding_idx, RawMut(out), w.stream()));
  // the reference CPU kernel throws on an out-of-range id; keep that behaviour
  int64_t slot = -1, pos = -1, bad = -1;
  Check(ctrfe_check_ids(ctx, w.stream(), &slot, &pos, &bad));
  return {out};
}

std::vector<std::vector<int64_t>> B200LookupTableV2InferShape(const std::vector<int64_t>& w_shape,
                                                              const std::vector<int64_t>& ids_shape,
                                                              int64_t padding_idx) {
  auto o = ids_shape;
  o.push_back(w_shape[1]);
  return {o};
}

std::vector<paddle::DataType> B200LookupTableV2InferDtype(const paddle::DataType& w_dtype,
                                                          const paddle::DataType& ids_dtype) {
  return {w_dtype};
}

PD_BUILD_OP(b200_lookup_table_v2)
    .Inputs({"W", "Ids"})
    .Outputs({"Out"})
    .Attrs({"padding_idx: int64_t"})
    .SetKernelFn(PD_KERNEL(B200LookupTableV2))
    .SetInferShapeFn(PD_INFER_SHAPE(B200LookupTableV2InferShape))
    .SetInferDtypeFn(PD_INFER_DTYPE(B200LookupTableV2InferDtype));

// W@GRAD (dense) of b200_lookup_table_v2: LookupTableV2GradOpMaker takes W, Ids, Out@GRAD
// (lookup_table_v2_op.cc:107-125); fp32 only here, SelectedRows (is_sparse) not provided.
std::vector<paddle::Tensor> B200LookupTableV2Grad(const paddle::Tensor& w, const paddle::Tensor& ids,
                                                  const paddle::Tensor& dout, int64_t padding_idx) {
  NeedGpu(dout, "Out@GRAD");
  PD_CHECK(dout.dtype() == paddle::DataType::FLOAT32, "b200_lookup_table_v2_grad: float32 only");
  const auto wshape = w.shape();
  auto dw = paddle::empty(wshape, dout.dtype(), dout.place());
  ctrfe_ctx* ctx = CtxFor(dout);
  Check(ctrfe_embedding_grad(ctx, static_cast<const float*>(Raw(dout)), Raw(ids), IdCode(ids.dtype()), ids.numel(),
                             wshape[0], wshape[1], padding_idx, static_cast<float*>(RawMut(dw)), dout.stream()));
  Check(ctrfe_check_ids(ctx, dout.stream(), nullptr, nullptr, nullptr));
  return {dw};
}

PD_BUILD_GRAD_OP(b200_lookup_table_v2)
    .Inputs({"W", "Ids", paddle::Grad("Out")})
    .Outputs({paddle::Grad("W")})
    .Attrs({"padding_idx: int64_t"})
    .SetKernelFn(PD_KERNEL(B200LookupTableV2Grad));

// ---------------------------------------------------------------------------------------
// b200_sequence_pool  (SequencePoolOpMaker, sequence_pool_op.cc:45-112).  The custom-op
// maker has no "intermediate" marker, so MaxIndex is a plain second output: int32 indices
// for MAX when !is_test, else a 0-size tensor.
// ---------------------------------------------------------------------------------------
std::vector<paddle::Tensor> B200SequencePool(const paddle::Tensor& x, const std::string& pooltype, float pad_value,
                                             bool is_test) {
  NeedGpu(x, "X");
  const auto& lod = LodOf(x);
  PD_CHECK(!lod.empty(), "Input(X) Tensor of SequencePoolOp does not contain LoD information.");
  PD_CHECK(lod.size() <= 2, "The lod level of input shall be no more than 2.");
  if (lod.size() == 2) PD_CHECK(lod[0].back() == lod[1].size() - 1, "The input lod information is illegal.");
  const auto& last = lod.back();
  const int64_t n_seg = static_cast<int64_t>(last.size()) - 1;
  auto shape = x.shape();
  PD_CHECK(shape[0] >= n_seg, "The first dimension of Input(X) must be large than batch size.");
  int64_t width = 1;
  for (size_t i = 1; i < shape.size(); ++i) width *= shape[i];
  shape[0] = n_seg;
  auto out = paddle::empty(shape, x.dtype(), x.place());
  const bool want_index = pooltype == "MAX" && !is_test;
  auto index = want_index ? paddle::empty(shape, paddle::DataType::INT32, x.place())
                          : paddle::empty({0}, paddle::DataType::INT32, x.place());
  Check(ctrfe_sequence_pool(CtxFor(x), Raw(x), DtypeCode(x.dtype()), width,
                            reinterpret_cast<const uint64_t*>(last.data()), n_seg, PoolCode(pooltype), pad_value,
                            RawMut(out), want_index ? index.data<int32_t>() : nullptr, x.stream()));
  return {out, index};
}

std::vector<std::vector<int64_t>> B200SequencePoolInferShape(const std::vector<int64_t>& x_shape,
                                                             const std::string& pooltype, float pad_value,
                                                             bool is_test) {
  auto o = x_shape;
  o[0] = -1;
  return {o, o};
}

std::vector<paddle::DataType> B200SequencePoolInferDtype(const paddle::DataType& x_dtype) {
  return {x_dtype, paddle::DataType::INT32};
}

PD_BUILD_OP(b200_sequence_pool)
    .Inputs({"X"})
    .Outputs({"Out", "MaxIndex"})
    .Attrs({"pooltype: std::string", "pad_value: float", "is_test: bool"})
    .SetKernelFn(PD_KERNEL(B200SequencePool))
    .SetInferShapeFn(PD_INFER_SHAPE(B200SequencePoolInferShape))
    .SetInferDtypeFn(PD_INFER_DTYPE(B200SequencePoolInferDtype));

// X@GRAD of b200_sequence_pool (SequencePoolGradOpMaker: X, MaxIndex, Out@GRAD,
// sequence_pool_op.cc:150-176); fp32 only.
std::vector<paddle::Tensor> B200SequencePoolGrad(const paddle::Tensor& x, const paddle::Tensor& max_index,
                                                 const paddle::Tensor& dout, const std::string& pooltype,
                                                 float pad_value, bool is_test) {
  NeedGpu(dout, "Out@GRAD");
  PD_CHECK(dout.dtype() == paddle::DataType::FLOAT32, "b200_sequence_pool_grad: float32 only");
  const auto& lod = LodOf(x);
  PD_CHECK(!lod.empty(), "Input(X) Tensor of SequencePoolGradOp does not contain LoD information.");
  const auto& last = lod.back();
  const int64_t n_seg = static_cast<int64_t>(last.size()) - 1;
  const auto shape = x.shape();
  int64_t width = 1;
  for (size_t i = 1; i < shape.size(); ++i) width *= shape[i];
  auto dx = paddle::empty(shape, dout.dtype(), dout.place());
  const bool is_max = pooltype == "MAX";
  Check(ctrfe_sequence_pool_grad(CtxFor(dout), static_cast<const float*>(Raw(dout)), width,
                                 reinterpret_cast<const uint64_t*>(last.data()), n_seg, shape[0], PoolCode(pooltype),
                                 is_max ? max_index.data<int32_t>() : nullptr, static_cast<float*>(RawMut(dx)),
                                 dout.stream()));
  return {dx};
}

PD_BUILD_GRAD_OP(b200_sequence_pool)
    .Inputs({"X", "MaxIndex", paddle::Grad("Out")})
    .Outputs({paddle::Grad("X")})
    .Attrs({"pooltype: std::string", "pad_value: float", "is_test: bool"})
    .SetKernelFn(PD_KERNEL(B200SequencePoolGrad));

// ---------------------------------------------------------------------------------------
// b200_cvm  (CVMOpMaker, cvm_op.cc:127-150).  The CVM input is declared because the
// reference op has it; forward never reads it (cvm_op.h:60-99).
// ---------------------------------------------------------------------------------------
std::vector<paddle::Tensor> B200Cvm(const paddle::Tensor& x, const paddle::Tensor& cvm, bool use_cvm) {
  NeedGpu(x, "X");
  const auto shape = x.shape();
  PD_CHECK(shape.size() == 2, "Input(X)'s rank should be 2, but got ", shape.size());
  auto y = paddle::empty({shape[0], use_cvm ? shape[1] : shape[1] - 2}, x.dtype(), x.place());
  Check(ctrfe_cvm(CtxFor(x), Raw(x), DtypeCode(x.dtype()), shape[0], shape[1], use_cvm ? 1 : 0, RawMut(y),
                  x.stream()));
  return {y};
}

std::vector<std::vector<int64_t>> B200CvmInferShape(const std::vector<int64_t>& x_shape,
                                                    const std::vector<int64_t>& cvm_shape, bool use_cvm) {
  return {{x_shape[0], use_cvm ? x_shape[1] : x_shape[1] - 2}};
}

std::vector<paddle::DataType> B200CvmInferDtype(const paddle::DataType& x_dtype, const paddle::DataType& cvm_dtype) {
  return {x_dtype};
}

PD_BUILD_OP(b200_cvm)
    .Inputs({"X", "CVM"})
    .Outputs({"Y"})
    .Attrs({"use_cvm: bool"})
    .SetKernelFn(PD_KERNEL(B200Cvm))
    .SetInferShapeFn(PD_INFER_SHAPE(B200CvmInferShape))
    .SetInferDtypeFn(PD_INFER_DTYPE(B200CvmInferDtype));

// X@GRAD of b200_cvm (CVMGradOpMaker: X, CVM, Y@GRAD, cvm_op.cc:152-170): columns 0-1 take
// the CVM input of the row's sequence, the rest Y@GRAD (cvm_op.h:100-140); fp32 only.
std::vector<paddle::Tensor> B200CvmGrad(const paddle::Tensor& x, const paddle::Tensor& cvm, const paddle::Tensor& dy,
                                        bool use_cvm) {
  NeedGpu(dy, "Y@GRAD");
  PD_CHECK(dy.dtype() == paddle::DataType::FLOAT32, "b200_cvm_grad: float32 only");
  const auto shape = x.shape();
  auto dx = paddle::empty(shape, dy.dtype(), dy.place());
  const auto& lod = LodOf(x);
  const uint64_t* lod0 = lod.empty() ? nullptr : reinterpret_cast<const uint64_t*>(lod[0].data());
  const int64_t n_seq = lod.empty() ? cvm.shape()[0] : static_cast<int64_t>(lod[0].size()) - 1;
  Check(ctrfe_cvm_grad(CtxFor(dy), static_cast<const float*>(Raw(dy)), static_cast<const float*>(Raw(cvm)), lod0, n_seq,
                       shape[0], shape[1], use_cvm ? 1 : 0, static_cast<float*>(RawMut(dx)), dy.stream()));
  return {dx};
}

PD_BUILD_GRAD_OP(b200_cvm)
    .Inputs({"X", "CVM", paddle::Grad("Y")})
    .Outputs({paddle::Grad("X")})
    .Attrs({"use_cvm: bool"})
    .SetKernelFn(PD_KERNEL(B200CvmGrad));
