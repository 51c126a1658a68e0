// ref_shim.cc -- extern "C" doorway onto the REFERENCE's own jit::refer kernels.
// TEST INFRASTRUCTURE ONLY (see oracle/ctr_oracle.c header).
//
// This TU #includes paddle/fluid/operators/jit/refer/refer.h straight from
// /root/reference (never copied into this repo) and instantiates
//   refer::SeqPool<T>     refer.h:370-390   (SUM / AVERAGE / SQRT over one [h,w] block)
//   refer::EmbSeqPool<T>  refer.h:477-528   (gather + SUM in id order)
// so the C restatement in ctr_oracle.c can be validated against the real code, and so
// bench.py can time the reference's own arithmetic (cpu_baseline.kind = "reference").
// Built by oracle/ref_build/Makefile into oracle/_ref/libref_jit.so.
#include <cstdint>

#include "paddle/fluid/operators/jit/refer/refer.h"

namespace jit = paddle::operators::jit;

extern "C" {

// type: 1 = kSum, 2 = kAvg, 3 = kSqrt   (SeqPoolType, kernel_base.h)
int ref_seqpool_f32(const float* x, float* y, int h, int w, int type) {
  try {
    jit::seq_pool_attr_t attr(w, static_cast<jit::SeqPoolType>(type), h);
    jit::refer::SeqPool<float>(x, y, &attr);
    return 0;
  } catch (...) { return 1; }
}

int ref_seqpool_f64(const double* x, double* y, int h, int w, int type) {
  try {
    jit::seq_pool_attr_t attr(w, static_cast<jit::SeqPoolType>(type), h);
    jit::refer::SeqPool<double>(x, y, &attr);
    return 0;
  } catch (...) { return 1; }
}

int ref_embseqpool_f32(const float* table, int64_t tbl_h, int64_t tbl_w, const int64_t* idx,
                       int64_t idx_h, int64_t idx_w, float* out) {
  try {
    jit::emb_seq_pool_attr_t attr(tbl_h, tbl_w, idx_h, idx_w, tbl_w * idx_w);
    jit::refer::EmbSeqPool<float>(table, idx, out, &attr);
    return 0;
  } catch (...) { return 1; }
}

// Whole-slot driver over a LoD, looping exactly as
// fusion_seqpool_cvm_concat_op.cc:140-156 does around the jit kernel (without the CVM
// lines, which are not part of refer.h): used to time the reference arithmetic.
int ref_seqpool_lod_f32(const float* x, const uint64_t* lod, int64_t n_seg, int w, int type,
                        float* y, int64_t y_ld) {
  try {
    jit::seq_pool_attr_t attr(w, static_cast<jit::SeqPoolType>(type));
    const float* src = x + lod[0] * w;
    for (int64_t j = 0; j < n_seg; ++j) {
      attr.h = static_cast<int>(lod[j + 1] - lod[j]);
      jit::refer::SeqPool<float>(src, y + j * y_ld, &attr);
      src += static_cast<int64_t>(attr.h) * attr.w;
    }
    return 0;
  } catch (...) { return 1; }
}

// Gather + SUM over a LoD, looping as EmbeddingVSumFunctor does
// (fused_embedding_seq_pool_op.h:116-126); segments must be non-empty, as there.
int ref_embseqpool_lod_f32(const float* table, int64_t tbl_h, int64_t tbl_w, const int64_t* ids,
                           const uint64_t* lod, int64_t n_seg, float* out, int64_t out_ld) {
  try {
    jit::emb_seq_pool_attr_t attr(tbl_h, tbl_w, 0, 1, tbl_w);
    for (int64_t i = 0; i < n_seg; ++i) {
      attr.index_height = static_cast<int64_t>(lod[i + 1] - lod[i]);
      jit::refer::EmbSeqPool<float>(table, ids + lod[i], out + i * out_ld, &attr);
    }
    return 0;
  } catch (...) { return 1; }
}

// The whole CPU path for one slot, as a CPU predictor runs it after
// seqpool_cvm_concat_fuse_pass with the gather folded in: per sequence the REFERENCE's
// refer::EmbSeqPool (gather + SUM in id order), an empty sequence -> pad_value
// (sequence_pooling.cc:395-398), then the two CVM lines of
// fusion_seqpool_cvm_concat_op.cc:152-153 (log in double, narrowed).  The loop and the
// CVM lines are this shim's; the arithmetic that dominates is the reference's.
// Threads split the sequences (OpenMP), each under FTZ/DAZ as naive_executor.cc:66 sets.
int ref_fused_lod_f32(const float* table, int64_t tbl_h, int64_t tbl_w, const int64_t* ids,
                      const uint64_t* lod, int64_t n_seg, float pad_value, int use_cvm, float* out,
                      int64_t out_ld, int n_threads) {
  int bad = 0;
  if (n_threads < 1) n_threads = 1;
#pragma omp parallel num_threads(n_threads) if (n_threads > 1)
  {
    const unsigned old_csr = __builtin_ia32_stmxcsr();
    __builtin_ia32_ldmxcsr(old_csr | 0x8040u);
    jit::emb_seq_pool_attr_t attr(tbl_h, tbl_w, 0, 1, tbl_w);
#pragma omp for schedule(static)
    for (int64_t i = 0; i < n_seg; ++i) {
      float* dst = out + i * out_ld;
      attr.index_height = static_cast<int64_t>(lod[i + 1] - lod[i]);
      if (attr.index_height == 0) {
        for (int64_t k = 0; k < tbl_w; ++k) dst[k] = pad_value;
      } else {
        try {
          jit::refer::EmbSeqPool<float>(table, ids + lod[i], dst, &attr);
        } catch (...) { bad = 1; continue; }
      }
      if (use_cvm) {
        dst[0] = static_cast<float>(std::log(static_cast<double>(dst[0] + 1)));
        dst[1] = static_cast<float>(std::log(static_cast<double>(dst[1] + 1)) - static_cast<double>(dst[0]));
      }
    }
    __builtin_ia32_ldmxcsr(old_csr);
  }
  return bad;
}

int ref_seqpool_type_sum() { return static_cast<int>(jit::SeqPoolType::kSum); }
int ref_seqpool_type_avg() { return static_cast<int>(jit::SeqPoolType::kAvg); }
int ref_seqpool_type_sqrt() { return static_cast<int>(jit::SeqPoolType::kSqrt); }

}  // extern "C"
