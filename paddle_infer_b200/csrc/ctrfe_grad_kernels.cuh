// ctrfe_grad_kernels.cuh -- backward of the three ops, fp32.
// Inference never runs these; they exist so that a training-time program that was rewritten to
// the b200_* ops still has its gradient ops.  Plain HBM-bound kernels, no fusion.
#pragma once

#include <stdint.h>

#include "../../include/ctrfe.h"

namespace ctrfe {

// sequence_pool_grad (math/sequence_pooling.cc:438-496, :179-239, :309-343): one warp per
// segment, lanes stride the columns, rows written with coalesced stores.  dx has been zeroed by
// the caller for LAST / FIRST / MAX (and when rows lie outside every segment).
__global__ void seqpool_grad_kernel(const float* __restrict__ dout, const unsigned long long* __restrict__ lod,
                                    long long n_seg, int w, int pooltype, const int* __restrict__ max_index,
                                    float* __restrict__ dx) {
  const long long seg = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  if (seg >= n_seg) return;
  const int lane = threadIdx.x & 31;
  const long long a = static_cast<long long>(lod[seg]), e = static_cast<long long>(lod[seg + 1]);
  if (a == e) return;
  const float h = static_cast<float>(e - a);
  for (int c = lane; c < w; c += 32) {
    const float g = dout[seg * w + c];
    switch (pooltype) {
      case CTRFE_POOL_SUM:
        for (long long r = a; r < e; ++r) dx[r * w + c] = g;
        break;
      case CTRFE_POOL_AVERAGE: {
        const float v = g / h;
        for (long long r = a; r < e; ++r) dx[r * w + c] = v;
        break;
      }
      case CTRFE_POOL_SQRT: {
        const float v = g / sqrtf(h);
        for (long long r = a; r < e; ++r) dx[r * w + c] = v;
        break;
      }
      case CTRFE_POOL_LAST: dx[(e - 1) * w + c] = g; break;
      case CTRFE_POOL_FIRST: dx[a * w + c] = g; break;
      default: {  // MAX
        const int idx = max_index[seg * w + c];
        if (idx != -1) dx[static_cast<long long>(idx) * w + c] = g;
      }
    }
  }
}

// cvm_grad (cvm_op.h:43-58,100-140; cvm_op.cu:49-84): dx[r, 0:2] = cvm[seq(r)], dx[r, 2:] = the
// tail of dy.  seq(r) = r without LoD, else the sequence that holds row r (binary search).
__global__ void cvm_grad_kernel(const float* __restrict__ dy, const float* __restrict__ cvm,
                                const unsigned long long* __restrict__ lod, long long n_seq, long long rows, int w,
                                int use_cvm, float* __restrict__ dx) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= rows * w) return;
  const long long r = i / w;
  const int c = static_cast<int>(i - r * w);
  const int off = use_cvm ? 0 : 2;
  if (c >= 2) {
    dx[i] = dy[r * (w - off) + c - off];
    return;
  }
  long long seq = r;
  if (lod) {  // largest s with lod[s] <= r
    long long lo = 0, hi = n_seq;
    while (hi - lo > 1) {
      const long long mid = (lo + hi) >> 1;
      if (static_cast<long long>(lod[mid]) <= r) lo = mid; else hi = mid;
    }
    seq = lo;
  }
  dx[i] = cvm[seq * 2 + c];
}

// lookup_table_v2_grad, dense W@GRAD (phi/kernels/gpu/embedding_grad_kernel.cu:38-60 with the
// CPU kernel's padding_idx rule, cpu/embedding_grad_kernel.cc:63-66): dw (zeroed by the caller)
// += dout rows, one atomic add per element.  A bad id is flagged, not followed.
__global__ void embedding_grad_kernel(const float* __restrict__ dout, const void* __restrict__ ids, int ids_i32,
                                      long long k, long long n_rows, int d, long long padding_idx, float* __restrict__ dw,
                                      unsigned long long* err) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= k * d) return;
  const long long r = i / d;
  const int c = static_cast<int>(i - r * d);
  const long long id = ids_i32 ? static_cast<long long>(static_cast<const int*>(ids)[r]) : static_cast<const long long*>(ids)[r];
  if (padding_idx != -1 && id == padding_idx) return;
  if (id < 0 || id >= n_rows) {
    if (c == 0) atomicMin(err, static_cast<unsigned long long>(r));
    return;
  }
  atomicAdd(dw + id * d + c, dout[i]);
}

}  // namespace ctrfe
