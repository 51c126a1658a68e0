/* A plain-C caller of libctrfe.so: the drop-in boundary needs nothing but include/ctrfe.h, the
 * CUDA runtime for device memory, and a C compiler.  Builds a small CTR batch (3 slots, ragged
 * sequences with empty ones, SUM / AVERAGE / MAX), runs it through ctrfe_fused_forward with host
 * LoD offsets, and checks the result against loops written out here in the reference's order
 * (jit/refer/refer.h:372-380 sum from 0 in row order; sequence_pooling.cc:166-174 MAX; empty ->
 * pad_value).  Also checks the error path: an id outside the table is reported with its slot
 * and position.  Exit code 0 = pass.  Driven by tests/test_c_abi.py. */
#include <cuda_runtime_api.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "ctrfe.h"

#define CK(call)                                                                 \
  do {                                                                           \
    cudaError_t e_ = (call);                                                     \
    if (e_ != cudaSuccess) {                                                     \
      fprintf(stderr, "cuda: %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
      return 2;                                                                  \
    }                                                                            \
  } while (0)
#define RC(call)                                                                           \
  do {                                                                                     \
    int rc_ = (call);                                                                      \
    if (rc_ != CTRFE_OK) {                                                                 \
      fprintf(stderr, "ctrfe rc=%d: %s at %s:%d\n", rc_, ctrfe_last_error(), __FILE__, __LINE__); \
      return 3;                                                                            \
    }                                                                                      \
  } while (0)

enum { N = 1000, D = 16, S = 3, B = 40 };

static uint32_t rng_state = 12345u;
static uint32_t rnd(void) { return rng_state = rng_state * 1664525u + 1013904223u; }

int main(void) {
  static float table[N * D];
  for (int i = 0; i < N * D; ++i) table[i] = (float)((int)(rnd() >> 8) % 2001 - 1000) / 1000.0f;
  const int pooltypes[S] = {CTRFE_POOL_SUM, CTRFE_POOL_AVERAGE, CTRFE_POOL_MAX};
  static uint64_t lod[S][B + 1];
  static int64_t ids[S][B * 8];
  for (int s = 0; s < S; ++s) {
    lod[s][0] = 0;
    for (int b = 0; b < B; ++b) {
      int len = (int)(rnd() >> 16) % 8;
      if (b % 7 == 3) len = 0; /* empty sequences take pad_value */
      lod[s][b + 1] = lod[s][b] + (uint64_t)len;
    }
    for (uint64_t k = 0; k < lod[s][B]; ++k) ids[s][k] = (int64_t)((rnd() >> 8) % N);
  }
  const float pad_value = 0.25f;

  /* expected, in the reference's order */
  static float want[S][B][D];
  for (int s = 0; s < S; ++s)
    for (int b = 0; b < B; ++b) {
      const uint64_t a = lod[s][b], e = lod[s][b + 1];
      for (int c = 0; c < D; ++c) {
        float acc = 0.0f;
        if (a == e) {
          acc = pad_value;
        } else if (pooltypes[s] == CTRFE_POOL_MAX) {
          acc = table[ids[s][a] * D + c];
          for (uint64_t k = a + 1; k < e; ++k)
            if (table[ids[s][k] * D + c] > acc) acc = table[ids[s][k] * D + c];
        } else {
          for (uint64_t k = a; k < e; ++k) acc = acc + table[ids[s][k] * D + c];
          if (pooltypes[s] == CTRFE_POOL_AVERAGE) acc = acc / (float)(e - a);
        }
        want[s][b][c] = acc;
      }
    }

  if (ctrfe_version() < 100) return 4;
  float* d_table;
  int64_t* d_ids[S];
  float* d_out[S];
  CK(cudaSetDevice(0));
  CK(cudaMalloc((void**)&d_table, sizeof(table)));
  CK(cudaMemcpy(d_table, table, sizeof(table), cudaMemcpyHostToDevice));
  ctrfe_slot slots[S];
  memset(slots, 0, sizeof(slots));
  for (int s = 0; s < S; ++s) {
    CK(cudaMalloc((void**)&d_ids[s], sizeof(ids[s])));
    CK(cudaMemcpy(d_ids[s], ids[s], sizeof(ids[s]), cudaMemcpyHostToDevice));
    CK(cudaMalloc((void**)&d_out[s], sizeof(float) * B * D));
    slots[s].table = d_table;
    slots[s].table_rows = N;
    slots[s].ids = d_ids[s];
    slots[s].lod = lod[s];
    slots[s].n_seg = B;
    slots[s].out = d_out[s];
    slots[s].out_ld = D;
    slots[s].pooltype = pooltypes[s];
    slots[s].ids_dtype = CTRFE_I64;
    slots[s].n_ids = (int64_t)(sizeof(ids[s]) / sizeof(ids[s][0]));
  }
  ctrfe_fused_desc desc;
  memset(&desc, 0, sizeof(desc));
  desc.n_slots = S;
  desc.dtype = CTRFE_F32;
  desc.slots = slots;
  desc.dim = D;
  desc.padding_idx = -1;
  desc.pad_value = pad_value;
  desc.cvm_mode = CTRFE_CVM_NONE;
  desc.cvm_offset = 2;
  desc.scale_mode = CTRFE_SCALE_DIVIDE;
  desc.init_mode = CTRFE_INIT_ZERO;
  desc.lod_location = CTRFE_LOD_HOST;

  ctrfe_ctx* ctx = NULL;
  RC(ctrfe_ctx_create(0, &ctx));
  cudaStream_t st;
  CK(cudaStreamCreate(&st));
  RC(ctrfe_fused_forward(ctx, &desc, st));
  int64_t bad_slot = -1, bad_pos = -1, bad_id = -1;
  RC(ctrfe_check_ids(ctx, st, &bad_slot, &bad_pos, &bad_id));

  static float got[B * D];
  int mism = 0;
  for (int s = 0; s < S; ++s) {
    CK(cudaMemcpy(got, d_out[s], sizeof(got), cudaMemcpyDeviceToHost));
    for (int b = 0; b < B; ++b)
      for (int c = 0; c < D; ++c)
        if (memcmp(&got[b * D + c], &want[s][b][c], 4) != 0) {
          if (mism < 5) fprintf(stderr, "slot %d seq %d col %d: got %.9g want %.9g\n", s, b, c, got[b * D + c], want[s][b][c]);
          ++mism;
        }
  }
  if (mism) {
    fprintf(stderr, "%d mismatching elements\n", mism);
    return 5;
  }

  /* error path: one id outside the table */
  if (lod[1][B] > 2) {
    int64_t bad = N + 17;
    CK(cudaMemcpy(d_ids[1] + 2, &bad, 8, cudaMemcpyHostToDevice));
    RC(ctrfe_fused_forward(ctx, &desc, st));
    const int rc = ctrfe_check_ids(ctx, st, &bad_slot, &bad_pos, &bad_id);
    if (rc != CTRFE_ERR_BAD_ID || bad_slot != 1 || bad_pos != 2 || bad_id != N + 17) {
      fprintf(stderr, "bad id not reported: rc=%d slot=%lld pos=%lld id=%lld (%s)\n", rc, (long long)bad_slot,
              (long long)bad_pos, (long long)bad_id, ctrfe_last_error());
      return 6;
    }
  }
  /* argument errors come back as status codes, never as a crash */
  desc.dim = 0;
  if (ctrfe_fused_forward(ctx, &desc, st) == CTRFE_OK) return 7;
  ctrfe_ctx_destroy(ctx);
  printf("ctrfe C smoke: ok (%d slots x %d sequences, bit-exact)\n", S, B);
  return 0;
}
