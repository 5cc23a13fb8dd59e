// Host/device interface of the cooperative fp32 decode kernel for one or two sequences (decode_f32_persistent.cu).
#pragma once
#include "kernels.h"

struct PersistF32Layer {
  const float *ln1_g, *ln1_b, *ln2_g, *ln2_b;
  const float *w_qkv, *b_qkv;      // [3d, d], [3d]
  const float *w_proj, *b_proj;    // [d, d]
  const float *w_fc, *b_fc;        // [4d, d]
  const float *w_proj2, *b_proj2;  // [d, 4d]
  float *kc, *vc;                  // this layer's K / V cache, [B][H][Smax][64] fp32
};

struct PersistF32Params {
  int L, Smax;
  int nsplit;                      // key ranges per (sequence, head) in the attention phase
  const PersistF32Layer* layers;   // [L], device
  const int32_t* ids;
  int ids_ld;
  const int* len;
  const int* npad;
  const float* wte;
  const float* wpe;
  float* x;            // [B, d] residual stream (in/out: the LM head reads it)
  float* q;            // [B, d]
  float* att;          // [B, d]
  float* h;            // [B, 4d]
  float* att_scratch;  // [B*H][nsplit][66] partial softmax states
  int* counters;       // [2 + B*H], zero at launch and again at exit: barrier arrivals, finished CTAs, per-head arrivals
  int* err;            // [8], zero unless a barrier wait timed out
  float eps;
  long long* trace;    // null, or [1 + 4 * barriers] clock64 stamps of thread 0 of CTA trace_cta (tutor_gpt2_decode_trace)
  int trace_cta;
};

bool k_decode_f32_persistent_supported(int B, int d, int ncta);
int k_decode_f32_persistent(const PersistF32Params& p, int B, int d, int ncta, cudaStream_t st);
