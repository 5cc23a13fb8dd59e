// Fused LM head + repetition penalty + greedy argmax + stopping bookkeeping: the logits tensor
// (HF/models/gpt2/modeling_gpt2.py:705, [B,1,50257] fp32) is never materialised.
//   phase 1: every CTA streams a slice of the tied wte [V,d] once, scores it against all batch rows,
//            applies RepetitionPenaltyLogitsProcessor (HF/generation/logits_process.py:404-410) from a
//            per-sequence seen-bitmap and keeps a running (max, lowest-index) pair  -> partials in scratch
//   phase 2: one small CTA reduces the partials, pads finished rows, appends the ids, updates the bitmap,
//            the finished mask and the device-side length (HF/generation/utils.py:2793-2805).
// fp32 / small-batch flavour here (SIMT, HBM-bound); the bf16 batched flavour is the tcgen05 GEMM with
// the EPI_ARGMAX epilogue in gemm_bf16_tc.cu.  Both feed the same phase 2.
#include <math_constants.h>

#include "kernels.h"

int k_gemm_bf16_argmax(const bf16* wte, int V, const bf16* X, int B, int K, const uint32_t* seen, int seen_words,
                       float penalty, float* part_val, int* part_idx, cudaStream_t st);

namespace {

constexpr int LM_BLOCKS = 148 * 4;
constexpr int LM_MAX_PARTS = 1024;

template <typename WT, int MB>
__global__ void __launch_bounds__(256) lmhead_argmax_simt_kernel(const float* __restrict__ x,
                                                                  const WT* __restrict__ wte, int b0, int B, int V,
                                                                  int d, const uint32_t* __restrict__ seen,
                                                                  int seen_words, float penalty,
                                                                  float* __restrict__ part_val,
                                                                  int* __restrict__ part_idx,
                                                                  const float* __restrict__ ln_gamma,
                                                                  const float* __restrict__ ln_beta, float ln_eps,
                                                                  float* __restrict__ xn_out) {
  extern __shared__ __align__(16) float xs[];  // [MB][d] input rows, normalised here when ln_gamma is given (ln_f)
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_launch_dependents();
  pdl_wait();
  for (int m = warp; m < MB; m += 8) {
    const float* xr = x + (size_t)(b0 + m) * d;
    float s = 0.f;
    for (int k = lane * 4; k < d; k += 128) {
      const float4 v = load4(xr + k);
      *reinterpret_cast<float4*>(xs + m * d + k) = v;
      s += (v.x + v.y) + (v.z + v.w);
    }
    if (ln_gamma) {
      const float mean = warp_sum(s) / d;
      float q = 0.f;
      for (int k = lane * 4; k < d; k += 128) {
        const float4 v = *reinterpret_cast<const float4*>(xs + m * d + k);
        const float a = v.x - mean, b = v.y - mean, c = v.z - mean, e = v.w - mean;
        q += (a * a + b * b) + (c * c + e * e);
      }
      const float rstd = 1.0f / sqrtf(warp_sum(q) / d + ln_eps);
      for (int k = lane * 4; k < d; k += 128) {
        const float4 v = *reinterpret_cast<const float4*>(xs + m * d + k);
        const float4 g = load4(ln_gamma + k), bb = load4(ln_beta + k);
        float4 o;
        o.x = (v.x - mean) * rstd * g.x + bb.x; o.y = (v.y - mean) * rstd * g.y + bb.y;
        o.z = (v.z - mean) * rstd * g.z + bb.z; o.w = (v.w - mean) * rstd * g.w + bb.w;
        *reinterpret_cast<float4*>(xs + m * d + k) = o;
        if (xn_out && blockIdx.x == 0) *reinterpret_cast<float4*>(xn_out + (size_t)(b0 + m) * d + k) = o;
      }
    }
  }
  __syncthreads();
  float best[MB];
  int besti[MB];
#pragma unroll
  for (int m = 0; m < MB; ++m) { best[m] = -CUDART_INF_F; besti[m] = 0x7fffffff; }
  for (int v = blockIdx.x * 8 + warp; v < V; v += gridDim.x * 8) {
    const WT* w = wte + (size_t)v * d;
    float acc[MB];
#pragma unroll
    for (int m = 0; m < MB; ++m) acc[m] = 0.f;
    for (int k = lane * 4; k < d; k += 128) {
      float4 wv = load4(w + k);
#pragma unroll
      for (int m = 0; m < MB; ++m) {
        const float4 a = *reinterpret_cast<const float4*>(xs + m * d + k);
        acc[m] += wv.x * a.x + wv.y * a.y + wv.z * a.z + wv.w * a.w;
      }
    }
#pragma unroll
    for (int m = 0; m < MB; ++m) {
      float s = warp_sum(acc[m]);
      if (penalty != 1.0f) {
        uint32_t word = seen[(size_t)(b0 + m) * seen_words + (v >> 5)];
        if ((word >> (v & 31)) & 1u) s = (s < 0.f) ? s * penalty : s / penalty;
      }
      if (s > best[m] || (s == best[m] && v < besti[m])) { best[m] = s; besti[m] = v; }
    }
  }
  __shared__ float sv[8][MB];
  __shared__ int si[8][MB];
  if (lane == 0) {
#pragma unroll
    for (int m = 0; m < MB; ++m) { sv[warp][m] = best[m]; si[warp][m] = besti[m]; }
  }
  __syncthreads();
  if (threadIdx.x < MB) {
    const int m = threadIdx.x;
    float bv = sv[0][m];
    int bi = si[0][m];
    for (int w = 1; w < 8; ++w)
      if (sv[w][m] > bv || (sv[w][m] == bv && si[w][m] < bi)) { bv = sv[w][m]; bi = si[w][m]; }
    part_val[(size_t)blockIdx.x * B + b0 + m] = bv;
    part_idx[(size_t)blockIdx.x * B + b0 + m] = bi;
  }
}

// One 1024-thread CTA: thread (slice, row) scans every 16th tile partial of its row - consecutive threads read
// consecutive rows, so each warp load is two full lines and 16 independent loads are in flight per thread - then the
// 16 slices of a row meet in shared memory and one thread per row does the bookkeeping of
// HF/generation/utils.py:2793-2805.  (The first version gave 16 neighbouring threads one row: every load touched 16
// lines, 17 us per call under ncu.)
__global__ void __launch_bounds__(1024) argmax_finalize_kernel(const float* __restrict__ part_val,
                                                               const int* __restrict__ part_idx, int nparts, int B,
                                                               DecodeState ds) {
  __shared__ int all_finished;
  __shared__ float sv[16][64];
  __shared__ int si[16][64];
  pdl_launch_dependents();
  pdl_wait();
  if (threadIdx.x == 0) all_finished = 1;
  const int len = ds.plain_out ? 0 : *ds.len;
  const int bl = threadIdx.x & 63, slice = threadIdx.x >> 6;
  for (int b0 = 0; b0 < B; b0 += 64) {
    const int b = b0 + bl;
    float bv = -CUDART_INF_F;
    int bi = 0x7fffffff;
    if (b < B) {
#pragma unroll 16
      for (int p = slice; p < nparts; p += 16) {
        const float v = part_val[(size_t)p * B + b];
        const int i = part_idx[(size_t)p * B + b];
        if (v > bv || (v == bv && i < bi)) { bv = v; bi = i; }
      }
    }
    __syncthreads();  // the previous pass has finished reading sv / si
    sv[slice][bl] = bv;
    si[slice][bl] = bi;
    __syncthreads();
    if (slice != 0 || b >= B) continue;
#pragma unroll
    for (int z = 1; z < 16; ++z) {
      const float ov = sv[z][bl];
      const int oi = si[z][bl];
      if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }  // ties go to the lowest id, like torch.argmax
    }
    if (bi == 0x7fffffff) bi = 0;  // all-NaN row: stay in range
    if (ds.plain_out) { ds.plain_out[b] = bi; continue; }
    int fin = ds.finished[b];
    const int tok = fin ? ds.pad_id : bi;  // finished rows emit the pad token (utils.py:2796-2797)
    ds.ids[(size_t)b * ds.ids_ld + len] = tok;
    ds.seen[(size_t)b * ds.seen_words + (tok >> 5)] |= 1u << (tok & 31);  // row-private word: no atomics needed
    if (!fin && tok == ds.eos_id) { fin = 1; ds.finished[b] = 1; }
    if (!fin) all_finished = 0;
  }
  __syncthreads();
  if (threadIdx.x == 0 && !ds.plain_out) {
    *ds.len = len + 1;
    if (all_finished && *ds.stop_len == 0) *ds.stop_len = len + 1;
  }
}

template <typename WT>
int launch_simt(const float* x, const WT* wte, int B, int V, int d, const DecodeState& ds, float* pv, int* pi,
                int nblocks, const float* ln_g, const float* ln_b, float ln_eps, float* xn_out, cudaStream_t st) {
  ARG_CHECK((size_t)8 * d * sizeof(float) <= 48 * 1024, "lmhead: d=%d too wide for the shared-memory row cache", d);
  for (int b0 = 0; b0 < B; b0 += 8) {
    const int mb = (B - b0 < 8) ? (B - b0) : 8;
#define LM_CASE(n)                                                                                              \
  case n:                                                                                                       \
    CUDA_TRY(tutor_launch(lmhead_argmax_simt_kernel<WT, n>, dim3(nblocks), dim3(256), (size_t)n * d * sizeof(float), \
                          st, x, wte, b0, B, V, d, (const uint32_t*)ds.seen, ds.seen_words, ds.penalty, pv, pi, ln_g,  \
                          ln_b, ln_eps, xn_out));                                                                    \
    break;
    switch (mb) {
      LM_CASE(1) LM_CASE(2) LM_CASE(3) LM_CASE(4) LM_CASE(5) LM_CASE(6) LM_CASE(7) LM_CASE(8)
    }
#undef LM_CASE
    KERNEL_LAUNCHED();
  }
  return TUTOR_OK;
}

}  // namespace

size_t k_lmhead_scratch_bytes(int B, int V) {
  (void)V;
  return (size_t)LM_MAX_PARTS * B * (sizeof(float) + sizeof(int));
}

int k_lmhead_argmax(const void* xn, int x_dtype, const void* wte, int w_dtype, int B, int V, int d,
                    const DecodeState& ds, void* scratch, cudaStream_t st) {
  return k_lmhead_argmax_ln(xn, x_dtype, wte, w_dtype, B, V, d, ds, scratch, nullptr, nullptr, 0.f, nullptr, st);
}

// ln_g != null (fp32 activations only): xn is the un-normalised residual stream and ln_f is applied inside the kernel
int k_lmhead_argmax_ln(const void* xn, int x_dtype, const void* wte, int w_dtype, int B, int V, int d,
                       const DecodeState& ds, void* scratch, const float* ln_g, const float* ln_b, float ln_eps,
                       float* xn_out, cudaStream_t st) {
  ARG_CHECK(!ln_g || x_dtype == DT_F32, "lmhead: fused ln_f needs fp32 activations");
  ARG_CHECK(B >= 1 && d % 128 == 0, "lmhead: unsupported B=%d d=%d", B, d);
  float* pv = reinterpret_cast<float*>(scratch);
  int* pi = reinterpret_cast<int*>(pv + (size_t)LM_MAX_PARTS * B);
  int nparts;
  if (x_dtype == DT_F32) {
    nparts = ceil_div(V, 8) < LM_BLOCKS ? ceil_div(V, 8) : LM_BLOCKS;
    if (w_dtype == DT_F32) TRY(launch_simt<float>((const float*)xn, (const float*)wte, B, V, d, ds, pv, pi, nparts, ln_g, ln_b, ln_eps, xn_out, st));
    else TRY(launch_simt<bf16>((const float*)xn, (const bf16*)wte, B, V, d, ds, pv, pi, nparts, ln_g, ln_b, ln_eps, xn_out, st));
  } else {
    ARG_CHECK(w_dtype == DT_BF16, "lmhead: bf16 activations need bf16 weights");
    nparts = ceil_div(V, 128);
    ARG_CHECK(nparts <= LM_MAX_PARTS, "lmhead: vocabulary %d too large for the partial buffer", V);
    TRY(k_gemm_bf16_argmax((const bf16*)wte, V, (const bf16*)xn, B, d, ds.seen, ds.seen_words, ds.penalty, pv, pi, st));
  }
  CUDA_TRY(tutor_launch(argmax_finalize_kernel, dim3(1), dim3(1024), 0, st, (const float*)pv, (const int*)pi, nparts, B, ds));
  KERNEL_LAUNCHED();
  return TUTOR_OK;
}

int k_lmhead_logits(const void* xn, int x_dtype, const void* wte, int w_dtype, int B, int V, int Vpad, int d,
                    float* logits, cudaStream_t st) {
  (void)Vpad;
  GemmEpilogue e{};
  e.out_f32 = logits;
  e.ldo = V;
  if (x_dtype == DT_F32 && w_dtype == DT_F32) return k_gemm_f32((const float*)xn, (const float*)wte, B, V, d, e, st);
  if (x_dtype == DT_BF16 && w_dtype == DT_BF16) return k_gemm_bf16((const bf16*)xn, (const bf16*)wte, B, V, d, e, st);
  tutor_set_error("lmhead_logits: unsupported dtype combination");
  return TUTOR_ERR_ARG;
}
