// fp32 (parity-mode) contractions  C[m,n] = sum_k A[m,k] * W[n,k]  with the shared fused epilogue.
// Both operands are K-contiguous: GPT-2 Conv1D weights ([in,out], HF/pytorch_utils.py:112) are
// transposed once at load; BERT nn.Linear weights and the tied LM head already are [out,in].
//   M <= 8  : weight-streaming GEMV, one warp per output column, 128-bit coalesced loads (HBM-bound)
//   M  > 8  : 64x64x16 register-tiled SIMT GEMM (fp32 FMA only - TF32 would break the 1e-3 logit gate)
#include <cstdlib>

#include "kernels.h"

namespace {

template <int MB>
__global__ void __launch_bounds__(256) gemv_f32_kernel(const float* __restrict__ A, const float* __restrict__ W,
                                                        int M0, int N, int K, GemmEpilogue epi) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n = blockIdx.x * 8 + warp;
  pdl_launch_dependents();
  if (n >= N) { pdl_wait(); return; }
  const float* w = W + (size_t)n * K;
  // PDL: the weight row does not depend on the predecessor kernel - pull it towards L2 before waiting for it
  for (int off = lane * 32; off < K; off += 1024) asm volatile("prefetch.global.L2 [%0];" ::"l"(w + off));
  pdl_wait();
  float acc[MB];
#pragma unroll
  for (int m = 0; m < MB; ++m) acc[m] = 0.f;
  // 4 independent 16-byte weight loads in flight per lane
  int k = lane * 4;
  for (; k + 384 < K; k += 512) {
    float4 w0 = load4(w + k), w1 = load4(w + k + 128), w2 = load4(w + k + 256), w3 = load4(w + k + 384);
#pragma unroll
    for (int m = 0; m < MB; ++m) {
      const float* a = A + (size_t)(M0 + m) * K + k;
      float4 a0 = __ldcg((const float4*)a), a1 = __ldcg((const float4*)(a + 128));
      float4 a2 = __ldcg((const float4*)(a + 256)), a3 = __ldcg((const float4*)(a + 384));
      acc[m] += (w0.x * a0.x + w0.y * a0.y + w0.z * a0.z + w0.w * a0.w) +
                (w1.x * a1.x + w1.y * a1.y + w1.z * a1.z + w1.w * a1.w) +
                (w2.x * a2.x + w2.y * a2.y + w2.z * a2.z + w2.w * a2.w) +
                (w3.x * a3.x + w3.y * a3.y + w3.z * a3.z + w3.w * a3.w);
    }
  }
  for (; k < K; k += 128) {
    float4 w0 = load4(w + k);
#pragma unroll
    for (int m = 0; m < MB; ++m) {
      float4 a0 = __ldcg((const float4*)(A + (size_t)(M0 + m) * K + k));
      acc[m] += w0.x * a0.x + w0.y * a0.y + w0.z * a0.z + w0.w * a0.w;
    }
  }
#pragma unroll
  for (int m = 0; m < MB; ++m) {
    float s = warp_sum(acc[m]);
    if (lane == 0) epilogue_store(epi, M0 + m, n, s);
  }
}

// Short prompts (9..64 rows, e.g. the 32-token prompt of the reference's own configuration).  A 64x64-tiled SGEMM has only
// N/64 = 12..48 CTAs here and spends ~100 us per layer GEMM in its k-loop.  Weight-streaming form with register blocking:
// a warp owns 4 weight rows x 8 input rows; per k-step a lane loads 4 + 8 float4 and issues 128 FMAs (the first version
// of this kernel - one weight row per warp, 9 loads per 32 FMAs - was bound by its loads: 1.97 ms for the 32-token fp32
// prefill of GPT-2 small).  The 32 lane-partial sums are combined by a butterfly that halves the values a lane holds at
// every step (31 shuffles, fixed order: deterministic), after which lane l owns (weight row l / 8, input row l % 8).
// Weights cross HBM once and are re-read from L2 by the other M/8 row chunks.
__global__ void __launch_bounds__(256) gemv_f32_rows_kernel(const float* __restrict__ A, const float* __restrict__ W,
                                                             int M, int N, int K, GemmEpilogue epi) {
  constexpr int RN = 4, MB = 8;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = (blockIdx.x * 8 + warp) * RN, m0 = blockIdx.y * MB;
  if (n0 >= N) return;
  const float* w[RN];
  const float* a[MB];
#pragma unroll
  for (int r = 0; r < RN; ++r) w[r] = W + (size_t)min(n0 + r, N - 1) * K;  // rows past N / M are clamped and never stored
#pragma unroll
  for (int m = 0; m < MB; ++m) a[m] = A + (size_t)min(m0 + m, M - 1) * K;
  float acc[RN * MB];
#pragma unroll
  for (int i = 0; i < RN * MB; ++i) acc[i] = 0.f;
  for (int k = lane * 4; k < K; k += 128) {
    float4 wv[RN], av[MB];
#pragma unroll
    for (int r = 0; r < RN; ++r) wv[r] = load4(w[r] + k);
#pragma unroll
    for (int m = 0; m < MB; ++m) av[m] = load4(a[m] + k);
#pragma unroll
    for (int r = 0; r < RN; ++r)
#pragma unroll
      for (int m = 0; m < MB; ++m)
        acc[r * MB + m] += wv[r].x * av[m].x + wv[r].y * av[m].y + wv[r].z * av[m].z + wv[r].w * av[m].w;
  }
  // butterfly: at distance off a lane keeps the half of its values selected by its own bit and sends the other half
#pragma unroll
  for (int off = 16, cnt = RN * MB; off >= 1; off >>= 1, cnt >>= 1) {
    const bool upper = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < cnt / 2; ++i) {
      const float lo = acc[i], hi = acc[i + cnt / 2];
      acc[i] = (upper ? hi : lo) + __shfl_xor_sync(0xffffffffu, upper ? lo : hi, off);
    }
  }
  const int r = lane >> 3, m = lane & 7;
  if (n0 + r < N && m0 + m < M) epilogue_store(epi, m0 + m, n0 + r, acc[0]);
}

// Same GEMV with the preceding LayerNorm folded in (decode, M <= 8): every CTA normalises the M input rows into
// shared memory itself - M*K floats, a few microseconds of redundant work per CTA - which removes one whole kernel
// (and one dependent-launch boundary) per LayerNorm from the 88-kernel decode chain of the parity path.
// xn_out (optional) receives the normalised rows from CTA 0 (the debug-logits entry point reads them).
template <int MB>
__global__ void __launch_bounds__(256) gemv_f32_ln_kernel(const float* __restrict__ X, const float* __restrict__ gamma,
                                                           const float* __restrict__ beta, float eps,
                                                           const float* __restrict__ W, int N, int K, GemmEpilogue epi,
                                                           float* __restrict__ xn_out) {
  extern __shared__ __align__(16) float xs[];  // [MB][K]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n = blockIdx.x * 8 + warp;
  pdl_launch_dependents();
  const float* w = W + (size_t)min(n, N - 1) * K;
  for (int off = lane * 32; off < K; off += 1024) asm volatile("prefetch.global.L2 [%0];" ::"l"(w + off));
  pdl_wait();
  for (int m = warp; m < MB; m += 8) {  // one warp per row, two-pass like layernorm_kernel
    const float* xr = X + (size_t)m * K;
    float s = 0.f;
    for (int k = lane * 4; k < K; k += 128) {
      const float4 v = load4(xr + k);
      *reinterpret_cast<float4*>(xs + m * K + k) = v;
      s += (v.x + v.y) + (v.z + v.w);
    }
    const float mean = warp_sum(s) / K;
    float q = 0.f;
    for (int k = lane * 4; k < K; k += 128) {
      const float4 v = *reinterpret_cast<const float4*>(xs + m * K + k);
      const float a = v.x - mean, b = v.y - mean, c = v.z - mean, d = v.w - mean;
      q += (a * a + b * b) + (c * c + d * d);
    }
    const float rstd = 1.0f / sqrtf(warp_sum(q) / K + eps);
    for (int k = lane * 4; k < K; k += 128) {
      const float4 v = *reinterpret_cast<const float4*>(xs + m * K + k);
      const float4 g = load4(gamma + k), b = load4(beta + k);
      float4 o;
      o.x = (v.x - mean) * rstd * g.x + b.x; o.y = (v.y - mean) * rstd * g.y + b.y;
      o.z = (v.z - mean) * rstd * g.z + b.z; o.w = (v.w - mean) * rstd * g.w + b.w;
      *reinterpret_cast<float4*>(xs + m * K + k) = o;
      if (xn_out && blockIdx.x == 0) *reinterpret_cast<float4*>(xn_out + (size_t)m * K + k) = o;
    }
  }
  __syncthreads();
  if (n >= N) return;
  float acc[MB];
#pragma unroll
  for (int m = 0; m < MB; ++m) acc[m] = 0.f;
  for (int k = lane * 4; k < K; k += 128) {
    const float4 w0 = load4(w + k);
#pragma unroll
    for (int m = 0; m < MB; ++m) {
      const float4 a0 = *reinterpret_cast<const float4*>(xs + m * K + k);
      acc[m] += w0.x * a0.x + w0.y * a0.y + w0.z * a0.z + w0.w * a0.w;
    }
  }
#pragma unroll
  for (int m = 0; m < MB; ++m) {
    const float s = warp_sum(acc[m]);
    if (lane == 0) epilogue_store(epi, m, n, s);
  }
}

// 16-byte read-only load the compiler may not move: the kernels below want their weights requested BEFORE
// griddepcontrol.wait, and a plain load of a const __restrict__ pointer gets sunk below it (checked in SASS: every
// LDG.CONSTANT sat after ACQBULK) - volatile asm statements keep their order relative to the wait.
__device__ __forceinline__ float4 ldg_pinned(const float* p) {
  float4 v;
  asm volatile("ld.volatile.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}

// Decode GEMV for one or two rows (the reference's own configuration is batch 1) built for the dependent-launch
// chain: the whole weight row of this warp's output column - NV float4 per lane - and the LayerNorm parameters are
// loaded into registers BEFORE griddepcontrol.wait (weights never change), so what remains on the critical path after
// the predecessor finishes is one round trip for the input row, the FMAs and a warp reduction.  With LN the input row is
// normalised in registers: a warp holds the full row (NV float4 per lane), so mean and variance are two warp
// reductions - no shared memory, no block barrier, and no separate LayerNorm kernel (25 of the 88 launches of a
// GPT-2-small decode forward).  xn_out (optional) receives the normalised rows from CTA 0.
// The input row is read with ld.global.cg: under programmatic dependent launch this grid is alive while its predecessor
// still writes that buffer, which rules out the read-only (.nc) path.
template <int MB, int NV, bool LN>
__global__ void __launch_bounds__(256) gemv_f32_reg_kernel(const float* __restrict__ X, const float* __restrict__ gamma,
                                                            const float* __restrict__ beta, float eps,
                                                            const float* __restrict__ W, int N, GemmEpilogue epi,
                                                            float* __restrict__ xn_out) {
  constexpr int K = NV * 128;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n = blockIdx.x * 8 + warp;
  pdl_launch_dependents();
  const float* w = W + (size_t)min(n, N - 1) * K + lane * 4;
  float4 wv[NV], gv[LN ? NV : 1], bv[LN ? NV : 1];
#pragma unroll
  for (int j = 0; j < NV; ++j) wv[j] = ldg_pinned(w + j * 128);
  if (LN) {
#pragma unroll
    for (int j = 0; j < NV; ++j) { gv[j] = ldg_pinned(gamma + lane * 4 + j * 128); bv[j] = ldg_pinned(beta + lane * 4 + j * 128); }
  }
  // (hoisting the epilogue's bias / residual loads up here measured slower: 381 vs 336 us per decode forward)
  pdl_wait();
#pragma unroll
  for (int m = 0; m < MB; ++m) {
    const float* xr = X + (size_t)m * K + lane * 4;
    float acc = 0.f;
    if (LN) {
      float4 xv[NV];
      float s = 0.f;
#pragma unroll
      for (int j = 0; j < NV; ++j) { xv[j] = __ldcg((const float4*)(xr + j * 128)); s += (xv[j].x + xv[j].y) + (xv[j].z + xv[j].w); }
      const float mean = warp_sum(s) / K;  // two-pass, like layernorm_kernel
      float q = 0.f;
#pragma unroll
      for (int j = 0; j < NV; ++j) {
        const float a = xv[j].x - mean, b = xv[j].y - mean, c = xv[j].z - mean, d = xv[j].w - mean;
        q += (a * a + b * b) + (c * c + d * d);
      }
      const float rstd = 1.0f / sqrtf(warp_sum(q) / K + eps);
#pragma unroll
      for (int j = 0; j < NV; ++j) {
        float4 o;
        o.x = (xv[j].x - mean) * rstd * gv[j].x + bv[j].x; o.y = (xv[j].y - mean) * rstd * gv[j].y + bv[j].y;
        o.z = (xv[j].z - mean) * rstd * gv[j].z + bv[j].z; o.w = (xv[j].w - mean) * rstd * gv[j].w + bv[j].w;
        if (xn_out && blockIdx.x == 0 && warp == 0) *reinterpret_cast<float4*>(xn_out + (size_t)m * K + lane * 4 + j * 128) = o;
        acc += wv[j].x * o.x + wv[j].y * o.y + wv[j].z * o.z + wv[j].w * o.w;
      }
    } else {
#pragma unroll
      for (int j0 = 0; j0 < NV; j0 += 8) {  // 8 independent 16-byte input loads in flight per lane
        float4 xv[8];
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (j0 + j < NV) xv[j] = __ldcg((const float4*)(xr + (j0 + j) * 128));
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (j0 + j < NV)
            acc += wv[j0 + j].x * xv[j].x + wv[j0 + j].y * xv[j].y + wv[j0 + j].z * xv[j].z + wv[j0 + j].w * xv[j].w;
      }
    }
    const float s = warp_sum(acc);
    if (lane == 0 && n < N) epilogue_store(epi, m, n, s);
  }
}

// Two shapes of the batch-1 chain do not fill the machine with one warp per output column:
//  * mlp c_proj (N = 768, K = 3072): 96 CTAs on 148 SMs, 24 float4 of weights per lane.  WPC = 2 warps share a column
//    (half of K each, partial sums meet in shared memory): 192 CTAs, 12 float4 per lane.
//  * ln_2 + c_fc (N = 3072, K = 768): 384 CTAs of which 296 are resident - a second wave whose weight loads overlap
//    nothing.  NC = 2 columns per warp: 192 CTAs, and LayerNorm is computed once for both columns.
template <int MB, int NV, bool LN, int NC, int WPC>
__global__ void __launch_bounds__(256) gemv_f32_reg2_kernel(const float* __restrict__ X, const float* __restrict__ gamma,
                                                             const float* __restrict__ beta, float eps,
                                                             const float* __restrict__ W, int N, GemmEpilogue epi) {
  constexpr int K = NV * 128, NVW = NV / WPC;  // float4 per lane of one column that this warp owns
  static_assert(NV % WPC == 0 && (WPC == 1 || !LN), "LayerNorm needs the whole row in one warp");
  __shared__ float part[8][MB * NC];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int col0 = (blockIdx.x * (8 / WPC) + warp / WPC) * NC;  // first of this warp's NC columns
  const int j0 = (warp % WPC) * NVW;                            // first float4 group of this warp's K range
  pdl_launch_dependents();
  float4 wv[NC][NVW], gv[LN ? NV : 1], bv[LN ? NV : 1];
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    const float* w = W + (size_t)min(col0 + c, N - 1) * K + lane * 4;
#pragma unroll
    for (int j = 0; j < NVW; ++j) wv[c][j] = ldg_pinned(w + (j0 + j) * 128);
  }
  if (LN) {
#pragma unroll
    for (int j = 0; j < NV; ++j) { gv[j] = ldg_pinned(gamma + lane * 4 + j * 128); bv[j] = ldg_pinned(beta + lane * 4 + j * 128); }
  }
  pdl_wait();
#pragma unroll
  for (int m = 0; m < MB; ++m) {
    const float* xr = X + (size_t)m * K + lane * 4;
    float4 xv[NVW];
#pragma unroll
    for (int j = 0; j < NVW; ++j) xv[j] = __ldcg((const float4*)(xr + (j0 + j) * 128));  // (.cg: see gemv_f32_reg_kernel)
    if (LN) {
      float s = 0.f;
#pragma unroll
      for (int j = 0; j < NVW; ++j) s += (xv[j].x + xv[j].y) + (xv[j].z + xv[j].w);
      const float mean = warp_sum(s) / K;
      float q = 0.f;
#pragma unroll
      for (int j = 0; j < NVW; ++j) {
        const float a = xv[j].x - mean, b = xv[j].y - mean, c = xv[j].z - mean, d = xv[j].w - mean;
        q += (a * a + b * b) + (c * c + d * d);
      }
      const float rstd = 1.0f / sqrtf(warp_sum(q) / K + eps);
#pragma unroll
      for (int j = 0; j < NVW; ++j) {
        xv[j].x = (xv[j].x - mean) * rstd * gv[j].x + bv[j].x; xv[j].y = (xv[j].y - mean) * rstd * gv[j].y + bv[j].y;
        xv[j].z = (xv[j].z - mean) * rstd * gv[j].z + bv[j].z; xv[j].w = (xv[j].w - mean) * rstd * gv[j].w + bv[j].w;
      }
    }
#pragma unroll
    for (int c = 0; c < NC; ++c) {
      float acc = 0.f;
#pragma unroll
      for (int j = 0; j < NVW; ++j)
        acc += wv[c][j].x * xv[j].x + wv[c][j].y * xv[j].y + wv[c][j].z * xv[j].z + wv[c][j].w * xv[j].w;
      const float s = warp_sum(acc);
      if (WPC == 1) {
        if (lane == 0 && col0 + c < N) epilogue_store(epi, m, col0 + c, s);
      } else if (lane == 0) {
        part[warp][m * NC + c] = s;
      }
    }
  }
  if (WPC > 1) {
    __syncthreads();
    if (warp % WPC == 0 && lane < MB * NC) {
      float s = 0.f;
#pragma unroll
      for (int h = 0; h < WPC; ++h) s += part[warp + h][lane];  // fixed order: deterministic
      const int m = lane / NC, c = lane % NC;
      if (col0 + c < N) epilogue_store(epi, m, col0 + c, s);
    }
  }
}

template <int MB, bool LN>
int launch_gemv_reg(const float* X, const float* gamma, const float* beta, float eps, const float* W, int N, int K,
                    const GemmEpilogue& epi, float* xn_out, cudaStream_t st) {
  const dim3 block(256);
  static const bool plain = tutor_experiment("TUTOR_GEMV_PLAIN") != nullptr;  // A/B: one warp per column everywhere
  if constexpr (LN) {
    if (!plain && K == 768 && N >= 2048 && !xn_out) {  // ln_2 + c_fc: two columns per warp
      CUDA_TRY(tutor_launch(gemv_f32_reg2_kernel<MB, 6, true, 2, 1>, dim3(ceil_div(N, 16)), block, 0, st, X, gamma, beta, eps,
                            W, N, epi));
      KERNEL_LAUNCHED();
      return TUTOR_OK;
    }
  } else {
    if (!plain && K == 3072) {  // mlp c_proj: two warps per column
      CUDA_TRY(tutor_launch(gemv_f32_reg2_kernel<MB, 24, false, 1, 2>, dim3(ceil_div(N, 4)), block, 0, st, X, gamma, beta, eps,
                            W, N, epi));
      KERNEL_LAUNCHED();
      return TUTOR_OK;
    }
  }
  const dim3 grid(ceil_div(N, 8));
#define GEMV_REG(nv)                                                                                                 \
  case nv:                                                                                                           \
    CUDA_TRY(tutor_launch(gemv_f32_reg_kernel<MB, nv, LN>, grid, block, 0, st, X, gamma, beta, eps, W, N, epi, xn_out)); \
    break;
  switch (K / 128) {
    GEMV_REG(1) GEMV_REG(4) GEMV_REG(6) GEMV_REG(10)
    case 24:  // mlp c_proj of GPT-2 small: never behind a LayerNorm (an LN instance would hold 3 x 24 float4 per lane)
      if constexpr (!LN) {
        CUDA_TRY(tutor_launch(gemv_f32_reg_kernel<MB, 24, false>, grid, block, 0, st, X, gamma, beta, eps, W, N, epi, xn_out));
        break;
      }
      return TUTOR_ERR_ARG;
    default:
      tutor_set_error("gemv_f32: K=%d has no register-resident instance", K);
      return TUTOR_ERR_ARG;
  }
#undef GEMV_REG
  KERNEL_LAUNCHED();
  return TUTOR_OK;
}
bool gemv_reg_ok(int M, int K, bool ln) {
  const int nv = K / 128;
  return M >= 1 && M <= 2 && K % 128 == 0 && (nv == 1 || nv == 4 || nv == 6 || nv == 10 || (nv == 24 && !ln));
}

constexpr int BM = 64, BN = 64, BK = 16, PAD = 4;

__global__ void __launch_bounds__(256) sgemm_f32_kernel(const float* __restrict__ A, const float* __restrict__ W,
                                                         int M, int N, int K, GemmEpilogue epi) {
  __shared__ __align__(16) float As[2][BK][BM + PAD];
  __shared__ __align__(16) float Ws[2][BK][BN + PAD];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  const int lr = tid >> 2, lk = (tid & 3) * 4;  // each thread stages one float4 of A and one of W per k-tile
  const int am = min(m0 + lr, M - 1), wn = min(n0 + lr, N - 1);
  const float* ap = A + (size_t)am * K + lk;
  const float* wp = W + (size_t)wn * K + lk;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  float4 ra = load4(ap), rw = load4(wp);
  auto stage = [&](int buf) {
    As[buf][lk][lr] = ra.x; As[buf][lk + 1][lr] = ra.y; As[buf][lk + 2][lr] = ra.z; As[buf][lk + 3][lr] = ra.w;
    Ws[buf][lk][lr] = rw.x; Ws[buf][lk + 1][lr] = rw.y; Ws[buf][lk + 2][lr] = rw.z; Ws[buf][lk + 3][lr] = rw.w;
  };
  stage(0);
  __syncthreads();
  const int nk = K / BK;
  for (int kt = 0; kt < nk; ++kt) {
    const int buf = kt & 1;
    if (kt + 1 < nk) { ra = load4(ap + (kt + 1) * BK); rw = load4(wp + (kt + 1) * BK); }
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      float4 a = *reinterpret_cast<const float4*>(&As[buf][k][ty * 4]);
      float4 w = *reinterpret_cast<const float4*>(&Ws[buf][k][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w}, wv[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], wv[j], acc[i][j]);
    }
    if (kt + 1 < nk) stage(buf ^ 1);
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int m = m0 + ty * 4 + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int n = n0 + tx * 4 + j;
      if (n < N) epilogue_store(epi, m, n, acc[i][j]);
    }
  }
}

}  // namespace

int k_gemv_f32_ln(const float* X, const float* gamma, const float* beta, float eps, const float* W, int M, int N, int K,
                  const GemmEpilogue& epi, float* xn_out, cudaStream_t st) {
  ARG_CHECK(M >= 1 && M <= 8 && K % 128 == 0, "gemv_f32_ln: unsupported M=%d K=%d", M, K);
  if (gemv_reg_ok(M, K, true))
    return M == 1 ? launch_gemv_reg<1, true>(X, gamma, beta, eps, W, N, K, epi, xn_out, st)
                  : launch_gemv_reg<2, true>(X, gamma, beta, eps, W, N, K, epi, xn_out, st);
  const dim3 grid(ceil_div(N, 8)), block(256);
  const size_t smem = (size_t)M * K * sizeof(float);
  ARG_CHECK(smem <= 48 * 1024, "gemv_f32_ln: %d rows of %d do not fit in shared memory", M, K);
#define GEMVLN_CASE(n) case n: CUDA_TRY(tutor_launch(gemv_f32_ln_kernel<n>, grid, block, smem, st, X, gamma, beta, eps, W, N, K, epi, xn_out)); break;
  switch (M) {
    GEMVLN_CASE(1) GEMVLN_CASE(2) GEMVLN_CASE(3) GEMVLN_CASE(4) GEMVLN_CASE(5) GEMVLN_CASE(6) GEMVLN_CASE(7)
    GEMVLN_CASE(8)
  }
#undef GEMVLN_CASE
  KERNEL_LAUNCHED();
  return TUTOR_OK;
}

int k_gemm_f32(const float* A, const float* W, int M, int N, int K, const GemmEpilogue& epi, cudaStream_t st) {
  ARG_CHECK(K % 128 == 0 || (M > 8 && K % BK == 0), "gemm_f32: K=%d unsupported (M=%d)", K, M);
  if (M == 0 || N == 0) return TUTOR_OK;
  if (gemv_reg_ok(M, K, false))
    return M == 1 ? launch_gemv_reg<1, false>(A, nullptr, nullptr, 0.f, W, N, K, epi, nullptr, st)
                  : launch_gemv_reg<2, false>(A, nullptr, nullptr, 0.f, W, N, K, epi, nullptr, st);
  if (M <= 8 && K % 128 == 0) {
    dim3 grid(ceil_div(N, 8)), block(256);
#define GEMV_CASE(n) case n: CUDA_TRY(tutor_launch(gemv_f32_kernel<n>, grid, block, 0, st, A, W, 0, N, K, epi)); break;
    switch (M) {
      GEMV_CASE(1) GEMV_CASE(2) GEMV_CASE(3) GEMV_CASE(4) GEMV_CASE(5) GEMV_CASE(6) GEMV_CASE(7)
      default: CUDA_TRY(tutor_launch(gemv_f32_kernel<8>, grid, block, 0, st, A, W, 0, N, K, epi)); break;
    }
#undef GEMV_CASE
  } else if (M <= 64 && K % 128 == 0) {
    gemv_f32_rows_kernel<<<dim3(ceil_div(N, 32), ceil_div(M, 8)), 256, 0, st>>>(A, W, M, N, K, epi);
  } else {
    ARG_CHECK(K % BK == 0, "gemm_f32: K=%d must be a multiple of %d", K, BK);
    dim3 grid(ceil_div(N, BN), ceil_div(M, BM)), block(256);
    sgemm_f32_kernel<<<grid, block, 0, st>>>(A, W, M, N, K, epi);
  }
  KERNEL_LAUNCHED();
  return TUTOR_OK;
}
