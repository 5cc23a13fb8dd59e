// Row-wise kernels of the tutoring path: embeddings, LayerNorm (+residual), mean-pool, cosine.
// All HBM-bound: one warp per row, 128-bit loads, row cached in registers, shuffle reductions.
#include "kernels.h"

namespace {

constexpr int LN_V4 = 16;               // float4 slots per lane: supports d <= 2048 (GPT-2 large: 1280)
constexpr int LN_MAX_PER_LANE = 4 * LN_V4;  // fully unrolled so the row stays in registers

// ---- GPT-2 embedding: x[m,:] = wte[id] + wpe[pos]   (HF/models/gpt2/modeling_gpt2.py:577-585)
template <typename WT>
__global__ void gpt2_embed_kernel(const int32_t* __restrict__ ids, int ids_ld, const int* __restrict__ len_dev,
                                  int T, const WT* __restrict__ wte, const float* __restrict__ wpe, int d,
                                  float* __restrict__ x, int M, const int* __restrict__ npad) {
  pdl_launch_dependents();
  pdl_wait();
  int m = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (m >= M) return;
  int lane = threadIdx.x & 31;
  int b = m / T, t = m - b * T;
  int pos = len_dev ? (*len_dev - 1) : t;  // decode: the last token written; prefill: 0..T-1
  int id = ids[(size_t)b * ids_ld + pos];
  const WT* e = wte + (size_t)id * d;
  // left-padded rows: position_ids = cumsum(attention_mask) - 1 (HF/generation/utils.py:718-721), i.e. the index
  // among the real tokens; pad slots themselves are never attended to, any in-range position does for them
  const int pe = npad ? max(pos - npad[b], 0) : pos;
  const float* p = wpe + (size_t)pe * d;
  float* o = x + (size_t)m * d;
  for (int i = lane * 4; i < d; i += 128) {
    float4 a = load4(e + i), c = load4(p + i);
    *reinterpret_cast<float4*>(o + i) = make_float4(a.x + c.x, a.y + c.y, a.z + c.z, a.w + c.w);
  }
}

// ---- LayerNorm over the last axis, biased variance, two-pass on a register-resident row.
// in row r is read from x[(r*in_stride + in_offset)*d]; optional fp32 and bf16 outputs.
// NV4 = float4 slots per lane (d <= 128*NV4); every load (row, gamma, beta) is issued up front so the kernel
// pays one memory round trip, not one per slot (the first version loaded gamma/beta inside the output loop:
// 3.0 us for 64 rows; see profiles/README.md).
// STREAM (thousands of rows, prefill): gamma / beta are loaded after the statistics instead (L1 hits by then) - held from
// the start they cost 2 * 4 * NV4 registers per thread and most of the resident warps, and there the kernel is bound by
// the rows it streams (GPT-2 large, 16384 rows: 27 us for 126 MB).
template <int NV4, bool STREAM>
__global__ void layernorm_kernel(const float* __restrict__ x, int in_stride, int in_offset,
                                 const float* __restrict__ gamma, const float* __restrict__ beta, float eps,
                                 float* __restrict__ out_f32, bf16* __restrict__ out_bf16, int M, int d) {
  pdl_launch_dependents();
  pdl_wait();
  int r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= M) return;
  int lane = threadIdx.x & 31;
  const float* xr = x + ((size_t)r * in_stride + in_offset) * d;
  float4 v[NV4], g[NV4], b[NV4];
  const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int j = 0; j < NV4; ++j) {
    const int i = lane * 4 + j * 128;
    const bool ok = i < d;
    v[j] = ok ? load4(xr + i) : zero;
    if (!STREAM) {
      g[j] = ok ? load4(gamma + i) : zero;
      b[j] = ok ? load4(beta + i) : zero;
    }
  }
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < NV4; ++j) s += (v[j].x + v[j].y) + (v[j].z + v[j].w);
  const float mean = warp_sum(s) / d;
  float q = 0.f;
#pragma unroll
  for (int j = 0; j < NV4; ++j) {
    if (lane * 4 + j * 128 < d) {
      const float a0 = v[j].x - mean, a1 = v[j].y - mean, a2 = v[j].z - mean, a3 = v[j].w - mean;
      q += (a0 * a0 + a1 * a1) + (a2 * a2 + a3 * a3);
    }
  }
  const float rstd = 1.0f / sqrtf(warp_sum(q) / d + eps);
#pragma unroll
  for (int j = 0; j < NV4; ++j) {
    const int i = lane * 4 + j * 128;
    if (i < d) {
      if (STREAM) { g[j] = load4(gamma + i); b[j] = load4(beta + i); }
      float4 o;
      o.x = (v[j].x - mean) * rstd * g[j].x + b[j].x;
      o.y = (v[j].y - mean) * rstd * g[j].y + b[j].y;
      o.z = (v[j].z - mean) * rstd * g[j].z + b[j].z;
      o.w = (v[j].w - mean) * rstd * g[j].w + b[j].w;
      if (out_f32) *reinterpret_cast<float4*>(out_f32 + (size_t)r * d + i) = o;
      if (out_bf16) {
        __nv_bfloat162 lo = __floats2bfloat162_rn(o.x, o.y), hi = __floats2bfloat162_rn(o.z, o.w);
        *reinterpret_cast<uint2*>(out_bf16 + (size_t)r * d + i) =
            make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
      }
    }
  }
}

// The same LayerNorm over rows held in bf16 (BERT in bf16 mode keeps its activation stream in bf16 between the linear
// layers, like the reference's bf16 modules do): NV8 16-byte vectors of 8 elements per lane, statistics in fp32.
template <int NV8>
__global__ void layernorm_bf16in_kernel(const bf16* __restrict__ x, const float* __restrict__ gamma,
                                        const float* __restrict__ beta, float eps, float* __restrict__ out_f32,
                                        bf16* __restrict__ out_bf16, int M, int d) {
  pdl_launch_dependents();
  pdl_wait();
  const int r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= M) return;
  const int lane = threadIdx.x & 31;
  const bf16* xr = x + (size_t)r * d;
  float v[NV8][8];
  uint4 raw[NV8];
#pragma unroll
  for (int j = 0; j < NV8; ++j) {  // the whole row first: one memory round trip
    const int i = lane * 8 + j * 256;
    raw[j] = i < d ? *reinterpret_cast<const uint4*>(xr + i) : make_uint4(0, 0, 0, 0);
  }
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < NV8; ++j) {
    unpack_bf16x8(raw[j], v[j]);
    s += ((v[j][0] + v[j][1]) + (v[j][2] + v[j][3])) + ((v[j][4] + v[j][5]) + (v[j][6] + v[j][7]));
  }
  const float mean = warp_sum(s) / d;
  float q = 0.f;
#pragma unroll
  for (int j = 0; j < NV8; ++j) {
    if (lane * 8 + j * 256 < d) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float a = v[j][k] - mean;
        q = fmaf(a, a, q);
      }
    }
  }
  const float rstd = 1.0f / sqrtf(warp_sum(q) / d + eps);
#pragma unroll
  for (int j = 0; j < NV8; ++j) {
    const int i = lane * 8 + j * 256;
    if (i < d) {
      // gamma / beta only now (L1 hits after the first rows): held from the start they cost 48 registers per thread and
      // two thirds of the resident warps, and the kernel is there to stream rows
      const float4 g0 = load4(gamma + i), g1 = load4(gamma + i + 4), b0 = load4(beta + i), b1 = load4(beta + i + 4);
      const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
      const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
      float o[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) o[k] = (v[j][k] - mean) * rstd * gg[k] + bb[k];
      if (out_f32) {
        *reinterpret_cast<float4*>(out_f32 + (size_t)r * d + i) = make_float4(o[0], o[1], o[2], o[3]);
        *reinterpret_cast<float4*>(out_f32 + (size_t)r * d + i + 4) = make_float4(o[4], o[5], o[6], o[7]);
      }
      if (out_bf16) {
        uint32_t w[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          __nv_bfloat162 h2 = __floats2bfloat162_rn(o[2 * k], o[2 * k + 1]);
          w[k] = *reinterpret_cast<uint32_t*>(&h2);
        }
        *reinterpret_cast<uint4*>(out_bf16 + (size_t)r * d + i) = make_uint4(w[0], w[1], w[2], w[3]);
      }
    }
  }
}

// ---- BERT embeddings: word + token_type(0) + position, then LayerNorm eps 1e-12
// (HF/models/bert/modeling_bert.py:102-111).
template <typename WT>
__global__ void bert_embed_ln_kernel(const int32_t* __restrict__ ids, int T, const WT* __restrict__ word,
                                     const float* __restrict__ pos, const float* __restrict__ type0,
                                     const float* __restrict__ gamma, const float* __restrict__ beta, float eps,
                                     float* __restrict__ out_f32, bf16* __restrict__ out_bf16, int M, int d) {
  int r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= M) return;
  int lane = threadIdx.x & 31;
  int t = r % T;
  const WT* w = word + (size_t)ids[r] * d;
  const float* p = pos + (size_t)t * d;
  float v[LN_MAX_PER_LANE];
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < LN_V4; ++j) {
    const int i = lane * 4 + j * 128, nv = 4 * j;
    v[nv] = v[nv + 1] = v[nv + 2] = v[nv + 3] = 0.f;
    if (i < d) {
      float4 a = load4(w + i), c = load4(type0 + i), e = load4(p + i);
      v[nv] = (a.x + c.x) + e.x; v[nv + 1] = (a.y + c.y) + e.y;
      v[nv + 2] = (a.z + c.z) + e.z; v[nv + 3] = (a.w + c.w) + e.w;
    }
    s += (v[nv] + v[nv + 1]) + (v[nv + 2] + v[nv + 3]);
  }
  float mean = warp_sum(s) / d;
  float q = 0.f;
#pragma unroll
  for (int j = 0; j < LN_V4; ++j) {
    if (lane * 4 + j * 128 < d) {
#pragma unroll
      for (int k = 0; k < 4; ++k) { float c = v[4 * j + k] - mean; q += c * c; }
    }
  }
  float rstd = 1.0f / sqrtf(warp_sum(q) / d + eps);
#pragma unroll
  for (int j = 0; j < LN_V4; ++j) {
    const int i = lane * 4 + j * 128, nv = 4 * j;
    if (i >= d) break;
    float4 g = load4(gamma + i), b = load4(beta + i);
    float4 o;
    o.x = (v[nv] - mean) * rstd * g.x + b.x;
    o.y = (v[nv + 1] - mean) * rstd * g.y + b.y;
    o.z = (v[nv + 2] - mean) * rstd * g.z + b.z;
    o.w = (v[nv + 3] - mean) * rstd * g.w + b.w;
    if (out_f32) *reinterpret_cast<float4*>(out_f32 + (size_t)r * d + i) = o;
    if (out_bf16) {
      __nv_bfloat162 lo = __floats2bfloat162_rn(o.x, o.y), hi = __floats2bfloat162_rn(o.z, o.w);
      *reinterpret_cast<uint2*>(out_bf16 + (size_t)r * d + i) =
          make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
    }
  }
}

// ---- unmasked mean over T (REF/lms_server.py:101). grid (d/128, B), block (128, MP_GROUPS): thread (c, g) owns one
// column and the rows g, g + MP_GROUPS, ... (coalesced 512 B per warp per row, 8 loads in flight per thread: the first
// version's 128-thread blocks kept 0.4 MB in flight in all and ran at 1.2 TB/s); the groups meet in shared memory and
// are added in a fixed order -> deterministic.
constexpr int MP_GROUPS = 8;
__global__ void __launch_bounds__(128 * MP_GROUPS) meanpool_kernel(const float* __restrict__ x, float* __restrict__ out,
                                                                   int T, int d) {
  __shared__ float part[MP_GROUPS][128];
  const int b = blockIdx.y, g = threadIdx.y;
  const int c = blockIdx.x * 128 + threadIdx.x;
  float s[8];
#pragma unroll
  for (int u = 0; u < 8; ++u) s[u] = 0.f;
  if (c < d) {
    const float* p = x + (size_t)b * T * d + c;
    int t = g;
    for (; t + 7 * MP_GROUPS < T; t += 8 * MP_GROUPS) {
#pragma unroll
      for (int u = 0; u < 8; ++u) s[u] += p[(size_t)(t + u * MP_GROUPS) * d];
    }
    for (; t < T; t += MP_GROUPS) s[0] += p[(size_t)t * d];
  }
  part[g][threadIdx.x] = ((s[0] + s[1]) + (s[2] + s[3])) + ((s[4] + s[5]) + (s[6] + s[7]));
  __syncthreads();
  if (g == 0 && c < d) {
    float tot = 0.f;
#pragma unroll
    for (int k = 0; k < MP_GROUPS; ++k) tot += part[k][threadIdx.x];
    out[(size_t)b * d + c] = tot / T;
  }
}

// ---- F.cosine_similarity(dim=1, eps): x.y / (max(|x|,eps) * max(|y|,eps))  (REF/lms_server.py:104)
__global__ void cosine_pairs_kernel(const float* __restrict__ a, const float* __restrict__ b,
                                    float* __restrict__ out, int B, int d, float eps) {
  int r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= B) return;
  int lane = threadIdx.x & 31;
  float dot = 0.f, na = 0.f, nb = 0.f;
  for (int i = lane; i < d; i += 32) {
    float x = a[(size_t)r * d + i], y = b[(size_t)r * d + i];
    dot += x * y; na += x * x; nb += y * y;
  }
  dot = warp_sum(dot); na = warp_sum(na); nb = warp_sum(nb);
  if (lane == 0) out[r] = dot / (fmaxf(sqrtf(na), eps) * fmaxf(sqrtf(nb), eps));
}

// ---- mark every prompt token in the per-sequence seen-bitmap (for the fused repetition penalty)
__global__ void seen_bitmap_kernel(const int32_t* __restrict__ ids, int ids_ld, int T, uint32_t* __restrict__ seen,
                                   int words_per_row, int B) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * T) return;
  int b = i / T, t = i - b * T;
  int id = ids[(size_t)b * ids_ld + t];
  atomicOr(&seen[(size_t)b * words_per_row + (id >> 5)], 1u << (id & 31));
}

}  // namespace

int k_gpt2_embed(const int32_t* ids, int ids_ld, const int* len_dev, int T, const void* wte, int wdtype,
                 const float* wpe, int d, float* x, int M, const int* npad, cudaStream_t st) {
  ARG_CHECK(d % 4 == 0, "embed: d %% 4 != 0");
  dim3 grid(ceil_div(M, 8)), block(256);
  if (wdtype == DT_F32)
    CUDA_TRY(tutor_launch(gpt2_embed_kernel<float>, grid, block, 0, st, ids, ids_ld, len_dev, T, (const float*)wte, wpe,
                          d, x, M, npad));
  else
    CUDA_TRY(tutor_launch(gpt2_embed_kernel<bf16>, grid, block, 0, st, ids, ids_ld, len_dev, T, (const bf16*)wte, wpe,
                          d, x, M, npad));
  KERNEL_LAUNCHED();
  return TUTOR_OK;
}

int k_layernorm(const float* x, int in_stride, int in_offset, const float* gamma, const float* beta, float eps,
                float* out_f32, bf16* out_bf16, int M, int d, cudaStream_t st) {
  ARG_CHECK(d % 4 == 0 && d <= 32 * LN_MAX_PER_LANE, "layernorm: unsupported d=%d", d);
  if (M == 0) return TUTOR_OK;
  const dim3 grid(ceil_div(M, 4)), block(128);
#define LN_LAUNCH(NV4)                                                                                                  \
  do {                                                                                                                  \
    if (M >= 1024)                                                                                                      \
      CUDA_TRY(tutor_launch(layernorm_kernel<NV4, true>, grid, block, 0, st, x, in_stride, in_offset, gamma, beta, eps,  \
                            out_f32, out_bf16, M, d));                                                                  \
    else                                                                                                                \
      CUDA_TRY(tutor_launch(layernorm_kernel<NV4, false>, grid, block, 0, st, x, in_stride, in_offset, gamma, beta, eps, \
                            out_f32, out_bf16, M, d));                                                                  \
  } while (0)
  if (d <= 128) LN_LAUNCH(1);
  else if (d <= 768) LN_LAUNCH(6);
  else if (d <= 1280) LN_LAUNCH(10);
  else LN_LAUNCH(16);
#undef LN_LAUNCH
  KERNEL_LAUNCHED();
  return TUTOR_OK;
}

int k_layernorm_bf16in(const bf16* x, const float* gamma, const float* beta, float eps, float* out_f32, bf16* out_bf16,
                       int M, int d, cudaStream_t st) {
  ARG_CHECK(d % 8 == 0 && d <= 2048 && (reinterpret_cast<uintptr_t>(x) & 15) == 0, "layernorm (bf16 rows): unsupported d=%d", d);
  if (M == 0) return TUTOR_OK;
  const dim3 grid(ceil_div(M, 4)), block(128);
#define LN_LAUNCH(NV8) \
  CUDA_TRY(tutor_launch(layernorm_bf16in_kernel<NV8>, grid, block, 0, st, x, gamma, beta, eps, out_f32, out_bf16, M, d))
  if (d <= 256) LN_LAUNCH(1);
  else if (d <= 768) LN_LAUNCH(3);
  else if (d <= 1280) LN_LAUNCH(5);
  else LN_LAUNCH(8);
#undef LN_LAUNCH
  KERNEL_LAUNCHED();
  return TUTOR_OK;
}

int k_bert_embed_ln(const int32_t* ids, int T, const void* word, int wdtype, const float* pos, const float* type0,
                    const float* gamma, const float* beta, float eps, float* out_f32, bf16* out_bf16, int M, int d,
                    cudaStream_t st) {
  ARG_CHECK(d % 4 == 0 && d <= 32 * LN_MAX_PER_LANE, "bert_embed: unsupported d=%d", d);
  if (M == 0) return TUTOR_OK;
  dim3 grid(ceil_div(M, 4)), block(128);
  if (wdtype == DT_F32)
    bert_embed_ln_kernel<float><<<grid, block, 0, st>>>(ids, T, (const float*)word, pos, type0, gamma, beta, eps,
                                                        out_f32, out_bf16, M, d);
  else
    bert_embed_ln_kernel<bf16><<<grid, block, 0, st>>>(ids, T, (const bf16*)word, pos, type0, gamma, beta, eps,
                                                       out_f32, out_bf16, M, d);
  KERNEL_LAUNCHED();
  return TUTOR_OK;
}

int k_meanpool(const float* x, float* out, int B, int T, int d, cudaStream_t st) {
  if (B == 0) return TUTOR_OK;
  meanpool_kernel<<<dim3(ceil_div(d, 128), B), dim3(128, MP_GROUPS), 0, st>>>(x, out, T, d);
  KERNEL_LAUNCHED();
  return TUTOR_OK;
}

int k_cosine_pairs(const float* a, const float* b, float* out, int B, int d, float eps, cudaStream_t st) {
  if (B == 0) return TUTOR_OK;
  cosine_pairs_kernel<<<ceil_div(B, 4), 128, 0, st>>>(a, b, out, B, d, eps);
  KERNEL_LAUNCHED();
  return TUTOR_OK;
}

int k_seen_bitmap(const int32_t* ids, int ids_ld, int T, uint32_t* seen, int words_per_row, int B, cudaStream_t st) {
  if (B * T == 0) return TUTOR_OK;
  seen_bitmap_kernel<<<ceil_div(B * T, 256), 256, 0, st>>>(ids, ids_ld, T, seen, words_per_row, B);
  KERNEL_LAUNCHED();
  return TUTOR_OK;
}
