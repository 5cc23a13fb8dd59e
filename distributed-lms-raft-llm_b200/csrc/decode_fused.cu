// Reductions that close the split-K decode GEMMs (gemm_bf16_tc.cu, EPI_PARTIAL), each fused with the
// elementwise work that follows it in the GPT-2 block so the partials are read exactly once:
//   reduce + bias + residual add + LayerNorm   (after attn c_proj and mlp c_proj; HF/models/gpt2/modeling_gpt2.py:282,304,307)
//   reduce + bias + gelu_new -> bf16           (after c_fc; :239-240)
// The partials live in L2 (a few MB); summation order over the splits is fixed, so results are deterministic.
#include "kernels.h"

namespace {

__device__ __forceinline__ float block_sum(float v, float* sm) {
  v = warp_sum(v);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  __syncthreads();  // sm may still be read by the previous reduction
  if (lane == 0) sm[warp] = v;
  __syncthreads();
  float t = 0.f;
  for (int w = 0; w < nw; ++w) t += sm[w];  // same order in every thread
  return t;
}

// one CTA per row m; N <= 4 * 2 * blockDim.x (each thread owns up to two float4 column groups)
__global__ void __launch_bounds__(256) reduce_bias_residual_ln_kernel(
    const float* __restrict__ partial, int ks, int M, int N, const float* __restrict__ bias, float* __restrict__ x,
    const float* __restrict__ gamma, const float* __restrict__ beta, float eps, float* __restrict__ xn_f32,
    bf16* __restrict__ xn_bf16) {
  __shared__ float sm[8];
  pdl_launch_dependents();
  pdl_wait();
  const int m = blockIdx.x;
  float4 v[2], gm[2], bt[2];  // gamma/beta are fetched with the first wave of loads, not after the reductions
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const int c = (threadIdx.x + j * 256) * 4;
    if (c < N) {
      gm[j] = *reinterpret_cast<const float4*>(gamma + c);
      bt[j] = *reinterpret_cast<const float4*>(beta + c);
    }
  }
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const int c = (threadIdx.x + j * 256) * 4;
    v[j] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c < N) {
      float4 acc = *reinterpret_cast<const float4*>(bias + c);
      const float4 r = *reinterpret_cast<const float4*>(x + (size_t)m * N + c);
      acc.x += r.x; acc.y += r.y; acc.z += r.z; acc.w += r.w;
#pragma unroll 8  // keep 8 independent L2 loads in flight; the sum stays in ascending-z order
      for (int z = 0; z < ks; ++z) {
        const float4 p = *reinterpret_cast<const float4*>(partial + ((size_t)z * M + m) * N + c);
        acc.x += p.x; acc.y += p.y; acc.z += p.z; acc.w += p.w;
      }
      *reinterpret_cast<float4*>(x + (size_t)m * N + c) = acc;
      v[j] = acc;
      s += (acc.x + acc.y) + (acc.z + acc.w);
    }
  }
  const float mean = block_sum(s, sm) / N;
  float q = 0.f;
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    if ((threadIdx.x + j * 256) * 4 < N) {
      const float a = v[j].x - mean, b = v[j].y - mean, c = v[j].z - mean, d = v[j].w - mean;
      q += (a * a + b * b) + (c * c + d * d);
    }
  }
  const float rstd = 1.0f / sqrtf(block_sum(q, sm) / N + eps);
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const int c = (threadIdx.x + j * 256) * 4;
    if (c < N) {
      const float4 g = gm[j], b = bt[j];
      float4 o;
      o.x = (v[j].x - mean) * rstd * g.x + b.x;
      o.y = (v[j].y - mean) * rstd * g.y + b.y;
      o.z = (v[j].z - mean) * rstd * g.z + b.z;
      o.w = (v[j].w - mean) * rstd * g.w + b.w;
      if (xn_f32) *reinterpret_cast<float4*>(xn_f32 + (size_t)m * N + c) = o;
      if (xn_bf16) {
        __nv_bfloat162 lo = __floats2bfloat162_rn(o.x, o.y), hi = __floats2bfloat162_rn(o.z, o.w);
        *reinterpret_cast<uint2*>(xn_bf16 + (size_t)m * N + c) =
            make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
      }
    }
  }
}

__global__ void __launch_bounds__(256) reduce_bias_act_kernel(const float* __restrict__ partial, int ks, int M, int N,
                                                              const float* __restrict__ bias, int act,
                                                              bf16* __restrict__ out) {
  pdl_launch_dependents();
  pdl_wait();
  const size_t i4 = (size_t)blockIdx.x * blockDim.x + threadIdx.x;  // one float4 of the [M,N] output per thread
  const size_t total4 = (size_t)M * N / 4;
  if (i4 >= total4) return;
  const size_t e = i4 * 4;
  const int n = (int)(e % N);
  float4 acc = *reinterpret_cast<const float4*>(bias + n);
#pragma unroll 8
  for (int z = 0; z < ks; ++z) {
    const float4 p = *reinterpret_cast<const float4*>(partial + (size_t)z * M * N + e);
    acc.x += p.x; acc.y += p.y; acc.z += p.z; acc.w += p.w;
  }
  acc.x = apply_act(acc.x, act); acc.y = apply_act(acc.y, act);
  acc.z = apply_act(acc.z, act); acc.w = apply_act(acc.w, act);
  __nv_bfloat162 lo = __floats2bfloat162_rn(acc.x, acc.y), hi = __floats2bfloat162_rn(acc.z, acc.w);
  *reinterpret_cast<uint2*>(out + e) = make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
}

}  // namespace

int k_reduce_bias_residual_ln(const float* partial, int ks, int M, int N, const float* bias, float* x,
                              const float* gamma, const float* beta, float eps, float* xn_f32, bf16* xn_bf16,
                              cudaStream_t st) {
  ARG_CHECK(N % 4 == 0 && N <= 2048, "reduce_ln: unsupported N=%d", N);
  if (M == 0) return TUTOR_OK;
  CUDA_TRY(tutor_launch(reduce_bias_residual_ln_kernel, dim3(M), dim3(256), 0, st, partial, ks, M, N, bias, x, gamma, beta,
                        eps, xn_f32, xn_bf16));
  KERNEL_LAUNCHED();
  return TUTOR_OK;
}

int k_reduce_bias_act(const float* partial, int ks, int M, int N, const float* bias, int act, bf16* out,
                      cudaStream_t st) {
  ARG_CHECK(N % 4 == 0, "reduce_act: unsupported N=%d", N);
  if (M == 0) return TUTOR_OK;
  const size_t total4 = (size_t)M * N / 4;
  CUDA_TRY(tutor_launch(reduce_bias_act_kernel, dim3((unsigned)((total4 + 255) / 256)), dim3(256), 0, st, partial, ks, M,
                        N, bias, act, out));
  KERNEL_LAUNCHED();
  return TUTOR_OK;
}
