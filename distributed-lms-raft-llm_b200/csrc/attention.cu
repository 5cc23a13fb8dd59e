// Attention for the tutoring path; head_dim = 64 everywhere (GPT-2 small/large, BERT-base).
//   attn_prefill : T queries per sequence, causal (GPT-2, HF/integrations/sdpa_attention.py:77: is_causal iff T>1)
//                  or bidirectional (BERT, HF/models/bert/modeling_bert.py:192-205); scale 1/8.
//   attn_decode  : one query per (b,h) streaming the KV cache [B][H][Smax][64] - the HBM-bound kernel that
//                  carries 86 % of the bytes at bs=64.
#include <cuda.h>
#include <math_constants.h>

#include <cstdlib>

#include "kernels.h"

namespace {

template <bool FAST> __device__ __forceinline__ float exp_sel(float x) { return FAST ? __expf(x) : expf(x); }

template <typename T> struct Vec16;  // one 16-byte vector of T -> floats
template <> struct Vec16<float> {
  static constexpr int N = 4;
  __device__ static void load(const float* p, float* f) {
    float4 v = *reinterpret_cast<const float4*>(p);
    f[0] = v.x; f[1] = v.y; f[2] = v.z; f[3] = v.w;
  }
};
template <> struct Vec16<bf16> {
  static constexpr int N = 8;
  __device__ static void load(const bf16* p, float* f) {
    uint4 v = *reinterpret_cast<const uint4*>(p);
    unpack_bf16x8(v, f);
  }
};

// ------------------------------------------------------------------------------------ decode
// grid (H, B), 128 threads.  A row of K (64 elements) is covered by LPR lanes holding one 16-byte
// vector each, so one warp-wide load instruction fetches KPW whole keys (fully coalesced 512 B).
// Every lane group runs its own online softmax over its share of the keys (4-way unrolled for
// memory-level parallelism); the 4*KPW partial states are merged once at the end through smem.
//
// FUSED (split-K decode path): the c_attn GEMM left raw partial sums [ks][B][3*H*64] in L2; this kernel finishes
// them for its own (b,h) - bias + fixed-order sum - appends the new K/V row to the cache (DynamicLayer.update,
// HF/cache_utils.py:119-120, without the torch.cat) and attends over positions [0,len) with the fresh row taken
// from shared memory, so the q/k/v of the current token never make a round trip through HBM.
// One element of q / k / v of the current token from the c_attn split-K partials (FUSED kernels).  With a folded LayerNorm
// (gemm_decode_fused.cu) the partial sums are x . W' and the LayerNorm enters here: rstd (sum - mean s[n]) + c[n] + bias[n].
// FOLD is a template parameter so that the default kernels stay exactly the code they were before the folded path existed
// (as a run-time branch it cost the split-attention kernel of GPT-2 large 20 %: registers, a spill).
template <bool FOLD>
__device__ __forceinline__ float qkv_from_partials(const float* __restrict__ qkv_partial, int ks, const float* __restrict__ qkv_bias,
                                                   const QkvFold& fold, int B, int b, int n, int d) {
  if (!FOLD) {
    float acc = qkv_bias[n];
#pragma unroll 8
    for (int zz = 0; zz < ks; ++zz) acc += qkv_partial[((size_t)zz * B + b) * (3 * d) + n];
    return acc;
  }
  float mean = 0.f, rstd = 1.f, sn = 0.f, add = qkv_bias[n];
  if (fold.s) {
    const float2* st2 = reinterpret_cast<const float2*>(fold.stats) + b;
    float2 tv[16];  // every pair is requested before the first is used
#pragma unroll
    for (int i = 0; i < 16; ++i) tv[i] = i < fold.nst ? st2[(size_t)i * B] : make_float2(0.f, 0.f);
    sn = fold.s[n];
    add += fold.c[n];
    float sm = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) sm += tv[i].x;
    mean = sm / (float)fold.nst;
    const float n_per = (float)(d / fold.nst);
    float m2 = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const float dm = tv[i].x - mean;
      if (i < fold.nst) m2 += tv[i].y + n_per * dm * dm;
    }
    rstd = 1.0f / sqrtf(m2 / (float)d + fold.eps);
  }
  float acc = fold.s ? 0.f : add;
#pragma unroll 8
  for (int zz = 0; zz < ks; ++zz) acc += qkv_partial[((size_t)zz * B + b) * (3 * d) + n];
  return fold.s ? fmaf(rstd, acc - mean * sn, add) : acc;
}

template <typename QT, typename KT, typename OT, bool FAST, bool FUSED, bool FOLD = false>
__global__ void __launch_bounds__(128) attn_decode_kernel(const QT* __restrict__ q, KT* kc, KT* vc,
                                                          OT* __restrict__ out, int H, int Smax,
                                                          const int* __restrict__ len_dev, int len_host,
                                                          const float* __restrict__ qkv_partial, int ks,
                                                          const float* __restrict__ qkv_bias, int B, const QkvFold fold,
                                                          float* split_scratch, int* split_counters,
                                                          const int* __restrict__ npad) {
  constexpr int EPL = Vec16<KT>::N;  // elements per lane
  constexpr int LPR = 64 / EPL;      // lanes per key row
  constexpr int KPW = 32 / LPR;      // keys per warp-wide load
  constexpr int NG = 4 * KPW;        // lane groups per block
  constexpr int UNR = 4;
  pdl_launch_dependents();
  pdl_wait();
  // gridDim.z > 1: the key range is cut into gridDim.z splits so that B*H*splits CTAs spread evenly over the SMs
  // (768 CTAs on 148 SMs leave some SMs with 6 and some with 5: a 14 % tail); the last CTA to finish a (b,h)
  // merges the split states in split order, so the result does not depend on which one that is.
  const int h = blockIdx.x, b = blockIdx.y, z = blockIdx.z, nsplit = gridDim.z;
  const int len = len_dev ? *len_dev : len_host;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = warp * KPW + lane / LPR;  // this lane's group
  const int e0 = (lane % LPR) * EPL;      // first element of the row this lane owns
  const size_t base = ((size_t)b * H + h) * Smax * 64;
  const int kp = npad ? npad[b] : 0;  // left padding: cache rows [0, kp) of this sequence are never attended to
  const KT* kb = kc + base + (size_t)kp * 64 + e0;
  const KT* vb = vc + base + (size_t)kp * 64 + e0;

  __shared__ float s_new[3][64];  // FUSED: q, and the K/V row of the current token (as stored, i.e. rounded to KT)
  float qv[EPL];
  if (FUSED) {
    const int d = H * 64;
    for (int e = threadIdx.x; e < 192; e += 128) {
      const int which = e >> 6, i = e & 63;
      const int n = which * d + h * 64 + i;
      const float acc = qkv_from_partials<FOLD>(qkv_partial, ks, qkv_bias, fold, B, b, n, d);
      if (which == 0) {
        s_new[0][i] = acc;
      } else {
        const KT st = from_f32<KT>(acc);
        if (z == 0) (which == 1 ? kc : vc)[base + (size_t)(len - 1) * 64 + i] = st;  // one split appends the row
        s_new[which][i] = to_f32(st);
      }
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < EPL; ++i) qv[i] = s_new[0][e0 + i] * 0.125f;
  } else {
    const QT* qp = q + ((size_t)b * H + h) * 64 + e0;
#pragma unroll
    for (int i = 0; i < EPL; ++i) qv[i] = to_f32(qp[i]) * 0.125f;  // scale = head_dim^-0.5
  }
  float m = -CUDART_INF_F, l = 0.f, o[EPL];
#pragma unroll
  for (int i = 0; i < EPL; ++i) o[i] = 0.f;

  // FUSED: the row this CTA just appended (position len-1) is taken from shared memory after the loop
  const int nk_all = (FUSED ? len - 1 : len) - kp;  // rows are counted from the first real token
  const int per = (nk_all + nsplit - 1) / nsplit;
  const int kbeg = z * per;
  const int nkeys = min(nk_all, kbeg + per);  // this split covers keys [kbeg, nkeys)
  // trip count is uniform per warp (the shuffles below need every lane); groups past the end are masked
  for (int jw = kbeg + warp * KPW; jw < nkeys; jw += NG * UNR) {
    const int j0 = jw + lane / LPR;
    float kf[UNR][EPL], s[UNR];
#pragma unroll
    for (int u = 0; u < UNR; ++u) {
      int j = min(j0 + u * NG, nkeys - 1);  // clamped: masked below, never out of bounds
      Vec16<KT>::load(kb + (size_t)j * 64, kf[u]);
    }
    float vf[UNR][EPL];
#pragma unroll
    for (int u = 0; u < UNR; ++u) {
      int j = min(j0 + u * NG, nkeys - 1);
      Vec16<KT>::load(vb + (size_t)j * 64, vf[u]);
    }

#pragma unroll
    for (int u = 0; u < UNR; ++u) {
      float d = 0.f;
#pragma unroll
      for (int i = 0; i < EPL; ++i) d = fmaf(qv[i], kf[u][i], d);
#pragma unroll
      for (int off = LPR / 2; off > 0; off >>= 1) d += __shfl_xor_sync(0xffffffffu, d, off);
      s[u] = (j0 + u * NG < nkeys) ? d : -CUDART_INF_F;
    }
    float mn = m;
#pragma unroll
    for (int u = 0; u < UNR; ++u) mn = fmaxf(mn, s[u]);
    if (mn == -CUDART_INF_F) continue;  // this group has seen no valid key yet (no shuffles below)
    float corr = exp_sel<FAST>(m - mn);
    l *= corr;
#pragma unroll
    for (int i = 0; i < EPL; ++i) o[i] *= corr;
#pragma unroll
    for (int u = 0; u < UNR; ++u) {
      float p = exp_sel<FAST>(s[u] - mn);
      l += p;
#pragma unroll
      for (int i = 0; i < EPL; ++i) o[i] = fmaf(p, vf[u][i], o[i]);
    }
    m = mn;
  }

  if (FUSED && z == nsplit - 1) {  // the current token's own key/value, folded into group 0 of the last split
    float d = 0.f;
#pragma unroll
    for (int i = 0; i < EPL; ++i) d = fmaf(qv[i], s_new[1][e0 + i], d);
#pragma unroll
    for (int off = LPR / 2; off > 0; off >>= 1) d += __shfl_xor_sync(0xffffffffu, d, off);
    if (g == 0) {
      const float mn = fmaxf(m, d);
      const float corr = exp_sel<FAST>(m - mn), p = exp_sel<FAST>(d - mn);
      l = l * corr + p;
#pragma unroll
      for (int i = 0; i < EPL; ++i) o[i] = fmaf(p, s_new[2][e0 + i], o[i] * corr);
      m = mn;
    }
  }

  __shared__ float sm_m[NG], sm_l[NG], sm_o[NG][64];
  if (lane % LPR == 0) { sm_m[g] = m; sm_l[g] = l; }
#pragma unroll
  for (int i = 0; i < EPL; ++i) sm_o[g][e0 + i] = o[i];
  __syncthreads();
  if (threadIdx.x < 64) {
    float M = -CUDART_INF_F;
#pragma unroll
    for (int s2 = 0; s2 < NG; ++s2) M = fmaxf(M, sm_m[s2]);
    float num = 0.f, den = 0.f;
#pragma unroll
    for (int s2 = 0; s2 < NG; ++s2) {
      float w = (sm_m[s2] == -CUDART_INF_F) ? 0.f : exp_sel<FAST>(sm_m[s2] - M);
      num = fmaf(w, sm_o[s2][threadIdx.x], num);
      den = fmaf(w, sm_l[s2], den);
    }
    if (nsplit == 1) {
      out[((size_t)b * H + h) * 64 + threadIdx.x] = from_f32<OT>(num / den);
    } else {  // publish this split's state: [max, denominator, 64 numerators]
      float* st = split_scratch + (((size_t)b * H + h) * nsplit + z) * 66;
      if (threadIdx.x == 0) { st[0] = M; st[1] = den; }
      st[2 + threadIdx.x] = num;
      __threadfence();
    }
  }
  if (nsplit > 1) {
    __shared__ int is_last;
    __syncthreads();
    if (threadIdx.x == 0) {
      const int done = atomicAdd(&split_counters[b * H + h], 1);
      is_last = (done == nsplit - 1);
      if (is_last) split_counters[b * H + h] = 0;  // ready for the next launch
    }
    __syncthreads();
    if (is_last && threadIdx.x < 64) {
      __threadfence();
      const float* st = split_scratch + ((size_t)b * H + h) * nsplit * 66;
      float M = -CUDART_INF_F;
      for (int s2 = 0; s2 < nsplit; ++s2) M = fmaxf(M, __ldcg(st + s2 * 66));
      float num = 0.f, den = 0.f;
      for (int s2 = 0; s2 < nsplit; ++s2) {  // fixed split order: deterministic
        const float ms = __ldcg(st + s2 * 66);
        const float w = (ms == -CUDART_INF_F) ? 0.f : exp_sel<FAST>(ms - M);
        num = fmaf(w, __ldcg(st + s2 * 66 + 2 + threadIdx.x), num);
        den = fmaf(w, __ldcg(st + s2 * 66 + 1), den);
      }
      out[((size_t)b * H + h) * 64 + threadIdx.x] = from_f32<OT>(num / den);
    }
  }
}

// ------------------------------------------------------------------- decode, bulk-copy pipelined (bf16)
// Same math and the same 16 independent lane-group softmax states as attn_decode_kernel, but the KV stream is
// decoupled from the arithmetic: a producer warp moves 64-key chunks (8 KB of K + 8 KB of V, contiguous in the
// [S][64] cache) into a two-stage shared-memory ring with cp.async.bulk (UBLKCP) signalled through mbarriers,
// and the four consumer warps read 16-byte vectors from shared memory.  The producer starts before the
// programmatic-launch wait: cache rows [0, len-1) were written by earlier steps, so they stream in while the
// preceding c_attn GEMM is still finishing.  grid (H, B), 160 threads, 32 KB smem -> one wave at B*H = 768.
constexpr int DK_THREADS = 160;

__device__ __forceinline__ uint32_t dk_smem(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ bool dk_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok) : "r"(dk_smem(bar)), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ void dk_wait(uint64_t* bar, uint32_t parity) {
  if (dk_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  while (!dk_try_wait(bar, parity)) {
    if (clock64() - t0 > 4000000000LL) {  // a protocol bug must abort the launch, not hang the GPU
      printf("tutor_b200 attn_decode: mbarrier wait timed out (block %d,%d)\n", blockIdx.x, blockIdx.y);
      __trap();
    }
  }
}

// DK_CHUNK keys per stage, DK_STAGES stages; MATH=false (microbench only) drains the ring without the arithmetic
template <bool FUSED, int DK_CHUNK, int DK_STAGES, bool MATH, bool FOLD = false>
__global__ void __launch_bounds__(DK_THREADS, 6) attn_decode_tma_kernel(const bf16* __restrict__ q, bf16* kc, bf16* vc,
                                                                    bf16* __restrict__ out, int H, int Smax,
                                                                    const int* __restrict__ len_dev, int len_host,
                                                                    const float* __restrict__ qkv_partial, int ks,
                                                                    const float* __restrict__ qkv_bias, int B,
                                                                    const QkvFold fold, const int* __restrict__ npad,
                                                                    int safe_rows) {
  constexpr int EPL = 8, LPR = 8, KPW = 4, NG = 16, UNR = DK_CHUNK / NG;
  // the ring is dynamic shared memory (2 * DK_STAGES * DK_CHUNK * 128 bytes: four stages exceed the static limit)
  extern __shared__ __align__(128) uint8_t dk_ring[];
  bf16(*sK)[DK_CHUNK * 64] = reinterpret_cast<bf16(*)[DK_CHUNK * 64]>(dk_ring);
  bf16(*sV)[DK_CHUNK * 64] = sK + DK_STAGES;
  __shared__ __align__(8) uint64_t full_bar[DK_STAGES], empty_bar[DK_STAGES];
  __shared__ float s_new[3][64];
  __shared__ float sm_m[NG], sm_l[NG], sm_o[NG][64];
  const int h = blockIdx.x, b = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // The length counter is bumped by the previous token's finalize kernel, which under programmatic dependent
  // launch may still be running when this kernel starts: it is read only AFTER the wait.  What may be streamed
  // before the wait are rows the host knows to be final: [0, safe_rows) (the prompt, written by prefill).
  const size_t base = ((size_t)b * H + h) * Smax * 64;
  const int kp = npad ? npad[b] : 0;  // left padding: the stream starts at the first real token's row
  const int early = min(DK_STAGES, max(0, safe_rows - kp) / DK_CHUNK);  // whole chunks that are safe to prefetch
  if (threadIdx.x == 0) {
    for (int s = 0; s < DK_STAGES; ++s) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(dk_smem(&full_bar[s])), "r"(1));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(dk_smem(&empty_bar[s])), "r"(4));
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  pdl_launch_dependents();

  if (warp == 4) {
    // ===== producer: one lane streams the cache rows of this (b,h) =====
    if (lane == 0) {
      auto copy_chunk = [&](int c, int rows) {
        const int s = c % DK_STAGES;
        const uint32_t bytes = (uint32_t)rows * 128;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(dk_smem(&full_bar[s])),
                     "r"(2 * bytes) : "memory");
        const size_t off = base + ((size_t)kp + (size_t)c * DK_CHUNK) * 64;
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                         dk_smem(&sK[s][0])), "l"(kc + off), "r"(bytes), "r"(dk_smem(&full_bar[s])) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                         dk_smem(&sV[s][0])), "l"(vc + off), "r"(bytes), "r"(dk_smem(&full_bar[s])) : "memory");
      };
      for (int c = 0; c < early; ++c) copy_chunk(c, DK_CHUNK);  // the ring is empty: no wait needed
      pdl_wait();
      const int len = len_dev ? *len_dev : len_host;
      const int nkeys = (FUSED ? len - 1 : len) - kp;
      const int nchunks = (nkeys + DK_CHUNK - 1) / DK_CHUNK;
      for (int c = early; c < nchunks; ++c) {
        dk_wait(&empty_bar[c % DK_STAGES], ((c / DK_STAGES) & 1) ^ 1);
        copy_chunk(c, min(DK_CHUNK, nkeys - c * DK_CHUNK));
      }
    }
  } else {
    // ===== consumers =====
    pdl_wait();  // q (or the c_attn partials) come from the preceding kernel; the length counter is final now
    const int len = len_dev ? *len_dev : len_host;
    const int nkeys = (FUSED ? len - 1 : len) - kp;
    const int nchunks = (nkeys + DK_CHUNK - 1) / DK_CHUNK;
    const int g = warp * KPW + lane / LPR;
    const int e0 = (lane % LPR) * EPL;
    float qv[EPL];
    if (FUSED) {
      // 192 values (q, k, v of this head) over 128 threads: both rounds' loads are issued before the first store, so
      // the prologue costs one L2 round trip, not two (it sits on the critical path right after the wait)
      const int d = H * 64;
      float accv[2];
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        const int e = threadIdx.x + r * 128;
        accv[r] = 0.f;
        if (e < 192) {
          const int n = (e >> 6) * d + h * 64 + (e & 63);
          accv[r] = qkv_from_partials<FOLD>(qkv_partial, ks, qkv_bias, fold, B, b, n, d);
        }
      }
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        const int e = threadIdx.x + r * 128;
        if (e < 192) {
          const int which = e >> 6, i = e & 63;
          if (which == 0) {
            s_new[0][i] = accv[r];
          } else {
            const bf16 st = __float2bfloat16_rn(accv[r]);
            (which == 1 ? kc : vc)[base + (size_t)(len - 1) * 64 + i] = st;  // cache append (no torch.cat)
            s_new[which][i] = __bfloat162float(st);
          }
        }
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");  // consumer warps only
#pragma unroll
      for (int i = 0; i < EPL; ++i) qv[i] = s_new[0][e0 + i] * 0.125f;
    } else {
      const bf16* qp = q + ((size_t)b * H + h) * 64 + e0;
#pragma unroll
      for (int i = 0; i < EPL; ++i) qv[i] = __bfloat162float(qp[i]) * 0.125f;
    }
    float m = -CUDART_INF_F, l = 0.f, o[EPL];
#pragma unroll
    for (int i = 0; i < EPL; ++i) o[i] = 0.f;

    for (int c = 0; c < nchunks; ++c) {
      const int s = c % DK_STAGES;
      dk_wait(&full_bar[s], (c / DK_STAGES) & 1);
      const int valid_rows = min(DK_CHUNK, nkeys - c * DK_CHUNK);
      float kf[UNR][EPL], vf[UNR][EPL], sc[UNR];
#pragma unroll
      for (int u = 0; u < UNR; ++u) {
        const int kk = u * NG + g;
        const int row = kk < valid_rows ? kk : 0;  // rows past the end are masked; row 0 of a chunk always exists
        unpack_bf16x8(*reinterpret_cast<const uint4*>(&sK[s][row * 64 + e0]), kf[u]);
        unpack_bf16x8(*reinterpret_cast<const uint4*>(&sV[s][row * 64 + e0]), vf[u]);
      }
      __syncwarp();
      if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(dk_smem(&empty_bar[s])) : "memory");
      if (!MATH) {
#pragma unroll
        for (int u = 0; u < UNR; ++u) o[u % EPL] += kf[u][0] + vf[u][1];
        continue;
      }
#pragma unroll
      for (int u = 0; u < UNR; ++u) {
        float d = 0.f;
#pragma unroll
        for (int i = 0; i < EPL; ++i) d = fmaf(qv[i], kf[u][i], d);
#pragma unroll
        for (int off = LPR / 2; off > 0; off >>= 1) d += __shfl_xor_sync(0xffffffffu, d, off);
        sc[u] = (u * NG + g < valid_rows) ? d : -CUDART_INF_F;
      }
      float mn = m;
#pragma unroll
      for (int u = 0; u < UNR; ++u) mn = fmaxf(mn, sc[u]);
      if (mn == -CUDART_INF_F) continue;
      const float corr = __expf(m - mn);
      l *= corr;
#pragma unroll
      for (int i = 0; i < EPL; ++i) o[i] *= corr;
#pragma unroll
      for (int u = 0; u < UNR; ++u) {
        const float p = __expf(sc[u] - mn);
        l += p;
#pragma unroll
        for (int i = 0; i < EPL; ++i) o[i] = fmaf(p, vf[u][i], o[i]);
      }
      m = mn;
    }
    if (FUSED) {  // the current token's own key/value, folded into group 0
      float d = 0.f;
#pragma unroll
      for (int i = 0; i < EPL; ++i) d = fmaf(qv[i], s_new[1][e0 + i], d);
#pragma unroll
      for (int off = LPR / 2; off > 0; off >>= 1) d += __shfl_xor_sync(0xffffffffu, d, off);
      if (g == 0) {
        const float mn = fmaxf(m, d);
        const float corr = __expf(m - mn), p = __expf(d - mn);
        l = l * corr + p;
#pragma unroll
        for (int i = 0; i < EPL; ++i) o[i] = fmaf(p, s_new[2][e0 + i], o[i] * corr);
        m = mn;
      }
    }
    if (lane % LPR == 0) { sm_m[g] = m; sm_l[g] = l; }
#pragma unroll
    for (int i = 0; i < EPL; ++i) sm_o[g][e0 + i] = o[i];
    asm volatile("bar.sync 1, 128;" ::: "memory");
    if (threadIdx.x < 64) {
      float M = -CUDART_INF_F;
#pragma unroll
      for (int s2 = 0; s2 < NG; ++s2) M = fmaxf(M, sm_m[s2]);
      float num = 0.f, den = 0.f;
#pragma unroll
      for (int s2 = 0; s2 < NG; ++s2) {
        const float w = (sm_m[s2] == -CUDART_INF_F) ? 0.f : __expf(sm_m[s2] - M);
        num = fmaf(w, sm_o[s2][threadIdx.x], num);
        den = fmaf(w, sm_l[s2], den);
      }
      out[((size_t)b * H + h) * 64 + threadIdx.x] = __float2bfloat16_rn(num / den);
    }
  }
}

// ----------------------------------------------------------------------------------- prefill
// grid (ceil(T/64), H, B), 64 threads: one query row per thread (q and the output accumulator live in
// registers), K/V tiles of 32 keys staged in shared memory as fp32 and read by broadcast.
// fp32 math throughout; exact expf in parity mode.
constexpr int PF_Q = 64, PF_KT = 32;

template <typename QT, typename KT, typename OT, bool FAST>
__global__ void __launch_bounds__(PF_Q) attn_prefill_kernel(const QT* __restrict__ q, const KT* __restrict__ kc,
                                                            const KT* __restrict__ vc, OT* __restrict__ out, int H,
                                                            int T, int Smax, int causal, const int* __restrict__ npad) {
  __shared__ __align__(16) float Ks[PF_KT][64];
  __shared__ __align__(16) float Vs[PF_KT][64];
  const int qt = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int tid = threadIdx.x;
  const int qi = qt * PF_Q + tid;
  const int qc = min(qi, T - 1);
  const int d = H * 64;
  // left padding: a real token (index >= kp) never sees the pad slots [0, kp); a pad slot keeps plain causal
  // visibility so that its (unused) row stays finite
  const int kp = npad ? npad[b] : 0;
  const int kmin = qc >= kp ? kp : 0;
  float qv[64], o[64];
  {
    const QT* qp = q + ((size_t)b * T + qc) * d + h * 64;
#pragma unroll
    for (int i = 0; i < 64; ++i) { qv[i] = to_f32(qp[i]) * 0.125f; o[i] = 0.f; }
  }
  float m = -CUDART_INF_F, l = 0.f;
  const size_t base = ((size_t)b * H + h) * Smax * 64;
  const int kend = causal ? min(T, (qt + 1) * PF_Q) : T;  // keys this block can ever need
  for (int k0 = 0; k0 < kend; k0 += PF_KT) {
    __syncthreads();
    constexpr int EPL = Vec16<KT>::N;
    for (int v = tid; v < PF_KT * 64 / EPL; v += PF_Q) {
      int r = v / (64 / EPL), c = (v % (64 / EPL)) * EPL;
      int j = min(k0 + r, T - 1);
      float f[EPL];
      Vec16<KT>::load(kc + base + (size_t)j * 64 + c, f);
#pragma unroll
      for (int i = 0; i < EPL; ++i) Ks[r][c + i] = f[i];
      Vec16<KT>::load(vc + base + (size_t)j * 64 + c, f);
#pragma unroll
      for (int i = 0; i < EPL; ++i) Vs[r][c + i] = f[i];
    }
    __syncthreads();
#pragma unroll
    for (int sub = 0; sub < PF_KT; sub += 16) {
      float s[16];
      float mn = m;
#pragma unroll
      for (int jj = 0; jj < 16; ++jj) {
        float acc = 0.f;
#pragma unroll
        for (int i = 0; i < 64; i += 4) {
          float4 kk = *reinterpret_cast<const float4*>(&Ks[sub + jj][i]);
          acc = fmaf(qv[i], kk.x, acc); acc = fmaf(qv[i + 1], kk.y, acc);
          acc = fmaf(qv[i + 2], kk.z, acc); acc = fmaf(qv[i + 3], kk.w, acc);
        }
        int j = k0 + sub + jj;
        bool ok = (j < T) && (!causal || j <= qc) && j >= kmin;
        s[jj] = ok ? acc : -CUDART_INF_F;
        mn = fmaxf(mn, s[jj]);
      }
      if (mn == -CUDART_INF_F) continue;  // nothing visible yet for this row in this sub-tile
      float corr = exp_sel<FAST>(m - mn);
      l *= corr;
#pragma unroll
      for (int i = 0; i < 64; ++i) o[i] *= corr;
#pragma unroll
      for (int jj = 0; jj < 16; ++jj) {
        float p = exp_sel<FAST>(s[jj] - mn);
        l += p;
#pragma unroll
        for (int i = 0; i < 64; i += 4) {
          float4 vv = *reinterpret_cast<const float4*>(&Vs[sub + jj][i]);
          o[i] = fmaf(p, vv.x, o[i]); o[i + 1] = fmaf(p, vv.y, o[i + 1]);
          o[i + 2] = fmaf(p, vv.z, o[i + 2]); o[i + 3] = fmaf(p, vv.w, o[i + 3]);
        }
      }
      m = mn;
    }
  }
  if (qi < T) {
    OT* op = out + ((size_t)b * T + qi) * d + h * 64;
    float inv = 1.0f / l;
#pragma unroll
    for (int i = 0; i < 64; ++i) op[i] = from_f32<OT>(o[i] * inv);
  }
}

// ------------------------------------------------------------------- prefill on tensor cores (bf16)
// Flash-attention-2 style: grid (ceil(T/64), H, B), 4 warps x 16 query rows.  Q lives in registers as
// mma.m16n8k16 A-fragments; K/V tiles of 64 keys are staged with cp.async into XOR-swizzled shared memory
// (double buffered) and read with ldmatrix (K) / ldmatrix.trans (V); S = QK^T and O += PV accumulate in fp32,
// the online softmax runs on the accumulator fragments and P is re-packed to bf16 A-fragments in registers.
// (The dense projections around it are the tcgen05 kernel; attention at these sizes - T <= 1024, hd 64 - is
// ~5 % of the prefill FLOPs, so the warp-level tensor path is enough to take it off the critical path.)
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem));
}
__device__ __forceinline__ void ldmatrix_x4(uint32_t* r, const void* smem) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"((uint32_t)__cvta_generic_to_shared(smem)));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t* r, const void* smem) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"((uint32_t)__cvta_generic_to_shared(smem)));
}
__device__ __forceinline__ void mma_bf16_16816(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ float fast_exp2(float x) {  // one MUFU.EX2; exp2(-inf) = 0
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 p = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&p);
}

template <bool CAUSAL, bool PADDED>
__global__ void __launch_bounds__(128) attn_prefill_mma_kernel(const bf16* __restrict__ q, const bf16* __restrict__ kc,
                                                                const bf16* __restrict__ vc, bf16* __restrict__ out,
                                                                int H, int T, int Smax, const int* __restrict__ npad) {
  // 128 queries per CTA, 32 per warp (two m16 row blocks): every K/V fragment fetched from shared memory with
  // ldmatrix feeds two MMAs, which halves the shared-memory traffic per FLOP of the 16-rows-per-warp layout
  constexpr int RB = 2;
  __shared__ __align__(128) uint8_t tiles[2][2][64 * 128];  // [stage][K|V][64 keys x 128 B], 16-byte chunks XOR-swizzled
  const int qt = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int d = H * 64;
  const int q0 = qt * 128 + warp * 32;
  uint32_t qf[RB][4][4];
#pragma unroll
  for (int rb = 0; rb < RB; ++rb) {
    const bf16* p0 = q + ((size_t)b * T + min(q0 + rb * 16 + g, T - 1)) * d + h * 64 + 2 * t;
    const bf16* p1 = q + ((size_t)b * T + min(q0 + rb * 16 + g + 8, T - 1)) * d + h * 64 + 2 * t;
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      qf[rb][kk][0] = *reinterpret_cast<const uint32_t*>(p0 + kk * 16);
      qf[rb][kk][1] = *reinterpret_cast<const uint32_t*>(p1 + kk * 16);
      qf[rb][kk][2] = *reinterpret_cast<const uint32_t*>(p0 + kk * 16 + 8);
      qf[rb][kk][3] = *reinterpret_cast<const uint32_t*>(p1 + kk * 16 + 8);
    }
  }
  float o[RB][8][4];
  float mrow[RB][2], lrow[RB][2];
#pragma unroll
  for (int rb = 0; rb < RB; ++rb) {
#pragma unroll
    for (int j = 0; j < 8; ++j) { o[rb][j][0] = o[rb][j][1] = o[rb][j][2] = o[rb][j][3] = 0.f; }
    mrow[rb][0] = mrow[rb][1] = -CUDART_INF_F;
    lrow[rb][0] = lrow[rb][1] = 0.f;
  }
  const size_t base = ((size_t)b * H + h) * Smax * 64;
  const int nkt_all = (T + 63) / 64;
  const int nkt = CAUSAL ? min(2 * (qt + 1), nkt_all) : nkt_all;
  const int kp = PADDED ? npad[b] : 0;  // left padding, see attn_prefill_kernel (compiled out when unpadded)

  auto load_tile = [&](int kt, int stage) {
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int idx = threadIdx.x + 128 * r;
      const int row = idx >> 3, ch = idx & 7;
      const int key = min(kt * 64 + row, T - 1);
      const size_t src = base + (size_t)key * 64 + ch * 8;
      const int dst = row * 128 + ((ch ^ (row & 7)) << 4);
      cp_async16(&tiles[stage][0][dst], kc + src);
      cp_async16(&tiles[stage][1][dst], vc + src);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };

  // Raw scores stay unscaled; the 1/8 scale and log2(e) are folded into one FFMA in front of ex2.  Masking
  // (sequence end, causal diagonal, left padding) is only evaluated on tiles that can contain a masked key -
  // the per-element scalar work of the softmax, not the MMAs, bounds this kernel.
  constexpr float C = 0.125f * 1.4426950408889634f;
  load_tile(0, 0);
  for (int kt = 0; kt < nkt; ++kt) {
    const int stage = kt & 1;
    if (kt + 1 < nkt) {
      load_tile(kt + 1, stage ^ 1);
      asm volatile("cp.async.wait_group 1;" ::: "memory");
    } else {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncthreads();
    const uint8_t* sK = tiles[stage][0];
    const uint8_t* sV = tiles[stage][1];
    // causal: this warp's rows are q0..q0+31; tiles entirely above its diagonal contribute nothing
    const bool tile_dead = CAUSAL && kt * 64 > q0 + 31;
    if (!tile_dead) {
      float s[RB][8][4];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        uint32_t kb[8];
        const int row = j * 8 + (lane & 7);
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          const int ch = half * 4 + (lane >> 3);
          ldmatrix_x4(&kb[half * 4], sK + row * 128 + ((ch ^ (row & 7)) << 4));
        }
#pragma unroll
        for (int rb = 0; rb < RB; ++rb) {
          s[rb][j][0] = s[rb][j][1] = s[rb][j][2] = s[rb][j][3] = 0.f;
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) mma_bf16_16816(s[rb][j], qf[rb][kk], kb[2 * kk], kb[2 * kk + 1]);
        }
      }
      const bool tile_masked = (kt * 64 + 63 >= T) || (CAUSAL && kt * 64 + 63 > q0) || (PADDED && kt * 64 < kp);
      uint32_t pa[RB][4][4];
#pragma unroll
      for (int rb = 0; rb < RB; ++rb) {
        if (tile_masked) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
#pragma unroll
            for (int r = 0; r < 4; ++r) {
              const int key = kt * 64 + j * 8 + 2 * t + (r & 1);
              const int qrow = q0 + rb * 16 + g + (r >> 1) * 8;
              if (key >= T || (CAUSAL && key > qrow) || (PADDED && key < kp && qrow >= kp)) s[rb][j][r] = -CUDART_INF_F;
            }
          }
        }
        float mx0 = -CUDART_INF_F, mx1 = -CUDART_INF_F;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          mx0 = fmaxf(mx0, fmaxf(s[rb][j][0], s[rb][j][1]));
          mx1 = fmaxf(mx1, fmaxf(s[rb][j][2], s[rb][j][3]));
        }
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1)); mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1)); mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
        // key 0 is visible to every row in tile 0, so the running maxima are finite from the first tile on
        const float mn0 = fmaxf(mrow[rb][0], mx0), mn1 = fmaxf(mrow[rb][1], mx1);
        const float c0 = fast_exp2((mrow[rb][0] - mn0) * C), c1 = fast_exp2((mrow[rb][1] - mn1) * C);
        const float b0 = mn0 * C, b1 = mn1 * C;
        float l0 = lrow[rb][0] * c0, l1 = lrow[rb][1] * c1;
#pragma unroll
        for (int j = 0; j < 8; ++j) { o[rb][j][0] *= c0; o[rb][j][1] *= c0; o[rb][j][2] *= c1; o[rb][j][3] *= c1; }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float p0 = fast_exp2(fmaf(s[rb][j][0], C, -b0)), p1 = fast_exp2(fmaf(s[rb][j][1], C, -b0));
          const float p2 = fast_exp2(fmaf(s[rb][j][2], C, -b1)), p3 = fast_exp2(fmaf(s[rb][j][3], C, -b1));
          l0 += p0 + p1; l1 += p2 + p3;
          pa[rb][j >> 1][(j & 1) * 2] = pack_bf16x2(p0, p1);
          pa[rb][j >> 1][(j & 1) * 2 + 1] = pack_bf16x2(p2, p3);
        }
        lrow[rb][0] = l0; lrow[rb][1] = l1;
        mrow[rb][0] = mn0; mrow[rb][1] = mn1;
      }
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
        for (int jp = 0; jp < 4; ++jp) {
          const int mi = lane >> 3;
          const int krow = kk * 16 + (mi & 1) * 8 + (lane & 7);
          const int ch = 2 * jp + (mi >> 1);
          uint32_t vb[4];
          ldmatrix_x4_trans(vb, sV + krow * 128 + ((ch ^ (krow & 7)) << 4));
#pragma unroll
          for (int rb = 0; rb < RB; ++rb) {
            mma_bf16_16816(o[rb][2 * jp], pa[rb][kk], vb[0], vb[1]);
            mma_bf16_16816(o[rb][2 * jp + 1], pa[rb][kk], vb[2], vb[3]);
          }
        }
      }
    }
    __syncthreads();  // everyone is done with this stage before the next iteration's cp.async refills it
  }
#pragma unroll
  for (int rb = 0; rb < RB; ++rb) {
    float l0 = lrow[rb][0], l1 = lrow[rb][1];
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    const float i0 = 1.0f / l0, i1 = 1.0f / l1;
    const int r0 = q0 + rb * 16 + g, r1 = r0 + 8;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (r0 < T)
        *reinterpret_cast<uint32_t*>(out + ((size_t)b * T + r0) * d + h * 64 + j * 8 + 2 * t) =
            pack_bf16x2(o[rb][j][0] * i0, o[rb][j][1] * i0);
      if (r1 < T)
        *reinterpret_cast<uint32_t*>(out + ((size_t)b * T + r1) * d + h * 64 + j * 8 + 2 * t) =
            pack_bf16x2(o[rb][j][2] * i1, o[rb][j][3] * i1);
    }
  }
}


#ifdef TUTOR_EXPERIMENTS  // measured no faster than the SIMT consumers (see use_mma_decode): not in the shipped library
// ------------------------------------------------------------------- decode on tensor cores (bf16)
// Same producer / ring structure as attn_decode_tma_kernel, with the arithmetic moved to mma.sync m16n8k16: the
// SIMT consumers needed ~50 instructions per key per lane (unpack, FMA, shuffles) and held the kernel at 80 % of
// the measured HBM peak while the same ring drained without arithmetic reached 94 %.  Here the query is row 0 of
// a 16-row A operand (rows 1..15 are zero - the tensor pipe is idle anyway), each 64-key chunk costs a warp
// 16 ldmatrix + 32 MMAs, and the softmax touches 4 scores per lane.  K/V chunks arrive through
// cp.async.bulk.tensor with the 128-byte swizzle, so ldmatrix (K) and ldmatrix.trans (V) are conflict-free.
// 96 threads: warps 0-1 consume (32 keys of every chunk each), warp 2 produces.  grid (H, B), 6 CTAs per SM.
constexpr int DM_THREADS = 96, DM_CHUNK = 32, DM_STAGES = 4;

// D rows 0-7 += A x B; the accumulator rows 8-15 (always zero here: only row 0 of A is populated) are fed as
// constants and their results dropped, so they cost no registers
__device__ __forceinline__ void mma_bf16_16816_top(float& c0, float& c1, uint32_t a0, uint32_t a2, uint32_t b0, uint32_t b1) {
  float d2, d3;
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%8,%5,%8}, {%6,%7}, {%0,%1,%9,%9};"
      : "+f"(c0), "+f"(c1), "=f"(d2), "=f"(d3)
      : "r"(a0), "r"(a2), "r"(b0), "r"(b1), "r"(0u), "f"(0.f));
}

// MATH=false (microbench only) drains the ring without the arithmetic
template <bool FUSED, bool MATH>
__global__ void __launch_bounds__(DM_THREADS, 6) attn_decode_mma_kernel(
    const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV, const bf16* __restrict__ q, bf16* kc,
    bf16* vc, bf16* __restrict__ out, int H, int Smax, const int* __restrict__ len_dev, int len_host,
    const float* __restrict__ qkv_partial, int ks, const float* __restrict__ qkv_bias, int B,
    const int* __restrict__ npad, int safe_rows, int split_rows, float* split_scratch, int* split_counters) {
  __shared__ __align__(1024) uint8_t sK[DM_STAGES][DM_CHUNK * 128];
  __shared__ __align__(1024) uint8_t sV[DM_STAGES][DM_CHUNK * 128];
  __shared__ __align__(8) uint64_t full_bar[DM_STAGES], empty_bar[DM_STAGES];
  __shared__ float s_new[3][64];
  __shared__ float sm_m[2], sm_l[2], sm_o[2][64], s_self;
  const int h = blockIdx.x, b = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row_base = (b * H + h) * Smax;  // first cache row of this (b,h) in the [B*H*Smax, 64] view
  const size_t base = (size_t)row_base * 64;
  const int kp = npad ? npad[b] : 0;
  // gridDim.z > 1: the key range of every (b,h) is cut at multiples of split_rows (a host-known constant, so the
  // rows a split may prefetch ahead of the programmatic-launch wait do not depend on the device-side length).
  // 768 equal CTAs on 148 SMs leave 28 SMs with six streams and 120 with five - a 16 % tail; with 2-4 splits the
  // block scheduler hands out 1536-3072 smaller units as slots free up.  The last CTA to finish a (b,h) merges the
  // split states in split order (deterministic).
  const int z = blockIdx.z, nsplit = gridDim.z;
  const int kbeg = z * split_rows;
  const int early = min(DM_STAGES, max(0, safe_rows - kp - kbeg) / DM_CHUNK);  // see attn_decode_tma_kernel
  if (threadIdx.x == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmK) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmV) : "memory");
    for (int s = 0; s < DM_STAGES; ++s) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(dk_smem(&full_bar[s])), "r"(1));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(dk_smem(&empty_bar[s])), "r"(2));
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();
  pdl_launch_dependents();

  if (warp == 2) {
    // ===== producer: one lane streams the cache rows of this (b,h), 32 keys (4 KB of K + 4 KB of V) per stage =====
    if (lane == 0) {
      auto copy_chunk = [&](int c) {
        const int s = c % DM_STAGES;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(dk_smem(&full_bar[s])),
                     "r"(2 * DM_CHUNK * 128) : "memory");
        const int row = row_base + kp + kbeg + c * DM_CHUNK;  // rows past the end of the tensor are zero-filled
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
                         dk_smem(&sK[s][0])), "l"(&tmK), "r"(dk_smem(&full_bar[s])), "r"(0), "r"(row) : "memory");
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
                         dk_smem(&sV[s][0])), "l"(&tmV), "r"(dk_smem(&full_bar[s])), "r"(0), "r"(row) : "memory");
      };
      for (int c = 0; c < early; ++c) copy_chunk(c);  // the ring is empty: no wait needed
      pdl_wait();
      const int len = len_dev ? *len_dev : len_host;
      const int nk_all = (FUSED ? len - 1 : len) - kp;
      const int nkeys = max(0, (z == nsplit - 1 ? nk_all : min(nk_all, kbeg + split_rows)) - kbeg);
      const int nchunks = (nkeys + DM_CHUNK - 1) / DM_CHUNK;
      for (int c = early; c < nchunks; ++c) {
        dk_wait(&empty_bar[c % DM_STAGES], ((c / DM_STAGES) & 1) ^ 1);
        copy_chunk(c);
      }
    }
    return;
  }

  // ===== consumers (64 threads): warp w owns keys [16w, 16w+16) of every chunk =====
  pdl_wait();  // q (or the c_attn partials) come from the preceding kernel; the length counter is final now
  const int len = len_dev ? *len_dev : len_host;
  const int nk_all = (FUSED ? len - 1 : len) - kp;
  const int nkeys = max(0, (z == nsplit - 1 ? nk_all : min(nk_all, kbeg + split_rows)) - kbeg);
  const int nchunks = (nkeys + DM_CHUNK - 1) / DM_CHUNK;
  const int g = lane >> 2, t = lane & 3;
  const bool self_here = FUSED && z == nsplit - 1;  // the split that folds in the current token's own key / value
  if (FUSED) {
    const int d = H * 64;
    for (int e = threadIdx.x; e < 192; e += 64) {
      const int which = e >> 6, i = e & 63;
      const int n = which * d + h * 64 + i;
      float acc = qkv_bias[n];
#pragma unroll 8
      for (int zz = 0; zz < ks; ++zz) acc += qkv_partial[((size_t)zz * B + b) * (3 * d) + n];
      const bf16 st = __float2bfloat16_rn(which == 0 ? acc * 0.125f : acc);  // q carries the 1/sqrt(64) scale (exact)
      if (which != 0 && z == 0) (which == 1 ? kc : vc)[base + (size_t)(len - 1) * 64 + i] = st;  // cache append (no torch.cat)
      s_new[which][i] = __bfloat162float(st);
    }
  } else {
    const int i = threadIdx.x;
    s_new[0][i] = __bfloat162float(__float2bfloat16_rn(__bfloat162float(q[((size_t)b * H + h) * 64 + i]) * 0.125f));
  }
  asm volatile("bar.sync 1, 64;" ::: "memory");  // consumer warps only
  // A fragments of the query: row 0 of the 16 x 16 tile is q, rows 1..15 are zero
  uint32_t qa0[4], qa2[4];
#pragma unroll
  for (int kk = 0; kk < 4; ++kk) {
    qa0[kk] = g == 0 ? pack_bf16x2(s_new[0][kk * 16 + 2 * t], s_new[0][kk * 16 + 2 * t + 1]) : 0u;
    qa2[kk] = g == 0 ? pack_bf16x2(s_new[0][kk * 16 + 8 + 2 * t], s_new[0][kk * 16 + 8 + 2 * t + 1]) : 0u;
  }
  constexpr float LOG2E = 1.4426950408889634f;
  float m = -CUDART_INF_F, l = 0.f, o[8][2];
#pragma unroll
  for (int j = 0; j < 8; ++j) o[j][0] = o[j][1] = 0.f;
  const int mi = lane >> 3, lr = lane & 7;
  const int key0 = warp * 16;
  const int krow = key0 + (mi >> 1) * 8 + lr, vrow = key0 + (mi & 1) * 8 + lr;

  for (int c = 0; c < nchunks; ++c) {
    const int s = c % DM_STAGES;
    dk_wait(&full_bar[s], (c / DM_STAGES) & 1);
    const int valid_rows = min(DM_CHUNK, nkeys - c * DM_CHUNK);
    uint32_t kb[16], vb[16];
    const uint8_t* kp8 = sK[s] + krow * 128;
    const uint8_t* vp8 = sV[s] + vrow * 128;
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) ldmatrix_x4(&kb[kk * 4], kp8 + (((2 * kk + (mi & 1)) ^ (krow & 7)) << 4));
#pragma unroll
    for (int p = 0; p < 4; ++p) ldmatrix_x4_trans(&vb[p * 4], vp8 + (((2 * p + (mi >> 1)) ^ (vrow & 7)) << 4));
    __syncwarp();  // this warp's 16 keys are in registers: hand the stage back to the producer
    if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(dk_smem(&empty_bar[s])) : "memory");
    if (!MATH) {
#pragma unroll
      for (int i = 0; i < 16; ++i) o[i & 7][0] += __uint_as_float(kb[i] ^ vb[i]);
      continue;
    }
    if (key0 >= valid_rows) continue;  // warp-uniform: nothing but masked keys for this warp in the last chunk
    float sc[2][2];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      sc[j][0] = sc[j][1] = 0.f;
#pragma unroll
      for (int kk = 0; kk < 4; ++kk)
        mma_bf16_16816_top(sc[j][0], sc[j][1], qa0[kk], qa2[kk], kb[kk * 4 + 2 * j], kb[kk * 4 + 2 * j + 1]);
    }
    if (key0 + 16 > valid_rows) {
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        if (key0 + j * 8 + 2 * t >= valid_rows) sc[j][0] = -CUDART_INF_F;
        if (key0 + j * 8 + 2 * t + 1 >= valid_rows) sc[j][1] = -CUDART_INF_F;
      }
    }
    float mx = fmaxf(fmaxf(sc[0][0], sc[0][1]), fmaxf(sc[1][0], sc[1][1]));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
    const float mn = fmaxf(m, mx);  // finite: key0 < valid_rows, so at least one key of the row is unmasked
    const float corr = fast_exp2((m - mn) * LOG2E), nb = mn * LOG2E;
    const float p00 = fast_exp2(fmaf(sc[0][0], LOG2E, -nb)), p01 = fast_exp2(fmaf(sc[0][1], LOG2E, -nb));
    const float p10 = fast_exp2(fmaf(sc[1][0], LOG2E, -nb)), p11 = fast_exp2(fmaf(sc[1][1], LOG2E, -nb));
    l = l * corr + (p00 + p01) + (p10 + p11);
    m = mn;
    const uint32_t pa0 = pack_bf16x2(p00, p01), pa2 = pack_bf16x2(p10, p11);
#pragma unroll
    for (int j = 0; j < 8; ++j) { o[j][0] *= corr; o[j][1] *= corr; }
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      mma_bf16_16816_top(o[2 * p][0], o[2 * p][1], pa0, pa2, vb[p * 4], vb[p * 4 + 1]);
      mma_bf16_16816_top(o[2 * p + 1][0], o[2 * p + 1][1], pa0, pa2, vb[p * 4 + 2], vb[p * 4 + 3]);
    }
  }
  // row 0 (the query) lives in lanes 0..3 of each consumer warp
  l += __shfl_xor_sync(0xffffffffu, l, 1);
  l += __shfl_xor_sync(0xffffffffu, l, 2);
  if (g == 0) {
    if (t == 0) { sm_m[warp] = m; sm_l[warp] = l; }
#pragma unroll
    for (int j = 0; j < 8; ++j) { sm_o[warp][j * 8 + 2 * t] = o[j][0]; sm_o[warp][j * 8 + 2 * t + 1] = o[j][1]; }
  }
  if (self_here && warp == 0) {  // score of the current token's own key (its row is not part of the streamed range)
    const float d = warp_sum(s_new[0][2 * lane] * s_new[1][2 * lane] + s_new[0][2 * lane + 1] * s_new[1][2 * lane + 1]);
    if (lane == 0) s_self = d;
  }
  asm volatile("bar.sync 1, 64;" ::: "memory");
  const int i = threadIdx.x;  // output dim
  float M = fmaxf(sm_m[0], sm_m[1]);
  if (self_here) M = fmaxf(M, s_self);
  float num = 0.f, den = 0.f;
  if (M != -CUDART_INF_F) {  // (a split past the end of a short sequence has seen no key at all)
    const float w0 = sm_m[0] == -CUDART_INF_F ? 0.f : fast_exp2((sm_m[0] - M) * LOG2E);
    const float w1 = sm_m[1] == -CUDART_INF_F ? 0.f : fast_exp2((sm_m[1] - M) * LOG2E);
    num = w0 * sm_o[0][i] + w1 * sm_o[1][i];
    den = w0 * sm_l[0] + w1 * sm_l[1];
    if (self_here) {
      const float ws = fast_exp2((s_self - M) * LOG2E);
      num = fmaf(ws, s_new[2][i], num);
      den += ws;
    }
  }
  if (nsplit == 1) {
    out[((size_t)b * H + h) * 64 + i] = __float2bfloat16_rn(num / den);
    return;
  }
  // publish this split's state [max, denominator, 64 numerators]; the last split to arrive merges them all
  float* st = split_scratch + (((size_t)b * H + h) * nsplit + z) * 66;
  if (i == 0) { st[0] = M; st[1] = den; }
  st[2 + i] = num;
  __threadfence();
  asm volatile("bar.sync 1, 64;" ::: "memory");
  if (i == 0) {
    const int done = atomicAdd(&split_counters[b * H + h], 1);
    const int last = done == nsplit - 1;
    if (last) split_counters[b * H + h] = 0;  // ready for the next launch
    sm_m[0] = last ? 1.f : 0.f;
  }
  asm volatile("bar.sync 1, 64;" ::: "memory");
  if (sm_m[0] != 0.f) {
    __threadfence();
    const float* sp = split_scratch + ((size_t)b * H + h) * nsplit * 66;
    float MM = -CUDART_INF_F;
    for (int s2 = 0; s2 < nsplit; ++s2) MM = fmaxf(MM, __ldcg(sp + s2 * 66));
    float nn = 0.f, dd = 0.f;
    for (int s2 = 0; s2 < nsplit; ++s2) {  // fixed split order: deterministic
      const float ms = __ldcg(sp + s2 * 66);
      const float w = (ms == -CUDART_INF_F) ? 0.f : fast_exp2((ms - MM) * LOG2E);
      nn = fmaf(w, __ldcg(sp + s2 * 66 + 2 + i), nn);
      dd = fmaf(w, __ldcg(sp + s2 * 66 + 1), dd);
    }
    out[((size_t)b * H + h) * 64 + i] = __float2bfloat16_rn(nn / dd);
  }
}

// TUTOR_ATTN_MMA=1 selects the tensor-core consumers (attn_decode_mma_kernel).  Measured on B200 at B=64, H=12,
// S=640: 24.4 us per launch against 23.4 us for the SIMT consumers of attn_decode_tma_kernel, and 0.638 against
// 0.612 ms per decode forward; with 2 / 3 / 4 key-range splits 0.689 / 0.729 / 0.776 ms.  Both kernels drain the
// same ring in 20.4-20.6 us without arithmetic and take 27.1 us under ncu: what is left is the per-CTA prologue and
// epilogue latency of a one-wave launch, which neither fewer instructions nor more (smaller) CTAs remove.
bool use_mma_decode() {
  static const bool mma = tutor_experiment("TUTOR_ATTN_MMA") != nullptr;
  return mma;
}

#endif  // TUTOR_EXPERIMENTS
}  // namespace

// number of key-range splits: enough CTAs for ~8 per SM, at least 64 keys per split
static int num_sms_cached() {
  static const int n = [] {
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    return sms;
  }();
  return n;
}

int k_attn_decode_splits(int B, int H, int len_hint, int num_sms) {
  // fewer (b,h) streams than SMs: split until every SM has work
  int n = 1;
  while (n < 8 && B * H * n < num_sms && len_hint / (2 * n) >= 64) n *= 2;
  // more streams than SMs: one ring CTA per (sequence, head), whose ring is four stages deep while at most three of them
  // share an SM (k_attn_decode_fused).  GPT-2 large, cache ~960, decode forward with that against the best key-range split
  // of the register kernel (3 splits at B = 16): B = 8: 1440 vs 1465 us, 12: 1499 vs 1552, 16: 1580 vs 1722, 20: 1733 vs
  // 2005, 24 (two stages): 1963 vs 2089.  B*H = 768 (GPT-2 small, B = 64) is 5.2 two-stage CTAs per SM and HBM-bound.
  if ((long)B * H > num_sms) return 1;
  return n;
}
size_t k_attn_decode_scratch_bytes(int B, int H) { return (size_t)B * H * (8 * 66 * sizeof(float) + sizeof(int)); }

int k_attn_decode(const void* q, int q_dtype, const void* k, const void* v, int kv_dtype, void* out, int out_dtype,
                  int B, int H, int Smax, const int* len_dev, int len_host, const int* npad, cudaStream_t st) {
  if (B == 0) return TUTOR_OK;
  ARG_CHECK(len_dev || (len_host >= 1 && len_host <= Smax), "attn_decode: len %d outside [1,%d]", len_host, Smax);
  dim3 grid(H, B), block(128);
  // the fp32 kernel is launched without the programmatic-launch attribute unless TUTOR_ATTN_PDL=1: with it the batch-1
  // decode forward measured 392 us against 381 us (its dependents' early weight loads compete with this kernel)
  static const bool f32_plain = tutor_experiment("TUTOR_ATTN_PDL") == nullptr;
  if (q_dtype == DT_F32 && kv_dtype == DT_F32 && out_dtype == DT_F32 && f32_plain)
    attn_decode_kernel<float, float, float, false, false><<<grid, block, 0, st>>>(
        (const float*)q, (float*)k, (float*)v, (float*)out, H, Smax, len_dev, len_host, nullptr, 0, nullptr, B, QkvFold{}, nullptr,
        nullptr, npad);
  else if (q_dtype == DT_F32 && kv_dtype == DT_F32 && out_dtype == DT_F32)
    // (through tutor_launch: part of the fp32 decode chain's programmatic dependent launches; the kernel waits first)
    CUDA_TRY(tutor_launch(attn_decode_kernel<float, float, float, false, false>, grid, block, 0, st, (const float*)q,
                          (float*)k, (float*)v, (float*)out, H, Smax, len_dev, len_host, (const float*)nullptr, 0,
                          (const float*)nullptr, B, QkvFold{}, (float*)nullptr, (int*)nullptr, npad));
  else if (q_dtype == DT_BF16 && kv_dtype == DT_BF16 && out_dtype == DT_BF16) {
#ifdef TUTOR_EXPERIMENTS
    if (use_mma_decode()) {
      CUtensorMap tk, tv;  // the cache as a [B*H*Smax, 64] bf16 matrix, 64-row boxes, 128-byte swizzle
      TRY(k_make_map_bf16(&tk, (const bf16*)k, B * H * Smax, 64, DM_CHUNK));
      TRY(k_make_map_bf16(&tv, (const bf16*)v, B * H * Smax, 64, DM_CHUNK));
      attn_decode_mma_kernel<false, true><<<grid, DM_THREADS, 0, st>>>(tk, tv, (const bf16*)q, (bf16*)k, (bf16*)v, (bf16*)out, H,
                                                                 Smax, len_dev, len_host, nullptr, 0, nullptr, B, npad, 0,
                                                                 0, nullptr, nullptr);
    } else
#endif
    {
      attn_decode_tma_kernel<false, 64, 2, true><<<grid, DK_THREADS, 2 * 2 * 64 * 128, st>>>((const bf16*)q, (bf16*)k, (bf16*)v, (bf16*)out, H, Smax,
                                                                 len_dev, len_host, nullptr, 0, nullptr, B, QkvFold{}, npad, 0);
    }
  } else {
    tutor_set_error("attn_decode: unsupported dtype combination q=%d kv=%d out=%d", q_dtype, kv_dtype, out_dtype);
    return TUTOR_ERR_ARG;
  }
  KERNEL_LAUNCHED();
  return TUTOR_OK;
}

int k_attn_decode_fused(const float* qkv_partial, int ks, const float* qkv_bias, const QkvFold& fold, void* k, void* v,
                        int kv_dtype, void* out, int out_dtype, int B, int H, int Smax, const int* len_dev, int len_host,
                        int nsplit, void* split_scratch, const int* npad, int safe_rows, int len_hint,
                        cudaStream_t st) {
  if (B == 0) return TUTOR_OK;
  ARG_CHECK(kv_dtype == DT_BF16 && out_dtype == DT_BF16, "attn_decode_fused: bf16 cache and output only");
  ARG_CHECK(qkv_partial && qkv_bias && ks >= 1, "attn_decode_fused: bad partials");
  ARG_CHECK(!fold.s || (fold.c && fold.stats && fold.nst >= 1 && fold.nst <= 16), "attn_decode_fused: bad folded-LayerNorm operands");
  ARG_CHECK(len_dev || (len_host >= 1 && len_host <= Smax), "attn_decode_fused: len %d outside [1,%d]", len_host, Smax);
  ARG_CHECK(nsplit >= 1 && nsplit <= 8 && (nsplit == 1 || split_scratch), "attn_decode_fused: bad split setup");
  float* sc = reinterpret_cast<float*>(split_scratch);
  int* counters = reinterpret_cast<int*>(sc + (size_t)B * H * 8 * 66);  // zeroed by the owner; self-resetting
#ifdef TUTOR_EXPERIMENTS
  if (use_mma_decode()) {
    ARG_CHECK(!fold.s, "attn_decode_fused: the mma.sync experiment has no folded-LayerNorm path");
    CUtensorMap tk, tv;
    TRY(k_make_map_bf16(&tk, (const bf16*)k, B * H * Smax, 64, DM_CHUNK));
    TRY(k_make_map_bf16(&tv, (const bf16*)v, B * H * Smax, 64, DM_CHUNK));
    // split boundaries are fixed multiples of the chunk size derived from the expected length, not the live one
    const int split_rows = (int)round_up((size_t)ceil_div(len_hint > 0 ? len_hint : Smax, nsplit), DM_CHUNK);
    CUDA_TRY(tutor_launch(attn_decode_mma_kernel<true, true>, dim3(H, B, nsplit), dim3(DM_THREADS), 0, st, tk, tv,
                          (const bf16*)nullptr, (bf16*)k, (bf16*)v, (bf16*)out, H, Smax, len_dev, len_host, qkv_partial, ks,
                          qkv_bias, B, npad, safe_rows, split_rows, sc, counters));
    KERNEL_LAUNCHED();
    return TUTOR_OK;
  }
#endif
  if (nsplit == 1) {
    // few (sequence, head) streams per SM (GPT-2 large at B = 16: 320 CTAs, 2.2 per SM): each CTA is bound by the latency
    // of its own ring, not by HBM, so the ring is as deep as the number of CTAs that must share an SM allows: four stages
    // up to 3 CTAs per SM, three up to 4, else two (6 per SM)
    const long bh = (long)B * H, sms = num_sms_cached();
    const int stages = bh <= 3 * sms ? 4 : bh <= 4 * sms ? 3 : 2;
    auto kern = fold.s ? attn_decode_tma_kernel<true, 64, 2, true, true> : attn_decode_tma_kernel<true, 64, 2, true, false>;
    if (stages == 4) kern = fold.s ? attn_decode_tma_kernel<true, 64, 4, true, true> : attn_decode_tma_kernel<true, 64, 4, true, false>;
    if (stages == 3) kern = fold.s ? attn_decode_tma_kernel<true, 64, 3, true, true> : attn_decode_tma_kernel<true, 64, 3, true, false>;
    const int ring_bytes = 2 * stages * 64 * 128;
    if (stages > 2) TRY(tutor_set_max_smem(reinterpret_cast<const void*>(kern), ring_bytes));
    CUDA_TRY(tutor_launch(kern, dim3(H, B), dim3(DK_THREADS), (size_t)ring_bytes, st, (const bf16*)nullptr, (bf16*)k, (bf16*)v,
                          (bf16*)out, H, Smax, len_dev, len_host, qkv_partial, ks, qkv_bias, B, fold, npad, safe_rows));
    KERNEL_LAUNCHED();
    return TUTOR_OK;
  }
  auto kern = fold.s ? attn_decode_kernel<bf16, bf16, bf16, true, true, true> : attn_decode_kernel<bf16, bf16, bf16, true, true, false>;
  CUDA_TRY(tutor_launch(kern, dim3(H, B, nsplit), dim3(128), 0, st, (const bf16*)nullptr, (bf16*)k, (bf16*)v, (bf16*)out, H,
                        Smax, len_dev, len_host, qkv_partial, ks, qkv_bias, B, fold, sc, counters, npad));
  KERNEL_LAUNCHED();
  return TUTOR_OK;
}

int k_attn_prefill(const void* q, int q_dtype, const void* k, const void* v, int kv_dtype, void* out, int out_dtype,
                   int B, int H, int T, int Smax, int causal, const int* npad, cudaStream_t st) {
  if (B == 0 || T == 0) return TUTOR_OK;
  ARG_CHECK(T <= Smax, "attn_prefill: T=%d > Smax=%d", T, Smax);
  dim3 grid(ceil_div(T, PF_Q), H, B), block(PF_Q);
  if (q_dtype == DT_F32 && kv_dtype == DT_F32 && out_dtype == DT_F32)
    attn_prefill_kernel<float, float, float, false><<<grid, block, 0, st>>>(
        (const float*)q, (const float*)k, (const float*)v, (float*)out, H, T, Smax, causal, npad);
  else if (q_dtype == DT_BF16 && kv_dtype == DT_BF16 && out_dtype == DT_BF16) {
    // tensor cores of the 5th generation by default; TUTOR_ATTN_PREFILL_MMA=1 keeps the mma.sync kernel as a cross-check
    static const bool legacy = getenv("TUTOR_ATTN_PREFILL_MMA") != nullptr;
    if (!legacy) return k_attn_prefill_tc((const bf16*)q, (const bf16*)k, (const bf16*)v, (bf16*)out, B, H, T, Smax, causal, npad, st);
    dim3 g2(ceil_div(T, 128), H, B);
#define PF_LAUNCH(C, P)                                                                                          \
  attn_prefill_mma_kernel<C, P><<<g2, 128, 0, st>>>((const bf16*)q, (const bf16*)k, (const bf16*)v, (bf16*)out, H, T, \
                                                    Smax, npad)
    if (causal && npad) PF_LAUNCH(true, true);
    else if (causal) PF_LAUNCH(true, false);
    else if (npad) PF_LAUNCH(false, true);
    else PF_LAUNCH(false, false);
#undef PF_LAUNCH
  }
  else {
    tutor_set_error("attn_prefill: unsupported dtype combination q=%d kv=%d out=%d", q_dtype, kv_dtype, out_dtype);
    return TUTOR_ERR_ARG;
  }
  KERNEL_LAUNCHED();
  return TUTOR_OK;
}
