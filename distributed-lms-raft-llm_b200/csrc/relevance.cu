// Corpus relevance scan (configs[4], an extension of REF/lms_server.py:1266 from one assignment text to a
// corpus of chunk embeddings): cosine of Q query embeddings against N corpus rows, then top-k.
//   1. cosine_scan : streams the corpus once (HBM-bound on its N*d*4 bytes); queries live in shared
//                    memory, each warp holds 4 corpus rows in registers so one query vector read from
//                    smem feeds 16 FMAs; scores[Q][N] go to scratch
//   2. topk_chunk  : per (query, 4096-score chunk) exact top-k by k rounds of block arg-max
//   3. topk_merge  : the same selection over the chunk candidates (and, after the NCCL all-gather of the
//                    per-rank candidates, over the world_size*k gathered ones)
// Ties always resolve to the lower corpus index, like torch.topk on the CPU oracle for distinct scores.
#include <math_constants.h>

#include <cstdlib>

#include "kernels.h"

namespace {

constexpr int CHUNK = 4096;

template <int DV4>  // d = 128 * DV4
__global__ void __launch_bounds__(256, 1) cosine_scan_kernel(const float* __restrict__ queries,
                                                              const float* __restrict__ corpus,
                                                              const float* __restrict__ inv_norm, int Q, int N,
                                                              float eps, float* __restrict__ scores) {
  constexpr int D = 128 * DV4;
  extern __shared__ float sq[];           // [Q][D] queries, then [Q] inverse norms
  float* qinv = sq + (size_t)Q * D;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < Q * D / 4; i += blockDim.x)
    reinterpret_cast<float4*>(sq)[i] = reinterpret_cast<const float4*>(queries)[i];
  __syncthreads();
  for (int q = warp; q < Q; q += 8) {
    float s = 0.f;
    for (int i = lane; i < D; i += 32) s += sq[q * D + i] * sq[q * D + i];
    s = warp_sum(s);
    if (lane == 0) qinv[q] = 1.0f / fmaxf(sqrtf(s), eps);
  }
  __syncthreads();
  const int ngroups = (N + 3) / 4;
  for (int g = blockIdx.x * 8 + warp; g < ngroups; g += gridDim.x * 8) {
    float4 c[4][DV4];
    float cinv[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int row = min(4 * g + r, N - 1);
      cinv[r] = inv_norm[row];
#pragma unroll
      for (int j = 0; j < DV4; ++j)
        c[r][j] = *reinterpret_cast<const float4*>(corpus + (size_t)row * D + j * 128 + lane * 4);
    }
    for (int q = 0; q < Q; ++q) {
      float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int j = 0; j < DV4; ++j) {
        const float4 x = *reinterpret_cast<const float4*>(sq + (size_t)q * D + j * 128 + lane * 4);
#pragma unroll
        for (int r = 0; r < 4; ++r)
          acc[r] += (c[r][j].x * x.x + c[r][j].y * x.y) + (c[r][j].z * x.z + c[r][j].w * x.w);
      }
#pragma unroll
      for (int r = 0; r < 4; ++r) acc[r] = warp_sum(acc[r]);
      if (lane < 4) {
        const int row = 4 * g + lane;
        const float a = lane == 0 ? acc[0] : lane == 1 ? acc[1] : lane == 2 ? acc[2] : acc[3];
        const float ci = lane == 0 ? cinv[0] : lane == 1 ? cinv[1] : lane == 2 ? cinv[2] : cinv[3];
        if (row < N) scores[(size_t)q * N + row] = a * qinv[q] * ci;
      }
    }
  }
}

// Second-generation scan: no cross-lane reduction at all.  Every lane owns two corpus rows and 2x32 accumulators;
// a warp stages [64 rows][16 dims] tiles of its rows in shared memory with cp.async (3 stages).  fp32 FMA only - the
// index sets are gated bit-for-bit against the CPU oracle, so no TF32.  QMAX = 32 queries per pass.
// Third revision: packed FFMA2 (fma.rn.f32x2, new on sm_100).  The scalar version ran at exactly the scalar-FFMA issue
// rate of the SMs (24.6 G FMA in 1.38 ms = 64 lanes per clock per SM).  The queries now sit in shared memory dim-major
// ([D][32]), so one broadcast LDS.128 delivers dim k of four queries as two register pairs, and each FFMA2 updates the
// accumulators of two queries for one row with {row[k], row[k]}: half the FMA-pipe cycles, and every lane still does
// the same fma.rn sequence in ascending k, so the scores are bit-identical to the scalar kernel's.
constexpr int SC_ROWS = 64, SC_K = 16, SC_LD = 20, SC_STAGES = 3, SC_QMAX = 32;

__device__ __forceinline__ void ffma2(unsigned long long& acc, unsigned long long a, unsigned long long b) {
  asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc) : "l"(a), "l"(b));
}
__device__ __forceinline__ unsigned long long pack2(float lo, float hi) {
  return (unsigned long long)__float_as_uint(lo) | ((unsigned long long)__float_as_uint(hi) << 32);
}

__global__ void __launch_bounds__(256, 1) cosine_scan2_kernel(const float* __restrict__ queries,
                                                               const float* __restrict__ corpus,
                                                               const float* __restrict__ inv_norm, int Q, int N, int D,
                                                               float eps, float* __restrict__ scores) {
  extern __shared__ __align__(16) float sm2[];
  float* sq = sm2;                                  // [D][SC_QMAX] queries, dim-major (columns >= Q zero-filled)
  float* qinv = sq + (size_t)SC_QMAX * D;           // [SC_QMAX]
  float* tiles = qinv + SC_QMAX;                    // [8 warps][SC_STAGES][SC_ROWS][SC_LD]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < SC_QMAX * D; i += blockDim.x) {
    const int k = i / SC_QMAX, q = i % SC_QMAX;
    sq[i] = q < Q ? queries[(size_t)q * D + k] : 0.f;
  }
  __syncthreads();
  for (int q = warp; q < SC_QMAX; q += 8) {
    float s = 0.f;
    for (int i = lane; i < D; i += 32) s += sq[i * SC_QMAX + q] * sq[i * SC_QMAX + q];
    s = warp_sum(s);
    if (lane == 0) qinv[q] = 1.0f / fmaxf(sqrtf(s), eps);
  }
  __syncthreads();
  float* wt = tiles + (size_t)warp * SC_STAGES * SC_ROWS * SC_LD;
  const int nk = D / SC_K;
  const int ngroups = (N + SC_ROWS - 1) / SC_ROWS;
  for (int g = blockIdx.x * 8 + warp; g < ngroups; g += gridDim.x * 8) {
    const int row0 = g * SC_ROWS;
    auto issue = [&](int kt) {  // one [64][16] tile: 8 x 16-byte cp.async per lane, 4 lanes per 64-byte row segment
      float* dst = wt + (size_t)(kt % SC_STAGES) * SC_ROWS * SC_LD;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int r = i * 8 + (lane >> 2), c = (lane & 3) * 4;
        const int row = min(row0 + r, N - 1);
        const float* src = corpus + (size_t)row * D + kt * SC_K + c;
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst + r * SC_LD + c)),
                     "l"(src));
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
    };
    unsigned long long accA[SC_QMAX / 2], accB[SC_QMAX / 2];  // {query 2j, query 2j+1} for rows lane and lane + 32
#pragma unroll
    for (int j = 0; j < SC_QMAX / 2; ++j) { accA[j] = 0ull; accB[j] = 0ull; }
    issue(0);
    if (nk > 1) issue(1);
    for (int kt = 0; kt < nk; ++kt) {
      if (kt + 2 < nk) { issue(kt + 2); asm volatile("cp.async.wait_group 2;" ::: "memory"); }
      else if (kt + 1 < nk) asm volatile("cp.async.wait_group 1;" ::: "memory");
      else asm volatile("cp.async.wait_group 0;" ::: "memory");
      __syncwarp();
      const float* t = wt + (size_t)(kt % SC_STAGES) * SC_ROWS * SC_LD;
#pragma unroll
      for (int c = 0; c < SC_K / 4; ++c) {
        const float4 ra = *reinterpret_cast<const float4*>(t + lane * SC_LD + c * 4);
        const float4 rb = *reinterpret_cast<const float4*>(t + (lane + 32) * SC_LD + c * 4);
        const float rav[4] = {ra.x, ra.y, ra.z, ra.w}, rbv[4] = {rb.x, rb.y, rb.z, rb.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const unsigned long long a2 = pack2(rav[e], rav[e]), b2 = pack2(rbv[e], rbv[e]);
          const float* qp = sq + (size_t)(kt * SC_K + c * 4 + e) * SC_QMAX;
#pragma unroll
          for (int q4 = 0; q4 < SC_QMAX / 4; ++q4) {
            // dim k of queries 4 q4 .. 4 q4 + 3: the same address in every lane (broadcast)
            const ulonglong2 x = *reinterpret_cast<const ulonglong2*>(qp + q4 * 4);
            ffma2(accA[2 * q4], a2, x.x); ffma2(accA[2 * q4 + 1], a2, x.y);
            ffma2(accB[2 * q4], b2, x.x); ffma2(accB[2 * q4 + 1], b2, x.y);
          }
        }
      }
      __syncwarp();  // the stage is refilled by the cp.async issued two iterations later
    }
    const int ra_row = row0 + lane, rb_row = row0 + lane + 32;
    const float ca = ra_row < N ? inv_norm[ra_row] : 0.f, cb = rb_row < N ? inv_norm[rb_row] : 0.f;
#pragma unroll
    for (int q = 0; q < SC_QMAX; ++q) {
      if (q < Q) {
        const float va = __uint_as_float((unsigned)(accA[q >> 1] >> ((q & 1) * 32)));
        const float vb = __uint_as_float((unsigned)(accB[q >> 1] >> ((q & 1) * 32)));
        if (ra_row < N) scores[(size_t)q * N + ra_row] = va * qinv[q] * ca;
        if (rb_row < N) scores[(size_t)q * N + rb_row] = vb * qinv[q] * cb;
      }
    }
  }
}

template <typename T>
__global__ void row_inv_norm_kernel(const T* __restrict__ x, float* __restrict__ out, int N, int d, float eps) {
  int r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= N) return;
  int lane = threadIdx.x & 31;
  float s = 0.f;
  for (int i = lane * 4; i < d; i += 128) {
    float4 v = load4(x + (size_t)r * d + i);
    s += (v.x * v.x + v.y * v.y) + (v.z * v.z + v.w * v.w);
  }
  s = warp_sum(s);
  if (lane == 0) out[r] = 1.0f / fmaxf(sqrtf(s), eps);
}

// block-wide (value desc, index asc) arg-max of up to CHUNK candidates held in smem, repeated k times
__device__ void select_topk(float* sv, const int32_t* sidx, int n, int k, float* out_v, int32_t* out_i,
                            int idx_base) {
  __shared__ float rv[32];
  __shared__ int ri[32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  for (int round = 0; round < k; ++round) {
    float bv = -CUDART_INF_F;
    int bp = 0x7fffffff;  // position in smem; positions are in ascending index order
    for (int p = threadIdx.x; p < n; p += blockDim.x) {
      float v = sv[p];
      if (v > bv) { bv = v; bp = p; }  // p ascending per thread: strict > keeps the lowest position
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      float ov = __shfl_xor_sync(0xffffffffu, bv, o);
      int op = __shfl_xor_sync(0xffffffffu, bp, o);
      if (ov > bv || (ov == bv && op < bp)) { bv = ov; bp = op; }
    }
    if (lane == 0) { rv[warp] = bv; ri[warp] = bp; }
    __syncthreads();
    if (warp == 0) {
      bv = lane < nwarps ? rv[lane] : -CUDART_INF_F;
      bp = lane < nwarps ? ri[lane] : 0x7fffffff;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        float ov = __shfl_xor_sync(0xffffffffu, bv, o);
        int op = __shfl_xor_sync(0xffffffffu, bp, o);
        if (ov > bv || (ov == bv && op < bp)) { bv = ov; bp = op; }
      }
      if (lane == 0) {
        if (bp != 0x7fffffff) {
          out_v[round] = bv;
          out_i[round] = sidx ? sidx[bp] : idx_base + bp;
          sv[bp] = -CUDART_INF_F;
        } else {  // fewer than k candidates
          out_v[round] = -CUDART_INF_F;
          out_i[round] = -1;
        }
      }
    }
    __syncthreads();
  }
}

// grid (nchunks, Q)
__global__ void __launch_bounds__(256) topk_chunk_kernel(const float* __restrict__ scores, int N, int k,
                                                          float* __restrict__ cand_v, int32_t* __restrict__ cand_i,
                                                          int idx_base) {
  __shared__ float sv[CHUNK];
  const int chunk = blockIdx.x, q = blockIdx.y, nchunks = gridDim.x;
  const int p0 = chunk * CHUNK;
  const int n = min(CHUNK, N - p0);
  for (int p = threadIdx.x; p < n; p += blockDim.x) sv[p] = scores[(size_t)q * N + p0 + p];
  __syncthreads();
  select_topk(sv, nullptr, n, k, cand_v + ((size_t)q * nchunks + chunk) * k, cand_i + ((size_t)q * nchunks + chunk) * k,
              idx_base + p0);
}

// k <= 32: the same exact selection without block-wide rounds.  Each warp first selects the top k of its own 512
// scores - a lane keeps the best of its 16 (strided) elements cached, a round is one warp arg-max (5 shuffles) and only
// the winning lane rescans - then warp 0 selects the top k of the 8 x k survivors.  Ties go to the lower index.
// (The block-wide version spent k rounds x (16 smem reads per thread + 2 barriers): 155 us of a 0.84 ms search.)
__device__ __forceinline__ void warp_argmax(float& v, int& p) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, v, o);
    const int op = __shfl_xor_sync(0xffffffffu, p, o);
    if (ov > v || (ov == v && op < p)) { v = ov; p = op; }
  }
}

__global__ void __launch_bounds__(256) topk_chunk_fast_kernel(const float* __restrict__ scores, int N, int k,
                                                               float* __restrict__ cand_v, int32_t* __restrict__ cand_i,
                                                               int idx_base) {
  __shared__ float sv[CHUNK];
  __shared__ float wv[8][32];
  __shared__ int wp[8][32];
  const int chunk = blockIdx.x, q = blockIdx.y, nchunks = gridDim.x;
  const int p0 = chunk * CHUNK;
  const int n = min(CHUNK, N - p0);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int wbase = warp * 512;  // this warp's 512 positions; lane owns wbase + lane + 32 j
  auto rescan = [&](float& bv, int& bp) {
    bv = -CUDART_INF_F; bp = 0x7fffffff;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const int p = wbase + lane + 32 * j;
      const float v = sv[p];
      if (v > bv) { bv = v; bp = p; }  // ascending p: strict > keeps the lowest position
    }
  };
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    const int p = wbase + lane + 32 * j;
    sv[p] = p < n ? scores[(size_t)q * N + p0 + p] : -CUDART_INF_F;
  }
  __syncwarp();
  float bv; int bp;
  rescan(bv, bp);
  for (int round = 0; round < k; ++round) {
    float v = bv; int p = bp;
    warp_argmax(v, p);
    if (lane == 0) { wv[warp][round] = v; wp[warp][round] = p; }
    if (p != 0x7fffffff && ((p - wbase) & 31) == lane) {  // the owner drops the winner and refreshes its cache
      sv[p] = -CUDART_INF_F;
      rescan(bv, bp);
    }
    __syncwarp();
  }
  __syncthreads();
  if (warp == 0) {
    // 8 x k survivors, lane l holds those of warp l % 8 at rounds l / 8, l / 8 + 4, ...
    const int sw = lane & 7;
    float cv[8]; int cp[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int r = (lane >> 3) + 4 * j;
      const bool ok = r < k;
      cv[j] = ok ? wv[sw][r] : -CUDART_INF_F;
      cp[j] = ok ? wp[sw][r] : 0x7fffffff;
      if (cv[j] == -CUDART_INF_F) cp[j] = 0x7fffffff;  // exhausted warp: not a candidate
    }
    float* out_v = cand_v + ((size_t)q * nchunks + chunk) * k;
    int32_t* out_i = cand_i + ((size_t)q * nchunks + chunk) * k;
    for (int round = 0; round < k; ++round) {
      float v = -CUDART_INF_F; int p = 0x7fffffff;
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (cv[j] > v || (cv[j] == v && cp[j] < p)) { v = cv[j]; p = cp[j]; }
      warp_argmax(v, p);
      if (lane == 0) {
        out_v[round] = p != 0x7fffffff ? v : -CUDART_INF_F;
        out_i[round] = p != 0x7fffffff ? idx_base + p0 + p : -1;
      }
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (cp[j] == p) { cv[j] = -CUDART_INF_F; cp[j] = 0x7fffffff; }  // positions are unique
    }
  }
}

// grid (Q): candidates [Q][ncand] (value, index); candidates must be ordered by ascending index among equal values
__global__ void __launch_bounds__(256) topk_merge_kernel(const float* __restrict__ cand_v,
                                                          const int32_t* __restrict__ cand_i, int ncand, int k,
                                                          float* __restrict__ out_v, int32_t* __restrict__ out_i) {
  extern __shared__ float dyn[];
  float* sv = dyn;
  int32_t* si = reinterpret_cast<int32_t*>(dyn + ncand);
  const int q = blockIdx.x;
  for (int p = threadIdx.x; p < ncand; p += blockDim.x) {
    sv[p] = cand_v[(size_t)q * ncand + p];
    si[p] = cand_i[(size_t)q * ncand + p];
    if (si[p] < 0) sv[p] = -CUDART_INF_F;
  }
  __syncthreads();
  select_topk(sv, si, ncand, k, out_v + (size_t)q * k, out_i + (size_t)q * k, 0);
}

}  // namespace

size_t k_cosine_topk_scratch_bytes(int Q, int N, int k) {
  size_t nchunks = (size_t)ceil_div(N, CHUNK);
  return k_cosine_scan_tc_scratch_bytes(768, N) + round_up((size_t)Q * N * sizeof(float), 256) +
         round_up((size_t)Q * nchunks * k * sizeof(float), 256) + round_up((size_t)Q * nchunks * k * sizeof(int32_t), 256);
}

int k_row_inv_norm(const void* corpus, int corpus_dtype, float* inv_norm, int N, int d, float eps, cudaStream_t st) {
  ARG_CHECK(d % 4 == 0, "row_inv_norm: d %% 4 != 0");
  if (N == 0) return TUTOR_OK;
  if (corpus_dtype == DT_F32)
    row_inv_norm_kernel<float><<<ceil_div(N, 8), 256, 0, st>>>((const float*)corpus, inv_norm, N, d, eps);
  else
    row_inv_norm_kernel<bf16><<<ceil_div(N, 8), 256, 0, st>>>((const bf16*)corpus, inv_norm, N, d, eps);
  KERNEL_LAUNCHED();
  return TUTOR_OK;
}

int k_topk_merge(const float* cand_scores, const int32_t* cand_idx, int Q, int ncand, int k, float* out_scores,
                 int32_t* out_idx, cudaStream_t st) {
  ARG_CHECK(k >= 1 && ncand >= 1, "topk_merge: bad k=%d ncand=%d", k, ncand);
  size_t smem = (size_t)ncand * 8;
  ARG_CHECK(smem <= 200 * 1024, "topk_merge: %d candidates do not fit in shared memory", ncand);
  TRY(tutor_set_max_smem(reinterpret_cast<const void*>(topk_merge_kernel), 200 * 1024));
  topk_merge_kernel<<<Q, 256, smem, st>>>(cand_scores, cand_idx, ncand, k, out_scores, out_idx);
  KERNEL_LAUNCHED();
  return TUTOR_OK;
}

int k_cosine_topk(const float* queries, const void* corpus, int corpus_dtype, const float* corpus_inv_norm, int Q,
                  int N, int d, int k, float* out_scores, int32_t* out_idx, int idx_base, void* scratch,
                  cudaStream_t st) {
  ARG_CHECK(corpus_dtype == DT_F32, "cosine_topk: corpus must be fp32 (scores are gated at 1e-4)");
  ARG_CHECK(d == 768 || d == 128, "cosine_topk: d=%d unsupported (768 or 128)", d);
  ARG_CHECK(Q >= 1 && (size_t)Q * d * 4 + Q * 4 <= 200 * 1024, "cosine_topk: %d queries do not fit in shared memory", Q);
  ARG_CHECK(k >= 1 && k <= CHUNK && N >= 1, "cosine_topk: bad k=%d N=%d", k, N);
  const int nchunks = ceil_div(N, CHUNK);
  uint8_t* p = reinterpret_cast<uint8_t*>(scratch);
  float* scores = reinterpret_cast<float*>(p);
  p += round_up((size_t)Q * N * sizeof(float), 256);
  float* cand_v = reinterpret_cast<float*>(p);
  p += round_up((size_t)Q * nchunks * k * sizeof(float), 256);
  int32_t* cand_i = reinterpret_cast<int32_t*>(p);
  p += round_up((size_t)Q * nchunks * k * sizeof(int32_t), 256);
  void* tc_scratch = p;
  const float eps = 1e-8f;
  static const bool use_tc = tutor_experiment("TUTOR_SCAN_SIMT") == nullptr;  // A/B: the fp32 FFMA2 scan below
  if (use_tc && Q <= 32 && (d == 128 || d == 768) && k <= 32 && (size_t)ceil_div(N, 128) * 16 <= 200 * 1024)
    return k_cosine_scan_tc(queries, (const float*)corpus, corpus_inv_norm, Q, N, d, eps, scores, tc_scratch, k, out_scores,
                            out_idx, idx_base, st);
  const int grid = 148;
  const size_t smem2 = ((size_t)SC_QMAX * d + SC_QMAX + (size_t)8 * SC_STAGES * SC_ROWS * SC_LD) * sizeof(float);
  if (Q <= SC_QMAX && d % SC_K == 0 && smem2 <= 220 * 1024) {
    TRY(tutor_set_max_smem(reinterpret_cast<const void*>(cosine_scan2_kernel), 220 * 1024));
    cosine_scan2_kernel<<<grid, 256, smem2, st>>>(queries, (const float*)corpus, corpus_inv_norm, Q, N, d, eps, scores);
  } else {
    const size_t smem = (size_t)Q * d * 4 + (size_t)Q * 4;
    if (d == 768) {
      TRY(tutor_set_max_smem(reinterpret_cast<const void*>(cosine_scan_kernel<6>), 200 * 1024));
      cosine_scan_kernel<6><<<grid, 256, smem, st>>>(queries, (const float*)corpus, corpus_inv_norm, Q, N, eps, scores);
    } else {
      TRY(tutor_set_max_smem(reinterpret_cast<const void*>(cosine_scan_kernel<1>), 200 * 1024));
      cosine_scan_kernel<1><<<grid, 256, smem, st>>>(queries, (const float*)corpus, corpus_inv_norm, Q, N, eps, scores);
    }
  }
  KERNEL_LAUNCHED();
  if (k <= 32) topk_chunk_fast_kernel<<<dim3(nchunks, Q), 256, 0, st>>>(scores, N, k, cand_v, cand_i, idx_base);
    else topk_chunk_kernel<<<dim3(nchunks, Q), 256, 0, st>>>(scores, N, k, cand_v, cand_i, idx_base);
  KERNEL_LAUNCHED();
  // chunk candidates are in ascending chunk order, i.e. ascending index among equal values
  return k_topk_merge(cand_v, cand_i, Q, nchunks * k, k, out_scores, out_idx, st);
}
