// Corpus cosine scan on the 5th-generation tensor cores (configs[4]; extends the single comparison at
// REF/lms_server.py:1266 to a corpus of chunk embeddings).  scores[q][n] = <query q, corpus row n> / (|q| |row n|).
//
// The fp32 SIMT scan (relevance.cu) is bound by the FMA pipes (24.6 G FMA per 1M x 768 x 32 search), 2.4x above the
// time the 3.07 GB corpus needs to stream from HBM.  The index sets are gated bit-for-bit against a CPU fp32 oracle, so
// plain TF32 (10-bit mantissas) is not an option; this kernel uses the error-compensated "3xTF32" product instead:
//     x = x_hi + x_lo,  x_hi = x with the low 13 mantissa bits cleared (exactly a TF32 number), x_lo = x - x_hi (exact)
//     <c, q> ~= c_hi.q_hi + c_lo.q_hi + c_hi.q_lo          (the dropped c_lo.q_lo term is ~2^-22 relative)
// The tensor core accumulates by truncation: a chain of n accumulations loses up to n * 2^-23 of the running sum, all in
// one direction (measured: 1.5e-5 relative after the 288 MMAs of a 768-dim row in one accumulator).  So the main term
// c_hi.q_hi and the two small cross terms go to separate TMEM accumulators, and the accumulators are drained every
// 6 k-blocks (192 dims, 24 MMAs on the main one) into fp32 registers that the epilogue warps add with round-to-nearest:
// the result is within ~1e-6 of the exact cosine, like an fp32 FMA chain.
//
// One persistent CTA per SM walks 128-row corpus tiles:
//   warp 0      TMA producer: per 32-dim k-block the fp32 corpus tile [128 x 32] and the pre-split query blocks
//               q_hi / q_lo [32 x 32] (L2-resident) into a 5-stage ring, 128-byte swizzle
//   warps 2-5   splitters: write c_lo beside the corpus tile (elementwise, so the swizzled layout is preserved; the MMA
//               itself ignores the low mantissa bits of the raw tile, i.e. reads it as c_hi), then hand the stage over
//   warp 1      one thread issues 2 x 4 tcgen05.mma kind::tf32 per k-block (K=8 each): c x [q_hi; q_lo] (N=64) and
//               c_lo x q_hi (N=32), into one of two TMEM accumulator sets
//   warps 6-9   epilogue: tcgen05.ld, scale by the inverse norms, coalesced stores (32 consecutive rows per query)
#include <cuda.h>
#include <math_constants.h>

#include <cstdlib>

#include "kernels.h"
#include "tc_ptx.cuh"

using namespace tcptx;

namespace {

constexpr int TS_BM = 128, TS_BK = 32, TS_NQ = 32, TS_STAGES = 5, TS_SEG = 6;  // k-blocks per accumulator segment
// kind::tf32 reads fp32 words from shared memory and ignores their low 13 mantissa bits, i.e. it sees exactly c_hi when
// given the raw corpus tile (verified by the fp64 accuracy test: a rounding conversion would show as ~1e-4 errors);
// set to true to store c_hi explicitly (costs 16 KB of shared-memory writes per k-block)
constexpr bool TS_WRITE_HI = false;
constexpr int TS_A_BYTES = TS_BM * TS_BK * 4;   // 16 KB
constexpr int TS_Q_BYTES = TS_NQ * TS_BK * 4;   // 4 KB
constexpr int TS_STAGE_BYTES = 2 * TS_A_BYTES + 2 * TS_Q_BYTES;  // c_hi, c_lo, q_hi, q_lo
constexpr int TS_THREADS = 320;
constexpr int TS_SMEM = TS_STAGES * TS_STAGE_BYTES + 256 + 1024;

// UMMA instruction descriptor, kind::tf32: D = f32 (bits 4-5 = 1), A = B = TF32 (format 2 in bits 7-9 and 10-12),
// both K-major, N >> 3 in [17,23), M >> 4 in [24,29)
constexpr uint32_t ts_idesc(int n) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(TS_BM >> 4) << 24);
}
__device__ __forceinline__ void tc_mma_tf32(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// queries [Q, D] -> q_hi, q_lo [32, D] (rows >= Q zero) and 1 / max(|q|, eps); one block per query row
__global__ void __launch_bounds__(256) split_queries_kernel(const float* __restrict__ queries, int Q, int D, float eps,
                                                           float* __restrict__ q_hi, float* __restrict__ q_lo,
                                                           float* __restrict__ qinv, int* __restrict__ counters) {
  __shared__ float red[8];
  const int q = blockIdx.x;
  if (threadIdx.x == 0) counters[q] = 0;  // arrival counters of the re-scoring CTAs (topk_rescore_kernel)
  if (threadIdx.x == 0 && q == 0) counters[TS_NQ] = 0;  // "the filter pass could not be proven exact"
  float s = 0.f;
  for (int i = threadIdx.x; i < D; i += blockDim.x) {
    const float x = q < Q ? queries[(size_t)q * D + i] : 0.f;
    const float hi = __uint_as_float(__float_as_uint(x) & 0xffffe000u);
    q_hi[(size_t)q * D + i] = hi;
    q_lo[(size_t)q * D + i] = __uint_as_float(__float_as_uint(x - hi) & 0xffffe000u);
    s += x * x;
  }
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < 8; ++w) t += red[w];
    qinv[q] = 1.0f / fmaxf(sqrtf(t), eps);
  }
}

// FILTER: one plain TF32 product per k-block (no split, no cross terms): scores good to ~2e-3, used only to pick the
// candidate groups that are then re-scored in fp32 (see k_cosine_scan_tc).  run_flag != null: a pass that only runs when
// the flag is set (the exact pass behind a filter pass whose result could not be proven exact).
template <bool FILTER>
__global__ void __launch_bounds__(TS_THREADS) cosine_scan_tf32_kernel(const __grid_constant__ CUtensorMap tmC,
                                                                    const __grid_constant__ CUtensorMap tmQh,
                                                                    const __grid_constant__ CUtensorMap tmQl,
                                                                    const float* __restrict__ inv_norm,
                                                                    const float* __restrict__ qinv, int Q, int N, int D,
                                                                    float* __restrict__ scores, float* __restrict__ gmax,
                                                                    const int* __restrict__ run_flag) {
  if (run_flag && *reinterpret_cast<const volatile int*>(run_flag) == 0) return;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* tiles = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(tiles + TS_STAGES * TS_STAGE_BYTES);
  uint64_t* split_bar = full_bar + TS_STAGES;
  uint64_t* empty_bar = split_bar + TS_STAGES;
  uint64_t* tmem_full_bar = empty_bar + TS_STAGES;  // [2]
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;     // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nk = D / TS_BK;
  const int ntiles = (N + TS_BM - 1) / TS_BM;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmC) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmQh) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmQl) : "memory");
    for (int s = 0; s < TS_STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&split_bar[s], 4); mbar_init(&empty_bar[s], 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(&tmem_full_bar[a], 1); mbar_init(&tmem_empty_bar[a], 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(256)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      uint32_t it = 0;
      for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
        for (int kb = 0; kb < nk; ++kb, ++it) {
          const int s = it % TS_STAGES;
          mbar_wait(&empty_bar[s], ((it / TS_STAGES) & 1) ^ 1, 0);
          mbar_expect_tx(&full_bar[s], TS_A_BYTES + (FILTER ? 1 : 2) * TS_Q_BYTES);
          uint8_t* st = tiles + s * TS_STAGE_BYTES;
          tma_load_2d(st, &tmC, &full_bar[s], kb * TS_BK, t * TS_BM);  // rows past N are zero-filled
          tma_load_2d(st + 2 * TS_A_BYTES, &tmQh, &full_bar[s], kb * TS_BK, 0);
          if (!FILTER) tma_load_2d(st + 2 * TS_A_BYTES + TS_Q_BYTES, &tmQl, &full_bar[s], kb * TS_BK, 0);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc64 = ts_idesc(2 * TS_NQ), idesc32 = ts_idesc(TS_NQ);
      uint32_t it = 0, si = 0;  // k-block and segment counters over the whole kernel
      for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
        for (int kb0 = 0; kb0 < nk; kb0 += TS_SEG, ++si) {
          // two accumulator sets of 96 columns: [c_hi.q_hi | c_hi.q_lo] from one N=64 MMA (the query tile is q_hi
          // stacked on q_lo, so the corpus tile is read from shared memory once for both) and [c_lo.q_hi]
          const uint32_t buf = si & 1;
          mbar_wait(&tmem_empty_bar[buf], ((si >> 1) & 1) ^ 1, 3);
          tc_fence_after();
          const uint32_t d_main = tmem_base + buf * 128, d_cross = d_main + 2 * TS_NQ;
          const int kb1 = kb0 + TS_SEG < nk ? kb0 + TS_SEG : nk;
          for (int kb = kb0; kb < kb1; ++kb, ++it) {
            const int s = it % TS_STAGES;
            mbar_wait(&full_bar[s], (it / TS_STAGES) & 1, 1);   // the query blocks (TMA) have landed
            if (!FILTER) mbar_wait(&split_bar[s], (it / TS_STAGES) & 1, 1);  // c_hi / c_lo have been written
            tc_fence_after();
            const uint32_t c_hi = smem_u32(tiles + s * TS_STAGE_BYTES);
            const uint32_t c_lo = c_hi + TS_A_BYTES, q_hi = c_hi + 2 * TS_A_BYTES;  // q_lo follows q_hi: one 64-row tile
#pragma unroll
            for (int k = 0; k < TS_BK / 8; ++k) {  // 8 TF32 = 32 bytes along K per MMA, inside the 128-byte swizzle atom
              const uint32_t first = (kb != kb0 || k != 0);
              if (FILTER) {  // the raw tile read as TF32 against q_hi only
                tc_mma_tf32(d_main, umma_desc_sw128(c_hi + k * 32), umma_desc_sw128(q_hi + k * 32), idesc32, first);
              } else {
                tc_mma_tf32(d_main, umma_desc_sw128(c_hi + k * 32), umma_desc_sw128(q_hi + k * 32), idesc64, first);
                tc_mma_tf32(d_cross, umma_desc_sw128(c_lo + k * 32), umma_desc_sw128(q_hi + k * 32), idesc32, first);
              }
            }
            tc_commit(&empty_bar[s]);
          }
          tc_commit(&tmem_full_bar[buf]);
        }
      }
    }
  } else if (warp < 6) {
    // ===== splitters: c -> c_hi (in place) and c_lo; 128 threads, consecutive 16-byte vectors =====
    const int tid = threadIdx.x - 64;
    uint32_t it = 0;
    for (int t = blockIdx.x; !FILTER && t < ntiles; t += gridDim.x) {
      for (int kb = 0; kb < nk; ++kb, ++it) {
        const int s = it % TS_STAGES;
        mbar_wait(&full_bar[s], (it / TS_STAGES) & 1, 4);
        const uint32_t hi_base = smem_u32(tiles + s * TS_STAGE_BYTES), lo_base = hi_base + TS_A_BYTES;
#pragma unroll
        for (int j = 0; j < TS_A_BYTES / 16 / 128; ++j) {
          const uint32_t off = (uint32_t)(tid + j * 128) * 16u;
          uint32_t x[4], h[4], l[4];
          asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(x[0]), "=r"(x[1]), "=r"(x[2]), "=r"(x[3]) : "r"(hi_base + off));
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            h[i] = x[i] & 0xffffe000u;
            l[i] = __float_as_uint(__uint_as_float(x[i]) - __uint_as_float(h[i])) & 0xffffe000u;
          }
          if (TS_WRITE_HI)
            asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(hi_base + off), "r"(h[0]), "r"(h[1]), "r"(h[2]), "r"(h[3]) : "memory");
          asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(lo_base + off), "r"(l[0]), "r"(l[1]), "r"(l[2]), "r"(l[3]) : "memory");
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) mbar_arrive(&split_bar[s]);
      }
    }
  } else {
    // ===== epilogue: lane = corpus row of the tile, TMEM column = query =====
    const int quad = warp & 3;
    uint32_t si = 0;
    for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
      const int row = t * TS_BM + quad * 32 + lane;
      const float ci = row < N ? inv_norm[row] : 0.f;
      float sum[32];
#pragma unroll
      for (int q = 0; q < 32; ++q) sum[q] = 0.f;
      for (int kb0 = 0; kb0 < nk; kb0 += TS_SEG, ++si) {
        const uint32_t buf = si & 1;
        mbar_wait(&tmem_full_bar[buf], (si >> 1) & 1, 2);
        tc_fence_after();
        float vm[32], vc[32], vd[32];
        const uint32_t taddr = tmem_base + buf * 128 + ((uint32_t)(quad * 32) << 16);
        if (FILTER) {
          tc_ld32(taddr, vm);
        } else {
          tc_ld32x2(taddr, taddr + TS_NQ, vm, vc);  // c_hi.q_hi, c_hi.q_lo
          tc_ld32(taddr + 2 * TS_NQ, vd);           // c_lo.q_hi
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tmem_empty_bar[buf]);
#pragma unroll
        for (int q = 0; q < 32; ++q) sum[q] += FILTER ? vm[q] : vm[q] + (vc[q] + vd[q]);  // round-to-nearest adds of the segment sums
      }
      // scores out (32 consecutive rows per query: coalesced), and the maximum of this warp's 32 rows per query:
      // the k best rows of a query lie in at most k such groups, so the top-k only has to look at k x 32 scores
      float gv[TS_NQ];
#pragma unroll
      for (int q = 0; q < TS_NQ; ++q) {
        gv[q] = (row < N && q < Q) ? sum[q] * qinv[q] * ci : -CUDART_INF_F;
        if (scores && row < N && q < Q) scores[(size_t)q * N + row] = gv[q];
      }
      // transposing butterfly: 31 shuffles leave the maximum over the warp's 32 rows of query q in lane q
      // (a warp_max per query would take 160)
#pragma unroll
      for (int step = 16; step >= 1; step >>= 1) {
        const bool upper = (lane & step) != 0;
#pragma unroll
        for (int i = 0; i < step; ++i) {
          const float send = upper ? gv[i] : gv[i + step];
          const float keep = upper ? gv[i + step] : gv[i];
          gv[i] = fmaxf(keep, __shfl_xor_sync(0xffffffffu, send, step));
        }
      }
      const float mine = gv[0];
      if (gmax && lane < Q) gmax[(size_t)lane * ((size_t)ntiles * 4) + (size_t)t * 4 + quad] = mine;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(256) : "memory");
  }
}

// warp-wide arg-max of (value desc, key asc); every lane gets the winner
__device__ __forceinline__ void warp_argmax(float& v, int& key) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, v, o);
    const int ok = __shfl_xor_sync(0xffffffffu, key, o);
    if (ov > v || (ov == v && ok < key)) { v = ov; key = ok; }
  }
}
// k best of 32 per-warp lists, each sorted (value desc, key asc) in lv / lk [32][32]: lane l walks list l; run by ONE warp.
// The winners go to out(round, value, key).
template <typename Out>
__device__ __forceinline__ void merge_sorted_lists(const float (*lv)[32], const int (*lk)[32], int k, int lane, Out out) {
  int head = 0;
  for (int round = 0; round < k; ++round) {
    float v = head < k ? lv[lane][head] : -CUDART_INF_F;
    int key = head < k ? lk[lane][head] : 0x7fffffff;
    const int mine = key;
    warp_argmax(v, key);
    if (key != 0x7fffffff && key == mine) ++head;  // keys are unique
    out(round, v, key);
  }
}

// Exact top-k of one query from the 32-row group maxima the scan left behind, in two launches:
// (1) topk_select_groups_kernel: the k groups with the largest maxima (ties to the lower group, i.e. the lower rows);
// (2) topk_rescore_kernel: the k x 32 rows of those groups are scored again in fp32 straight from the corpus - the scan
//     does not write its 4 Q N bytes of scores at all, they were 4 % of its traffic and only these few were ever read -
//     one CTA per (group, query) with every load of its 32 rows in flight at once; the last CTA of a query to arrive
//     takes the k best of the k x 32 scores (ties to the lower row).
// The selections are two-level: every warp extracts the k best of its own slice with shuffles only, then one warp merges
// the sorted lists (k block-wide arg-max rounds with two barriers each took 10x longer).  k <= 32.
constexpr int TK_THREADS = 1024;
__global__ void __launch_bounds__(TK_THREADS) topk_select_groups_kernel(const float* __restrict__ gmax, int N, int G, int k,
                                                                       int* __restrict__ sel_out, float* __restrict__ last_out,
                                                                       const int* __restrict__ run_flag) {
  if (run_flag && *reinterpret_cast<const volatile int*>(run_flag) == 0) return;
  extern __shared__ float sg[];  // [G] group maxima of this query
  __shared__ float lv[32][32];   // per-warp sorted lists
  __shared__ int lk[32][32];
  const int q = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int ngroups = (N + 31) / 32;  // groups past this are padding of the last tile
  for (int g = tid; g < G; g += TK_THREADS) sg[g] = g < ngroups ? gmax[(size_t)q * G + g] : -CUDART_INF_F;
  __syncthreads();
  // thread t owns the groups t, t + 1024, ...; its warp extracts the k best of the warp's 32 slices.  A thread keeps its
  // three best in registers, so that dropping a winner costs a register shift and only every third win of the same thread
  // a walk over its slice (with one cached value the owner re-walked up to 31 elements in every round it won, and its
  // warp waited: 0.85 us per round)
  float cv[3];
  int ck[3];
  auto rescan = [&]() {
#pragma unroll
    for (int i = 0; i < 3; ++i) { cv[i] = -CUDART_INF_F; ck[i] = 0x7fffffff; }
    for (int g = tid; g < G; g += TK_THREADS) {
      const float v = sg[g];
      if (v == -CUDART_INF_F) continue;
      if (v > cv[2]) {  // ascending g: strict > keeps the lower group ahead among equals
        if (v > cv[1]) {
          cv[2] = cv[1]; ck[2] = ck[1];
          if (v > cv[0]) { cv[1] = cv[0]; ck[1] = ck[0]; cv[0] = v; ck[0] = g; }
          else { cv[1] = v; ck[1] = g; }
        } else { cv[2] = v; ck[2] = g; }
      }
    }
  };
  rescan();
  for (int round = 0; round < k; ++round) {
    float v = cv[0]; int key = ck[0];
    warp_argmax(v, key);
    if (lane == 0) { lv[warp][round] = v; lk[warp][round] = (v == -CUDART_INF_F) ? 0x7fffffff : key; }
    if (key != 0x7fffffff && key == ck[0]) {  // the owner drops the winner
      sg[key] = -CUDART_INF_F;
      cv[0] = cv[1]; ck[0] = ck[1]; cv[1] = cv[2]; ck[1] = ck[2]; cv[2] = -CUDART_INF_F; ck[2] = 0x7fffffff;
      if (ck[0] == 0x7fffffff) rescan();  // cache empty: whatever is left of the slice
    }
  }
  __syncthreads();
  if (warp == 0)
    merge_sorted_lists(lv, lk, k, lane, [&](int round, float v, int key) {
      if (lane == 0) {
        sel_out[q * 32 + round] = (key == 0x7fffffff || v == -CUDART_INF_F) ? -1 : key;
        // the smallest selected maximum: every group that was NOT selected has a maximum <= this one
        if (last_out && round == k - 1) last_out[q] = (key == 0x7fffffff) ? -CUDART_INF_F : v;
      }
    });
}

// grid (k, Q), 256 threads: warp w scores rows 4w .. 4w+3 of group sel[q][blockIdx.x]
template <int NV>  // float4 per lane and row: d = 128 NV
__global__ void __launch_bounds__(256) topk_rescore_kernel(const float* __restrict__ queries, const float* __restrict__ qinv,
                                                          const float* __restrict__ corpus,
                                                          const float* __restrict__ inv_norm, const int* __restrict__ sel,
                                                          int N, int ng, int k, float* __restrict__ cand,
                                                          int* __restrict__ counters, float* __restrict__ out_v,
                                                          int32_t* __restrict__ out_i, int idx_base,
                                                          const float* __restrict__ unselected_max, float margin,
                                                          int* __restrict__ inexact, const int* __restrict__ run_flag) {
  if (run_flag && *reinterpret_cast<const volatile int*>(run_flag) == 0) return;
  constexpr int D = 128 * NV;
  __shared__ int s_last;
  const int r = blockIdx.x, q = blockIdx.y, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = sel[q * 32 + r];
  float4 cv[4][NV], qv[NV];
#pragma unroll
  for (int i = 0; i < NV; ++i) qv[i] = *reinterpret_cast<const float4*>(queries + (size_t)q * D + (i * 32 + lane) * 4);
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int row = g >= 0 ? g * 32 + warp * 4 + j : N;
#pragma unroll
    for (int i = 0; i < NV; ++i)
      cv[j][i] = row < N ? *reinterpret_cast<const float4*>(corpus + (size_t)row * D + (i * 32 + lane) * 4)
                         : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  const float qi = qinv[q];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int row = g >= 0 ? g * 32 + warp * 4 + j : N;
    float dot = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i)
      dot = fmaf(cv[j][i].x, qv[i].x, fmaf(cv[j][i].y, qv[i].y, fmaf(cv[j][i].z, qv[i].z, fmaf(cv[j][i].w, qv[i].w, dot))));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
    if (lane == 0) cand[(size_t)q * 1024 + r * 32 + warp * 4 + j] = row < N ? dot * qi * inv_norm[row] : -CUDART_INF_F;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = atomicAdd(&counters[q], 1) == ng - 1;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  // the query's last CTA takes the k best of the k x 32 scores, two-level: warp w owns the groups 4w .. 4w+3 (lane l their
  // rows l), extracts the k best of its 128 candidates with shuffles, and warp 0 merges the eight sorted lists
  __shared__ float lv[32][32];
  __shared__ int lk[32][32];
  float v[4];
  int rk[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int gi = warp * 4 + j;
    const int gj = gi < ng ? sel[q * 32 + gi] : -1;
    const int row = gj >= 0 ? gj * 32 + lane : N;
    rk[j] = row < N ? row : 0x7fffffff;
    v[j] = row < N ? __ldcg(cand + (size_t)q * 1024 + gi * 32 + lane) : -CUDART_INF_F;
  }
  for (int round = 0; round < k; ++round) {
    float bv = -CUDART_INF_F;
    int bk = 0x7fffffff;
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (rk[j] != 0x7fffffff && (v[j] > bv || (v[j] == bv && rk[j] < bk))) { bv = v[j]; bk = rk[j]; }
    warp_argmax(bv, bk);
    if (lane == 0) { lv[warp][round] = bv; lk[warp][round] = bk; }
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (rk[j] == bk) rk[j] = 0x7fffffff;  // rows are unique
  }
  __syncthreads();
  if (warp == 0) {
    if (lane >= 8)  // only eight lists exist
      for (int j = 0; j < k; ++j) { lv[lane][j] = -CUDART_INF_F; lk[lane][j] = 0x7fffffff; }
    __syncwarp();
    merge_sorted_lists(lv, lk, k, lane, [&](int round, float bv, int bk) {
      if (lane == 0) {
        out_v[(size_t)q * k + round] = bk != 0x7fffffff ? bv : -CUDART_INF_F;
        out_i[(size_t)q * k + round] = bk != 0x7fffffff ? idx_base + bk : -1;
        // filter pass: a row outside the candidate groups scored at most unselected_max in the TF32 scan, i.e. at most
        // unselected_max + margin exactly; unless the k-th exact score is above that, the exact pass has to run
        if (inexact && round == k - 1 && unselected_max[q] != -CUDART_INF_F &&
            !(bk != 0x7fffffff && bv > unselected_max[q] + margin))
          atomicExch(inexact, 1);
      }
    });
  }
  if (threadIdx.x == 0) counters[q] = 0;
}

}  // namespace

size_t k_cosine_scan_tc_scratch_bytes(int d, int N) {
  return round_up((size_t)2 * TS_NQ * d * sizeof(float) + TS_NQ * sizeof(float), 256) +
         round_up((size_t)TS_NQ * ceil_div(N, TS_BM) * 4 * sizeof(float), 256) +  // split queries + group maxima
         // selected groups, re-scored rows, counters + the "inexact" flag, smallest selected maximum per query
         (size_t)TS_NQ * (32 * sizeof(int) + 1024 * sizeof(float) + sizeof(int) + sizeof(float)) + 256;
}

// TF32 reads operands truncated to 10 mantissa bits: each product is off by < 2^-9 relative, so a cosine by < 2^-9
// (Cauchy-Schwarz on the normalised vectors), plus the truncating accumulation (< 2e-5 over 768 dims)
constexpr float TS_FILTER_MARGIN = 3e-3f;

int k_cosine_scan_tc(const float* queries, const float* corpus, const float* corpus_inv_norm, int Q, int N, int d,
                     float eps, float* scores, void* scratch, int k, float* out_scores, int32_t* out_idx, int idx_base,
                     cudaStream_t st) {
  ARG_CHECK(Q >= 1 && Q <= TS_NQ && d % TS_BK == 0 && N >= 1, "cosine_scan_tc: unsupported Q=%d d=%d N=%d", Q, d, N);
  float* q_hi = reinterpret_cast<float*>(scratch);
  float* q_lo = q_hi + (size_t)TS_NQ * d;
  float* qinv = q_lo + (size_t)TS_NQ * d;
  float* gmax = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(scratch) +
                                         round_up((size_t)2 * TS_NQ * d * sizeof(float) + TS_NQ * sizeof(float), 256));
  int* sel = reinterpret_cast<int*>(reinterpret_cast<uint8_t*>(gmax) +
                                    round_up((size_t)TS_NQ * ceil_div(N, TS_BM) * 4 * sizeof(float), 256));
  float* cand = reinterpret_cast<float*>(sel + TS_NQ * 32);
  int* counters = reinterpret_cast<int*>(cand + (size_t)TS_NQ * 1024);  // [TS_NQ] arrivals, [TS_NQ] = inexact flag
  int* inexact = counters + TS_NQ;
  float* sel_min = reinterpret_cast<float*>(counters + TS_NQ + 1);
  split_queries_kernel<<<TS_NQ, 256, 0, st>>>(queries, Q, d, eps, q_hi, q_lo, qinv, counters);
  KERNEL_LAUNCHED();
  CUtensorMap tc, th, tl;
  TRY(k_make_map_2d(&tc, corpus, 4, N, d, TS_BM, TS_BK));
  TRY(k_make_map_2d(&th, q_hi, 4, TS_NQ, d, TS_NQ, TS_BK));
  TRY(k_make_map_2d(&tl, q_lo, 4, TS_NQ, d, TS_NQ, TS_BK));
  TRY(tutor_set_max_smem(reinterpret_cast<const void*>(cosine_scan_tf32_kernel<false>), TS_SMEM));
  TRY(tutor_set_max_smem(reinterpret_cast<const void*>(cosine_scan_tf32_kernel<true>), TS_SMEM));
  int dev = 0, sms = 148;
  CUDA_TRY(cudaGetDevice(&dev));
  CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int ntiles = ceil_div(N, TS_BM), nctas = ntiles < sms ? ntiles : sms;
  // top-k straight from 32-row group maxima (k <= 32, and the maxima of one query must fit in shared memory); the score
  // matrix itself is never written (the top-k re-scores the few rows it looks at)
  const int G = ntiles * 4;
  ARG_CHECK(k >= 1 && k <= 32 && (d == 128 || d == 768) && (size_t)G * 4 <= 200 * 1024,
            "cosine_scan_tc: k=%d / %d row groups / d=%d unsupported", k, G, d);
  TRY(tutor_set_max_smem(reinterpret_cast<const void*>(topk_select_groups_kernel), 200 * 1024));
  (void)scores;
  auto topk = [&](int ng, bool filter_pass, const int* run_flag) {
    topk_select_groups_kernel<<<Q, TK_THREADS, (size_t)G * 4, st>>>(gmax, N, G, ng, sel, filter_pass ? sel_min : nullptr, run_flag);
    KERNEL_LAUNCHED();
    if (d == 768)
      topk_rescore_kernel<6><<<dim3(ng, Q), 256, 0, st>>>(queries, qinv, corpus, corpus_inv_norm, sel, N, ng, k, cand, counters,
                                                          out_scores, out_idx, idx_base, sel_min, TS_FILTER_MARGIN,
                                                          filter_pass ? inexact : nullptr, run_flag);
    else
      topk_rescore_kernel<1><<<dim3(ng, Q), 256, 0, st>>>(queries, qinv, corpus, corpus_inv_norm, sel, N, ng, k, cand, counters,
                                                          out_scores, out_idx, idx_base, sel_min, TS_FILTER_MARGIN,
                                                          filter_pass ? inexact : nullptr, run_flag);
    KERNEL_LAUNCHED();
  };
  // Filter pass: plain TF32 scan (a third of the MMA work, no operand split: the scan is then bound by HBM, not by
  // shared memory), the 32 best groups re-scored in fp32.  The result is exact when the k-th exact score clears every
  // unselected group's TF32 maximum by the TF32 error bound; when it does not (near-duplicate rows around rank k), a flag
  // lets the error-compensated 3xTF32 pass below run and overwrite the result - otherwise its kernels return at once.
  static const bool filter_off = getenv("TUTOR_SCAN_NO_FILTER") != nullptr;  // cross-check switch for the tests
  // (selecting and re-scoring 32 groups instead of k costs ~35 us more; below ~200 k rows the shorter scan does not pay for it)
  const bool filter = !filter_off && 3 * k <= 32 && N >= 200000;
  if (filter) {
    cosine_scan_tf32_kernel<true><<<nctas, TS_THREADS, TS_SMEM, st>>>(tc, th, tl, corpus_inv_norm, qinv, Q, N, d, nullptr, gmax, nullptr);
    KERNEL_LAUNCHED();
    topk(32, true, nullptr);
  }
  const int* flag = filter ? inexact : nullptr;
  cosine_scan_tf32_kernel<false><<<nctas, TS_THREADS, TS_SMEM, st>>>(tc, th, tl, corpus_inv_norm, qinv, Q, N, d, nullptr, gmax, flag);
  KERNEL_LAUNCHED();
  topk(k, false, flag);
  return TUTOR_OK;
}
