// Decode-time (one token per sequence) linear layers of the GPT-2 block with the work around them folded in, so a
// transformer layer is 5 launches instead of 8 (HF/models/gpt2/modeling_gpt2.py:262-309, Conv1D at
// HF/pytorch_utils.py:119-123):
//
//   * "swap-AB" split-K tcgen05 GEMM as in gemm_bf16_tc.cu: the weight matrix [N,K] is the 128-row UMMA A operand
//     (TMA, SWIZZLE_128B), the batch (<= 128 rows) is the UMMA N dimension, ks CTAs share one weight tile and each
//     streams K/ks of it so that a 1-5 MB weight covers ~all SMs;
//   * LNB: the B operand is LayerNorm(x) computed in the kernel - the four epilogue warps read the fp32 residual
//     stream and per-row statistics, normalise, and write bf16 straight into the swizzled shared-memory tiles the
//     MMA reads (no TMA for B, no separate LayerNorm kernel, no bf16 copy of the activations in HBM);
//   * MODE_ACT / MODE_RESID: the ks CTAs of one output tile form a thread-block cluster; each leaves its fp32
//     partial tile in its own shared memory and, after a cluster barrier, reduces a 1/ks slice of the batch rows by
//     reading all peers over DSMEM in a fixed order (deterministic), then applies bias + gelu_new -> bf16, or
//     bias + residual add -> x (fp32, in place) plus the per-row (mean, M2) of its 128 columns, from which the next
//     LNB consumer rebuilds the LayerNorm statistics (Chan's parallel-variance combination).
//   * MODE_PARTIAL keeps the raw-partials output (c_attn: the attention kernel finishes that reduction itself).
//   * Folded LayerNorm (opt-in, TUTOR_FOLD=1): W' = bf16(gamma (.) W), s[n] = sum_k W'[n,k],
//     c[n] = sum_k beta[k] W[n,k] are made at load time, so that  LayerNorm(x) . W[n] = rstd (x . W'[n] - mean s[n]) + c[n].
//     The B operand is then the raw residual stream as bf16 (written by the producer of x next to the fp32 copy) and
//     arrives by TMA like any activation; rstd and mean enter once per output element in whoever closes the split-K sum
//     (the cluster reduction here, the attention kernel for c_attn).  Meant to take the statistics + normalise +
//     smem-write pass off the critical path of c_attn and c_fc; measured 614 vs 605 us per decode forward at B = 64
//     (inside the chain x is L2-resident and the hand-written operand is ready as early as the TMA tile), so the LNB
//     flavour stays the default.
//
// Every mbarrier wait is bounded (tc_ptx.cuh).  PDL: weight tiles are requested before griddepcontrol.wait, anything
// the predecessor kernel produces (x, statistics, activations) only after it.
#include <cuda.h>

#include "kernels.h"
#include "tc_ptx.cuh"

using namespace tcptx;

namespace {

constexpr int NTHREADS = 192;  // warp 0 TMA producer, warp 1 TMEM allocator + MMA issuer, warps 2..5 B-producers + epilogue
enum { MODE_PARTIAL = 0, MODE_ACT = 1, MODE_RESID = 2 };
constexpr int MAX_STAT_TILES = 16;  // LayerNorm rows of up to 16 * 128 = 2048 features

// Phase timestamps of CTA (0,0,0) for tools/microbench_fused.cu (compiled in only with -DTUTOR_TRACE)
#ifdef TUTOR_TRACE
__device__ long long g_fused_trace[16];
#define FTRACE(i)                                                                                \
  do {                                                                                           \
    if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) g_fused_trace[i] = clock64();     \
  } while (0)
#else
#define FTRACE(i)
#endif

struct LnOperand {
  const float* x;      // [M, K] fp32 residual stream
  const float* stats;  // [nst][M][2]: (mean, M2) of each row over columns [i*K/nst, (i+1)*K/nst)
  int nst;
  const float* gamma;  // [K]
  const float* beta;
  float eps;
};

struct FusedEpi {
  const float* bias;  // [N]
  int act;            // MODE_ACT
  bf16* out_bf16;     // MODE_ACT: [M, N]
  float* x;           // MODE_RESID: [M, N] residual stream, updated in place
  float* stats;       // MODE_RESID: [N/128][M][2]
  bf16* xb;           // MODE_RESID: optional bf16 copy of the new x (the operand of a folded-LayerNorm consumer)
  // MODE_ACT with a folded LayerNorm in front (fold_s != null): out = act(rstd (acc - mean s[n]) + c[n] + bias[n])
  const float* fold_s;      // [N]
  const float* fold_c;      // [N]
  const float* fold_stats;  // [fold_nst][M][2] (mean, M2) of the rows of x, fold_k features in all
  int fold_nst, fold_k;
  float fold_eps;
};

// mean and 1/sqrt(var + eps) of row m from its per-tile (mean, M2) pairs (Chan et al. combination, equal counts)
__device__ __forceinline__ void row_mean_rstd(const float* stats, int nst, int M, int m, int K, float eps, float& mean,
                                              float& rstd) {
  const float2* st2 = reinterpret_cast<const float2*>(stats) + m;
  float2 tv[MAX_STAT_TILES];  // every pair is requested before the first is used
#pragma unroll
  for (int i = 0; i < MAX_STAT_TILES; ++i) tv[i] = i < nst ? st2[(size_t)i * M] : make_float2(0.f, 0.f);
  float sm = 0.f;
#pragma unroll
  for (int i = 0; i < MAX_STAT_TILES; ++i) sm += tv[i].x;  // padding entries are zero
  mean = sm / (float)nst;
  const float n_per = (float)(K / nst);
  float m2 = 0.f;
#pragma unroll
  for (int i = 0; i < MAX_STAT_TILES; ++i) {
    const float dm = tv[i].x - mean;
    if (i < nst) m2 += tv[i].y + n_per * dm * dm;
  }
  rstd = 1.0f / sqrtf(m2 / (float)K + eps);
}

__device__ __forceinline__ void st_shared_v2(uint32_t addr, uint32_t a, uint32_t b) {
  asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(addr), "r"(a), "r"(b) : "memory");
}

template <int BN, bool LNB, int MODE>
__global__ void __launch_bounds__(NTHREADS) gemm_decode_fused_kernel(const __grid_constant__ CUtensorMap tmA,
                                                                   const __grid_constant__ CUtensorMap tmB,
                                                                   const __grid_constant__ CUtensorMap tmC, int rowsA,
                                                                   int rowsB, int K, LnOperand ln, FusedEpi ep) {
  using L = SmemLayout<BN>;
  constexpr int TMEM_COLS = (BN < 32) ? 32 : BN;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* tiles = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(tiles + L::TILE_BYTES);
  uint64_t* empty_bar = full_bar + L::STAGES;
  uint64_t* tmem_full_bar = empty_bar + L::STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);
  float* ln_smem = reinterpret_cast<float*>(tiles + L::TILE_BYTES + L::BAR_BYTES);  // LNB: gamma, beta of the K slice

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int a_row0 = blockIdx.x * BM, b_row0 = blockIdx.y * BN;
  const int nk = (K / BK) / (int)gridDim.z;  // k-blocks of this CTA's K slice
  const int kb0 = (int)blockIdx.z * nk;
  if (threadIdx.x == 64) FTRACE(0);

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
    if (!LNB) asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
    // full: the TMA thread's expect_tx arrival, plus one arrival per B-producer warp when B is written by hand
    for (int s = 0; s < L::STAGES; ++s) { mbar_init(&full_bar[s], LNB ? 5 : 1); mbar_init(&empty_bar[s], 1); }
    mbar_init(tmem_full_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"(TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (threadIdx.x == 64) FTRACE(1);

  // cluster epilogue state of the reducing warps: this CTA's slice of the batch rows, and the global operands
  // (bias, residual rows) requested before the cluster barrier so that their latency hides behind it
  const uint32_t ksz = (MODE == MODE_PARTIAL) ? 1u : cluster_nctarank();
  const int rows_per = (BN + (int)ksz - 1) / (int)ksz;
  const int r0 = (MODE == MODE_PARTIAL) ? 0 : (int)cluster_ctarank() * rows_per;
  const int r1 = r0 + rows_per < BN ? r0 + rows_per : BN;
  const int n = a_row0 + lane * 4;  // this lane's 4 output features in the reduction
  const bool n_ok = n < rowsA;      // rowsA % 4 == 0 (checked by the launcher)
  float4 bias4 = make_float4(0.f, 0.f, 0.f, 0.f), res_pre[4];
  float4 s4 = make_float4(0.f, 0.f, 0.f, 0.f);  // folded LayerNorm: s[n] of this lane's features
  float mean_pre[4], rstd_pre[4];               // ... and the statistics of the rows this warp reduces first
#pragma unroll
  for (int i = 0; i < 4; ++i) { res_pre[i] = make_float4(0.f, 0.f, 0.f, 0.f); mean_pre[i] = 0.f; rstd_pre[i] = 1.f; }
  const bool fold = MODE == MODE_ACT && !LNB && ep.fold_s != nullptr;

  pdl_launch_dependents();
  if (warp == 0) {
    // ===== TMA producer: weight tiles (static data) are requested ahead of the programmatic-launch wait =====
    if (lane == 0) {
      constexpr uint32_t TX = LNB ? L::A_BYTES : L::STAGE_BYTES;
      const int pre = nk < L::STAGES ? nk : L::STAGES;
      for (int kb = 0; kb < pre; ++kb) {
        mbar_expect_tx(&full_bar[kb], TX);
        tma_load_2d(tiles + kb * L::STAGE_BYTES, &tmA, &full_bar[kb], (kb0 + kb) * BK, a_row0);
      }
      if (!LNB) {
        pdl_wait();
        for (int kb = 0; kb < pre; ++kb)
          tma_load_2d(tiles + kb * L::STAGE_BYTES + L::A_BYTES, &tmB, &full_bar[kb], (kb0 + kb) * BK, b_row0);
      }
      for (int kb = pre; kb < nk; ++kb) {
        const int s = kb % L::STAGES;
        mbar_wait(&empty_bar[s], ((kb / L::STAGES) & 1) ^ 1, 0);
        mbar_expect_tx(&full_bar[s], TX);
        uint8_t* sa = tiles + s * L::STAGE_BYTES;
        tma_load_2d(sa, &tmA, &full_bar[s], (kb0 + kb) * BK, a_row0);
        if (!LNB) tma_load_2d(sa + L::A_BYTES, &tmB, &full_bar[s], (kb0 + kb) * BK, b_row0);
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread) =====
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(BM, BN);
      for (int kb = 0; kb < nk; ++kb) {
        const int s = kb % L::STAGES;
        mbar_wait(&full_bar[s], (kb / L::STAGES) & 1, 1);
        tc_fence_after();
        const uint32_t sa = smem_u32(tiles + s * L::STAGE_BYTES);
        // one descriptor pair per stage; a k-step of 16 bf16 = 32 bytes moves the 16-byte-granular address field by 2
        const uint64_t da = umma_desc_sw128(sa), db = umma_desc_sw128(sa + L::A_BYTES);
#pragma unroll
        for (int k = 0; k < BK / UMMA_K; ++k)
          tc_mma_bf16(tmem_base, da + (uint64_t)(k * (UMMA_K * 2 >> 4)), db + (uint64_t)(k * (UMMA_K * 2 >> 4)), idesc,
                      (kb | k) != 0);
        tc_commit(&empty_bar[s]);
      }
      tc_commit(tmem_full_bar);
    }
  } else {
    const int t = threadIdx.x - 64;  // 0..127
    const int quad = warp & 3;
    if (LNB) {
      // ===== B producer: LayerNorm(x)[b_row0 .. +BN, this CTA's K slice] -> bf16, 128-byte-swizzled K-major tiles.
      // 16 consecutive threads read one 256-byte row segment (4 full lines per warp load); a thread keeps the same
      // float4 column group c4 in every k-block and walks the rows (t / 16) + 8 * it.
      constexpr int V4 = BN / 8;              // float4 loads per thread per k-block
      // k-blocks loaded together: 16 float4 in flight per thread (8 at BN <= 32; 8 k-blocks at once at BN = 16
      // measured slower: 0.621 vs 0.612 ms per decode forward)
      constexpr int KG = BN >= 64 ? 128 / BN : 4;
      const int c4 = t & 15, rsub = t >> 4;
      // gamma / beta of the K slice are weights: staged in shared memory ahead of the programmatic-launch wait
      float* sgam = ln_smem;
      float* sbet = sgam + nk * BK;
      float* row_mean = sbet + nk * BK;  // [BN]
      float* row_rstd = row_mean + BN;   // [BN]
      for (int i = t; i < nk * BK / 4; i += 128) {
        reinterpret_cast<float4*>(sgam)[i] = reinterpret_cast<const float4*>(ln.gamma + (size_t)kb0 * BK)[i];
        reinterpret_cast<float4*>(sbet)[i] = reinterpret_cast<const float4*>(ln.beta + (size_t)kb0 * BK)[i];
      }
      pdl_wait();  // x and its statistics belong to the predecessor until here
      if (threadIdx.x == 64) FTRACE(2);
      const float* xcol = ln.x + (size_t)kb0 * BK + c4 * 4;
      for (int kg = 0; kg < nk; kg += KG) {
        float4 xv[KG][V4];  // every global load of the group is in flight before the first use
#pragma unroll
        for (int j = 0; j < KG; ++j)
          if (kg + j < nk) {
#pragma unroll
            for (int it = 0; it < V4; ++it) {
              const int m = b_row0 + rsub + it * 8;
              xv[j][it] = m < rowsB ? *reinterpret_cast<const float4*>(xcol + (size_t)m * K + (kg + j) * BK)
                                    : make_float4(0.f, 0.f, 0.f, 0.f);
            }
          }
        if (kg == 0) {
          // LayerNorm statistics of each row from the per-tile (mean, M2) pairs (Chan et al. combination, equal counts)
          if (t < BN) {
            const int m = b_row0 + t;
            float mean = 0.f, rstd = 0.f;
            if (m < rowsB) row_mean_rstd(ln.stats, ln.nst, rowsB, m, K, ln.eps, mean, rstd);
            row_mean[t] = mean;
            row_rstd[t] = rstd;
          }
          asm volatile("bar.sync 1, 128;" ::: "memory");  // also orders the gamma / beta staging above
          if (threadIdx.x == 64) FTRACE(3);
        }
#pragma unroll
        for (int j = 0; j < KG; ++j) {
          const int kb = kg + j;
          if (kb < nk) {
            const int s = kb % L::STAGES;
            if (kb >= L::STAGES) mbar_wait(&empty_bar[s], ((kb / L::STAGES) & 1) ^ 1, 4);
            const uint32_t sb = smem_u32(tiles + s * L::STAGE_BYTES + L::A_BYTES);
            const float4 g = *reinterpret_cast<const float4*>(sgam + kb * BK + c4 * 4);
            const float4 be = *reinterpret_cast<const float4*>(sbet + kb * BK + c4 * 4);
#pragma unroll
            for (int it = 0; it < V4; ++it) {
              const int row = rsub + it * 8;
              const float mean = row_mean[row], rstd = row_rstd[row];
              const float4 a = xv[j][it];
              // rows past the batch have rstd = mean = 0 and x = 0: they come out as beta, which only ever reaches
              // accumulator columns nobody reads (TMA's out-of-bounds fill would have written zeros)
              __nv_bfloat162 lo = __floats2bfloat162_rn((a.x - mean) * rstd * g.x + be.x, (a.y - mean) * rstd * g.y + be.y);
              __nv_bfloat162 hi = __floats2bfloat162_rn((a.z - mean) * rstd * g.z + be.z, (a.w - mean) * rstd * g.w + be.w);
              // SWIZZLE_128B: 16-byte chunk index XOR (row mod 8) inside each 128-byte row of the 1 KB-aligned tile
              const uint32_t addr = sb + row * 128 + ((((uint32_t)c4 >> 1) ^ ((uint32_t)row & 7u)) << 4) + ((c4 & 1) << 3);
              st_shared_v2(addr, *reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> visible to the MMA
            __syncwarp();
            if (lane == 0) mbar_arrive(&full_bar[s]);
          }
        }
      }
      if (threadIdx.x == 64) FTRACE(4);
    } else {
      pdl_wait();  // the output buffers belong to the predecessor until here
      if (fold) {  // row statistics of the folded LayerNorm: requested now, they arrive under the mainloop
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int row = r0 + (warp - 2) + 4 * i;
          if (row < r1 && b_row0 + row < rowsB)
            row_mean_rstd(ep.fold_stats, ep.fold_nst, rowsB, b_row0 + row, ep.fold_k, ep.fold_eps, mean_pre[i], rstd_pre[i]);
        }
      }
    }
    // ===== epilogue part 1: accumulator -> fp32 partial tile [batch][128 n] in the (now idle) pipeline smem =====
    mbar_wait(tmem_full_bar, 0, 2);
    tc_fence_after();
    if (threadIdx.x == 64) FTRACE(5);
    const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16);
    float* stage = reinterpret_cast<float*>(tiles);
    for (int c0 = 0; c0 < BN; c0 += 32) {
      float v[32];
      tc_ld32(taddr + c0, v);
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (c0 + i < BN) stage[(c0 + i) * BM + quad * 32 + lane] = v[i];
    }
    if (threadIdx.x == 64) FTRACE(6);
    if (MODE != MODE_PARTIAL && n_ok) {
      bias4 = *reinterpret_cast<const float4*>(ep.bias + n);
      if (fold) {
        s4 = *reinterpret_cast<const float4*>(ep.fold_s + n);
        const float4 c4 = *reinterpret_cast<const float4*>(ep.fold_c + n);
        bias4.x += c4.x; bias4.y += c4.y; bias4.z += c4.z; bias4.w += c4.w;
      }
      if (MODE == MODE_RESID) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int row = r0 + (warp - 2) + 4 * i;
          if (row < r1 && b_row0 + row < rowsB)
            res_pre[i] = *reinterpret_cast<const float4*>(ep.x + (size_t)(b_row0 + row) * rowsA + n);
        }
      }
    }
    if (MODE == MODE_PARTIAL) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (threadIdx.x == 64) tma_store_3d(&tmC, tiles, a_row0, b_row0, (int)blockIdx.z);  // OOB rows/cols clipped
    }
  }

  if (MODE != MODE_PARTIAL) {
    // ===== epilogue part 2: cluster-wide reduction over the K splits, straight out of the peers' shared memory =====
    __syncwarp();
    cluster_sync_all();  // every CTA of the cluster has staged its partial tile
    if (threadIdx.x == 64) FTRACE(7);
    if (warp >= 2) {
      const uint32_t stage_u32 = smem_u32(tiles);
      // two rows per pass: 2 * ks remote 16-byte loads in flight per lane; sums in ascending split order (deterministic)
      auto reduce_pair = [&](int i, float4 resA, float4 resB, float meanA, float rstdA, float meanB, float rstdB) {
        const int rowA = r0 + (warp - 2) + 8 * i, rowB = rowA + 4;
        const bool okA = rowA < r1 && b_row0 + rowA < rowsB, okB = rowB < r1 && b_row0 + rowB < rowsB;
        if (!okA) return;
        float4 pa[8], pb[8];
        const uint32_t offA = stage_u32 + (uint32_t)(rowA * BM + lane * 4) * 4u;
        const uint32_t offB = stage_u32 + (uint32_t)(rowB * BM + lane * 4) * 4u;
#pragma unroll
        for (int z = 0; z < 8; ++z)
          if (z < (int)ksz) {
            pa[z] = ld_dsmem_f4(dsmem_addr(offA, (uint32_t)z));
            if (okB) pb[z] = ld_dsmem_f4(dsmem_addr(offB, (uint32_t)z));
          }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          if (h == 1 && !okB) break;
          const int m = b_row0 + (h ? rowB : rowA);
          float4 acc = fold ? make_float4(0.f, 0.f, 0.f, 0.f) : bias4;
#pragma unroll
          for (int z = 0; z < 8; ++z)
            if (z < (int)ksz) {
              const float4 q = h ? pb[z] : pa[z];
              acc.x += q.x; acc.y += q.y; acc.z += q.z; acc.w += q.w;
            }
          if (fold) {  // rstd (x . W' - mean s) + c + bias
            const float mean = h ? meanB : meanA, rstd = h ? rstdB : rstdA;
            acc.x = fmaf(rstd, acc.x - mean * s4.x, bias4.x); acc.y = fmaf(rstd, acc.y - mean * s4.y, bias4.y);
            acc.z = fmaf(rstd, acc.z - mean * s4.z, bias4.z); acc.w = fmaf(rstd, acc.w - mean * s4.w, bias4.w);
          }
          if (MODE == MODE_ACT) {
            if (n_ok) {
              __nv_bfloat162 lo = __floats2bfloat162_rn(apply_act(acc.x, ep.act), apply_act(acc.y, ep.act));
              __nv_bfloat162 hi = __floats2bfloat162_rn(apply_act(acc.z, ep.act), apply_act(acc.w, ep.act));
              *reinterpret_cast<uint2*>(ep.out_bf16 + (size_t)m * rowsA + n) =
                  make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
            }
          } else {
            const float4 res = h ? resB : resA;
            acc.x += res.x; acc.y += res.y; acc.z += res.z; acc.w += res.w;
            if (n_ok) {
              *reinterpret_cast<float4*>(ep.x + (size_t)m * rowsA + n) = acc;
              if (ep.xb) {
                __nv_bfloat162 lo = __floats2bfloat162_rn(acc.x, acc.y), hi = __floats2bfloat162_rn(acc.z, acc.w);
                *reinterpret_cast<uint2*>(ep.xb + (size_t)m * rowsA + n) =
                    make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
              }
            }
            // (mean, M2) of this row over the tile's 128 columns (rowsA % 128 == 0 in this mode)
            const float mean_t = warp_sum((acc.x + acc.y) + (acc.z + acc.w)) * (1.0f / BM);
            const float a = acc.x - mean_t, b = acc.y - mean_t, c = acc.z - mean_t, d = acc.w - mean_t;
            const float m2_t = warp_sum((a * a + b * b) + (c * c + d * d));
            if (lane == 0)
              *reinterpret_cast<float2*>(ep.stats + ((size_t)blockIdx.x * rowsB + m) * 2) = make_float2(mean_t, m2_t);
          }
        }
      };
      reduce_pair(0, res_pre[0], res_pre[1], mean_pre[0], rstd_pre[0], mean_pre[1], rstd_pre[1]);
      reduce_pair(1, res_pre[2], res_pre[3], mean_pre[2], rstd_pre[2], mean_pre[3], rstd_pre[3]);
      for (int i = 2; r0 + (warp - 2) + 8 * i < r1; ++i) {  // more than 16 rows per CTA (large batch tile, few splits)
        float4 ra = make_float4(0.f, 0.f, 0.f, 0.f), rb = ra;
        if (MODE == MODE_RESID && n_ok) {
          const int rowA = r0 + (warp - 2) + 8 * i, rowB = rowA + 4;
          if (rowA < r1 && b_row0 + rowA < rowsB) ra = *reinterpret_cast<const float4*>(ep.x + (size_t)(b_row0 + rowA) * rowsA + n);
          if (rowB < r1 && b_row0 + rowB < rowsB) rb = *reinterpret_cast<const float4*>(ep.x + (size_t)(b_row0 + rowB) * rowsA + n);
        }
        float ma = 0.f, sa = 1.f, mb = 0.f, sb = 1.f;
        if (fold) {
          const int rowA = r0 + (warp - 2) + 8 * i, rowB = rowA + 4;
          if (rowA < r1 && b_row0 + rowA < rowsB) row_mean_rstd(ep.fold_stats, ep.fold_nst, rowsB, b_row0 + rowA, ep.fold_k, ep.fold_eps, ma, sa);
          if (rowB < r1 && b_row0 + rowB < rowsB) row_mean_rstd(ep.fold_stats, ep.fold_nst, rowsB, b_row0 + rowB, ep.fold_k, ep.fold_eps, mb, sb);
        }
        reduce_pair(i, ra, rb, ma, sa, mb, sb);
      }
    }
    if (threadIdx.x == 64) FTRACE(8);
    __syncwarp();
    cluster_sync_all();  // peers may have been reading this CTA's shared memory until here
    if (threadIdx.x == 64) FTRACE(9);
  }

  tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 64) FTRACE(10);
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
  }
}

// per-row (mean, M2) over each 128-column tile of x [M, d]: seeds the statistics chain (after the embedding)
__global__ void __launch_bounds__(128) row_tile_stats_kernel(const float* __restrict__ x, float* __restrict__ stats, int M,
                                                           int d, bf16* __restrict__ xb) {
  pdl_launch_dependents();
  pdl_wait();
  const int m = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (m >= M) return;
  for (int tile = 0; tile < d / BM; ++tile) {
    const float4 v = *reinterpret_cast<const float4*>(x + (size_t)m * d + tile * BM + lane * 4);
    const float mean_t = warp_sum((v.x + v.y) + (v.z + v.w)) * (1.0f / BM);
    const float a = v.x - mean_t, b = v.y - mean_t, c = v.z - mean_t, e = v.w - mean_t;
    const float m2_t = warp_sum((a * a + b * b) + (c * c + e * e));
    if (lane == 0) *reinterpret_cast<float2*>(stats + ((size_t)tile * M + m) * 2) = make_float2(mean_t, m2_t);
    if (xb) {  // bf16 copy of x: the operand of a folded-LayerNorm GEMM
      __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
      *reinterpret_cast<uint2*>(xb + (size_t)m * d + tile * BM + lane * 4) =
          make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
    }
  }
}

template <int BN, bool LNB, int MODE>
int launch(const bf16* W, int N, const bf16* X, int M, int K, int ks, const LnOperand& ln, const FusedEpi& ep,
           float* partial, cudaStream_t st) {
  using L = SmemLayout<BN>;
  // LNB: gamma and beta of one CTA's K slice and the row statistics are staged behind the barriers
  const int SMEM = L::TOTAL + (LNB ? 2 * (K / ks + BN) * (int)sizeof(float) : 0);
  ARG_CHECK(SMEM <= 200 * 1024, "decode gemm: K slice %d too long for the shared-memory staging", K / ks);
  auto kern = gemm_decode_fused_kernel<BN, LNB, MODE>;
  TRY(tutor_set_max_smem(reinterpret_cast<const void*>(kern), 200 * 1024));
  CUtensorMap ta, tb, tc;
  TRY(k_make_map_bf16(&ta, W, N, K, BM));
  if (!LNB) TRY(k_make_map_bf16(&tb, X, M, K, BN));
  else tb = ta;  // unused
  if (MODE == MODE_PARTIAL) TRY(k_make_map_partial(&tc, partial, N, M, ks, BN));
  else tc = ta;  // unused
  dim3 grid(ceil_div(N, BM), ceil_div(M, BN), ks);
  CUDA_TRY(tutor_launch_cluster(kern, grid, dim3(NTHREADS), (size_t)SMEM, st, MODE == MODE_PARTIAL ? 1 : ks, ta, tb, tc, N,
                                M, K, ln, ep));
  KERNEL_LAUNCHED();
  return TUTOR_OK;
}

template <bool LNB, int MODE>
int launch_bn(int bn, const bf16* W, int N, const bf16* X, int M, int K, int ks, const LnOperand& ln, const FusedEpi& ep,
              float* partial, cudaStream_t st) {
  if (bn <= 16) return launch<16, LNB, MODE>(W, N, X, M, K, ks, ln, ep, partial, st);
  if (bn <= 32) return launch<32, LNB, MODE>(W, N, X, M, K, ks, ln, ep, partial, st);
  if (bn <= 64) return launch<64, LNB, MODE>(W, N, X, M, K, ks, ln, ep, partial, st);
  return launch<128, LNB, MODE>(W, N, X, M, K, ks, ln, ep, partial, st);
}

int check_common(int M, int N, int K, int ks, int bn) {
  ARG_CHECK(K % BK == 0 && ks >= 1 && (K / BK) % ks == 0, "decode gemm: K=%d not divisible into %d splits", K, ks);
  ARG_CHECK(M >= 1 && M <= 128 && N % 4 == 0, "decode gemm: batch %d outside 1..128 or N=%d not a multiple of 4", M, N);
  ARG_CHECK(bn == 16 || bn == 32 || bn == 64 || bn == 128, "decode gemm: batch tile %d", bn);
  return TUTOR_OK;
}

}  // namespace

// smallest batch tile (UMMA N) that covers M in one tile
int k_decode_batch_tile(int M) { return M <= 16 ? 16 : M <= 32 ? 32 : M <= 64 ? 64 : 128; }

// largest split count <= 8 (portable cluster size) that divides K/64 and keeps tiles * batch_tiles * ks <= max_ctas
int k_pick_cluster_splitk(int N, int K, int batch_tiles, int max_ctas) {
  const int tiles = ceil_div(N, BM) * batch_tiles, nk = K / BK;
  int best = 1;
  for (int ks = 1; ks <= nk && ks <= 8; ++ks)
    if (nk % ks == 0 && tiles * ks <= max_ctas) best = ks;
  return best;
}

int k_row_tile_stats(const float* x, float* stats, int M, int d, bf16* xb, cudaStream_t st) {
  ARG_CHECK(d % BM == 0, "row_tile_stats: d=%d must be a multiple of %d", d, BM);
  if (M == 0) return TUTOR_OK;
  CUDA_TRY(tutor_launch(row_tile_stats_kernel, dim3(ceil_div(M, 4)), dim3(128), 0, st, x, stats, M, d, xb));
  KERNEL_LAUNCHED();
  return TUTOR_OK;
}

// partial[z][m][n] = sum over K slice z of LayerNorm(x)[m,:] . W[n,:]      (c_attn; the attention kernel reduces)
int k_gemm_ln_splitk(const float* x, const float* stats, int nst, const float* gamma, const float* beta, float eps,
                     const bf16* W, int M, int N, int K, int ks, int bn, float* partial, cudaStream_t st) {
  TRY(check_common(M, N, K, ks, bn));
  ARG_CHECK(nst >= 1 && K % nst == 0, "decode gemm: %d statistic tiles do not divide K=%d", nst, K);
  LnOperand ln{x, stats, nst, gamma, beta, eps};
  FusedEpi ep{};
  return launch_bn<true, MODE_PARTIAL>(bn, W, N, nullptr, M, K, ks, ln, ep, partial, st);
}

// out[m,n] = act(bias[n] + LayerNorm(x)[m,:] . W[n,:]) as bf16                       (c_fc + gelu_new)
int k_gemm_ln_act(const float* x, const float* stats, int nst, const float* gamma, const float* beta, float eps,
                  const bf16* W, const float* bias, int act, int M, int N, int K, int ks, int bn, bf16* out,
                  cudaStream_t st) {
  TRY(check_common(M, N, K, ks, bn));
  ARG_CHECK(ks <= 8, "decode gemm: cluster of %d CTAs exceeds the portable size 8", ks);
  ARG_CHECK(nst >= 1 && K % nst == 0, "decode gemm: %d statistic tiles do not divide K=%d", nst, K);
  LnOperand ln{x, stats, nst, gamma, beta, eps};
  FusedEpi ep{};
  ep.bias = bias; ep.act = act; ep.out_bf16 = out;
  return launch_bn<true, MODE_ACT>(bn, W, N, nullptr, M, K, ks, ln, ep, nullptr, st);
}

// x[m,n] += bias[n] + A[m,:] . W[n,:]  (fp32, in place);  stats[n/128][m] = (mean, M2) of the new x over that tile
int k_gemm_resid_stats(const bf16* A, const bf16* W, const float* bias, int M, int N, int K, int ks, int bn, float* x,
                       float* stats, bf16* xb, cudaStream_t st) {
  TRY(check_common(M, N, K, ks, bn));
  ARG_CHECK(ks <= 8, "decode gemm: cluster of %d CTAs exceeds the portable size 8", ks);
  ARG_CHECK(N % BM == 0, "decode gemm: N=%d must be a multiple of %d for the statistics", N, BM);
  LnOperand ln{};
  FusedEpi ep{};
  ep.bias = bias; ep.act = ACT_NONE; ep.x = x; ep.stats = stats; ep.xb = xb;
  return launch_bn<false, MODE_RESID>(bn, W, N, A, M, K, ks, ln, ep, nullptr, st);
}

// ---- folded LayerNorm (see the head of this file): xb is the raw residual stream as bf16, Wf = bf16(gamma (.) W)

// partial[z][m][n] = sum over K slice z of xb[m,:] . Wf[n,:]        (c_attn; the attention kernel closes sum + LayerNorm)
int k_gemm_fold_splitk(const bf16* xb, const bf16* Wf, int M, int N, int K, int ks, int bn, float* partial, cudaStream_t st) {
  TRY(check_common(M, N, K, ks, bn));
  LnOperand ln{};
  FusedEpi ep{};
  return launch_bn<false, MODE_PARTIAL>(bn, Wf, N, xb, M, K, ks, ln, ep, partial, st);
}

// out[m,n] = act(rstd[m] (xb[m,:] . Wf[n,:] - mean[m] s[n]) + c[n] + bias[n]) as bf16                 (c_fc + gelu_new)
int k_gemm_fold_act(const bf16* xb, const float* stats, int nst, float eps, const bf16* Wf, const float* s, const float* c,
                    const float* bias, int act, int M, int N, int K, int ks, int bn, bf16* out, cudaStream_t st) {
  TRY(check_common(M, N, K, ks, bn));
  ARG_CHECK(ks <= 8, "decode gemm: cluster of %d CTAs exceeds the portable size 8", ks);
  ARG_CHECK(nst >= 1 && nst <= MAX_STAT_TILES && K % nst == 0, "decode gemm: %d statistic tiles do not divide K=%d", nst, K);
  LnOperand ln{};
  FusedEpi ep{};
  ep.bias = bias; ep.act = act; ep.out_bf16 = out;
  ep.fold_s = s; ep.fold_c = c; ep.fold_stats = stats; ep.fold_nst = nst; ep.fold_k = K; ep.fold_eps = eps;
  return launch_bn<false, MODE_ACT>(bn, Wf, N, xb, M, K, ks, ln, ep, nullptr, st);
}
