// Prefill / bidirectional attention on the 5th-generation tensor cores (bf16, head_dim 64).
//   causal:        GPT-2 prefill, HF/integrations/sdpa_attention.py:77 (is_causal iff T > 1)
//   bidirectional: BERT, HF/models/bert/modeling_bert.py:192-205 (through SDPA)
//
// One CTA per (128 queries, head, sequence), flash-style over 128-key blocks:
//   S = Q K^T        tcgen05.mma 128x128x64, Q and K tiles K-major from TMA (SWIZZLE_128B), S in TMEM (128 columns)
//   P = softmax tile four warps, one query row per thread: two passes over the S row in TMEM (row maximum, then
//                    exp2 / row sum), P written as bf16 into a swizzled shared-memory tile that is the A operand of
//   O += P V         tcgen05.mma 128x64x128, V tile straight from the cache layout [key][dim] as an MN-major B operand,
//                    O accumulated in TMEM (64 columns).
// The running maximum is only moved when a block's maximum exceeds it by more than 8 (in log2 units): P then stays
// below 2^8, and O / l is exact for any reference maximum.  Moving it means rescaling O in TMEM (load, scale, store),
// which happens for the first few blocks of a row only.
// Warp roles: 0 = TMA producer, 1 = MMA issuer + TMEM allocation, 2..5 = softmax / epilogue.  Two CTAs share an SM
// (112 KB of shared memory and 256 TMEM columns each): while one CTA's softmax warps work on a block - the exponentials
// bound this kernel, not the MMAs - the other CTA's MMAs run.
#include <math_constants.h>

#include "kernels.h"
#include "tc_ptx.cuh"

namespace {
using namespace tcptx;

constexpr int AT_THREADS = 192;
constexpr int AT_TILE = 128 * 128;             // bytes of a 128-row x 64-column bf16 tile
constexpr int AT_SMEM = 7 * AT_TILE + 128;     // Q, K[2], V[2], P[2 k-blocks] + barriers
constexpr uint32_t AT_TMEM_COLS = 256;         // S: columns 0..127, O: 128..191

// tcgen05.st 32 lanes x 32 columns
__device__ __forceinline__ void tc_st32(uint32_t taddr, const float* v) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
      "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]), "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7]), "f"(v[8]), "f"(v[9]), "f"(v[10]),
      "f"(v[11]), "f"(v[12]), "f"(v[13]), "f"(v[14]), "f"(v[15]), "f"(v[16]), "f"(v[17]), "f"(v[18]), "f"(v[19]), "f"(v[20]),
      "f"(v[21]), "f"(v[22]), "f"(v[23]), "f"(v[24]), "f"(v[25]), "f"(v[26]), "f"(v[27]), "f"(v[28]), "f"(v[29]), "f"(v[30]),
      "f"(v[31])
      : "memory");
}
__device__ __forceinline__ void tc_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// B operand given as [k][n] with n contiguous (MN-major), 64 n-elements = one 128-byte swizzle atom per k row:
// 8 k-rows form a 1024-byte group (stride byte offset), the leading byte offset (between 64-element n groups) is unused
__device__ __forceinline__ uint64_t umma_desc_sw128_mn(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
  d |= (uint64_t)(1024 >> 4) << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  __nv_bfloat162 p = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&p);
}

template <bool CAUSAL, bool PADDED>
__global__ void __launch_bounds__(AT_THREADS, 2)
    attn_prefill_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                           const __grid_constant__ CUtensorMap tmV, bf16* __restrict__ out, int H, int T, int Smax,
                           const int* __restrict__ npad) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sQ = smem;
  uint8_t* sK = smem + AT_TILE;      // [2]
  uint8_t* sV = smem + 3 * AT_TILE;  // [2]
  uint8_t* sP = smem + 5 * AT_TILE;  // [2 k-blocks of 64 keys]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 7 * AT_TILE);
  uint64_t* q_full = bars;
  uint64_t* k_full = bars + 1;    // [2]
  uint64_t* v_full = bars + 3;    // [2]
  uint64_t* kv_empty = bars + 5;  // [2]
  uint64_t* s_full = bars + 7;
  uint64_t* p_ready = bars + 8;
  uint64_t* pv_done = bars + 9;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 10);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int qt = (int)gridDim.x - 1 - (int)blockIdx.x;  // the longest (causal) tiles first
  const int h = blockIdx.y, b = blockIdx.z;
  const int q0 = qt * 128;
  const int nblk_all = (T + 127) / 128;
  const int nblk = CAUSAL ? min(qt + 1, nblk_all) : nblk_all;

  if (threadIdx.x == 0) {
    if (smem_u32(smem) & 1023) __trap();  // the swizzled tiles need 1 KB alignment and there is no room for slack
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmQ) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmK) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmV) : "memory");
    for (int i = 0; i < 8; ++i) mbar_init(&bars[i], 1);
    mbar_init(p_ready, 128);
    mbar_init(pv_done, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(AT_TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tS = tmem_base, tO = tmem_base + 128;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      mbar_expect_tx(q_full, AT_TILE);
      tma_load_2d(sQ, &tmQ, q_full, h * 64, b * T + q0);
      const int kv_row0 = (b * H + h) * Smax;
      for (int j = 0; j < nblk; ++j) {
        const int st = j & 1;
        mbar_wait(&kv_empty[st], ((j >> 1) & 1) ^ 1, 0);
        mbar_expect_tx(&k_full[st], AT_TILE);
        tma_load_2d(sK + st * AT_TILE, &tmK, &k_full[st], 0, kv_row0 + j * 128);
        mbar_expect_tx(&v_full[st], AT_TILE);
        tma_load_2d(sV + st * AT_TILE, &tmV, &v_full[st], 0, kv_row0 + j * 128);
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread) =====
    if (lane == 0) {
      constexpr uint32_t idesc_s = umma_idesc_bf16(128, 128);
      constexpr uint32_t idesc_o = umma_idesc_bf16(128, 64) | (1u << 16);  // B (= V) is MN-major
      auto issue_pv = [&](int j) {  // O (+)= P_j V_j
        const int st = j & 1;
        mbar_wait(&v_full[st], (j >> 1) & 1, 1);
        tc_fence_after();
        const uint32_t sp = smem_u32(sP), sv = smem_u32(sV + st * AT_TILE);
#pragma unroll
        for (int k = 0; k < 8; ++k)  // 16 keys per MMA: k-block k/4 of P, 32 bytes along its rows; two 8-key groups of V
          tc_mma_bf16(tO, umma_desc_sw128(sp + (k >> 2) * AT_TILE + (k & 3) * 32), umma_desc_sw128_mn(sv + k * 2048), idesc_o,
                      (j | k) != 0);
        tc_commit(pv_done);
        tc_commit(&kv_empty[st]);
      };
      mbar_wait(q_full, 0, 1);
      for (int j = 0; j < nblk; ++j) {
        const int st = j & 1;
        mbar_wait(&k_full[st], (j >> 1) & 1, 1);
        if (j > 0) mbar_wait(p_ready, (j - 1) & 1, 1);  // S has been read (and P_{j-1} written)
        tc_fence_after();
        const uint32_t sq = smem_u32(sQ), sk = smem_u32(sK + st * AT_TILE);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          tc_mma_bf16(tS, umma_desc_sw128(sq + k * 32), umma_desc_sw128(sk + k * 32), idesc_s, k != 0);
        tc_commit(s_full);
        if (j > 0) issue_pv(j - 1);  // behind S_j in the tensor pipe: the softmax of block j overlaps it
      }
      mbar_wait(p_ready, (nblk - 1) & 1, 1);
      issue_pv(nblk - 1);
    }
  } else {
    // ===== softmax + epilogue: thread = query row; a warp reaches the 32 TMEM lanes of its quadrant =====
    const int quad = warp & 3;
    const int r = quad * 32 + lane;
    const int qrow = q0 + r;
    const uint32_t lane_off = (uint32_t)(quad * 32) << 16;
    const int kp = PADDED ? npad[b] : 0;  // left padding: pad keys are invisible to real queries
    constexpr float C = 0.125f * 1.4426950408889634f;  // 1/sqrt(64) and log2(e) in one factor
    float m_run = -CUDART_INF_F, l_run = 0.f;
    uint8_t* prow = sP + r * 128;
    for (int j = 0; j < nblk; ++j) {
      const int key0 = j * 128;
      const bool edge = (key0 + 127 >= T) || (CAUSAL && key0 + 127 > q0) || (PADDED && key0 < kp);
      mbar_wait(s_full, j & 1, 2);
      tc_fence_after();
      // ---- pass 1: row maximum
      float mx = -CUDART_INF_F;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        float s[32];
        tc_ld32(tS + lane_off + c * 32, s);
        if (edge) {
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const int key = key0 + c * 32 + i;
            if (key >= T || (CAUSAL && key > qrow) || (PADDED && key < kp && qrow >= kp)) s[i] = -CUDART_INF_F;
          }
        }
#pragma unroll
        for (int i = 0; i < 32; ++i) mx = fmaxf(mx, s[i]);
      }
      const float mb = mx * C;
      // ---- move the reference maximum (rarely): O and l are rescaled
      const bool move = mb > m_run + 8.0f;
      float alpha = 1.0f;
      if (move) {
        alpha = ex2(m_run - mb);  // 0 for the first block
        m_run = mb;
        l_run *= alpha;
      }
      if (j > 0) {
        mbar_wait(pv_done, (j - 1) & 1, 3);  // P V of the previous block has read P and finished with O
        tc_fence_after();
        if (__any_sync(0xffffffffu, move)) {
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            float o[32];
            tc_ld32(tO + lane_off + c * 32, o);
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] *= alpha;
            tc_st32(tO + lane_off + c * 32, o);
          }
          tc_wait_st();
        }
      }
      // ---- pass 2: P = exp2(S C - m), row sum, bf16 into the swizzled A tile
      const float mref = (m_run == -CUDART_INF_F) ? 0.f : m_run;  // a fully masked row stays at zero weight
      float lsum = 0.f;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        float s[32];
        tc_ld32(tS + lane_off + c * 32, s);
        if (edge) {
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const int key = key0 + c * 32 + i;
            if (key >= T || (CAUSAL && key > qrow) || (PADDED && key < kp && qrow >= kp)) s[i] = -CUDART_INF_F;
          }
        }
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          s[i] = ex2(fmaf(s[i], C, -mref));
          lsum += s[i];
        }
        uint8_t* blk = prow + (c >> 1) * AT_TILE;  // k-block of 64 keys
#pragma unroll
        for (int g = 0; g < 4; ++g) {              // 8 keys = one 16-byte chunk
          const int ch = (c & 1) * 4 + g;
          uint4 v;
          v.x = pack2(s[g * 8 + 0], s[g * 8 + 1]); v.y = pack2(s[g * 8 + 2], s[g * 8 + 3]);
          v.z = pack2(s[g * 8 + 4], s[g * 8 + 5]); v.w = pack2(s[g * 8 + 6], s[g * 8 + 7]);
          *reinterpret_cast<uint4*>(blk + ((ch ^ (r & 7)) << 4)) = v;
        }
      }
      l_run += lsum;
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy smem writes -> the MMA's async proxy
      tc_fence_before();
      mbar_arrive(p_ready);
    }
    // ---- epilogue: O / l -> bf16 [B*T, d]
    mbar_wait(pv_done, (nblk - 1) & 1, 4);
    tc_fence_after();
    const float inv = 1.0f / l_run;
    bf16* dst = out + ((size_t)b * T + qrow) * (H * 64) + h * 64;
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      float o[32];
      tc_ld32(tO + lane_off + c * 32, o);
      if (qrow < T) {
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint4 v;
          v.x = pack2(o[g * 8 + 0] * inv, o[g * 8 + 1] * inv); v.y = pack2(o[g * 8 + 2] * inv, o[g * 8 + 3] * inv);
          v.z = pack2(o[g * 8 + 4] * inv, o[g * 8 + 5] * inv); v.w = pack2(o[g * 8 + 6] * inv, o[g * 8 + 7] * inv);
          *reinterpret_cast<uint4*>(dst + c * 32 + g * 8) = v;
        }
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(AT_TMEM_COLS) : "memory");
  }
}

}  // namespace

// q [B*T, H*64] bf16, k / v [B][H][Smax][64] bf16 (cache layout), out [B*T, H*64] bf16
int k_attn_prefill_tc(const bf16* q, const bf16* k, const bf16* v, bf16* out, int B, int H, int T, int Smax, int causal,
                      const int* npad, cudaStream_t st) {
  CUtensorMap tq, tk, tv;
  TRY(k_make_map_bf16(&tq, q, B * T, H * 64, 128));
  TRY(k_make_map_bf16(&tk, k, B * H * Smax, 64, 128));
  TRY(k_make_map_bf16(&tv, v, B * H * Smax, 64, 128));
  dim3 grid(ceil_div(T, 128), H, B);
#define AT_LAUNCH(C, P)                                                                                      \
  do {                                                                                                       \
    auto kern = attn_prefill_tc_kernel<C, P>;                                                                \
    TRY(tutor_set_max_smem(reinterpret_cast<const void*>(kern), AT_SMEM));                                   \
    cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared); \
    kern<<<grid, AT_THREADS, AT_SMEM, st>>>(tq, tk, tv, out, H, T, Smax, npad);                              \
  } while (0)
  if (causal && npad) AT_LAUNCH(true, true);
  else if (causal) AT_LAUNCH(true, false);
  else if (npad) AT_LAUNCH(false, true);
  else AT_LAUNCH(false, false);
#undef AT_LAUNCH
  CUDA_TRY(cudaGetLastError());
  KERNEL_LAUNCHED();
  return TUTOR_OK;
}
