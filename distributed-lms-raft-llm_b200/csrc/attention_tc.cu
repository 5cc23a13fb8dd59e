// Prefill / bidirectional attention on the 5th-generation tensor cores (bf16, head_dim 64).
//   causal:        GPT-2 prefill, HF/integrations/sdpa_attention.py:77 (is_causal iff T > 1)
//   bidirectional: BERT, HF/models/bert/modeling_bert.py:192-205 (through SDPA)
//
// A work item is (128 queries, head, sequence), processed flash-style over 64-key blocks:
//   S = Q K^T        tcgen05.mma 128x64x64, Q and K tiles K-major from TMA (SWIZZLE_128B), S in TMEM (two 64-column
//                    buffers: the MMA of block g + 1 runs while the softmax warps work on block g)
//   P = softmax tile four warps, one query row per thread, the 64 scores of the row in registers; P written as bf16
//                    into a swizzled shared-memory tile (two buffers) that is the A operand of
//   O += P V         tcgen05.mma 128x64x64, V tile straight from the cache layout [key][dim] as an MN-major B operand,
//                    O accumulated in TMEM (64 columns).
// The running maximum is only moved when a block's maximum exceeds it by more than 8 (in log2 units): P then stays
// below 2^8, and O / l is exact for any reference maximum.  Moving it means rescaling O in TMEM (load, scale, store),
// which happens for the first few blocks of a row only.
// Persistent: two CTAs per SM walk the items, longest first; barriers, TMEM, the K/V ring (three stages) and a second Q
// buffer carry over from item to item, so an item's first tiles arrive while the previous one finishes.
// Warp roles: 0 = TMA producer, 1 = MMA issuer + TMEM allocation, 2..5 = softmax / epilogue.
//
// Measured (BERT 32 x 512 x 12 heads, 25.8 GFLOP; mma.sync kernel 97.7 us): 128-key blocks, one CTA per item, two CTAs per
// SM 71.7 us; persistent 66.0 us; two threads per row (eight softmax warps) 92.8 us; 64-key blocks, three CTAs per SM
// 68.9 us; this version 70 us, and 97 / 113 us against 135 / 180 us on the causal GPT-2 shapes (64 x 512 x 12, 16 x 1024 x 20).
// Per block a softmax warp spends ~110 cycles finding S ready, ~130 loading it from TMEM, ~700-900 on maximum + 64 exp2 per
// thread, ~250 on a barrier test, ~130 packing and storing P (profiles/r2c_attn_prefill_tc.txt); the MUFU pipe is 40 % busy,
// the tensor pipe 19 %.  Two co-resident CTAs slow each other by ~1.5x, so the SM, not a single chain, is the limit.
#include <math_constants.h>

#include "kernels.h"
#include "tc_ptx.cuh"

namespace {
using namespace tcptx;

constexpr int AT_THREADS = 192;               // TMA warp, MMA warp, 4 softmax warps
constexpr int AT_TILE = 128 * 128;             // bytes of a 128-row x 64-column bf16 tile (Q, P)
constexpr int AT_KV = 64 * 128;                // bytes of a 64-key K or V tile
constexpr int AT_BN = 64;                      // keys per block
constexpr int AT_ST = 3;                       // K / V stages
constexpr int AT_SMEM = 4 * AT_TILE + 2 * AT_ST * AT_KV + 256;  // Q[2], P[2], K[AT_ST], V[AT_ST], barriers
constexpr uint32_t AT_TMEM_COLS = 256;         // S[2]: columns 0..63 and 64..127, O: 128..191

// 32 lanes x 8 columns, multiplied in place (the rare rescale of O: narrow on purpose, the score row is live around it)
__device__ __forceinline__ void tc_scale8(uint32_t taddr, float f) {
  uint32_t r[8];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 8; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * f);
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr), "r"(r[0]), "r"(r[1]),
               "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
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
                           const __grid_constant__ CUtensorMap tmV, bf16* __restrict__ out, int B, int H, int T, int Smax,
                           const int* __restrict__ npad) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sQ = smem;                                    // [2]: the next item's queries arrive during this item
  uint8_t* sP = smem + 2 * AT_TILE;                      // [2]
  uint8_t* sK = smem + 4 * AT_TILE;                      // [AT_ST]
  uint8_t* sV = smem + 4 * AT_TILE + AT_ST * AT_KV;      // [AT_ST]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 4 * AT_TILE + 2 * AT_ST * AT_KV);
  uint64_t* k_full = bars;                 // [AT_ST]
  uint64_t* v_full = bars + AT_ST;         // [AT_ST]
  uint64_t* k_empty = bars + 2 * AT_ST;    // [AT_ST] freed by S = Q K^T
  uint64_t* v_empty = bars + 3 * AT_ST;    // [AT_ST] freed by O += P V
  uint64_t* s_full = bars + 4 * AT_ST;     // [2]
  uint64_t* q_full = s_full + 2;           // [2]
  uint64_t* q_empty = s_full + 4;          // [2]
  uint64_t* pv_done = s_full + 6;          // [2]: block g completes pv_done[g & 1]
  uint64_t* p_ready = s_full + 8;          // [2], 128 arrivals each
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(s_full + 10);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nqt = (T + 127) / 128, nitems = nqt * H * B;
  // Work items (query tile, head, sequence), the longest (causal) tiles first; CTA c takes items c, c + grid, ...
  // Barriers, TMEM and the K/V ring live across items: the next item's Q / K / V arrive while this one finishes.
  auto item = [&](int w, int& qt, int& h, int& b) {
    qt = nqt - 1 - w / (H * B);
    const int rest = w - (nqt - 1 - qt) * (H * B);
    h = rest % H;
    b = rest / H;
  };
  const int nkb = (T + AT_BN - 1) / AT_BN;
  auto blocks_of = [&](int qt) { return CAUSAL ? min(2 * (qt + 1), nkb) : nkb; };

  if (threadIdx.x == 0) {
    if (smem_u32(smem) & 1023) __trap();  // the swizzled tiles need 1 KB alignment and there is no room for slack
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmQ) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmK) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmV) : "memory");
    for (int i = 0; i < 4 * AT_ST + 8; ++i) mbar_init(&bars[i], 1);
    mbar_init(&p_ready[0], 128);
    mbar_init(&p_ready[1], 128);
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
  const uint32_t tS = tmem_base, tO = tmem_base + 128;  // S buffer i at tS + 64 i

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      int g = 0, it = 0;  // K/V blocks and items this CTA has started
      for (int w = blockIdx.x; w < nitems; w += gridDim.x, ++it) {
        int qt, h, b;
        item(w, qt, h, b);
        mbar_wait(&q_empty[it & 1], ((it >> 1) & 1) ^ 1, 0);  // the last S = Q K^T of two items ago has read this buffer
        mbar_expect_tx(&q_full[it & 1], AT_TILE);
        tma_load_2d(sQ + (it & 1) * AT_TILE, &tmQ, &q_full[it & 1], h * 64, b * T + qt * 128);
        const int kv_row0 = (b * H + h) * Smax, nblk = blocks_of(qt);
        for (int j = 0; j < nblk; ++j, ++g) {
          const int st = g % AT_ST;
          const uint32_t use = (uint32_t)(g / AT_ST) & 1;
          mbar_wait(&k_empty[st], use ^ 1, 0);
          mbar_expect_tx(&k_full[st], AT_KV);
          tma_load_2d(sK + st * AT_KV, &tmK, &k_full[st], 0, kv_row0 + j * AT_BN);
          mbar_wait(&v_empty[st], use ^ 1, 0);
          mbar_expect_tx(&v_full[st], AT_KV);
          tma_load_2d(sV + st * AT_KV, &tmV, &v_full[st], 0, kv_row0 + j * AT_BN);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread) =====
    if (lane == 0) {
      constexpr uint32_t idesc_s = umma_idesc_bf16(128, 64);
      constexpr uint32_t idesc_o = umma_idesc_bf16(128, 64) | (1u << 16);  // B (= V) is MN-major
      auto issue_pv = [&](int g, bool first) {  // O (+)= P V of block g (first: the item's first block starts O afresh)
        const int st = g % AT_ST;
        mbar_wait(&p_ready[g & 1], (g >> 1) & 1, 1);  // P of block g is in shared memory (and S buffer g & 1 has been read)
        mbar_wait(&v_full[st], (g / AT_ST) & 1, 1);
        tc_fence_after();
        const uint32_t sp = smem_u32(sP + (g & 1) * AT_TILE), sv = smem_u32(sV + st * AT_KV);
#pragma unroll
        for (int k = 0; k < 4; ++k)  // 16 keys per MMA: 32 bytes along the rows of P; two 8-key groups (2 x 1024 bytes) of V
          tc_mma_bf16(tO, umma_desc_sw128(sp + k * 32), umma_desc_sw128_mn(sv + k * 2048), idesc_o, !first || k != 0);
        tc_commit(&pv_done[g & 1]);
        tc_commit(&v_empty[st]);
      };
      int g = 0, it = 0;
      bool prev_first = false;
      for (int w = blockIdx.x; w < nitems; w += gridDim.x, ++it) {
        int qt, h, b;
        item(w, qt, h, b);
        const int nblk = blocks_of(qt);
        mbar_wait(&q_full[it & 1], (it >> 1) & 1, 1);
        for (int j = 0; j < nblk; ++j, ++g) {
          // S of block g goes into TMEM buffer g & 1 as soon as the softmax of block g - 2 has read it: one block ahead
          // of the softmax warps, which therefore find their scores waiting
          const int st = g % AT_ST;
          mbar_wait(&k_full[st], (g / AT_ST) & 1, 1);
          if (g > 1) mbar_wait(&p_ready[g & 1], ((g - 2) >> 1) & 1, 1);
          tc_fence_after();
          const uint32_t sq = smem_u32(sQ + (it & 1) * AT_TILE), sk = smem_u32(sK + st * AT_KV);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            tc_mma_bf16(tS + (g & 1) * 64, umma_desc_sw128(sq + k * 32), umma_desc_sw128(sk + k * 32), idesc_s, k != 0);
          tc_commit(&s_full[g & 1]);
          tc_commit(&k_empty[st]);
          if (j == nblk - 1) tc_commit(&q_empty[it & 1]);
          // P V of the previous block, behind S in the tensor pipe.  When that block opened a new item it overwrites O:
          // the softmax threads read the old O (their epilogue) before they signalled p_ready for it.
          if (g > 0) issue_pv(g - 1, prev_first);
          prev_first = j == 0;
        }
      }
      if (g > 0) issue_pv(g - 1, prev_first);
    }
  } else {
    // ===== softmax + epilogue: thread = query row; a warp reaches the 32 TMEM lanes of its quadrant (warp index mod 4) =====
    const int quad = warp & 3;
    const int r = quad * 32 + lane;
    const uint32_t lane_off = (uint32_t)(quad * 32) << 16;
    constexpr float C = 0.125f * 1.4426950408889634f;  // 1/sqrt(64) and log2(e) in one factor
    int g = 0;
#ifdef TUTOR_EXPERIMENTS
#define AT_STAMP(i) if (tracing && g >= 2 && g < 6) stamp[g - 2][i] = clock64()
    const bool tracing = blockIdx.x == 0 && threadIdx.x == 64;
    long long stamp[4][8];
#else
#define AT_STAMP(i)
#endif
    for (int w = blockIdx.x; w < nitems; w += gridDim.x) {
      int qt, h, b;
      item(w, qt, h, b);
      const int nblk = blocks_of(qt), q0 = qt * 128, qrow = q0 + r;
      const int kp = PADDED ? npad[b] : 0;  // left padding: pad keys are invisible to real queries
      float m_run = -CUDART_INF_F, l_run = 0.f;
#ifdef TUTOR_EXPERIMENTS
      const long long t_item = clock64();
      long long t_first = 0;
#endif
      for (int j = 0; j < nblk; ++j, ++g) {
        const int key0 = j * AT_BN;
        const bool edge = (key0 + AT_BN - 1 >= T) || (CAUSAL && key0 + AT_BN - 1 > q0) || (PADDED && key0 < kp);
        AT_STAMP(0);
        mbar_wait(&s_full[g & 1], (g >> 1) & 1, 2);
        tc_fence_after();
        AT_STAMP(1);
#ifdef TUTOR_EXPERIMENTS
        if (j == 0) t_first = clock64();
#endif
        float s[64];
        tc_ld32x2(tS + (g & 1) * 64 + lane_off, tS + (g & 1) * 64 + lane_off + 32, s, s + 32);
        if (edge) {
#pragma unroll
          for (int i = 0; i < 64; ++i) {
            const int key = key0 + i;
            if (key >= T || (CAUSAL && key > qrow) || (PADDED && key < kp && qrow >= kp)) s[i] = -CUDART_INF_F;
          }
        }
        AT_STAMP(2);
        float mx0 = fmaxf(s[0], s[1]), mx1 = fmaxf(s[2], s[3]);
#pragma unroll
        for (int i = 4; i < 64; i += 2) { mx0 = fmaxf(mx0, s[i]); mx1 = fmaxf(mx1, s[i + 1]); }
        const float mb = fmaxf(mx0, mx1) * C;
        // ---- move the reference maximum (rarely): O and l are rescaled
        const bool move = mb > m_run + 8.0f;
        float alpha = 1.0f;
        if (move) {
          alpha = ex2(m_run - mb);  // 0 for the first block
          m_run = mb;
          l_run *= alpha;
        }
        const float mref = (m_run == -CUDART_INF_F) ? 0.f : m_run;  // a fully masked row stays at zero weight
        // ---- P = exp2(S C - m) and the row sum; independent of the previous block's P V, so it runs before the wait
        float l0 = 0.f, l1 = 0.f;
#pragma unroll
        for (int i = 0; i < 64; i += 2) {
          s[i] = ex2(fmaf(s[i], C, -mref));
          s[i + 1] = ex2(fmaf(s[i + 1], C, -mref));
          l0 += s[i];
          l1 += s[i + 1];
        }
        l_run += l0 + l1;
        AT_STAMP(3);
        // P buffer g & 1 was read by the P V of block g - 2 (one barrier per buffer: never more than one phase behind)
        if (g > 1) mbar_wait(&pv_done[g & 1], ((g - 2) >> 1) & 1, 3);
        if (j > 0 && __any_sync(0xffffffffu, move)) {
          mbar_wait(&pv_done[(g - 1) & 1], ((g - 1) >> 1) & 1, 3);  // O is rescaled between two P V
          tc_fence_after();
          {
#pragma unroll 1
            for (int c = 0; c < 8; ++c) tc_scale8(tO + lane_off + c * 8, alpha);
            tc_wait_st();
          }
        }
        AT_STAMP(4);
        uint8_t* prow = sP + (g & 1) * AT_TILE + r * 128;  // P buffer g & 1: block g - 2's P V finished a block ago
#pragma unroll
        for (int ch = 0; ch < 8; ++ch) {  // 8 keys = one 16-byte chunk of the swizzled A tile
          uint4 v;
          v.x = pack2(s[ch * 8 + 0], s[ch * 8 + 1]); v.y = pack2(s[ch * 8 + 2], s[ch * 8 + 3]);
          v.z = pack2(s[ch * 8 + 4], s[ch * 8 + 5]); v.w = pack2(s[ch * 8 + 6], s[ch * 8 + 7]);
          *reinterpret_cast<uint4*>(prow + ((ch ^ (r & 7)) << 4)) = v;
        }
        AT_STAMP(5);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy smem writes -> the MMA's async proxy
        tc_fence_before();
        AT_STAMP(6);
        mbar_arrive(&p_ready[g & 1]);
        AT_STAMP(7);
      }
#ifdef TUTOR_EXPERIMENTS
      if (tracing && w == (int)blockIdx.x)
        for (int i = 0; i < 4; ++i)
          printf("block %d: wait S %lld | TMEM load %lld | max + exp2 %lld | wait PV %lld | pack + store %lld | fence %lld | arrive %lld | to next %lld\n",
                 i + 2, stamp[i][1] - stamp[i][0], stamp[i][2] - stamp[i][1], stamp[i][3] - stamp[i][2], stamp[i][4] - stamp[i][3],
                 stamp[i][5] - stamp[i][4], stamp[i][6] - stamp[i][5], stamp[i][7] - stamp[i][6], i < 3 ? stamp[i + 1][0] - stamp[i][7] : 0LL);
#endif
#ifdef TUTOR_EXPERIMENTS
      const long long t_epi = clock64();
#endif
      // ---- epilogue: O / l -> bf16 [B*T, d]
      mbar_wait(&pv_done[(g - 1) & 1], ((g - 1) >> 1) & 1, 4);
      tc_fence_after();
      const float inv = 1.0f / l_run;
      bf16* dst = out + ((size_t)b * T + qrow) * (H * 64) + h * 64;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        float o[32];
        tc_ld32(tO + lane_off + c * 32, o);
        if (qrow < T) {
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            uint4 v;
            v.x = pack2(o[q * 8 + 0] * inv, o[q * 8 + 1] * inv); v.y = pack2(o[q * 8 + 2] * inv, o[q * 8 + 3] * inv);
            v.z = pack2(o[q * 8 + 4] * inv, o[q * 8 + 5] * inv); v.w = pack2(o[q * 8 + 6] * inv, o[q * 8 + 7] * inv);
            *reinterpret_cast<uint4*>(dst + c * 32 + q * 8) = v;
          }
        }
      }
      tc_fence_before();
#ifdef TUTOR_EXPERIMENTS
      if (tracing)
        printf("item %d (%d blocks): first S after %lld, blocks %lld (%lld each), epilogue %lld\n", w, nblk, t_first - t_item,
               t_epi - t_first, (t_epi - t_first) / nblk, clock64() - t_epi);
#endif
    }
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
  TRY(k_make_map_bf16(&tk, k, B * H * Smax, 64, AT_BN));
  TRY(k_make_map_bf16(&tv, v, B * H * Smax, 64, AT_BN));
  int dev = 0, sms = 0;
  CUDA_TRY(cudaGetDevice(&dev));
  CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int nitems = ceil_div(T, 128) * H * B;
  dim3 grid(nitems < 2 * sms ? nitems : 2 * sms);  // two resident CTAs per SM
#define AT_LAUNCH(C, P)                                                                                      \
  do {                                                                                                       \
    auto kern = attn_prefill_tc_kernel<C, P>;                                                                \
    TRY(tutor_set_max_smem(reinterpret_cast<const void*>(kern), AT_SMEM));                                   \
    cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared); \
    kern<<<grid, AT_THREADS, AT_SMEM, st>>>(tq, tk, tv, out, B, H, T, Smax, npad);                              \
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
