// Prefill / bidirectional attention on the 5th-generation tensor cores (bf16, head_dim 64).
//   causal:        GPT-2 prefill, HF/integrations/sdpa_attention.py:77 (is_causal iff T > 1)
//   bidirectional: BERT, HF/models/bert/modeling_bert.py:192-205 (through SDPA)
//
// A work item is (128 queries, head, sequence), processed flash-style over 64-key blocks:
//   S = Q K^T        tcgen05.mma 128x64x64 with Q as a tensor-memory A operand (the softmax threads store their query row
//                    there once per item, bf16 pairs packed into 32 columns) and the K tile K-major from TMA (SWIZZLE_128B);
//                    S in TMEM, two 64-column buffers: the MMA of block g + 1 runs while the softmax warps work on block g
//   P = softmax tile four warps, one query row per thread, the 64 scores of the row in registers; P is stored as packed
//                    bf16 over the first 32 columns of the score buffer it came from and is the A operand of
//   O += P V         tcgen05.mma 128x64x64, V tile straight from the cache layout [key][dim] as an MN-major B operand,
//                    O accumulated in TMEM (64 columns).
// Only K and V tiles live in shared memory (four stages of 16 KB).  The running maximum is only moved when a block's
// maximum exceeds it by more than 8 (in log2 units): P then stays below 2^8, and O / l is exact for any reference maximum.
// Moving it means rescaling O in TMEM (load, scale, store), which happens for the first few blocks of a row only.
// Persistent: two CTAs per SM (256 TMEM columns each) walk the items, longest first; barriers, TMEM and the K/V ring carry
// over from item to item, and the next item's queries are stored an item ahead.
// Warp roles: 0 = TMA producer, 1 = issues S = Q K^T, 6 = issues O += P V, 2..5 = softmax / epilogue.
//
// Measured (BERT 32 x 512 x 12 heads, 25.8 GFLOP; mma.sync kernel 97.7 us): 128-key blocks with Q, K, V, P in shared memory,
// one CTA per item 71.7 us; persistent 66.0 us; two threads per row (eight softmax warps) 92.8 us; 64-key blocks, three CTAs
// per SM 68.9 us; S / P / Q double-buffered 64.3 us; two softmax warpgroups on alternating blocks with separate O
// accumulators 77.9 us; Q and P in TMEM 64.3 us; + one barrier per K/V stage, two commits per block, descriptors built once
// per stage 59.4 us; + separate issuing threads for S and P V 58.6 us (93 / 104 us on the causal GPT-2 shapes 64 x 512 x 12
// and 16 x 1024 x 20, against 135 / 180 us).  Timing experiments on the last version: exp2 replaced by an FMA 56.7 us, score
// loads from TMEM removed as well 60.5 us, six K/V stages 60.9 us - neither the softmax arithmetic nor TMEM reads nor the
// tile supply bound it; what remains is the per-block hand-off chain (commit -> barrier -> TMEM load ... TMEM store ->
// barrier -> MMA) of two co-resident CTAs and the 5-or-6 items per CTA quantisation (profiles/r2c_attn_prefill_tc.txt).
#include <math_constants.h>

#include "kernels.h"
#include "tc_ptx.cuh"

namespace {
using namespace tcptx;

constexpr int AT_THREADS = 224;               // TMA warp, S-MMA warp, 4 softmax warps, PV-MMA warp
constexpr int AT_TILE = 128 * 128;             // bytes of a 128-row x 64-column bf16 tile (Q, P)
constexpr int AT_KV = 64 * 128;                // bytes of a 64-key K or V tile
constexpr int AT_BN = 64;                      // keys per block
constexpr int AT_ST = 4;                       // K / V stages
constexpr int AT_SMEM = 2 * AT_ST * AT_KV + 256;  // K[AT_ST], V[AT_ST], barriers
constexpr uint32_t AT_TMEM_COLS = 256;         // S[2]: columns 0..127, O: 128..191, Q[2] (packed bf16): 192..255

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
#ifdef TUTOR_EXPERIMENTS_NOMUFU
  return fmaf(x, 1e-9f, 1.0f);  // timing experiment only: no MUFU
#else
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
#endif
}
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  __nv_bfloat162 p = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&p);
}

// tcgen05.mma with the A operand in tensor memory (bf16 pairs packed into 32-bit columns, lane = row)
__device__ __forceinline__ void tc_mma_bf16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d_tmem),
      "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// 32 lanes x 32 columns of 32-bit words into tensor memory
__device__ __forceinline__ void tc_st32(uint32_t taddr, const uint32_t* v) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
      "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]),
      "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]),
      "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]),
      "r"(v[31])
      : "memory");
}

template <bool CAUSAL, bool PADDED>
__global__ void __launch_bounds__(AT_THREADS, 2)
    attn_prefill_tc_kernel(const bf16* __restrict__ q, const __grid_constant__ CUtensorMap tmK,
                           const __grid_constant__ CUtensorMap tmV, bf16* __restrict__ out, int B, int H, int T, int Smax,
                           const int* __restrict__ npad) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sK = smem;                      // [AT_ST]
  uint8_t* sV = smem + AT_ST * AT_KV;      // [AT_ST]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 2 * AT_ST * AT_KV);
  uint64_t* kv_full = bars;                // [AT_ST] K and V tile of a block have landed
  uint64_t* kv_empty = bars + AT_ST;       // [AT_ST] ... and O += P V of that block has completed (it read V last)
  uint64_t* s_full = bars + 2 * AT_ST;     // [2]
  uint64_t* p_ready = s_full + 2;          // [2], one arrival per softmax warp
  uint64_t* q_full = s_full + 4;           // [2], one arrival per softmax warp
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(s_full + 6);

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
    if (smem_u32(smem) & 1023) __trap();  // the swizzled tiles need 1 KB alignment
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmK) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmV) : "memory");
    for (int i = 0; i < 2 * AT_ST + 2; ++i) mbar_init(&bars[i], 1);
    for (int i = 0; i < 2; ++i) { mbar_init(&p_ready[i], 4); mbar_init(&q_full[i], 4); }  // one arrival per softmax warp
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
  // S buffer i at tS + 64 i (P of the block overwrites its first 32 columns), Q buffer i at tQ + 32 i
  const uint32_t tS = tmem_base, tO = tmem_base + 128, tQ = tmem_base + 192;

  if (warp == 0) {
    // ===== TMA producer: K and V tiles =====
    if (lane == 0) {
      int g = 0;
      for (int w = blockIdx.x; w < nitems; w += gridDim.x) {
        int qt, h, b;
        item(w, qt, h, b);
        const int kv_row0 = (b * H + h) * Smax, nblk = blocks_of(qt);
        for (int j = 0; j < nblk; ++j, ++g) {
          const int st = g % AT_ST;
          const uint32_t use = (uint32_t)(g / AT_ST) & 1;
          mbar_wait(&kv_empty[st], use ^ 1, 0);
          mbar_expect_tx(&kv_full[st], 2 * AT_KV);
          tma_load_2d(sK + st * AT_KV, &tmK, &kv_full[st], 0, kv_row0 + j * AT_BN);
          tma_load_2d(sV + st * AT_KV, &tmV, &kv_full[st], 0, kv_row0 + j * AT_BN);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread).  Both A operands live in tensor memory: Q (stored there once per item by the softmax
    // threads) and P (stored over the S buffer they have just read).  Only K and V are read from shared memory - with A in
    // shared memory too the kernel was bound by shared-memory bandwidth (Q and P re-read for every 64 keys: 32 of the 48 KB
    // a block moved through the MMA's operand path, plus 16 KB of P stores). =====
    if (lane == 0) {
      // The issuing thread is the pacemaker of the whole CTA: with three barrier tests (~130 cycles each even when the
      // barrier is long complete), four commits (~65 each) and descriptors rebuilt for every MMA it spent ~1600 cycles per
      // 64-key block - as long as the softmax warps need for the block.  Hence one barrier per K/V stage, two commits per
      // block, and descriptors made once per stage.
      constexpr uint32_t idesc_s = umma_idesc_bf16(128, 64);
      uint64_t kdesc[AT_ST];
#pragma unroll
      for (int st = 0; st < AT_ST; ++st) kdesc[st] = umma_desc_sw128(smem_u32(sK + st * AT_KV));
      auto stage_desc = [&](const uint64_t* d, int st) {  // register array indexed by a run-time stage without local memory
        uint64_t r = d[0];
#pragma unroll
        for (int i = 1; i < AT_ST; ++i) r = st == i ? d[i] : r;
        return r;
      };
      // This thread issues S = Q K^T only; O += P V is issued by warp 6, so that neither waits behind the other's barriers.
      // Without a common issue order the reuse of a score buffer needs its own edge: S of block g overwrites the buffer whose
      // P the P V of block g - 2 reads, so it waits for that P V to complete (kv_empty of its stage).
      int g = 0, it = 0;
      for (int w = blockIdx.x; w < nitems; w += gridDim.x, ++it) {
        int qt, h, b;
        item(w, qt, h, b);
        const int nblk = blocks_of(qt);
        mbar_wait(&q_full[it & 1], (it >> 1) & 1, 1);
        for (int j = 0; j < nblk; ++j, ++g) {
          const int st = g % AT_ST;
          mbar_wait(&kv_full[st], (g / AT_ST) & 1, 1);
          if (g >= 2) mbar_wait(&kv_empty[(g - 2) % AT_ST], ((g - 2) / AT_ST) & 1, 1);
          tc_fence_after();
          const uint32_t tq = tQ + (it & 1) * 32;
          const uint64_t kd = stage_desc(kdesc, st);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            tc_mma_bf16_ts(tS + (g & 1) * 64, tq + k * 8, kd + (uint64_t)(k * (32 >> 4)), idesc_s, k != 0);
          tc_commit(&s_full[g & 1]);
        }
      }
    }
  } else if (warp == 6) {
    // ===== second MMA issuer: O += P V =====
    if (lane == 0) {
      constexpr uint32_t idesc_o = umma_idesc_bf16(128, 64) | (1u << 16);  // B (= V) is MN-major
      uint64_t vdesc[AT_ST];
#pragma unroll
      for (int st = 0; st < AT_ST; ++st) vdesc[st] = umma_desc_sw128_mn(smem_u32(sV + st * AT_KV));
      int g = 0;
      for (int w = blockIdx.x; w < nitems; w += gridDim.x) {
        int qt, h, b;
        item(w, qt, h, b);
        const int nblk = blocks_of(qt);
        for (int j = 0; j < nblk; ++j, ++g) {
          const int st = g % AT_ST;
          mbar_wait(&p_ready[g & 1], (g >> 1) & 1, 1);  // P of block g is in tensor memory (its K/V stage landed long ago)
          tc_fence_after();
          const uint32_t tp = tS + (g & 1) * 64;
          uint64_t vd = vdesc[0];
#pragma unroll
          for (int i = 1; i < AT_ST; ++i) vd = st == i ? vdesc[i] : vd;
#pragma unroll
          for (int k = 0; k < 4; ++k)  // 16 keys per MMA: 8 columns of packed P; two 8-key groups (2 x 1024 bytes) of V
            tc_mma_bf16_ts(tO, tp + k * 8, vd + (uint64_t)(k * (2048 >> 4)), idesc_o, j != 0 || k != 0);
          tc_commit(&kv_empty[st]);  // frees the stage, and tells the softmax warps that O is complete up to this block
        }
      }
    }
  } else {
    // ===== softmax + epilogue: thread = query row; a warp reaches the 32 TMEM lanes of its quadrant (warp index mod 4) =====
    const int quad = warp & 3;
    const int r = quad * 32 + lane;
    const uint32_t lane_off = (uint32_t)(quad * 32) << 16;
    constexpr float C = 0.125f * 1.4426950408889634f;  // 1/sqrt(64) and log2(e) in one factor
    // this thread's query row of item w -> Q buffer `buf` (64 bf16 = 32 packed columns); rows past the tensor are clamped
    auto store_q = [&](int w, int buf) {
      int qt, h, b;
      item(w, qt, h, b);
      const size_t row = min((size_t)b * T + qt * 128 + r, (size_t)B * T - 1);
      const uint4* src = reinterpret_cast<const uint4*>(q + row * (H * 64) + h * 64);
      uint32_t v[32];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const uint4 x = src[i];
        v[4 * i] = x.x; v[4 * i + 1] = x.y; v[4 * i + 2] = x.z; v[4 * i + 3] = x.w;
      }
      tc_st32(tQ + buf * 32 + lane_off, v);
      tc_wait_st();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&q_full[buf]);
    };
    int g = 0, it = 0;
    store_q(blockIdx.x, 0);
    for (int w = blockIdx.x; w < nitems; w += gridDim.x, ++it) {
      int qt, h, b;
      item(w, qt, h, b);
      // the next item's queries, an item ahead: that buffer was last read by the S MMAs of the previous item, all of
      // which this thread has seen completed (it consumed their results)
      if (w + (int)gridDim.x < nitems) store_q(w + (int)gridDim.x, (it + 1) & 1);
      const int nblk = blocks_of(qt), q0 = qt * 128, qrow = q0 + r;
      const int kp = PADDED ? npad[b] : 0;  // left padding: pad keys are invisible to real queries
      float m_run = -CUDART_INF_F, l_run = 0.f;
      for (int j = 0; j < nblk; ++j, ++g) {
        const int key0 = j * AT_BN;
        const bool edge = (key0 + AT_BN - 1 >= T) || (CAUSAL && key0 + AT_BN - 1 > q0) || (PADDED && key0 < kp);
        const uint32_t tSb = tS + (g & 1) * 64 + lane_off;
        mbar_wait(&s_full[g & 1], (g >> 1) & 1, 2);
        tc_fence_after();
        float s[64];
#ifdef TUTOR_EXPERIMENTS_NOLDTM
#pragma unroll
        for (int i = 0; i < 64; ++i) s[i] = (float)(i + r) * 0.01f;  // timing experiment only: the scores are not read
#else
        tc_ld32x2(tSb, tSb + 32, s, s + 32);
#endif
        if (edge) {
#pragma unroll
          for (int i = 0; i < 64; ++i) {
            const int key = key0 + i;
            if (key >= T || (CAUSAL && key > qrow) || (PADDED && key < kp && qrow >= kp)) s[i] = -CUDART_INF_F;
          }
        }
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
        // ---- P = exp2(S C - m) as packed bf16 pairs, and the row sum
        float l0 = 0.f, l1 = 0.f;
        uint32_t pk[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const float p0 = ex2(fmaf(s[2 * i], C, -mref)), p1 = ex2(fmaf(s[2 * i + 1], C, -mref));
          l0 += p0;
          l1 += p1;
          pk[i] = pack2(p0, p1);
        }
        l_run += l0 + l1;
        if (j > 0 && __any_sync(0xffffffffu, move)) {
          mbar_wait(&kv_empty[(g - 1) % AT_ST], ((g - 1) / AT_ST) & 1, 3);  // O is rescaled between two P V
          tc_fence_after();
#pragma unroll 1
          for (int c = 0; c < 8; ++c) tc_scale8(tO + lane_off + c * 8, alpha);
        }
        tc_st32(tSb, pk);  // over the scores this thread has just read: the A operand of this block's P V
        tc_wait_st();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&p_ready[g & 1]);  // one arrival per warp (128 arrivals on one barrier serialise)
      }
      // ---- epilogue: O / l -> bf16 [B*T, d]
      mbar_wait(&kv_empty[(g - 1) % AT_ST], ((g - 1) / AT_ST) & 1, 4);  // the item's last P V has completed
      tc_fence_after();
      const float inv = 1.0f / l_run;
      bf16* dst = out + ((size_t)b * T + qrow) * (H * 64) + h * 64;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        float o[32];
        tc_ld32(tO + lane_off + c * 32, o);
        if (qrow < T) {
#pragma unroll
          for (int qq = 0; qq < 4; ++qq) {
            uint4 v;
            v.x = pack2(o[qq * 8 + 0] * inv, o[qq * 8 + 1] * inv); v.y = pack2(o[qq * 8 + 2] * inv, o[qq * 8 + 3] * inv);
            v.z = pack2(o[qq * 8 + 4] * inv, o[qq * 8 + 5] * inv); v.w = pack2(o[qq * 8 + 6] * inv, o[qq * 8 + 7] * inv);
            *reinterpret_cast<uint4*>(dst + c * 32 + qq * 8) = v;
          }
        }
      }
      tc_fence_before();
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
  CUtensorMap tk, tv;
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
    kern<<<grid, AT_THREADS, AT_SMEM, st>>>(q, tk, tv, out, B, H, T, Smax, npad);                              \
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
