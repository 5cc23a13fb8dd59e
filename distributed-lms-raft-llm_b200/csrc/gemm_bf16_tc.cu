// bf16 contraction  D[i,j] = sum_k A[i,k] * B[j,k]  on the 5th-gen tensor cores (sm_100a):
// TMA (cp.async.bulk.tensor, SWIZZLE_128B) stages K-major tiles in shared memory, one elected thread
// issues tcgen05.mma (UMMA 128 x BN x 16, kind::f16, bf16 in / fp32 accumulate) into TMEM, four
// epilogue warps read the accumulator back with tcgen05.ld and apply the shared fused epilogue.
//
//   prefill / BERT  (M = tokens, large):  A = activations [M,K], B = weights [N,K]   -> EPI_ROWMAJOR
//   decode (M = batch <= 256, "swap-AB"): A = weights [N,K],     B = activations [M,K] -> EPI_SWAPPED
//   LM head at decode:                    A = wte [V,K],         B = ln_f(x) [batch,K] -> EPI_ARGMAX
//                                         (penalty + per-tile argmax; no logits tensor is written)
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer,
// warps 2..5 = epilogue (TMEM lane quadrant = warp_id % 4).  Every mbarrier wait is bounded
// (clock64 deadline -> __trap) so a protocol bug aborts the launch instead of hanging the GPU.
#include <cuda.h>
#include <math_constants.h>

#include <cstdlib>
#include <mutex>

#include "kernels.h"
#include "tc_ptx.cuh"

namespace {

constexpr int NTHREADS = 192;        // decode flavours: TMA warp, MMA warp, 4 epilogue warps
constexpr int NTHREADS_WIDE = 320;   // one-tile row-major flavour: 8 epilogue warps, two per TMEM lane quadrant
constexpr int PERSIST_EPI_WARPS = 16;  // persistent flavour: four warps per TMEM lane quadrant (epilogue issue-bound)
constexpr int NTHREADS_PERSIST = 64 + 32 * PERSIST_EPI_WARPS;
enum { EPI_ROWMAJOR = 0, EPI_SWAPPED = 1, EPI_ARGMAX = 2, EPI_PARTIAL = 3 };

// Phase timestamps of CTA (0,0,0) for tools/microbench.cu (compiled in only with -DTUTOR_TRACE)
#ifdef TUTOR_TRACE
__device__ int g_epi_skip;  // tools/microbench_gemm.cu: 1 = persistent epilogue only hands the accumulator back
__device__ long long g_gemm_trace[8];
#define TRACE(i)                                                              \
  do {                                                                        \
    if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) g_gemm_trace[i] = clock64(); \
  } while (0)
#else
#define TRACE(i)
#endif

struct ArgmaxArgs {
  const uint32_t* seen;  // [B][seen_words]
  int seen_words;
  float penalty;
  int V;                 // real vocabulary size (rows >= V are padding)
  float* part_val;       // [ntiles][B]
  int* part_idx;
  int B;
};

using namespace tcptx;


// 256-bit global accesses (LDG/STG.E.256, new on sm_100): in the row-major epilogue every lane owns a different
// output row, so each warp-level access touches 32 separate lines; moving a whole 32-byte sector per lane halves
// both the instruction count and the sector operations of the 128-bit form (which used half of every sector).
__device__ __forceinline__ void ldg256(const float* p, float* v) {
  asm volatile("ld.global.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7])
               : "l"(p));
}
__device__ __forceinline__ void stg256(float* p, const float* v) {
  asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]),
               "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7])
               : "memory");
}
__device__ __forceinline__ void stg256_bf16(bf16* p, const float* v) {  // 16 floats -> 16 bf16 = 32 bytes
  uint32_t w[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
    w[i] = *reinterpret_cast<uint32_t*>(&h);
  }
  asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]),
               "r"(w[4]), "r"(w[5]), "r"(w[6]), "r"(w[7])
               : "memory");
}

// vectorised row run: 32 consecutive columns n0.. of row m (EPI_ROWMAJOR)
__device__ __forceinline__ void store_run32(const GemmEpilogue& e, int m, int n0, int N, float* v) {
  if (n0 + 32 > N || (e.ldo & 15)) {  // ragged edge / unaligned rows: scalar path
    for (int i = 0; i < 32 && n0 + i < N; ++i) epilogue_store(e, m, n0 + i, v[i]);
    return;
  }
  if (e.bias) {
#pragma unroll
    for (int i = 0; i < 32; i += 4) {
      float4 b = *reinterpret_cast<const float4*>(e.bias + n0 + i);
      v[i] += b.x; v[i + 1] += b.y; v[i + 2] += b.z; v[i + 3] += b.w;
    }
  }
  if (e.act != ACT_NONE) {
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = apply_act_fast(v[i], e.act);  // bf16 mode: see common.cuh
  }
  float* of = nullptr;
  bf16* ob = nullptr;
  if (!e.qkv_mode || n0 < e.d) {
    size_t o = (size_t)m * e.ldo + n0;
    if (!e.qkv_mode && e.residual) {
#pragma unroll
      for (int i = 0; i < 32; i += 8) {
        float r[8];
        ldg256(e.residual + o + i, r);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[i + j] += r[j];
      }
    }
    if (!e.qkv_mode && e.residual_bf16) {
#pragma unroll
      for (int i = 0; i < 32; i += 8) {
        float r[8];
        unpack_bf16x8(*reinterpret_cast<const uint4*>(e.residual_bf16 + o + i), r);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[i + j] += r[j];
      }
    }
    if (e.out_f32) of = e.out_f32 + o;
    if (e.out_bf16) ob = e.out_bf16 + o;
  } else {  // 32 columns sit inside one head's 64-wide row of the K or V cache
    int which = (n0 >= 2 * e.d);
    int c = n0 - (which ? 2 * e.d : e.d);
    int h = c >> 6, i0 = c & 63;
    int b = m / e.T, t = m - b * e.T;
    int s = e.pos_dev ? (*e.pos_dev - 1) : (e.pos0 + t);
    size_t o = (((size_t)b * e.H + h) * e.Smax + s) * 64 + i0;
    void* base = which ? e.vcache : e.kcache;
    if (e.kv_dtype == DT_F32) of = reinterpret_cast<float*>(base) + o;
    else ob = reinterpret_cast<bf16*>(base) + o;
  }
  if (of) {
#pragma unroll
    for (int i = 0; i < 32; i += 8) stg256(of + i, v + i);
  }
  if (ob) {
#pragma unroll
    for (int i = 0; i < 32; i += 16) stg256_bf16(ob + i, v + i);
  }
}

template <int BN, int EPI>
__global__ void __launch_bounds__(EPI == EPI_ROWMAJOR ? NTHREADS_WIDE : NTHREADS) gemm_bf16_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA,
                                                                      const __grid_constant__ CUtensorMap tmB,
                                                                      const __grid_constant__ CUtensorMap tmC,
                                                                      int rowsA, int rowsB, int K, GemmEpilogue epi,
                                                                      ArgmaxArgs am) {
  // LM head at batch 64: three stages (72 KB) so that three CTAs share an SM - 444 slots hold all 393 vocabulary
  // tiles of GPT-2 in one wave instead of 296 + a 97-CTA tail
  using L = SmemLayout<BN, (EPI == EPI_ARGMAX && BN == 64) ? 3 : 0>;
  constexpr int TMEM_COLS = (BN < 32) ? 32 : BN;  // power of two >= 32 (BN in {16,32,64,128,256})
  extern __shared__ uint8_t smem_raw[];
  uint8_t* tiles = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(tiles + L::TILE_BYTES);
  uint64_t* empty_bar = full_bar + L::STAGES;
  uint64_t* tmem_full_bar = empty_bar + L::STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int a_row0 = blockIdx.x * BM, b_row0 = blockIdx.y * BN;
  // split-K: blockIdx.z owns k-blocks [kb0, kb0 + nk); gridDim.z == 1 for the unsplit flavours
  const int nk = (K / BK) / (int)gridDim.z;
  const int kb0 = (int)blockIdx.z * nk;
  if (threadIdx.x == 0) TRACE(0);

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
    for (int s = 0; s < L::STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
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
  if (threadIdx.x == 0) TRACE(1);

  pdl_launch_dependents();
  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      // PDL: in the swapped (decode) flavours operand A is a weight matrix nobody writes, so its first tiles are
      // requested before waiting for the predecessor kernel; operand B (activations) only after the wait.
      const bool a_is_static = (EPI != EPI_ROWMAJOR);
      const int pre = a_is_static ? (nk < L::STAGES ? nk : L::STAGES) : 0;
      for (int kb = 0; kb < pre; ++kb) {
        mbar_expect_tx(&full_bar[kb], L::STAGE_BYTES);
        tma_load_2d(tiles + kb * L::STAGE_BYTES, &tmA, &full_bar[kb], (kb0 + kb) * BK, a_row0);
      }
      pdl_wait();
      for (int kb = 0; kb < pre; ++kb)
        tma_load_2d(tiles + kb * L::STAGE_BYTES + L::A_BYTES, &tmB, &full_bar[kb], (kb0 + kb) * BK, b_row0);
      TRACE(2);
      for (int kb = pre; kb < nk; ++kb) {
        const int s = kb % L::STAGES;
        const uint32_t ph = (kb / L::STAGES) & 1;
        mbar_wait(&empty_bar[s], ph ^ 1, 0);
        mbar_expect_tx(&full_bar[s], L::STAGE_BYTES);
        uint8_t* sa = tiles + s * L::STAGE_BYTES;
        tma_load_2d(sa, &tmA, &full_bar[s], (kb0 + kb) * BK, a_row0);
        tma_load_2d(sa + L::A_BYTES, &tmB, &full_bar[s], (kb0 + kb) * BK, b_row0);
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread) =====
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(BM, BN);
      for (int kb = 0; kb < nk; ++kb) {
        const int s = kb % L::STAGES;
        const uint32_t ph = (kb / L::STAGES) & 1;
        mbar_wait(&full_bar[s], ph, 1);
        tc_fence_after();
        if (kb == 0) TRACE(3);
        const uint32_t sa = smem_u32(tiles + s * L::STAGE_BYTES);
        // one descriptor pair per stage; advancing 16 bf16 = 32 bytes along K inside the 128-byte swizzle atom moves the
        // 16-byte-granular address field by 2
        const uint64_t da = umma_desc_sw128(sa), db = umma_desc_sw128(sa + L::A_BYTES);
#pragma unroll
        for (int k = 0; k < BK / UMMA_K; ++k)
          tc_mma_bf16(tmem_base, da + (uint64_t)(k * (UMMA_K * 2 >> 4)), db + (uint64_t)(k * (UMMA_K * 2 >> 4)), idesc,
                      (kb | k) != 0);
        tc_commit(&empty_bar[s]);  // frees the smem stage once these MMAs have read it
      }
      tc_commit(tmem_full_bar);    // accumulator complete
      TRACE(4);
    }
  } else {
    // ===== epilogue: 4 warps, each owns the 32 TMEM lanes of its quadrant =====
    const int quad = warp & 3;
    const int row = a_row0 + quad * 32 + lane;  // index along A's rows
    mbar_wait(tmem_full_bar, 0, 2);
    tc_fence_after();
    if (threadIdx.x == 64) TRACE(5);
    pdl_wait();  // outputs may overwrite buffers the predecessor kernel still reads (already satisfied by now)
    const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16);
    if (EPI == EPI_ARGMAX) {
      // per-tile argmax over the 128 vocabulary rows for every batch column: penalised scores are transposed
      // through the (now idle) pipeline smem as [column][128 rows]; then each warp reduces whole columns -
      // 4 conflict-free loads + one 5-round shuffle arg-max per column instead of one per (row-quadrant, column)
      float* stage = reinterpret_cast<float*>(tiles);
      const int wslot = warp - 2;
      for (int c0 = 0; c0 < BN; c0 += 32) {
        float v[32];
        tc_ld32(taddr + c0, v);
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          if (c0 + i >= BN) break;  // BN = 16: only half of the 32-column load is real
          const int b = b_row0 + c0 + i;
          float val = v[i];
          if (b < am.B && row < am.V && am.penalty != 1.0f) {
            // 32 lanes = 32 consecutive vocabulary rows = one word of the seen-bitmap
            uint32_t w = am.seen[(size_t)b * am.seen_words + (row >> 5)];
            if ((w >> (row & 31)) & 1u) val = (val < 0.f) ? val * am.penalty : val / am.penalty;
          }
          if (row >= am.V) val = -CUDART_INF_F;
          stage[(c0 + i) * BM + quad * 32 + lane] = val;
        }
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");  // the four epilogue warps only
      for (int c = wslot; c < BN; c += 4) {
        if (b_row0 + c >= am.B) break;
        float val = stage[c * BM + lane];
        int idx = lane;
#pragma unroll
        for (int r = 1; r < 4; ++r) {
          const float ov = stage[c * BM + r * 32 + lane];
          if (ov > val) { val = ov; idx = r * 32 + lane; }  // ascending rows: strict > keeps the lowest index
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          const float ov = __shfl_xor_sync(0xffffffffu, val, o);
          const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
          if (ov > val || (ov == val && oi < idx)) { val = ov; idx = oi; }
        }
        if (lane == 0) {
          am.part_val[(size_t)blockIdx.x * am.B + b_row0 + c] = val;
          am.part_idx[(size_t)blockIdx.x * am.B + b_row0 + c] = a_row0 + idx;
        }
      }
    } else {
      // row-major flavour: warps 2-5 take the first half of the columns, warps 6-9 the second half of the same
      // TMEM lanes (the epilogue, not the MMA, bounds a 128xBN tile: ~25 instructions per output element)
      const int c_begin = (EPI == EPI_ROWMAJOR) ? ((warp - 2) >> 2) * (BN / 2) : 0;
      const int c_end = (EPI == EPI_ROWMAJOR) ? c_begin + BN / 2 : BN;
      for (int c0 = c_begin; c0 < c_end; c0 += 32) {
        float v[32];
        tc_ld32(taddr + c0, v);  // warp-collective: executed by all lanes even for out-of-range rows
        if (EPI == EPI_ROWMAJOR) {
          // (a per-warp smem transpose for row-coalesced stores was tried and measured 2.4x slower than these
          // per-thread 16-byte vector stores: 32 scalar row passes cost more than the uncoalesced sectors save)
          if (row < rowsA && b_row0 + c0 < rowsB) store_run32(epi, row, b_row0 + c0, rowsB, v);
        } else if (EPI == EPI_PARTIAL) {
          // raw fp32 partial sums: staged [batch][128 n] in the (now idle) pipeline smem, then ONE bulk tensor
          // store per CTA - 64 scattered STG per thread cost ~4000 cycles here, the TMA store a few hundred
          float* stage = reinterpret_cast<float*>(tiles);
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (c0 + i < BN) stage[(c0 + i) * BM + quad * 32 + lane] = v[i];
        } else {  // EPI_SWAPPED: lane = output column n, TMEM column = batch row m
          if (row < rowsA) {
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              const int m = b_row0 + c0 + i;
              if (c0 + i < BN && m < rowsB) epilogue_store(epi, m, row, v[i]);
            }
          }
        }
      }
      if (EPI == EPI_PARTIAL) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy smem writes -> async proxy
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (threadIdx.x == 64) tma_store_3d(&tmC, tiles, a_row0, b_row0, (int)blockIdx.z);  // OOB rows/cols clipped
      }
    }
  }

  if (threadIdx.x == 64) TRACE(6);
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 0) TRACE(7);
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
  }
}

// ------------------------------------------------------------------------- persistent row-major GEMM
// Prefill / BERT flavour (M = tokens).  One CTA per SM walks a static list of 128 x BN output tiles; the TMA
// producer and the MMA issuer run ahead into the next tile while the 8 epilogue warps are still draining the
// previous one: the accumulator is double-buffered in TMEM (2 x BN columns) and handed over through
// tmem_full / tmem_empty mbarriers.  The one-tile-per-CTA kernel above spends most of a tile's life in its epilogue
// (~25 instructions per output element on 4-8 warps) with the tensor pipe idle; here the two overlap.
// Tiles are ordered m-fastest, so the CTAs running at any moment share one weight tile in L2.
template <int BN> struct PersistLayout {
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int B_BYTES = BN * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = (192 * 1024) / STAGE_BYTES;  // 4 stages at BN=256, 6 at BN=128
  static constexpr int TILE_BYTES = STAGES * STAGE_BYTES;
  static constexpr int TOTAL = TILE_BYTES + 256 + 1024;
};

template <int BN>
__global__ void __launch_bounds__(NTHREADS_PERSIST) gemm_bf16_persistent_kernel(const __grid_constant__ CUtensorMap tmA,
                                                                            const __grid_constant__ CUtensorMap tmB,
                                                                            int M, int N, int K, GemmEpilogue epi) {
  using L = PersistLayout<BN>;
  constexpr int TMEM_COLS = 2 * BN;  // two accumulators; 256 or 512 columns (power of two)
  extern __shared__ uint8_t smem_raw[];
  uint8_t* tiles = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(tiles + L::TILE_BYTES);
  uint64_t* empty_bar = full_bar + L::STAGES;
  uint64_t* tmem_full_bar = empty_bar + L::STAGES;  // [2]
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;     // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nk = K / BK;
  const int m_tiles = (M + BM - 1) / BM, n_tiles = (N + BN - 1) / BN;
  const int ntiles = m_tiles * n_tiles;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
    for (int s = 0; s < L::STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(&tmem_full_bar[a], 1); mbar_init(&tmem_empty_bar[a], PERSIST_EPI_WARPS); }
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

  if (warp == 0) {
    // ===== TMA producer: runs through all of this CTA's tiles back to back =====
    if (lane == 0) {
      uint32_t it = 0;
      for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
        const int a_row0 = (t % m_tiles) * BM, b_row0 = (t / m_tiles) * BN;
        for (int kb = 0; kb < nk; ++kb, ++it) {
          const int s = it % L::STAGES;
          mbar_wait(&empty_bar[s], ((it / L::STAGES) & 1) ^ 1, 0);
          mbar_expect_tx(&full_bar[s], L::STAGE_BYTES);
          uint8_t* sa = tiles + s * L::STAGE_BYTES;
          tma_load_2d(sa, &tmA, &full_bar[s], kb * BK, a_row0);
          tma_load_2d(sa + L::A_BYTES, &tmB, &full_bar[s], kb * BK, b_row0);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: accumulator (ti & 1) once the epilogue has drained its previous contents =====
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(BM, BN);
      uint32_t it = 0, ti = 0;
      for (int t = blockIdx.x; t < ntiles; t += gridDim.x, ++ti) {
        const uint32_t acc = ti & 1;
        mbar_wait(&tmem_empty_bar[acc], ((ti >> 1) & 1) ^ 1, 3);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = 0; kb < nk; ++kb, ++it) {
          const int s = it % L::STAGES;
          mbar_wait(&full_bar[s], (it / L::STAGES) & 1, 1);
          tc_fence_after();
          const uint32_t sa = smem_u32(tiles + s * L::STAGE_BYTES);
          const uint64_t da = umma_desc_sw128(sa), db = umma_desc_sw128(sa + L::A_BYTES);  // k-step: address field + 2
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k)
            tc_mma_bf16(d_tmem, da + (uint64_t)(k * (UMMA_K * 2 >> 4)), db + (uint64_t)(k * (UMMA_K * 2 >> 4)), idesc,
                        (kb | k) != 0);
          tc_commit(&empty_bar[s]);
        }
        tc_commit(&tmem_full_bar[acc]);
      }
    }
  } else {
    // ===== epilogue: warp w reads TMEM lanes of quadrant w % 4 and column slice (w - 2) / 4 of the tile =====
    constexpr int SLICE = BN / (PERSIST_EPI_WARPS / 4);
    const int quad = warp & 3, half = (warp - 2) >> 2;
    uint32_t ti = 0;
    for (int t = blockIdx.x; t < ntiles; t += gridDim.x, ++ti) {
      const int a_row0 = (t % m_tiles) * BM, b_row0 = (t / m_tiles) * BN;
      const uint32_t acc = ti & 1;
      mbar_wait(&tmem_full_bar[acc], (ti >> 1) & 1, 2);
      tc_fence_after();
      const int row = a_row0 + quad * 32 + lane;
      const uint32_t taddr = tmem_base + acc * BN + ((uint32_t)(quad * 32) << 16);
      const int c_begin = half * SLICE, c_end = c_begin + SLICE;
#ifdef TUTOR_TRACE
      if (g_epi_skip == 1) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tmem_empty_bar[acc]);
        continue;
      }
#endif
      for (int c0 = c_begin; c0 < c_end; c0 += 32) {
        float v[32];
#ifdef TUTOR_TRACE
        if (g_epi_skip == 3) {  // no TMEM read: the stores see zeros
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] = 0.f;
        } else
#endif
        tc_ld32(taddr + c0, v);
        if (c0 + 32 >= c_end) {  // last read of this accumulator by this warp: hand it back before the stores
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tmem_empty_bar[acc]);
        }
#ifdef TUTOR_TRACE
        if (g_epi_skip == 4) {  // TMEM read only: no epilogue math, no global traffic
          float acc = 0.f;
#pragma unroll
          for (int i = 0; i < 32; ++i) acc += v[i];
          if (acc == 123.456f) epi.out_f32[0] = acc;
          continue;
        }
#endif
        if (row < M && b_row0 + c0 < N) store_run32(epi, row, b_row0 + c0, N, v);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
  }
}

// ---------------------------------------------------- persistent row-major GEMM, TMA-staged epilogue
// Same mainloop as gemm_bf16_persistent_kernel.  Measured on B200 (tools/microbench_gemm.cu, BERT-base layer shapes at
// 16384 tokens): the mainloop alone runs at 1.3-1.7 PFLOP/s, while the epilogue above - every lane owning one output
// row, bias fetched from global inside the loop, 32 scattered sectors per store instruction - held the kernels at
// 290-760 TFLOP/s (stalls: long scoreboard + LSU queue throttle).  Here each epilogue warp owns a 32-row x 64-column
// piece of the tile and a private 4 KB staging buffer: bias comes from shared memory, the fp32 residual arrives by a
// TMA tile load into that buffer, results are written back in place in the 128-byte-swizzled layout (conflict-free
// for row-per-lane 16-byte accesses) and leave through one TMA tile store per 32 x 128-byte piece - full lines,
// asynchronous, no LSU traffic.  Warps never synchronise with each other.
struct TmaEpilogue {
  const float* bias;  // [N] or null
  int act;
  int has_res;        // residual tile (fp32 with an fp32 output, bf16 with a bf16 output) is read through tmRes
  int out_f32;        // 1: fp32 output in 32-column pieces, 0: bf16 output in 64-column pieces
  // QKV scatter (qkv != 0, bf16): columns [0,d) -> q [M,d] through tmOut; [d,2d) / [2d,3d) -> the K / V cache,
  // addressed as a [(B*H*Smax), 64] matrix through tmRes / tmV: a piece is one head's 64 dims of 32 consecutive tokens
  // of one sequence (T % 32 == 0), i.e. 32 consecutive cache rows
  int qkv, d, H, T, Smax, pos0;
  // tile order of the persistent schedule.  1: the column tiles of one row block are neighbours (CTAs of one round read
  // a row block of A together, so A crosses HBM once: at K = 3072-5120 the 100-170 MB of A do not survive in L2 between
  // passes, and one pass per column tile measured 292 MB of DRAM reads for a 130 MB GEMM); 0: row blocks first
  int raster_n;
};

// EW epilogue warps.  16: every warp owns one 64-column slice of a tile, bias in shared memory, 3 mainloop stages.
// 8: two slices per warp and the bias in registers (fetched by shuffle), which frees the room for a 4th stage - for
// long-K GEMMs (K >= 2048) whose mainloop hides the longer epilogue and pays for a short ring instead.
// CG = 2: CTA pair (cta_group::2) - a 256 x BN tile per pair, each CTA stages its own 128 rows of A and HALF of the B tile
template <int BN, int EW, int CG = 1> struct PersistTmaLayout {
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int B_BYTES = (BN / CG) * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int EPI_BYTES = EW * 4096;                      // one swizzled 32 x 128 B buffer per epilogue warp
  static constexpr int BIAS_BYTES = EW == 16 ? EW * 256 : 0;       // 64 floats per epilogue warp
  static constexpr int BAR_BYTES = 256;
  static constexpr int STAGES_FIT = (227 * 1024 - EPI_BYTES - BIAS_BYTES - BAR_BYTES - 1024) / STAGE_BYTES;  // 3 / 4 at BN=256
  static constexpr int STAGES = STAGES_FIT > 6 ? 6 : STAGES_FIT;
  static constexpr int TILE_BYTES = STAGES * STAGE_BYTES;
  static constexpr int TOTAL = TILE_BYTES + EPI_BYTES + BIAS_BYTES + BAR_BYTES + 1024;
};

// ---- CTA pair (cta_group::2) plumbing: the leader CTA (cluster rank 0) issues the MMAs for both SMs; operands are staged
// by both CTAs, whose TMA loads complete on the LEADER's barrier; commits are multicast to both CTAs' barriers.
__device__ __forceinline__ uint32_t leader_addr(uint32_t local_smem_addr) {  // same offset in the cluster's CTA 0
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, 0;" : "=r"(r) : "r"(local_smem_addr));
  return r;
}
__device__ __forceinline__ void mbar_expect_tx_cluster(uint32_t cluster_addr, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cluster.b64 _, [%0], %1;" ::"r"(cluster_addr), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// TMA tile load of a CTA pair: data into this CTA's shared memory, completion on the leader's barrier `leader_bar`
__device__ __forceinline__ void tma_load_2d_pair(void* dst, const CUtensorMap* map, uint32_t leader_bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(dst)),
      "l"(map), "r"(leader_bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tc_mma_bf16_pair(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tc_commit_pair(uint64_t* bar) {  // arrives on `bar` in both CTAs of the pair
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                   smem_u32(bar)),
               "h"((uint16_t)3)
               : "memory");
}

template <int BN, int EW, int CG>
__global__ void __launch_bounds__(64 + 32 * EW) gemm_bf16_persistent_tma_kernel(
    const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
    const __grid_constant__ CUtensorMap tmOut, const __grid_constant__ CUtensorMap tmRes,
    const __grid_constant__ CUtensorMap tmV, int M, int N, int K, TmaEpilogue epi) {
  using L = PersistTmaLayout<BN, EW, CG>;
  static_assert(BN == 256 && (EW == 16 || EW == 8), "four 64-column slices per tile row block, 1 or 2 per warp");
  static_assert(CG == 1 || CG == 2, "one CTA or a CTA pair");
  constexpr int SPW = 16 / EW;  // 64-column slices per epilogue warp
  constexpr int TMEM_COLS = 2 * BN;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* tiles = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* stage_all = tiles + L::TILE_BYTES;  // 1 KB aligned: TILE_BYTES is a multiple of 1024
  float* bias_all = reinterpret_cast<float*>(stage_all + L::EPI_BYTES);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(stage_all + L::EPI_BYTES + L::BIAS_BYTES);
  uint64_t* empty_bar = full_bar + L::STAGES;
  uint64_t* tmem_full_bar = empty_bar + L::STAGES;  // [2]
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;     // [2]
  uint64_t* res_bar = tmem_empty_bar + 2;           // [EW]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(res_bar + EW);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nk = K / BK;
  // a tile is (BM * CG) x BN; CTA `rank` of a pair owns rows [rank * BM, +BM) of it and stages B rows [rank * BN/2, +BN/2)
  const int rank = CG == 2 ? (int)cluster_ctarank() : 0;
  const int m_tiles = (M + BM * CG - 1) / (BM * CG), n_tiles = (N + BN - 1) / BN;
  const int ntiles = m_tiles * n_tiles;
  const int tile0 = (int)blockIdx.x / CG, tile_step = (int)gridDim.x / CG;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmOut) : "memory");
    // pair: both CTAs' producers announce their bytes on the leader's full barrier; both CTAs' epilogue warps free the accumulator
    for (int s = 0; s < L::STAGES; ++s) { mbar_init(&full_bar[s], CG); mbar_init(&empty_bar[s], 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(&tmem_full_bar[a], 1); mbar_init(&tmem_empty_bar[a], EW * CG); }
    for (int w = 0; w < EW; ++w) mbar_init(&res_bar[w], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 1) {
    if (CG == 2) {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(TMEM_COLS)
                   : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    } else {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(TMEM_COLS)
                   : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
  }
  tc_fence_before();
  __syncthreads();
  if (CG == 2) cluster_sync_all();  // the peer's barriers are initialised before anybody arrives on them remotely
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===== TMA producer (in both CTAs of a pair) =====
    if (lane == 0) {
      uint32_t it = 0;
      for (int t = tile0; t < ntiles; t += tile_step) {
        const int tm = epi.raster_n ? t / n_tiles : t % m_tiles, tn = epi.raster_n ? t % n_tiles : t / m_tiles;
        const int a_row0 = tm * BM * CG + rank * BM, b_row0 = tn * BN + rank * (BN / CG);
        for (int kb = 0; kb < nk; ++kb, ++it) {
          const int s = it % L::STAGES;
          mbar_wait(&empty_bar[s], ((it / L::STAGES) & 1) ^ 1, 0);
          uint8_t* sa = tiles + s * L::STAGE_BYTES;
          if (CG == 2) {
            const uint32_t lbar = leader_addr(smem_u32(&full_bar[s]));
            mbar_expect_tx_cluster(lbar, L::STAGE_BYTES);
            tma_load_2d_pair(sa, &tmA, lbar, kb * BK, a_row0);
            tma_load_2d_pair(sa + L::A_BYTES, &tmB, lbar, kb * BK, b_row0);
          } else {
            mbar_expect_tx(&full_bar[s], L::STAGE_BYTES);
            tma_load_2d(sa, &tmA, &full_bar[s], kb * BK, a_row0);
            tma_load_2d(sa + L::A_BYTES, &tmB, &full_bar[s], kb * BK, b_row0);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (pair: the leader CTA only; its MMAs run on both SMs and read both CTAs' shared memory) =====
    if (lane == 0 && rank == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(BM * CG, BN);
      uint32_t it = 0, ti = 0;
      for (int t = tile0; t < ntiles; t += tile_step, ++ti) {
        const uint32_t acc = ti & 1;
        mbar_wait(&tmem_empty_bar[acc], ((ti >> 1) & 1) ^ 1, 3);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = 0; kb < nk; ++kb, ++it) {
          const int s = it % L::STAGES;
          mbar_wait(&full_bar[s], (it / L::STAGES) & 1, 1);
          tc_fence_after();
          const uint32_t sa = smem_u32(tiles + s * L::STAGE_BYTES);
          const uint64_t da = umma_desc_sw128(sa), db = umma_desc_sw128(sa + L::A_BYTES);  // k-step: address field + 2
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            if (CG == 2)
              tc_mma_bf16_pair(d_tmem, da + (uint64_t)(k * (UMMA_K * 2 >> 4)), db + (uint64_t)(k * (UMMA_K * 2 >> 4)), idesc,
                               (kb | k) != 0);
            else
              tc_mma_bf16(d_tmem, da + (uint64_t)(k * (UMMA_K * 2 >> 4)), db + (uint64_t)(k * (UMMA_K * 2 >> 4)), idesc,
                          (kb | k) != 0);
          }
          if (CG == 2) tc_commit_pair(&empty_bar[s]); else tc_commit(&empty_bar[s]);
        }
        if (CG == 2) tc_commit_pair(&tmem_full_bar[acc]); else tc_commit(&tmem_full_bar[acc]);
      }
    }
  } else {
    // ===== epilogue: warp w owns TMEM lanes of quadrant w % 4 (32 output rows) and columns [64 q, 64 q + 64) =====
    const int ew = warp - 2, quad = warp & 3;
    const uint32_t stage = smem_u32(stage_all + ew * 4096);
    float* sbias = bias_all + ew * 64;  // (EW == 16 only)
    float2 breg = make_float2(0.f, 0.f);  // EW == 8: this lane's two bias values of the current slice
    uint64_t* rbar = &res_bar[ew];
    uint32_t ti = 0, res_uses = 0;
    const uint32_t swz = (uint32_t)(lane & 7);
    const uint32_t my_row = stage + lane * 128;
    // the accumulator is handed back to the MMA issuer, i.e. (pair) to the leader CTA's barrier
    auto free_acc = [&](uint32_t acc) {
      if (CG == 2) mbar_arrive_cluster(leader_addr(smem_u32(&tmem_empty_bar[acc])));
      else mbar_arrive(&tmem_empty_bar[acc]);
    };
    for (int t = tile0; t < ntiles; t += tile_step, ++ti) {
      const int tm = epi.raster_n ? t / n_tiles : t % m_tiles, tn = epi.raster_n ? t % n_tiles : t / m_tiles;
      const int a_row0 = tm * BM * CG + rank * BM, b_row0 = tn * BN;
      const uint32_t acc = ti & 1;
      const int row0 = a_row0 + quad * 32;
#pragma unroll 1
      for (int rep = 0; rep < SPW; ++rep) {
      const int cslice = (ew >> 2) * SPW + rep, col0 = b_row0 + cslice * 64;
      const bool last_slice = rep == SPW - 1;
      const bool live = row0 < M && col0 < N;  // warp-uniform; a dead piece only hands the accumulator back
      if (live) {
        // ahead of the accumulator: bias of these 64 columns (shared memory or registers), first residual piece on its way
        if (EW == 16) {
          if (lane < 16) {
            const int n = col0 + lane * 4;
            float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f);
            if (epi.bias && n < N) b4 = *reinterpret_cast<const float4*>(epi.bias + n);  // N % 4 == 0
            *reinterpret_cast<float4*>(sbias + lane * 4) = b4;
          }
        } else {
          const int n = col0 + lane * 2;
          breg = (epi.bias && n < N) ? *reinterpret_cast<const float2*>(epi.bias + n) : make_float2(0.f, 0.f);
        }
        if (epi.has_res && lane == 0) {
          tma_store_wait_read();  // the previous tile's store has left the buffer
          mbar_expect_tx(rbar, 4096);
          tma_load_2d(reinterpret_cast<void*>(stage_all + ew * 4096), &tmRes, rbar, col0, row0);
        }
        __syncwarp();
      }
      if (rep == 0) {
        mbar_wait(&tmem_full_bar[acc], (ti >> 1) & 1, 2);
        tc_fence_after();
      }
      const uint32_t taddr = tmem_base + acc * BN + ((uint32_t)(quad * 32) << 16) + cslice * 64;
#ifdef TUTOR_TRACE
      if (g_epi_skip == 1) {  // tools/microbench_gemm.cu: mainloop only
        if (epi.has_res && live) { mbar_wait(rbar, res_uses & 1, 5); ++res_uses; }
        tc_fence_before();
        __syncwarp();
        if (lane == 0 && last_slice) free_acc(acc);
        continue;
      }
#endif
      if (!live) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0 && last_slice) free_acc(acc);
        continue;
      }
      // bias of element i (0..63) of this slice
      auto bias_at = [&](int i) -> float {
        if (EW == 16) return sbias[i];
        return __shfl_sync(0xffffffffu, (i & 1) ? breg.y : breg.x, i >> 1);
      };
      if (epi.out_f32) {
        // two pieces of 32 fp32 columns (128-byte rows)
#pragma unroll 1
        for (int sub = 0; sub < 2; ++sub) {
          float v[32];
          tc_ld32(taddr + sub * 32, v);
          if (sub == 1 && last_slice) {  // last read of this accumulator by this warp
            tc_fence_before();
            __syncwarp();
            if (lane == 0) free_acc(acc);
          }
          const int c0 = col0 + sub * 32;
          if (c0 >= N) continue;  // (the TMA store would clip it anyway)
          if (sub == 0) {
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = apply_act_fast(v[i] + bias_at(i), epi.act);
          } else {
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = apply_act_fast(v[i] + bias_at(32 + i), epi.act);
          }
          if (epi.has_res) {
            mbar_wait(rbar, res_uses & 1, 5);
            ++res_uses;
#pragma unroll
            for (int c = 0; c < 8; ++c) {
              float4 r;
              asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                           : "r"(my_row + (((uint32_t)c ^ swz) << 4)));
              v[4 * c] += r.x; v[4 * c + 1] += r.y; v[4 * c + 2] += r.z; v[4 * c + 3] += r.w;
            }
          } else {
            if (lane == 0) tma_store_wait_read();  // the buffer's previous contents have left
            __syncwarp();
          }
#pragma unroll
          for (int c = 0; c < 8; ++c)
            asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(my_row + (((uint32_t)c ^ swz) << 4)), "f"(v[4 * c]),
                         "f"(v[4 * c + 1]), "f"(v[4 * c + 2]), "f"(v[4 * c + 3])
                         : "memory");
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          __syncwarp();
          if (lane == 0) {
            tma_store_2d(&tmOut, stage, c0, row0);
            if (epi.has_res && sub == 0 && c0 + 32 < N) {  // second residual piece into the same buffer
              tma_store_wait_read();
              mbar_expect_tx(rbar, 4096);
              tma_load_2d(reinterpret_cast<void*>(stage_all + ew * 4096), &tmRes, rbar, c0 + 32, row0);
            }
          }
          __syncwarp();
        }
      } else if (epi.has_res) {
        // one piece of 64 bf16 columns added to the bf16 residual piece that sits where the output will go; the
        // accumulator is read in two halves, which keeps this path inside the register budget of the 64-column one
        mbar_wait(rbar, res_uses & 1, 5);
        ++res_uses;
#pragma unroll 1
        for (int sub = 0; sub < 2; ++sub) {
          float v[32];
          tc_ld32(taddr + sub * 32, v);
          if (sub == 1) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0 && last_slice) free_acc(acc);
          }
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            const uint32_t addr = my_row + (((uint32_t)(sub * 4 + c) ^ swz) << 4);
            uint4 rv;
            asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(rv.x), "=r"(rv.y), "=r"(rv.z), "=r"(rv.w) : "r"(addr));
            float r[8];
            unpack_bf16x8(rv, r);
            uint32_t w[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const int i = 8 * c + 2 * j;
              __nv_bfloat162 h2 =
                  __floats2bfloat162_rn(apply_act_fast(v[i] + bias_at(sub * 32 + i), epi.act) + r[2 * j],
                                        apply_act_fast(v[i + 1] + bias_at(sub * 32 + i + 1), epi.act) + r[2 * j + 1]);
              w[j] = *reinterpret_cast<uint32_t*>(&h2);
            }
            asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]) : "memory");
          }
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) tma_store_2d(&tmOut, stage, col0, row0);
        __syncwarp();
      } else {
        // one piece of 64 bf16 columns (128-byte rows)
        float v[64];
#ifdef TUTOR_TRACE
        if (g_epi_skip == 3) {  // tools/microbench_gemm.cu: no TMEM read
#pragma unroll
          for (int i = 0; i < 64; ++i) v[i] = 0.f;
        } else
#endif
        tc_ld64(taddr, v);
        tc_fence_before();
        __syncwarp();
#ifdef TUTOR_TRACE
        if (g_epi_skip == 4) {  // TMEM read only: no epilogue math, no shared-memory or global traffic
          if (lane == 0 && last_slice) free_acc(acc);
          if (v[lane] == 12345.f) sbias[0] = v[0];  // keep the load
          continue;
        }
#endif
        if (lane == 0) {
          if (last_slice) free_acc(acc);
          tma_store_wait_read();  // the buffer's previous contents have left
        }
        __syncwarp();
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          uint32_t w[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int i = 8 * c + 2 * j;
            __nv_bfloat162 h2 = __floats2bfloat162_rn(apply_act_fast(v[i] + bias_at(i), epi.act),
                                                     apply_act_fast(v[i + 1] + bias_at(i + 1), epi.act));
            w[j] = *reinterpret_cast<uint32_t*>(&h2);
          }
          asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(my_row + (((uint32_t)c ^ swz) << 4)), "r"(w[0]),
                       "r"(w[1]), "r"(w[2]), "r"(w[3])
                       : "memory");
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) {
          if (!epi.qkv || col0 < epi.d) {
            tma_store_2d(&tmOut, stage, col0, row0);
          } else {  // this head's rows of the K or V cache (replaces DynamicLayer.update's torch.cat)
            const int which = col0 >= 2 * epi.d, c = col0 - (which ? 2 * epi.d : epi.d);
            const int b = row0 / epi.T, t0 = row0 - b * epi.T;
            const int crow = (b * epi.H + (c >> 6)) * epi.Smax + epi.pos0 + t0;
            tma_store_2d(which ? &tmV : &tmRes, stage, 0, crow);
          }
        }
        __syncwarp();
      }
      }  // slices of this warp
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // stores complete before the CTA exits
  }

  tc_fence_before();
  __syncthreads();
  if (CG == 2) cluster_sync_all();  // neither CTA of a pair leaves (or frees TMEM) while the other may still use it
  if (warp == 1) {
    if (CG == 2) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
    else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
  }
}

// ------------------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn g_encode = nullptr;
std::once_flag g_encode_once;

int get_encoder() {
  std::call_once(g_encode_once, [] {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      g_encode = reinterpret_cast<EncodeTiledFn>(fn);
  });
  if (!g_encode) {
    tutor_set_error("cuTensorMapEncodeTiled is not available from the CUDA driver");
    return TUTOR_ERR_CUDA;
  }
  return TUTOR_OK;
}

// 2-D bf16 tensor [rows, K] row-major; box = [box_rows, 64] with the 128-byte swizzle the UMMA descriptor expects
int make_map(CUtensorMap* map, const bf16* ptr, int rows, int K, int box_rows) {
  TRY(get_encoder());
  ARG_CHECK((reinterpret_cast<uintptr_t>(ptr) & 15) == 0, "gemm_bf16: operand not 16-byte aligned");
  cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)K * sizeof(bf16)};
  cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = g_encode(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<bf16*>(ptr), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    tutor_set_error("cuTensorMapEncodeTiled failed (%d) rows=%d K=%d box_rows=%d", (int)r, rows, K, box_rows);
    return TUTOR_ERR_CUDA;
  }
  return TUTOR_OK;
}

// fp32 partial tensor [ks][M][N] for the EPI_PARTIAL bulk store; box = [1][box_m][128], no swizzle
int make_map_out(CUtensorMap* map, float* ptr, int N, int M, int ks, int box_m) {
  TRY(get_encoder());
  ARG_CHECK((reinterpret_cast<uintptr_t>(ptr) & 15) == 0 && N % 4 == 0, "gemm_bf16: partial buffer misaligned");
  cuuint64_t dims[3] = {(cuuint64_t)N, (cuuint64_t)M, (cuuint64_t)ks};
  cuuint64_t strides[2] = {(cuuint64_t)N * 4, (cuuint64_t)M * N * 4};
  cuuint32_t box[3] = {(cuuint32_t)BM, (cuuint32_t)box_m, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = g_encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, ptr, dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    tutor_set_error("cuTensorMapEncodeTiled (partials) failed (%d) N=%d M=%d ks=%d", (int)r, N, M, ks);
    return TUTOR_ERR_CUDA;
  }
  return TUTOR_OK;
}

// [rows, cols] row-major tensor of fp32 or bf16 cut into 32-row x 128-byte boxes (the epilogue warps' staging tiles)
int make_map_epi(CUtensorMap* map, const void* ptr, bool f32, int rows, int cols) {
  TRY(get_encoder());
  const size_t es = f32 ? 4 : 2;
  ARG_CHECK((reinterpret_cast<uintptr_t>(ptr) & 15) == 0 && (cols * es) % 16 == 0, "gemm_bf16: output not 16-byte aligned");
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)cols * es};
  cuuint32_t box[2] = {(cuuint32_t)(128 / es), 32};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = g_encode(map, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                        const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    tutor_set_error("cuTensorMapEncodeTiled (epilogue) failed (%d) rows=%d cols=%d", (int)r, rows, cols);
    return TUTOR_ERR_CUDA;
  }
  return TUTOR_OK;
}

template <int BN, int EPI>
int launch(const bf16* A, int rowsA, const bf16* B, int rowsB, int K, const GemmEpilogue& epi, const ArgmaxArgs& am,
           cudaStream_t st, int ks = 1) {
  using L = SmemLayout<BN, (EPI == EPI_ARGMAX && BN == 64) ? 3 : 0>;
  TRY(tutor_set_max_smem(reinterpret_cast<const void*>(gemm_bf16_tcgen05_kernel<BN, EPI>), L::TOTAL));
  CUtensorMap ta, tb, tc;
  TRY(make_map(&ta, A, rowsA, K, BM));
  TRY(make_map(&tb, B, rowsB, K, BN));
  if (EPI == EPI_PARTIAL) TRY(make_map_out(&tc, epi.out_f32, rowsA, rowsB, ks, BN));
  else tc = ta;  // unused
  dim3 grid(ceil_div(rowsA, BM), ceil_div(rowsB, BN), ks);
  CUDA_TRY(tutor_launch(gemm_bf16_tcgen05_kernel<BN, EPI>, grid, dim3(EPI == EPI_ROWMAJOR ? NTHREADS_WIDE : NTHREADS),
                        (size_t)L::TOTAL, st, ta, tb, tc, rowsA,
                        rowsB, K, epi, am));
  KERNEL_LAUNCHED();
  return TUTOR_OK;
}

template <int BN>
int launch_persistent(const bf16* A, int M, const bf16* W, int N, int K, const GemmEpilogue& epi, cudaStream_t st) {
  using L = PersistLayout<BN>;
  TRY(tutor_set_max_smem(reinterpret_cast<const void*>(gemm_bf16_persistent_kernel<BN>), L::TOTAL));
  CUtensorMap ta, tb;
  TRY(make_map(&ta, A, M, K, BM));
  TRY(make_map(&tb, W, N, K, BN));
  int dev = 0, sms = 148;
  CUDA_TRY(cudaGetDevice(&dev));
  CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int ntiles = ceil_div(M, BM) * ceil_div(N, BN);
  CUDA_TRY(tutor_launch(gemm_bf16_persistent_kernel<BN>, dim3(ntiles < sms ? ntiles : sms), dim3(NTHREADS_PERSIST),
                        (size_t)L::TOTAL, st, ta, tb, M, N, K, epi));
  KERNEL_LAUNCHED();
  return TUTOR_OK;
}

// the TMA-staged epilogue covers plain row-major outputs: bias -> activation -> (+ fp32 residual) -> fp32, or
// bias -> activation -> (+ bf16 residual) -> bf16
bool tma_epilogue_ok(const GemmEpilogue& e, int M, int N) {
  if (e.qkv_mode) {  // q [M,d] + scatter into the K / V cache: whole 32-token pieces of one sequence, bf16 everywhere
    return e.out_bf16 && !e.out_f32 && !e.residual && !e.pos_dev && e.kv_dtype == DT_BF16 && e.T % 32 == 0 &&
           e.d % 64 == 0 && N == 3 * e.d && e.ldo == e.d && M % e.T == 0 && e.Smax >= e.pos0 + e.T &&
           !((reinterpret_cast<uintptr_t>(e.out_bf16) | reinterpret_cast<uintptr_t>(e.kcache) |
              reinterpret_cast<uintptr_t>(e.vcache) | reinterpret_cast<uintptr_t>(e.bias)) & 15);
  }
  if (e.ldo != N) return false;
  if ((e.out_f32 != nullptr) == (e.out_bf16 != nullptr)) return false;  // exactly one output
  if (e.residual && !e.out_f32) return false;
  if (e.residual_bf16 && (!e.out_bf16 || e.residual)) return false;
  const void* out = e.out_f32 ? (const void*)e.out_f32 : (const void*)e.out_bf16;
  if ((reinterpret_cast<uintptr_t>(out) & 15) || (e.residual && (reinterpret_cast<uintptr_t>(e.residual) & 15)) ||
      (e.residual_bf16 && (reinterpret_cast<uintptr_t>(e.residual_bf16) & 15)))
    return false;
  if (e.bias && (reinterpret_cast<uintptr_t>(e.bias) & 15)) return false;
  return N % 8 == 0;
}

template <int EW, int CG>
int launch_persistent_tma(const bf16* A, int M, const bf16* W, int N, int K, const GemmEpilogue& e, cudaStream_t st) {
  using L = PersistTmaLayout<256, EW, CG>;
  static_assert(L::TOTAL <= 227 * 1024, "shared-memory budget");
  TRY(tutor_set_max_smem(reinterpret_cast<const void*>(gemm_bf16_persistent_tma_kernel<256, EW, CG>), L::TOTAL));
  CUtensorMap ta, tb, to, tr, tv;
  TRY(make_map(&ta, A, M, K, BM));
  TRY(make_map(&tb, W, N, K, 256 / CG));
  const bool f32 = e.out_f32 != nullptr;
  TmaEpilogue te{e.bias, e.act, e.residual != nullptr || e.residual_bf16 != nullptr, f32 ? 1 : 0, 0, 0, 0, 0, 0, 0, 0};
  static const int raster = getenv("TUTOR_GEMM_RASTER_M") ? 0 : 1;  // TUTOR_GEMM_RASTER_M=1: row blocks first (A/B measurements)
  te.raster_n = raster;
  if (e.qkv_mode) {
    const int cache_rows = (M / e.T) * e.H * e.Smax;
    TRY(make_map_epi(&to, e.out_bf16, false, M, e.d));
    TRY(make_map_epi(&tr, e.kcache, false, cache_rows, 64));
    TRY(make_map_epi(&tv, e.vcache, false, cache_rows, 64));
    te.qkv = 1; te.d = e.d; te.H = e.H; te.T = e.T; te.Smax = e.Smax; te.pos0 = e.pos0;
  } else {
    TRY(make_map_epi(&to, f32 ? (const void*)e.out_f32 : (const void*)e.out_bf16, f32, M, N));
    if (e.residual) TRY(make_map_epi(&tr, e.residual, true, M, N));
    else if (e.residual_bf16) TRY(make_map_epi(&tr, e.residual_bf16, false, M, N));
    else tr = to;  // unused
    tv = to;       // unused
  }
  int dev = 0, sms = 148;
  CUDA_TRY(cudaGetDevice(&dev));
  CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int ntiles = ceil_div(M, BM * CG) * ceil_div(N, 256);
  if (CG == 2) {
    // CTA pairs: a cluster of two along x; as many pairs as there are tiles, at most one CTA per SM
    const int pairs = ntiles < sms / 2 ? ntiles : sms / 2;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * pairs); cfg.blockDim = dim3(64 + 32 * EW); cfg.dynamicSmemBytes = (size_t)L::TOTAL; cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = tutor_pdl_enabled ? 2 : 1;
    CUDA_TRY(cudaLaunchKernelEx(&cfg, gemm_bf16_persistent_tma_kernel<256, EW, CG>, ta, tb, to, tr, tv, M, N, K, te));
  } else {
    CUDA_TRY(tutor_launch(gemm_bf16_persistent_tma_kernel<256, EW, CG>, dim3(ntiles < sms ? ntiles : sms), dim3(64 + 32 * EW),
                          (size_t)L::TOTAL, st, ta, tb, to, tr, tv, M, N, K, te));
  }
  KERNEL_LAUNCHED();
  return TUTOR_OK;
}

template <int EPI>
int launch_swapped(const bf16* W, int N, const bf16* X, int M, int K, const GemmEpilogue& epi, const ArgmaxArgs& am,
                   cudaStream_t st, int ks = 1) {
  // batch is the UMMA N dimension: pick the smallest tile that covers it, loop over batch tiles beyond 128
  if (M <= 16) return launch<16, EPI>(W, N, X, M, K, epi, am, st, ks);
  if (M <= 32) return launch<32, EPI>(W, N, X, M, K, epi, am, st, ks);
  if (M <= 64) return launch<64, EPI>(W, N, X, M, K, epi, am, st, ks);
  return launch<128, EPI>(W, N, X, M, K, epi, am, st, ks);
}

}  // namespace

// tensor-map builders for the other tcgen05 kernels of the library (gemm_decode_fused.cu)
int k_make_map_bf16(void* map, const bf16* ptr, int rows, int K, int box_rows) {
  return make_map(reinterpret_cast<CUtensorMap*>(map), ptr, rows, K, box_rows);
}
// generic row-major 2-D tensor [rows, cols] of 4-byte (fp32) or 2-byte (bf16) elements, boxes of box_rows x box_cols
// (box_cols * element size must be 128 bytes: one swizzle atom per box row)
int k_make_map_2d(void* map, const void* ptr, int elem_bytes, int rows, int cols, int box_rows, int box_cols) {
  TRY(get_encoder());
  ARG_CHECK((elem_bytes == 4 || elem_bytes == 2) && box_cols * elem_bytes == 128 && box_rows >= 1 && box_rows <= 256,
            "make_map_2d: unsupported box %d x %d of %d-byte elements", box_rows, box_cols, elem_bytes);
  ARG_CHECK((reinterpret_cast<uintptr_t>(ptr) & 15) == 0 && ((size_t)cols * elem_bytes) % 16 == 0,
            "make_map_2d: tensor not 16-byte aligned");
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)cols * elem_bytes};
  cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = g_encode(reinterpret_cast<CUtensorMap*>(map),
                        elem_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                        const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    tutor_set_error("cuTensorMapEncodeTiled failed (%d) rows=%d cols=%d box=%dx%d", (int)r, rows, cols, box_rows, box_cols);
    return TUTOR_ERR_CUDA;
  }
  return TUTOR_OK;
}
int k_make_map_partial(void* map, float* ptr, int N, int M, int ks, int box_m) {
  return make_map_out(reinterpret_cast<CUtensorMap*>(map), ptr, N, M, ks, box_m);
}

int k_gemm_bf16(const bf16* A, const bf16* W, int M, int N, int K, const GemmEpilogue& epi, cudaStream_t st) {
  ARG_CHECK(K % BK == 0, "gemm_bf16: K=%d must be a multiple of %d", K, BK);
  if (M == 0 || N == 0) return TUTOR_OK;
  ArgmaxArgs am{};
  // few rows: the weight matrix as the 128-row UMMA operand, the rows as its N dimension.  Above TUTOR_SWAP_MAX_M rows
  // (default 32) the row-major kernel's single row tile is faster: a 128-token BERT query took 2.7 ms through the
  // swapped kernel's per-element epilogue.
  static const int swap_max = tutor_experiment("TUTOR_SWAP_MAX_M") ? atoi(tutor_experiment("TUTOR_SWAP_MAX_M")) : 32;
  if (M <= (swap_max < 128 ? swap_max : 128)) return launch_swapped<EPI_SWAPPED>(W, N, A, M, K, epi, am, st);
  // 128x256 tiles halve the operand bytes per FLOP of 128x128 ones (the 128x128 kernel measured L2-bandwidth
  // bound: 32 KB of operands per 4 MMAs = 6.2 TB/s at 410 TFLOP/s); two 97 KB CTAs still share an SM
  static const bool wide = tutor_experiment("TUTOR_GEMM_BN128") == nullptr;
  static const bool persistent = tutor_experiment("TUTOR_GEMM_NONPERSISTENT") == nullptr;  // debugging aid / A-B measurements
  static const bool tma_epi = tutor_experiment("TUTOR_GEMM_NO_TMA_EPILOGUE") == nullptr;  // A/B measurements
  if (persistent && tma_epi && N >= 128 && tma_epilogue_ok(epi, M, N)) {
    // long K: the mainloop hides an 8-warp epilogue and gains a 4th pipeline stage (TUTOR_GEMM_EW16=1 forces 16 warps)
    static const bool ew16 = tutor_experiment("TUTOR_GEMM_EW16") != nullptr;

    // CTA pairs (cta_group::2, 256 x 256 tiles, each CTA stages half of B: a third less shared-memory traffic per FLOP,
    // 5-9 % faster at the BERT / GPT-2 large shapes); TUTOR_GEMM_PAIR=0 keeps one CTA per tile (cross-check)
    static const bool pair = getenv("TUTOR_GEMM_PAIR") == nullptr || atoi(getenv("TUTOR_GEMM_PAIR")) != 0;
    if (pair && M >= 256) {
      // pairs stage half of B per CTA, so 8 epilogue warps leave room for 6 stages instead of 4: faster at every K
      // (QKV 16384 x 2304 x 768: 49.5 vs 54.4 us, intermediate + GELU 71.3 vs 77.2 us)
      if (!ew16) return launch_persistent_tma<8, 2>(A, M, W, N, K, epi, st);
      return launch_persistent_tma<16, 2>(A, M, W, N, K, epi, st);
    }
    if (K >= 2048 && !ew16) return launch_persistent_tma<8, 1>(A, M, W, N, K, epi, st);
    return launch_persistent_tma<16, 1>(A, M, W, N, K, epi, st);
  }
  if (persistent) {
    if (wide && N >= 256) return launch_persistent<256>(A, M, W, N, K, epi, st);
    return launch_persistent<128>(A, M, W, N, K, epi, st);
  }
  if (wide && N >= 256) return launch<256, EPI_ROWMAJOR>(A, M, W, N, K, epi, am, st);
  return launch<128, EPI_ROWMAJOR>(A, M, W, N, K, epi, am, st);
}

// Decode-time split-K: ks CTAs share one 128-row weight tile, each streaming K/ks of it, so that a GEMM whose
// whole weight is only a few MB still spreads over ~all SMs.  Raw fp32 partials [ks][M][N] go to `partial`
// (L2-resident); the fused reduce kernels in decode_fused.cu sum them in a fixed order (deterministic).
int k_gemm_bf16_splitk(const bf16* X, const bf16* W, int M, int N, int K, int ks, float* partial, cudaStream_t st) {
  ARG_CHECK(K % BK == 0 && ks >= 1 && (K / BK) % ks == 0, "gemm_bf16_splitk: K=%d not divisible into %d splits", K, ks);
  ARG_CHECK(M >= 1 && M <= 128, "gemm_bf16_splitk: batch %d outside 1..128", M);
  GemmEpilogue epi{};
  epi.out_f32 = partial;
  ArgmaxArgs am{};
  return launch_swapped<EPI_PARTIAL>(W, N, X, M, K, epi, am, st, ks);
}

// number of K-splits for a decode GEMM: the largest divisor of K/64 that keeps tiles*ks within one wave of SMs
int k_pick_splitk(int N, int K, int num_sms) {
  const int tiles = ceil_div(N, BM), nk = K / BK;
  int best = 1;
  // capped at 12: beyond that the fp32 partials (ks*M*N*4 B written, then read) outweigh the shorter weight stream
  for (int ks = 1; ks <= nk && ks <= 12; ++ks)
    if (nk % ks == 0 && tiles * ks <= num_sms) best = ks;
  return best;
}

// fused LM head + repetition penalty + per-tile argmax; partials [ceil(V/128)][B]
int k_gemm_bf16_argmax(const bf16* wte, int V, const bf16* X, int B, int K, const uint32_t* seen, int seen_words,
                       float penalty, float* part_val, int* part_idx, cudaStream_t st) {
  ARG_CHECK(K % BK == 0, "lmhead: K=%d must be a multiple of %d", K, BK);
  ARG_CHECK(B <= 128, "lmhead: batch %d > 128 per call", B);
  GemmEpilogue epi{};
  ArgmaxArgs am{seen, seen_words, penalty, V, part_val, part_idx, B};
  // rows past V are zero-filled by TMA and masked to -inf in the epilogue: no padded copy of wte needed
  return launch_swapped<EPI_ARGMAX>(wte, V, X, B, K, epi, am, st);
}
