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
  using L = SmemLayout<BN>;
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
        const uint32_t sb = sa + L::A_BYTES;
#pragma unroll
        for (int k = 0; k < BK / UMMA_K; ++k) {
          // advance 16 bf16 = 32 bytes along K inside the 128-byte swizzle atom
          tc_mma_bf16(tmem_base, umma_desc_sw128(sa + k * UMMA_K * 2), umma_desc_sw128(sb + k * UMMA_K * 2), idesc,
                      (kb | k) != 0);
        }
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
          if (b < am.B && am.penalty != 1.0f) {
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
          const uint32_t sb = sa + L::A_BYTES;
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k)
            tc_mma_bf16(d_tmem, umma_desc_sw128(sa + k * UMMA_K * 2), umma_desc_sw128(sb + k * UMMA_K * 2), idesc,
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
      for (int c0 = c_begin; c0 < c_end; c0 += 32) {
        float v[32];
        tc_ld32(taddr + c0, v);
        if (c0 + 32 >= c_end) {  // last read of this accumulator by this warp: hand it back before the stores
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tmem_empty_bar[acc]);
        }
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

template <int BN, int EPI>
int launch(const bf16* A, int rowsA, const bf16* B, int rowsB, int K, const GemmEpilogue& epi, const ArgmaxArgs& am,
           cudaStream_t st, int ks = 1) {
  using L = SmemLayout<BN>;
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
int k_make_map_partial(void* map, float* ptr, int N, int M, int ks, int box_m) {
  return make_map_out(reinterpret_cast<CUtensorMap*>(map), ptr, N, M, ks, box_m);
}

int k_gemm_bf16(const bf16* A, const bf16* W, int M, int N, int K, const GemmEpilogue& epi, cudaStream_t st) {
  ARG_CHECK(K % BK == 0, "gemm_bf16: K=%d must be a multiple of %d", K, BK);
  if (M == 0 || N == 0) return TUTOR_OK;
  ArgmaxArgs am{};
  if (M <= 128) return launch_swapped<EPI_SWAPPED>(W, N, A, M, K, epi, am, st);
  // 128x256 tiles halve the operand bytes per FLOP of 128x128 ones (the 128x128 kernel measured L2-bandwidth
  // bound: 32 KB of operands per 4 MMAs = 6.2 TB/s at 410 TFLOP/s); two 97 KB CTAs still share an SM
  static const bool wide = getenv("TUTOR_GEMM_BN128") == nullptr;
  static const bool persistent = getenv("TUTOR_GEMM_NONPERSISTENT") == nullptr;  // debugging aid / A-B measurements
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
