// One persistent kernel for the whole transformer stack of a bf16 decode step (one token per sequence):
//   embedding -> L x [ ln_1 + c_attn -> attention over the KV cache -> c_proj + residual -> ln_2 + c_fc + gelu_new
//                      -> mlp c_proj + residual ] -> ln_f
// (HF/models/gpt2/modeling_gpt2.py:262-309 per block, :577-607 around them; Conv1D at HF/pytorch_utils.py:119-123).
//
// Why: as separate launches the four linears of a block cost ~6 us each against ~0.6 us of weight streaming - a chain
// of launch gaps, prologues and cluster barriers (profiles/r1f_microbench_fused_decode_gemms.txt) - and nothing overlaps
// the HBM-bound attention.  Here one CTA per SM stays resident for the whole step and runs two independent engines:
//
//   * GEMM engine: walks a host-built list of steps.  A GEMM step is one linear layer for one row group (half of the
//     batch): "swap-AB" split-K tcgen05 tiles, the weight tile [128 x 64k] as the UMMA A operand streamed by TMA through a
//     ring that runs ahead of the data dependencies (weights are static), the activations [rows x 64k] as the UMMA N
//     operand, also by TMA, as soon as the step's input counter is complete.  LayerNorm is folded into the weights
//     (W' = gamma * W, see the engine's fold_ln): the operand is the raw residual stream in bf16 and the normalisation
//     becomes  rstd * (acc - mean * s[n]) + c[n]  when the split-K sum is closed, so no thread ever touches an operand.
//     Partial tiles go to L2; per-tile arrival counters close the split-K sum in a fixed order (deterministic) fused
//     with bias + residual + LayerNorm statistics, or LayerNorm correction + bias + gelu_new.  Steps are chained by
//     device-scope counters in L2 (release / relaxed poll), never by kernel boundaries or grid-wide barriers.  The CTAs
//     are dealt round-robin to the row groups, so the groups' chains advance side by side.
//   * attention engine (4 streams x 5 warps): every stream pulls (layer, row, head) units from one global queue, streams
//     the unit's K/V rows with cp.async.bulk through its own shared-memory ring (rows of earlier tokens are final, so they
//     are requested before the unit's q/k/v exist), finishes the c_attn split-K sums for its head (with the LayerNorm
//     correction), appends the new K/V row and runs the online softmax; streams of one CTA are out of phase, so
//     prologues and merges hide behind other streams' copies.
//
// With two row groups the GEMM chain of one group runs under the attention of the other, which keeps HBM busy.
#include <cuda.h>
#include <math_constants.h>

#include <type_traits>

#include "decode_mega.h"
#include "tc_ptx.cuh"

using namespace tcptx;

namespace {

constexpr int MG_NSLOT = 4;    // attention streams per CTA
constexpr int MG_CHUNK = 64;   // keys per ring stage (8 KB of K + 8 KB of V)
constexpr int MG_RING = 2;     // ring stages per stream
// Warp layout, in warpgroups of four so that registers can be moved between roles with setmaxnreg:
//   warps 0-3    epilogue + split-K closer (TMEM lane quadrant = warp & 3)                              -> 96 registers
//   warp 4 weight TMA producer, 5 MMA issuer, 6-9 K/V stream producers, 10 activation TMA producer      -> 40 registers
//   warps 12-27  attention consumers, four per stream                                                   -> 72 (as launched)
constexpr int MG_W_TMA = 4, MG_W_MMA = 5, MG_W_APROD = 6, MG_W_ACT = 10, MG_W_ACONS = 12;
constexpr int MG_THREADS = (MG_W_ACONS + 4 * MG_NSLOT) * 32;  // 896
constexpr int MG_REGS_EPI = 96, MG_REGS_LIGHT = 40;
constexpr int MG_MAX_STAT_TILES = 10;  // d <= 1280

struct __align__(128) AttnSlot {
  bf16 sK[MG_RING][MG_CHUNK * 64];
  bf16 sV[MG_RING][MG_CHUNK * 64];
  float s_new[3][64];
  float w_o[4][64];
  float w_m[4], w_l[4];
  uint64_t full[MG_RING], empty[MG_RING], dfull[2], dempty[2];
  int desc[2];
};

// A wait that does not end is a protocol bug.  It must neither hang the GPU nor call a function: any call (printf
// included) makes ptxas allocate the whole kernel under the smallest setmaxnreg value, and the epilogue warps then
// spill.  So a wait that exceeds ~2 s records (what, a, b) in p.err and returns; from then on every wait gives up after
// ~0.5 ms, the launch drains with garbage results, and the engine reports the code on its next host read.
__device__ __forceinline__ bool mega_give_up(int* err, long long t0, int what, int a, int b) {
  const long long dt = clock64() - t0;
  if (dt > (1LL << 20) && *reinterpret_cast<volatile int*>(err) != 0) return true;
  if (dt > 4000000000LL) {
    if (atomicCAS(err, 0, what) == 0) { err[1] = a; err[2] = b; err[3] = (int)blockIdx.x; err[4] = (int)threadIdx.x; }
    return true;
  }
  return false;
}
__device__ __forceinline__ int ld_relaxed(const int* p) {
  int v;
  asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// device-scope counter in L2 reaches `target` (red.release on the producing side).  Polled by ONE thread with relaxed
// loads; the caller then joins its group with a CTA barrier (or, for TMA, issues a proxy fence), and every load of data
// the counter guards goes to L2 (ld.global.cg / TMA / cp.async.bulk), never through L1.  No acquire fence:
// fence.acq_rel.gpu costs ~750 cycles per wait (tools/microbench_sync.cu) and, like ld.acquire, invalidates the SM's
// whole L1 (CCTL.IVALL), which turns the other warps' cached parameter loads and register spills into L2 round trips.
// The producer's release (membar.gpu before the reduction) makes the data visible in L2 before the counter moves, and a
// load issued after the poll cannot be served from anything older than L2.
__device__ __forceinline__ void wait_counter(int* err, const int* f, int target, int a, int b) {
  if (ld_relaxed(f) >= target) return;
  const long long t0 = clock64();
  while (ld_relaxed(f) < target) {
    if (mega_give_up(err, t0, 1, a, b)) break;
  }
}
// the same for waits that are long by design (an attention unit waiting for its c_attn): sleep between polls so that
// hundreds of waiting threads do not load the L2 slice the producers have to reach
__device__ __forceinline__ void wait_counter_backoff(int* err, const int* f, int target, int a, int b) {
  if (ld_relaxed(f) >= target) return;
  const long long t0 = clock64();
  while (ld_relaxed(f) < target) {
    __nanosleep(256);
    if (mega_give_up(err, t0, 1, a, b)) break;
  }
}
__device__ __forceinline__ void mg_wait(int* err, uint64_t* bar, uint32_t parity, int which) {
  if (mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (mega_give_up(err, t0, 2, which, (int)parity)) break;
  }
}
// Every counter sits in a 128-byte line of its own (MG_FLAG_STRIDE ints apart): hundreds of threads poll them, and
// counters packed into one line turn that line's L2 slice into a hot spot that delays every hand-off in the kernel.
#define MG_FLAG(i) (p.flags + (size_t)(i) * MG_FLAG_STRIDE)
// keeps a value in a register (opaque to the optimiser) instead of being re-derived from the kernel parameters
__device__ __forceinline__ int pin(int v) { asm volatile("" : "+r"(v)); return v; }
template <typename T> __device__ __forceinline__ T* pin(T* v) { asm volatile("" : "+l"(v)); return v; }
// counter += v with release semantics at device scope: everything this thread observed before it (its own writes and,
// through the CTA barrier that precedes every call, the writes of the other threads of its group) is visible to a
// thread that then sees the counter.  One thread does this after the barrier; a fence in every thread costs far more.
__device__ __forceinline__ void red_release(int* p, int v) {
  asm volatile("red.release.gpu.global.add.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void bar_named(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ uint2 pack_bf16x4(float4 v) {
  __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
  return make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
}
// LayerNorm (mean, rstd) of one row from its per-128-column (mean, M2) pairs: lane i < 16 with `has` holds pair i
// (Chan et al. combination, equal counts); every lane of the 16-lane half returns the row's values
__device__ __forceinline__ float2 row_mean_rstd(float2 tv, bool has, float inv_nt, float inv_d, float eps) {
  float sm = has ? tv.x : 0.f;
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) sm += __shfl_xor_sync(0xffffffffu, sm, o);
  const float mean = sm * inv_nt;
  const float dm = tv.x - mean;
  float m2 = has ? tv.y + (float)BM * dm * dm : 0.f;
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) m2 += __shfl_xor_sync(0xffffffffu, m2, o);
  return make_float2(mean, rsqrtf(m2 * inv_d + eps));
}

#define MG_TRACE_STEP(si, slot)                                                                         \
  do {                                                                                                  \
    if (tracing && t == 0 && (si) - p.step_begin < MG_TRACE_MAX_STEPS)                                   \
      p.trace[((si) - p.step_begin) * MG_TRACE_STEP_SLOTS + (slot)] = clock64();                         \
  } while (0)
#define MG_TRACE_UNIT(slot_id)                                                                          \
  do {                                                                                                  \
    if (tracing && slot == 0 && ct == 0 && tun < MG_TRACE_MAX_UNITS)                                     \
      p.trace[MG_TRACE_MAX_STEPS * MG_TRACE_STEP_SLOTS + tun * MG_TRACE_UNIT_SLOTS + (slot_id)] = clock64(); \
  } while (0)

template <int GN>
struct MegaSmem {
  static constexpr int NST = GN >= 64 ? 3 : 4;  // weight/activation pipeline stages
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int B_BYTES = GN * BK * 2;
  static constexpr int STAGE = A_BYTES + B_BYTES;
  static constexpr int TILES = NST * STAGE;
  static constexpr int CTRL = 1024;  // barriers, TMEM slot
  static constexpr int TOTAL = 1024 + TILES + CTRL + MG_NSLOT * (int)sizeof(AttnSlot);
};

template <int GN>
__global__ void __launch_bounds__(MG_THREADS, 1) decode_mega_kernel(const __grid_constant__ MegaParams p) {
  using SM = MegaSmem<GN>;
  constexpr int NST = SM::NST;
  constexpr int TSTRIDE = GN < 32 ? 32 : GN;  // TMEM columns per accumulator buffer
  constexpr int TMEM_COLS = 2 * TSTRIDE;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* tiles = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* fullA = reinterpret_cast<uint64_t*>(tiles + SM::TILES);
  uint64_t* fullB = fullA + NST;
  uint64_t* empty = fullB + NST;
  uint64_t* tmem_full = empty + NST;   // [2]
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);
  AttnSlot* slots = reinterpret_cast<AttnSlot*>(tiles + SM::TILES + SM::CTRL);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool attn_on = p.attn_begin < p.attn_end;

  if (threadIdx.x == 0) {
    for (int s = 0; s < NST; ++s) { mbar_init(&fullA[s], 1); mbar_init(&fullB[s], 1); mbar_init(&empty[s], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], 4); }
    for (int sl = 0; sl < MG_NSLOT; ++sl) {
      for (int s = 0; s < MG_RING; ++s) { mbar_init(&slots[sl].full[s], 1); mbar_init(&slots[sl].empty[s], 4); }
      for (int i = 0; i < 2; ++i) { mbar_init(&slots[sl].dfull[i], 1); mbar_init(&slots[sl].dempty[i], 4); }
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == MG_W_MMA) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"(TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  pdl_wait();  // the previous token's bookkeeping kernel owns ids / the length counter until here
  const int len = *p.len;  // tokens so far, the one being fed included: its position is len - 1
  const int nctas = (int)gridDim.x;
  const int cta = (int)blockIdx.x;
  // The CTAs are dealt round-robin to the row groups: the GEMM engine of CTA c only runs the steps of group c % G, as
  // member c / G of that group's team.  The groups' chains of dependent steps then advance side by side on disjoint
  // CTAs instead of queueing in every CTA's single engine.  (The attention engine serves every group.)
  const int my_g = cta % p.G, ci = cta / p.G;
  const int nct = p.team;                          // CTAs of the team with GEMM work; the others only run attention
  const bool tracing = p.trace != nullptr && cta == p.trace_cta;
  if (tracing && threadIdx.x == 0) p.trace[MG_TRACE_WORDS - 1] = clock64();

  if (warp >= 4 && warp < MG_W_ACONS) {
   asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(MG_REGS_LIGHT));
   if (warp == MG_W_TMA) {
    // ===================== weight producer: TMA tiles in program order, as far ahead as the ring allows ==========
    if (lane == 0) {
      uint32_t kbc = 0;
      for (int si = p.step_begin; si < p.step_end; ++si) {
        const MegaStep st = p.steps[si];
        if (st.kind != MG_GEMM || st.g != my_g || ci >= nct) continue;
        const int ph = st.p, kpu = p.kpu[ph], ntl = p.ntiles[ph];
        const CUtensorMap* map = p.maps + (size_t)st.l * 4 + ph;
        // the per-column vectors the closing epilogue adds are read once per token: cold in L2.  Ask for them now,
        // the epilogue warps need them a few microseconds from now.
        const MegaLayerW* lwp = p.lw + st.l;
        const float* pf0 = ph == 1 ? lwp->b_proj : (ph == 2 ? lwp->b_fc : (ph == 3 ? lwp->b_proj2 : nullptr));
        for (int u = (ci + nct - p.rot[ph] % nct) % nct; u < p.nunits[ph]; u += nct) {
          const int t = u % ntl, z = u / ntl;
          if (pf0)
            for (int q = 0; q < BM; q += 32) {
              asm volatile("prefetch.global.L2 [%0];" ::"l"(pf0 + t * BM + q));
              if (ph == 2) {
                asm volatile("prefetch.global.L2 [%0];" ::"l"(lwp->s_fc + t * BM + q));
                asm volatile("prefetch.global.L2 [%0];" ::"l"(lwp->c_fc + t * BM + q));
              }
            }
          for (int kb = 0; kb < kpu; ++kb, ++kbc) {
            const int s = kbc % NST;
            if (kbc >= NST) mg_wait(p.err, &empty[s], ((kbc / NST) & 1) ^ 1, 10);
            mbar_expect_tx(&fullA[s], SM::A_BYTES);
            tma_load_2d(tiles + s * SM::STAGE, map, &fullA[s], (z * kpu + kb) * BK, t * BM);
          }
        }
      }
    }
   } else if (warp == MG_W_ACT) {
    // ===================== activation producer: the UMMA N operand of every unit, once the step's inputs exist ======
    if (lane == 0) {
      uint32_t kbc = 0;
      for (int si = p.step_begin; si < p.step_end; ++si) {
        const MegaStep st = p.steps[si];
        if (st.kind != MG_GEMM || st.g != my_g || ci >= nct) continue;
        const int ph = st.p, kpu = p.kpu[ph], ntl = p.ntiles[ph];
        const int u_first = (ci + nct - p.rot[ph] % nct) % nct;
        if (u_first >= p.nunits[ph]) continue;
        if (st.in_flag >= 0) wait_counter(p.err, MG_FLAG(st.in_flag), st.in_target, si, st.in_flag);
        // the rows were written with ordinary stores by other SMs; TMA reads them through the async proxy
        asm volatile("fence.proxy.async.global;" ::: "memory");
        if (tracing && si - p.step_begin < MG_TRACE_MAX_STEPS)
          p.trace[(si - p.step_begin) * MG_TRACE_STEP_SLOTS + 0] = clock64();
        const CUtensorMap* map = p.maps + (size_t)p.L * 4 + (ph == 1 ? 1 : (ph == 3 ? 2 : 0));
        const int row0 = st.g * p.rpg;
        for (int u = u_first; u < p.nunits[ph]; u += nct) {
          const int z = u / ntl;
          for (int kb = 0; kb < kpu; ++kb, ++kbc) {
            const int s = kbc % NST;
            if (kbc >= NST) mg_wait(p.err, &empty[s], ((kbc / NST) & 1) ^ 1, 17);
            mbar_expect_tx(&fullB[s], SM::B_BYTES);
            tma_load_2d(tiles + s * SM::STAGE + SM::A_BYTES, map, &fullB[s], (z * kpu + kb) * BK, row0);
          }
        }
      }
    }
   } else if (warp == MG_W_MMA) {
    // ===================== MMA issuer ================================================================================
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(BM, GN);
      uint32_t kbc = 0, uc = 0;
      for (int si = p.step_begin; si < p.step_end; ++si) {
        const MegaStep st = p.steps[si];
        if (st.kind != MG_GEMM || st.g != my_g || ci >= nct) continue;
        const int ph = st.p, kpu = p.kpu[ph];
        for (int u = (ci + nct - p.rot[ph] % nct) % nct; u < p.nunits[ph]; u += nct, ++uc) {
          const uint32_t buf = uc & 1;
          if (uc >= 2) mg_wait(p.err, &tmem_empty[buf], ((uc >> 1) & 1) ^ 1, 11);
          tc_fence_after();
          const uint32_t dtm = tmem_base + buf * TSTRIDE;
          for (int kb = 0; kb < kpu; ++kb, ++kbc) {
            const int s = kbc % NST;
            mg_wait(p.err, &fullA[s], (kbc / NST) & 1, 12);
            mg_wait(p.err, &fullB[s], (kbc / NST) & 1, 13);
            tc_fence_after();
            const uint32_t sa = smem_u32(tiles + s * SM::STAGE);
            const uint32_t sb = sa + SM::A_BYTES;
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k)
              tc_mma_bf16(dtm, umma_desc_sw128(sa + k * UMMA_K * 2), umma_desc_sw128(sb + k * UMMA_K * 2), idesc,
                          (kb | k) != 0);
            tc_commit(&empty[s]);
          }
          tc_commit(&tmem_full[buf]);
        }
      }
    }
   } else if (warp >= MG_W_APROD && warp < MG_W_APROD + MG_NSLOT) {
    // ===================== attention: K/V stream producer of one slot ================================================
    if (attn_on && lane == 0) {
      AttnSlot& S = slots[warp - MG_W_APROD];
      const int BH = p.B * p.H;
      uint32_t cc = 0, un = 0;
      while (true) {
        const int q = p.attn_begin + atomicAdd(MG_FLAG(0), 1);
        const int di = un & 1;
        if (un >= 2) mg_wait(p.err, &S.dempty[di], ((un >> 1) & 1) ^ 1, 20);
        S.desc[di] = q < p.attn_end ? q : -1;
        mbar_arrive(&S.dfull[di]);
        ++un;
        if (q >= p.attn_end) break;
        const int l = q / BH, rem = q - l * BH, b = rem / p.H, h = rem - b * p.H;
        const int kp = p.npad ? p.npad[b] : 0;
        const int nkeys = len - 1 - kp;  // rows of earlier tokens: final before this launch
        const int nchunks = (nkeys + MG_CHUNK - 1) / MG_CHUNK;
        const size_t base = ((size_t)b * p.H + h) * p.Smax * 64 + (size_t)kp * 64;
        const bf16* kc = p.lw[l].kc + base;
        const bf16* vc = p.lw[l].vc + base;
        // The ring holds two chunks; the unit usually waits for its q/k/v long enough (the c_attn of its layer is still
        // being computed, HBM is idle) to pull the rest of its rows into L2, from where the ring refills at L2 latency.
        if (p.l2_prefetch)
          for (int c = MG_RING; c < nchunks; ++c) {
            const uint32_t bytes = (uint32_t)min(MG_CHUNK, nkeys - c * MG_CHUNK) * 128;
            asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(kc + (size_t)c * MG_CHUNK * 64), "r"(bytes) : "memory");
            asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(vc + (size_t)c * MG_CHUNK * 64), "r"(bytes) : "memory");
          }
        for (int c = 0; c < nchunks; ++c, ++cc) {
          const int s = cc % MG_RING;
          if (cc >= MG_RING) mg_wait(p.err, &S.empty[s], ((cc / MG_RING) & 1) ^ 1, 21);
          const uint32_t bytes = (uint32_t)min(MG_CHUNK, nkeys - c * MG_CHUNK) * 128;
          mbar_expect_tx(&S.full[s], 2 * bytes);
          asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                           smem_u32(&S.sK[s][0])), "l"(kc + (size_t)c * MG_CHUNK * 64), "r"(bytes), "r"(smem_u32(&S.full[s]))
                       : "memory");
          asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                           smem_u32(&S.sV[s][0])), "l"(vc + (size_t)c * MG_CHUNK * 64), "r"(bytes), "r"(smem_u32(&S.full[s]))
                       : "memory");
        }
      }
    }
   }
  } else if (warp < 4) {
    // ===================== epilogue + split-K closer (128 threads) ===================================================
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(MG_REGS_EPI));
    const int t = threadIdx.x, ew = warp, quad = warp & 3;
    const int B = p.B, d = p.d;
    const int nstt = d / BM;
    const float inv_nst = 1.0f / (float)nstt, inv_d = 1.0f / (float)d;
    uint32_t uc = 0;
    for (int si = p.step_begin; si < p.step_end; ++si) {
      const MegaStep st = p.steps[si];
      if (st.g != my_g || ci >= nct) continue;
      const int g0 = st.g * p.rpg;
      const int rows_g = min(p.rpg, B - g0);
      if (st.kind == MG_EMBED) {
        // x[m,:] = wte[id] + wpe[pos] (modeling_gpt2.py:577-585), its bf16 copy (the c_attn operand) and the
        // per-128-column (mean, M2) of the row
        if (ci >= rows_g) continue;
        int mine = 0;
        for (int j = ci; j < rows_g; j += nct) {
          if ((mine++ & 3) != ew) continue;
          const int m = g0 + j;
          const int id = p.ids[(size_t)m * p.ids_ld + len - 1];
          const int pe = p.npad ? max(len - 1 - p.npad[m], 0) : len - 1;
          const bf16* e = p.wte + (size_t)id * d;
          const float* pp = p.wpe + (size_t)pe * d;
          for (int tile = 0; tile < nstt; ++tile) {
            const int i = tile * BM + lane * 4;
            const float4 a = load4(e + i), c = load4(pp + i);
            const float4 v = make_float4(a.x + c.x, a.y + c.y, a.z + c.z, a.w + c.w);
            *reinterpret_cast<float4*>(p.x + (size_t)m * d + i) = v;
            *reinterpret_cast<uint2*>(p.xb + (size_t)m * d + i) = pack_bf16x4(v);
            const float mean_t = warp_sum((v.x + v.y) + (v.z + v.w)) * (1.0f / BM);
            const float a0 = v.x - mean_t, a1 = v.y - mean_t, a2 = v.z - mean_t, a3 = v.w - mean_t;
            const float m2_t = warp_sum((a0 * a0 + a1 * a1) + (a2 * a2 + a3 * a3));
            if (lane == 0) *reinterpret_cast<float2*>(p.stats + ((size_t)tile * B + m) * 2) = make_float2(mean_t, m2_t);
          }
        }
        bar_named(1, 128);
        if (t == 0 && mine > 0) red_release(MG_FLAG(st.out_flag), mine);
        continue;
      }
      if (st.kind == MG_FINAL) {
        // ln_f of every row of the group -> bf16 operand of the LM head (logits_to_keep = 1: the only position)
        if (ci >= rows_g) continue;
        if (t == 0) wait_counter(p.err, MG_FLAG(st.in_flag), st.in_target, si, st.in_flag);
        bar_named(1, 128);
        int mine = 0;
        for (int j = ci; j < rows_g; j += nct) {
          if ((mine++ & 3) != ew) continue;
          const int m = g0 + j;
          const int li = lane & 15;
          const float2 tv = li < nstt ? __ldcg(reinterpret_cast<const float2*>(p.stats) + (size_t)li * B + m)
                                      : make_float2(0.f, 0.f);
          const float2 mr = row_mean_rstd(tv, li < nstt, inv_nst, inv_d, p.eps);
          for (int q = 0; q < nstt; ++q) {
            const int i = q * BM + lane * 4;
            const float4 v = __ldcg(reinterpret_cast<const float4*>(p.x + (size_t)m * d + i));
            const float4 gg = load4(p.lnf_g + i), bb = load4(p.lnf_b + i);
            const float4 o = make_float4((v.x - mr.x) * mr.y * gg.x + bb.x, (v.y - mr.x) * mr.y * gg.y + bb.y,
                                         (v.z - mr.x) * mr.y * gg.z + bb.z, (v.w - mr.x) * mr.y * gg.w + bb.w);
            *reinterpret_cast<uint2*>(p.xn_last + (size_t)m * d + i) = pack_bf16x4(o);
          }
        }
        continue;
      }
      // ----- GEMM step: phase ph of layer st.l for rows [g0, g0 + rows_g).  The operands arrive by TMA (other warps);
      // this group drains the accumulators and closes the split-K sums.
      // Everything the inner loops index by phase is copied into registers once (pin()): left to itself the compiler
      // re-reads the parameter arrays with a dynamically indexed constant load in front of every store.
      const int ph = st.p;
      const int ntl = pin(p.ntiles[ph]), ks = pin(p.ks[ph]), N = pin(p.nout[ph]), nun = pin(p.nunits[ph]);
      const int u_first = (ci + nct - p.rot[ph] % nct) % nct;
      if (u_first >= nun) continue;
      float* part = pin(p.partial[ph]);
      int* tcnt = pin(MG_FLAG(p.tc0 + ((st.l * p.G + st.g) * 3 + ph - 1) * p.maxtiles));
      const size_t zstride = (size_t)B * N;  // one K split of the partial buffer
      for (int u = u_first; u < nun; u += nct, ++uc) {
        const int tl = u % ntl, z = u / ntl;
        // ---- accumulator -> fp32 partial tile in L2: partial[z][row][N]
        const uint32_t buf = uc & 1;
        mg_wait(p.err, &tmem_full[buf], (uc >> 1) & 1, 16);
        tc_fence_after();
        MG_TRACE_STEP(si, 2);
        const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + buf * TSTRIDE;
        float* dst = part + (size_t)z * zstride + (size_t)g0 * N + tl * BM + quad * 32 + lane;
#pragma unroll
        for (int c0 = 0; c0 < GN; c0 += 32) {
          float v[32];
          tc_ld32(taddr + c0, v);
          if (rows_g >= c0 + 32) {
#pragma unroll
            for (int i = 0; i < 32; ++i) dst[(size_t)(c0 + i) * N] = v[i];
          } else {
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (c0 + i < GN && c0 + i < rows_g) dst[(size_t)(c0 + i) * N] = v[i];
          }
        }
        MG_TRACE_STEP(si, 6);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tmem_empty[buf]);
        bar_named(1, 128);
        if (t == 0) red_release(ph == 0 ? MG_FLAG(st.out_flag) : tcnt + (size_t)tl * MG_FLAG_STRIDE, 1);
        MG_TRACE_STEP(si, 3);
      }
      if (ph == 0) continue;  // c_attn partials are summed by the attention unit of each head
      // ---- close the split-K sums: this CTA's units reduce the rows r == z (mod ks) of their tile
      const MegaLayerW* lwp = p.lw + st.l;
      const float* bias = pin(ph == 1 ? lwp->b_proj : (ph == 2 ? lwp->b_fc : lwp->b_proj2));
      float* xg = pin(p.x + (size_t)g0 * d);  // rows of this group
      const int li = lane & 15;
      for (int u = u_first; u < nun; u += nct) {
        const int tl = u % ntl, z = u / ntl;
        // whatever does not depend on the other splits is requested before the wait
        const int n = tl * BM + lane * 4;
        const float4 bias4 = load4(bias + n);
        float4 s4 = make_float4(0.f, 0.f, 0.f, 0.f), c4 = s4;
        if (ph == 2) { s4 = load4(lwp->s_fc + n); c4 = load4(lwp->c_fc + n); }
        const int r_first = z + ew * ks, r_step = 4 * ks;
        if (t == 0) wait_counter(p.err, tcnt + (size_t)tl * MG_FLAG_STRIDE, ks, si, -tl - 1);
        bar_named(1, 128);
        MG_TRACE_STEP(si, 4);
        // One warp owns the rows r_first, r_first + 4 ks, ...  Under load an L2 round trip costs 1-2 thousand cycles,
        // so every load of a batch of rows (RPB rows x up to KSB splits, 12 vectors in all, plus the residual or the
        // LayerNorm statistics of each row) is in flight before the first is used.
        auto close_rows = [&](auto rpb_c, auto ksb_c) {
          constexpr int RPB = decltype(rpb_c)::value, KSB = decltype(ksb_c)::value;
          for (int rb = r_first; rb < rows_g; rb += RPB * r_step) {
            float4 pv[RPB][KSB], res[RPB];
            float2 tv[RPB];
            const float* pr = part + (size_t)(g0 + rb) * N + n;
#pragma unroll
            for (int j = 0; j < RPB; ++j) {
              const bool ok = rb + j * r_step < rows_g;
#pragma unroll
              for (int zz = 0; zz < KSB; ++zz)
                pv[j][zz] = (ok && zz < ks) ? __ldcg(reinterpret_cast<const float4*>(pr + (size_t)(j * r_step) * N + (size_t)zz * zstride))
                                            : make_float4(0.f, 0.f, 0.f, 0.f);
              res[j] = make_float4(0.f, 0.f, 0.f, 0.f);
              tv[j] = make_float2(0.f, 0.f);
              if (ok) {
                if (ph != 2) res[j] = __ldcg(reinterpret_cast<const float4*>(xg + (size_t)(rb + j * r_step) * N + n));
                else if (li < nstt) tv[j] = __ldcg(reinterpret_cast<const float2*>(p.stats) + (size_t)li * B + g0 + rb + j * r_step);
              }
            }
#pragma unroll
            for (int j = 0; j < RPB; ++j) {
              const int r = rb + j * r_step, m = g0 + r;
              if (r >= rows_g) break;
              float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
              for (int zz = 0; zz < KSB; ++zz) {  // fixed order: deterministic
                acc.x += pv[j][zz].x; acc.y += pv[j][zz].y; acc.z += pv[j][zz].z; acc.w += pv[j][zz].w;
              }
              if (j == 0 && rb == r_first && tracing && t == 0 && si - p.step_begin < MG_TRACE_MAX_STEPS && acc.x != 1234.5f)
                p.trace[(si - p.step_begin) * MG_TRACE_STEP_SLOTS + 1] = clock64();
              if (ph == 2) {
                // ln_2 folded into c_fc: LayerNorm(x) W^T = rstd (x W'^T - mean s) + c, then bias + gelu_new
                const float2 mr = row_mean_rstd(tv[j], li < nstt, inv_nst, inv_d, p.eps);
                acc.x = mr.y * (acc.x - mr.x * s4.x) + c4.x + bias4.x;
                acc.y = mr.y * (acc.y - mr.x * s4.y) + c4.y + bias4.y;
                acc.z = mr.y * (acc.z - mr.x * s4.z) + c4.z + bias4.z;
                acc.w = mr.y * (acc.w - mr.x * s4.w) + c4.w + bias4.w;
                const float4 o = make_float4(gelu_tanh_fast(acc.x), gelu_tanh_fast(acc.y), gelu_tanh_fast(acc.z),
                                             gelu_tanh_fast(acc.w));
                *reinterpret_cast<uint2*>(p.hbuf + (size_t)m * N + n) = pack_bf16x4(o);
              } else {
                acc.x += bias4.x + res[j].x; acc.y += bias4.y + res[j].y; acc.z += bias4.z + res[j].z; acc.w += bias4.w + res[j].w;
                *reinterpret_cast<float4*>(xg + (size_t)r * N + n) = acc;
                *reinterpret_cast<uint2*>(p.xb + (size_t)m * N + n) = pack_bf16x4(acc);  // operand of the next linear
                const float mean_t = warp_sum((acc.x + acc.y) + (acc.z + acc.w)) * (1.0f / BM);
                const float a0 = acc.x - mean_t, a1 = acc.y - mean_t, a2 = acc.z - mean_t, a3 = acc.w - mean_t;
                const float m2_t = warp_sum((a0 * a0 + a1 * a1) + (a2 * a2 + a3 * a3));
                if (lane == 0) *reinterpret_cast<float2*>(p.stats + ((size_t)tl * B + m) * 2) = make_float2(mean_t, m2_t);
              }
            }
          }
        };
        using std::integral_constant;
        if (ks <= 3) close_rows(integral_constant<int, 4>{}, integral_constant<int, 3>{});
        else if (ks <= 4) close_rows(integral_constant<int, 3>{}, integral_constant<int, 4>{});
        else if (ks <= 6) close_rows(integral_constant<int, 2>{}, integral_constant<int, 6>{});
        else close_rows(integral_constant<int, 1>{}, integral_constant<int, 12>{});  // the plan never splits deeper than 12
        MG_TRACE_STEP(si, 7);
        bar_named(1, 128);
        if (t == 0) red_release(MG_FLAG(st.out_flag), 1);
        MG_TRACE_STEP(si, 5);
      }
    }
  } else if (attn_on && warp >= MG_W_ACONS) {
    // ===================== attention: the four consumer warps of one slot ============================================
    constexpr int EPL = 8, LPR = 8, KPW = 4, NG = 16, UNR = MG_CHUNK / NG;
    const int cwarp = warp - MG_W_ACONS;
    const int slot = cwarp >> 2, cw = cwarp & 3;
    const int ct = cw * 32 + lane;  // thread within the slot
    AttnSlot& S = slots[slot];
    const int BH = p.B * p.H, d = p.d;
    const int grp = cw * KPW + lane / LPR;  // lane group: one key at a time
    const int e0 = (lane % LPR) * EPL;
    const int ks = p.ks[0];
    const int nstt = d / BM, li = lane & 15;
    const float inv_nst = 1.0f / (float)nstt, inv_d = 1.0f / (float)d;
    uint32_t cc = 0, un = 0;
    int tun = -1;
    while (true) {
      const int di = un & 1;
      mg_wait(p.err, &S.dfull[di], (un >> 1) & 1, 22);
      const int q = S.desc[di];
      __syncwarp();
      if (lane == 0) mbar_arrive(&S.dempty[di]);
      ++un;
      if (q < 0) break;
      ++tun;
      MG_TRACE_UNIT(0);
      if (tracing && slot == 0 && ct == 0 && tun < MG_TRACE_MAX_UNITS)
        p.trace[MG_TRACE_MAX_STEPS * MG_TRACE_STEP_SLOTS + tun * MG_TRACE_UNIT_SLOTS + 5] = q;
      const int l = q / BH, rem = q - l * BH, b = rem / p.H, h = rem - b * p.H;
      const int g = b / p.rpg;
      const MegaLayerW* lwp = p.lw + l;
      const int kp = p.npad ? p.npad[b] : 0;
      const int nkeys = len - 1 - kp;
      const int nchunks = (nkeys + MG_CHUNK - 1) / MG_CHUNK;
      const size_t base = ((size_t)b * p.H + h) * p.Smax * 64;
      // per-column constants of this head's q, k, v: static, requested before the wait
      float sv[2], cv[2];
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        const int e = ct + r * 128;
        sv[r] = cv[r] = 0.f;
        if (e < 192) {
          const int n = (e >> 6) * d + h * 64 + (e & 63);
          sv[r] = lwp->s_qkv[n];
          cv[r] = lwp->c_qkv[n] + lwp->b_qkv[n];
        }
      }
      if (p.attn_wait && ct == 0) wait_counter_backoff(p.err, MG_FLAG(p.flag_qkv0 + l * p.G + g), p.nunits[0], -2, q);
      bar_named(2 + slot, 128);
      MG_TRACE_UNIT(1);
      {
        // q, k, v of this head: ln_1 folded into c_attn (modeling_gpt2.py:185-196): rstd (sum of the split-K partials
        // - mean s) + c + bias, the partials summed in a fixed order
        const float2 tv = li < nstt ? __ldcg(reinterpret_cast<const float2*>(p.stats) + (size_t)li * p.B + b)
                                    : make_float2(0.f, 0.f);
        const float* part = p.partial[0];
        float accv[2];
#pragma unroll
        for (int r = 0; r < 2; ++r) {
          const int e = ct + r * 128;
          accv[r] = 0.f;
          if (e < 192) {
            const int n = (e >> 6) * d + h * 64 + (e & 63);
            float acc = 0.f;
            for (int zz = 0; zz < ks; ++zz) acc += __ldcg(part + ((size_t)zz * p.B + b) * (3 * d) + n);
            accv[r] = acc;
          }
        }
        const float2 mr = row_mean_rstd(tv, li < nstt, inv_nst, inv_d, p.eps);
#pragma unroll
        for (int r = 0; r < 2; ++r) {
          const int e = ct + r * 128;
          if (e < 192) {
            const float val = mr.y * (accv[r] - mr.x * sv[r]) + cv[r];
            const int which = e >> 6, i = e & 63;
            if (which == 0) {
              S.s_new[0][i] = val;
            } else {
              const bf16 stv = __float2bfloat16_rn(val);
              (which == 1 ? lwp->kc : lwp->vc)[base + (size_t)(len - 1) * 64 + i] = stv;  // cache append (no torch.cat)
              S.s_new[which][i] = __bfloat162float(stv);
            }
          }
        }
      }
      bar_named(2 + slot, 128);
      MG_TRACE_UNIT(2);
      float qv[EPL];
#pragma unroll
      for (int i = 0; i < EPL; ++i) qv[i] = S.s_new[0][e0 + i] * 0.125f;
      float m = -CUDART_INF_F, lsum = 0.f, o[EPL];
#pragma unroll
      for (int i = 0; i < EPL; ++i) o[i] = 0.f;
      for (int c = 0; c < nchunks; ++c, ++cc) {
        const int s = cc % MG_RING;
        mg_wait(p.err, &S.full[s], (cc / MG_RING) & 1, 23);
        const int valid_rows = min(MG_CHUNK, nkeys - c * MG_CHUNK);
        float kf[UNR][EPL], vf[UNR][EPL], sc[UNR];
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
          const int kk = u * NG + grp;
          const int row = kk < valid_rows ? kk : 0;
          unpack_bf16x8(*reinterpret_cast<const uint4*>(&S.sK[s][row * 64 + e0]), kf[u]);
          unpack_bf16x8(*reinterpret_cast<const uint4*>(&S.sV[s][row * 64 + e0]), vf[u]);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&S.empty[s]);
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
          float dd = 0.f;
#pragma unroll
          for (int i = 0; i < EPL; ++i) dd = fmaf(qv[i], kf[u][i], dd);
#pragma unroll
          for (int off = LPR / 2; off > 0; off >>= 1) dd += __shfl_xor_sync(0xffffffffu, dd, off);
          sc[u] = (u * NG + grp < valid_rows) ? dd : -CUDART_INF_F;
        }
        float mn = m;
#pragma unroll
        for (int u = 0; u < UNR; ++u) mn = fmaxf(mn, sc[u]);
        if (mn == -CUDART_INF_F) continue;
        const float corr = __expf(m - mn);
        lsum *= corr;
#pragma unroll
        for (int i = 0; i < EPL; ++i) o[i] *= corr;
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
          const float pr = __expf(sc[u] - mn);
          lsum += pr;
#pragma unroll
          for (int i = 0; i < EPL; ++i) o[i] = fmaf(pr, vf[u][i], o[i]);
        }
        m = mn;
      }
      MG_TRACE_UNIT(3);
      {  // the current token's own key/value, folded into lane group 0
        float dd = 0.f;
#pragma unroll
        for (int i = 0; i < EPL; ++i) dd = fmaf(qv[i], S.s_new[1][e0 + i], dd);
#pragma unroll
        for (int off = LPR / 2; off > 0; off >>= 1) dd += __shfl_xor_sync(0xffffffffu, dd, off);
        if (grp == 0) {
          const float mn = fmaxf(m, dd);
          const float corr = __expf(m - mn), pr = __expf(dd - mn);
          lsum = lsum * corr + pr;
#pragma unroll
          for (int i = 0; i < EPL; ++i) o[i] = fmaf(pr, S.s_new[2][e0 + i], o[i] * corr);
          m = mn;
        }
      }
      // merge the warp's four lane groups with shuffles, then the four warps through shared memory
#pragma unroll
      for (int off = LPR; off < 32; off <<= 1) {
        const float m2 = __shfl_xor_sync(0xffffffffu, m, off);
        const float l2 = __shfl_xor_sync(0xffffffffu, lsum, off);
        const float mn = fmaxf(m, m2);
        const float wa = (m == -CUDART_INF_F) ? 0.f : __expf(m - mn);
        const float wb = (m2 == -CUDART_INF_F) ? 0.f : __expf(m2 - mn);
        lsum = lsum * wa + l2 * wb;
#pragma unroll
        for (int i = 0; i < EPL; ++i) {
          const float o2 = __shfl_xor_sync(0xffffffffu, o[i], off);
          o[i] = o[i] * wa + o2 * wb;
        }
        m = mn;
      }
      if (lane < LPR) {
#pragma unroll
        for (int i = 0; i < EPL; ++i) S.w_o[cw][e0 + i] = o[i];
        if (lane == 0) { S.w_m[cw] = m; S.w_l[cw] = lsum; }
      }
      bar_named(2 + slot, 128);
      if (ct < 64) {
        float M = -CUDART_INF_F;
#pragma unroll
        for (int w = 0; w < 4; ++w) M = fmaxf(M, S.w_m[w]);
        float num = 0.f, den = 0.f;
#pragma unroll
        for (int w = 0; w < 4; ++w) {
          const float wgt = (S.w_m[w] == -CUDART_INF_F) ? 0.f : __expf(S.w_m[w] - M);
          num = fmaf(wgt, S.w_o[w][ct], num);
          den = fmaf(wgt, S.w_l[w], den);
        }
        p.att[((size_t)b * p.H + h) * 64 + ct] = __float2bfloat16_rn(num / den);
      }
      bar_named(2 + slot, 128);
      if (ct == 0) red_release(MG_FLAG(p.flag_att0 + l * p.G + g), 1);
      MG_TRACE_UNIT(4);
    }
  }

  pdl_launch_dependents();
  tc_fence_before();
  __syncthreads();
  if (tracing && threadIdx.x == 0) p.trace[MG_TRACE_WORDS - 2] = clock64();
  if (warp == MG_W_MMA) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
  }
  // the last CTA to finish clears every counter for the next launch (nothing reads them after this point)
  __shared__ int s_last;
  if (threadIdx.x == 0) {
    __threadfence();
    s_last = (atomicAdd(MG_FLAG(1), 1) == nctas - 1);
  }
  __syncthreads();
  if (s_last) {
    for (int i = threadIdx.x; i < p.nflags; i += MG_THREADS) *MG_FLAG(i) = 0;
  }
}

template <int GN>
int launch_mega(const MegaParams& p, int nctas, cudaStream_t st) {
  auto kern = decode_mega_kernel<GN>;
  const int smem = MegaSmem<GN>::TOTAL;
  TRY(tutor_set_max_smem(reinterpret_cast<const void*>(kern), smem));
  CUDA_TRY(tutor_launch(kern, dim3(nctas), dim3(MG_THREADS), (size_t)smem, st, p));
  KERNEL_LAUNCHED();
  return TUTOR_OK;
}

}  // namespace

int k_decode_mega_gn(int rows_per_group) { return rows_per_group <= 16 ? 16 : rows_per_group <= 32 ? 32 : 64; }

int k_decode_mega(const MegaParams& p, int gn, int nctas, cudaStream_t st) {
  ARG_CHECK(p.d % BM == 0 && p.d <= MG_MAX_STAT_TILES * BM, "decode_mega: n_embd %d must be a multiple of 128, at most %d",
            p.d, MG_MAX_STAT_TILES * BM);
  ARG_CHECK(p.rpg >= 1 && p.rpg <= gn, "decode_mega: %d rows per group exceed the %d-row tile", p.rpg, gn);
  if (gn == 16) return launch_mega<16>(p, nctas, st);
  if (gn == 32) return launch_mega<32>(p, nctas, st);
  ARG_CHECK(gn == 64, "decode_mega: row tile %d", gn);
  return launch_mega<64>(p, nctas, st);
}
