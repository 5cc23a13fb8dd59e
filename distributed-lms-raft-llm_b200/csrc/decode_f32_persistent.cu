// The transformer stack of a parity-mode (fp32) decode step for one or two sequences as ONE cooperative kernel
// (BASELINE configs[0]: GPT-2 small, batch 1 - the reference's own configuration, REF/tutoring_server.py:21-29).
//   embedding -> L x [ ln_1 + c_attn -> attention -> c_proj + residual -> ln_2 + c_fc + gelu_new -> mlp c_proj + residual ]
// (HF/models/gpt2/modeling_gpt2.py:262-309, :577-607).  ln_f and the LM head follow as their own launch.
//
// As separate launches this was 63 dependent kernels of ~5 us each for 494 MB of weights (0.32 ms per forward, 22 % of the
// HBM roofline): a latency chain.  Here one CTA per SM stays resident and the five phases of a layer are separated by a
// grid-wide barrier (one release-add + relaxed polls on a counter in L2).
//
// Weights never change, so they do not have to wait for anything: a producer warp per CTA streams the weight rows its CTA
// will need - in the order it will need them - with cp.async.bulk into a 56-row ring in shared memory, up to a whole layer
// ahead of the consumers and already during the previous token's LM head (before the programmatic-launch wait).  What is
// left on the critical path of a phase is the L2 round trip for the phase's input vector, a warp reduction and the
// barrier itself.  (The first version requested the next phase's rows into registers between two barriers: the HBM burst
// of every phase then sat on the critical path and slowed the barrier's own L2 traffic - 4.8 us per phase.)
//
// Exact arithmetic throughout (fp32 FMA, expf, tanhf; no TF32): the greedy tokens must stay bit-identical to the oracle.
// Work split: a GEMV phase with N outputs and K = C*d inputs is N*C units (row, d-wide chunk of K); unit u runs on warp
// u mod (warps in the grid), a warp owns at most two units; the chunks of one row sit on consecutive warps of one CTA and
// meet in shared memory in a fixed order (deterministic).  Attention: (sequence, head, key range) per CTA, partial softmax
// states merged by the last CTA of a head to arrive.
#include <math_constants.h>

#include "decode_f32_persistent.h"
#include "tc_ptx.cuh"

namespace {

constexpr int PF_WARPS = 16, PF_CONSUMERS = PF_WARPS * 32;  // consumer warps 0..15
constexpr int PF_THREADS = PF_CONSUMERS + 32;                // + the producer warp
constexpr int PF_MAXU = 2;                                   // units per warp and phase
constexpr int PF_SLOTS = 56;                                 // weight rows in the shared-memory ring

__device__ __forceinline__ int ld_relaxed(const int* p) {
  int v;
  asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release(int* p, int v) {
  asm volatile("red.release.gpu.global.add.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void consumer_sync() { asm volatile("bar.sync 1, %0;" ::"n"(PF_CONSUMERS) : "memory"); }

// A wait of more than ~2 s is a bug: it records itself in err[] and falls through instead of hanging the GPU.
__device__ __forceinline__ void record_timeout(int* err, int code, int which, int target) {
  if (atomicCAS(err, 0, code) == 0) { err[1] = which; err[2] = target; err[3] = (int)blockIdx.x; err[4] = (int)threadIdx.x; }
}
__device__ __forceinline__ void ring_wait(uint64_t* bar, uint32_t parity, int* err, int code, int which) {
  if (tcptx::mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  while (!tcptx::mbar_try_wait(bar, parity)) {
    const long long dt = clock64() - t0;
    if ((dt > (1LL << 20) && *reinterpret_cast<volatile int*>(err) != 0) || dt > 4000000000LL) {
      if (dt > 4000000000LL) record_timeout(err, code, which, (int)parity);
      return;
    }
  }
}

// Every CTA arrives, every CTA waits for `target` arrivals in all (consumer warps only).  Data written before the barrier
// by other CTAs is read after it with ld.global.cg only (L2), so no acquire fence - which would flush the SM's L1 - is
// needed (decode_mega.cu).
__device__ __forceinline__ void grid_barrier(int* count, int target, int* err, int which, long long* tr = nullptr) {
  if (tr && threadIdx.x == 0) tr[4 * which - 3] = clock64();  // thread 0 has finished its part of the phase
  consumer_sync();
  if (threadIdx.x == 0) {
    if (tr) tr[4 * which - 2] = clock64();                    // ... and so have the other warps of the CTA
    red_release(count, 1);
    if (tr) tr[4 * which - 1] = clock64();
    if (ld_relaxed(count) < target) {
      const long long t0 = clock64();
      while (ld_relaxed(count) < target) {
        const long long dt = clock64() - t0;
        if ((dt > (1LL << 20) && *reinterpret_cast<volatile int*>(err) != 0) || dt > 4000000000LL) {
          if (dt > 4000000000LL) record_timeout(err, 3, which, target);
          break;
        }
      }
    }
    if (tr) tr[4 * which] = clock64();                        // every CTA has arrived
  }
  consumer_sync();
}

enum { PH_QKV = 0, PH_PROJ = 1, PH_FC = 2, PH_PROJ2 = 3 };

// units of round j (unit u = cta*16 + warp + j*GW) that exist in a phase of `nunits` units: a prefix of the CTA's warps
__device__ __forceinline__ int live_units(int nunits, int j, int cta, int GW) {
  const int c = nunits - j * GW - cta * PF_WARPS;
  return c < 0 ? 0 : (c > PF_WARPS ? PF_WARPS : c);
}

template <int NV, int MB>
__global__ void __launch_bounds__(PF_THREADS, 1) decode_f32_persistent_kernel(const __grid_constant__ PersistF32Params p) {
  constexpr int D = NV * 128;
  constexpr uint32_t ROW_BYTES = D * sizeof(float);
  extern __shared__ __align__(16) float smem[];
  constexpr int SLOT = D + 4;                 // a weight row (one d-wide chunk) + the aligned 16 bytes holding its bias
  float* ring = smem;                         // [PF_SLOTS][SLOT], filled by the producer
  float* s_ln = ring + PF_SLOTS * SLOT;       // [2][4][D]: ln_1 / ln_2 gamma, beta of the current and the next layer
  float* s_in = s_ln + 2 * 4 * D;             // [MB][4 D]: the input vector(s) of the current phase
  float* s_part = s_in + MB * 4 * D;          // [PF_WARPS][MB]: chunk sums of mlp c_proj
  float* s_att = s_part + PF_WARPS * MB;      // attention: per-warp (m, l, o[64])
  uint64_t* full = reinterpret_cast<uint64_t*>(s_att + PF_WARPS * 66);  // [PF_SLOTS] row has landed
  uint64_t* empty = full + PF_SLOTS;                                    // [PF_SLOTS] row has been consumed
  // the per-layer pointer table, copied once: after 494 MB of weights have gone through L2 every first touch of a small
  // array (this table, a bias) is an HBM round trip, and inside a phase nothing hides it
  PersistF32Layer* s_layers = reinterpret_cast<PersistF32Layer*>(empty + PF_SLOTS);  // [L]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ncta = (int)gridDim.x, cta = (int)blockIdx.x;
  const int GW = ncta * PF_WARPS, gw = cta * PF_WARPS + warp;
  const int H = D / 64;

  if (threadIdx.x == 0) {
    for (int s = 0; s < PF_SLOTS; ++s) { tcptx::mbar_init(&full[s], 1); tcptx::mbar_init(&empty[s], 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  for (int i = threadIdx.x; i < p.L * (int)(sizeof(PersistF32Layer) / 8); i += PF_THREADS)
    reinterpret_cast<uint64_t*>(s_layers)[i] = reinterpret_cast<const uint64_t*>(p.layers)[i];
  __syncthreads();

  if (warp == PF_WARPS) {
    // ---- producer: every weight row this CTA will use, in consumption order.  Nothing here depends on the previous
    // kernel, so the ring fills before the programmatic-launch wait of the consumers.
    if (lane == 0) {
      uint32_t seq = 0;
      for (int l = 0; l < p.L; ++l) {
        const PersistF32Layer& L = s_layers[l];
        const float* W[4] = {L.w_qkv, L.w_proj, L.w_fc, L.w_proj2};
        const float* Bs[4] = {L.b_qkv, L.b_proj, L.b_fc, L.b_proj2};
        const int NU[4] = {3 * D, D, 4 * D, 4 * D};
#pragma unroll
        for (int ph = 0; ph < 4; ++ph) {
#pragma unroll
          for (int j = 0; j < PF_MAXU; ++j) {
            const int cnt = live_units(NU[ph], j, cta, GW);
            for (int w = 0; w < cnt; ++w, ++seq) {
              const uint32_t slot = seq % PF_SLOTS, fill = seq / PF_SLOTS;
              ring_wait(&empty[slot], (fill & 1) ^ 1, p.err, 4, (int)seq);
              tcptx::mbar_expect_tx(&full[slot], ROW_BYTES + 16);
              const int u = cta * PF_WARPS + w + j * GW;
              const float* src = W[ph] + (size_t)u * D;
              const float* bsrc = Bs[ph] + ((ph == PH_PROJ2 ? u >> 2 : u) & ~3);  // the unit's bias is element (row & 3)
              asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                               tcptx::smem_u32(ring + (size_t)slot * SLOT)), "l"(src), "r"(ROW_BYTES), "r"(tcptx::smem_u32(&full[slot]))
                           : "memory");
              asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                               tcptx::smem_u32(ring + (size_t)slot * SLOT + D)), "l"(bsrc), "r"(16), "r"(tcptx::smem_u32(&full[slot]))
                           : "memory");
            }
          }
        }
      }
    }
    return;
  }

  // ---- consumers
  pdl_wait();  // ids / the length counter belong to the previous token's bookkeeping kernel until here
  const int len = *p.len;  // tokens so far including the one being fed: its position is len - 1
  int nbar = 0;
  long long* tr = (p.trace && cta == p.trace_cta) ? p.trace : nullptr;
  if (tr && threadIdx.x == 0) tr[0] = clock64();
  uint32_t seq_base = 0;  // rows of the ring this CTA has consumed before the current round

  auto stage_ln = [&](int l, int buf) {  // LayerNorm parameters of layer l -> shared memory (read after a consumer_sync)
    const PersistF32Layer& L = s_layers[l];
    const float* src[4] = {L.ln1_g, L.ln1_b, L.ln2_g, L.ln2_b};
    for (int i = threadIdx.x; i < 4 * D / 4; i += PF_CONSUMERS) {
      const int which = i / (D / 4), off = i - which * (D / 4);
      reinterpret_cast<float4*>(s_ln + (buf * 4 + which) * D)[off] = reinterpret_cast<const float4*>(src[which])[off];
    }
  };
  // the weight row of the warp's unit in this round: wait for it / hand its slot back to the producer
  auto wait_row = [&]() -> const float4* {
    const uint32_t seq = seq_base + warp, slot = seq % PF_SLOTS;
    ring_wait(&full[slot], (seq / PF_SLOTS) & 1, p.err, 5, (int)seq);
    return reinterpret_cast<const float4*>(ring + (size_t)slot * SLOT) + lane;
  };
  auto release_row = [&]() {
    __syncwarp();
    if (lane == 0) tcptx::mbar_arrive(&empty[(seq_base + warp) % PF_SLOTS]);
  };

  // ---- embedding (modeling_gpt2.py:577-585): CTA b writes row b
  if (cta < MB) {
    const int b = cta;
    const int id = p.ids[(size_t)b * p.ids_ld + len - 1];
    const int pe = p.npad ? max(len - 1 - p.npad[b], 0) : len - 1;
    for (int i = threadIdx.x; i < D; i += PF_CONSUMERS) p.x[(size_t)b * D + i] = p.wte[(size_t)id * D + i] + p.wpe[(size_t)pe * D + i];
  }
  stage_ln(0, 0);
  grid_barrier(p.counters, ++nbar * ncta, p.err, nbar, tr);

  // The input vector(s) of a phase go through shared memory once per CTA.  (Every warp reading them from L2 itself meant
  // 2368 requests for each of the same 24 lines at the same moment: ~4000 cycles of queueing at the L2 slices that own
  // them, the largest item of the first version's phase time.)
  auto stage_in = [&](const float* in, int ld, int n) {
    const int per = n / 4;
    for (int i = threadIdx.x; i < MB * per; i += PF_CONSUMERS) {
      const int m = i / per, off = i - m * per;
      reinterpret_cast<float4*>(s_in + m * 4 * D)[off] = __ldcg(reinterpret_cast<const float4*>(in + (size_t)m * ld) + off);
    }
    consumer_sync();
  };

  // one GEMV phase with K = d
  auto gemv = [&](int phase, int l, const float* in, int ld_in, int nrows, const float* lng, const float* lnb) {
    const PersistF32Layer& L = s_layers[l];
    if (live_units(nrows, 0, cta, GW) == 0) return;  // nothing for this CTA in this phase (uniform over the CTA)
    stage_in(in, ld_in, D);
#pragma unroll
    for (int j = 0; j < PF_MAXU; ++j) {
      const int cnt = live_units(nrows, j, cta, GW);
      if (warp < cnt) {
        const int row = gw + j * GW;
        float bv = 0.f, res[MB];  // lane 0: bias and the residual the row is added to, requested before the reductions
#pragma unroll
        for (int m = 0; m < MB; ++m) res[m] = 0.f;
        if (lane == 0 && phase == PH_PROJ)
#pragma unroll
          for (int m = 0; m < MB; ++m) res[m] = __ldcg(p.x + (size_t)m * D + row);
        const float4* w = nullptr;
        float out[MB];
#pragma unroll
        for (int m = 0; m < MB; ++m) {  // one sequence at a time: its input row in registers, the weight row read from the ring
          float4 xv[NV];
#pragma unroll
          for (int i = 0; i < NV; ++i) xv[i] = *reinterpret_cast<const float4*>(s_in + m * 4 * D + lane * 4 + i * 128);
          if (lng) {  // LayerNorm of the row (two-pass, like layernorm_kernel), on the fly
            float s = 0.f;
#pragma unroll
            for (int i = 0; i < NV; ++i) s += (xv[i].x + xv[i].y) + (xv[i].z + xv[i].w);
            const float mean = warp_sum(s) / D;
            float q = 0.f;
#pragma unroll
            for (int i = 0; i < NV; ++i) {
              const float a0 = xv[i].x - mean, a1 = xv[i].y - mean, a2 = xv[i].z - mean, a3 = xv[i].w - mean;
              q += (a0 * a0 + a1 * a1) + (a2 * a2 + a3 * a3);
            }
            const float rstd = 1.0f / sqrtf(warp_sum(q) / D + p.eps);
#pragma unroll
            for (int i = 0; i < NV; ++i) {
              const float4 g = *reinterpret_cast<const float4*>(lng + lane * 4 + i * 128);
              const float4 be = *reinterpret_cast<const float4*>(lnb + lane * 4 + i * 128);
              xv[i].x = (xv[i].x - mean) * rstd * g.x + be.x; xv[i].y = (xv[i].y - mean) * rstd * g.y + be.y;
              xv[i].z = (xv[i].z - mean) * rstd * g.z + be.z; xv[i].w = (xv[i].w - mean) * rstd * g.w + be.w;
            }
          }
          if (m == 0) {
            w = wait_row();
            bv = reinterpret_cast<const float*>(w - lane)[D + (row & 3)];
          }
          float acc = 0.f;
#pragma unroll
          for (int i = 0; i < NV; ++i) {
            const float4 wv = w[i * 32];
            acc += wv.x * xv[i].x + wv.y * xv[i].y + wv.z * xv[i].z + wv.w * xv[i].w;
          }
          out[m] = warp_sum(acc);
        }
        release_row();
        if (lane == 0) {
#pragma unroll
          for (int m = 0; m < MB; ++m) {
            const float v = out[m] + bv;
            if (phase == PH_QKV) {  // c_attn + split + cache append (modeling_gpt2.py:185-196, cache_utils.py:102-121)
              if (row < D) {
                p.q[(size_t)m * D + row] = v;
              } else {
                const int cidx = row < 2 * D ? row - D : row - 2 * D;
                float* cache = row < 2 * D ? L.kc : L.vc;
                cache[(((size_t)m * H + (cidx >> 6)) * p.Smax + (len - 1)) * 64 + (cidx & 63)] = v;
              }
            } else if (phase == PH_FC) {
              p.h[(size_t)m * 4 * D + row] = gelu_tanh(v);  // NewGELUActivation, exact tanhf
            } else {  // c_proj + residual
              p.x[(size_t)m * D + row] = res[m] + v;
            }
          }
        }
      }
      seq_base += cnt;
    }
  };
  // mlp c_proj: rows of 4d split into 4 chunks of d, one chunk per warp (chunk = warp mod 4), partial sums combined
  // through shared memory in chunk order.  A round's units exist for all 16 warps of the CTA or for none, so the two
  // CTA barriers of a round are taken by everybody or skipped by everybody.
  auto gemv_chunked = [&](int l) {
    const PersistF32Layer& L = s_layers[l];
    if (live_units(4 * D, 0, cta, GW) == 0) return;
    stage_in(p.h, 4 * D, 4 * D);
#pragma unroll
    for (int j = 0; j < PF_MAXU; ++j) {
      const int cnt = live_units(4 * D, j, cta, GW);
      if (cnt == 0) continue;
      const int u = gw + j * GW, row = u >> 2, c = u & 3;
      float bv = 0.f, res[MB];
#pragma unroll
      for (int m = 0; m < MB; ++m) res[m] = 0.f;
      if (lane == 0 && c == 0)
#pragma unroll
        for (int m = 0; m < MB; ++m) res[m] = __ldcg(p.x + (size_t)m * D + row);
      const float4* w = nullptr;
#pragma unroll
      for (int m = 0; m < MB; ++m) {
        float4 xv[NV];
#pragma unroll
        for (int i = 0; i < NV; ++i)
          xv[i] = *reinterpret_cast<const float4*>(s_in + m * 4 * D + c * D + lane * 4 + i * 128);
        if (m == 0) {
          w = wait_row();
          bv = reinterpret_cast<const float*>(w - lane)[D + (row & 3)];
        }
        float acc = 0.f;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
          const float4 wv = w[i * 32];
          acc += wv.x * xv[i].x + wv.y * xv[i].y + wv.z * xv[i].z + wv.w * xv[i].w;
        }
        acc = warp_sum(acc);
        if (lane == 0) s_part[warp * MB + m] = acc;
      }
      release_row();
      consumer_sync();
      if (c == 0 && lane == 0) {
#pragma unroll
        for (int m = 0; m < MB; ++m) {
          float s = 0.f;
#pragma unroll
          for (int cc = 0; cc < 4; ++cc) s += s_part[(warp + cc) * MB + m];  // fixed order: deterministic
          p.x[(size_t)m * D + row] = res[m] + (s + bv);
        }
      }
      consumer_sync();
      seq_base += cnt;
    }
  };

  for (int l = 0; l < p.L; ++l) {
    const PersistF32Layer& L = s_layers[l];
    const float* ln = s_ln + (l & 1) * 4 * D;
    // the cached K / V rows this CTA will attend over were written tokens ago and have left L2 since: ask for them now,
    // a phase early (one 128-byte line = half a row per thread)
    if (cta < MB * H * p.nsplit) {
      const int bh = cta / p.nsplit, sp = cta - bh * p.nsplit, b = bh / H;
      const int kp = p.npad ? p.npad[b] : 0;
      const int nkeys = len - kp, per = (nkeys + p.nsplit - 1) / p.nsplit;
      const int k0 = kp + sp * per, k1 = min(kp + nkeys - 1, k0 + per);  // not the newest row: it is written in this phase
      for (int i = threadIdx.x; i < (k1 - k0) * 4; i += PF_CONSUMERS) {
        const float* base = (i & 1 ? L.vc : L.kc) + ((size_t)bh * p.Smax + k0 + (i >> 2)) * 64 + ((i >> 1) & 1) * 32;
        asm volatile("prefetch.global.L2 [%0];" ::"l"(base));
      }
    }
    // ---- ln_1 + c_attn
    gemv(PH_QKV, l, p.x, D, 3 * D, ln, ln + D);
    if (l + 1 < p.L) stage_ln(l + 1, (l + 1) & 1);  // next layer's LayerNorm parameters, four barriers ahead of their use
    grid_barrier(p.counters, ++nbar * ncta, p.err, nbar, tr);

    // ---- attention (sdpa_attention.py:42-104 at T = 1): CTA t < MB*H*nsplit owns (sequence, head, key range)
    {
      const int nsplit = p.nsplit;
      if (cta < MB * H * nsplit) {
        const int bh = cta / nsplit, sp = cta - bh * nsplit, b = bh / H, h = bh - b * H;
        const int kp = p.npad ? p.npad[b] : 0;
        const int nkeys = len - kp;  // the new row is already in the cache
        const int per = (nkeys + nsplit - 1) / nsplit;
        const int k0 = kp + sp * per, k1 = min(kp + nkeys, k0 + per);
        const float* kc = L.kc + ((size_t)b * H + h) * p.Smax * 64;
        const float* vc = L.vc + ((size_t)b * H + h) * p.Smax * 64;
        const float2 qv = __ldcg(reinterpret_cast<const float2*>(p.q + (size_t)b * D + h * 64 + lane * 2));
        float m = -CUDART_INF_F, lsum = 0.f, o0 = 0.f, o1 = 0.f;
        for (int kb = k0 + warp; kb < k1; kb += 4 * PF_WARPS) {  // four keys per trip: their loads are in flight together
          float2 kk[4], vv[4];
          float s[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int k = kb + u * PF_WARPS;
            kk[u] = vv[u] = make_float2(0.f, 0.f);
            if (k < k1) {
              kk[u] = __ldcg(reinterpret_cast<const float2*>(kc + (size_t)k * 64 + lane * 2));
              vv[u] = __ldcg(reinterpret_cast<const float2*>(vc + (size_t)k * 64 + lane * 2));
            }
          }
#pragma unroll
          for (int u = 0; u < 4; ++u) s[u] = qv.x * kk[u].x + qv.y * kk[u].y;
#pragma unroll
          for (int o = 16; o > 0; o >>= 1)
#pragma unroll
            for (int u = 0; u < 4; ++u) s[u] += __shfl_xor_sync(0xffffffffu, s[u], o);
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            if (kb + u * PF_WARPS >= k1) break;
            const float sc = s[u] * 0.125f;
            const float mn = fmaxf(m, sc);
            const float corr = expf(m - mn), pr = expf(sc - mn);
            lsum = lsum * corr + pr;
            o0 = o0 * corr + pr * vv[u].x;
            o1 = o1 * corr + pr * vv[u].y;
            m = mn;
          }
        }
        float* mine = s_att + warp * 66;
        if (lane == 0) { mine[0] = m; mine[1] = lsum; }
        mine[2 + lane * 2] = o0;
        mine[3 + lane * 2] = o1;
        consumer_sync();
        if (warp == 0) {  // merge the 16 warps (fixed order), two output dims per lane
          const float wm = lane < PF_WARPS ? s_att[lane * 66] : -CUDART_INF_F;  // lane w weighs warp w
          float M = wm;
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) M = fmaxf(M, __shfl_xor_sync(0xffffffffu, M, o));
          const float wgt_mine = wm == -CUDART_INF_F ? 0.f : expf(wm - M);
          float den = 0.f, n0 = 0.f, n1 = 0.f;
#pragma unroll
          for (int w = 0; w < PF_WARPS; ++w) {
            const float wgt = __shfl_sync(0xffffffffu, wgt_mine, w);
            den += wgt * s_att[w * 66 + 1];
            n0 += wgt * s_att[w * 66 + 2 + lane * 2];
            n1 += wgt * s_att[w * 66 + 3 + lane * 2];
          }
          if (nsplit == 1) {
            *reinterpret_cast<float2*>(p.att + (size_t)b * D + h * 64 + lane * 2) = make_float2(n0 / den, n1 / den);
          } else {
            float* ps = p.att_scratch + ((size_t)bh * nsplit + sp) * 66;
            if (lane == 0) { ps[0] = M; ps[1] = den; }
            ps[2 + lane * 2] = n0;
            ps[3 + lane * 2] = n1;
            __threadfence();
            __syncwarp();
            int last = 0;
            if (lane == 0) last = atomicAdd(p.counters + 2 + bh, 1) == nsplit - 1;
            last = __shfl_sync(0xffffffffu, last, 0);
            if (last) {  // the head's last key range to arrive merges all of them, in range order
              __threadfence();
              float MM = -CUDART_INF_F;
              for (int s2 = 0; s2 < nsplit; ++s2) MM = fmaxf(MM, __ldcg(p.att_scratch + ((size_t)bh * nsplit + s2) * 66));
              float dd = 0.f, a0 = 0.f, a1 = 0.f;
              for (int s2 = 0; s2 < nsplit; ++s2) {
                const float* q2 = p.att_scratch + ((size_t)bh * nsplit + s2) * 66;
                const float wm = __ldcg(q2);
                const float wgt = wm == -CUDART_INF_F ? 0.f : expf(wm - MM);
                dd += wgt * __ldcg(q2 + 1);
                a0 += wgt * __ldcg(q2 + 2 + lane * 2);
                a1 += wgt * __ldcg(q2 + 3 + lane * 2);
              }
              *reinterpret_cast<float2*>(p.att + (size_t)b * D + h * 64 + lane * 2) = make_float2(a0 / dd, a1 / dd);
              if (lane == 0) p.counters[2 + bh] = 0;
            }
          }
        }
      }
    }
    grid_barrier(p.counters, ++nbar * ncta, p.err, nbar, tr);

    // ---- attention c_proj + residual
    gemv(PH_PROJ, l, p.att, D, D, nullptr, nullptr);
    grid_barrier(p.counters, ++nbar * ncta, p.err, nbar, tr);

    // ---- ln_2 + c_fc + gelu_new
    gemv(PH_FC, l, p.x, D, 4 * D, ln + 2 * D, ln + 3 * D);
    grid_barrier(p.counters, ++nbar * ncta, p.err, nbar, tr);

    // ---- mlp c_proj + residual (K = 4 d: four chunk warps per row)
    gemv_chunked(l);
    grid_barrier(p.counters, ++nbar * ncta, p.err, nbar, tr);
  }

  pdl_launch_dependents();
  // the last CTA to get here clears the barrier counter for the next launch (nobody polls it any more)
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(p.counters + 1, 1) == ncta - 1) { p.counters[0] = 0; p.counters[1] = 0; }
  }
}

template <int NV, int MB>
int launch(const PersistF32Params& p, int ncta, cudaStream_t st) {
  auto kern = decode_f32_persistent_kernel<NV, MB>;
  const int smem = (PF_SLOTS * (NV * 128 + 4) + 2 * 4 * NV * 128 + MB * 4 * NV * 128 + PF_WARPS * MB + PF_WARPS * 66) * (int)sizeof(float) +
                   2 * PF_SLOTS * (int)sizeof(uint64_t) + p.L * (int)sizeof(PersistF32Layer);
  ARG_CHECK(smem <= 227 * 1024, "decode_f32_persistent: %d layers do not fit the shared-memory pointer table", p.L);
  TRY(tutor_set_max_smem(reinterpret_cast<const void*>(kern), smem));
  CUDA_TRY(tutor_launch(kern, dim3(ncta), dim3(PF_THREADS), (size_t)smem, st, p));
  KERNEL_LAUNCHED();
  return TUTOR_OK;
}

}  // namespace

bool k_decode_f32_persistent_supported(int B, int d, int ncta) {
  // a warp holds at most PF_MAXU units per phase: 4d rows (c_fc) and 4d chunk units (mlp c_proj) over ncta * 16 warps
  return (B == 1 || B == 2) && (d == 128 || d == 768) && 4 * d <= PF_MAXU * ncta * PF_WARPS && (ncta * PF_WARPS) % 4 == 0;
}

int k_decode_f32_persistent(const PersistF32Params& p, int B, int d, int ncta, cudaStream_t st) {
  ARG_CHECK(k_decode_f32_persistent_supported(B, d, ncta), "decode_f32_persistent: B=%d d=%d on %d CTAs unsupported", B, d, ncta);
  ARG_CHECK(p.nsplit >= 1 && B * (d / 64) * p.nsplit <= ncta, "decode_f32_persistent: %d key ranges do not fit %d CTAs", p.nsplit, ncta);
  if (d == 768) return B == 1 ? launch<6, 1>(p, ncta, st) : launch<6, 2>(p, ncta, st);
  return B == 1 ? launch<1, 1>(p, ncta, st) : launch<1, 2>(p, ncta, st);
}
