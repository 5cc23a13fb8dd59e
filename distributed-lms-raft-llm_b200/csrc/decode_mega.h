// Host/device interface of the persistent decode-step kernel (decode_mega.cu).
#pragma once
#include <cuda.h>

#include "kernels.h"

enum { MG_EMBED = 0, MG_GEMM = 1, MG_FINAL = 2 };
constexpr int MG_FLAG_STRIDE = 32;  // ints between two counters: one 128-byte line each

// One entry of the GEMM engine's program.  p: 0 = ln_1 + c_attn, 1 = attention c_proj + residual,
// 2 = ln_2 + c_fc + gelu_new, 3 = mlp c_proj + residual.  in_flag / out_flag index MegaParams::flags (-1: none).
struct MegaStep {
  int kind, l, g, p;
  int in_flag, in_target, out_flag, pad;
};

// LayerNorm folded into the linear that follows it: W' = gamma (.) W is what the tensor map of c_attn / c_fc points at,
// s[n] = sum_k W'[n,k] and c[n] = sum_k beta[k] W[n,k], so that LayerNorm(x) W^T = rstd (x W'^T - mean s) + c.
struct MegaLayerW {
  const float *s_qkv, *c_qkv, *s_fc, *c_fc;
  const float *b_qkv, *b_proj, *b_fc, *b_proj2;
  bf16 *kc, *vc;  // this layer's K / V cache, [B][H][Smax][64]
};

struct MegaParams {
  int B, d, H, L, Smax;
  int G, rpg;  // row groups and rows per group (the last group may be shorter)
  int team;    // CTAs per group whose GEMM engine takes part (CTA c serves group c % G as member c / G, if c / G < team)
  // per phase: K splits, 128-row weight tiles, 64-wide k-blocks per unit, units = ntiles * ks, output width, and the
  // CTA the phase's unit 0 lands on
  int ks[4], ntiles[4], kpu[4], nunits[4], nout[4], rot[4];
  // [L][4] weight tensor maps (bf16 [N, K], 128 x 64 boxes, 128-byte swizzle), then the three activation operands
  // xb, att, hbuf (boxes of one row group x 64), all in global memory
  const CUtensorMap* maps;
  const MegaLayerW* lw;     // [L]
  const MegaStep* steps;
  int step_begin, step_end;
  float* x;         // [B, d] residual stream
  bf16* xb;         // [B, d] the same in bf16: operand of c_attn / c_fc
  float* stats;     // [d/128][B][2] per-row (mean, M2) of x over each 128-column tile
  bf16* att;        // [B, d] attention output
  bf16* hbuf;       // [B, 4d] gelu(c_fc)
  float* partial[4];  // per phase: [ks][B][N] split-K partial sums
  // counters in L2, all zero at launch and cleared again by the last CTA: [0] attention unit queue, [1] finished CTAs,
  // then the step flags and the per-tile arrival counters
  int* flags;
  int nflags;
  int* err;  // [8], zero unless a wait inside the kernel timed out: (what, a, b, block, thread)
  int flag_qkv0, flag_att0;  // QKV[l][g] = flag_qkv0 + l*G + g (c_attn partials complete); ATT likewise (heads done)
  int tc0, maxtiles;         // tile counter of (l, g, p, tile) = tc0 + ((l*G + g)*3 + p-1)*maxtiles + tile
  int attn_begin, attn_end;  // attention units [begin, end) of the (layer, row, head) order; empty: none in this launch
  int attn_wait;             // 1: units wait for QKV[l][g] (c_attn runs in this launch)
  int l2_prefetch;           // 1: a unit's K/V rows beyond its ring are prefetched into L2 when the unit is taken
  const int32_t* ids;
  int ids_ld;
  const int* len;
  const int* npad;
  const bf16* wte;
  const float* wpe;
  const float *lnf_g, *lnf_b;
  bf16* xn_last;  // [B, d] ln_f(x) for the LM head
  float eps;
  // profiling aid (null in production): CTA trace_cta stamps clock64() at fixed points of every step / attention unit
  long long* trace;
  int trace_cta;
};
constexpr int MG_TRACE_STEP_SLOTS = 8;    // per step: 0 inputs ready (activation TMA issued), 2 accumulator ready, 6 partials stored, 3 flag out,
                                          // 4 tile complete, 7 reduced, 5 flag out
constexpr int MG_TRACE_MAX_STEPS = 128;
constexpr int MG_TRACE_UNIT_SLOTS = 8;    // per attention unit of slot 0: 0 taken, 1 q/k/v ready, 2 prologue done, 3 streamed, 4 done, 5 unit id
constexpr int MG_TRACE_MAX_UNITS = 64;
constexpr int MG_TRACE_WORDS = MG_TRACE_MAX_STEPS * MG_TRACE_STEP_SLOTS + MG_TRACE_MAX_UNITS * MG_TRACE_UNIT_SLOTS + 8;

int k_decode_mega_gn(int rows_per_group);  // UMMA N (16 / 32 / 64) that covers a row group
int k_decode_mega(const MegaParams& p, int gn, int nctas, cudaStream_t st);
