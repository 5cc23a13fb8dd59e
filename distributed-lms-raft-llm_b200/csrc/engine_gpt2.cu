// GPT-2 greedy-generation engine behind tutor_gpt2_* (include/tutor_b200.h).
// Replaces GPT2LMHeadModel.generate as the reference calls it (REF/tutoring_server.py:21-29):
//   prefill   = GPT2LMHeadModel.forward over the prompt          (HF/models/gpt2/modeling_gpt2.py:522-706)
//   decode    = the `while unfinished` loop of GenerationMixin._sample, greedy branch
//               (HF/generation/utils.py:2743-2805), one captured CUDA graph replayed per token with a
//               device-side length counter - no per-step host sync, no torch.cat cache regrowth
//               (HF/cache_utils.py:119-120): the KV cache is preallocated [L][2][B][H][Smax][64].
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "decode_f32_persistent.h"
#include "decode_mega.h"
#include "engine_common.h"

struct Gpt2Layer {
  float *ln1_g = nullptr, *ln1_b = nullptr, *ln2_g = nullptr, *ln2_b = nullptr;
  Linear qkv, proj, fc, proj2;
  // bf16 decode: ln_1 / ln_2 folded into c_attn / c_fc (decode_mega.h).  Built when the Conv1D weight is loaded after
  // its LayerNorm's parameters (the order of a transformers state_dict); a LayerNorm tensor loaded later invalidates it.
  std::vector<float> h_ln[4];  // host copies of ln_1.weight, ln_1.bias, ln_2.weight, ln_2.bias
  void *qkv_f = nullptr, *fc_f = nullptr;
  float *qkv_s = nullptr, *qkv_c = nullptr, *fc_s = nullptr, *fc_c = nullptr;
  bool fold_qkv = false, fold_fc = false;
};

struct tutor_gpt2 {
  tutor_gpt2_config cfg{};
  int device = 0, dt = DT_F32;
  int d = 0, H = 0, L = 0, V = 0, P = 0;
  std::mutex mu;
  cudaStream_t stream = nullptr;
  // weights
  void* wte = nullptr;
  float *wpe = nullptr, *lnf_g = nullptr, *lnf_b = nullptr;
  std::vector<Gpt2Layer> layers;
  std::map<std::string, bool> loaded;
  // KV cache [L][2][Bcap][H][Smax][64]
  void* kv = nullptr;
  int Bcap = 0, Smax = 0;
  // activations for up to Mcap tokens
  int Mcap = 0;
  float* x = nullptr;
  void *xn = nullptr, *q = nullptr, *att = nullptr, *hbuf = nullptr, *xn_last = nullptr;
  // decode state
  int32_t* ids = nullptr;
  int ids_ld = 0;
  int* counters = nullptr;  // per decode lane i: [2i]=len [2i+1]=stop_len (lane 0 is also what prefill updates)
  uint32_t* seen = nullptr;
  int seen_words = 0;
  int* finished = nullptr;
  int* npad = nullptr;     // [B] leading pad slots per row (left-padded batches); all zero otherwise
  bool padded = false;
  void* lm_scratch = nullptr;
  void* attn_scratch = nullptr;  // split states + arrival counters of the decode attention, one slice per lane
  size_t attn_scratch_lane_bytes = 0;
  int attn_splits = 1;
  float* partial = nullptr;   // split-K partial sums of the decode GEMMs (bf16 mode), [ks][B][N] fp32, one slice per lane
  size_t partial_lane_elems = 0, lm_scratch_lane_bytes = 0;
  int num_sms = 148;
  bool use_fused = true, use_pdl = true;
  bool use_cluster = true;   // TUTOR_NO_CLUSTER=1: the 8-launch-per-layer chain (separate reduce kernels) for A/B runs
  float* ln_stats = nullptr;  // [d/128][B][2] per-row (mean, M2) of x over each 128-column tile, one slice per lane
  // TUTOR_FOLD=1: LayerNorm folded into c_attn / c_fc (gemm_decode_fused.cu): their operand is xb = bf16(x), kept next
  // to x by whoever writes x.  Measured 614 vs 605 us per forward at B = 64 (the TMA operand arrives no earlier than the
  // hand-normalised one, and the attention prologue and the residual epilogues do more): opt-in, default off.
  bool use_fold = false;
  bf16* xb = nullptr;         // [B, d]
  size_t ln_stats_lane_elems = 0;
  bool no_ln_gemv = false;   // TUTOR_NO_LN_GEMV=1: separate LayerNorm kernels even for 1-2 rows (A/B measurements)
  bool use_ln_gemv = false;  // TUTOR_LN_GEMV=1: LayerNorm folded into the fp32 decode GEMVs - measured slower (2213 vs 2560 tok/s)
  // Decode lanes: the batch is cut into micro-batches whose kernel chains run concurrently (own stream, own
  // length counter), so one lane's latency-bound GEMM/reduce chain overlaps the other's bandwidth-bound attention.
  static constexpr int MAX_LANES = 4;
  int cfg_lanes = 1, nlanes = 1;  // TUTOR_LANES=2|4 measured slower than 1 on B200 (profiles/README.md)
  cudaStream_t lane_stream[MAX_LANES] = {nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t ev_fork = nullptr, ev_join[MAX_LANES] = {nullptr, nullptr, nullptr, nullptr};
  int lane_b0(int i) const { return (int)((long long)cur_B * i / nlanes); }
  int* h_counters = nullptr;  // pinned
  // current generation
  int cur_B = 0, cur_T = 0, cur_total = 0;
  // prompt rows the decode attention may stream ahead of its programmatic-launch wait.  The value is baked into the
  // captured graph, so it is quantised to what the kernels can use (whole chunks, at most the ring depth = 128 rows)
  // and is part of the graph key: a shorter prompt after a longer one re-captures instead of replaying a stale bound.
  int safe_rows = 0;
  float cur_penalty = 1.f;
  int cur_eos = -1, cur_pad = 0;
  // persistent decode-step kernel (decode_mega.cu): the whole transformer stack of a step in one launch.  Opt-in
  // (TUTOR_MEGA=1): on B200 it measured slower than the 5-launches-per-layer chain of gemm_decode_fused.cu at every
  // shape tried (configs[2]: 640 vs 594 us per forward; profiles/r2a_*) - every hand-off between SMs through L2 costs
  // 1-2.5 us under load, and a step of the chain needs two of them
  int mega_mode = 0;
  uint64_t alloc_epoch = 1;  // bumped whenever a buffer or weight the plan points at is (re)allocated
  struct Mega {
    uint64_t epoch = 0;
    int B = 0, gn = 0, nsteps = 0;
    MegaParams mp{};
    void* maps = nullptr;
    MegaLayerW* lw = nullptr;
    MegaStep* steps = nullptr;
    int* flags = nullptr;
    float* partial[4] = {nullptr, nullptr, nullptr, nullptr};
    bf16* xb = nullptr;
    long long* trace = nullptr;  // tutor_gpt2_decode_trace
    int trace_cta = -1;
  } mega;
  // cooperative fp32 decode kernel for one or two sequences (decode_f32_persistent.cu); TUTOR_F32_PERSISTENT=0 selects the
  // launch chain of gemv kernels instead (A/B baseline and cross-check)
  bool use_persist_f32 = true;
  struct PersistF32 {
    uint64_t epoch = 0;
    int B = 0, nsplit = 1;
    PersistF32Layer* layers = nullptr;
    float* att_scratch = nullptr;
    int* counters = nullptr;  // [2 + B*H] counters, then err[8]
  } pf;
  // decode graph
  cudaGraphExec_t graph_exec = nullptr;
  uint64_t graph_kernels = 0;
  bool use_graph = true;
  bool prefill_pad = true;  // TUTOR_NO_PREFILL_PAD=1: ragged prompts keep their row count (cross-check)

  size_t esz() const { return dt == DT_F32 ? 4 : 2; }
  void* kcache(int l) const {
    return (uint8_t*)kv + (size_t)(2 * l) * Bcap * H * Smax * 64 * esz();
  }
  void* vcache(int l) const {
    return (uint8_t*)kv + (size_t)(2 * l + 1) * Bcap * H * Smax * 64 * esz();
  }
  DecodeState dstate(int lane = 0, int b0 = 0) const {
    DecodeState ds{};
    ds.ids = ids + (size_t)b0 * ids_ld; ds.ids_ld = ids_ld; ds.len = counters + 2 * lane; ds.stop_len = counters + 2 * lane + 1;
    ds.seen = seen + (size_t)b0 * seen_words; ds.seen_words = seen_words; ds.finished = finished + b0;
    ds.penalty = cur_penalty; ds.eos_id = cur_eos; ds.pad_id = cur_pad;
    return ds;
  }
};

namespace {

void drop_graph(tutor_gpt2* e) {
  if (e->graph_exec) { cudaGraphExecDestroy(e->graph_exec); e->graph_exec = nullptr; }
}

int gemm(tutor_gpt2* e, const void* A, const Linear& w, int M, const GemmEpilogue& epi) {
  if (e->dt == DT_F32) return k_gemm_f32((const float*)A, (const float*)w.w, M, w.N, w.K, epi, e->stream);
  return k_gemm_bf16((const bf16*)A, (const bf16*)w.w, M, w.N, w.K, epi, e->stream);
}

void persist_free(tutor_gpt2* e) {
  dev_free(e->pf.layers); dev_free(e->pf.att_scratch); dev_free(e->pf.counters);
  e->pf.epoch = 0;
}

bool persist_applies(const tutor_gpt2* e, int B) {
  return e->dt == DT_F32 && e->use_persist_f32 && k_decode_f32_persistent_supported(B, e->d, e->num_sms);
}

// key ranges per (sequence, head) of the cooperative kernel's attention phase: ~64 keys per CTA, as many CTAs as there are
int persist_nsplit(const tutor_gpt2* e, int B, int total) {
  const int most = e->num_sms / (B * e->H);
  int n = ceil_div(total, 64);
  if (n > most) n = most;
  return n < 1 ? 1 : n;
}

// per-layer pointers, scratch and counters of the cooperative fp32 decode kernel (rebuilt when an allocation changed)
int persist_prepare(tutor_gpt2* e, int B) {
  tutor_gpt2::PersistF32& f = e->pf;
  if (f.epoch == e->alloc_epoch && f.B == B) return TUTOR_OK;
  drop_graph(e);
  CUDA_TRY(cudaStreamSynchronize(e->stream));
  persist_free(e);
  std::vector<PersistF32Layer> lw(e->L);
  for (int l = 0; l < e->L; ++l) {
    const Gpt2Layer& w = e->layers[l];
    lw[l] = PersistF32Layer{w.ln1_g, w.ln1_b, w.ln2_g, w.ln2_b, (const float*)w.qkv.w, w.qkv.b, (const float*)w.proj.w, w.proj.b,
                            (const float*)w.fc.w, w.fc.b, (const float*)w.proj2.w, w.proj2.b, (float*)e->kcache(l),
                            (float*)e->vcache(l)};
  }
  TRY(dev_alloc(&f.layers, lw.size() * sizeof(PersistF32Layer)));
  CUDA_TRY(cudaMemcpy(f.layers, lw.data(), lw.size() * sizeof(PersistF32Layer), cudaMemcpyHostToDevice));
  TRY(dev_alloc(&f.att_scratch, (size_t)e->num_sms * 66 * sizeof(float)));
  const size_t nc = (size_t)2 + B * e->H + 8;
  TRY(dev_alloc(&f.counters, nc * sizeof(int)));
  CUDA_TRY(cudaMemset(f.counters, 0, nc * sizeof(int)));
  f.B = B;
  f.epoch = e->alloc_epoch;
  return TUTOR_OK;
}

int forward_decode_persist_f32(tutor_gpt2* e, int B) {
  tutor_gpt2::PersistF32& f = e->pf;
  PersistF32Params p{};
  p.L = e->L; p.Smax = e->Smax; p.nsplit = f.nsplit; p.layers = f.layers;
  p.ids = e->ids; p.ids_ld = e->ids_ld; p.len = e->counters; p.npad = e->padded ? e->npad : nullptr;
  p.wte = (const float*)e->wte; p.wpe = e->wpe;
  p.x = e->x; p.q = (float*)e->q; p.att = (float*)e->att; p.h = (float*)e->hbuf;
  p.att_scratch = f.att_scratch; p.counters = f.counters; p.err = f.counters + 2 + B * e->H;
  p.eps = e->cfg.layer_norm_epsilon;
  p.trace = e->mega.trace_cta >= 0 ? e->mega.trace : nullptr;
  p.trace_cta = e->mega.trace_cta;
  TRY(k_decode_f32_persistent(p, B, e->d, e->num_sms, e->stream));
  // ln_f is folded into the LM-head kernel; CTA 0 also leaves ln_f(x) in xn_last for tutor_gpt2_logits
  return k_lmhead_argmax_ln(e->x, e->dt, e->wte, e->dt, B, e->V, e->d, e->dstate(), e->lm_scratch, e->lnf_g, e->lnf_b,
                            e->cfg.layer_norm_epsilon, (float*)e->xn_last, e->stream);
}

// Rows per sequence of a prefill forward.  bf16 prompts whose length is not a multiple of 32 run as if right-padded to
// the next one (pad slots hold token 0): causal attention keeps the real positions' results what they were, the pad
// positions' cache rows sit past the current length (masked, and overwritten by the decode steps that reach them), and the
// c_attn GEMM keeps its TMA-staged cache scatter, which needs whole 32-token pieces of one sequence (GPT-2 large,
// 16 x 1023: 182 us per layer through the per-lane-store epilogue against 105 us).  TUTOR_NO_PREFILL_PAD=1 turns it off.
int prefill_rows(const tutor_gpt2* e, int B, int T) {
  const int Tp = (T + 31) / 32 * 32;
  if (!e->prefill_pad || e->dt != DT_BF16 || Tp == T || T < 64 || B * T <= 128 || Tp > e->P || Tp > e->Smax || Tp > e->ids_ld) return T;
  return Tp;
}

// One forward over M = B*T tokens.  decode=true: T == 1 and every position comes from the device counter.
int forward(tutor_gpt2* e, int B, int T_real, bool decode) {
  const int T = decode ? T_real : prefill_rows(e, B, T_real);
  const int M = B * T, d = e->d, dt = e->dt;
  cudaStream_t st = e->stream;
  if (T != T_real)  // pad slots of the id rows: any valid token
    CUDA_TRY(cudaMemset2DAsync(e->ids + T_real, (size_t)e->ids_ld * 4, 0, (size_t)(T - T_real) * 4, B, st));
  const int* len_dev = decode ? e->counters : nullptr;
  float* xn_f = dt == DT_F32 ? (float*)e->xn : nullptr;
  bf16* xn_b = dt == DT_BF16 ? (bf16*)e->xn : nullptr;
  const int* npad = e->padded ? e->npad : nullptr;
  TRY(k_gpt2_embed(e->ids, e->ids_ld, len_dev, T, e->wte, dt, e->wpe, d, e->x, M, npad, st));
  // parity-mode decode of a few rows: LayerNorms are folded into the GEMVs that consume them (gemv_f32_ln_kernel)
  // (M <= 2: the register GEMV, where LayerNorm costs two warp reductions; 3..8 rows: the shared-memory variant, which
  // measured slower than separate LayerNorm kernels and stays opt-in behind TUTOR_LN_GEMV=1)
  const bool ln_in_gemv = decode && dt == DT_F32 && d % 128 == 0 &&
                          ((M <= 2 && d <= 1280 && !e->no_ln_gemv) || (e->use_ln_gemv && M <= 8 && (size_t)M * d * 4 <= 48 * 1024));
  for (int l = 0; l < e->L; ++l) {
    const Gpt2Layer& w = e->layers[l];
    if (!ln_in_gemv) TRY(k_layernorm(e->x, 1, 0, w.ln1_g, w.ln1_b, e->cfg.layer_norm_epsilon, xn_f, xn_b, M, d, st));
    GemmEpilogue qe{};  // c_attn + split + cache append (modeling_gpt2.py:185-196)
    qe.bias = w.qkv.b;
    qe.out_f32 = dt == DT_F32 ? (float*)e->q : nullptr;
    qe.out_bf16 = dt == DT_BF16 ? (bf16*)e->q : nullptr;
    qe.ldo = d; qe.qkv_mode = 1; qe.kv_dtype = dt;
    qe.kcache = e->kcache(l); qe.vcache = e->vcache(l);
    qe.d = d; qe.H = e->H; qe.Smax = e->Smax; qe.T = T; qe.pos0 = 0; qe.pos_dev = len_dev;
    if (ln_in_gemv)
      TRY(k_gemv_f32_ln(e->x, w.ln1_g, w.ln1_b, e->cfg.layer_norm_epsilon, (const float*)w.qkv.w, M, w.qkv.N, w.qkv.K, qe,
                        nullptr, st));
    else
      TRY(gemm(e, e->xn, w.qkv, M, qe));
    if (decode)
      TRY(k_attn_decode(e->q, dt, e->kcache(l), e->vcache(l), dt, e->att, dt, B, e->H, e->Smax, e->counters, 0, npad, st));
    else
      TRY(k_attn_prefill(e->q, dt, e->kcache(l), e->vcache(l), dt, e->att, dt, B, e->H, T, e->Smax, 1, npad, st));
    GemmEpilogue pe{};  // attn c_proj + residual (modeling_gpt2.py:223, 282)
    pe.bias = w.proj.b; pe.residual = e->x; pe.out_f32 = e->x; pe.ldo = d;
    TRY(gemm(e, e->att, w.proj, M, pe));
    if (!ln_in_gemv) TRY(k_layernorm(e->x, 1, 0, w.ln2_g, w.ln2_b, e->cfg.layer_norm_epsilon, xn_f, xn_b, M, d, st));
    GemmEpilogue fe{};  // c_fc + gelu_new (modeling_gpt2.py:239-240)
    fe.bias = w.fc.b; fe.act = ACT_GELU_TANH; fe.ldo = 4 * d;
    fe.out_f32 = dt == DT_F32 ? (float*)e->hbuf : nullptr;
    fe.out_bf16 = dt == DT_BF16 ? (bf16*)e->hbuf : nullptr;
    if (ln_in_gemv)
      TRY(k_gemv_f32_ln(e->x, w.ln2_g, w.ln2_b, e->cfg.layer_norm_epsilon, (const float*)w.fc.w, M, w.fc.N, w.fc.K, fe,
                        nullptr, st));
    else
      TRY(gemm(e, e->xn, w.fc, M, fe));
    GemmEpilogue oe{};  // mlp c_proj + residual (modeling_gpt2.py:241, 307)
    oe.bias = w.proj2.b; oe.residual = e->x; oe.out_f32 = e->x; oe.ldo = d;
    TRY(gemm(e, e->hbuf, w.proj2, M, oe));
  }
  // ln_f on the last position of every row only (logits_to_keep = 1, HF/generation/utils.py:2490-2491)
  if (ln_in_gemv)  // folded into the LM-head kernel; CTA 0 also leaves ln_f(x) in xn_last for tutor_gpt2_logits
    return k_lmhead_argmax_ln(e->x, dt, e->wte, dt, B, e->V, d, e->dstate(), e->lm_scratch, e->lnf_g, e->lnf_b,
                              e->cfg.layer_norm_epsilon, (float*)e->xn_last, st);
  TRY(k_layernorm(e->x, T, T_real - 1, e->lnf_g, e->lnf_b, e->cfg.layer_norm_epsilon,
                  dt == DT_F32 ? (float*)e->xn_last : nullptr, dt == DT_BF16 ? (bf16*)e->xn_last : nullptr, B, d, st));
  TRY(k_lmhead_argmax(e->xn_last, dt, e->wte, dt, B, e->V, d, e->dstate(), e->lm_scratch, st));
  return TUTOR_OK;
}

// One decode forward (T == 1) in bf16 mode for the rows [b0, b0+B) of one lane: every GEMM is split along K over
// ~all SMs and leaves raw partials in L2; the reductions are fused with the work that follows (QKV -> attention
// prologue + cache append, c_proj -> bias + residual + next LayerNorm, c_fc -> bias + gelu_new).
// 8 launches per layer, all deterministic.
int forward_decode_fused(tutor_gpt2* e, int lane, int b0, int B) {
  const int d = e->d;
  cudaStream_t st = e->lane_stream[lane];
  const float eps = e->cfg.layer_norm_epsilon;
  const int* len = e->counters + 2 * lane;
  float* x = e->x + (size_t)b0 * d;
  bf16* xn = (bf16*)e->xn + (size_t)b0 * d;
  bf16* att = (bf16*)e->att + (size_t)b0 * d;
  bf16* hb = (bf16*)e->hbuf + (size_t)b0 * 4 * d;
  bf16* xl = (bf16*)e->xn_last + (size_t)b0 * d;
  float* partial = e->partial + (size_t)lane * e->partial_lane_elems;
  const size_t kv_off = (size_t)b0 * e->H * e->Smax * 64 * e->esz();
  const int* npad = e->padded ? e->npad + b0 : nullptr;
  TRY(k_gpt2_embed(e->ids + (size_t)b0 * e->ids_ld, e->ids_ld, len, 1, e->wte, e->dt, e->wpe, d, x, B, npad, st));
  TRY(k_layernorm(x, 1, 0, e->layers[0].ln1_g, e->layers[0].ln1_b, eps, nullptr, xn, B, d, st));
  for (int l = 0; l < e->L; ++l) {
    const Gpt2Layer& w = e->layers[l];
    void* kc = (uint8_t*)e->kcache(l) + kv_off;
    void* vc = (uint8_t*)e->vcache(l) + kv_off;
    int ks = k_pick_splitk(3 * d, d, e->num_sms);
    TRY(k_gemm_bf16_splitk(xn, (const bf16*)w.qkv.w, B, 3 * d, d, ks, partial, st));
    TRY(k_attn_decode_fused(partial, ks, w.qkv.b, QkvFold{}, kc, vc, DT_BF16, att, DT_BF16, B, e->H, e->Smax, len, 0,
                            e->attn_splits, (uint8_t*)e->attn_scratch + (size_t)lane * e->attn_scratch_lane_bytes, npad,
                            e->safe_rows, (e->cur_T + e->cur_total) / 2, st));
    ks = k_pick_splitk(d, d, e->num_sms);
    TRY(k_gemm_bf16_splitk(att, (const bf16*)w.proj.w, B, d, d, ks, partial, st));
    TRY(k_reduce_bias_residual_ln(partial, ks, B, d, w.proj.b, x, w.ln2_g, w.ln2_b, eps, nullptr, xn, st));
    ks = k_pick_splitk(4 * d, d, e->num_sms);
    TRY(k_gemm_bf16_splitk(xn, (const bf16*)w.fc.w, B, 4 * d, d, ks, partial, st));
    TRY(k_reduce_bias_act(partial, ks, B, 4 * d, w.fc.b, ACT_GELU_TANH, hb, st));
    ks = k_pick_splitk(d, 4 * d, e->num_sms);
    TRY(k_gemm_bf16_splitk(hb, (const bf16*)w.proj2.w, B, d, 4 * d, ks, partial, st));
    const bool last = l + 1 == e->L;  // the last block's residual feeds ln_f, every other one the next ln_1
    TRY(k_reduce_bias_residual_ln(partial, ks, B, d, w.proj2.b, x, last ? e->lnf_g : e->layers[l + 1].ln1_g,
                                  last ? e->lnf_b : e->layers[l + 1].ln1_b, eps, nullptr, last ? xl : xn, st));
  }
  return k_lmhead_argmax(xl, e->dt, e->wte, e->dt, B, e->V, d, e->dstate(lane, b0),
                         (uint8_t*)e->lm_scratch + (size_t)lane * e->lm_scratch_lane_bytes, st);
}

// Tile plan of the fused decode GEMMs.  Measured on B200 (B=64, GPT-2 small; tools/gpu_sweep.sh): cutting the batch into
// 16-row tiles and splitting K only until tiles x batch tiles x splits reaches ~4/3 of the SM count beats wide batch
// tiles with deep K splits (0.611 vs 0.652 ms per forward) - fewer fp32 partials cross L2 / DSMEM, and the weight tile
// shared by the batch tiles is served by L2.
struct DecodePlan { int bn_qkv, bn_fc, bn_res, ks_qkv, ks_proj, ks_fc, ks_proj2; };
DecodePlan decode_plan(const tutor_gpt2* e, int B) {
  const int d = e->d, cap = e->num_sms * 4 / 3;
  DecodePlan p;
  p.bn_qkv = p.bn_fc = p.bn_res = 16;
  auto env = [](const char* name, int& v) { if (const char* s = tutor_experiment(name)) v = atoi(s); };
  env("TUTOR_BN_QKV", p.bn_qkv); env("TUTOR_BN_FC", p.bn_fc); env("TUTOR_BN_RES", p.bn_res);  // A/B knobs
  p.ks_qkv = k_pick_cluster_splitk(3 * d, d, ceil_div(B, p.bn_qkv), cap);
  p.ks_proj = k_pick_cluster_splitk(d, d, ceil_div(B, p.bn_res), cap);
  p.ks_fc = k_pick_cluster_splitk(4 * d, d, ceil_div(B, p.bn_fc), cap);
  p.ks_proj2 = k_pick_cluster_splitk(d, 4 * d, ceil_div(B, p.bn_res), cap);
  env("TUTOR_KS_QKV", p.ks_qkv); env("TUTOR_KS_PROJ", p.ks_proj); env("TUTOR_KS_FC", p.ks_fc);
  env("TUTOR_KS_PROJ2", p.ks_proj2);
  return p;
}

// The same forward in 5 launches per layer (gemm_decode_fused.cu): LayerNorm is applied while the consuming GEMM
// stages its operand (from x and per-tile row statistics), and the split-K sums of c_proj / c_fc / mlp c_proj are
// closed inside the GEMM by a thread-block cluster reading its peers' shared memory.
int forward_decode_cluster(tutor_gpt2* e, int lane, int b0, int B) {
  const int d = e->d, nst = d / 128;
  cudaStream_t st = e->lane_stream[lane];
  const float eps = e->cfg.layer_norm_epsilon;
  const int* len = e->counters + 2 * lane;
  float* x = e->x + (size_t)b0 * d;
  bf16* att = (bf16*)e->att + (size_t)b0 * d;
  bf16* hb = (bf16*)e->hbuf + (size_t)b0 * 4 * d;
  bf16* xl = (bf16*)e->xn_last + (size_t)b0 * d;
  float* partial = e->partial + (size_t)lane * e->partial_lane_elems;
  float* stats = e->ln_stats + (size_t)lane * e->ln_stats_lane_elems;
  const size_t kv_off = (size_t)b0 * e->H * e->Smax * 64 * e->esz();
  const int* npad = e->padded ? e->npad + b0 : nullptr;
  const DecodePlan pl = decode_plan(e, B);
  const char* skip = tutor_experiment("TUTOR_SKIP") ? tutor_experiment("TUTOR_SKIP") : "";
  const int bn_qkv = pl.bn_qkv, bn_fc = pl.bn_fc, bn_res = pl.bn_res;
  const int ks_qkv = pl.ks_qkv, ks_proj = pl.ks_proj, ks_fc = pl.ks_fc, ks_proj2 = pl.ks_proj2;
  // folded LayerNorm when every layer has its folded weights (they are made at load time when the LayerNorm parameters
  // arrive before the Conv1D weights, as in a state_dict) and the statistic tiles fit the kernels' register arrays
  bool fold = e->use_fold && nst <= 16;
  for (int l = 0; l < e->L; ++l) fold = fold && e->layers[l].fold_qkv && e->layers[l].fold_fc;
  bf16* xb = fold ? e->xb + (size_t)b0 * d : nullptr;
  TRY(k_gpt2_embed(e->ids + (size_t)b0 * e->ids_ld, e->ids_ld, len, 1, e->wte, e->dt, e->wpe, d, x, B, npad, st));
  TRY(k_row_tile_stats(x, stats, B, d, xb, st));
  for (int l = 0; l < e->L; ++l) {
    const Gpt2Layer& w = e->layers[l];
    void* kc = (uint8_t*)e->kcache(l) + kv_off;
    void* vc = (uint8_t*)e->vcache(l) + kv_off;
    if (fold) {
      const QkvFold qf{w.qkv_s, w.qkv_c, stats, nst, eps};
      if (!strchr(skip, 'q')) TRY(k_gemm_fold_splitk(xb, (const bf16*)w.qkv_f, B, 3 * d, d, ks_qkv, bn_qkv, partial, st));
      if (!strchr(skip, 'a'))
        TRY(k_attn_decode_fused(partial, ks_qkv, w.qkv.b, qf, kc, vc, DT_BF16, att, DT_BF16, B, e->H, e->Smax, len, 0,
                                e->attn_splits, (uint8_t*)e->attn_scratch + (size_t)lane * e->attn_scratch_lane_bytes, npad,
                                e->safe_rows, (e->cur_T + e->cur_total) / 2, st));
      if (!strchr(skip, 'p'))
        TRY(k_gemm_resid_stats(att, (const bf16*)w.proj.w, w.proj.b, B, d, d, ks_proj, bn_res, x, stats, xb, st));
      if (!strchr(skip, 'f'))
        TRY(k_gemm_fold_act(xb, stats, nst, eps, (const bf16*)w.fc_f, w.fc_s, w.fc_c, w.fc.b, ACT_GELU_TANH, B, 4 * d, d, ks_fc,
                            bn_fc, hb, st));
      if (!strchr(skip, 'm'))
        TRY(k_gemm_resid_stats(hb, (const bf16*)w.proj2.w, w.proj2.b, B, d, 4 * d, ks_proj2, bn_res, x, stats, xb, st));
      continue;
    }
    // TUTOR_SKIP=<letters> leaves kernels out of the step (q: c_attn, a: attention, p: c_proj, f: c_fc, m: mlp c_proj,
    // h: LM head) - timing ablation only (tools/gpu_ablate.sh): the tokens are meaningless then
    if (!strchr(skip, 'q'))
      TRY(k_gemm_ln_splitk(x, stats, nst, w.ln1_g, w.ln1_b, eps, (const bf16*)w.qkv.w, B, 3 * d, d, ks_qkv, bn_qkv, partial,
                           st));
    if (!strchr(skip, 'a'))
      TRY(k_attn_decode_fused(partial, ks_qkv, w.qkv.b, QkvFold{}, kc, vc, DT_BF16, att, DT_BF16, B, e->H, e->Smax, len, 0,
                              e->attn_splits, (uint8_t*)e->attn_scratch + (size_t)lane * e->attn_scratch_lane_bytes, npad,
                              e->safe_rows, (e->cur_T + e->cur_total) / 2, st));
    if (!strchr(skip, 'p'))
      TRY(k_gemm_resid_stats(att, (const bf16*)w.proj.w, w.proj.b, B, d, d, ks_proj, bn_res, x, stats, nullptr, st));
    if (!strchr(skip, 'f'))
      TRY(k_gemm_ln_act(x, stats, nst, w.ln2_g, w.ln2_b, eps, (const bf16*)w.fc.w, w.fc.b, ACT_GELU_TANH, B, 4 * d, d, ks_fc,
                        bn_fc, hb, st));
    if (!strchr(skip, 'm'))
      TRY(k_gemm_resid_stats(hb, (const bf16*)w.proj2.w, w.proj2.b, B, d, 4 * d, ks_proj2, bn_res, x, stats, nullptr, st));
  }
  TRY(k_layernorm(x, 1, 0, e->lnf_g, e->lnf_b, eps, nullptr, xl, B, d, st));
  if (strchr(skip, 'h')) return TUTOR_OK;
  return k_lmhead_argmax(xl, e->dt, e->wte, e->dt, B, e->V, d, e->dstate(lane, b0),
                         (uint8_t*)e->lm_scratch + (size_t)lane * e->lm_scratch_lane_bytes, st);
}


bool mega_applies(const tutor_gpt2* e, int B) {
  if (!(e->dt == DT_BF16 && e->use_fused && e->mega_mode > 0 && B <= 128 && e->d % 128 == 0 && e->d <= 1280)) return false;
  for (const Gpt2Layer& l : e->layers)
    if (!l.fold_qkv || !l.fold_fc) return false;  // weights were loaded ahead of their LayerNorm: use the launch chain
  return true;
}

void mega_free(tutor_gpt2* e) {
  tutor_gpt2::Mega& m = e->mega;
  if (m.maps) cudaFree(m.maps);
  m.maps = nullptr;
  dev_free(m.lw); dev_free(m.steps); dev_free(m.flags); dev_free(m.xb);
  for (auto& q : m.partial) dev_free(q);
  m.epoch = 0;
}


// Split plan of one linear: ks K-splits of the N/128 weight tiles; a CTA runs ceil(units / CTAs) units one after the other.
// Cost of a round ~ k-blocks per unit + a fixed ~3 k-block-equivalents for operand load, drain, partial store and flags;
// ties go to fewer splits (fewer partials through L2).
int mega_pick_ks(int ntiles, int nkb, int nctas) {
  int best = 1, best_cost = 1 << 30;
  for (int ks = 1; ks <= nkb && ks <= 12; ++ks) {  // 12: what the closing epilogue keeps in flight per row
    if (nkb % ks) continue;
    const int rounds = ceil_div(ntiles * ks, nctas);
    const int cost = rounds * (nkb / ks + 3);
    if (cost < best_cost) { best_cost = cost; best = ks; }
  }
  return best;
}

// (Re)builds the plan of the persistent decode kernel for batch B: tensor maps, per-layer pointers, the step program,
// counters and partial-sum buffers.  Cheap to call: does nothing unless B or an underlying allocation changed.
int mega_prepare(tutor_gpt2* e, int B) {
  tutor_gpt2::Mega& m = e->mega;
  if (m.epoch == e->alloc_epoch && m.B == B) return TUTOR_OK;
  drop_graph(e);
  CUDA_TRY(cudaStreamSynchronize(e->stream));
  mega_free(e);
  const int d = e->d, L = e->L, nctas = e->num_sms;
  MegaParams mp{};
  mp.B = B; mp.d = d; mp.H = e->H; mp.L = L; mp.Smax = e->Smax;
  int G = B >= 32 ? 2 : 1;
  if (const char* s = tutor_experiment("TUTOR_MEGA_GROUPS")) G = atoi(s) < 1 ? 1 : (atoi(s) > 4 ? 4 : atoi(s));
  if (G > B) G = B;
  while (ceil_div(B, G) > 64) ++G;  // a row group is one UMMA N tile: at most 64 rows
  mp.G = G; mp.rpg = ceil_div(B, G);
  mp.G = ceil_div(B, mp.rpg);
  G = mp.G;
  m.gn = k_decode_mega_gn(mp.rpg);
  const int nout[4] = {3 * d, d, 4 * d, d}, kin[4] = {d, d, d, 4 * d};
  int rot = 0;
  int team = nctas / G;  // CTAs whose GEMM engine serves one row group
  if (const char* s = tutor_experiment("TUTOR_MEGA_TEAM")) team = atoi(s) < 1 ? 1 : (atoi(s) < team ? atoi(s) : team);
  mp.team = team;
  for (int ph = 0; ph < 4; ++ph) {
    mp.nout[ph] = nout[ph];
    mp.ntiles[ph] = nout[ph] / 128;
    const int nkb = kin[ph] / 64;
    mp.ks[ph] = mega_pick_ks(mp.ntiles[ph], nkb, team);
    mp.kpu[ph] = nkb / mp.ks[ph];
    mp.nunits[ph] = mp.ntiles[ph] * mp.ks[ph];
    mp.rot[ph] = rot;  // phases with few units land on different CTAs
    rot = (rot + mp.nunits[ph]) % team;
    TRY(dev_alloc(&m.partial[ph], (size_t)mp.ks[ph] * B * nout[ph] * sizeof(float)));
    mp.partial[ph] = m.partial[ph];
  }
  // weight tensor maps and per-layer pointers
  std::vector<CUtensorMap> maps((size_t)L * 4 + 3);
  std::vector<MegaLayerW> lw(L);
  for (int l = 0; l < L; ++l) {
    const Gpt2Layer& w = e->layers[l];
    const Linear* lin[4] = {&w.qkv, &w.proj, &w.fc, &w.proj2};
    const void* wp[4] = {w.qkv_f, w.proj.w, w.fc_f, w.proj2.w};  // c_attn / c_fc with their LayerNorm folded in
    for (int ph = 0; ph < 4; ++ph) TRY(k_make_map_bf16(&maps[(size_t)l * 4 + ph], (const bf16*)wp[ph], lin[ph]->N, lin[ph]->K, 128));
    lw[l] = MegaLayerW{w.qkv_s, w.qkv_c, w.fc_s, w.fc_c, w.qkv.b, w.proj.b, w.fc.b, w.proj2.b,
                       (bf16*)e->kcache(l), (bf16*)e->vcache(l)};
  }
  // activation operands: one row group x 64 columns per box; rows past the batch read as zeros
  TRY(dev_alloc(&m.xb, (size_t)B * d * sizeof(bf16)));
  TRY(k_make_map_bf16(&maps[(size_t)L * 4 + 0], m.xb, B, d, m.gn));
  TRY(k_make_map_bf16(&maps[(size_t)L * 4 + 1], (const bf16*)e->att, B, d, m.gn));
  TRY(k_make_map_bf16(&maps[(size_t)L * 4 + 2], (const bf16*)e->hbuf, B, 4 * d, m.gn));
  CUDA_TRY(cudaMalloc(&m.maps, maps.size() * sizeof(CUtensorMap)));
  CUDA_TRY(cudaMemcpy(m.maps, maps.data(), maps.size() * sizeof(CUtensorMap), cudaMemcpyHostToDevice));
  TRY(dev_alloc(&m.lw, lw.size() * sizeof(MegaLayerW)));
  CUDA_TRY(cudaMemcpy(m.lw, lw.data(), lw.size() * sizeof(MegaLayerW), cudaMemcpyHostToDevice));
  mp.maps = reinterpret_cast<const CUtensorMap*>(m.maps);
  mp.lw = m.lw;
  // counters: [0] queue, [1] finished CTAs, then XR[(L+1)*G], QKV, ATT, X1, HH [L*G each], tile counters
  mp.maxtiles = 4 * d / 128;
  const int xr0 = 2, qkv0 = xr0 + (L + 1) * G, att0 = qkv0 + L * G, x10 = att0 + L * G, hh0 = x10 + L * G;
  mp.tc0 = hh0 + L * G;
  mp.nflags = mp.tc0 + L * G * 3 * mp.maxtiles;
  mp.flag_qkv0 = qkv0; mp.flag_att0 = att0;
  TRY(dev_alloc(&m.flags, ((size_t)mp.nflags * MG_FLAG_STRIDE + 8) * sizeof(int)));
  CUDA_TRY(cudaMemset(m.flags, 0, ((size_t)mp.nflags * MG_FLAG_STRIDE + 8) * sizeof(int)));
  mp.flags = m.flags;
  mp.err = m.flags + (size_t)mp.nflags * MG_FLAG_STRIDE;  // behind the counters: the kernel's clearing pass stops short of it
  // the step program of the GEMM engines
  std::vector<MegaStep> steps;
  auto rows_of = [&](int g) { return (mp.rpg < B - g * mp.rpg) ? mp.rpg : B - g * mp.rpg; };
  for (int g = 0; g < G; ++g) steps.push_back(MegaStep{MG_EMBED, 0, g, 0, -1, 0, xr0 + g, 0});
  for (int g = 0; g < G; ++g) steps.push_back(MegaStep{MG_GEMM, 0, g, 0, xr0 + g, rows_of(g), qkv0 + g, 0});
  for (int l = 0; l < L; ++l)
    for (int g = 0; g < G; ++g) {
      const int lg = l * G + g;
      steps.push_back(MegaStep{MG_GEMM, l, g, 1, att0 + lg, rows_of(g) * e->H, x10 + lg, 0});
      steps.push_back(MegaStep{MG_GEMM, l, g, 2, x10 + lg, mp.nunits[1], hh0 + lg, 0});
      steps.push_back(MegaStep{MG_GEMM, l, g, 3, hh0 + lg, mp.nunits[2], xr0 + (l + 1) * G + g, 0});
      if (l + 1 < L)
        steps.push_back(MegaStep{MG_GEMM, l + 1, g, 0, xr0 + (l + 1) * G + g, mp.nunits[3], qkv0 + (l + 1) * G + g, 0});
      else
        steps.push_back(MegaStep{MG_FINAL, l, g, 0, xr0 + (l + 1) * G + g, mp.nunits[3], -1, 0});
    }
  m.nsteps = (int)steps.size();
  TRY(dev_alloc(&m.steps, steps.size() * sizeof(MegaStep)));
  CUDA_TRY(cudaMemcpy(m.steps, steps.data(), steps.size() * sizeof(MegaStep), cudaMemcpyHostToDevice));
  mp.steps = m.steps;
  mp.x = e->x; mp.xb = m.xb; mp.stats = e->ln_stats; mp.att = (bf16*)e->att; mp.hbuf = (bf16*)e->hbuf;
  mp.ids = e->ids; mp.ids_ld = e->ids_ld; mp.len = e->counters; mp.npad = nullptr;
  mp.wte = (const bf16*)e->wte; mp.wpe = e->wpe; mp.lnf_g = e->lnf_g; mp.lnf_b = e->lnf_b;
  mp.xn_last = (bf16*)e->xn_last; mp.eps = e->cfg.layer_norm_epsilon;
  m.mp = mp;
  m.B = B;
  m.epoch = e->alloc_epoch;
  return TUTOR_OK;
}

// One decode forward: the persistent kernel runs the transformer stack, the LM head follows as its own launches.
int forward_decode_mega(tutor_gpt2* e, int B) {
  tutor_gpt2::Mega& m = e->mega;
  cudaStream_t st = e->stream;
  MegaParams mp = m.mp;
  mp.npad = e->padded ? e->npad : nullptr;
  mp.trace = m.trace_cta >= 0 ? m.trace : nullptr;
  mp.trace_cta = m.trace_cta;
  mp.step_begin = 0; mp.step_end = m.nsteps;
  mp.attn_begin = 0; mp.attn_end = e->L * B * e->H; mp.attn_wait = 1;
  mp.l2_prefetch = tutor_experiment("TUTOR_MEGA_L2PF") != nullptr;  // measured slower: 774 vs 662 us (the prefetch traffic delays the chain)
  TRY(k_decode_mega(mp, m.gn, e->num_sms, st));
  return k_lmhead_argmax(e->xn_last, e->dt, e->wte, e->dt, B, e->V, e->d, e->dstate(), e->lm_scratch, st);
}

int decode_forward(tutor_gpt2* e, int B) {
  if (!(e->dt == DT_BF16 && e->use_fused && B <= 128)) {
    if (persist_applies(e, B) && e->pf.epoch == e->alloc_epoch && e->pf.B == B) {
      tutor_pdl_enabled = e->use_pdl;  // both kernels wait (griddepcontrol.wait) before touching dependent data
      int rc = forward_decode_persist_f32(e, B);
      tutor_pdl_enabled = false;
      return rc;
    }
    // parity (fp32) path: same kernel chain as prefill, with PDL when every GEMM is the weight-streaming GEMV
    // (M <= 8), all of whose kernels wait before touching dependent data
    tutor_pdl_enabled = e->use_pdl && e->dt == DT_F32 && B <= 8;
    int rc = forward(e, B, 1, true);
    tutor_pdl_enabled = false;
    return rc;
  }
  if (mega_applies(e, B) && e->nlanes == 1) {
    tutor_pdl_enabled = e->use_pdl;
    int rc = forward_decode_mega(e, B);
    tutor_pdl_enabled = false;
    return rc;
  }
  // fork: lanes 1.. wait for everything already queued on the main stream, then all chains run concurrently
  if (e->nlanes > 1) {
    CUDA_TRY(cudaEventRecord(e->ev_fork, e->stream));
    for (int i = 1; i < e->nlanes; ++i) CUDA_TRY(cudaStreamWaitEvent(e->lane_stream[i], e->ev_fork, 0));
  }
  tutor_pdl_enabled = e->use_pdl;  // every kernel on this path waits (griddepcontrol.wait) before dependent accesses
  int rc = TUTOR_OK;
  for (int i = 0; i < e->nlanes && rc == TUTOR_OK; ++i) {
    const int b0 = e->lane_b0(i), b1 = e->lane_b0(i + 1);
    rc = e->use_cluster ? forward_decode_cluster(e, i, b0, b1 - b0) : forward_decode_fused(e, i, b0, b1 - b0);
  }
  tutor_pdl_enabled = false;
  if (rc != TUTOR_OK) return rc;
  for (int i = 1; i < e->nlanes; ++i) {  // join
    CUDA_TRY(cudaEventRecord(e->ev_join[i], e->lane_stream[i]));
    CUDA_TRY(cudaStreamWaitEvent(e->stream, e->ev_join[i], 0));
  }
  return TUTOR_OK;
}

int ensure_capacity(tutor_gpt2* e, int B, int T, int total) {
  const int Smax = (int)round_up(total, 64);
  const size_t es = e->esz();
  if (B > e->Bcap || Smax > e->Smax) {
    drop_graph(e);
    ++e->alloc_epoch;
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    const int nb = B > e->Bcap ? B : e->Bcap, ns = Smax > e->Smax ? Smax : e->Smax;
    // From here until the last allocation succeeds the engine has NO capacity: if one of them fails, the next request -
    // however small - comes back through this branch and allocates everything again, instead of running on freed or
    // half-replaced buffers.
    e->Bcap = 0; e->Smax = 0;
    dev_free(e->kv);
    const size_t bytes = (size_t)e->L * 2 * nb * e->H * ns * 64 * es;
    TRY(dev_alloc(&e->kv, bytes));
    // rows past the current length are read (masked) by the attention kernels: they must hold finite values
    CUDA_TRY(cudaMemsetAsync(e->kv, 0, bytes, e->stream));
    dev_free(e->ids); dev_free(e->seen); dev_free(e->finished); dev_free(e->lm_scratch); dev_free(e->xn_last);
    dev_free(e->partial); dev_free(e->npad);
    TRY(dev_alloc(&e->npad, (size_t)nb * sizeof(int))); dev_free(e->attn_scratch);
    e->attn_scratch_lane_bytes = round_up(k_attn_decode_scratch_bytes(nb, e->H), 256);
    TRY(dev_alloc(&e->attn_scratch, tutor_gpt2::MAX_LANES * e->attn_scratch_lane_bytes));
    CUDA_TRY(cudaMemsetAsync(e->attn_scratch, 0, tutor_gpt2::MAX_LANES * e->attn_scratch_lane_bytes, e->stream));
    if (e->dt == DT_BF16) {  // largest [ks][B][N] any decode GEMM of this model writes
      const int d = e->d;
      const int shapes[4][2] = {{3 * d, d}, {d, d}, {4 * d, d}, {d, 4 * d}};
      size_t elems = 0;
      for (auto& s : shapes) {
        const size_t n = (size_t)k_pick_splitk(s[0], s[1], e->num_sms) * nb * s[0];
        if (n > elems) elems = n;
      }
      // the cluster path's c_attn partials: a live batch below capacity has fewer batch tiles and may split K deeper
      // than decode_plan(nb) does, so size for the deepest split (8) of the full capacity
      const size_t qkv_elems = (size_t)8 * nb * 3 * d;
      if (qkv_elems > elems) elems = qkv_elems;
      e->partial_lane_elems = elems;  // every lane slice can hold the whole batch (lanes only ever hold a part of it)
      TRY(dev_alloc(&e->partial, tutor_gpt2::MAX_LANES * elems * sizeof(float)));
      dev_free(e->ln_stats);
      e->ln_stats_lane_elems = (size_t)(d / 128) * nb * 2;
      TRY(dev_alloc(&e->ln_stats, tutor_gpt2::MAX_LANES * e->ln_stats_lane_elems * sizeof(float)));
      dev_free(e->xb);
      TRY(dev_alloc(&e->xb, (size_t)nb * d * sizeof(bf16)));
    }
    e->ids_ld = ns;
    e->seen_words = ceil_div(e->V, 32);
    TRY(dev_alloc(&e->ids, (size_t)nb * ns * sizeof(int32_t)));
    TRY(dev_alloc(&e->seen, (size_t)nb * e->seen_words * sizeof(uint32_t)));
    TRY(dev_alloc(&e->finished, (size_t)nb * sizeof(int)));
    e->lm_scratch_lane_bytes = k_lmhead_scratch_bytes(nb, e->V);
    TRY(dev_alloc(&e->lm_scratch, tutor_gpt2::MAX_LANES * e->lm_scratch_lane_bytes));
    TRY(dev_alloc(&e->xn_last, (size_t)nb * e->d * 4));
    e->Bcap = nb; e->Smax = ns;
  }
  const int M = B * ((T + 31) / 32 * 32);  // room for a prefill that rounds its rows up (prefill_rows)
  if (M > e->Mcap) {
    drop_graph(e);
    ++e->alloc_epoch;
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    e->Mcap = 0;  // same rule as above
    dev_free(e->x); dev_free(e->xn); dev_free(e->q); dev_free(e->att); dev_free(e->hbuf);
    const size_t d = e->d;
    TRY(dev_alloc(&e->x, (size_t)M * d * 4));
    TRY(dev_alloc(&e->xn, (size_t)M * d * es));
    TRY(dev_alloc(&e->q, (size_t)M * d * es));
    TRY(dev_alloc(&e->att, (size_t)M * d * es));
    TRY(dev_alloc(&e->hbuf, (size_t)M * 4 * d * es));
    e->Mcap = M;
  }
  return TUTOR_OK;
}

int check_args(tutor_gpt2* e, int B, int T, int total) {
  ARG_CHECK(e != nullptr, "gpt2: null handle");
  ARG_CHECK(B >= 1 && T >= 1, "gpt2.generate: empty input (B=%d, T=%d)", B, T);
  // transformers raises ValueError here (generation/utils.py:1575-1581)
  ARG_CHECK(T < total, "Input length of input_ids is %d, but `max_length` is set to %d. This can lead to unexpected "
            "behavior. You should consider increasing `max_length` or, better yet, setting `max_new_tokens`.", T, total);
  // position ids index wpe (modeling_gpt2.py:584): the last forward embeds position total-2
  ARG_CHECK(total - 1 <= e->P, "gpt2.generate: total length %d exceeds n_positions %d (+1)", total, e->P);
  ARG_CHECK(e->dt == DT_F32 || B <= 128, "gpt2.generate: bf16 mode supports batch <= 128 per call (got %d)", B);
  return TUTOR_OK;
}

// common prologue once the prompt is in e->ids (device, int32, row stride ids_ld)
int start_generation(tutor_gpt2* e, int B, int T, int total, float penalty, int eos, int pad,
                     const int32_t* npad_host = nullptr) {
  const int nl = (e->dt == DT_BF16 && e->use_fused && B <= 128) ? (B >= 2 * e->cfg_lanes ? e->cfg_lanes : 1) : 1;
  bool padded = false;
  if (npad_host) {
    for (int b = 0; b < B; ++b) {
      ARG_CHECK(npad_host[b] >= 0 && npad_host[b] < T, "gpt2.generate: row %d has %d pad slots in a %d-token prompt", b,
                npad_host[b], T);
      padded = padded || npad_host[b] > 0;
    }
    if (padded) CUDA_TRY(cudaMemcpy(e->npad, npad_host, (size_t)B * sizeof(int), cudaMemcpyHostToDevice));
  }
  if (penalty != e->cur_penalty || eos != e->cur_eos || pad != e->cur_pad || B != e->cur_B || nl != e->nlanes ||
      padded != e->padded)
    drop_graph(e);
  e->padded = padded;
  e->nlanes = nl;
  int splits = k_attn_decode_splits((B + nl - 1) / nl, e->H, (T + total) / 2, e->num_sms);
  if (const char* sv = getenv("TUTOR_ATTN_SPLITS")) splits = atoi(sv) < 1 ? 1 : (atoi(sv) > 8 ? 8 : atoi(sv));
  if (splits != e->attn_splits) drop_graph(e);
  e->attn_splits = splits;
  const int safe = (T < 128 ? T : 128) / 64 * 64;
  if (safe != e->safe_rows) drop_graph(e);
  e->safe_rows = safe;
  e->cur_B = B; e->cur_T = T; e->cur_total = total; e->cur_penalty = penalty; e->cur_eos = eos; e->cur_pad = pad;
  if (mega_applies(e, B) && nl == 1) TRY(mega_prepare(e, B));
  if (persist_applies(e, B)) {
    TRY(persist_prepare(e, B));
    const int ns = persist_nsplit(e, B, total);
    if (ns != e->pf.nsplit) drop_graph(e);  // a captured kernel argument
    e->pf.nsplit = ns;
  }
  cudaStream_t st = e->stream;
  e->h_counters[0] = T; e->h_counters[1] = 0;
  CUDA_TRY(cudaMemcpyAsync(e->counters, e->h_counters, 2 * sizeof(int), cudaMemcpyHostToDevice, st));
  CUDA_TRY(cudaMemsetAsync(e->finished, 0, (size_t)B * sizeof(int), st));
  CUDA_TRY(cudaMemsetAsync(e->seen, 0, (size_t)B * e->seen_words * sizeof(uint32_t), st));
  // the penalty covers every id so far, prompt included (logits_process.py:404-410)
  TRY(k_seen_bitmap(e->ids, e->ids_ld, T, e->seen, e->seen_words, B, st));
  TRY(forward(e, B, T, false));  // selects token T, len becomes T+1
  for (int i = 1; i < e->nlanes; ++i)  // every lane continues from the same length with its own counter pair
    CUDA_TRY(cudaMemcpyAsync(e->counters + 2 * i, e->counters, 2 * sizeof(int), cudaMemcpyDeviceToDevice, st));
  return TUTOR_OK;
}

int read_counters(tutor_gpt2* e, int* len, int* stop_len);

int run_steps(tutor_gpt2* e, int n_steps, bool poll_eos) {
  const int B = e->cur_B;
  cudaStream_t st = e->stream;
  for (int i = 0; i < n_steps; ++i) {
    if (e->use_graph && e->graph_exec) {
      CUDA_TRY(cudaGraphLaunch(e->graph_exec, st));
      tutor_count_launches(e->graph_kernels);
    } else {
      TRY(decode_forward(e, B));  // first step runs eagerly (also sets every kernel attribute) ...
      if (e->use_graph && !e->graph_exec) {  // ... then one step is captured and replayed from here on
        cudaGraph_t g = nullptr;
        const uint64_t before = tutor_launch_count();
        CUDA_TRY(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
        int rc = decode_forward(e, B);
        cudaError_t ce = cudaStreamEndCapture(st, &g);
        const uint64_t n = tutor_launch_count() - before;
        tutor_count_launches((uint64_t)0 - n);  // captured launches did not run
        if (rc != TUTOR_OK) { if (g) cudaGraphDestroy(g); return rc; }
        if (ce != cudaSuccess) return tutor_cuda_fail(ce, "cudaStreamEndCapture", __FILE__, __LINE__);
        ce = cudaGraphInstantiate(&e->graph_exec, g, 0);
        cudaGraphDestroy(g);
        if (ce != cudaSuccess) return tutor_cuda_fail(ce, "cudaGraphInstantiate", __FILE__, __LINE__);
        e->graph_kernels = n;
      }
    }
    if (poll_eos && e->cur_eos >= 0 && (i % 32) == 31 && i + 1 < n_steps) {
      int len, stop;
      TRY(read_counters(e, &len, &stop));
      if (stop != 0) break;  // every row has produced EOS (utils.py:2804-2805)
    }
  }
  return TUTOR_OK;
}

// len is common to all lanes; generation has stopped once every lane has recorded its stop length, and the
// sequence ends where the last of them finished
int read_counters(tutor_gpt2* e, int* len, int* stop_len) {
  CUDA_TRY(cudaMemcpyAsync(e->h_counters, e->counters, 2 * tutor_gpt2::MAX_LANES * sizeof(int), cudaMemcpyDeviceToHost,
                           e->stream));
  CUDA_TRY(cudaStreamSynchronize(e->stream));
  if (e->mega.flags && e->mega.epoch) {  // a wait inside the persistent decode kernel gave up: the ids are garbage
    int err[5] = {0, 0, 0, 0, 0};
    CUDA_TRY(cudaMemcpy(err, e->mega.mp.err, sizeof(err), cudaMemcpyDeviceToHost));
    if (err[0] != 0) {
      tutor_set_error("decode kernel: a wait timed out (kind %d, %d %d) in block %d thread %d", err[0], err[1], err[2],
                      err[3], err[4]);
      CUDA_TRY(cudaMemset(e->mega.mp.err, 0, 8 * sizeof(int)));
      return TUTOR_ERR_CUDA;
    }
  }
  if (e->pf.counters && e->pf.epoch) {  // a grid barrier of the cooperative fp32 kernel gave up
    int err[5] = {0, 0, 0, 0, 0};
    int* dev = e->pf.counters + 2 + e->pf.B * e->H;
    CUDA_TRY(cudaMemcpy(err, dev, sizeof(err), cudaMemcpyDeviceToHost));
    if (err[0] != 0) {
      tutor_set_error("fp32 decode kernel: grid barrier %d timed out (target %d) in block %d", err[1], err[2], err[3]);
      CUDA_TRY(cudaMemset(e->pf.counters, 0, ((size_t)2 + e->pf.B * e->H + 8) * sizeof(int)));
      return TUTOR_ERR_CUDA;
    }
  }
  *len = e->h_counters[0];
  int stop = 0;
  for (int i = 0; i < e->nlanes; ++i) {
    if (e->h_counters[2 * i + 1] == 0) { stop = 0; break; }
    if (e->h_counters[2 * i + 1] > stop) stop = e->h_counters[2 * i + 1];
  }
  *stop_len = stop;
  return TUTOR_OK;
}

int upload_prompt(tutor_gpt2* e, const int64_t* ids_host, int B, int T) {
  std::vector<int32_t> tmp((size_t)B * T);
  for (size_t i = 0; i < tmp.size(); ++i) {
    ARG_CHECK(ids_host[i] >= 0 && ids_host[i] < e->V, "gpt2.generate: token id %lld outside [0,%d)",
              (long long)ids_host[i], e->V);
    tmp[i] = (int32_t)ids_host[i];
  }
  CUDA_TRY(cudaMemcpy2DAsync(e->ids, (size_t)e->ids_ld * 4, tmp.data(), (size_t)T * 4, (size_t)T * 4, B,
                             cudaMemcpyHostToDevice, e->stream));
  CUDA_TRY(cudaStreamSynchronize(e->stream));  // tmp is pageable and about to go out of scope
  return TUTOR_OK;
}

struct NameParts { int layer; std::string rest; };
NameParts split_name(const std::string& full) {
  std::string n = full;
  if (n.rfind("transformer.", 0) == 0) n = n.substr(12);
  NameParts p{-1, n};
  if (n.rfind("h.", 0) == 0) {
    size_t dot = n.find('.', 2);
    p.layer = atoi(n.substr(2, dot - 2).c_str());
    p.rest = n.substr(dot + 1);
  }
  return p;
}

}  // namespace

extern "C" {

int tutor_gpt2_create(const tutor_gpt2_config* cfg, int device, tutor_gpt2_t** out) {
  ARG_CHECK(cfg && out, "gpt2_create: null argument");
  ARG_CHECK(cfg->n_embd == cfg->n_head * 64, "gpt2_create: head_dim must be 64 (n_embd=%d n_head=%d)", cfg->n_embd,
            cfg->n_head);
  ARG_CHECK(cfg->n_embd % 128 == 0, "gpt2_create: n_embd must be a multiple of 128");
  ARG_CHECK(cfg->dtype == DT_F32 || cfg->dtype == DT_BF16, "gpt2_create: bad dtype %d", cfg->dtype);
  TRY(tutor_require_device(device));
  tutor_gpt2* e = new tutor_gpt2();
  e->cfg = *cfg; e->device = device; e->dt = cfg->dtype;
  e->d = cfg->n_embd; e->H = cfg->n_head; e->L = cfg->n_layer; e->V = cfg->vocab_size; e->P = cfg->n_positions;
  e->layers.resize(e->L);
  e->use_graph = getenv("TUTOR_NO_GRAPH") == nullptr;
  e->prefill_pad = getenv("TUTOR_NO_PREFILL_PAD") == nullptr;
  e->use_fused = getenv("TUTOR_NO_FUSED") == nullptr;  // debugging aids: unfused decode kernels / no PDL
  e->use_pdl = getenv("TUTOR_NO_PDL") == nullptr;
  e->use_cluster = getenv("TUTOR_NO_CLUSTER") == nullptr;
  if (const char* f = getenv("TUTOR_FOLD")) e->use_fold = atoi(f) != 0;
  if (const char* mm = getenv("TUTOR_MEGA")) e->mega_mode = atoi(mm);
  if (const char* pf = getenv("TUTOR_F32_PERSISTENT")) e->use_persist_f32 = atoi(pf) != 0;
  if (!e->use_cluster) e->mega_mode = 0;
  e->use_ln_gemv = tutor_experiment("TUTOR_LN_GEMV") != nullptr;
  e->no_ln_gemv = tutor_experiment("TUTOR_NO_LN_GEMV") != nullptr;
  {
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) e->num_sms = prop.multiProcessorCount;
  }
  if (const char* nl = tutor_experiment("TUTOR_LANES")) e->cfg_lanes = atoi(nl);
  if (e->cfg_lanes < 1) e->cfg_lanes = 1;
  if (e->cfg_lanes > tutor_gpt2::MAX_LANES) e->cfg_lanes = tutor_gpt2::MAX_LANES;
  bool ok = cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking) == cudaSuccess;
  e->lane_stream[0] = e->stream;
  for (int i = 1; i < tutor_gpt2::MAX_LANES && ok; ++i)
    ok = cudaStreamCreateWithFlags(&e->lane_stream[i], cudaStreamNonBlocking) == cudaSuccess &&
         cudaEventCreateWithFlags(&e->ev_join[i], cudaEventDisableTiming) == cudaSuccess;
  ok = ok && cudaEventCreateWithFlags(&e->ev_fork, cudaEventDisableTiming) == cudaSuccess;
  if (!ok || cudaMalloc(&e->counters, 2 * tutor_gpt2::MAX_LANES * sizeof(int)) != cudaSuccess ||
      cudaMallocHost(&e->h_counters, 2 * tutor_gpt2::MAX_LANES * sizeof(int)) != cudaSuccess) {
    int rc = tutor_cuda_fail(cudaGetLastError(), "gpt2_create", __FILE__, __LINE__);
    delete e;
    return rc;
  }
  *out = e;
  return TUTOR_OK;
}

void tutor_gpt2_destroy(tutor_gpt2_t* e) {
  if (!e) return;
  cudaSetDevice(e->device);
  cudaStreamSynchronize(e->stream);
  drop_graph(e);
  mega_free(e);
  persist_free(e);
  dev_free(e->mega.trace);
  dev_free(e->wte); dev_free(e->wpe); dev_free(e->lnf_g); dev_free(e->lnf_b);
  for (auto& l : e->layers) {
    dev_free(l.ln1_g); dev_free(l.ln1_b); dev_free(l.ln2_g); dev_free(l.ln2_b);
    dev_free(l.qkv.w); dev_free(l.qkv.b); dev_free(l.proj.w); dev_free(l.proj.b);
    dev_free(l.fc.w); dev_free(l.fc.b); dev_free(l.proj2.w); dev_free(l.proj2.b);
    dev_free(l.qkv_f); dev_free(l.fc_f); dev_free(l.qkv_s); dev_free(l.qkv_c); dev_free(l.fc_s); dev_free(l.fc_c);
  }
  dev_free(e->kv); dev_free(e->x); dev_free(e->xn); dev_free(e->q); dev_free(e->att); dev_free(e->hbuf);
  dev_free(e->xn_last); dev_free(e->ids); dev_free(e->counters); dev_free(e->seen); dev_free(e->finished);
  dev_free(e->lm_scratch); dev_free(e->partial); dev_free(e->attn_scratch); dev_free(e->npad); dev_free(e->ln_stats); dev_free(e->xb);
  if (e->h_counters) cudaFreeHost(e->h_counters);
  for (int i = 1; i < tutor_gpt2::MAX_LANES; ++i) {
    if (e->lane_stream[i]) cudaStreamDestroy(e->lane_stream[i]);
    if (e->ev_join[i]) cudaEventDestroy(e->ev_join[i]);
  }
  if (e->ev_fork) cudaEventDestroy(e->ev_fork);
  cudaStreamDestroy(e->stream);
  delete e;
}

int tutor_gpt2_load_tensor(tutor_gpt2_t* e, const char* name, const float* data, const int64_t* shape, int ndim) {
  ARG_CHECK(e && name && data && shape, "gpt2_load_tensor: null argument");
  std::lock_guard<std::mutex> lock(e->mu);
  CUDA_TRY(cudaSetDevice(e->device));
  // weight buffers are about to be replaced: the captured decode step holds their addresses
  CUDA_TRY(cudaStreamSynchronize(e->stream));
  drop_graph(e);
  ++e->alloc_epoch;
  const std::string full(name);
  if (full == "lm_head.weight") return TUTOR_OK;  // tied to wte
  NameParts p = split_name(full);
  if (p.rest == "attn.bias" || p.rest == "attn.masked_bias") return TUTOR_OK;  // causal-mask buffers
  const int d = e->d;
  auto want = [&](int64_t a, int64_t b) -> int {
    bool ok = (b < 0) ? (ndim == 1 && shape[0] == a) : (ndim == 2 && shape[0] == a && shape[1] == b);
    ARG_CHECK(ok, "gpt2_load_tensor: %s has the wrong shape", name);
    return TUTOR_OK;
  };
  int rc = TUTOR_ERR_ARG;
  if (p.layer < 0) {
    if (p.rest == "wte.weight") { TRY(want(e->V, d)); rc = upload_matrix(&e->wte, data, e->V, d, false, e->dt); }
    else if (p.rest == "wpe.weight") { TRY(want(e->P, d)); rc = upload_f32(&e->wpe, data, (size_t)e->P * d); }
    else if (p.rest == "ln_f.weight") { TRY(want(d, -1)); rc = upload_f32(&e->lnf_g, data, d); }
    else if (p.rest == "ln_f.bias") { TRY(want(d, -1)); rc = upload_f32(&e->lnf_b, data, d); }
    else { tutor_set_error("gpt2_load_tensor: unknown tensor %s", name); return TUTOR_ERR_ARG; }
  } else {
    ARG_CHECK(p.layer < e->L, "gpt2_load_tensor: layer %d out of range in %s", p.layer, name);
    Gpt2Layer& l = e->layers[p.layer];
    struct { const char* n; Linear* lin; int K, N; } lins[] = {
        {"attn.c_attn", &l.qkv, d, 3 * d}, {"attn.c_proj", &l.proj, d, d},
        {"mlp.c_fc", &l.fc, d, 4 * d}, {"mlp.c_proj", &l.proj2, 4 * d, d}};
    struct { const char* n; float** ptr; } vecs[] = {
        {"ln_1.weight", &l.ln1_g}, {"ln_1.bias", &l.ln1_b}, {"ln_2.weight", &l.ln2_g}, {"ln_2.bias", &l.ln2_b}};
    bool found = false;
    for (int vi = 0; vi < 4; ++vi)
      if (p.rest == vecs[vi].n) {
        TRY(want(d, -1));
        rc = upload_f32(vecs[vi].ptr, data, d);
        found = true;
        if (e->dt == DT_BF16) {
          l.h_ln[vi].assign(data, data + d);
          (vi < 2 ? l.fold_qkv : l.fold_fc) = false;  // a fold made with the old parameters is stale
        }
      }
    for (auto& li : lins) {
      if (p.rest == std::string(li.n) + ".weight") {
        TRY(want(li.K, li.N));  // Conv1D stores [in,out]; transpose to [out,in]
        rc = upload_matrix(&li.lin->w, data, li.K, li.N, true, e->dt);
        li.lin->N = li.N; li.lin->K = li.K; found = true;
        const bool is_qkv = li.lin == &l.qkv, is_fc = li.lin == &l.fc;
        if (rc == TUTOR_OK && e->dt == DT_BF16 && (is_qkv || is_fc)) {
          const int g = is_qkv ? 0 : 2;
          bool& ok = is_qkv ? l.fold_qkv : l.fold_fc;
          ok = false;
          if ((int)l.h_ln[g].size() == d && (int)l.h_ln[g + 1].size() == d) {
            rc = upload_ln_folded(is_qkv ? &l.qkv_f : &l.fc_f, is_qkv ? &l.qkv_s : &l.fc_s, is_qkv ? &l.qkv_c : &l.fc_c,
                                  data, li.K, li.N, l.h_ln[g].data(), l.h_ln[g + 1].data());
            ok = rc == TUTOR_OK;
          }
        }
      } else if (p.rest == std::string(li.n) + ".bias") {
        TRY(want(li.N, -1));
        rc = upload_f32(&li.lin->b, data, li.N); found = true;
      }
    }
    if (!found) { tutor_set_error("gpt2_load_tensor: unknown tensor %s", name); return TUTOR_ERR_ARG; }
  }
  if (rc == TUTOR_OK) e->loaded[(p.layer < 0 ? std::string() : "h." + std::to_string(p.layer) + ".") + p.rest] = true;
  return rc;
}

int tutor_gpt2_check_loaded(tutor_gpt2_t* e) {
  ARG_CHECK(e, "gpt2: null handle");
  std::vector<std::string> need = {"wte.weight", "wpe.weight", "ln_f.weight", "ln_f.bias"};
  const char* per[] = {"ln_1.weight", "ln_1.bias", "ln_2.weight", "ln_2.bias", "attn.c_attn.weight", "attn.c_attn.bias",
                       "attn.c_proj.weight", "attn.c_proj.bias", "mlp.c_fc.weight", "mlp.c_fc.bias",
                       "mlp.c_proj.weight", "mlp.c_proj.bias"};
  for (int l = 0; l < e->L; ++l)
    for (const char* n : per) need.push_back("h." + std::to_string(l) + "." + n);
  for (auto& n : need)
    if (!e->loaded.count(n)) {
      tutor_set_error("gpt2: weight transformer.%s has not been loaded", n.c_str());
      return TUTOR_ERR_STATE;
    }
  return TUTOR_OK;
}

int tutor_gpt2_prefill(tutor_gpt2_t* e, const int64_t* ids_host, int B, int T, int total, float penalty, int eos,
                       int pad) {
  TRY(check_args(e, B, T, total));
  ARG_CHECK(ids_host, "gpt2: null ids");
  std::lock_guard<std::mutex> lock(e->mu);
  CUDA_TRY(cudaSetDevice(e->device));
  TRY(tutor_gpt2_check_loaded(e));
  TRY(ensure_capacity(e, B, T, total));
  TRY(upload_prompt(e, ids_host, B, T));
  return start_generation(e, B, T, total, penalty, eos, pad);
}

int tutor_gpt2_step(tutor_gpt2_t* e, int n_steps) {
  ARG_CHECK(e && e->cur_B > 0, "gpt2_step: no generation in progress");
  std::lock_guard<std::mutex> lock(e->mu);
  CUDA_TRY(cudaSetDevice(e->device));
  int len, stop;
  TRY(read_counters(e, &len, &stop));
  ARG_CHECK(len + n_steps <= e->cur_total, "gpt2_step: %d more steps would exceed max length %d (len %d)", n_steps,
            e->cur_total, len);
  return run_steps(e, n_steps, false);
}

static int read_ids_locked(tutor_gpt2_t* e, int64_t* out_host, int cap, int* out_len, int* stop_len) {
  int len, stop;
  TRY(read_counters(e, &len, &stop));
  const int n = stop ? stop : len;
  ARG_CHECK((size_t)cap >= (size_t)e->cur_B * n, "gpt2_read_ids: output capacity %d < %d", cap, e->cur_B * n);
  std::vector<int32_t> tmp((size_t)e->cur_B * n);
  CUDA_TRY(cudaMemcpy2D(tmp.data(), (size_t)n * 4, e->ids, (size_t)e->ids_ld * 4, (size_t)n * 4, e->cur_B,
                        cudaMemcpyDeviceToHost));
  for (size_t i = 0; i < tmp.size(); ++i) out_host[i] = tmp[i];
  *out_len = n;
  if (stop_len) *stop_len = stop;
  return TUTOR_OK;
}

int tutor_gpt2_read_ids(tutor_gpt2_t* e, int64_t* out_host, int cap, int* out_len, int* stop_len) {
  ARG_CHECK(e && e->cur_B > 0 && out_host && out_len, "gpt2_read_ids: bad argument or no generation in progress");
  std::lock_guard<std::mutex> lock(e->mu);
  CUDA_TRY(cudaSetDevice(e->device));
  return read_ids_locked(e, out_host, cap, out_len, stop_len);
}

int tutor_gpt2_generate(tutor_gpt2_t* e, const int64_t* ids_host, int B, int T, int total, float penalty, int eos,
                        int pad, int64_t* out_host, int* out_len) {
  return tutor_gpt2_generate_padded(e, ids_host, nullptr, B, T, total, penalty, eos, pad, out_host, out_len);
}

int tutor_gpt2_generate_padded(tutor_gpt2_t* e, const int64_t* ids_host, const int32_t* npad_host, int B, int T,
                               int total, float penalty, int eos, int pad, int64_t* out_host, int* out_len) {
  TRY(check_args(e, B, T, total));
  ARG_CHECK(ids_host && out_host && out_len, "gpt2.generate: null argument");
  std::lock_guard<std::mutex> lock(e->mu);  // held to the end: 10 gRPC handler threads share one engine
  CUDA_TRY(cudaSetDevice(e->device));
  TRY(tutor_gpt2_check_loaded(e));
  TRY(ensure_capacity(e, B, T, total));
  TRY(upload_prompt(e, ids_host, B, T));
  TRY(start_generation(e, B, T, total, penalty, eos, pad, npad_host));
  TRY(run_steps(e, total - T - 1, true));
  return read_ids_locked(e, out_host, B * total, out_len, nullptr);
}

int tutor_gpt2_generate_dev(tutor_gpt2_t* e, const int32_t* ids_dev, int B, int T, int total, float penalty, int eos,
                            int pad, int32_t* out_dev, int* out_len) {
  TRY(check_args(e, B, T, total));
  ARG_CHECK(ids_dev && out_dev && out_len, "gpt2.generate_dev: null argument");
  std::lock_guard<std::mutex> lock(e->mu);
  CUDA_TRY(cudaSetDevice(e->device));
  TRY(tutor_gpt2_check_loaded(e));
  TRY(ensure_capacity(e, B, T, total));
  CUDA_TRY(cudaMemcpy2DAsync(e->ids, (size_t)e->ids_ld * 4, ids_dev, (size_t)T * 4, (size_t)T * 4, B,
                             cudaMemcpyDeviceToDevice, e->stream));
  TRY(start_generation(e, B, T, total, penalty, eos, pad));
  TRY(run_steps(e, total - T - 1, true));
  int len, stop;
  TRY(read_counters(e, &len, &stop));
  CUDA_TRY(cudaMemcpy2DAsync(out_dev, (size_t)total * 4, e->ids, (size_t)e->ids_ld * 4, (size_t)total * 4, B,
                             cudaMemcpyDeviceToDevice, e->stream));
  CUDA_TRY(cudaStreamSynchronize(e->stream));
  *out_len = stop ? stop : len;
  return TUTOR_OK;
}

int tutor_gpt2_logits(tutor_gpt2_t* e, float* logits_dev) {
  ARG_CHECK(e && e->cur_B > 0 && logits_dev, "gpt2_logits: bad argument or no forward has run");
  std::lock_guard<std::mutex> lock(e->mu);
  CUDA_TRY(cudaSetDevice(e->device));
  TRY(k_lmhead_logits(e->xn_last, e->dt, e->wte, e->dt, e->cur_B, e->V, e->V, e->d, logits_dev, e->stream));
  CUDA_TRY(cudaStreamSynchronize(e->stream));
  return TUTOR_OK;
}

double tutor_gpt2_decode_bytes(tutor_gpt2_t* e, int B, int S) {
  if (!e) return 0.0;
  const double d = e->d, es = (double)e->esz();
  // parameters streamed per forward: 12 d^2 + 13 d per layer, 2 d for ln_f, V*d for the tied head
  const double params = e->L * (12.0 * d * d + 13.0 * d) + 2.0 * d + (double)e->V * d;
  const double kv_read = (double)B * S * e->L * 2.0 * d * es;
  const double kv_write = (double)B * e->L * 2.0 * d * es;
  return params * es + kv_read + kv_write;
}

void* tutor_gpt2_stream(tutor_gpt2_t* e) { return e ? (void*)e->stream : nullptr; }

int tutor_gpt2_decode_trace(tutor_gpt2_t* e, int cta, long long* out_host, int capacity) {
  ARG_CHECK(e, "gpt2_decode_trace: null handle");
  std::lock_guard<std::mutex> lock(e->mu);
  CUDA_TRY(cudaSetDevice(e->device));
  CUDA_TRY(cudaStreamSynchronize(e->stream));
  tutor_gpt2::Mega& m = e->mega;
  if (out_host) {
    ARG_CHECK(m.trace, "gpt2_decode_trace: no trace has been armed");
    const int n = capacity < MG_TRACE_WORDS ? capacity : MG_TRACE_WORDS;
    CUDA_TRY(cudaMemcpy(out_host, m.trace, (size_t)n * sizeof(long long), cudaMemcpyDeviceToHost));
  }
  if (cta >= 0 || !out_host) {
    if (cta >= 0 && !m.trace) TRY(dev_alloc(&m.trace, MG_TRACE_WORDS * sizeof(long long)));
    if (m.trace) CUDA_TRY(cudaMemset(m.trace, 0, MG_TRACE_WORDS * sizeof(long long)));
    if (cta != m.trace_cta) drop_graph(e);  // the trace pointer is a captured kernel argument
    m.trace_cta = cta;
  }
  return MG_TRACE_WORDS;
}

}  // extern "C"
