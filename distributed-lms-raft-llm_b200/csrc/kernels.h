// Internal launcher declarations (host side). Every launcher enqueues on `st`, never syncs,
// returns TUTOR_OK or a negative code with the text available through tutor_last_error().
#pragma once
#include "common.cuh"

// ---- elementwise.cu
int k_gpt2_embed(const int32_t* ids, int ids_ld, const int* len_dev, int T, const void* wte, int wdtype,
                 const float* wpe, int d, float* x, int M, const int* npad, cudaStream_t st);
int k_layernorm(const float* x, int in_stride, int in_offset, const float* gamma, const float* beta, float eps,
                float* out_f32, bf16* out_bf16, int M, int d, cudaStream_t st);
// the same over bf16 rows [M, d] (d % 8 == 0, d <= 2048)
int k_layernorm_bf16in(const bf16* x, const float* gamma, const float* beta, float eps, float* out_f32, bf16* out_bf16,
                       int M, int d, cudaStream_t st);
int k_bert_embed_ln(const int32_t* ids, int T, const void* word, int wdtype, const float* pos, const float* type0,
                    const float* gamma, const float* beta, float eps, float* out_f32, bf16* out_bf16, int M, int d,
                    cudaStream_t st);
int k_meanpool(const float* x, float* out, int B, int T, int d, cudaStream_t st);
int k_cosine_pairs(const float* a, const float* b, float* out, int B, int d, float eps, cudaStream_t st);
int k_seen_bitmap(const int32_t* ids, int ids_ld, int T, uint32_t* seen, int words_per_row, int B, cudaStream_t st);

// ---- gemm_f32.cu : C[m,n] = sum_k A[m,k] * W[n,k]   (A [M,K] fp32, W [N,K] fp32, both K-contiguous)
int k_gemm_f32(const float* A, const float* W, int M, int N, int K, const GemmEpilogue& epi, cudaStream_t st);

// decode GEMV (M <= 8) with the preceding LayerNorm of X folded in; xn_out (nullable) also receives LN(X)
int k_gemv_f32_ln(const float* X, const float* gamma, const float* beta, float eps, const float* W, int M, int N, int K,
                  const GemmEpilogue& epi, float* xn_out, cudaStream_t st);

// ---- gemm_bf16_tc.cu : same contraction on bf16 operands, fp32 accumulation in TMEM (tcgen05 + TMA)
int k_gemm_bf16(const bf16* A, const bf16* W, int M, int N, int K, const GemmEpilogue& epi, cudaStream_t st);
int k_gemm_bf16_splitk(const bf16* X, const bf16* W, int M, int N, int K, int ks, float* partial, cudaStream_t st);
int k_pick_splitk(int N, int K, int num_sms);
int k_make_map_bf16(void* tensor_map, const bf16* ptr, int rows, int K, int box_rows);
int k_make_map_partial(void* tensor_map, float* ptr, int N, int M, int ks, int box_m);
int k_make_map_2d(void* tensor_map, const void* ptr, int elem_bytes, int rows, int cols, int box_rows, int box_cols);

// ---- gemm_decode_fused.cu : decode-step linears with LayerNorm folded into the operand load and the split-K
// reduction done across a thread-block cluster (DSMEM) with bias + gelu / bias + residual fused behind it.
// stats: [d/128][M][2] = per-row (mean, M2) of x over each 128-column tile.
int k_decode_batch_tile(int M);
int k_pick_cluster_splitk(int N, int K, int batch_tiles, int max_ctas);
int k_row_tile_stats(const float* x, float* stats, int M, int d, bf16* xb, cudaStream_t st);  // xb: optional bf16 copy of x
int k_gemm_ln_splitk(const float* x, const float* stats, int nst, const float* gamma, const float* beta, float eps,
                     const bf16* W, int M, int N, int K, int ks, int bn, float* partial, cudaStream_t st);
int k_gemm_ln_act(const float* x, const float* stats, int nst, const float* gamma, const float* beta, float eps,
                  const bf16* W, const float* bias, int act, int M, int N, int K, int ks, int bn, bf16* out,
                  cudaStream_t st);
int k_gemm_resid_stats(const bf16* A, const bf16* W, const float* bias, int M, int N, int K, int ks, int bn, float* x,
                       float* stats, bf16* xb, cudaStream_t st);  // xb: optional bf16 copy of the new x
// LayerNorm folded into the weights (Wf = bf16(gamma (.) W), s = row sums of Wf, c = beta . W): operand xb = bf16(x)
int k_gemm_fold_splitk(const bf16* xb, const bf16* Wf, int M, int N, int K, int ks, int bn, float* partial, cudaStream_t st);
int k_gemm_fold_act(const bf16* xb, const float* stats, int nst, float eps, const bf16* Wf, const float* s, const float* c,
                    const float* bias, int act, int M, int N, int K, int ks, int bn, bf16* out, cudaStream_t st);
// c_attn with a folded LayerNorm: q/k/v[n] = rstd (sum of partials - mean s[n]) + c[n] + bias[n], closed by the attention
// kernel; s == null: plain partial sums + bias
struct QkvFold {
  const float* s;      // [3d]
  const float* c;      // [3d]
  const float* stats;  // [nst][B][2] per-tile (mean, M2) of the rows of x
  int nst;
  float eps;
};
// legacy mma.sync kernel kept only as an on-GPU cross-check of the tcgen05 kernel (tests)
int k_gemm_bf16_legacy(const bf16* A, const bf16* W, int M, int N, int K, const GemmEpilogue& epi, cudaStream_t st);

// ---- attention.cu : head_dim is 64 for every model on this path
// q [B*T, H*64] (fp32 or bf16), K/V [B][H][Smax][64] of kv_dtype, out [B*T, H*64]
// npad (nullable, [B]): leading pad slots of each left-padded sequence - real tokens never attend to them
// tcgen05 flavour (attention_tc.cu): bf16, head_dim 64, q / out [B*T, H*64], k / v in the cache layout [B][H][Smax][64]
int k_attn_prefill_tc(const bf16* q, const bf16* k, const bf16* v, bf16* out, int B, int H, int T, int Smax, int causal,
                      const int* npad, cudaStream_t st);
int k_attn_prefill(const void* q, int q_dtype, const void* k, const void* v, int kv_dtype, void* out, int out_dtype,
                   int B, int H, int T, int Smax, int causal, const int* npad, cudaStream_t st);
// one query per (b,h) against cache positions [0, len): len = *len_dev if non-null else len_host
int k_attn_decode(const void* q, int q_dtype, const void* k, const void* v, int kv_dtype, void* out, int out_dtype,
                  int B, int H, int Smax, const int* len_dev, int len_host, const int* npad, cudaStream_t st);

// fused QKV reduce + cache append + attention for the split-K decode path (bf16): qkv_partial [ks][B][3*H*64]
// safe_rows: cache rows [0, safe_rows) are final before this launch chain began (the prompt) and may be streamed
// ahead of the programmatic-launch wait; nsplit > 1 cuts the key range over gridDim.z CTAs per (b,h) (better SM balance); split_scratch holds
// k_attn_decode_scratch_bytes(B,H) bytes laid out for B rows, zero-initialised once by the owner
int k_attn_decode_fused(const float* qkv_partial, int ks, const float* qkv_bias, const QkvFold& fold, void* k, void* v,
                        int kv_dtype, void* out, int out_dtype, int B, int H, int Smax, const int* len_dev, int len_host,
                        int nsplit, void* split_scratch, const int* npad, int safe_rows, int len_hint,
                        cudaStream_t st);
int k_attn_decode_splits(int B, int H, int len_hint, int num_sms);
size_t k_attn_decode_scratch_bytes(int B, int H);

// ---- decode_fused.cu : deterministic split-K reductions fused with what follows them
// x[m,:] += bias + sum_z partial[z][m][:]  (x fp32 in/out), then LayerNorm(x[m,:]) -> xn
int k_reduce_bias_residual_ln(const float* partial, int ks, int M, int N, const float* bias, float* x,
                              const float* gamma, const float* beta, float eps, float* xn_f32, bf16* xn_bf16,
                              cudaStream_t st);
// out[m,n] = act(bias[n] + sum_z partial[z][m][n]) as bf16
int k_reduce_bias_act(const float* partial, int ks, int M, int N, const float* bias, int act, bf16* out,
                      cudaStream_t st);

// ---- lmhead.cu : fused LM head + repetition penalty + argmax (no logits tensor)
struct DecodeState {
  int32_t* ids;       // [B][ids_ld] prompt + generated
  int ids_ld;
  int* len;           // tokens currently in ids (same for all rows)
  uint32_t* seen;     // [B][seen_words] bitmap of every id so far (prompt included)
  int seen_words;
  int* finished;      // [B]
  int* stop_len;      // 0 until every row is finished, then the length at that moment
  float penalty;      // 1.0 = off
  int eos_id, pad_id;
  int32_t* plain_out;   // unit-test form: if non-null write the arg-max ids here [B] and skip all bookkeeping
};
size_t k_lmhead_scratch_bytes(int B, int V);
int k_lmhead_argmax(const void* xn, int x_dtype, const void* wte, int w_dtype, int B, int V, int d,
                    const DecodeState& ds, void* scratch, cudaStream_t st);
int k_lmhead_argmax_ln(const void* x, int x_dtype, const void* wte, int w_dtype, int B, int V, int d,
                       const DecodeState& ds, void* scratch, const float* ln_g, const float* ln_b, float ln_eps,
                       float* xn_out, cudaStream_t st);
// debug / parity: materialise logits [B,V] fp32
int k_lmhead_logits(const void* xn, int x_dtype, const void* wte, int w_dtype, int B, int V, int Vpad, int d,
                    float* logits, cudaStream_t st);

// ---- relevance_tc.cu : the corpus scan on the tensor cores (3xTF32, fp32-equivalent accuracy); scratch holds the
// split queries.  scores[q][n] = cos(query q, corpus row n), the same contract as the SIMT scan in relevance.cu
// and the exact top-k [Q,k] taken from per-32-row group maxima the scan leaves behind (k <= 32, N <= 1.6 M rows)
size_t k_cosine_scan_tc_scratch_bytes(int d, int N);
int k_cosine_scan_tc(const float* queries, const float* corpus, const float* corpus_inv_norm, int Q, int N, int d,
                     float eps, float* scores, void* scratch, int k, float* out_scores, int32_t* out_idx, int idx_base,
                     cudaStream_t st);

// ---- relevance.cu : corpus scan
size_t k_cosine_topk_scratch_bytes(int Q, int N, int k);
int k_cosine_topk(const float* queries, const void* corpus, int corpus_dtype, const float* corpus_inv_norm, int Q,
                  int N, int d, int k, float* out_scores, int32_t* out_idx, int idx_base, void* scratch,
                  cudaStream_t st);
int k_row_inv_norm(const void* corpus, int corpus_dtype, float* inv_norm, int N, int d, float eps, cudaStream_t st);
int k_topk_merge(const float* cand_scores, const int32_t* cand_idx, int Q, int ncand, int k, float* out_scores,
                 int32_t* out_idx, cudaStream_t st);
