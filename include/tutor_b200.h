/* tutor_b200 — C ABI of the B200-native tutoring-inference engine.
 *
 * Drop-in arithmetic for the two call sites of naggender2/distributed-lms-raft-llm's hot path
 * (REF = GUI_RAFT_LLM_SourceCode):
 *   - GPT-2 greedy generation   REF/tutoring_server.py:21-29   (model.generate)
 *   - BERT relevance gate       REF/lms_server.py:97-104       (compute_embeddings, cosine_similarity)
 * The reference has no FFI of its own (it is pure Python calling transformers/torch); these entry points are what
 * a ctypes/cffi binding placed behind those Python call sites binds — see INTEGRATION.md.
 *
 * Conventions: every function returns 0 on success or a negative TUTOR_ERR_* code; the message is available from
 * tutor_last_error() (thread-local).  No C++ types or exceptions cross the ABI.  "host" pointers are ordinary
 * (ideally pinned) host memory; "dev" pointers are CUDA device pointers owned by the caller.  Engine handles own
 * their weights, KV cache and workspaces in device memory and serialise concurrent callers with an internal mutex
 * (the reference's gRPC servers run 10 handler threads on one model, REF/tutoring_server.py:40).
 * There is no CPU fallback: every entry point fails with TUTOR_ERR_CUDA when no sm_100 device is usable.
 */
#ifndef TUTOR_B200_H
#define TUTOR_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TUTOR_OK 0
#define TUTOR_ERR_ARG (-1)
#define TUTOR_ERR_CUDA (-2)
#define TUTOR_ERR_STATE (-3)
#define TUTOR_ERR_LIMIT (-4)

#define TUTOR_DT_F32 0  /* parity mode: fp32 weights, fp32 KV cache, fp32 math (no TF32) */
#define TUTOR_DT_BF16 1 /* throughput mode: bf16 weights + KV cache, fp32 accumulation, fp32 residual stream */

#define TUTOR_ACT_NONE 0
#define TUTOR_ACT_GELU_TANH 1 /* NewGELUActivation, transformers/activations.py:59-66 */
#define TUTOR_ACT_GELU_ERF 2  /* exact GELU, transformers/activations.py:70-89 */

const char* tutor_last_error(void);
int tutor_version(void);
/* number of kernels this library has launched in the calling process (bench.py's gpu_launches) */
uint64_t tutor_launch_count(void);

/* ----------------------------------------------------------------------------------------------------------
 * GPT-2: replaces GPT2LMHeadModel(...).generate as called at REF/tutoring_server.py:21-29
 * (transformers/generation/utils.py:2131-2553 resolved to greedy + RepetitionPenalty + MaxLength + EOS).      */
typedef struct tutor_gpt2 tutor_gpt2_t;

typedef struct {
  int32_t n_layer, n_head, n_embd, vocab_size, n_positions; /* transformers GPT2Config; head_dim must be 64 */
  float layer_norm_epsilon;
  int32_t dtype; /* TUTOR_DT_* */
} tutor_gpt2_config;

int tutor_gpt2_create(const tutor_gpt2_config* cfg, int device, tutor_gpt2_t** out);
void tutor_gpt2_destroy(tutor_gpt2_t* h);
/* One call per entry of the transformers state_dict (names as in GPT2LMHeadModel.state_dict(), e.g.
 * "transformer.h.3.attn.c_attn.weight"); data is host fp32, row-major, `shape[ndim]`.  Conv1D weights
 * ([in,out], transformers/pytorch_utils.py:108-123) are transposed to K-contiguous on load; lm_head.weight is
 * tied to wte (modeling_gpt2.py:646-651) and ignored.  Unknown names return TUTOR_ERR_ARG.                   */
int tutor_gpt2_load_tensor(tutor_gpt2_t* h, const char* name, const float* host_data, const int64_t* shape,
                           int ndim);
/* Fails with TUTOR_ERR_STATE (listing the first missing tensor) unless every weight has been loaded. */
int tutor_gpt2_check_loaded(tutor_gpt2_t* h);

/* The reference-facing call: host int64 ids in, host int64 ids out (REF/tutoring_server.py:20-30 hands CPU
 * LongTensors to and from generate).  ids_host [B,T]; out_host [B, max_total_len] receives B rows of *out_len
 * tokens each, packed with row stride *out_len (prompt included, finished rows right-padded with pad_id).
 * max_total_len counts the prompt (MaxLengthCriteria, transformers/generation/stopping_criteria.py:74-83);
 * T >= max_total_len is an error like transformers' ValueError (utils.py:1575-1581).  repetition_penalty 1.0
 * disables the penalty; eos_id < 0 disables the EOS criterion.  One stream sync at the end (plus one poll of
 * the device-side finished flag every 32 steps).                                                              */
int tutor_gpt2_generate(tutor_gpt2_t* h, const int64_t* ids_host, int B, int T, int max_total_len,
                        float repetition_penalty, int eos_id, int pad_id, int64_t* out_host, int* out_len);
/* Left-padded batches (what a tokenizer with padding_side="left" produces for prompts of different lengths, and what
 * the micro-batching queue in serving.py builds): npad_host[b] = number of leading pad slots of row b (its
 * attention_mask is npad zeros followed by ones).  Real tokens never attend to the pad slots and take
 * position_ids = cumsum(mask)-1 (transformers/generation/utils.py:718-721); everything else as tutor_gpt2_generate.
 * npad_host may be null (no padding).                                                                           */
int tutor_gpt2_generate_padded(tutor_gpt2_t* h, const int64_t* ids_host, const int32_t* npad_host, int B, int T,
                               int max_total_len, float repetition_penalty, int eos_id, int pad_id,
                               int64_t* out_host, int* out_len);
/* Same as tutor_gpt2_generate, with the prompt and the result resident in device memory (int32): ids_dev [B,T], out_dev
 * [B, max_total_len] written with row stride max_total_len.                                                   */
int tutor_gpt2_generate_dev(tutor_gpt2_t* h, const int32_t* ids_dev, int B, int T, int max_total_len,
                            float repetition_penalty, int eos_id, int pad_id, int32_t* out_dev, int* out_len);

/* Stepwise form of the same loop (GenerationMixin._sample, utils.py:2736-2805) for parity tests and profiling:
 * prefill runs the prompt and selects token T; step runs n single-token forwards; read_ids copies the current
 * ids; logits materialises the fp32 logits of the most recent forward (the fused path never does).            */
int tutor_gpt2_prefill(tutor_gpt2_t* h, const int64_t* ids_host, int B, int T, int max_total_len,
                       float repetition_penalty, int eos_id, int pad_id);
int tutor_gpt2_step(tutor_gpt2_t* h, int n_steps);
int tutor_gpt2_read_ids(tutor_gpt2_t* h, int64_t* out_host, int out_capacity, int* out_len, int* stop_len);
int tutor_gpt2_logits(tutor_gpt2_t* h, float* logits_dev /* [B, vocab] */);
/* Algorithmic HBM bytes of one decode forward at cached length S for batch B (weights + KV read + KV write),
 * the figure bench.py's roofline uses.                                                        */
double tutor_gpt2_decode_bytes(tutor_gpt2_t* h, int B, int S);
/* Stream the engine enqueues on (a cudaStream_t), so callers can bracket it with CUDA events. */
void* tutor_gpt2_stream(tutor_gpt2_t* h);
/* Profiling aid of the persistent decode kernel (bf16 mode; no counterpart in the reference): cta >= 0 arms a
 * per-phase clock64() trace of that CTA for the following decode steps (re-captures the step graph); out_host, if
 * non-null, receives up to `capacity` 64-bit words of the trace recorded by the most recent step.  Layout:
 * 128 steps x 8 stamps of the GEMM engine, then 64 units x 8 stamps of attention stream 0 (csrc/decode_mega.h).
 * cta < 0 with out_host == NULL disarms it.  Returns the number of words available (> 0) or a negative error.  */
int tutor_gpt2_decode_trace(tutor_gpt2_t* h, int cta, long long* out_host, int capacity);

/* ----------------------------------------------------------------------------------------------------------
 * BERT: replaces BertModel.forward + mean(dim=1) inside compute_embeddings (REF/lms_server.py:97-101).         */
typedef struct tutor_bert tutor_bert_t;

typedef struct {
  int32_t num_hidden_layers, num_attention_heads, hidden_size, intermediate_size, vocab_size,
      max_position_embeddings; /* transformers BertConfig; head_dim must be 64 */
  float layer_norm_eps;
  int32_t dtype;
} tutor_bert_config;

int tutor_bert_create(const tutor_bert_config* cfg, int device, tutor_bert_t** out);
void tutor_bert_destroy(tutor_bert_t* h);
/* Names as in BertModel.state_dict() ("encoder.layer.0.attention.self.query.weight", ...). The pooler
 * (modeling_bert.py:456-468) is accepted and dropped: the reference never reads its output.                    */
int tutor_bert_load_tensor(tutor_bert_t* h, const char* name, const float* host_data, const int64_t* shape,
                           int ndim);
int tutor_bert_check_loaded(tutor_bert_t* h);
/* ids_host [B,T] int64 (all-ones attention mask, zero token types, as compute_embeddings produces for one
 * text or an equal-length batch) -> emb_host [B, hidden] fp32 = last_hidden_state.mean(dim=1).
 * hidden_host, if non-null, also receives last_hidden_state [B,T,hidden].                                      */
int tutor_bert_embed(tutor_bert_t* h, const int64_t* ids_host, int B, int T, float* emb_host, float* hidden_host);
/* Device-resident variant: ids_dev int32 [B,T] -> emb_dev fp32 [B, hidden]. */
int tutor_bert_embed_dev(tutor_bert_t* h, const int32_t* ids_dev, int B, int T, float* emb_dev);
void* tutor_bert_stream(tutor_bert_t* h);

/* ----------------------------------------------------------------------------------------------------------
 * Relevance scoring: replaces cosine_similarity (REF/lms_server.py:103-104, F.cosine_similarity eps 1e-8);
 * the corpus scan extends the single comparison at REF/lms_server.py:1266 to N assignment-chunk embeddings.     */
int tutor_cosine_pairs(const float* a_dev, const float* b_dev, float* out_dev, int B, int d, void* stream);
int tutor_cosine_pairs_host(const float* a_host, const float* b_host, float* out_host, int B, int d);
int tutor_row_inv_norm(const float* corpus_dev, float* inv_norm_dev, int N, int d, void* stream);
size_t tutor_cosine_topk_scratch_bytes(int Q, int N, int k);
/* queries_dev [Q,d], corpus_dev [N,d] fp32 with precomputed inverse norms -> out_scores_dev/out_idx_dev [Q,k],
 * indices offset by idx_base (the shard's first global row).  scratch_dev: tutor_cosine_topk_scratch_bytes.   */
int tutor_cosine_topk(const float* queries_dev, const float* corpus_dev, const float* inv_norm_dev, int Q, int N,
                      int d, int k, float* out_scores_dev, int32_t* out_idx_dev, int idx_base, void* scratch_dev,
                      void* stream);
/* Merge of gathered per-shard candidates [Q, ncand] (shards in ascending index order) into the global top-k. */
int tutor_topk_merge(const float* cand_scores_dev, const int32_t* cand_idx_dev, int Q, int ncand, int k,
                     float* out_scores_dev, int32_t* out_idx_dev, void* stream);

/* The exchange step of the sharded search on GPUs (SURVEY.md section 8b "topk_merge_allgather(ncclComm_t, ...)"; the
 * reference itself compares against one text, REF/lms_server.py:1254-1266, and has no collective): every rank's [Q,k]
 * candidates -> one ncclAllGather of 8-byte (score, row) elements on `stream` -> merge on every rank.  NCCL is bound at
 * run time: tutor_nccl_load(path or NULL) dlopens the library (the one PyTorch ships), rank 0 makes a unique id, every
 * rank creates its communicator from it (the id travels by any side channel, e.g. torch.distributed).  scratch holds
 * tutor_topk_merge_allgather_scratch_bytes(Q, k, world) bytes.  k <= 32, world * k <= 256.  Capturable in a CUDA graph. */
int tutor_nccl_load(const char* path);
int tutor_nccl_unique_id(void* id128 /* 128 bytes, host */);
int tutor_nccl_comm_create(const void* id128, int world, int rank, void** comm);
void tutor_nccl_comm_destroy(void* comm);
size_t tutor_topk_merge_allgather_scratch_bytes(int Q, int k, int world);
int tutor_topk_merge_allgather(void* comm, const float* scores, const int32_t* idx, int Q, int k, int world, void* scratch,
                               float* out_scores, int32_t* out_idx, void* stream);

/* ----------------------------------------------------------------------------------------------------------
 * Single kernels, for unit tests and ncu (device pointers, caller's stream).  Each stands in for the ATen op
 * listed in SURVEY.md §2.2.                                                                                   */
/* nn.LayerNorm (modeling_gpt2.py:252; modeling_bert.py:111): x [M,d] fp32 -> out_f32 and/or out_bf16. */
int tutor_layernorm(const float* x_dev, const float* gamma_dev, const float* beta_dev, float eps, float* out_f32_dev,
                    void* out_bf16_dev, int M, int d, void* stream);
/* Conv1D / nn.Linear (pytorch_utils.py:119-123): out[M,N] = act(A[M,K] · W[N,K]^T + bias) + residual.
 * dtype F32: A, W fp32, out fp32.  dtype BF16: A, W bf16 (tcgen05), out fp32 if out_is_f32 else bf16.
 * impl 0 = product kernel; impl 1 (bf16 only) = the mma.sync cross-check kernel used by the tests.            */
int tutor_linear(int dtype, const void* A_dev, const void* W_dev, const float* bias_dev, const float* residual_dev,
                 void* out_dev, int out_is_f32, int M, int N, int K, int act, int impl, void* stream);
/* The two kernels of the bf16 activation stream BERT runs in bf16 mode (BertSelfOutput / BertOutput,
 * modeling_bert.py:294-298, 352-356, with every tensor between the linear layers held in bf16 as the reference's bf16
 * modules hold it): out_bf16[M,N] = bf16(act(A · W^T + bias) + residual_bf16), fp32 accumulation and sum; and
 * nn.LayerNorm over bf16 rows (d % 8 == 0, d <= 2048), fp32 statistics.                                        */
int tutor_linear_bf16_residual(const void* A_dev, const void* W_dev, const float* bias_dev, const void* residual_bf16_dev,
                               void* out_bf16_dev, int M, int N, int K, int act, void* stream);
int tutor_layernorm_bf16(const void* x_bf16_dev, const float* gamma_dev, const float* beta_dev, float eps,
                         float* out_f32_dev, void* out_bf16_dev, int M, int d, void* stream);
/* SDPA (sdpa_attention.py:42-104), head_dim 64, scale 1/8. q/out [B*T, H*64]; k/v [B][H][Smax][64]. */
int tutor_attn_prefill(int dtype, const void* q_dev, const void* k_dev, const void* v_dev, void* out_dev, int B, int H,
                       int T, int Smax, int causal, void* stream);
int tutor_attn_decode(int dtype, const void* q_dev, const void* k_dev, const void* v_dev, void* out_dev, int B, int H,
                      int Smax, int len, void* stream);
/* The instance the bf16 decode step actually launches (GPT2Attention.forward at T = 1, modeling_gpt2.py:185-226 with
 * DynamicLayer.update, cache_utils.py:102-121): sums the ks split-K partials [ks][B][3*H*64] of c_attn in a fixed order,
 * adds the bias, appends the new K/V row at cache position len-1 (bf16 cache, written in place) and attends over rows
 * [0, len).  out bf16 [B, H*64].  Exposed for unit tests, ncu and bench.py's roofline leg.                      */
int tutor_attn_decode_fused(const float* qkv_partial_dev, int ks, const float* qkv_bias_dev, void* k_dev, void* v_dev,
                            void* out_dev, int B, int H, int Smax, int len, void* stream);
/* Fused LM head + repetition penalty + argmax (modeling_gpt2.py:705; logits_process.py:404-410;
 * utils.py:2793): x [B,d] (fp32 for F32, bf16 for BF16) against wte [V,d]; seen_dev [B, ceil(V/32)] bitmap of
 * ids to penalise (may be null when penalty == 1); out_idx_dev [B] int32.  scratch_dev of
 * tutor_lmhead_scratch_bytes(B) bytes holds the per-CTA partial maxima.                                       */
size_t tutor_lmhead_scratch_bytes(int B);
int tutor_lmhead_argmax(int dtype, const void* x_dev, const void* wte_dev, int B, int V, int d,
                        const uint32_t* seen_dev, float penalty, int32_t* out_idx_dev, void* scratch_dev,
                        void* stream);
int tutor_meanpool(const float* x_dev, float* out_dev, int B, int T, int d, void* stream);
/* The decode-step (one token per sequence, M <= 128 rows) linears of GPT2Block.forward (modeling_gpt2.py:262-309) as
 * the engine launches them in bf16 mode: LayerNorm folded into the operand load, split-K over a thread-block cluster
 * with the reduction done through distributed shared memory, bias + gelu_new or bias + residual fused behind it.
 * stats_dev [d/128][M][2] holds per-row (mean, M2) of the fp32 residual stream x over each 128-column tile:
 * tutor_row_tile_stats builds it from x, tutor_decode_linear_residual rewrites it for the x it produces.
 * ks = K splits (0 = pick; <= 8 for the cluster forms), bn = batch tile 16/32/64/128 (0 = pick).
 *   ln_linear_partial : partial_dev[ks][M][N] fp32 = per-split sums of LN(x) . W^T  (c_attn; modeling_gpt2.py:185)
 *   ln_linear_act     : out_bf16_dev[M,N] = act(LN(x) . W^T + bias)                (ln_2 + c_fc + gelu_new; :239-240,:304)
 *   linear_residual   : x_dev[M,N] += A[M,K] . W^T + bias, in place              (c_proj + residual; :223,:282,:241,:307) */
int tutor_row_tile_stats(const float* x_dev, float* stats_dev, int M, int d, void* stream);
int tutor_decode_ln_linear_partial(const float* x_dev, const float* stats_dev, const float* gamma_dev,
                                   const float* beta_dev, float eps, const void* W_dev, int M, int N, int K, int ks,
                                   float* partial_dev, void* stream);
int tutor_decode_ln_linear_act(const float* x_dev, const float* stats_dev, const float* gamma_dev, const float* beta_dev,
                               float eps, const void* W_dev, const float* bias_dev, int act, int M, int N, int K, int ks,
                               int bn, void* out_bf16_dev, void* stream);
int tutor_decode_linear_residual(const void* A_dev, const void* W_dev, const float* bias_dev, int M, int N, int K,
                                 int ks, int bn, float* x_dev, float* stats_dev, void* stream);

/* No counterpart in the reference: tutor_lmhead_argmax for callers that may not allocate device memory (a host
 * whose own pool owns the whole GPU heap, a latency-critical loop that must stay off the driver's allocation lock,
 * a plug-in with a fixed memory budget).  Same arguments and result as tutor_lmhead_argmax minus scratch_dev: the
 * partial maxima go to a 1 MiB __device__ array built into the library (resident on every GPU that loads it until
 * exit), which limits B to TUTOR_STATIC_MAX_BATCH; a larger B returns TUTOR_ERR_LIMIT.  It calls no cudaMalloc,
 * cudaFree or allocating library routine.  The first call on each device looks the array's address up
 * (cudaGetSymbolAddress) - make that call outside any stream capture.  All calls on a device share the array:
 * calls that can overlap on the GPU race; keep them on one stream.  Nothing in the array needs clearing between
 * calls (every slot that is read has been rewritten by the same call).                                          */
#define TUTOR_STATIC_MAX_BATCH 128
int tutor_lmhead_argmax_static(int dtype, const void* x_dev, const void* wte_dev, int B, int V, int d,
                               const uint32_t* seen_dev, float penalty, int32_t* out_idx_dev, void* stream);

/* ----------------------------------------------------------------------------------------------------------
 * Host tokenisation, the step either side of the GPU work (no CUDA involved; plain host code in the same library).
 * GPT-2 byte-level BPE: replaces tokenizer.encode / tokenizer.decode of REF/tutoring_server.py:20,30 (transformers'
 * GPT2Tokenizer = `tokenizers` BPE + ByteLevel).  Built from the reference's file formats: vocab.json + merges.txt.
 * encode adds no special tokens (GPT2Tokenizer's default); a literal "<|endoftext|>" in the text becomes its id.
 * decode returns raw bytes (the caller decodes UTF-8 with errors="replace", as GPT2Tokenizer does).            */
typedef struct tutor_bpe tutor_bpe_t;
int tutor_bpe_create(const char* vocab_json_path, const char* merges_txt_path, tutor_bpe_t** out);
void tutor_bpe_destroy(tutor_bpe_t* h);
int tutor_bpe_eos_id(tutor_bpe_t* h);
int tutor_bpe_vocab_size(tutor_bpe_t* h);
int tutor_bpe_encode(tutor_bpe_t* h, const char* utf8, int64_t nbytes, int32_t* ids_out, int capacity, int* n_out);
int tutor_bpe_decode(tutor_bpe_t* h, const int32_t* ids, int n, int skip_special_tokens, char* out, int capacity,
                     int* nbytes_out);
/* BERT WordPiece: replaces tokenizer(text, return_tensors="pt", padding=True, truncation=True, max_length=512) of
 * REF/lms_server.py:98 (transformers' BertTokenizer = BertNormalizer + BertPreTokenizer + WordPiece + "[CLS] $A [SEP]",
 * lower-casing and accent stripping on, as for bert-base-uncased).  Built from vocab.txt.  ids_out = [CLS] pieces [SEP],
 * cut to max_length ids in all when max_length > 0.                                                             */
typedef struct tutor_wordpiece tutor_wordpiece_t;
int tutor_wordpiece_create(const char* vocab_txt_path, tutor_wordpiece_t** out);
void tutor_wordpiece_destroy(tutor_wordpiece_t* h);
int tutor_wordpiece_encode(tutor_wordpiece_t* h, const char* utf8, int64_t nbytes, int max_length, int32_t* ids_out,
                           int capacity, int* n_out);

#ifdef __cplusplus
}
#endif
#endif /* TUTOR_B200_H */
