// BERT-base encoder + mean-pool behind tutor_bert_* (include/tutor_b200.h).
// Replaces `model(**inputs).last_hidden_state.mean(dim=1)` inside compute_embeddings
// (REF/lms_server.py:97-101): BertEmbeddings + 12 post-LN BertLayers (HF/models/bert/modeling_bert.py:
// 72-112, 379-421) with all-ones mask and zero token types; the pooler (:456-468) is dead work in the
// reference (it only reads last_hidden_state) and is skipped.  The three Q/K/V nn.Linear layers are fused
// into one [3d,d] GEMM at load; residual+LayerNorm and bias+erf-GELU ride in the GEMM epilogues / LN kernel.
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include <cstdlib>

#include "engine_common.h"

struct BertLayer {
  Linear qkv, attn_out, inter, out;
  float *ln1_g = nullptr, *ln1_b = nullptr, *ln2_g = nullptr, *ln2_b = nullptr;
};

struct tutor_bert {
  tutor_bert_config cfg{};
  int device = 0, dt = DT_F32;
  int d = 0, H = 0, L = 0, F = 0, V = 0, P = 0;
  std::mutex mu;
  cudaStream_t stream = nullptr;
  void* word = nullptr;
  float *pos = nullptr, *type0 = nullptr, *emb_g = nullptr, *emb_b = nullptr;
  std::vector<BertLayer> layers;
  std::map<std::string, bool> loaded;
  int Mcap = 0;
  int32_t* ids = nullptr;
  float *x = nullptr, *tmp = nullptr, *emb = nullptr;
  bf16* xb = nullptr;
  void *q = nullptr, *kbuf = nullptr, *vbuf = nullptr, *att = nullptr, *hbuf = nullptr;
  int Bcap = 0;
  // One captured CUDA graph per (B, T) shape seen twice: a forward is ~85 launches, and for a single query the host
  // side of those launches (tensor-map encoding + cudaLaunchKernelEx) was the whole 1.0 ms of its latency.
  struct ShapeGraph { cudaGraphExec_t exec = nullptr; uint64_t kernels = 0; int seen = 0; };
  std::map<std::pair<int, int>, ShapeGraph> graphs;
  bool use_graph = true;  // TUTOR_BERT_NO_GRAPH=1: eager launches (A/B measurements)
  // bf16 mode: the activations between the linear layers (dense + residual sums, LayerNorm outputs) live in bf16 only,
  // as in the reference's bf16 modules; statistics and accumulation stay fp32.  125 MB less traffic per sub-layer at
  // 16384 tokens than the fp32 stream (TUTOR_BERT_F32_STREAM=1 keeps that one: cross-check).
  bool bf16_stream = false;
  size_t esz() const { return dt == DT_F32 ? 4 : 2; }
};

namespace {

void drop_graphs(tutor_bert* e) {
  for (auto& g : e->graphs)
    if (g.second.exec) cudaGraphExecDestroy(g.second.exec);
  e->graphs.clear();
}

int ensure_capacity(tutor_bert* e, int B, int T) {
  const int M = B * T;
  if (M > e->Mcap || B > e->Bcap) drop_graphs(e);  // the captured launches hold the old buffer addresses
  if (M > e->Mcap) {
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    dev_free(e->ids); dev_free(e->x); dev_free(e->tmp); dev_free(e->xb); dev_free(e->q); dev_free(e->kbuf);
    dev_free(e->vbuf); dev_free(e->att); dev_free(e->hbuf);
    const size_t d = e->d, es = e->esz();
    TRY(dev_alloc(&e->ids, (size_t)M * 4));
    TRY(dev_alloc(&e->x, (size_t)M * d * 4));
    TRY(dev_alloc(&e->tmp, (size_t)M * d * 4));
    TRY(dev_alloc(&e->xb, (size_t)M * d * 2));
    TRY(dev_alloc(&e->q, (size_t)M * d * es));
    TRY(dev_alloc(&e->kbuf, (size_t)M * d * es));
    TRY(dev_alloc(&e->vbuf, (size_t)M * d * es));
    TRY(dev_alloc(&e->att, (size_t)M * d * es));
    TRY(dev_alloc(&e->hbuf, (size_t)M * e->F * es));
    e->Mcap = M;
  }
  if (B > e->Bcap) {
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    dev_free(e->emb);
    TRY(dev_alloc(&e->emb, (size_t)B * e->d * 4));
    e->Bcap = B;
  }
  return TUTOR_OK;
}

int gemm(tutor_bert* e, const Linear& w, int M, const GemmEpilogue& epi, const void* a_f32, const void* a_bf16) {
  if (e->dt == DT_F32) return k_gemm_f32((const float*)a_f32, (const float*)w.w, M, w.N, w.K, epi, e->stream);
  return k_gemm_bf16((const bf16*)a_bf16, (const bf16*)w.w, M, w.N, w.K, epi, e->stream);
}

// ids already in e->ids; leaves last_hidden_state in e->x and the mean-pooled embedding in e->emb
int forward(tutor_bert* e, int B, int T) {
  const int M = B * T, d = e->d, dt = e->dt;
  cudaStream_t st = e->stream;
  const float eps = e->cfg.layer_norm_eps;
  bf16* xb = dt == DT_BF16 ? e->xb : nullptr;
  // (the bf16 activation stream has no use for the fp32 copy of the embeddings unless they are the output)
  TRY(k_bert_embed_ln(e->ids, T, e->word, dt, e->pos, e->type0, e->emb_g, e->emb_b, eps,
                      e->bf16_stream && e->L > 0 ? nullptr : e->x, xb, M, d, st));
  for (int l = 0; l < e->L; ++l) {
    const BertLayer& w = e->layers[l];
    GemmEpilogue qe{};  // fused query/key/value (modeling_bert.py:158-160, 179-181)
    qe.bias = w.qkv.b;
    qe.out_f32 = dt == DT_F32 ? (float*)e->q : nullptr;
    qe.out_bf16 = dt == DT_BF16 ? (bf16*)e->q : nullptr;
    qe.ldo = d; qe.qkv_mode = 1; qe.kv_dtype = dt; qe.kcache = e->kbuf; qe.vcache = e->vbuf;
    qe.d = d; qe.H = e->H; qe.Smax = T; qe.T = T; qe.pos0 = 0;
    TRY(gemm(e, w.qkv, M, qe, e->x, e->xb));
    TRY(k_attn_prefill(e->q, dt, e->kbuf, e->vbuf, dt, e->att, dt, B, e->H, T, T, /*causal=*/0, nullptr, st));
    const bool bs = e->bf16_stream;
    bf16* tb = reinterpret_cast<bf16*>(e->tmp);
    GemmEpilogue oe{};  // BertSelfOutput: dense + residual, then LayerNorm (modeling_bert.py:294-298)
    oe.bias = w.attn_out.b; oe.ldo = d;
    if (bs) { oe.residual_bf16 = xb; oe.out_bf16 = tb; }
    else { oe.residual = e->x; oe.out_f32 = e->tmp; }
    TRY(gemm(e, w.attn_out, M, oe, e->att, e->att));
    if (bs) TRY(k_layernorm_bf16in(tb, w.ln1_g, w.ln1_b, eps, nullptr, xb, M, d, st));
    else TRY(k_layernorm(e->tmp, 1, 0, w.ln1_g, w.ln1_b, eps, e->x, xb, M, d, st));
    GemmEpilogue ie{};  // BertIntermediate: dense + exact GELU (modeling_bert.py:339-342)
    ie.bias = w.inter.b; ie.act = ACT_GELU_ERF; ie.ldo = e->F;
    ie.out_f32 = dt == DT_F32 ? (float*)e->hbuf : nullptr;
    ie.out_bf16 = dt == DT_BF16 ? (bf16*)e->hbuf : nullptr;
    TRY(gemm(e, w.inter, M, ie, e->x, e->xb));
    GemmEpilogue fe{};  // BertOutput: dense + residual, then LayerNorm (modeling_bert.py:352-356)
    fe.bias = w.out.b; fe.ldo = d;
    if (bs) { fe.residual_bf16 = xb; fe.out_bf16 = tb; }
    else { fe.residual = e->x; fe.out_f32 = e->tmp; }
    TRY(gemm(e, w.out, M, fe, e->hbuf, e->hbuf));
    // the last layer's output is also wanted in fp32: last_hidden_state and the mean-pool read e->x
    if (bs) TRY(k_layernorm_bf16in(tb, w.ln2_g, w.ln2_b, eps, l == e->L - 1 ? e->x : nullptr, xb, M, d, st));
    else TRY(k_layernorm(e->tmp, 1, 0, w.ln2_g, w.ln2_b, eps, e->x, xb, M, d, st));
  }
  return k_meanpool(e->x, e->emb, B, T, d, st);  // unmasked mean over all T positions (REF/lms_server.py:101)
}

// forward() through the shape's captured graph: the first call of a shape runs eagerly (it also sets every kernel
// attribute), the second is captured and replayed from then on.  At most 64 shapes are kept.
int run_forward(tutor_bert* e, int B, int T) {
  if (!e->use_graph) return forward(e, B, T);
  if (e->graphs.size() > 64) drop_graphs(e);
  tutor_bert::ShapeGraph& sg = e->graphs[{B, T}];
  if (sg.exec) {
    CUDA_TRY(cudaGraphLaunch(sg.exec, e->stream));
    tutor_count_launches(sg.kernels);
    return TUTOR_OK;
  }
  if (sg.seen++ == 0) return forward(e, B, T);
  cudaGraph_t g = nullptr;
  const uint64_t before = tutor_launch_count();
  CUDA_TRY(cudaStreamBeginCapture(e->stream, cudaStreamCaptureModeThreadLocal));
  const int rc = forward(e, B, T);
  const cudaError_t ce = cudaStreamEndCapture(e->stream, &g);
  const uint64_t n = tutor_launch_count() - before;
  tutor_count_launches((uint64_t)0 - n);  // captured launches did not run
  if (rc != TUTOR_OK) { if (g) cudaGraphDestroy(g); return rc; }
  if (ce != cudaSuccess) return tutor_cuda_fail(ce, "cudaStreamEndCapture", __FILE__, __LINE__);
  const cudaError_t ci = cudaGraphInstantiate(&sg.exec, g, 0);
  cudaGraphDestroy(g);
  if (ci != cudaSuccess) { sg.exec = nullptr; return tutor_cuda_fail(ci, "cudaGraphInstantiate", __FILE__, __LINE__); }
  sg.kernels = n;
  CUDA_TRY(cudaGraphLaunch(sg.exec, e->stream));
  tutor_count_launches(sg.kernels);
  return TUTOR_OK;
}

// copy a [rows, cols] fp32 host block into rows [row0, row0+rows) of a device matrix of dtype dt
int upload_rows(void* base, int64_t row0, const float* src, int64_t rows, int64_t cols, int dt) {
  void* tmp = nullptr;
  TRY(upload_matrix(&tmp, src, rows, cols, false, dt));
  const size_t es = dt == DT_F32 ? 4 : 2;
  cudaError_t ce = cudaMemcpy((uint8_t*)base + (size_t)row0 * cols * es, tmp, (size_t)rows * cols * es,
                              cudaMemcpyDeviceToDevice);
  cudaFree(tmp);
  if (ce != cudaSuccess) return tutor_cuda_fail(ce, "cudaMemcpy", __FILE__, __LINE__);
  return TUTOR_OK;
}

int check_args(tutor_bert* e, int B, int T) {
  ARG_CHECK(e, "bert: null handle");
  ARG_CHECK(B >= 1 && T >= 1, "bert: empty input (B=%d T=%d)", B, T);
  ARG_CHECK(T <= e->P, "bert: sequence length %d exceeds max_position_embeddings %d", T, e->P);
  return TUTOR_OK;
}

}  // namespace

extern "C" {

int tutor_bert_create(const tutor_bert_config* cfg, int device, tutor_bert_t** out) {
  ARG_CHECK(cfg && out, "bert_create: null argument");
  ARG_CHECK(cfg->hidden_size == cfg->num_attention_heads * 64, "bert_create: head_dim must be 64");
  ARG_CHECK(cfg->hidden_size % 128 == 0 && cfg->intermediate_size % 128 == 0,
            "bert_create: hidden and intermediate sizes must be multiples of 128");
  ARG_CHECK(cfg->dtype == DT_F32 || cfg->dtype == DT_BF16, "bert_create: bad dtype %d", cfg->dtype);
  TRY(tutor_require_device(device));
  tutor_bert* e = new tutor_bert();
  e->cfg = *cfg; e->device = device; e->dt = cfg->dtype;
  e->d = cfg->hidden_size; e->H = cfg->num_attention_heads; e->L = cfg->num_hidden_layers;
  e->F = cfg->intermediate_size; e->V = cfg->vocab_size; e->P = cfg->max_position_embeddings;
  e->layers.resize(e->L);
  e->use_graph = getenv("TUTOR_BERT_NO_GRAPH") == nullptr;
  e->bf16_stream = e->dt == DT_BF16 && !(getenv("TUTOR_BERT_F32_STREAM") && atoi(getenv("TUTOR_BERT_F32_STREAM")));
  if (cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking) != cudaSuccess) {
    int rc = tutor_cuda_fail(cudaGetLastError(), "bert_create", __FILE__, __LINE__);
    delete e;
    return rc;
  }
  *out = e;
  return TUTOR_OK;
}

void tutor_bert_destroy(tutor_bert_t* e) {
  if (!e) return;
  cudaSetDevice(e->device);
  cudaStreamSynchronize(e->stream);
  drop_graphs(e);
  dev_free(e->word); dev_free(e->pos); dev_free(e->type0); dev_free(e->emb_g); dev_free(e->emb_b);
  for (auto& l : e->layers) {
    for (Linear* w : {&l.qkv, &l.attn_out, &l.inter, &l.out}) { dev_free(w->w); dev_free(w->b); }
    dev_free(l.ln1_g); dev_free(l.ln1_b); dev_free(l.ln2_g); dev_free(l.ln2_b);
  }
  dev_free(e->ids); dev_free(e->x); dev_free(e->tmp); dev_free(e->emb); dev_free(e->xb); dev_free(e->q);
  dev_free(e->kbuf); dev_free(e->vbuf); dev_free(e->att); dev_free(e->hbuf);
  cudaStreamDestroy(e->stream);
  delete e;
}

int tutor_bert_load_tensor(tutor_bert_t* e, const char* name, const float* data, const int64_t* shape, int ndim) {
  ARG_CHECK(e && name && data && shape, "bert_load_tensor: null argument");
  std::lock_guard<std::mutex> lock(e->mu);
  CUDA_TRY(cudaSetDevice(e->device));
  // weight buffers are about to be replaced: captured forwards hold their addresses
  CUDA_TRY(cudaStreamSynchronize(e->stream));
  drop_graphs(e);
  std::string n(name);
  if (n.rfind("bert.", 0) == 0) n = n.substr(5);
  if (n.rfind("pooler.", 0) == 0 || n == "embeddings.position_ids" || n == "embeddings.token_type_ids")
    return TUTOR_OK;  // pooler output is never read by the reference; the id buffers are not weights
  const int d = e->d;
  auto want = [&](int64_t a, int64_t b) -> int {
    bool ok = (b < 0) ? (ndim == 1 && shape[0] == a) : (ndim == 2 && shape[0] == a && shape[1] == b);
    ARG_CHECK(ok, "bert_load_tensor: %s has the wrong shape", name);
    return TUTOR_OK;
  };
  int rc = TUTOR_ERR_ARG;
  if (n == "embeddings.word_embeddings.weight") { TRY(want(e->V, d)); rc = upload_matrix(&e->word, data, e->V, d, false, e->dt); }
  else if (n == "embeddings.position_embeddings.weight") { TRY(want(e->P, d)); rc = upload_f32(&e->pos, data, (size_t)e->P * d); }
  else if (n == "embeddings.token_type_embeddings.weight") {
    ARG_CHECK(ndim == 2 && shape[0] >= 1 && shape[1] == d, "bert_load_tensor: %s has the wrong shape", name);
    rc = upload_f32(&e->type0, data, d);  // token_type_ids are all zero on this path: row 0 only
  }
  else if (n == "embeddings.LayerNorm.weight") { TRY(want(d, -1)); rc = upload_f32(&e->emb_g, data, d); }
  else if (n == "embeddings.LayerNorm.bias") { TRY(want(d, -1)); rc = upload_f32(&e->emb_b, data, d); }
  else if (n.rfind("encoder.layer.", 0) == 0) {
    size_t dot = n.find('.', 14);
    int li = atoi(n.substr(14, dot - 14).c_str());
    ARG_CHECK(li >= 0 && li < e->L, "bert_load_tensor: layer out of range in %s", name);
    BertLayer& l = e->layers[li];
    std::string r = n.substr(dot + 1);
    const char* qkv_names[3] = {"attention.self.query", "attention.self.key", "attention.self.value"};
    bool found = false;
    for (int i = 0; i < 3 && !found; ++i) {
      if (r == std::string(qkv_names[i]) + ".weight") {
        TRY(want(d, d));
        if (!l.qkv.w) { TRY(dev_alloc(&l.qkv.w, (size_t)3 * d * d * e->esz())); l.qkv.N = 3 * d; l.qkv.K = d; }
        rc = upload_rows(l.qkv.w, (int64_t)i * d, data, d, d, e->dt);
        found = true;
      } else if (r == std::string(qkv_names[i]) + ".bias") {
        TRY(want(d, -1));
        if (!l.qkv.b) TRY(dev_alloc(&l.qkv.b, (size_t)3 * d * 4));
        CUDA_TRY(cudaMemcpy(l.qkv.b + (size_t)i * d, data, (size_t)d * 4, cudaMemcpyHostToDevice));
        rc = TUTOR_OK; found = true;
      }
    }
    struct { const char* n; Linear* lin; int N, K; } lins[] = {{"attention.output.dense", &l.attn_out, d, d},
                                                               {"intermediate.dense", &l.inter, e->F, d},
                                                               {"output.dense", &l.out, d, e->F}};
    for (auto& x : lins) {
      if (r == std::string(x.n) + ".weight") {
        TRY(want(x.N, x.K));  // nn.Linear stores [out,in]: already K-contiguous
        rc = upload_matrix(&x.lin->w, data, x.N, x.K, false, e->dt);
        x.lin->N = x.N; x.lin->K = x.K; found = true;
      } else if (r == std::string(x.n) + ".bias") {
        TRY(want(x.N, -1));
        rc = upload_f32(&x.lin->b, data, x.N); found = true;
      }
    }
    struct { const char* n; float** p; } vecs[] = {{"attention.output.LayerNorm.weight", &l.ln1_g},
                                                   {"attention.output.LayerNorm.bias", &l.ln1_b},
                                                   {"output.LayerNorm.weight", &l.ln2_g},
                                                   {"output.LayerNorm.bias", &l.ln2_b}};
    for (auto& v : vecs)
      if (r == v.n) { TRY(want(d, -1)); rc = upload_f32(v.p, data, d); found = true; }
    if (!found) { tutor_set_error("bert_load_tensor: unknown tensor %s", name); return TUTOR_ERR_ARG; }
  } else {
    tutor_set_error("bert_load_tensor: unknown tensor %s", name);
    return TUTOR_ERR_ARG;
  }
  if (rc == TUTOR_OK) e->loaded[n] = true;
  return rc;
}

int tutor_bert_check_loaded(tutor_bert_t* e) {
  ARG_CHECK(e, "bert: null handle");
  std::vector<std::string> need = {"embeddings.word_embeddings.weight", "embeddings.position_embeddings.weight",
                                   "embeddings.token_type_embeddings.weight", "embeddings.LayerNorm.weight",
                                   "embeddings.LayerNorm.bias"};
  const char* per[] = {"attention.self.query", "attention.self.key", "attention.self.value", "attention.output.dense",
                       "intermediate.dense", "output.dense"};
  for (int l = 0; l < e->L; ++l) {
    std::string p = "encoder.layer." + std::to_string(l) + ".";
    for (const char* n : per) { need.push_back(p + n + ".weight"); need.push_back(p + n + ".bias"); }
    for (const char* n : {"attention.output.LayerNorm", "output.LayerNorm"}) {
      need.push_back(p + n + ".weight"); need.push_back(p + n + ".bias");
    }
  }
  for (auto& n : need)
    if (!e->loaded.count(n)) {
      tutor_set_error("bert: weight %s has not been loaded", n.c_str());
      return TUTOR_ERR_STATE;
    }
  return TUTOR_OK;
}

int tutor_bert_embed(tutor_bert_t* e, const int64_t* ids_host, int B, int T, float* emb_host, float* hidden_host) {
  TRY(check_args(e, B, T));
  ARG_CHECK(ids_host && emb_host, "bert_embed: null argument");
  std::lock_guard<std::mutex> lock(e->mu);
  CUDA_TRY(cudaSetDevice(e->device));
  TRY(tutor_bert_check_loaded(e));
  TRY(ensure_capacity(e, B, T));
  std::vector<int32_t> tmp((size_t)B * T);
  for (size_t i = 0; i < tmp.size(); ++i) {
    ARG_CHECK(ids_host[i] >= 0 && ids_host[i] < e->V, "bert_embed: token id %lld outside [0,%d)",
              (long long)ids_host[i], e->V);
    tmp[i] = (int32_t)ids_host[i];
  }
  CUDA_TRY(cudaMemcpyAsync(e->ids, tmp.data(), tmp.size() * 4, cudaMemcpyHostToDevice, e->stream));
  TRY(run_forward(e, B, T));
  CUDA_TRY(cudaMemcpyAsync(emb_host, e->emb, (size_t)B * e->d * 4, cudaMemcpyDeviceToHost, e->stream));
  if (hidden_host)
    CUDA_TRY(cudaMemcpyAsync(hidden_host, e->x, (size_t)B * T * e->d * 4, cudaMemcpyDeviceToHost, e->stream));
  CUDA_TRY(cudaStreamSynchronize(e->stream));
  return TUTOR_OK;
}

int tutor_bert_embed_dev(tutor_bert_t* e, const int32_t* ids_dev, int B, int T, float* emb_dev) {
  TRY(check_args(e, B, T));
  ARG_CHECK(ids_dev && emb_dev, "bert_embed_dev: null argument");
  std::lock_guard<std::mutex> lock(e->mu);
  CUDA_TRY(cudaSetDevice(e->device));
  TRY(tutor_bert_check_loaded(e));
  TRY(ensure_capacity(e, B, T));
  CUDA_TRY(cudaMemcpyAsync(e->ids, ids_dev, (size_t)B * T * 4, cudaMemcpyDeviceToDevice, e->stream));
  TRY(run_forward(e, B, T));
  CUDA_TRY(cudaMemcpyAsync(emb_dev, e->emb, (size_t)B * e->d * 4, cudaMemcpyDeviceToDevice, e->stream));
  CUDA_TRY(cudaStreamSynchronize(e->stream));
  return TUTOR_OK;
}

void* tutor_bert_stream(tutor_bert_t* e) { return e ? (void*)e->stream : nullptr; }

}  // extern "C"
