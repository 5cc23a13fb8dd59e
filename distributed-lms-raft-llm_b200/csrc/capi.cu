// C-ABI glue: error text, launch counter, weight upload helpers and the single-kernel entry points
// declared in include/tutor_b200.h.
#include <atomic>
#include <cstring>
#include <mutex>
#include <vector>

#include "engine_common.h"

thread_local bool tutor_pdl_enabled = false;

namespace {
thread_local char g_err[1024] = "";
std::atomic<uint64_t> g_launches{0};

inline uint16_t f32_to_bf16_rne(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);  // NaN stays NaN
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}
}  // namespace

void tutor_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int tutor_cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
  cudaGetLastError();  // a failed call (e.g. an allocation that did not fit) must not resurface at the next launch check
  const char* base = strrchr(file, '/');
  tutor_set_error("CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), base ? base + 1 : file, line, what);
  return TUTOR_ERR_CUDA;
}

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-device attribute: remember (function, device) pairs so that a
// process driving several GPUs sets it on each of them, and only once
int tutor_set_max_smem(const void* func, int bytes) {
  static std::mutex mu;
  static std::vector<std::pair<const void*, int>> done;
  int dev = 0;
  CUDA_TRY(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lock(mu);
  for (auto& p : done)
    if (p.first == func && p.second == dev) return TUTOR_OK;
  CUDA_TRY(cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  done.emplace_back(func, dev);
  return TUTOR_OK;
}

void tutor_count_launches(uint64_t n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

int tutor_require_device(int device) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) {
    tutor_set_error("no CUDA device is available (%s); this library has no CPU fallback",
                    e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
    return TUTOR_ERR_CUDA;
  }
  ARG_CHECK(device >= 0 && device < n, "device %d out of range (have %d)", device, n);
  // the architecture check is remembered per device: cudaGetDeviceProperties costs ~2 ms a call, which was a third of a
  // host-path relevance batch (tutor_cosine_pairs_host comes through here on every request)
  static std::atomic<int> checked[64];
  if (device >= 64 || checked[device].load(std::memory_order_relaxed) == 0) {
    int major = 0, minor = 0;
    CUDA_TRY(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
    CUDA_TRY(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, device));
    if (major != 10) {
      tutor_set_error("device %d is sm_%d%d; this library is built for sm_100a (B200) only", device, major, minor);
      return TUTOR_ERR_CUDA;
    }
    if (device < 64) checked[device].store(1, std::memory_order_relaxed);
  }
  CUDA_TRY(cudaSetDevice(device));
  return TUTOR_OK;
}

int upload_f32(float** dst, const float* src, size_t n) {
  dev_free(*dst);
  TRY(dev_alloc(dst, n * sizeof(float)));
  CUDA_TRY(cudaMemcpy(*dst, src, n * sizeof(float), cudaMemcpyHostToDevice));
  return TUTOR_OK;
}

int upload_matrix(void** dst, const float* src, int64_t rows, int64_t cols, bool transpose, int dt) {
  const size_t n = (size_t)rows * cols;
  dev_free(*dst);
  std::vector<float> t;
  const float* s = src;
  if (transpose) {
    t.resize(n);
    constexpr int64_t TB = 32;  // blocked so both sides stay cache-friendly on the host
    for (int64_t r0 = 0; r0 < rows; r0 += TB)
      for (int64_t c0 = 0; c0 < cols; c0 += TB)
        for (int64_t r = r0; r < rows && r < r0 + TB; ++r)
          for (int64_t c = c0; c < cols && c < c0 + TB; ++c) t[(size_t)c * rows + r] = src[(size_t)r * cols + c];
    s = t.data();
  }
  if (dt == DT_F32) {
    TRY(dev_alloc(dst, n * 4));
    CUDA_TRY(cudaMemcpy(*dst, s, n * 4, cudaMemcpyHostToDevice));
  } else {
    std::vector<uint16_t> h(n);
    for (size_t i = 0; i < n; ++i) h[i] = f32_to_bf16_rne(s[i]);
    TRY(dev_alloc(dst, n * 2));
    CUDA_TRY(cudaMemcpy(*dst, h.data(), n * 2, cudaMemcpyHostToDevice));
  }
  return TUTOR_OK;
}

// LayerNorm folded into the Conv1D that consumes it (decode_mega.cu): src is the Conv1D weight [K, N] ([in, out]),
// gamma / beta the LayerNorm parameters [K].  Uploads W'[n,k] = bf16(gamma[k] * src[k,n]) as [N, K], s[n] = sum_k W'[n,k]
// (of the rounded values, so that subtracting mean * s removes exactly what the tensor core accumulated) and
// c[n] = sum_k beta[k] * src[k,n].
int upload_ln_folded(void** dst_w, float** dst_s, float** dst_c, const float* src, int64_t K, int64_t N, const float* gamma,
                     const float* beta) {
  std::vector<uint16_t> w((size_t)N * K);
  std::vector<double> s(N, 0.0), c(N, 0.0);
  constexpr int64_t TB = 32;
  for (int64_t k0 = 0; k0 < K; k0 += TB)
    for (int64_t n0 = 0; n0 < N; n0 += TB)
      for (int64_t k = k0; k < K && k < k0 + TB; ++k)
        for (int64_t n = n0; n < N && n < n0 + TB; ++n) {
          const float v = src[(size_t)k * N + n];
          const uint16_t h = f32_to_bf16_rne(gamma[k] * v);
          w[(size_t)n * K + k] = h;
          const uint32_t bits = (uint32_t)h << 16;
          float f;
          memcpy(&f, &bits, 4);
          s[n] += f;
          c[n] += (double)beta[k] * v;
        }
  std::vector<float> sf(N), cf(N);
  for (int64_t n = 0; n < N; ++n) { sf[n] = (float)s[n]; cf[n] = (float)c[n]; }
  dev_free(*dst_w);
  TRY(dev_alloc(dst_w, w.size() * 2));
  CUDA_TRY(cudaMemcpy(*dst_w, w.data(), w.size() * 2, cudaMemcpyHostToDevice));
  TRY(upload_f32(dst_s, sf.data(), (size_t)N));
  return upload_f32(dst_c, cf.data(), (size_t)N);
}

extern "C" {

const char* tutor_last_error(void) { return g_err; }
int tutor_version(void) { return 100; }
uint64_t tutor_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

int tutor_layernorm(const float* x, const float* gamma, const float* beta, float eps, float* out_f32, void* out_bf16,
                    int M, int d, void* stream) {
  ARG_CHECK(x && gamma && beta && (out_f32 || out_bf16), "layernorm: null argument");
  return k_layernorm(x, 1, 0, gamma, beta, eps, out_f32, (bf16*)out_bf16, M, d, (cudaStream_t)stream);
}

int tutor_linear(int dtype, const void* A, const void* W, const float* bias, const float* residual, void* out,
                 int out_is_f32, int M, int N, int K, int act, int impl, void* stream) {
  ARG_CHECK(A && W && out, "linear: null argument");
  GemmEpilogue e{};
  e.bias = bias; e.residual = residual; e.act = act; e.ldo = N;
  if (dtype == DT_F32) {
    e.out_f32 = (float*)out;
    return k_gemm_f32((const float*)A, (const float*)W, M, N, K, e, (cudaStream_t)stream);
  }
  ARG_CHECK(dtype == DT_BF16, "linear: bad dtype %d", dtype);
  if (out_is_f32) e.out_f32 = (float*)out; else e.out_bf16 = (bf16*)out;
  if (impl == 1) return k_gemm_bf16_legacy((const bf16*)A, (const bf16*)W, M, N, K, e, (cudaStream_t)stream);
  return k_gemm_bf16((const bf16*)A, (const bf16*)W, M, N, K, e, (cudaStream_t)stream);
}

int tutor_linear_bf16_residual(const void* A, const void* W, const float* bias, const void* residual_bf16, void* out_bf16,
                               int M, int N, int K, int act, void* stream) {
  ARG_CHECK(A && W && out_bf16, "linear_bf16_residual: null argument");
  GemmEpilogue e{};
  e.bias = bias; e.residual_bf16 = (const bf16*)residual_bf16; e.out_bf16 = (bf16*)out_bf16; e.act = act; e.ldo = N;
  return k_gemm_bf16((const bf16*)A, (const bf16*)W, M, N, K, e, (cudaStream_t)stream);
}

int tutor_layernorm_bf16(const void* x, const float* gamma, const float* beta, float eps, float* out_f32, void* out_bf16,
                         int M, int d, void* stream) {
  ARG_CHECK(x && gamma && beta && (out_f32 || out_bf16), "layernorm_bf16: null argument");
  return k_layernorm_bf16in((const bf16*)x, gamma, beta, eps, out_f32, (bf16*)out_bf16, M, d, (cudaStream_t)stream);
}

int tutor_attn_prefill(int dtype, const void* q, const void* k, const void* v, void* out, int B, int H, int T, int Smax,
                       int causal, void* stream) {
  ARG_CHECK(q && k && v && out, "attn_prefill: null argument");
  return k_attn_prefill(q, dtype, k, v, dtype, out, dtype, B, H, T, Smax, causal, nullptr, (cudaStream_t)stream);
}

int tutor_attn_decode(int dtype, const void* q, const void* k, const void* v, void* out, int B, int H, int Smax, int len,
                      void* stream) {
  ARG_CHECK(q && k && v && out, "attn_decode: null argument");
  return k_attn_decode(q, dtype, k, v, dtype, out, dtype, B, H, Smax, nullptr, len, nullptr, (cudaStream_t)stream);
}

int tutor_attn_decode_fused(const float* qkv_partial, int ks, const float* qkv_bias, void* k, void* v, void* out, int B,
                            int H, int Smax, int len, void* stream) {
  ARG_CHECK(qkv_partial && qkv_bias && k && v && out, "attn_decode_fused: null argument");
  ARG_CHECK(len >= 1 && len <= Smax, "attn_decode_fused: len %d outside [1,%d]", len, Smax);
  const int safe = ((len - 1 < 128 ? len - 1 : 128) / 64) * 64;  // what the engine allows to stream ahead of the wait
  return k_attn_decode_fused(qkv_partial, ks, qkv_bias, QkvFold{}, k, v, DT_BF16, out, DT_BF16, B, H, Smax, nullptr, len, 1, nullptr,
                             nullptr, safe, len, (cudaStream_t)stream);
}

size_t tutor_lmhead_scratch_bytes(int B) { return k_lmhead_scratch_bytes(B, 0); }

int tutor_lmhead_argmax(int dtype, const void* x, const void* wte, int B, int V, int d, const uint32_t* seen,
                        float penalty, int32_t* out_idx, void* scratch, void* stream) {
  ARG_CHECK(x && wte && out_idx && scratch, "lmhead_argmax: null argument");
  ARG_CHECK(penalty == 1.0f || seen, "lmhead_argmax: a penalty needs the seen bitmap");
  DecodeState ds{};
  ds.seen = const_cast<uint32_t*>(seen);
  ds.seen_words = ceil_div(V, 32);
  ds.penalty = penalty;
  ds.plain_out = out_idx;
  return k_lmhead_argmax(x, dtype, wte, dtype, B, V, d, ds, scratch, (cudaStream_t)stream);
}

// ---- decode-step linears with fused LayerNorm operand / cluster split-K reduction (gemm_decode_fused.cu)
namespace {
int sm_count() {
  int dev = 0, sms = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  return sms;
}
}  // namespace

int tutor_row_tile_stats(const float* x, float* stats, int M, int d, void* stream) {
  ARG_CHECK(x && stats, "row_tile_stats: null argument");
  return k_row_tile_stats(x, stats, M, d, nullptr, (cudaStream_t)stream);
}

int tutor_decode_ln_linear_partial(const float* x, const float* stats, const float* gamma, const float* beta, float eps,
                                   const void* W, int M, int N, int K, int ks, float* partial, void* stream) {
  ARG_CHECK(x && stats && gamma && beta && W && partial && K % 128 == 0, "decode_ln_linear_partial: bad argument");
  if (ks <= 0) ks = k_pick_splitk(N, K, sm_count());
  return k_gemm_ln_splitk(x, stats, K / 128, gamma, beta, eps, (const bf16*)W, M, N, K, ks, k_decode_batch_tile(M),
                          partial, (cudaStream_t)stream);
}

int tutor_decode_ln_linear_act(const float* x, const float* stats, const float* gamma, const float* beta, float eps,
                               const void* W, const float* bias, int act, int M, int N, int K, int ks, int bn,
                               void* out_bf16, void* stream) {
  ARG_CHECK(x && stats && gamma && beta && W && bias && out_bf16 && K % 128 == 0, "decode_ln_linear_act: bad argument");
  if (bn <= 0) bn = k_decode_batch_tile(M);
  if (ks <= 0) ks = k_pick_cluster_splitk(N, K, ceil_div(M, bn), sm_count());
  return k_gemm_ln_act(x, stats, K / 128, gamma, beta, eps, (const bf16*)W, bias, act, M, N, K, ks, bn, (bf16*)out_bf16,
                       (cudaStream_t)stream);
}

int tutor_decode_linear_residual(const void* A, const void* W, const float* bias, int M, int N, int K, int ks, int bn,
                                 float* x, float* stats, void* stream) {
  ARG_CHECK(A && W && bias && x && stats, "decode_linear_residual: null argument");
  if (bn <= 0) bn = k_decode_batch_tile(M);
  if (ks <= 0) ks = k_pick_cluster_splitk(N, K, ceil_div(M, bn), sm_count());
  return k_gemm_resid_stats((const bf16*)A, (const bf16*)W, bias, M, N, K, ks, bn, x, stats, nullptr, (cudaStream_t)stream);
}

int tutor_meanpool(const float* x, float* out, int B, int T, int d, void* stream) {
  ARG_CHECK(x && out && T >= 1, "meanpool: bad argument");
  return k_meanpool(x, out, B, T, d, (cudaStream_t)stream);
}

int tutor_cosine_pairs(const float* a, const float* b, float* out, int B, int d, void* stream) {
  ARG_CHECK(a && b && out, "cosine_pairs: null argument");
  return k_cosine_pairs(a, b, out, B, d, 1e-8f, (cudaStream_t)stream);
}

// grow-only device staging for the host-pointer cosine (cudaMalloc/cudaFree per call cost ~14 ms per relevance batch)
namespace {
std::mutex g_cos_mu;
float* g_cos_buf = nullptr;
size_t g_cos_floats = 0;
int g_cos_dev = -1;
}  // namespace

int tutor_cosine_pairs_host(const float* a, const float* b, float* out, int B, int d) {
  ARG_CHECK(a && b && out && B >= 1 && d >= 1, "cosine_pairs_host: bad argument");
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) dev = 0;
  TRY(tutor_require_device(dev));
  std::lock_guard<std::mutex> lock(g_cos_mu);
  const size_t n = (size_t)B * d, need = 2 * n + B;
  if (need > g_cos_floats || dev != g_cos_dev) {
    if (g_cos_buf) cudaFree(g_cos_buf);
    g_cos_buf = nullptr; g_cos_floats = 0;
    TRY(dev_alloc(&g_cos_buf, need * sizeof(float)));
    g_cos_floats = need; g_cos_dev = dev;
  }
  float *da = g_cos_buf, *db = g_cos_buf + n, *dout = g_cos_buf + 2 * n;
  CUDA_TRY(cudaMemcpyAsync(da, a, n * sizeof(float), cudaMemcpyHostToDevice, nullptr));
  CUDA_TRY(cudaMemcpyAsync(db, b, n * sizeof(float), cudaMemcpyHostToDevice, nullptr));
  TRY(k_cosine_pairs(da, db, dout, B, d, 1e-8f, nullptr));
  CUDA_TRY(cudaMemcpyAsync(out, dout, B * sizeof(float), cudaMemcpyDeviceToHost, nullptr));
  CUDA_TRY(cudaStreamSynchronize(nullptr));
  return TUTOR_OK;
}

int tutor_row_inv_norm(const float* corpus, float* inv_norm, int N, int d, void* stream) {
  ARG_CHECK(corpus && inv_norm, "row_inv_norm: null argument");
  return k_row_inv_norm(corpus, DT_F32, inv_norm, N, d, 1e-8f, (cudaStream_t)stream);
}

size_t tutor_cosine_topk_scratch_bytes(int Q, int N, int k) { return k_cosine_topk_scratch_bytes(Q, N, k); }

int tutor_cosine_topk(const float* queries, const float* corpus, const float* inv_norm, int Q, int N, int d, int k,
                      float* out_scores, int32_t* out_idx, int idx_base, void* scratch, void* stream) {
  ARG_CHECK(queries && corpus && inv_norm && out_scores && out_idx && scratch, "cosine_topk: null argument");
  return k_cosine_topk(queries, corpus, DT_F32, inv_norm, Q, N, d, k, out_scores, out_idx, idx_base, scratch,
                       (cudaStream_t)stream);
}

int tutor_topk_merge(const float* cand_scores, const int32_t* cand_idx, int Q, int ncand, int k, float* out_scores,
                     int32_t* out_idx, void* stream) {
  ARG_CHECK(cand_scores && cand_idx && out_scores && out_idx, "topk_merge: null argument");
  return k_topk_merge(cand_scores, cand_idx, Q, ncand, k, out_scores, out_idx, (cudaStream_t)stream);
}

}  // extern "C"
