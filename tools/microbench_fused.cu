// Per-kernel cost and phase timeline (cycles, CTA 0) of the fused decode linears inside a CUDA graph of dependent
// launches, with and without programmatic dependent launch; weights rotate over 12 "layers" so they come from HBM.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -DTUTOR_TRACE -I distributed-lms-raft-llm_b200/csrc \
//        tools/microbench_fused.cu -o gpurun_out/microbench_fused -L distributed-lms-raft-llm_b200 -ltutor_b200 -lcuda
#include <cstdio>
#include <functional>

#include "kernels.h"
#include "gemm_decode_fused.cu"

extern "C" const char* tutor_last_error(void);
#define DIE() do { printf("error: %s\n", tutor_last_error()); exit(2); } while (0)
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA %s at %d\n", cudaGetErrorString(e_), __LINE__); exit(1); } } while (0)

// the kernels of this translation unit belong to this executable's CUDA runtime, not the library's
int tutor_set_max_smem(const void* func, int bytes) {
  CK(cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  return 0;
}

static float time_graph(cudaStream_t st, int chain, bool pdl, const std::function<void(int)>& launch_one) {
  cudaGraph_t g; cudaGraphExec_t ge;
  launch_one(0);  // eager once: sets the function attributes outside the capture
  CK(cudaStreamSynchronize(st));
  tutor_pdl_enabled = pdl;
  CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
  for (int i = 0; i < chain; ++i) launch_one(i);
  CK(cudaStreamEndCapture(st, &g));
  tutor_pdl_enabled = false;
  CK(cudaGraphInstantiate(&ge, g, 0));
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  for (int i = 0; i < 3; ++i) CK(cudaGraphLaunch(ge, st));
  CK(cudaStreamSynchronize(st));
  const int reps = 20;
  CK(cudaEventRecord(e0, st));
  for (int i = 0; i < reps; ++i) CK(cudaGraphLaunch(ge, st));
  CK(cudaEventRecord(e1, st));
  CK(cudaStreamSynchronize(st));
  float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
  cudaGraphExecDestroy(ge); cudaGraphDestroy(g);
  return ms * 1e3f / (reps * chain);
}

static void show_trace() {
  long long tr[16];
  CK(cudaMemcpyFromSymbol(tr, g_fused_trace, sizeof(tr)));
  const char* names[] = {"start", "setup", "wait done", "stats", "B written", "tmem_full", "staged", "csync1", "reduced", "csync2", "end"};
  printf("     CTA0 cycles:");
  for (int i = 1; i <= 10; ++i) if (tr[i] > tr[0]) printf(" %s +%lld |", names[i], tr[i] - tr[0]);
  printf("\n");
  long long z[16] = {0};
  CK(cudaMemcpyToSymbol(g_fused_trace, z, sizeof(z)));
}

int main() {
  cudaStream_t st; CK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
  const int B = 64, d = 768;
  bf16 *a, *w, *h; float *partial, *xf, *bias, *gam, *stats;
  CK(cudaMalloc(&a, (size_t)B * 4 * d * 2)); CK(cudaMalloc(&h, (size_t)B * 4 * d * 2));
  CK(cudaMalloc(&w, (size_t)12 * 4 * d * d * 2));
  CK(cudaMalloc(&partial, (size_t)12 * B * 4 * d * 4)); CK(cudaMalloc(&xf, (size_t)B * d * 4));
  CK(cudaMalloc(&bias, 4 * d * 4)); CK(cudaMalloc(&gam, 4 * d * 4)); CK(cudaMalloc(&stats, 6 * B * 2 * 4));
  CK(cudaMemset(a, 0, (size_t)B * 4 * d * 2)); CK(cudaMemset(w, 0, (size_t)12 * 4 * d * d * 2));
  CK(cudaMemset(xf, 0, (size_t)B * d * 4)); CK(cudaMemset(bias, 0, 4 * d * 4)); CK(cudaMemset(gam, 0, 4 * d * 4));
  CK(cudaMemset(stats, 0, 6 * B * 2 * 4));
  auto W = [&](int i) { return w + (size_t)(i % 12) * 4 * d * d; };
  for (int pdl = 0; pdl < 2; ++pdl) {
    printf("--- PDL %s\n", pdl ? "on" : "off");
    float t;
    for (int ks : {6, 12}) {
      t = time_graph(st, 96, pdl, [&](int i) { if (k_gemm_ln_splitk(xf, stats, 6, gam, bias, 1e-5f, W(i), B, 3 * d, d, ks, 64, partial, st)) DIE(); });
      printf("c_attn LN -> partials   ks=%2d         %6.2f us/kernel\n", ks, t); show_trace();
    }
    for (auto cfg : {std::pair<int,int>{6, 64}, {3, 32}, {2, 16}}) {
      t = time_graph(st, 96, pdl, [&](int i) { if (k_gemm_ln_act(xf, stats, 6, gam, bias, 1e-5f, W(i), bias, ACT_GELU_TANH, B, 4 * d, d, cfg.first, cfg.second, h, st)) DIE(); });
      printf("c_fc LN -> cluster gelu ks=%2d bn=%3d  %6.2f us/kernel\n", cfg.first, cfg.second, t); show_trace();
    }
    for (auto cfg : {std::pair<int,int>{6, 32}, {6, 16}}) {
      t = time_graph(st, 96, pdl, [&](int i) { if (k_gemm_resid_stats(a, W(i), bias, B, d, d, cfg.first, cfg.second, xf, stats, nullptr, st)) DIE(); });
      printf("proj -> cluster resid   ks=%2d bn=%3d  %6.2f us/kernel\n", cfg.first, cfg.second, t); show_trace();
    }
    for (auto cfg : {std::pair<int,int>{8, 32}, {8, 16}}) {
      t = time_graph(st, 96, pdl, [&](int i) { if (k_gemm_resid_stats(h, W(i), bias, B, d, 4 * d, cfg.first, cfg.second, xf, stats, nullptr, st)) DIE(); });
      printf("proj2 -> cluster resid  ks=%2d bn=%3d  %6.2f us/kernel\n", cfg.first, cfg.second, t); show_trace();
    }
    // the layer's GEMM chain as the engine issues it (attention left out)
    t = time_graph(st, 96, pdl, [&](int i) {
      int rc = 0;
      switch (i & 3) {
        case 0: rc = k_gemm_ln_splitk(xf, stats, 6, gam, bias, 1e-5f, W(i), B, 3 * d, d, 6, 64, partial, st); break;
        case 1: rc = k_gemm_resid_stats(a, W(i), bias, B, d, d, 6, 32, xf, stats, nullptr, st); break;
        case 2: rc = k_gemm_ln_act(xf, stats, 6, gam, bias, 1e-5f, W(i), bias, ACT_GELU_TANH, B, 4 * d, d, 6, 64, h, st); break;
        default: rc = k_gemm_resid_stats(h, W(i), bias, B, d, 4 * d, 8, 32, xf, stats, nullptr, st);
      }
      if (rc) DIE();
    });
    printf("chain qkv/proj/fc/proj2                %6.2f us/kernel\n", t);
  }
  return 0;
}
