// Where does the row-major (prefill / BERT) tcgen05 GEMM spend its time?  Same shapes as one BERT-base layer at
// M = 16384 tokens, run (a) as shipped, (b) with the epilogue reduced to handing the accumulator back (mainloop only:
// TMA + MMA), so the difference is the epilogue's share.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -DTUTOR_TRACE -I distributed-lms-raft-llm_b200/csrc \
//        tools/microbench_gemm.cu -o /tmp/mbg -L distributed-lms-raft-llm_b200 -ltutor_b200 -lcuda
#include <cstdio>
#include <cstdlib>

#include "kernels.h"
#include "gemm_bf16_tc.cu"

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA %s at %d\n", cudaGetErrorString(e_), __LINE__); exit(1); } } while (0)
extern "C" const char* tutor_last_error(void);
int tutor_set_max_smem(const void* func, int bytes) {
  CK(cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  return 0;
}

int main() {
  cudaStream_t st; CK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
  const int M = 16384;
  bf16 *a, *w, *ob; float *of, *res, *bias;
  CK(cudaMalloc(&a, (size_t)M * 3072 * 2)); CK(cudaMalloc(&w, (size_t)3072 * 3072 * 2)); CK(cudaMalloc(&ob, (size_t)M * 3072 * 2));
  CK(cudaMalloc(&of, (size_t)M * 3072 * 4)); CK(cudaMalloc(&res, (size_t)M * 3072 * 4)); CK(cudaMalloc(&bias, 3072 * 4));
  CK(cudaMemset(a, 0, (size_t)M * 3072 * 2)); CK(cudaMemset(w, 0, (size_t)3072 * 3072 * 2)); CK(cudaMemset(res, 0, (size_t)M * 3072 * 4));
  CK(cudaMemset(bias, 0, 3072 * 4));
  struct { const char* n; int N, K, act, resid, f32out; } shapes[] = {
      {"qkv      N=2304 K=768  bf16 out", 2304, 768, ACT_NONE, 0, 0},
      {"attn_out N=768  K=768  +res f32", 768, 768, ACT_NONE, 1, 1},
      {"inter    N=3072 K=768  gelu bf16", 3072, 768, ACT_GELU_ERF, 0, 0},
      {"out      N=768  K=3072 +res f32", 768, 3072, ACT_NONE, 1, 1}};
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  const char* names[] = {"as shipped", "mainloop only (epilogue hands the accumulator straight back)",
                         "epilogue without global loads / stores (null output pointers)", "epilogue without the TMEM read", "TMEM read only (no epilogue math, no global traffic)"};
  for (int skip = 0; skip < 5; ++skip) {
    const int dev_skip = skip == 2 ? 0 : skip;
    CK(cudaMemcpyToSymbol(g_epi_skip, &dev_skip, sizeof(int)));
    printf("--- %s\n", names[skip]);
    for (auto& s : shapes) {
      GemmEpilogue e{};
      e.bias = bias; e.act = s.act; e.ldo = s.N;
      if (s.resid && skip != 2) e.residual = res;
      if (skip != 2) { if (s.f32out) e.out_f32 = of; else e.out_bf16 = ob; }
      for (int i = 0; i < 3; ++i) if (k_gemm_bf16(a, w, M, s.N, s.K, e, st)) { printf("error: %s\n", tutor_last_error()); return 2; }
      CK(cudaStreamSynchronize(st));
      CK(cudaEventRecord(e0, st));
      const int reps = 20;
      for (int i = 0; i < reps; ++i) k_gemm_bf16(a, w, M, s.N, s.K, e, st);
      CK(cudaEventRecord(e1, st));
      CK(cudaStreamSynchronize(st));
      float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
      const double us = ms * 1e3 / reps, tf = 2.0 * M * s.N * s.K / us * 1e-6;
      const double opnd_mb = (double)((M + 127) / 128) * ((s.N + 255) / 256) * (128 + 256) * s.K * 2 / 1e6;
      printf("%s  %7.1f us  %6.1f TFLOP/s   operand tiles %6.0f MB -> %5.2f TB/s into shared memory\n", s.n, us, tf, opnd_mb,
             opnd_mb / us);
    }
  }
  return 0;
}
