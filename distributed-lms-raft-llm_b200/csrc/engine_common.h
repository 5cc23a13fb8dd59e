// Host-side helpers shared by the GPT-2 and BERT engines.
#pragma once
#include <string>
#include <vector>

#include "../../include/tutor_b200.h"
#include "kernels.h"

struct Linear {
  void* w = nullptr;   // [N,K] K-contiguous, fp32 or bf16
  float* b = nullptr;  // [N]
  int N = 0, K = 0;
};

template <typename T> inline void dev_free(T*& p) {
  if (p) cudaFree(p);
  p = nullptr;
}
template <typename T> inline int dev_alloc(T** p, size_t bytes) {
  *p = nullptr;
  CUDA_TRY(cudaMalloc(reinterpret_cast<void**>(p), bytes ? bytes : 1));
  return TUTOR_OK;
}

int tutor_require_device(int device);       // sets the device; fails unless it is an sm_100 GPU
void tutor_count_launches(uint64_t n);      // bumps the counter behind tutor_launch_count()
int upload_f32(float** dst, const float* src, size_t n);
// rows x cols fp32 host matrix -> device matrix of `dt`; transpose=true stores it as [cols, rows]
int upload_matrix(void** dst, const float* src, int64_t rows, int64_t cols, bool transpose, int dt);
// LayerNorm folded into the linear that follows it (bf16 decode): W' = gamma (.) W as bf16 [N, K], s[n] = sum_k W'[n,k],
// c[n] = sum_k beta[k] W[n,k]; src is the Conv1D weight [K, N]
int upload_ln_folded(void** dst_w, float** dst_s, float** dst_c, const float* src, int64_t K, int64_t N, const float* gamma,
                     const float* beta);
