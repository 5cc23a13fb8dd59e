// Ground truth for the launch-bound decode chain: per-kernel cost of dependent launches inside a CUDA graph,
// with and without programmatic dependent launch, for an empty kernel and for the library's decode kernels.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I distributed-lms-raft-llm_b200/csrc \
//        tools/microbench.cu -o gpurun_out/microbench -L distributed-lms-raft-llm_b200 -ltutor_b200 -lcuda
// (-DTUTOR_TRACE -DTUTOR_EXPERIMENTS adds the decode-attention template variants)
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <vector>

#include "kernels.h"
#ifdef TUTOR_TRACE
#include "gemm_bf16_tc.cu"  // traced copy of the GEMM overrides the library's
#include "attention.cu"     // direct access to the decode-attention template variants
#endif

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA %s at %d\n", cudaGetErrorString(e_), __LINE__); exit(1); } } while (0)

__global__ void k_empty(int* p) {
  pdl_launch_dependents();
  pdl_wait();
  if (p && threadIdx.x == 0 && blockIdx.x == 0) p[0] += 1;
}

// n dependent global loads per thread (pointer chase through a small L2-resident table), then one store
__global__ void k_chase(const int* __restrict__ tab, int* out, int n) {
  pdl_launch_dependents();
  pdl_wait();
  int i = threadIdx.x + blockIdx.x * blockDim.x;
  for (int k = 0; k < n; ++k) i = tab[i & 0xffff];
  out[threadIdx.x + blockIdx.x * blockDim.x] = i;
}
__global__ void k_syncs(int* out, int n) {
  __shared__ int sm[8];
  int v = threadIdx.x;
  for (int k = 0; k < n; ++k) { if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v; __syncthreads(); v += sm[k & 7]; __syncthreads(); }
  out[threadIdx.x + blockIdx.x * blockDim.x] = v;
}

// pure streaming reads shaped like the decode attention: nstreams contiguous regions per CTA, 16 B per thread per load
__global__ void k_stream(const uint4* __restrict__ base, size_t region_vec, int nstreams, int nvec, uint4* sink) {
  const uint4* p = base + (size_t)blockIdx.x * nstreams * region_vec;
  uint4 acc = make_uint4(0, 0, 0, 0);
  for (int i = threadIdx.x; i < nvec; i += blockDim.x * 4) {
#pragma unroll
    for (int s = 0; s < 2; ++s) {
      if (s >= nstreams) break;
      const uint4* q = p + (size_t)s * region_vec;
      uint4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) { int j = i + u * blockDim.x; v[u] = j < nvec ? q[j] : make_uint4(0, 0, 0, 0); }
#pragma unroll
      for (int u = 0; u < 4; ++u) { acc.x ^= v[u].x; acc.y ^= v[u].y; acc.z ^= v[u].z; acc.w ^= v[u].w; }
    }
  }
  if (acc.x == 0x12345678u) sink[blockIdx.x] = acc;  // never true for our data; keeps the loads alive
}
__global__ void k_flat(const uint4* __restrict__ base, size_t nvec, uint4* sink) {
  uint4 acc = make_uint4(0, 0, 0, 0);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (size_t)gridDim.x * blockDim.x * 4) {
    uint4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) { size_t j = i + (size_t)u * gridDim.x * blockDim.x; v[u] = j < nvec ? base[j] : make_uint4(0, 0, 0, 0); }
#pragma unroll
    for (int u = 0; u < 4; ++u) { acc.x ^= v[u].x; acc.y ^= v[u].y; acc.z ^= v[u].z; acc.w ^= v[u].w; }
  }
  if (acc.x == 0x12345678u) sink[blockIdx.x] = acc;
}

static float time_graph(cudaStream_t st, int chain, bool pdl, const std::function<void(int)>& launch_one) {
  cudaGraph_t g; cudaGraphExec_t ge;
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

int main() {
  cudaStream_t st; CK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
  int* cnt; CK(cudaMalloc(&cnt, 4)); CK(cudaMemset(cnt, 0, 4));
  const int B = 64, d = 768;
  bf16 *x, *w, *h; float *partial, *xf, *bias, *gam; int* len;
  CK(cudaMalloc(&x, (size_t)B * 4 * d * 2)); CK(cudaMalloc(&h, (size_t)B * 4 * d * 2));
  CK(cudaMalloc(&w, (size_t)12 * 4 * d * d * 2));  // 12 "layers" of the largest weight
  CK(cudaMalloc(&partial, (size_t)32 * B * 4 * d * 4)); CK(cudaMalloc(&xf, (size_t)B * d * 4));
  CK(cudaMalloc(&bias, 4 * d * 4)); CK(cudaMalloc(&gam, 4 * d * 4)); CK(cudaMalloc(&len, 4));
  CK(cudaMemset(x, 0, (size_t)B * 4 * d * 2)); CK(cudaMemset(w, 0, (size_t)12 * 4 * d * d * 2));
  CK(cudaMemset(partial, 0, (size_t)32 * B * 4 * d * 4)); CK(cudaMemset(xf, 0, (size_t)B * d * 4));
  CK(cudaMemset(bias, 0, 4 * d * 4)); CK(cudaMemset(gam, 0, 4 * d * 4));
  {  // streaming-read patterns, 12 rotating "layers" of 2 x 98 KB x 768 regions (1.8 GB, larger than L2)
    const size_t region = 768 * 64 * 2 / 16;  // one (b,h) stream at Smax=768 in 16-byte vectors
    const int used = 640 * 64 * 2 / 16;        // S = 640 rows actually read
    uint4* big; CK(cudaMalloc(&big, (size_t)12 * 768 * 2 * region * 16)); CK(cudaMemset(big, 1, (size_t)12 * 768 * 2 * region * 16));
    uint4* sink; CK(cudaMalloc(&sink, 4096 * 16));
    const double mb = 768.0 * 2 * used * 16 / 1e6;
    struct { const char* n; int threads, nstreams; } pats[] = {{"768 CTAs x 128 thr, K and V streams", 128, 2}, {"768 CTAs x 256 thr, K and V streams", 256, 2}};
    for (auto& pt : pats) {
      float t = time_graph(st, 48, false, [&](int i) { k_stream<<<768, pt.threads, 0, st>>>(big + (size_t)(i % 12) * 768 * 2 * region, region, pt.nstreams, used, sink); });
      printf("stream %-40s %6.2f us  %7.1f GB/s\n", pt.n, t, mb / t * 1e-3 * 1e3);
    }
    float t = time_graph(st, 48, false, [&](int i) { k_stream<<<1536, 128, 0, st>>>(big + (size_t)(i % 12) * 768 * 2 * region, region, 1, used, sink); });
    printf("stream %-40s %6.2f us  %7.1f GB/s\n", "1536 CTAs x 128 thr, one stream each", t, mb / t * 1e-3 * 1e3);
    for (int ctas : {148 * 4, 148 * 8, 148 * 16}) {
      t = time_graph(st, 48, false, [&](int i) { k_flat<<<ctas, 256, 0, st>>>(big + (size_t)(i % 12) * 768 * 2 * region, (size_t)768 * 2 * used, sink); });
      printf("flat grid-stride read, %4d CTAs x 256 thr            %6.2f us  %7.1f GB/s\n", ctas, t, mb / t * 1e-3 * 1e3);
    }
#ifdef TUTOR_TRACE
    {
      bf16 *qq, *oo; CK(cudaMalloc(&qq, 64 * 768 * 2)); CK(cudaMalloc(&oo, 64 * 768 * 2)); CK(cudaMemset(qq, 0, 64 * 768 * 2));
      const size_t layer = (size_t)768 * 2 * region * 8;  // bf16 elements per "layer" (K block then V block)
      auto kp = [&](int i) { return (bf16*)big + (size_t)(i % 12) * layer; };
      auto vp = [&](int i) { return (bf16*)big + (size_t)(i % 12) * layer + layer / 2; };
#define ATT(NAME, ...)                                                                                              \
      t = time_graph(st, 48, false, [&](int i) { __VA_ARGS__; });                                                    \
      printf("attn %-42s %6.2f us  %7.2f TB/s\n", NAME, t, mb / t * 1e-3);
      ATT("register kernel (128 thr)", (attn_decode_kernel<bf16, bf16, bf16, true, false><<<dim3(12, 64), 128, 0, st>>>(qq, kp(i), vp(i), oo, 12, 768, nullptr, 640, nullptr, 0, nullptr, 64, QkvFold{}, nullptr, nullptr, nullptr)));
      ATT("ring chunk 64 x 2 stages", (attn_decode_tma_kernel<false, 64, 2, true><<<dim3(12, 64), 160, 32768, st>>>(qq, kp(i), vp(i), oo, 12, 768, nullptr, 640, nullptr, 0, nullptr, 64, QkvFold{}, nullptr, 0)));
      ATT("ring chunk 64 x 2 stages, no math", (attn_decode_tma_kernel<false, 64, 2, false><<<dim3(12, 64), 160, 32768, st>>>(qq, kp(i), vp(i), oo, 12, 768, nullptr, 640, nullptr, 0, nullptr, 64, QkvFold{}, nullptr, 0)));
      ATT("ring chunk 32 x 4 stages", (attn_decode_tma_kernel<false, 32, 4, true><<<dim3(12, 64), 160, 32768, st>>>(qq, kp(i), vp(i), oo, 12, 768, nullptr, 640, nullptr, 0, nullptr, 64, QkvFold{}, nullptr, 0)));
      ATT("ring chunk 32 x 4 stages, no math", (attn_decode_tma_kernel<false, 32, 4, false><<<dim3(12, 64), 160, 32768, st>>>(qq, kp(i), vp(i), oo, 12, 768, nullptr, 640, nullptr, 0, nullptr, 64, QkvFold{}, nullptr, 0)));
      {  // tensor-core consumers: the cache of every "layer" as a [768*768, 64] bf16 matrix behind a tensor map
        static CUtensorMap tk[12], tv[12];
        for (int i = 0; i < 12; ++i) {
          if (k_make_map_bf16(&tk[i], kp(i), 768 * 768, 64, DM_CHUNK) || k_make_map_bf16(&tv[i], vp(i), 768 * 768, 64, DM_CHUNK)) exit(3);
        }
        ATT("mma ring chunk 32 x 4 stages", (attn_decode_mma_kernel<false, true><<<dim3(12, 64), DM_THREADS, 0, st>>>(tk[i % 12], tv[i % 12], qq, kp(i), vp(i), oo, 12, 768, nullptr, 640, nullptr, 0, nullptr, 64, nullptr, 0, 0, nullptr, nullptr)));
        ATT("mma ring chunk 32 x 4 stages, no math", (attn_decode_mma_kernel<false, false><<<dim3(12, 64), DM_THREADS, 0, st>>>(tk[i % 12], tv[i % 12], qq, kp(i), vp(i), oo, 12, 768, nullptr, 640, nullptr, 0, nullptr, 64, nullptr, 0, 0, nullptr, nullptr)));
      }
      ATT("ring chunk 16 x 8 stages", (attn_decode_tma_kernel<false, 16, 8, true><<<dim3(12, 64), 160, 32768, st>>>(qq, kp(i), vp(i), oo, 12, 768, nullptr, 640, nullptr, 0, nullptr, 64, QkvFold{}, nullptr, 0)));
    }
#endif
    cudaFree(big); cudaFree(sink);
  }
  if (getenv("MB_ATTN_ONLY")) return 0;
  for (int pdl = 0; pdl < 2; ++pdl) {
    printf("--- PDL %s\n", pdl ? "on" : "off");
    float t = time_graph(st, 96, pdl, [&](int) { CK(tutor_launch(k_empty, dim3(64), dim3(128), 0, st, cnt)); });
    printf("empty kernel (64 CTAs)              %6.2f us/kernel\n", t);
    t = time_graph(st, 96, pdl, [&](int) { CK(tutor_launch(k_empty, dim3(1), dim3(32), 0, st, (int*)nullptr)); });
    printf("empty kernel (1 CTA)                %6.2f us/kernel\n", t);
    int *tab, *outb; CK(cudaMalloc(&tab, 65536 * 4)); CK(cudaMalloc(&outb, 65536 * 4));
    { std::vector<int> hh(65536); for (int i = 0; i < 65536; ++i) hh[i] = (i * 7919 + 13) & 0xffff; CK(cudaMemcpy(tab, hh.data(), 65536 * 4, cudaMemcpyHostToDevice)); }
    for (int n : {0, 1, 2, 4, 8}) {
      t = time_graph(st, 96, pdl, [&](int) { CK(tutor_launch(k_chase, dim3(64), dim3(256), 0, st, (const int*)tab, outb, n)); });
      printf("chase %d dependent L2 loads (64x256)  %6.2f us/kernel\n", n, t);
    }
    for (int n : {0, 2, 8}) {
      t = time_graph(st, 96, pdl, [&](int) { CK(tutor_launch(k_syncs, dim3(64), dim3(256), 0, st, outb, n)); });
      printf("%d sync pairs (64x256)               %6.2f us/kernel\n", n, t);
    }
    for (int ks : {1, 4, 12}) {
      t = time_graph(st, 96, pdl, [&](int) {
        if (k_reduce_bias_residual_ln(partial, ks, B, d, bias, xf, gam, bias, 1e-5f, nullptr, x, st)) exit(2); });
      printf("reduce+res+LN ks=%2d                 %6.2f us/kernel\n", ks, t);
    }
    t = time_graph(st, 96, pdl, [&](int) { if (k_layernorm(xf, 1, 0, gam, bias, 1e-5f, nullptr, x, B, d, st)) exit(2); });
    printf("layernorm 64 rows                   %6.2f us/kernel\n", t);
    struct { const char* n; int N, K; } shapes[] = {{"qkv  N=2304 K=768 ", 3 * d, d}, {"proj N=768  K=768 ", d, d},
                                                    {"fc   N=3072 K=768 ", 4 * d, d}, {"prj2 N=768  K=3072", d, 4 * d}};
    for (auto& s : shapes) {
      int ks = k_pick_splitk(s.N, s.K, 148);
      t = time_graph(st, 96, pdl, [&](int i) {
        if (k_gemm_bf16_splitk(x, w + (size_t)(i % 12) * 4 * d * d, B, s.N, s.K, ks, partial, st)) exit(2);
      });
      printf("splitk gemm %s ks=%2d     %6.2f us/kernel\n", s.n, ks, t);
#ifdef TUTOR_TRACE
      long long tr[8];
      CK(cudaMemcpyFromSymbol(tr, g_gemm_trace, sizeof(tr)));
      printf("   CTA0 cycles: setup %lld | tma issued +%lld | first full +%lld | mma committed +%lld | tmem_full seen +%lld | stores issued +%lld | end +%lld\n",
             tr[1] - tr[0], tr[2] - tr[0], tr[3] - tr[0], tr[4] - tr[0], tr[5] - tr[0], tr[6] - tr[0], tr[7] - tr[0]);
#endif
    }
    t = time_graph(st, 96, pdl, [&](int) {
      if (k_reduce_bias_residual_ln(partial, 12, B, d, bias, xf, gam, bias, 1e-5f, nullptr, x, st)) exit(2); });
    printf("reduce+res+LN ks=12                 %6.2f us/kernel\n", t);
    t = time_graph(st, 96, pdl, [&](int) {
      if (k_reduce_bias_act(partial, 6, B, 4 * d, bias, ACT_GELU_TANH, h, st)) exit(2); });
    printf("reduce+gelu ks=6                    %6.2f us/kernel\n", t);
    t = time_graph(st, 96, pdl, [&](int i) {
      if (i & 1) { if (k_reduce_bias_act(partial, 6, B, 4 * d, bias, ACT_GELU_TANH, h, st)) exit(2); }
      else if (k_gemm_bf16_splitk(x, w + (size_t)(i % 12) * 4 * d * d, B, 4 * d, d, 6, partial, st)) exit(2);
    });
    printf("alternating fc gemm / reduce+gelu   %6.2f us/kernel\n", t);
  }
  return 0;
}
