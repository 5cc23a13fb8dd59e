// Latencies that bound a flag-chained persistent kernel on B200: L2 round trip of data written by another SM,
// release-reduction + poll propagation between two CTAs (ping-pong), fences.  nvcc -arch=sm_100a -O3 -o mb tools/microbench_sync.cu
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ int ld_relaxed(const int* p) { int v; asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ int ld_acquire(const int* p) { int v; asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ void red_release(int* p, int v) { asm volatile("red.release.gpu.global.add.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ void red_relaxed(int* p, int v) { asm volatile("red.relaxed.gpu.global.add.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }

// ping-pong between block 0 and block `peer`: round trip of (release-add, relaxed poll, acquire fence) x 2
__global__ void pingpong(int* flags, float* data, long long* out, int peer, int iters, int mode) {
  if (blockIdx.x != 0 && blockIdx.x != peer) return;
  const bool a = blockIdx.x == 0;
  long long t0 = clock64();
  for (int i = 1; i <= iters; ++i) {
    if (a) {
      if (mode >= 1) for (int k = threadIdx.x; k < 1024; k += blockDim.x) data[k] = (float)i;  // payload: 4 KB
      __syncthreads();
      if (threadIdx.x == 0) { red_release(flags, 1); while (ld_relaxed(flags + 32) < i) {} asm volatile("fence.acq_rel.gpu;" ::: "memory"); }
      __syncthreads();
    } else {
      if (threadIdx.x == 0) { while (ld_relaxed(flags) < i) {} asm volatile("fence.acq_rel.gpu;" ::: "memory"); }
      __syncthreads();
      float s = 0.f;
      if (mode >= 1) for (int k = threadIdx.x; k < 1024; k += blockDim.x) s += __ldcg(data + k);
      if (s == -1.f) out[3] = 1;
      __syncthreads();
      if (threadIdx.x == 0) red_release(flags + 32, 1);
    }
  }
  if (a && threadIdx.x == 0) out[0] = (clock64() - t0) / iters;
}
// dependent chain of L2 loads (pointer chase over lines written by the host): pure L2 hit latency
__global__ void chase(const int* next, long long* out, int iters) {
  int idx = 0;
  long long t0 = clock64();
  for (int i = 0; i < iters; ++i) idx = __ldcg(next + idx);
  out[0] = (clock64() - t0) / iters + (idx == -1);
}
__global__ void fences(int* flags, long long* out) {
  long long t0 = clock64();
  for (int i = 0; i < 100; ++i) { flags[threadIdx.x] = i; asm volatile("fence.acq_rel.gpu;" ::: "memory"); }
  long long t1 = clock64();
  for (int i = 0; i < 100; ++i) { flags[threadIdx.x] = i; __threadfence(); }
  long long t2 = clock64();
  for (int i = 0; i < 100; ++i) { red_release(flags + 64, 1); }
  long long t3 = clock64();
  for (int i = 0; i < 100; ++i) { red_relaxed(flags + 64, 1); }
  long long t4 = clock64();
  for (int i = 0; i < 100; ++i) { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
  long long t5 = clock64();
  if (threadIdx.x == 0) { out[0] = (t1 - t0) / 100; out[1] = (t2 - t1) / 100; out[2] = (t3 - t2) / 100; out[3] = (t4 - t3) / 100; out[4] = (t5 - t4) / 100; }
}
int main() {
  int* flags; float* data; long long* out; int* next;
  cudaMalloc(&flags, 4096); cudaMalloc(&data, 1 << 20); cudaMalloc(&out, 256); cudaMalloc(&next, 1 << 22);
  long long h[8];
  int hn[1 << 20];
  static int perm[1 << 14];
  for (int i = 0; i < (1 << 14); ++i) perm[i] = i;
  srand(1); for (int i = (1 << 14) - 1; i > 0; --i) { int j = rand() % (i + 1); int t = perm[i]; perm[i] = perm[j]; perm[j] = t; }
  for (int i = 0; i < (1 << 20); ++i) hn[i] = 0;
  for (int i = 0; i < (1 << 14); ++i) hn[perm[i] * 64] = perm[(i + 1) % (1 << 14)] * 64;  // one hop per 256-byte line
  cudaMemcpy(next, hn, sizeof(hn), cudaMemcpyHostToDevice);
  chase<<<1, 1>>>(next, out, 4000); chase<<<1, 1>>>(next, out, 4000);
  cudaMemcpy(h, out, 64, cudaMemcpyDeviceToHost); printf("L2-hit dependent load: %lld cycles\n", h[0]);
  for (int threads : {1, 128}) { fences<<<1, threads>>>(flags, out); cudaMemcpy(h, out, 64, cudaMemcpyDeviceToHost);
    printf("%3d threads: st+fence.acq_rel.gpu %lld | st+__threadfence %lld | red.release %lld | red.relaxed %lld | fence.proxy.async %lld cycles\n", threads, h[0], h[1], h[2], h[3], h[4]); }
  for (int mode = 0; mode < 2; ++mode)
    for (int peer : {1, 2, 37, 74, 147}) {
      cudaMemset(flags, 0, 4096);
      pingpong<<<148, 128>>>(flags, data, out, peer, 2000, mode);
      cudaMemcpy(h, out, 64, cudaMemcpyDeviceToHost);
      printf("ping-pong block 0 <-> %3d, %s: %lld cycles per round trip (two hand-offs)\n", peer, mode ? "4 KB payload" : "flag only", h[0]);
    }
  int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0); printf("SM clock (max) %d kHz, err %s\n", clk, cudaGetErrorString(cudaDeviceSynchronize()));
}
