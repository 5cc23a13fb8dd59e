// Do 8 independent L1-bypassing loads from one thread overlap?  nvcc -arch=sm_100a -O3
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE> __device__ __forceinline__ float4 ld(const float4* p) {
  float4 v;
  if (MODE == 0) v = *p;
  else if (MODE == 1) v = __ldcg(p);
  else if (MODE == 2) asm volatile("ld.relaxed.gpu.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  else if (MODE == 3) asm volatile("ld.global.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  else asm volatile("ld.global.cv.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
template <int MODE, int N> __global__ void k(const float4* data, long long* out, int stride) {
  float acc = 0.f;
  long long best = 1 << 30;
  for (int rep = 0; rep < 20; ++rep) {
    const float4* p = data + (size_t)rep * 8192 + threadIdx.x;
    long long t0 = clock64();
    float4 v[N];
#pragma unroll
    for (int i = 0; i < N; ++i) v[i] = ld<MODE>(p + (size_t)i * stride);
#pragma unroll
    for (int i = 0; i < N; ++i) acc += v[i].x + v[i].w;
    long long t1 = clock64();
    if (acc == 123.f) out[1] = 1;
    if (t1 - t0 < best) best = t1 - t0;
  }
  if (threadIdx.x == 0) out[0] = best;
}
template <int MODE, int N> void run(const char* name, const float4* d, long long* out) {
  for (int threads : {1, 128}) {
    long long h;
    k<MODE, N><<<1, threads>>>(d, out, 48);
    cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost);
    printf("%-28s N=%d threads=%3d: %lld cycles\n", name, N, threads, h);
  }
}
int main() {
  float4* d; long long* out;
  cudaMalloc(&d, 64 << 20); cudaMemset(d, 0, 64 << 20); cudaMalloc(&out, 64);
  run<0, 1>("plain", d, out); run<0, 8>("plain", d, out);
  run<1, 1>("__ldcg", d, out); run<1, 8>("__ldcg", d, out);
  run<2, 8>("ld.relaxed.gpu", d, out);
  run<3, 8>("ld.L1::no_allocate", d, out);
  run<4, 8>("ld.cv", d, out);
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
}
