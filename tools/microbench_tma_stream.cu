// Per-SM streaming rate of 16 KB TMA tiles (128 x 64 bf16, SWIZZLE_128B) through an NST-deep ring, for N CTAs.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lcuda
#include <cstdio>
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
template <int NST> __global__ void __launch_bounds__(64) stream(const __grid_constant__ CUtensorMap map, int tiles_per_cta, int total_tiles, long long* out) {
  extern __shared__ uint8_t raw[];
  uint8_t* tiles = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  uint64_t* full = (uint64_t*)(tiles + NST * 16384);
  uint64_t* empty = full + NST;
  if (threadIdx.x == 0) {
    for (int s = 0; s < NST; ++s) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&full[s])));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&empty[s])));
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  long long t0 = clock64();
  if (threadIdx.x == 0) {
    for (int i = 0; i < tiles_per_cta; ++i) {
      int s = i % NST;
      if (i >= NST) { uint32_t ok = 0, par = ((i / NST) & 1) ^ 1; while (!ok) asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0,1,0,p;}" : "=r"(ok) : "r"(s32(&empty[s])), "r"(par) : "memory"); }
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(&full[s])), "r"(16384) : "memory");
      int tile = (blockIdx.x * tiles_per_cta + i) % total_tiles;
      asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(s32(tiles + s * 16384)), "l"(&map), "r"(s32(&full[s])), "r"(0), "r"(tile * 128) : "memory");
    }
  } else if (threadIdx.x == 32) {
    for (int i = 0; i < tiles_per_cta; ++i) {
      int s = i % NST; uint32_t ok = 0, par = (i / NST) & 1;
      while (!ok) asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0,1,0,p;}" : "=r"(ok) : "r"(s32(&full[s])), "r"(par) : "memory");
      asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(&empty[s])) : "memory");
    }
    out[blockIdx.x] = clock64() - t0;
  }
}
int main() {
  const int total_tiles = 8192;  // 128 MB of bf16 [8192*128, 64]
  __nv_bfloat16* d; cudaMalloc(&d, (size_t)total_tiles * 16384); cudaMemset(d, 0, (size_t)total_tiles * 16384);
  long long* out; cudaMalloc(&out, 148 * 8);
  CUtensorMap map; cuuint64_t dims[2] = {64, (cuuint64_t)total_tiles * 128}; cuuint64_t strides[1] = {128}; cuuint32_t box[2] = {64, 128}, es[2] = {1, 1};
  CUresult r = cuTensorMapEncodeTiled(&map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return 1; }
  cudaFuncSetAttribute(stream<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  cudaFuncSetAttribute(stream<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  for (int warm = 0; warm < 2; ++warm)
  for (int nst : {4, 8})
    for (int ctas : {1, 8, 16, 32, 64, 148}) {
      for (int tiles : {56, 512}) {  // 56 tiles = 0.9 MB per CTA (one layer's share of a 16-CTA cluster); 512 = 8 MB
        size_t smem = nst * 16384 + 1024 + 256;
        if (nst == 4) stream<4><<<ctas, 64, smem>>>(map, tiles, total_tiles, out); else stream<8><<<ctas, 64, smem>>>(map, tiles, total_tiles, out);
        long long h[148]; cudaMemcpy(h, out, ctas * 8, cudaMemcpyDeviceToHost);
        long long mx = 0; for (int i = 0; i < ctas; ++i) mx = h[i] > mx ? h[i] : mx;
        double us = mx / (clk / 1000.0);
        if (warm) printf("ring %d, %3d CTAs, %3d tiles/CTA (%s): %.1f us, %.0f GB/s per CTA, %.2f TB/s total\n", nst, ctas, tiles, tiles * ctas * 16384 > (100 << 20) ? "HBM" : "mostly L2 after warm-up", us, tiles * 16384 / us / 1e3, (double)tiles * ctas * 16384 / us / 1e6);
      }
    }
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
}
