// Test-only cross-check for the tcgen05 GEMM: the same bf16 contraction on the legacy warp-level
// tensor-core path (mma.sync m16n8k16, HMMA in SASS), fragments loaded straight from global memory.
// Deliberately simple and slow; never on the product path.
#include "kernels.h"

namespace {

__global__ void __launch_bounds__(128) gemm_bf16_mma_sync_kernel(const bf16* __restrict__ A,
                                                                  const bf16* __restrict__ W, int M, int N, int K,
                                                                  GemmEpilogue epi) {
  // CTA tile 64x64, 4 warps as 2x2, each warp 32x32 = 2 (m16) x 4 (n8) mma tiles
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.y * 64 + (warp >> 1) * 32, n0 = blockIdx.x * 64 + (warp & 1) * 32;
  const int gr = lane >> 2, gc = (lane & 3) * 2;
  float c[2][4][4];
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int r = 0; r < 4; ++r) c[i][j][r] = 0.f;
  for (int k0 = 0; k0 < K; k0 += 16) {
    uint32_t a[2][4], b[4][2];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const bf16* r0 = A + (size_t)min(m0 + i * 16 + gr, M - 1) * K + k0 + gc;
      const bf16* r1 = A + (size_t)min(m0 + i * 16 + gr + 8, M - 1) * K + k0 + gc;
      a[i][0] = *reinterpret_cast<const uint32_t*>(r0);
      a[i][1] = *reinterpret_cast<const uint32_t*>(r1);
      a[i][2] = *reinterpret_cast<const uint32_t*>(r0 + 8);
      a[i][3] = *reinterpret_cast<const uint32_t*>(r1 + 8);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const bf16* w = W + (size_t)min(n0 + j * 8 + gr, N - 1) * K + k0 + gc;
      b[j][0] = *reinterpret_cast<const uint32_t*>(w);
      b[j][1] = *reinterpret_cast<const uint32_t*>(w + 8);
    }
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j)
        asm volatile(
            "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
            "{%0,%1,%2,%3};"
            : "+f"(c[i][j][0]), "+f"(c[i][j][1]), "+f"(c[i][j][2]), "+f"(c[i][j][3])
            : "r"(a[i][0]), "r"(a[i][1]), "r"(a[i][2]), "r"(a[i][3]), "r"(b[j][0]), "r"(b[j][1]));
  }
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        int m = m0 + i * 16 + gr + (r >> 1) * 8, n = n0 + j * 8 + gc + (r & 1);
        if (m < M && n < N) epilogue_store(epi, m, n, c[i][j][r]);
      }
}

}  // namespace

int k_gemm_bf16_legacy(const bf16* A, const bf16* W, int M, int N, int K, const GemmEpilogue& epi, cudaStream_t st) {
  ARG_CHECK(K % 16 == 0, "gemm_bf16_legacy: K=%d must be a multiple of 16", K);
  if (M == 0 || N == 0) return TUTOR_OK;
  gemm_bf16_mma_sync_kernel<<<dim3(ceil_div(N, 64), ceil_div(M, 64)), 128, 0, st>>>(A, W, M, N, K, epi);
  KERNEL_LAUNCHED();
  return TUTOR_OK;
}
