// Shared device/host helpers for the tutoring-inference kernels (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdarg>
#include <cstdio>

#define TUTOR_OK 0
#define TUTOR_ERR_ARG (-1)
#define TUTOR_ERR_CUDA (-2)
#define TUTOR_ERR_STATE (-3)
#define TUTOR_ERR_LIMIT (-4)

// thread-local last-error text, read through tutor_last_error()
void tutor_set_error(const char* fmt, ...);
int tutor_cuda_fail(cudaError_t e, const char* what, const char* file, int line);

#define CUDA_TRY(expr)                                                        \
  do {                                                                        \
    cudaError_t _e = (expr);                                                  \
    if (_e != cudaSuccess) return tutor_cuda_fail(_e, #expr, __FILE__, __LINE__); \
  } while (0)
void tutor_count_launches(uint64_t n);
int tutor_set_max_smem(const void* func, int bytes);  // once per (kernel, device)
// after every <<<>>>: count the launch (tutor_launch_count) and surface launch-configuration errors
#define KERNEL_LAUNCHED()        \
  do {                           \
    tutor_count_launches(1);     \
    CUDA_TRY(cudaGetLastError()); \
  } while (0)
#define TRY(expr)               \
  do {                          \
    int _rc = (expr);           \
    if (_rc != TUTOR_OK) return _rc; \
  } while (0)
#define ARG_CHECK(cond, ...)      \
  do {                            \
    if (!(cond)) {                \
      tutor_set_error(__VA_ARGS__); \
      return TUTOR_ERR_ARG;       \
    }                             \
  } while (0)

typedef __nv_bfloat16 bf16;

// Programmatic dependent launch (PDL): while set (per host thread, by the decode engine), kernels launched through
// tutor_launch() may start before their stream predecessor has finished; each of them calls pdl_wait() before it
// touches anything another kernel produces or still reads, so only launch latency and prologues overlap.
extern thread_local bool tutor_pdl_enabled;
#ifdef __CUDACC__
template <typename... KArgs, typename... Args>
inline cudaError_t tutor_launch(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr;
  attr.id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr.val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = &attr;
  cfg.numAttrs = tutor_pdl_enabled ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
// same, with the grid cut into thread-block clusters of cluster_z CTAs along z (cluster_z == 1: plain launch)
template <typename... KArgs, typename... Args>
inline cudaError_t tutor_launch_cluster(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                        int cluster_z, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[2];
  int n = 0;
  if (cluster_z > 1) {
    attr[n].id = cudaLaunchAttributeClusterDimension;
    attr[n].val.clusterDim.x = 1; attr[n].val.clusterDim.y = 1; attr[n].val.clusterDim.z = (unsigned)cluster_z;
    ++n;
  }
  if (tutor_pdl_enabled) {
    attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  cfg.attrs = attr;
  cfg.numAttrs = n;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
// no-ops unless the kernel was launched with the PDL attribute
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
#endif

// A/B knobs that select paths measured slower than the default, or (TUTOR_SKIP) that drop kernels and produce meaningless
// tokens, exist only in builds made with -DTUTOR_EXPERIMENTS (TUTOR_EXPERIMENTS=1 python -m ...build, used by tools/):
// the shipped library has one path per shape.  Knobs that select another CORRECT path and are exercised by the tests
// (TUTOR_NO_FUSED, TUTOR_NO_CLUSTER, TUTOR_NO_PDL, TUTOR_NO_GRAPH, TUTOR_ATTN_SPLITS, TUTOR_MEGA) stay.
#ifdef TUTOR_EXPERIMENTS
#include <cstdlib>
static inline const char* tutor_experiment(const char* name) { return getenv(name); }
#else
static inline const char* tutor_experiment(const char*) { return nullptr; }
#endif

enum { ACT_NONE = 0, ACT_GELU_TANH = 1, ACT_GELU_ERF = 2 };
enum { DT_F32 = 0, DT_BF16 = 1 };

static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
static inline size_t round_up(size_t a, size_t b) { return (a + b - 1) / b * b; }

#ifdef __CUDACC__
__device__ __forceinline__ float to_f32(float v) { return v; }
__device__ __forceinline__ float to_f32(bf16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ bf16 from_f32<bf16>(float v) { return __float2bfloat16_rn(v); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// NewGELUActivation (HF/activations.py:59-66) and exact erf GELU (HF/activations.py:70-89);
// accurate tanhf/erff on purpose (parity mode forbids the .approx forms).
__device__ __forceinline__ float gelu_tanh(float x) {
  const float k = 0.7978845608028654f;  // sqrt(2/pi)
  return 0.5f * x * (1.0f + tanhf(k * (x + 0.044715f * x * x * x)));
}
__device__ __forceinline__ float gelu_erf(float x) { return 0.5f * x * (1.0f + erff(x * 0.7071067811865476f)); }
__device__ __forceinline__ float apply_act(float x, int act) {
  if (act == ACT_GELU_TANH) return gelu_tanh(x);
  if (act == ACT_GELU_ERF) return gelu_erf(x);
  return x;
}

// bf16 throughput mode only: one-instruction tanh (MUFU.TANH, |err| ~ 5e-4) for both GELU flavours; the results are
// rounded to bf16 (2^-9 relative) right after, so it does not show.
// The accurate tanhf/erff above cost 30-40 instructions per element - more than the MMA time of a 128x128 tile.
__device__ __forceinline__ float gelu_tanh_fast(float x) {
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(0.7978845608028654f * (x + 0.044715f * x * x * x)));
  return 0.5f * x * (1.0f + t);
}
// exact (erf) GELU in tanh form: 0.5 x (1 + tanh(x (a + b x^2 + c x^4))) with (a, b, c) fitted to x Phi(x) over
// [-7, 7] (max |err| 2.5e-5, 20x closer to the erf form than gelu_new's constants); one MUFU.TANH instead of
// MUFU.EX2 + MUFU.RCP + a 5-term polynomial - the intermediate GEMM's epilogue was MUFU/FMA-bound (90 us vs 73 us for
// the same shape without the activation).  bf16 mode only; the result is rounded to bf16 (2^-9 relative) right after.
__device__ __forceinline__ float gelu_erf_fast(float x) {
  const float x2 = fminf(x * x, 49.0f);  // beyond |x| = 7 the quartic would turn over; tanh is +-1 there anyway
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(x * fmaf(x2, fmaf(x2, -3.51516794e-4f, 3.70056460e-2f), 7.97507884e-1f)));
  return 0.5f * x * (1.0f + t);
}
__device__ __forceinline__ float apply_act_fast(float x, int act) {
  if (act == ACT_GELU_TANH) return gelu_tanh_fast(x);
  if (act == ACT_GELU_ERF) return gelu_erf_fast(x);
  return x;
}

// 8 packed bf16 (one 16-byte vector) -> 8 floats
__device__ __forceinline__ void unpack_bf16x8(const uint4& v, float* f) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    f[2 * i] = __uint_as_float(w[i] << 16);
    f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
  }
}

// Load 4 consecutive elements as floats (16 B for fp32, 8 B for bf16); p must be suitably aligned.
__device__ __forceinline__ float4 load4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ float4 load4(const bf16* p) {
  uint2 v = *reinterpret_cast<const uint2*>(p);
  float4 r;
  r.x = __uint_as_float(v.x << 16);
  r.y = __uint_as_float(v.x & 0xffff0000u);
  r.z = __uint_as_float(v.y << 16);
  r.w = __uint_as_float(v.y & 0xffff0000u);
  return r;
}

// -------------------------------------------------------------------------------------------
// The one epilogue every GEMM flavour shares: value (m,n) of  A·Wᵀ  ->  bias, activation,
// residual, then either a plain row-major store or the QKV scatter that writes K and V straight
// into the preallocated cache (replaces DynamicLayer.update's torch.cat, HF/cache_utils.py:119).
struct GemmEpilogue {
  const float* bias;      // [N] or null
  const float* residual;  // [M, ldo] fp32 or null (may alias out_f32)
  float* out_f32;         // [M, ldo] or null
  bf16* out_bf16;         // [M, ldo] or null
  int ldo;
  int act;
  // QKV scatter (qkv_mode != 0): columns [0,d) -> q (out_*, ldo=d), [d,2d) -> K cache, [2d,3d) -> V cache
  int qkv_mode;
  int kv_dtype;  // DT_F32 / DT_BF16
  void* kcache;  // this layer's K base: [B][H][Smax][64]
  void* vcache;
  int d, H, Smax, T;   // T = tokens per sequence in this GEMM (row m -> b = m / T, t = m % T)
  int pos0;            // first cache position of this GEMM's tokens (prefill: 0)
  const int* pos_dev;  // if non-null: position = *pos_dev - 1 (decode step; T must be 1)
  const bf16* residual_bf16;  // [M, ldo] bf16 or null: the residual of a bf16 activation stream (out_bf16 only, not with residual)
};

// everything after bias + activation: residual add and the store (plain or QKV scatter)
__device__ __forceinline__ void epilogue_finish(const GemmEpilogue& e, int m, int n, float v);
__device__ __forceinline__ void epilogue_store(const GemmEpilogue& e, int m, int n, float v) {
  if (e.bias) v += e.bias[n];
  epilogue_finish(e, m, n, apply_act(v, e.act));
}
__device__ __forceinline__ void epilogue_finish(const GemmEpilogue& e, int m, int n, float v) {
  if (!e.qkv_mode) {
    size_t o = (size_t)m * e.ldo + n;
    if (e.residual) v += e.residual[o];
    if (e.residual_bf16) v += __bfloat162float(e.residual_bf16[o]);
    if (e.out_f32) e.out_f32[o] = v;
    if (e.out_bf16) e.out_bf16[o] = __float2bfloat16_rn(v);
    return;
  }
  if (n < e.d) {
    size_t o = (size_t)m * e.ldo + n;
    if (e.out_f32) e.out_f32[o] = v;
    if (e.out_bf16) e.out_bf16[o] = __float2bfloat16_rn(v);
    return;
  }
  int which = (n >= 2 * e.d);
  int c = n - (which ? 2 * e.d : e.d);
  int h = c >> 6, i = c & 63;
  int b = m / e.T, t = m - b * e.T;
  int s = e.pos_dev ? (*e.pos_dev - 1) : (e.pos0 + t);
  size_t o = (((size_t)b * e.H + h) * e.Smax + s) * 64 + i;
  void* base = which ? e.vcache : e.kcache;
  if (e.kv_dtype == DT_F32) reinterpret_cast<float*>(base)[o] = v;
  else reinterpret_cast<bf16*>(base)[o] = __float2bfloat16_rn(v);
}
#endif  // __CUDACC__
