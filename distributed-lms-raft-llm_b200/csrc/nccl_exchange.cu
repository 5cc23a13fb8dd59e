// The one exchange step of the sharded corpus search (BASELINE configs[4]): every rank's [Q,k]
// (score, global row) candidates are packed into 8-byte elements, gathered with ONE ncclAllGather on the caller's stream
// straight into the layout the merge kernel reads ([rank][Q][k], no permute / contiguous copies), and merged on every
// rank.  Everything is enqueued on one stream and can be captured in a CUDA graph together with the scan.
// NCCL is the library PyTorch already has in the process: it is bound at run time (dlopen), so libtutor_b200.so has no
// link-time dependency on it and single-GPU users never load it.
#include <dlfcn.h>
#include <math_constants.h>

#include <mutex>

#include "../../include/tutor_b200.h"
#include "kernels.h"

namespace {

struct NcclId { char internal[128]; };
typedef void* NcclComm;
typedef int (*fn_get_unique_id)(NcclId*);
typedef int (*fn_comm_init_rank)(NcclComm*, int, NcclId, int);
typedef int (*fn_comm_destroy)(NcclComm);
typedef int (*fn_all_gather)(const void*, void*, size_t, int, NcclComm, cudaStream_t);
typedef const char* (*fn_error_string)(int);
constexpr int NCCL_UINT64 = 5;  // ncclDataType_t (nccl.h)

struct NcclApi {
  void* handle = nullptr;
  fn_get_unique_id get_unique_id = nullptr;
  fn_comm_init_rank comm_init_rank = nullptr;
  fn_comm_destroy comm_destroy = nullptr;
  fn_all_gather all_gather = nullptr;
  fn_error_string error_string = nullptr;
} g_nccl;
std::mutex g_nccl_mu;

int nccl_fail(int rc, const char* what) {
  tutor_set_error("%s failed: %s", what, g_nccl.error_string ? g_nccl.error_string(rc) : "NCCL error");
  return TUTOR_ERR_CUDA;
}
#define NCCL_TRY(expr, what)                \
  do {                                      \
    int _r = (expr);                        \
    if (_r != 0) return nccl_fail(_r, what); \
  } while (0)

__global__ void pack_candidates_kernel(const float* __restrict__ scores, const int32_t* __restrict__ idx, int n,
                                       uint2* __restrict__ packed) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) packed[i] = make_uint2(__float_as_uint(scores[i]), (uint32_t)idx[i]);
}

// gathered: [world][Q][k] (score, row).  One warp per query keeps its world*k candidates in registers (<= 8 per lane)
// and extracts the k best: larger score first, lower row among equal scores (what torch.topk over the whole corpus
// returns for distinct rows).
__global__ void __launch_bounds__(128) merge_gathered_kernel(const uint2* __restrict__ gathered, int Q, int k, int world,
                                                            float* __restrict__ out_v, int32_t* __restrict__ out_i) {
  const int q = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (q >= Q) return;
  const int ncand = world * k;
  float cv[8];
  int ck[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = lane + 32 * j;
    cv[j] = -CUDART_INF_F;
    ck[j] = 0x7fffffff;
    if (c < ncand) {
      const int r = c / k, jj = c - r * k;
      const uint2 e = gathered[((size_t)r * Q + q) * k + jj];
      if ((int)e.y >= 0) { cv[j] = __uint_as_float(e.x); ck[j] = (int)e.y; }
    }
  }
  for (int round = 0; round < k; ++round) {
    float v = -CUDART_INF_F;
    int key = 0x7fffffff;
#pragma unroll
    for (int j = 0; j < 8; ++j)
      if (ck[j] != 0x7fffffff && (cv[j] > v || (cv[j] == v && ck[j] < key))) { v = cv[j]; key = ck[j]; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, v, o);
      const int ok = __shfl_xor_sync(0xffffffffu, key, o);
      if (ok != 0x7fffffff && (key == 0x7fffffff || ov > v || (ov == v && ok < key))) { v = ov; key = ok; }
    }
    if (lane == 0) {
      out_v[(size_t)q * k + round] = key != 0x7fffffff ? v : -CUDART_INF_F;
      out_i[(size_t)q * k + round] = key != 0x7fffffff ? key : -1;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j)
      if (ck[j] == key) ck[j] = 0x7fffffff;  // rows are unique across ranks
  }
}

}  // namespace

extern "C" {

int tutor_nccl_load(const char* path) {
  std::lock_guard<std::mutex> lock(g_nccl_mu);
  if (g_nccl.handle) return TUTOR_OK;
  void* h = nullptr;
  if (path && *path) h = dlopen(path, RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) {
    tutor_set_error("cannot load NCCL (%s): %s", path && *path ? path : "libnccl.so.2", dlerror());
    return TUTOR_ERR_STATE;
  }
  g_nccl.get_unique_id = (fn_get_unique_id)dlsym(h, "ncclGetUniqueId");
  g_nccl.comm_init_rank = (fn_comm_init_rank)dlsym(h, "ncclCommInitRank");
  g_nccl.comm_destroy = (fn_comm_destroy)dlsym(h, "ncclCommDestroy");
  g_nccl.all_gather = (fn_all_gather)dlsym(h, "ncclAllGather");
  g_nccl.error_string = (fn_error_string)dlsym(h, "ncclGetErrorString");
  if (!g_nccl.get_unique_id || !g_nccl.comm_init_rank || !g_nccl.comm_destroy || !g_nccl.all_gather) {
    tutor_set_error("the NCCL library at %s lacks a required symbol", path ? path : "libnccl.so.2");
    return TUTOR_ERR_STATE;
  }
  g_nccl.handle = h;
  return TUTOR_OK;
}

int tutor_nccl_unique_id(void* id128) {
  ARG_CHECK(id128, "nccl_unique_id: null argument");
  ARG_CHECK(g_nccl.handle, "nccl_unique_id: call tutor_nccl_load first");
  NCCL_TRY(g_nccl.get_unique_id(reinterpret_cast<NcclId*>(id128)), "ncclGetUniqueId");
  return TUTOR_OK;
}

int tutor_nccl_comm_create(const void* id128, int world, int rank, void** comm) {
  ARG_CHECK(id128 && comm && world >= 1 && rank >= 0 && rank < world, "nccl_comm_create: bad argument");
  ARG_CHECK(g_nccl.handle, "nccl_comm_create: call tutor_nccl_load first");
  NcclId id = *reinterpret_cast<const NcclId*>(id128);
  NcclComm c = nullptr;
  NCCL_TRY(g_nccl.comm_init_rank(&c, world, id, rank), "ncclCommInitRank");
  *comm = c;
  return TUTOR_OK;
}

void tutor_nccl_comm_destroy(void* comm) {
  if (comm && g_nccl.comm_destroy) g_nccl.comm_destroy(comm);
}

size_t tutor_topk_merge_allgather_scratch_bytes(int Q, int k, int world) { return (size_t)(world + 1) * Q * k * 8; }

int tutor_topk_merge_allgather(void* comm, const float* scores, const int32_t* idx, int Q, int k, int world, void* scratch,
                               float* out_scores, int32_t* out_idx, void* stream) {
  ARG_CHECK(comm && scores && idx && scratch && out_scores && out_idx, "topk_merge_allgather: null argument");
  ARG_CHECK(Q >= 1 && k >= 1 && k <= 32 && world >= 1 && world * k <= 256,
            "topk_merge_allgather: Q=%d k=%d world=%d unsupported (k <= 32, world*k <= 256)", Q, k, world);
  ARG_CHECK(g_nccl.handle, "topk_merge_allgather: call tutor_nccl_load first");
  cudaStream_t st = (cudaStream_t)stream;
  uint2* mine = reinterpret_cast<uint2*>(scratch);
  uint2* gathered = mine + (size_t)Q * k;
  const int n = Q * k;
  pack_candidates_kernel<<<ceil_div(n, 256), 256, 0, st>>>(scores, idx, n, mine);
  KERNEL_LAUNCHED();
  NCCL_TRY(g_nccl.all_gather(mine, gathered, (size_t)n, NCCL_UINT64, comm, st), "ncclAllGather");
  merge_gathered_kernel<<<ceil_div(Q, 4), 128, 0, st>>>(gathered, Q, k, world, out_scores, out_idx);
  KERNEL_LAUNCHED();
  return TUTOR_OK;
}

}  // extern "C"
