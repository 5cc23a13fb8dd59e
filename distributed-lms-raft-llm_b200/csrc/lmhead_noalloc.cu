// tutor_lmhead_argmax_static: the fused LM-head + repetition-penalty + arg-max of lmhead.cu for callers
// that are not allowed to allocate device memory.
//
// Who that is: code that runs where cudaMalloc is off limits - a host that has handed the whole GPU heap to
// its own pool allocator and forbids outside allocations, a real-time serving loop that must not take the
// driver's allocation lock (cudaMalloc can synchronise the device), or a plug-in loaded into a process whose
// memory budget is fixed at start-up.  tutor_lmhead_argmax asks such a caller for a scratch buffer it would
// have to allocate; this variant brings its own.
//
// The scratch (the per-CTA partial maxima that phase 1 writes and phase 2 reduces) lives in the file-scope
// __device__ array below: 1 MiB, enough for a batch of up to TUTOR_STATIC_MAX_BATCH = 128 rows.  Every GPU
// that loads this library carries that 1 MiB from load until the process exits, whether or not the entry
// point is ever called.
//
// Things a user must know:
//  * The array's name is only a host-side handle here; the device address is looked up once per device with
//    cudaGetSymbolAddress, on the entry point's FIRST call on that device (with that device current).  Make
//    that first call outside any CUDA stream capture.
//  * All calls on a device share the one array.  Two calls whose kernels can overlap on the GPU (different
//    streams, or different host threads on different streams) race on it and can return wrong ids; keep
//    calls on one stream or order them yourself.
//  * The array keeps whatever the previous call left in it.  That is harmless: phase 1 overwrites every
//    slot phase 2 reads (each CTA stores its partial for every batch row, even when it scanned no
//    vocabulary row), so nothing in it has to start at zero and nothing is cleared.
//  * No cudaMalloc / cudaFree / allocating library routine is called, on the first call or later.
#include <mutex>

#include "../../include/tutor_b200.h"
#include "kernels.h"

#define TUTOR_STATIC_SCRATCH_BYTES (1u << 20)

static __device__ __align__(256) unsigned char g_static_scratch[TUTOR_STATIC_SCRATCH_BYTES];

namespace {
constexpr int MAX_DEVICES = 64;
std::mutex g_setup_mu;
void* g_scratch_addr[MAX_DEVICES] = {nullptr};

// device address of g_static_scratch on the current device (looked up once per device)
int static_scratch(void** out) {
  int dev = -1;
  CUDA_TRY(cudaGetDevice(&dev));
  ARG_CHECK(dev >= 0 && dev < MAX_DEVICES, "lmhead_argmax_static: device index %d is not supported", dev);
  std::lock_guard<std::mutex> lock(g_setup_mu);
  if (!g_scratch_addr[dev]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess) return tutor_cuda_fail(e, "cudaGetSymbolAddress(g_static_scratch)", __FILE__, __LINE__);
    g_scratch_addr[dev] = p;
  }
  *out = g_scratch_addr[dev];
  return TUTOR_OK;
}
}  // namespace

extern "C" int tutor_lmhead_argmax_static(int dtype, const void* x, const void* wte, int B, int V, int d,
                                          const uint32_t* seen, float penalty, int32_t* out_idx, void* stream) {
  ARG_CHECK(x && wte && out_idx, "lmhead_argmax_static: null argument");
  ARG_CHECK(penalty == 1.0f || seen, "lmhead_argmax_static: a penalty needs the seen bitmap");
  if (B < 1 || B > TUTOR_STATIC_MAX_BATCH || k_lmhead_scratch_bytes(B, V) > TUTOR_STATIC_SCRATCH_BYTES) {
    tutor_set_error("lmhead_argmax_static: batch %d is outside 1..%d, the most the built-in %u-byte scratch holds; "
                    "use tutor_lmhead_argmax with a caller-allocated scratch for larger batches",
                    B, TUTOR_STATIC_MAX_BATCH, TUTOR_STATIC_SCRATCH_BYTES);
    return TUTOR_ERR_LIMIT;
  }
  void* scratch = nullptr;
  TRY(static_scratch(&scratch));
  DecodeState ds{};
  ds.seen = const_cast<uint32_t*>(seen);
  ds.seen_words = ceil_div(V, 32);
  ds.penalty = penalty;
  ds.plain_out = out_idx;
  return k_lmhead_argmax(x, dtype, wte, dtype, B, V, d, ds, scratch, (cudaStream_t)stream);
}
