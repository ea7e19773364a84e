// ddnn_dense_backward_dw_sgd_static -- the weight-gradient / SGD step for callers that cannot
// hand the library a workspace.
//
// Who that is.  The regular entry point, ddnn_dense_backward_dw_sgd, reduces its split-K
// partial sums through a caller-owned device buffer, so its caller must be able to allocate
// device memory.  Some hosts of this library cannot:
//   * a node process whose framework grabbed (almost) all of HBM for its own caching pool at
//     start-up -- a raw cudaMalloc from a plug-in then fails with out-of-memory, and the pool
//     is not reachable through a C ABI;
//   * a thin FFI stub (ctypes, cgo, JNI) that can forward pointers it was given but has no
//     device allocator of its own to produce a workspace pointer from;
//   * a latency-sensitive serving loop that bans cudaMalloc/cudaFree on the request path because
//     both synchronise the whole device and stall every other stream.
// For them the scratch lives inside the library: one file-scope __device__ array of
// DDNN_STATIC_SCRATCH_BYTES.  Every GPU that loads libddnn.so carries that array from module
// load until the process exits, whether or not this entry point is ever called, which is why
// the size is kept modest.  Layers whose plan needs more scratch are refused (see below).
//
// What to know before calling.
//   * Set-up happens inside the first call made on each device: with that device current, the
//     device address of the array is looked up once (cudaGetSymbolAddress) and remembered.
//     Make that first call on every device you use before any stream capture is started.
//   * The array is shared by every call in the process and is never cleared: it still holds
//     the previous call's partial sums.  That is safe because a call only reads slots that the
//     same call has written first (each split writes its complete K x N / N partial slab before
//     the reducer runs), so nothing in it has to start at zero.
//   * Because the array is shared, two calls whose kernels can run at the same time on one GPU
//     (different streams, or different host threads on different streams) overwrite each other's
//     partial sums.  Issue these calls on one stream, or order them with events.
//   * No cudaMalloc, cudaFree or allocating library call is made here or below.
#include <mutex>

#include "ddnn_internal.h"

namespace ddnn {
namespace {

// The scratch itself.  In HOST code this name is only a handle for the runtime's symbol
// table -- it is not a device address and must never be passed to a kernel as one.
static __device__ __align__(256) unsigned char g_static_scratch[DDNN_STATIC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
std::mutex g_setup_mu;
void* g_scratch_addr[kMaxDevices] = {};  // per device ordinal: the array's DEVICE address

// Returns the device address of g_static_scratch on the current device (one-time look-up).
int static_scratch_ptr(void** out) {
  int dev = -1;
  DDNN_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) {
    set_error("ddnn_dense_backward_dw_sgd_static: device ordinal %d out of range", dev);
    return DDNN_ERR_ARG;
  }
  std::lock_guard<std::mutex> lock(g_setup_mu);
  if (!g_scratch_addr[dev]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess || !p) {
      set_error("ddnn_dense_backward_dw_sgd_static: cudaGetSymbolAddress failed on device %d: %s", dev,
                cudaGetErrorString(e));
      return DDNN_ERR_CUDA;
    }
    g_scratch_addr[dev] = p;
  }
  *out = g_scratch_addr[dev];
  return DDNN_OK;
}

}  // namespace
}  // namespace ddnn

using namespace ddnn;

extern "C" {

int ddnn_dense_backward_dw_sgd_static(const void* X, const void* dZ, float* W_master, void* W_shadow, float* bias,
                                      float* dW_out, float* db_out, float lr, int apply, int64_t B, int64_t K,
                                      int64_t N, int dtype, void* stream) {
  if (B < 0 || K <= 0 || N <= 0) {
    set_error("ddnn_dense_backward_dw_sgd_static: bad dimensions B=%lld K=%lld N=%lld", (long long)B, (long long)K,
              (long long)N);
    return DDNN_ERR_DIM;
  }
  if (B == 0) return DDNN_OK;
  const size_t need = ddnn_dw_workspace_bytes(B, K, N, dtype);
  if (need > (size_t)DDNN_STATIC_SCRATCH_BYTES) {
    set_error(
        "ddnn_dense_backward_dw_sgd_static: this layer (B=%lld K=%lld N=%lld) needs %zu bytes of split-K scratch "
        "but the built-in scratch holds %zu; use ddnn_dense_backward_dw_sgd with a caller-owned workspace",
        (long long)B, (long long)K, (long long)N, need, (size_t)DDNN_STATIC_SCRATCH_BYTES);
    return DDNN_ERR_WORKSPACE;
  }
  void* scratch = nullptr;
  int rc = static_scratch_ptr(&scratch);
  if (rc != DDNN_OK) return rc;
  // Same launches as the regular entry point, with the built-in array as the workspace.
  return ddnn_dense_backward_dw_sgd(X, dZ, W_master, W_shadow, bias, dW_out, db_out, lr, apply, B, K, N, dtype,
                                    scratch, (size_t)DDNN_STATIC_SCRATCH_BYTES, stream);
}

}  // extern "C"
