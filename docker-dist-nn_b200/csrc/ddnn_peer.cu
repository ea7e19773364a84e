// Peer-memory transport for the layer-per-GPU pipeline: no NCCL kernel and no copy in the data
// path.  A stage maps its neighbours' receive buffers (CUDA IPC) and passes the mapped pointer
// straight to the GEMM entry points, so the tcgen05 epilogue of the stage's last layer stores the
// activation tile by tile into the next GPU's HBM over NVLink (and the dX epilogue stores dZ into
// the previous GPU's); one 4-byte flag written after the kernel tells the neighbour's stream, which
// waits on it with a stream memory operation, that the micro-batch has landed.
#include <cuda.h>

#include <cstring>
#include <map>
#include <mutex>
#include <string>

#include "ddnn_internal.h"

namespace ddnn {
namespace {

__global__ void flag_write_kernel(volatile int32_t* flag, int32_t value) {
  __threadfence_system();  // everything this stream produced before is visible system-wide first
  *flag = value;
}

typedef CUresult (*WaitValue32Fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
typedef CUresult (*AddrRangeFn)(CUdeviceptr*, size_t*, CUdeviceptr);

void* driver_fn(const char* name) {
  void* f = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &f, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
    return nullptr;
  return f;
}

std::mutex g_ipc_mu;
std::map<std::string, void*> g_ipc_open;  // handle bytes -> mapped base pointer

}  // namespace
}  // namespace ddnn

using namespace ddnn;

extern "C" {

int ddnn_ipc_export(const void* ptr, void* handle64, int64_t* offset) {
  if (!ptr || !handle64 || !offset) {
    set_error("ddnn_ipc_export: null argument");
    return DDNN_ERR_ARG;
  }
  static AddrRangeFn range = (AddrRangeFn)driver_fn("cuMemGetAddressRange");
  if (!range) {
    set_error("ddnn_ipc_export: cuMemGetAddressRange is not available");
    return DDNN_ERR_CUDA;
  }
  CUdeviceptr base = 0;
  size_t size = 0;
  CUresult r = range(&base, &size, (CUdeviceptr)ptr);
  if (r != CUDA_SUCCESS) {
    set_error("ddnn_ipc_export: cuMemGetAddressRange failed (%d)", (int)r);
    return DDNN_ERR_CUDA;
  }
  cudaIpcMemHandle_t h;
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
  DDNN_CUDA_CHECK(cudaIpcGetMemHandle(&h, (void*)base));
  memcpy(handle64, &h, 64);
  *offset = (int64_t)((CUdeviceptr)ptr - base);
  return DDNN_OK;
}

int ddnn_ipc_import(const void* handle64, int64_t offset, void** ptr) {
  if (!handle64 || !ptr) {
    set_error("ddnn_ipc_import: null argument");
    return DDNN_ERR_ARG;
  }
  std::string key((const char*)handle64, 64);
  std::lock_guard<std::mutex> g(g_ipc_mu);
  auto it = g_ipc_open.find(key);
  void* base = nullptr;
  if (it != g_ipc_open.end()) {
    base = it->second;
  } else {
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    DDNN_CUDA_CHECK(cudaIpcOpenMemHandle(&base, h, cudaIpcMemLazyEnablePeerAccess));
    g_ipc_open[key] = base;
  }
  *ptr = (char*)base + offset;
  return DDNN_OK;
}

int ddnn_flag_write(int32_t* flag, int32_t value, void* stream) {
  if (!flag) {
    set_error("ddnn_flag_write: null flag");
    return DDNN_ERR_ARG;
  }
  flag_write_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(flag, value);
  count_launch();
  DDNN_LAUNCH_CHECK();
  return DDNN_OK;
}

int ddnn_flag_wait(const int32_t* flag, int32_t value, void* stream) {
  if (!flag) {
    set_error("ddnn_flag_wait: null flag");
    return DDNN_ERR_ARG;
  }
  static WaitValue32Fn wait32 = (WaitValue32Fn)driver_fn("cuStreamWaitValue32");
  if (!wait32) {
    set_error("ddnn_flag_wait: this driver has no cuStreamWaitValue32 (stream memory operations)");
    return DDNN_ERR_UNSUPPORTED;
  }
  // A polling kernel would be the alternative, but on one GPU it can deadlock behind the very stream that has
  // to produce the flag, so a driver that refuses the memory operation is reported instead of worked around.
  CUresult r = wait32((CUstream)stream, (CUdeviceptr)flag, (cuuint32_t)value, CU_STREAM_WAIT_VALUE_GEQ);
  if (r != CUDA_SUCCESS) {
    set_error("ddnn_flag_wait: cuStreamWaitValue32 failed (%d)", (int)r);
    return DDNN_ERR_CUDA;
  }
  return DDNN_OK;
}

}  // extern "C"
