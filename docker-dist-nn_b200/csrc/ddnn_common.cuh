// Shared device/host helpers for libddnn.so (sm_100a only).
#pragma once

#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>

#include "../../include/ddnn.h"

namespace ddnn {

typedef __nv_bfloat16 bf16;

// ---- host-side error plumbing -------------------------------------------------------
void set_error(const char* fmt, ...);
extern std::atomic<long long> g_launches;
inline void count_launch(int n = 1) { g_launches.fetch_add(n, std::memory_order_relaxed); }

#define DDNN_CUDA_CHECK(expr)                                                              \
  do {                                                                                     \
    cudaError_t _e = (expr);                                                               \
    if (_e != cudaSuccess) {                                                               \
      ::ddnn::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return DDNN_ERR_CUDA;                                                                \
    }                                                                                      \
  } while (0)

#define DDNN_LAUNCH_CHECK()                                                                \
  do {                                                                                     \
    cudaError_t _e = cudaGetLastError();                                                   \
    if (_e != cudaSuccess) {                                                               \
      ::ddnn::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e), __FILE__, __LINE__); \
      return DDNN_ERR_CUDA;                                                                \
    }                                                                                      \
  } while (0)

inline int cdiv(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }

// ---- programmatic dependent launch ----------------------------------------------------
// A train step is a chain of short kernels (MNIST: ten of 5-25 us), so the launch latency and the prologue of
// kernel i+1 are worth hiding behind kernel i.  Kernels launched through launch_pdl() may start as soon as every
// CTA of their predecessor in the stream has executed pdl_launch_dependents(); they MUST execute pdl_wait()
// before their first global-memory access (it returns once the predecessor grid has completed and flushed).
// Both instructions are no-ops in a kernel launched without the attribute.  ddnn_set_pdl(0) turns the attribute
// off (plain stream order).
extern std::atomic<int> g_pdl;
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                              Args... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = g_pdl.load(std::memory_order_relaxed) ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
#endif

// Opts a kernel into `bytes` of dynamic shared memory once per (kernel, device): function attributes
// belong to the device's context, and one process may drive several GPUs.
int ensure_dynamic_smem(const void* kernel, size_t bytes);

// ---- element access -----------------------------------------------------------------
__device__ __forceinline__ float to_f32(float v) { return v; }
__device__ __forceinline__ float to_f32(bf16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ bf16 from_f32<bf16>(float v) { return __float2bfloat16_rn(v); }

// 4 consecutive elements starting at p (p must be 4-element aligned).
__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ float4 ld4(const bf16* p) {
  uint2 r = *reinterpret_cast<const uint2*>(p);
  __nv_bfloat162 a = *reinterpret_cast<__nv_bfloat162*>(&r.x);
  __nv_bfloat162 b = *reinterpret_cast<__nv_bfloat162*>(&r.y);
  float2 fa = __bfloat1622float2(a), fb = __bfloat1622float2(b);
  return make_float4(fa.x, fa.y, fb.x, fb.y);
}
__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ void st4(bf16* p, float4 v) {
  __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y);
  __nv_bfloat162 b = __floats2bfloat162_rn(v.z, v.w);
  uint2 r;
  r.x = *reinterpret_cast<uint32_t*>(&a);
  r.y = *reinterpret_cast<uint32_t*>(&b);
  *reinterpret_cast<uint2*>(p) = r;
}

// ---- activations (grpc_node.py:62-73) -------------------------------------------------
__device__ __forceinline__ float act_fwd(float z, int act) {
  if (act == DDNN_ACT_RELU) return fmaxf(z, 0.0f);
  if (act == DDNN_ACT_SIGMOID) return 1.0f / (1.0f + expf(-z));
  return z;
}
// derivative expressed on the stored OUTPUT a = f(z)
__device__ __forceinline__ float act_bwd(float a, int act) {
  if (act == DDNN_ACT_RELU) return a > 0.0f ? 1.0f : 0.0f;
  if (act == DDNN_ACT_SIGMOID) return a * (1.0f - a);
  return 1.0f;
}

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

// ---- epilogue functors shared by the SIMT GEMM, the split-K reducer and tcgen05 ------
// Each handles up to 4 consecutive columns [col, col+nv) of row `row`.
template <typename TO>
struct EpiFwd {  // A = act(acc + bias)
  const float* bias;
  TO* out;
  int64_t ld;
  int act;
  __device__ __forceinline__ void apply4(int64_t row, int col, float4 v, int nv, bool vec) const {
    float x[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (j < nv) x[j] = act_fwd(x[j] + __ldg(bias + col + j), act);
    TO* p = out + row * ld + col;
    if (vec && nv == 4) {
      st4(p, make_float4(x[0], x[1], x[2], x[3]));
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (j < nv) p[j] = from_f32<TO>(x[j]);
    }
  }
};

template <typename TA>
struct EpiDx {  // dZprev = acc * f'(Aprev)
  const TA* aprev;
  TA* out;
  int64_t ld;
  int act;
  __device__ __forceinline__ void apply4(int64_t row, int col, float4 v, int nv, bool vec) const {
    float x[4] = {v.x, v.y, v.z, v.w};
    const TA* a = aprev + row * ld + col;
    TA* p = out + row * ld + col;
    if (vec && nv == 4) {
      float4 av = (act == DDNN_ACT_LINEAR) ? make_float4(1.f, 1.f, 1.f, 1.f) : ld4(a);
      st4(p, make_float4(x[0] * act_bwd(av.x, act), x[1] * act_bwd(av.y, act),
                         x[2] * act_bwd(av.z, act), x[3] * act_bwd(av.w, act)));
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (j < nv) {
          float g = (act == DDNN_ACT_LINEAR) ? 1.0f : act_bwd(to_f32(a[j]), act);
          p[j] = from_f32<TA>(x[j] * g);
        }
    }
  }
};

struct EpiStoreF32 {  // raw fp32 store (gradient output or split-K partial)
  float* out;
  int64_t ld;
  __device__ __forceinline__ void prefetch(int64_t, int) const {}
  __device__ __forceinline__ void apply4(int64_t row, int col, float4 v, int nv, bool vec) const {
    float* p = out + row * ld + col;
    if (vec && nv == 4) {
      st4(p, v);
    } else {
      float x[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (j < nv) p[j] = x[j];
    }
  }
};

struct EpiSgd {  // W_master -= lr * acc ; shadow = bf16(W_master)
  float* w;
  bf16* shadow;  // nullable
  int64_t ld;
  float lr;
  __device__ __forceinline__ void prefetch(int64_t row, int col) const {  // pulls the line towards L1 ahead of apply4
    asm volatile("prefetch.global.L1 [%0];" ::"l"(w + row * ld + col));
  }
  __device__ __forceinline__ void apply4(int64_t row, int col, float4 v, int nv, bool vec) const {
    float* p = w + row * ld + col;
    if (vec && nv == 4) {
      float4 o = ld4(p);
      o.x -= lr * v.x; o.y -= lr * v.y; o.z -= lr * v.z; o.w -= lr * v.w;
      st4(p, o);
      if (shadow) st4(shadow + row * ld + col, o);
    } else {
      float x[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (j < nv) {
          float o = p[j] - lr * x[j];
          p[j] = o;
          if (shadow) shadow[row * ld + col + j] = __float2bfloat16_rn(o);
        }
    }
  }
};

}  // namespace ddnn
