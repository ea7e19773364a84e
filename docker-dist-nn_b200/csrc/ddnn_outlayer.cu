// Fused training step of a skinny OUTPUT layer (N <= 16 classes, K <= 256 inputs): one pass over the
// layer's input activations A (B x K) does
//   logits = A.W + b  ->  softmax  ->  row loss, dZ = (p - onehot) * inv_denom          (phase 1)
//   dZprev = (dZ.W^T) * f'(A)      and the split partials of  dW = A^T.dZ, db = sum dZ   (phase 2)
// instead of five launches (forward, softmax-CE, loss partials, dX, dW) that each stream A or a
// same-sized tensor through HBM.  A is read from HBM once (phase 2 re-reads the 64-row tile it just
// touched from L1/L2); dZprev is the only large write.  The split partials go through the same
// deterministic reducer as every other dW (reduce_dw), which applies the SGD update.
//
// Phase 1 maps 4 lanes to a row and 8 rows to a warp (16-byte loads, W staged in shared memory in the
// permuted conflict-free order of skinny_fwd_vec_kernel); phase 2 maps a thread to an input feature k
// (W row in registers, dZ rows broadcast from shared memory, 16 A loads in flight).
#include "ddnn_internal.h"

namespace ddnn {
namespace {

constexpr int WPITCH = 20;
constexpr int TILE_ROWS = 64;  // 8 warps x 8 rows

template <typename T> struct Chunk;
template <> struct Chunk<float> { static constexpr int E = 4; };
template <> struct Chunk<bf16> { static constexpr int E = 8; };

template <typename T>
__device__ __forceinline__ void unpack16(const uint4& r, float* x) {
  if (Chunk<T>::E == 4) {
    x[0] = __uint_as_float(r.x); x[1] = __uint_as_float(r.y); x[2] = __uint_as_float(r.z); x[3] = __uint_as_float(r.w);
  } else {
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&r);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 f = __bfloat1622float2(h[e]);
      x[2 * e] = f.x; x[2 * e + 1] = f.y;
    }
  }
}

template <int NV4, int MAXC, typename T>
__global__ void __launch_bounds__(256)
outlayer_fused_kernel(const T* __restrict__ A, const T* __restrict__ W, const float* __restrict__ bias,
                      const int32_t* __restrict__ labels, T* __restrict__ dZ, float* __restrict__ row_loss,
                      T* __restrict__ dZprev, float* __restrict__ dw_part, float* __restrict__ db_part, int64_t B,
                      int K, int N, int np, int act_prev, float inv_denom, int rows_per_block) {
  constexpr int E = Chunk<T>::E;
  constexpr int NV = NV4 * 4;
  extern __shared__ float ws[];  // permuted W for phase 1: [Kp][WPITCH]
  __shared__ __align__(16) float dzs[TILE_ROWS][16];
  const int Kp = (K + 4 * E - 1) / (4 * E) * (4 * E);
  for (int i = threadIdx.x; i < Kp * WPITCH; i += blockDim.x) ws[i] = 0.f;
  __syncthreads();
  for (int i = threadIdx.x; i < K * 16; i += blockDim.x) {
    const int k = i >> 4, n = i & 15;
    const int srow = (k / (4 * E)) * (4 * E) + (k % E) * 4 + (k / E) % 4;
    if (n < N) ws[srow * WPITCH + n] = to_f32(W[(int64_t)k * N + n]);
  }
  // phase-2 ownership: thread k holds W[k][0..N) and its dW accumulators
  const int k = threadIdx.x;
  float wk[NV], dwacc[NV];
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    wk[j] = (k < K && j < N) ? to_f32(W[(int64_t)k * N + j]) : 0.f;
    dwacc[j] = 0.f;
  }
  float dbs = 0.f;
  float bj[NV];
#pragma unroll
  for (int j = 0; j < NV; ++j) bj[j] = j < N ? __ldg(bias + j) : 0.f;
  __syncthreads();

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane & 3, rsub = lane >> 2;
  const int nchunks = K / E, per_lane = (nchunks + 3) / 4;
  const int64_t r_begin = (int64_t)blockIdx.x * rows_per_block;
  const int64_t r_end = min(B, r_begin + rows_per_block);

  for (int64_t rb = r_begin; rb < r_end; rb += TILE_ROWS) {
    // ---------------- phase 1: logits, softmax, loss, dZ for 64 rows ----------------
    const int rt = warp * 8 + rsub;  // row inside the tile
    const int64_t row = rb + rt;
    const bool row_ok = row < r_end;
    uint4 raw[MAXC];
#pragma unroll
    for (int i = 0; i < MAXC; ++i) {
      raw[i] = make_uint4(0, 0, 0, 0);
      const int ch = g + 4 * i;
      if (row_ok && i < per_lane && ch < nchunks) raw[i] = *reinterpret_cast<const uint4*>(A + row * K + ch * E);
    }
    float z[NV];
#pragma unroll
    for (int j = 0; j < NV; ++j) z[j] = 0.f;
#pragma unroll
    for (int i = 0; i < MAXC; ++i) {
      if (i < per_lane) {
        float x[E];
        unpack16<T>(raw[i], x);
#pragma unroll
        for (int j = 0; j < E; ++j) {
          const float4* wr = reinterpret_cast<const float4*>(ws + (i * 4 * E + j * 4 + g) * WPITCH);
#pragma unroll
          for (int q = 0; q < NV4; ++q) {
            const float4 w = wr[q];
            z[4 * q + 0] = fmaf(x[j], w.x, z[4 * q + 0]);
            z[4 * q + 1] = fmaf(x[j], w.y, z[4 * q + 1]);
            z[4 * q + 2] = fmaf(x[j], w.z, z[4 * q + 2]);
            z[4 * q + 3] = fmaf(x[j], w.w, z[4 * q + 3]);
          }
        }
      }
    }
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      z[j] += __shfl_xor_sync(0xffffffffu, z[j], 1);
      z[j] += __shfl_xor_sync(0xffffffffu, z[j], 2);
      z[j] += bj[j];
    }
    // every lane of the row group now holds all N logits: softmax-CE in registers
    float m = -INFINITY;
#pragma unroll
    for (int j = 0; j < NV; ++j)
      if (j < N) m = fmaxf(m, z[j]);
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < NV; ++j)
      if (j < N) s += expf(z[j] - m);
    const float inv = 1.0f / s;
    const int y = row_ok ? labels[row] : -1;
    if (g == 0) {
      float d[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) d[j] = 0.f;
      float zy = 0.f;
#pragma unroll
      for (int j = 0; j < NV; ++j) {
        if (j < N) {
          const float pj = expf(z[j] - m) * inv;
          d[j] = row_ok ? (pj - (j == y ? 1.f : 0.f)) * inv_denom : 0.f;
          if (j == y) zy = z[j];
        }
      }
#pragma unroll
      for (int q = 0; q < 4; ++q)
        *reinterpret_cast<float4*>(&dzs[rt][4 * q]) = make_float4(d[4 * q], d[4 * q + 1], d[4 * q + 2], d[4 * q + 3]);
      if (row_ok) {
        row_loss[row] = (y >= 0 && y < N) ? -((zy - m) - logf(s)) : 0.f;
        if (dZ) {
#pragma unroll
          for (int j = 0; j < NV; ++j)
            if (j < N) dZ[row * N + j] = from_f32<T>(d[j]);
        }
      }
    }
    __syncthreads();
    // ---------------- phase 2: dZprev rows and dW / db partial sums ----------------
    if (threadIdx.x < 16) {
#pragma unroll 8
      for (int r = 0; r < TILE_ROWS; ++r) dbs += dzs[r][threadIdx.x];
    }
    if (k < K) {
#pragma unroll 1  // keep the body (16 rows x N FMAs) resident in the instruction cache: fully unrolled, ncu showed
                  // "no instruction" as a top stall and 2x the instruction count
      for (int rg = 0; rg < TILE_ROWS; rg += 16) {
        float a16[16];
#pragma unroll
        for (int u = 0; u < 16; ++u) a16[u] = (rb + rg + u < r_end) ? to_f32(A[(rb + rg + u) * K + k]) : 0.f;
#pragma unroll
        for (int u = 0; u < 16; ++u) {
          float dot = 0.f;
#pragma unroll
          for (int q = 0; q < NV4; ++q) {
            const float4 d = *reinterpret_cast<const float4*>(&dzs[rg + u][4 * q]);
            dot = fmaf(d.x, wk[4 * q], dot);
            dot = fmaf(d.y, wk[4 * q + 1], dot);
            dot = fmaf(d.z, wk[4 * q + 2], dot);
            dot = fmaf(d.w, wk[4 * q + 3], dot);
            dwacc[4 * q + 0] = fmaf(a16[u], d.x, dwacc[4 * q + 0]);
            dwacc[4 * q + 1] = fmaf(a16[u], d.y, dwacc[4 * q + 1]);
            dwacc[4 * q + 2] = fmaf(a16[u], d.z, dwacc[4 * q + 2]);
            dwacc[4 * q + 3] = fmaf(a16[u], d.w, dwacc[4 * q + 3]);
          }
          if (dZprev && rb + rg + u < r_end)
            dZprev[(rb + rg + u) * K + k] = from_f32<T>(dot * act_bwd(a16[u], act_prev));
        }
      }
    }
    __syncthreads();  // dzs is rewritten by the next tile
  }
  const int zb = blockIdx.x;
  if (k < K) {
    float* o = dw_part + ((int64_t)zb * K + k) * np;
#pragma unroll
    for (int j = 0; j < NV; ++j)
      if (j < np) o[j] = dwacc[j];
  }
  if (threadIdx.x < np) db_part[(int64_t)zb * np + threadIdx.x] = dbs;
}

int rows_per_block(int64_t B) {
  long want_blocks = 4 * 148;
  long rpb = (B + want_blocks - 1) / want_blocks;
  rpb = ((rpb + TILE_ROWS - 1) / TILE_ROWS) * TILE_ROWS;
  return (int)(rpb < TILE_ROWS ? TILE_ROWS : rpb);
}

}  // namespace

// usable when the layer is skinny (N <= 16), K <= 256 (one input feature per thread in phase 2) and the
// rows of A are 16-byte aligned vectors of at most 16 chunks per lane
bool outlayer_fused_ok(const void* A, int64_t K, int64_t N, int dtype) {
  if (N < 1 || N > 16 || K < 1 || K > 256) return false;
  const int e = dtype == DDNN_F32 ? 4 : 8;
  if (K % e != 0 || (K / e) > 4 * 16) return false;
  return (reinterpret_cast<uintptr_t>(A) % 16) == 0;
}

int outlayer_splits(int64_t B) { return cdiv(B, rows_per_block(B)); }

template <typename T>
static int outlayer_launch(const void* A, const void* W, const float* bias, const int32_t* labels, void* dZ,
                           float* row_loss, void* dZprev, float* dw_part, float* db_part, int64_t B, int64_t K,
                           int64_t N, int act_prev, float inv_denom, cudaStream_t st) {
  constexpr int E = Chunk<T>::E;
  const int np = (int)((N + 3) & ~(int64_t)3);
  const int rpb = rows_per_block(B);
  const int blocks = cdiv(B, rpb);
  const size_t smem = (size_t)((K + 4 * E - 1) / (4 * E) * (4 * E)) * WPITCH * sizeof(float);
  const int per_lane = (int)((K / E + 3) / 4);
  const int nv4 = (int)((N + 3) / 4);
#define DDNN_OUT_LAUNCH(NV, MC)                                                                              \
  outlayer_fused_kernel<NV, MC, T><<<blocks, 256, smem, st>>>(                                                \
      (const T*)A, (const T*)W, bias, labels, (T*)dZ, row_loss, (T*)dZprev, dw_part, db_part, B, (int)K, (int)N, \
      np, act_prev, inv_denom, rpb)
#define DDNN_OUT_NV(MC)                        \
  switch (nv4) {                               \
    case 1: DDNN_OUT_LAUNCH(1, MC); break;     \
    case 2: DDNN_OUT_LAUNCH(2, MC); break;     \
    case 3: DDNN_OUT_LAUNCH(3, MC); break;     \
    default: DDNN_OUT_LAUNCH(4, MC); break;    \
  }
  if (per_lane <= 4) { DDNN_OUT_NV(4) }
  else if (per_lane <= 8) { DDNN_OUT_NV(8) }
  else { DDNN_OUT_NV(16) }
#undef DDNN_OUT_NV
#undef DDNN_OUT_LAUNCH
  count_launch();
  DDNN_LAUNCH_CHECK();
  return DDNN_OK;
}

int outlayer_fused(const void* A, const void* W, const float* bias, const int32_t* labels, void* dZ, float* row_loss,
                   void* dZprev, float* dw_part, float* db_part, int64_t B, int64_t K, int64_t N, int act_prev,
                   float inv_denom, int dtype, cudaStream_t st) {
  if (dtype == DDNN_F32)
    return outlayer_launch<float>(A, W, bias, labels, dZ, row_loss, dZprev, dw_part, db_part, B, K, N, act_prev,
                                  inv_denom, st);
  if (dtype == DDNN_BF16)
    return outlayer_launch<bf16>(A, W, bias, labels, dZ, row_loss, dZprev, dw_part, db_part, B, K, N, act_prev,
                                 inv_denom, st);
  set_error("outlayer_fused: unsupported dtype %d", dtype);
  return DDNN_ERR_UNSUPPORTED;
}

}  // namespace ddnn
