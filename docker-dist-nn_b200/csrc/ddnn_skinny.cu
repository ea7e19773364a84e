// Skinny output layers (N <= 16, e.g. the 10-class softmax layer): HBM-bound kernels.
//
// A 64-wide GEMM tile wastes >80 % of its lanes at N = 10, so these layers get their own
// kernels whose cost is one coalesced pass over the big operand:
//   forward  out[b,:N] = act(x[b,:] . W + bias) : one warp per row, lane owns k = lane + 32 j,
//            W staged in shared memory (row pitch 20 floats: conflict-free float4 reads),
//            warp-shuffle reduction of the N partial sums
//   dX       out[b,k]  = (dz[b,:N] . W[k,:N]) * f'(a[b,k]) : thread owns k (W row in registers),
//            dz rows broadcast from shared memory, a/out coalesced over k
//   dW       part[z][k,:N] = sum_{b in chunk z} x[b,k] dz[b,:N] ; db partials alongside:
//            thread owns k, accumulators in registers, x coalesced over k, dz broadcast
#include "ddnn_internal.h"

namespace ddnn {
namespace {

constexpr int WPITCH = 20;   // floats per staged W row (16 used): bank-conflict-free float4 reads
constexpr int ROWS_TILE = 32;  // dz rows staged per shared-memory tile in dX / dW

template <typename T>
__device__ __forceinline__ float ld1(const T* p) { return to_f32(*p); }

template <int NV4, typename TI, typename TO>
__global__ void __launch_bounds__(256)
skinny_fwd_kernel(const TI* __restrict__ X, const TI* __restrict__ W, const float* __restrict__ bias,
                  TO* __restrict__ out, int64_t B, int K, int N, int act) {
  extern __shared__ float ws[];  // [K][WPITCH]
  for (int i = threadIdx.x; i < K * 16; i += blockDim.x) {
    const int k = i >> 4, n = i & 15;
    ws[k * WPITCH + n] = n < N ? ld1(W + (int64_t)k * N + n) : 0.f;
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  for (int64_t row = (int64_t)blockIdx.x * wpb + warp; row < B; row += (int64_t)gridDim.x * wpb) {
    float acc[NV4 * 4];
#pragma unroll
    for (int j = 0; j < NV4 * 4; ++j) acc[j] = 0.f;
    const TI* xr = X + row * K;
    for (int k = lane; k < K; k += 32) {
      const float x = ld1(xr + k);
      const float4* wr = reinterpret_cast<const float4*>(ws + k * WPITCH);
#pragma unroll
      for (int q = 0; q < NV4; ++q) {
        const float4 w = wr[q];
        acc[4 * q + 0] = fmaf(x, w.x, acc[4 * q + 0]);
        acc[4 * q + 1] = fmaf(x, w.y, acc[4 * q + 1]);
        acc[4 * q + 2] = fmaf(x, w.z, acc[4 * q + 2]);
        acc[4 * q + 3] = fmaf(x, w.w, acc[4 * q + 3]);
      }
    }
#pragma unroll
    for (int j = 0; j < NV4 * 4; ++j) acc[j] = warp_sum(acc[j]);
    // lane n writes output n
    float v = 0.f;
#pragma unroll
    for (int j = 0; j < NV4 * 4; ++j)
      if (lane == j) v = acc[j];
    if (lane < N) out[row * N + lane] = from_f32<TO>(act_fwd(v + __ldg(bias + lane), act));
  }
}

// Vector variant of the forward (rows 16-byte aligned: K % E == 0, E = elements per 16 bytes):
// 4 lanes share a row, a warp covers 8 rows per iteration.  Lane g of a row loads the 16-byte
// chunks g, g+4, g+8, ... (all of them up front: up to 8 loads in flight per lane), multiplies
// against W rows staged in shared memory in a permuted order (the 4 rows read together by the 4
// lanes of a group sit next to each other -> conflict-free), and the 4 partial sums per output are
// combined with 2 shuffle steps.
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

template <int NV4, int MAXC, typename TI, typename TO>
__global__ void __launch_bounds__(256)
skinny_fwd_vec_kernel(const TI* __restrict__ X, const TI* __restrict__ W, const float* __restrict__ bias,
                      TO* __restrict__ out, int64_t B, int K, int N, int act) {
  constexpr int E = Chunk<TI>::E;
  // [Kp][WPITCH] with Kp = K rounded up to 4E; row k is stored at s(k) = (k / 4E) * 4E + (k % E) * 4 + (k / E) % 4,
  // rows of the last partial group that no k maps to stay zero
  extern __shared__ float ws[];
  const int Kp = (K + 4 * E - 1) / (4 * E) * (4 * E);
  for (int i = threadIdx.x; i < Kp * WPITCH; i += blockDim.x) ws[i] = 0.f;
  __syncthreads();
  for (int i = threadIdx.x; i < K * 16; i += blockDim.x) {
    const int k = i >> 4, n = i & 15;
    const int srow = (k / (4 * E)) * (4 * E) + (k % E) * 4 + (k / E) % 4;
    if (n < N) ws[srow * WPITCH + n] = ld1(W + (int64_t)k * N + n);
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  const int g = lane & 3, rsub = lane >> 2;
  const int nchunks = K / E;               // 16-byte chunks per row
  const int per_lane = (nchunks + 3) / 4;  // <= MAXC
  for (int64_t r0 = ((int64_t)blockIdx.x * wpb + warp) * 8; r0 < B; r0 += (int64_t)gridDim.x * wpb * 8) {
    const int64_t row = r0 + rsub;
    const bool row_ok = row < B;
    uint4 raw[MAXC];
#pragma unroll
    for (int i = 0; i < MAXC; ++i) {
      raw[i] = make_uint4(0, 0, 0, 0);
      const int ch = g + 4 * i;
      if (row_ok && i < per_lane && ch < nchunks) raw[i] = *reinterpret_cast<const uint4*>(X + row * K + ch * E);
    }
    float acc[NV4 * 4];
#pragma unroll
    for (int j = 0; j < NV4 * 4; ++j) acc[j] = 0.f;
#pragma unroll
    for (int i = 0; i < MAXC; ++i) {
      if (i < per_lane) {  // warp-uniform
        float x[E];
        unpack16<TI>(raw[i], x);
        // chunk (g + 4i) covers k = (g + 4i) * E + j  ->  s = i * 4E + j * 4 + g  (< Kp; rows past K hold zeros)
#pragma unroll
        for (int j = 0; j < E; ++j) {
          const float4* wr = reinterpret_cast<const float4*>(ws + (i * 4 * E + j * 4 + g) * WPITCH);
#pragma unroll
          for (int q = 0; q < NV4; ++q) {
            const float4 w = wr[q];
            acc[4 * q + 0] = fmaf(x[j], w.x, acc[4 * q + 0]);
            acc[4 * q + 1] = fmaf(x[j], w.y, acc[4 * q + 1]);
            acc[4 * q + 2] = fmaf(x[j], w.z, acc[4 * q + 2]);
            acc[4 * q + 3] = fmaf(x[j], w.w, acc[4 * q + 3]);
          }
        }
      }
    }
#pragma unroll
    for (int j = 0; j < NV4 * 4; ++j) {
      acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], 1);
      acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], 2);
    }
    // lane g of the row writes outputs g, g+4, g+8, g+12
    if (row_ok) {
#pragma unroll
      for (int q = 0; q < NV4; ++q) {
        const int n = g + 4 * q;
        float v = acc[4 * q];
#pragma unroll
        for (int e = 1; e < 4; ++e)
          if (g == e) v = acc[4 * q + e];
        // acc[4q + e] holds output column 4q + e; lane g takes e = g  -> column 4q + g
        if (4 * q + g < N) out[row * N + 4 * q + g] = from_f32<TO>(act_fwd(v + __ldg(bias + 4 * q + g), act));
        (void)n;
      }
    }
  }
}

template <int NV4, typename T>
__global__ void __launch_bounds__(256)
skinny_dx_kernel(const T* __restrict__ dZ, const T* __restrict__ W, const T* __restrict__ Aprev, T* __restrict__ out,
                 int64_t B, int K, int N, int act, int rows_per_block) {
  __shared__ __align__(16) float dzs[ROWS_TILE][16];
  const int k = blockIdx.y * blockDim.x + threadIdx.x;
  float w[NV4 * 4];
#pragma unroll
  for (int j = 0; j < NV4 * 4; ++j) w[j] = (k < K && j < N) ? ld1(W + (int64_t)k * N + j) : 0.f;
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
  const int64_t r1 = min(B, r0 + rows_per_block);
  for (int64_t rb = r0; rb < r1; rb += ROWS_TILE) {
    const int nr = (int)min((int64_t)ROWS_TILE, r1 - rb);
    __syncthreads();
    for (int i = threadIdx.x; i < ROWS_TILE * 16; i += blockDim.x) {
      const int r = i >> 4, n = i & 15;
      dzs[r][n] = (r < nr && n < N) ? ld1(dZ + (rb + r) * N + n) : 0.f;
    }
    __syncthreads();
    if (k < K) {
#pragma unroll
      for (int rg = 0; rg < ROWS_TILE; rg += 16) {
        float a8[16];  // the 16 A_prev loads of this group are issued before the first use
#pragma unroll
        for (int u = 0; u < 16; ++u)
          a8[u] = (act != DDNN_ACT_LINEAR && rg + u < nr) ? ld1(Aprev + (rb + rg + u) * K + k) : 1.f;
#pragma unroll
        for (int u = 0; u < 16; ++u) {
          if (rg + u < nr) {
            float acc = 0.f;
#pragma unroll
            for (int q = 0; q < NV4; ++q) {
              const float4 d = *reinterpret_cast<const float4*>(&dzs[rg + u][4 * q]);
              acc = fmaf(d.x, w[4 * q], acc);
              acc = fmaf(d.y, w[4 * q + 1], acc);
              acc = fmaf(d.z, w[4 * q + 2], acc);
              acc = fmaf(d.w, w[4 * q + 3], acc);
            }
            const float gq = act == DDNN_ACT_LINEAR ? 1.f : act_bwd(a8[u], act);
            out[(rb + rg + u) * K + k] = from_f32<T>(acc * gq);
          }
        }
      }
    }
  }
}

template <int NV4, typename T>
__global__ void __launch_bounds__(256)
skinny_dw_kernel(const T* __restrict__ X, const T* __restrict__ dZ, float* __restrict__ dw_part,
                 float* __restrict__ db_part, int64_t B, int K, int N, int np, int rows_per_block) {
  __shared__ __align__(16) float dzs[ROWS_TILE][16];
  const int k = blockIdx.y * blockDim.x + threadIdx.x;
  const int z = blockIdx.x;
  float acc[NV4 * 4];
#pragma unroll
  for (int j = 0; j < NV4 * 4; ++j) acc[j] = 0.f;
  float dbs = 0.f;  // thread t < 16 of k-tile 0 owns column t of db
  const int64_t r0 = (int64_t)z * rows_per_block;
  const int64_t r1 = min(B, r0 + rows_per_block);
  for (int64_t rb = r0; rb < r1; rb += ROWS_TILE) {
    const int nr = (int)min((int64_t)ROWS_TILE, r1 - rb);
    __syncthreads();
    for (int i = threadIdx.x; i < ROWS_TILE * 16; i += blockDim.x) {
      const int r = i >> 4, n = i & 15;
      dzs[r][n] = (r < nr && n < N) ? ld1(dZ + (rb + r) * N + n) : 0.f;
    }
    __syncthreads();
    if (blockIdx.y == 0 && threadIdx.x < 16)
      for (int r = 0; r < nr; ++r) dbs += dzs[r][threadIdx.x];
    if (k < K) {
#pragma unroll
      for (int rg = 0; rg < ROWS_TILE; rg += 16) {
        float x8[16];  // 16 independent loads in flight before the FMAs (rows past nr contribute x = 0)
#pragma unroll
        for (int u = 0; u < 16; ++u) x8[u] = rg + u < nr ? ld1(X + (rb + rg + u) * K + k) : 0.f;
#pragma unroll
        for (int u = 0; u < 16; ++u) {
#pragma unroll
          for (int q = 0; q < NV4; ++q) {
            const float4 d = *reinterpret_cast<const float4*>(&dzs[rg + u][4 * q]);
            acc[4 * q + 0] = fmaf(x8[u], d.x, acc[4 * q + 0]);
            acc[4 * q + 1] = fmaf(x8[u], d.y, acc[4 * q + 1]);
            acc[4 * q + 2] = fmaf(x8[u], d.z, acc[4 * q + 2]);
            acc[4 * q + 3] = fmaf(x8[u], d.w, acc[4 * q + 3]);
          }
        }
      }
    }
  }
  if (k < K) {
    float* o = dw_part + ((int64_t)z * K + k) * np;
#pragma unroll
    for (int j = 0; j < NV4 * 4; ++j)
      if (j < np) o[j] = acc[j];
  }
  if (blockIdx.y == 0 && threadIdx.x < np) db_part[(int64_t)z * np + threadIdx.x] = dbs;
}

int nv4_of(int64_t N) { return (int)((N + 3) / 4); }

}  // namespace

bool skinny_ok(int64_t K, int64_t N) { return N >= 1 && N <= 16 && K >= 1 && (K + 32) * WPITCH * 4 <= 160 * 1024; }

template <typename TI, typename TO>
static int skinny_fwd_t(const void* X, const void* W, const float* bias, void* A, int64_t B, int64_t K, int64_t N,
                        int act, cudaStream_t st) {
  constexpr int E = Chunk<TI>::E;
  const size_t smem = (size_t)((K + 4 * E - 1) / (4 * E) * (4 * E)) * WPITCH * sizeof(float);
  const bool vec = (K % E == 0) && (K / E <= 4 * 8) && (reinterpret_cast<uintptr_t>(X) % 16 == 0);
  long blocks = vec ? (B + 63) / 64 : (B + 7) / 8;
  if (blocks > 148 * 4) blocks = 148 * 4;
#define DDNN_SK_LAUNCH(KERN)                                                                                  \
  do {                                                                                                        \
    auto kern = KERN;                                                                                         \
    if (smem > 48 * 1024)                                                                                     \
      if (int rc_ = ensure_dynamic_smem((const void*)kern, smem)) return rc_;                                 \
    kern<<<(unsigned)blocks, 256, smem, st>>>((const TI*)X, (const TI*)W, bias, (TO*)A, B, (int)K, (int)N, act); \
  } while (0)
  const int nv = nv4_of(N);
  if (vec) {
    const bool small = K / E <= 4 * 4;  // <= 4 chunks per lane: keep the register footprint small
    if (small) {
      switch (nv) {
        case 1: DDNN_SK_LAUNCH((skinny_fwd_vec_kernel<1, 4, TI, TO>)); break;
        case 2: DDNN_SK_LAUNCH((skinny_fwd_vec_kernel<2, 4, TI, TO>)); break;
        case 3: DDNN_SK_LAUNCH((skinny_fwd_vec_kernel<3, 4, TI, TO>)); break;
        default: DDNN_SK_LAUNCH((skinny_fwd_vec_kernel<4, 4, TI, TO>)); break;
      }
    } else {
      switch (nv) {
        case 1: DDNN_SK_LAUNCH((skinny_fwd_vec_kernel<1, 8, TI, TO>)); break;
        case 2: DDNN_SK_LAUNCH((skinny_fwd_vec_kernel<2, 8, TI, TO>)); break;
        case 3: DDNN_SK_LAUNCH((skinny_fwd_vec_kernel<3, 8, TI, TO>)); break;
        default: DDNN_SK_LAUNCH((skinny_fwd_vec_kernel<4, 8, TI, TO>)); break;
      }
    }
  } else {
    switch (nv) {
      case 1: DDNN_SK_LAUNCH((skinny_fwd_kernel<1, TI, TO>)); break;
      case 2: DDNN_SK_LAUNCH((skinny_fwd_kernel<2, TI, TO>)); break;
      case 3: DDNN_SK_LAUNCH((skinny_fwd_kernel<3, TI, TO>)); break;
      default: DDNN_SK_LAUNCH((skinny_fwd_kernel<4, TI, TO>)); break;
    }
  }
#undef DDNN_SK_LAUNCH
  count_launch();
  DDNN_LAUNCH_CHECK();
  return DDNN_OK;
}

int skinny_forward(const void* X, const void* W, const float* bias, void* A, int64_t B, int64_t K, int64_t N, int act,
                   int dtype, int out_dtype, cudaStream_t st) {
  if (dtype == DDNN_F32 && out_dtype == DDNN_F32) return skinny_fwd_t<float, float>(X, W, bias, A, B, K, N, act, st);
  if (dtype == DDNN_BF16 && out_dtype == DDNN_BF16) return skinny_fwd_t<bf16, bf16>(X, W, bias, A, B, K, N, act, st);
  if (dtype == DDNN_BF16 && out_dtype == DDNN_F32) return skinny_fwd_t<bf16, float>(X, W, bias, A, B, K, N, act, st);
  set_error("skinny_forward: unsupported dtype combination in=%d out=%d", dtype, out_dtype);
  return DDNN_ERR_UNSUPPORTED;
}

static int rows_per_block_for(int64_t B, int ktiles) {
  // >= 4 CTAs per SM (enough loads in flight to cover HBM latency), whole shared-memory tiles per block
  long want_blocks = (4 * 148 + ktiles - 1) / ktiles;
  long rpb = (B + want_blocks - 1) / want_blocks;
  rpb = ((rpb + ROWS_TILE - 1) / ROWS_TILE) * ROWS_TILE;
  if (rpb < ROWS_TILE) rpb = ROWS_TILE;
  return (int)rpb;
}

template <typename T>
static int skinny_dx_t(const void* dZ, const void* W, const void* Aprev, void* out, int64_t B, int64_t K, int64_t N,
                       int act, cudaStream_t st) {
  const int ktiles = cdiv(K, 256);
  const int rpb = ROWS_TILE;  // one shared-memory tile per CTA: B/32 CTAs keep enough loads in flight
  dim3 grid(cdiv(B, rpb), ktiles);
#define DDNN_SK_DX(NV) \
  skinny_dx_kernel<NV, T><<<grid, 256, 0, st>>>((const T*)dZ, (const T*)W, (const T*)Aprev, (T*)out, B, (int)K, (int)N, act, rpb)
  switch (nv4_of(N)) {
    case 1: DDNN_SK_DX(1); break;
    case 2: DDNN_SK_DX(2); break;
    case 3: DDNN_SK_DX(3); break;
    default: DDNN_SK_DX(4); break;
  }
#undef DDNN_SK_DX
  count_launch();
  DDNN_LAUNCH_CHECK();
  return DDNN_OK;
}

int skinny_dx(const void* dZ, const void* W, const void* Aprev, void* dZprev, int64_t B, int64_t K, int64_t N,
              int act_prev, int dtype, cudaStream_t st) {
  if (dtype == DDNN_F32) return skinny_dx_t<float>(dZ, W, Aprev, dZprev, B, K, N, act_prev, st);
  if (dtype == DDNN_BF16) return skinny_dx_t<bf16>(dZ, W, Aprev, dZprev, B, K, N, act_prev, st);
  set_error("skinny_dx: unsupported dtype %d", dtype);
  return DDNN_ERR_UNSUPPORTED;
}

int skinny_dw_splits(int64_t B, int64_t K, int64_t N) {
  const int ktiles = cdiv(K, 256);
  const int rpb = rows_per_block_for(B, ktiles);
  return cdiv(B, rpb);
}

template <typename T>
static int skinny_dw_t(const void* X, const void* dZ, float* dw_part, float* db_part, int splits, int64_t B, int64_t K,
                       int64_t N, cudaStream_t st) {
  const int ktiles = cdiv(K, 256);
  const int rpb = rows_per_block_for(B, ktiles);
  const int np = (int)((N + 3) & ~(int64_t)3);
  dim3 grid(splits, ktiles);
#define DDNN_SK_DW(NV) \
  skinny_dw_kernel<NV, T><<<grid, 256, 0, st>>>((const T*)X, (const T*)dZ, dw_part, db_part, B, (int)K, (int)N, np, rpb)
  switch (nv4_of(N)) {
    case 1: DDNN_SK_DW(1); break;
    case 2: DDNN_SK_DW(2); break;
    case 3: DDNN_SK_DW(3); break;
    default: DDNN_SK_DW(4); break;
  }
#undef DDNN_SK_DW
  count_launch();
  DDNN_LAUNCH_CHECK();
  return DDNN_OK;
}

int skinny_dw_partials(const void* X, const void* dZ, float* dw_part, float* db_part, int splits, int64_t B, int64_t K,
                       int64_t N, int dtype, cudaStream_t st) {
  if (dtype == DDNN_F32) return skinny_dw_t<float>(X, dZ, dw_part, db_part, splits, B, K, N, st);
  if (dtype == DDNN_BF16) return skinny_dw_t<bf16>(X, dZ, dw_part, db_part, splits, B, K, N, st);
  set_error("skinny_dw: unsupported dtype %d", dtype);
  return DDNN_ERR_UNSUPPORTED;
}

}  // namespace ddnn
