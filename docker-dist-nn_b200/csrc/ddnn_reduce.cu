// Deterministic split-K reduction fused with the dW/db epilogue (SGD step or fp32 store),
// plus the bias-gradient column reduction used beside the tcgen05 dW GEMM.
#include "ddnn_internal.h"

namespace ddnn {
namespace {

// dw_part: [splits][rows][np] fp32, db_part: [splits][np] (the bias row).  A group of G lanes
// (1, 8 or 32) owns 4 consecutive output columns: lane g sums splits g, g+G, ... (loads batched 4
// at a time), then the G partial sums are combined by a fixed xor-shuffle tree -- the order depends
// only on (splits, G), so the result is run-to-run deterministic.
template <int G, typename EpiW, typename EpiB>
__global__ void __launch_bounds__(256)
reduce_dw_kernel(const float* __restrict__ dw_part, const float* __restrict__ db_part, int splits_w, int splits_b,
                 int64_t rows,
                 int np, int N, EpiW ew, EpiB eb, bool vecw, bool vecb, int has_db) {
  pdl_launch_dependents();
  pdl_wait();  // the partials come from the previous kernel of the stream
  const int q = np >> 2;
  const int64_t total = (rows + (has_db ? 1 : 0)) * q;
  const int sub = threadIdx.x % G;
  const int64_t groups = (int64_t)gridDim.x * blockDim.x / G;
  const int64_t g0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / G;
  // uniform trip count for every thread: the shuffles below are executed by whole warps
  for (int64_t it = 0; it * groups < total; ++it) {
    const int64_t idx = g0 + it * groups;
    const bool active = idx < total;
    const int64_t row = active ? idx / q : 0;
    const int col = active ? (int)(idx - row * q) * 4 : 0;
    const bool is_w = row < rows;
    const float* base = is_w ? dw_part + row * np + col : db_part + col;
    const int64_t zs = is_w ? rows * np : np;
    const int splits = is_w ? splits_w : splits_b;  // the bias row may come from a different split plan
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    if (active) {
      // the lane that will apply the update starts fetching its W (or bias) operand now, not after the reduction
      if (sub == 0 && col < N) {
        if (is_w) ew.prefetch(row, col);
        else eb.prefetch(0, col);
      }
      // 4 split-loads per round, all issued before the first add (also in the last, partial round: a scalar tail
      // loop would serialise its loads behind the adds)
      for (int z = sub; z < splits; z += 4 * G) {
        float4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
          v[u] = (z + u * G < splits) ? ld4(base + (int64_t)(z + u * G) * zs) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int u = 0; u < 4; ++u) { s.x += v[u].x; s.y += v[u].y; s.z += v[u].z; s.w += v[u].w; }
      }
    }
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) {
      s.x += __shfl_xor_sync(0xffffffffu, s.x, o);
      s.y += __shfl_xor_sync(0xffffffffu, s.y, o);
      s.z += __shfl_xor_sync(0xffffffffu, s.z, o);
      s.w += __shfl_xor_sync(0xffffffffu, s.w, o);
    }
    if (active && sub == 0 && col < N) {
      const int nv = min(4, N - col);
      if (is_w) ew.apply4(row, col, s, nv, vecw);
      else eb.apply4(0, col, s, nv, vecb);
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
colsum_kernel(const T* __restrict__ dZ, float* __restrict__ part, int64_t B, int N, int np, int rlen, bool vec) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float4 red[8][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int col = blockIdx.x * 128 + lane * 4;
  const int z = blockIdx.y;
  const int64_t r0 = (int64_t)z * rlen;
  const int64_t r1 = min(B, r0 + rlen);
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
  if (col < N) {
    for (int64_t r = r0 + warp; r < r1; r += 8) {
      const T* p = dZ + r * N + col;
      float4 v;
      if (vec && col + 4 <= N) {
        v = ld4(p);
      } else {
        v.x = to_f32(p[0]);
        v.y = col + 1 < N ? to_f32(p[1]) : 0.f;
        v.z = col + 2 < N ? to_f32(p[2]) : 0.f;
        v.w = col + 3 < N ? to_f32(p[3]) : 0.f;
      }
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
  }
  red[warp][lane] = s;
  __syncthreads();
  if (warp == 0 && col < N) {
    float4 a = red[0][lane];
#pragma unroll
    for (int w = 1; w < 8; ++w) {
      float4 v = red[w][lane];
      a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
    }
    st4(part + (int64_t)z * np + col, a);  // np is a multiple of 4: the pad columns exist
  }
}

// bf16, N % 8 == 0: lane owns 8 consecutive columns (one 16-byte load per row), each warp keeps 8
// row-loads in flight; 8 warps stride the rows of the split and are combined through shared memory.
__global__ void __launch_bounds__(256)
colsum_bf16x8_kernel(const bf16* __restrict__ dZ, float* __restrict__ part, int64_t B, int N, int np, int rlen) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float red[8][32][9];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int col = blockIdx.x * 256 + lane * 8;
  const int z = blockIdx.y;
  const int64_t r0 = (int64_t)z * rlen;
  const int64_t r1 = min(B, r0 + rlen);
  float s[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) s[e] = 0.f;
  if (col < N) {
    for (int64_t r = r0 + warp * 8; r < r1; r += 64) {
      uint4 v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        v[u] = make_uint4(0, 0, 0, 0);
        if (r + u < r1) v[u] = *reinterpret_cast<const uint4*>(dZ + (r + u) * N + col);
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v[u]);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 f = __bfloat1622float2(h[e]);
          s[2 * e] += f.x;
          s[2 * e + 1] += f.y;
        }
      }
    }
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) red[warp][lane][e] = s[e];
  __syncthreads();
  if (warp == 0 && col < N) {
    float a[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) a[e] = red[0][lane][e];
#pragma unroll
    for (int w = 1; w < 8; ++w)
#pragma unroll
      for (int e = 0; e < 8; ++e) a[e] += red[w][lane][e];
    st4(part + (int64_t)z * np + col, make_float4(a[0], a[1], a[2], a[3]));
    st4(part + (int64_t)z * np + col + 4, make_float4(a[4], a[5], a[6], a[7]));
  }
}

// W -= lr * G (and the bf16 shadow), bias -= lr * gb: the update after the last gradient chunk.  Flat over the
// K*N elements (any N); 16-byte accesses when every pointer allows it.
__global__ void __launch_bounds__(256)
sgd_apply_kernel(float* __restrict__ W, bf16* __restrict__ shadow, float* __restrict__ bias, const float* __restrict__ G,
                 const float* __restrict__ gb, float lr, int64_t total, int N, bool vec) {
  pdl_launch_dependents();
  pdl_wait();
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (vec) {
    for (int64_t i = tid * 4; i < total; i += stride * 4) {
      float4 w = ld4(W + i);
      const float4 g = ld4(G + i);
      w.x -= lr * g.x; w.y -= lr * g.y; w.z -= lr * g.z; w.w -= lr * g.w;
      st4(W + i, w);
      if (shadow) st4(shadow + i, w);
    }
  } else {
    for (int64_t i = tid; i < total; i += stride) {
      const float w = W[i] - lr * G[i];
      W[i] = w;
      if (shadow) shadow[i] = __float2bfloat16_rn(w);
    }
  }
  for (int64_t i = tid; i < N; i += stride) bias[i] -= lr * gb[i];
}

template <typename T>
bool vec_ok(const T* p, int64_t ld) {
  return p && (ld % 4 == 0) && ((reinterpret_cast<uintptr_t>(p) % (4 * sizeof(T))) == 0);
}

}  // namespace

int reduce_dw(const float* dw_part, const float* db_part, int splits, int64_t K, int64_t N, float* W_master,
              void* W_shadow, float* bias, float* dW_out, float* db_out, float lr, int apply, cudaStream_t st) {
  return reduce_dw2(dw_part, db_part, splits, splits, K, N, W_master, W_shadow, bias, dW_out, db_out, lr, apply, st);
}

int reduce_dw2(const float* dw_part, const float* db_part, int splits, int splits_b, int64_t K, int64_t N,
               float* W_master,
              void* W_shadow, float* bias, float* dW_out, float* db_out, float lr, int apply, cudaStream_t st) {
  const int np = (int)((N + 3) & ~(int64_t)3);
  const int64_t rows = dw_part ? K : 0;  // no dW partials: only the bias row is reduced
  const int has_db = db_part != nullptr;
  const int64_t total = (rows + has_db) * (np / 4);
  if (total == 0) return DDNN_OK;
  const int smax = splits > splits_b ? splits : splits_b;
  const int G = smax <= 8 ? 1 : (smax <= 64 ? 8 : 32);
  // grid: whole groups per block; the grid-stride loop needs gridDim*256 to be a multiple of G (it is)
  long blocks = (long)((total * G + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  if (blocks < 1) blocks = 1;
#define DDNN_REDUCE(GG, EW, EB, ew, eb, vw, vb)                                                              \
  DDNN_CUDA_CHECK(launch_pdl(reduce_dw_kernel<GG, EW, EB>, dim3((unsigned)blocks), dim3(256), 0, st, dw_part, db_part, \
                             splits, splits_b, rows, np, (int)N, ew, eb, vw, vb, has_db))
  if (apply) {
    EpiSgd ew{W_master, (bf16*)W_shadow, N, lr};
    EpiSgd eb{bias, nullptr, N, lr};
    const bool vw = vec_ok(W_master, N) && (!W_shadow || vec_ok((bf16*)W_shadow, N));
    const bool vb = vec_ok(bias, N);
    if (G == 1) DDNN_REDUCE(1, EpiSgd, EpiSgd, ew, eb, vw, vb);
    else if (G == 8) DDNN_REDUCE(8, EpiSgd, EpiSgd, ew, eb, vw, vb);
    else DDNN_REDUCE(32, EpiSgd, EpiSgd, ew, eb, vw, vb);
  } else {
    EpiStoreF32 ew{dW_out, N};
    EpiStoreF32 eb{db_out, N};
    const bool vw = vec_ok(dW_out, N), vb = vec_ok(db_out, N);
    if (G == 1) DDNN_REDUCE(1, EpiStoreF32, EpiStoreF32, ew, eb, vw, vb);
    else if (G == 8) DDNN_REDUCE(8, EpiStoreF32, EpiStoreF32, ew, eb, vw, vb);
    else DDNN_REDUCE(32, EpiStoreF32, EpiStoreF32, ew, eb, vw, vb);
  }
#undef DDNN_REDUCE
  count_launch();
  return DDNN_OK;
}

int sgd_apply(float* W, void* W_shadow, float* bias, const float* G, const float* gb, float lr, int64_t K, int64_t N,
              cudaStream_t st) {
  const int64_t total = K * N;
  const bool vec = total % 4 == 0 && reinterpret_cast<uintptr_t>(W) % 16 == 0 && reinterpret_cast<uintptr_t>(G) % 16 == 0 &&
                   (!W_shadow || reinterpret_cast<uintptr_t>(W_shadow) % 8 == 0);
  long blocks = (long)((total / (vec ? 4 : 1) + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  if (blocks < 1) blocks = 1;
  DDNN_CUDA_CHECK(launch_pdl(sgd_apply_kernel, dim3((unsigned)blocks), dim3(256), 0, st, W, (bf16*)W_shadow, bias, G, gb,
                             lr, total, (int)N, vec));
  count_launch();
  return DDNN_OK;
}

int colsum_splits(int64_t B, int64_t N) {
  long strips = cdiv(N, 256);
  long want = (2 * 148 + strips - 1) / strips;
  long max_by_rows = B / 64;
  if (max_by_rows < 1) max_by_rows = 1;
  long s = want < max_by_rows ? want : max_by_rows;
  if (s > 512) s = 512;
  return (int)(s < 1 ? 1 : s);
}

int colsum_partials(const void* dZ, float* db_part, int splits, int64_t B, int64_t N, int dtype, cudaStream_t st) {
  const int np = (int)((N + 3) & ~(int64_t)3);
  const int rlen = cdiv(B, splits);
  dim3 grid(cdiv(N, 128), splits);
  if (dtype == DDNN_F32) {
    DDNN_CUDA_CHECK(launch_pdl(colsum_kernel<float>, grid, dim3(256), 0, st, (const float*)dZ, db_part, B, (int)N, np,
                               rlen, vec_ok((const float*)dZ, N)));
  } else if (dtype == DDNN_BF16 && N % 8 == 0 && reinterpret_cast<uintptr_t>(dZ) % 16 == 0) {
    dim3 grid8(cdiv(N, 256), splits);
    DDNN_CUDA_CHECK(launch_pdl(colsum_bf16x8_kernel, grid8, dim3(256), 0, st, (const bf16*)dZ, db_part, B, (int)N, np,
                               rlen));
  } else if (dtype == DDNN_BF16) {
    DDNN_CUDA_CHECK(launch_pdl(colsum_kernel<bf16>, grid, dim3(256), 0, st, (const bf16*)dZ, db_part, B, (int)N, np,
                               rlen, vec_ok((const bf16*)dZ, N)));
  } else {
    set_error("colsum: unsupported dtype %d", dtype);
    return DDNN_ERR_UNSUPPORTED;
  }
  count_launch();
  DDNN_LAUNCH_CHECK();
  return DDNN_OK;
}

}  // namespace ddnn
