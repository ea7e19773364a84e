// SIMT fp32-accumulate GEMM family for the narrow / misaligned layers and the fp32 path.
//
// One register-tiled kernel, three operand layouts (no materialised transposes):
//   forward  Z = X.W      : A = X  [M=B][K] k-contiguous,  B = W  [K][N] n-contiguous   (grpc_node.py:87)
//   dX       dA = dZ.W^T  : A = dZ [M=B][K'=N] contiguous, B = W  [N'=K][K'=N] k'-contiguous
//   dW       dW = X^T.dZ  : A = X  [K'=B][M=K] m-contiguous, B = dZ [K'=B][N] n-contiguous
// Tiles are staged k-major in shared memory ([BK][BM], [BK][BN]) with register prefetch of
// the next k-slab (double buffer, one __syncthreads per slab); every global access is a
// 16-byte (fp32) / 8-byte (bf16) vector along the contiguous dimension when the pitch
// allows, scalar with bounds checks otherwise (N = 10 output layers).
#include "ddnn_internal.h"

namespace ddnn {

namespace {

constexpr int BK = 16;
constexpr int NTHREADS = 256;

template <typename T>
__device__ __forceinline__ float4 gload4(const T* __restrict__ base, int64_t row, int col, int64_t rows,
                                         int cols, int64_t ld, bool vec) {
  float4 r = make_float4(0.f, 0.f, 0.f, 0.f);
  if (row >= rows || col >= cols) return r;
  const T* p = base + row * ld + col;
  if (vec && col + 4 <= cols) return ld4(p);
  r.x = to_f32(p[0]);
  if (col + 1 < cols) r.y = to_f32(p[1]);
  if (col + 2 < cols) r.z = to_f32(p[2]);
  if (col + 3 < cols) r.w = to_f32(p[3]);
  return r;
}

template <int BM, int BN, bool A_KMAJ, bool B_KMAJ, typename TA, typename TB, typename Epi, bool COLSUM>
__global__ void __launch_bounds__(NTHREADS, 2)
simt_gemm_kernel(const TA* __restrict__ A, const TB* __restrict__ Bm, int M, int N, int K, int64_t lda,
                 int64_t ldb, int klen, bool veca, bool vecb, bool veco, Epi epi, int64_t zstride,
                 float* __restrict__ colsum_part, int np) {
  constexpr int TM = BM / 16, TN = BN / 16;
  constexpr int LA = BM / 64, LB = BN / 64;  // float4 loads per thread per slab
  __shared__ __align__(16) float As[2][BK][BM + 4];
  __shared__ __align__(16) float Bs[2][BK][BN + 4];

  const int t = threadIdx.x, tx = t & 15, ty = t >> 4;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN, z = blockIdx.z;
  const int kbeg = z * klen;
  const int kend = min(K, kbeg + klen);
  const int nslab = (kend - kbeg + BK - 1) / BK;

  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;
  float cs[TN];
#pragma unroll
  for (int j = 0; j < TN; ++j) cs[j] = 0.f;
  const bool do_cs = COLSUM && blockIdx.y == 0 && ty == 0;

  float4 ra[LA], rb[LB];

  auto gload = [&](int k0) {
#pragma unroll
    for (int i = 0; i < LA; ++i) {
      int f = t + i * NTHREADS;
      if (A_KMAJ) {
        int m = f >> 2, kq = f & 3;
        ra[i] = gload4(A, (int64_t)m0 + m, k0 + kq * 4, M, kend, lda, veca);
      } else {
        int kk = f / (BM / 4), mq = f % (BM / 4);
        ra[i] = gload4(A, (int64_t)k0 + kk, m0 + mq * 4, kend, M, lda, veca);
      }
    }
#pragma unroll
    for (int i = 0; i < LB; ++i) {
      int f = t + i * NTHREADS;
      if (B_KMAJ) {
        int n = f >> 2, kq = f & 3;
        rb[i] = gload4(Bm, (int64_t)n0 + n, k0 + kq * 4, N, kend, ldb, vecb);
      } else {
        int kk = f / (BN / 4), nq = f % (BN / 4);
        rb[i] = gload4(Bm, (int64_t)k0 + kk, n0 + nq * 4, kend, N, ldb, vecb);
      }
    }
  };
  auto sstore = [&](int buf) {
#pragma unroll
    for (int i = 0; i < LA; ++i) {
      int f = t + i * NTHREADS;
      if (A_KMAJ) {
        int m = f >> 2, kq = f & 3;
        As[buf][kq * 4 + 0][m] = ra[i].x;
        As[buf][kq * 4 + 1][m] = ra[i].y;
        As[buf][kq * 4 + 2][m] = ra[i].z;
        As[buf][kq * 4 + 3][m] = ra[i].w;
      } else {
        int kk = f / (BM / 4), mq = f % (BM / 4);
        *reinterpret_cast<float4*>(&As[buf][kk][mq * 4]) = ra[i];
      }
    }
#pragma unroll
    for (int i = 0; i < LB; ++i) {
      int f = t + i * NTHREADS;
      if (B_KMAJ) {
        int n = f >> 2, kq = f & 3;
        Bs[buf][kq * 4 + 0][n] = rb[i].x;
        Bs[buf][kq * 4 + 1][n] = rb[i].y;
        Bs[buf][kq * 4 + 2][n] = rb[i].z;
        Bs[buf][kq * 4 + 3][n] = rb[i].w;
      } else {
        int kk = f / (BN / 4), nq = f % (BN / 4);
        *reinterpret_cast<float4*>(&Bs[buf][kk][nq * 4]) = rb[i];
      }
    }
  };

  if (nslab > 0) {
    gload(kbeg);
    sstore(0);
  }
  __syncthreads();

  for (int s = 0; s < nslab; ++s) {
    const int buf = s & 1;
    if (s + 1 < nslab) gload(kbeg + (s + 1) * BK);
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float a[TM], b[TN];
#pragma unroll
      for (int r = 0; r < TM / 4; ++r) {
        float4 v = *reinterpret_cast<const float4*>(&As[buf][kk][r * 64 + ty * 4]);
        a[r * 4 + 0] = v.x; a[r * 4 + 1] = v.y; a[r * 4 + 2] = v.z; a[r * 4 + 3] = v.w;
      }
#pragma unroll
      for (int c = 0; c < TN / 4; ++c) {
        float4 v = *reinterpret_cast<const float4*>(&Bs[buf][kk][c * 64 + tx * 4]);
        b[c * 4 + 0] = v.x; b[c * 4 + 1] = v.y; b[c * 4 + 2] = v.z; b[c * 4 + 3] = v.w;
      }
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
      if (COLSUM) {
        if (do_cs) {
#pragma unroll
          for (int j = 0; j < TN; ++j) cs[j] += b[j];
        }
      }
    }
    if (s + 1 < nslab) sstore(buf ^ 1);
    __syncthreads();
  }

  if (COLSUM) {
    // partial slab z of the split-K workspace
    if (do_cs) {
#pragma unroll
      for (int j = 0; j < TN; ++j) {
        int col = n0 + (j >> 2) * 64 + tx * 4 + (j & 3);
        if (col < N) colsum_part[(int64_t)z * np + col] = cs[j];
      }
    }
  }
  const int64_t zoff = (int64_t)z * zstride;
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    const int row = m0 + (i >> 2) * 64 + ty * 4 + (i & 3);
    if (row >= M) continue;
#pragma unroll
    for (int c = 0; c < TN / 4; ++c) {
      const int col = n0 + c * 64 + tx * 4;
      if (col >= N) continue;
      const int nv = min(4, N - col);
      epi.apply4(row + zoff, col, make_float4(acc[i][c * 4], acc[i][c * 4 + 1], acc[i][c * 4 + 2], acc[i][c * 4 + 3]),
                 nv, veco);
    }
  }
}

template <typename T>
bool vec_ok(const T* p, int64_t ld) {
  return (ld % 4 == 0) && ((reinterpret_cast<uintptr_t>(p) % (4 * sizeof(T))) == 0);
}

template <bool A_KMAJ, bool B_KMAJ, typename TA, typename TB, typename Epi, bool COLSUM>
int launch(const TA* A, const TB* Bm, int M, int N, int K, int64_t lda, int64_t ldb, int splits, bool veco,
           Epi epi, int64_t zstride, float* cs_part, int np, cudaStream_t st) {
  const bool veca = vec_ok(A, lda), vecb = vec_ok(Bm, ldb);
  int klen = cdiv(cdiv(K, splits), BK) * BK;
  if (klen <= 0) klen = BK;
  const long big_tiles = (long)cdiv(M, 128) * cdiv(N, 128) * splits;
  if (big_tiles >= 148 && N > 64) {
    dim3 grid(cdiv(N, 128), cdiv(M, 128), splits);
    simt_gemm_kernel<128, 128, A_KMAJ, B_KMAJ, TA, TB, Epi, COLSUM><<<grid, NTHREADS, 0, st>>>(
        A, Bm, M, N, K, lda, ldb, klen, veca, vecb, veco, epi, zstride, cs_part, np);
  } else {
    dim3 grid(cdiv(N, 64), cdiv(M, 64), splits);
    simt_gemm_kernel<64, 64, A_KMAJ, B_KMAJ, TA, TB, Epi, COLSUM><<<grid, NTHREADS, 0, st>>>(
        A, Bm, M, N, K, lda, ldb, klen, veca, vecb, veco, epi, zstride, cs_part, np);
  }
  count_launch();
  DDNN_LAUNCH_CHECK();
  return DDNN_OK;
}

template <typename TI, typename TO>
int fwd_t(const void* X, const void* W, const float* bias, void* A, int64_t B, int64_t K, int64_t N, int act,
          cudaStream_t st) {
  EpiFwd<TO> epi{bias, (TO*)A, N, act};
  return launch<true, false, TI, TI, EpiFwd<TO>, false>((const TI*)X, (const TI*)W, (int)B, (int)N, (int)K, K, N, 1,
                                                        vec_ok((TO*)A, N), epi, 0, nullptr, 0, st);
}

template <typename T>
int dx_t(const void* dZ, const void* W, const void* Aprev, void* out, int64_t B, int64_t K, int64_t N, int act,
         cudaStream_t st) {
  EpiDx<T> epi{(const T*)Aprev, (T*)out, K, act};
  // M = B, N' = K (layer input width), K' = N (layer output width)
  const bool veco = vec_ok((T*)out, K) && (act == DDNN_ACT_LINEAR || vec_ok((const T*)Aprev, K));
  return launch<true, true, T, T, EpiDx<T>, false>((const T*)dZ, (const T*)W, (int)B, (int)K, (int)N, N, N, 1, veco,
                                                   epi, 0, nullptr, 0, st);
}

}  // namespace

int simt_forward(const void* X, const void* W, const float* bias, void* A, int64_t B, int64_t K, int64_t N,
                 int act, int dtype, int out_dtype, cudaStream_t st) {
  if (dtype == DDNN_F32 && out_dtype == DDNN_F32) return fwd_t<float, float>(X, W, bias, A, B, K, N, act, st);
  if (dtype == DDNN_BF16 && out_dtype == DDNN_BF16) return fwd_t<bf16, bf16>(X, W, bias, A, B, K, N, act, st);
  if (dtype == DDNN_BF16 && out_dtype == DDNN_F32) return fwd_t<bf16, float>(X, W, bias, A, B, K, N, act, st);
  set_error("simt_forward: unsupported dtype combination in=%d out=%d", dtype, out_dtype);
  return DDNN_ERR_UNSUPPORTED;
}

int simt_dx(const void* dZ, const void* W, const void* Aprev, void* dZprev, int64_t B, int64_t K, int64_t N,
            int act_prev, int dtype, cudaStream_t st) {
  if (dtype == DDNN_F32) return dx_t<float>(dZ, W, Aprev, dZprev, B, K, N, act_prev, st);
  if (dtype == DDNN_BF16) return dx_t<bf16>(dZ, W, Aprev, dZprev, B, K, N, act_prev, st);
  set_error("simt_dx: unsupported dtype %d", dtype);
  return DDNN_ERR_UNSUPPORTED;
}

int simt_dw_splits(int64_t B, int64_t K, int64_t N) {
  // enough (tile x split) CTAs for >= 2 waves of 148 SMs, at least 4 k-slabs per split
  long tiles = (long)cdiv(K, 64) * cdiv(N, 64);
  long want = (2 * 148 + tiles - 1) / tiles;
  long max_by_k = B / (4 * BK);
  if (max_by_k < 1) max_by_k = 1;
  long s = want < max_by_k ? want : max_by_k;
  if (s < 1) s = 1;
  if (s > 256) s = 256;
  return (int)s;
}

int simt_dw_partials(const void* X, const void* dZ, float* dw_part, float* db_part, int splits, int64_t B,
                     int64_t K, int64_t N, int dtype, cudaStream_t st) {
  const int np = (int)((N + 3) & ~(int64_t)3);
  EpiStoreF32 epi{dw_part, np};
  // M = K (layer input width), N = N, K' = B (batch)
  if (dtype == DDNN_F32)
    return launch<false, false, float, float, EpiStoreF32, true>((const float*)X, (const float*)dZ, (int)K, (int)N,
                                                                (int)B, K, N, splits, true, epi, K, db_part, np, st);
  if (dtype == DDNN_BF16)
    return launch<false, false, bf16, bf16, EpiStoreF32, true>((const bf16*)X, (const bf16*)dZ, (int)K, (int)N, (int)B,
                                                              K, N, splits, true, epi, K, db_part, np, st);
  set_error("simt_dw: unsupported dtype %d", dtype);
  return DDNN_ERR_UNSUPPORTED;
}

}  // namespace ddnn
