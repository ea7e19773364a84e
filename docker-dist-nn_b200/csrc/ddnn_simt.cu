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

// ---- fp32, 16-byte aligned operands: 4-stage cp.async pipeline ----------------------------------
// The register-prefetch kernel above keeps ONE k-slab in flight: at the narrow-layer sizes (a slab of
// 16 k is ~300 issue cycles per warp) that is far less than the ~800-cycle HBM latency, and ncu showed
// it memory-latency bound (FMA pipe 36 % busy, 1 700 cycles per slab).  Here three slabs are always in
// flight as cp.async (LDGSTS) 16-byte copies straight into shared memory in their natural layout
// (zero-filled past the edges), so the loads of slab s+3 overlap the FMAs of slab s.
//   A_KMAJ  (forward: A = X [M][K])  -> As[BM][BK+4], thread reads 4 k of its rows as one float4
//   !A_KMAJ (dW:      A = X [K'][M]) -> As[BK][BM+4] as before
//   B is always [K][N] row-major      -> Bs[BK][BN+4]
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem, int src_bytes) {
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(gmem), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int BM, int BN, bool A_KMAJ, typename Epi, bool COLSUM>
__global__ void __launch_bounds__(NTHREADS, 2)
simt_gemm_async_kernel(const float* __restrict__ A, const float* __restrict__ Bm, int M, int N, int K, int64_t lda,
                       int64_t ldb, int klen, bool veco, Epi epi, int64_t zstride, float* __restrict__ colsum_part,
                       int np) {
  constexpr int ST = 4;
  constexpr int TM = BM / 16, TN = BN / 16;
  constexpr int A_FLOATS = A_KMAJ ? BM * (BK + 4) : BK * (BM + 4);
  constexpr int B_FLOATS = BK * (BN + 4);
  extern __shared__ __align__(16) float sm[];
  float* As = sm;                    // [ST][A_FLOATS]
  float* Bs = sm + ST * A_FLOATS;    // [ST][B_FLOATS]

  const int t = threadIdx.x, tx = t & 15, ty = t >> 4;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN, z = blockIdx.z;
  const int kbeg = z * klen;
  const int kend = min(K, kbeg + klen);
  const int nslab = (kend - kbeg + BK - 1) / BK;

  auto load_stage = [&](int st, int k0) {
    float* as = As + st * A_FLOATS;
    float* bs = Bs + st * B_FLOATS;
#pragma unroll
    for (int i = 0; i < BM / 64; ++i) {
      const int f = t + i * NTHREADS;
      if (A_KMAJ) {
        const int m = f >> 2, kq = (f & 3) * 4;
        const int rem = (m0 + m < M) ? (kend - (k0 + kq)) * 4 : 0;
        const int nb = rem < 0 ? 0 : (rem > 16 ? 16 : rem);
        cp_async16(as + m * (BK + 4) + kq, nb ? A + (int64_t)(m0 + m) * lda + k0 + kq : A, nb);
      } else {
        const int kk = f / (BM / 4), mq = (f % (BM / 4)) * 4;
        const int rem = (k0 + kk < kend) ? (M - (m0 + mq)) * 4 : 0;
        const int nb = rem < 0 ? 0 : (rem > 16 ? 16 : rem);
        cp_async16(as + kk * (BM + 4) + mq, nb ? A + (int64_t)(k0 + kk) * lda + m0 + mq : A, nb);
      }
    }
#pragma unroll
    for (int i = 0; i < BN / 64; ++i) {
      const int f = t + i * NTHREADS;
      const int kk = f / (BN / 4), nq = (f % (BN / 4)) * 4;
      const int rem = (k0 + kk < kend) ? (N - (n0 + nq)) * 4 : 0;
      const int nb = rem < 0 ? 0 : (rem > 16 ? 16 : rem);
      cp_async16(bs + kk * (BN + 4) + nq, nb ? Bm + (int64_t)(k0 + kk) * ldb + n0 + nq : Bm, nb);
    }
  };

  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;
  float cs[TN];
#pragma unroll
  for (int j = 0; j < TN; ++j) cs[j] = 0.f;
  const bool do_cs = COLSUM && blockIdx.y == 0 && ty == 0;

#pragma unroll
  for (int s = 0; s < ST - 1; ++s) {
    if (s < nslab) load_stage(s, kbeg + s * BK);
    cp_async_commit();
  }
  for (int s = 0; s < nslab; ++s) {
    cp_async_wait<ST - 2>();  // slab s has landed (for this thread's copies) ...
    __syncthreads();          // ... and for everyone's; all warps are also done with slab s-1
    if (s + ST - 1 < nslab) load_stage((s + ST - 1) % ST, kbeg + (s + ST - 1) * BK);
    cp_async_commit();
    const float* as = As + (s % ST) * A_FLOATS;
    const float* bs = Bs + (s % ST) * B_FLOATS;
#pragma unroll
    for (int kq = 0; kq < BK; kq += 4) {
      float a[4][TM], b[4][TN];  // [kk][row / col]
      if (A_KMAJ) {
#pragma unroll
        for (int i = 0; i < TM; ++i) {
          const int row = (i >> 2) * 64 + ty * 4 + (i & 3);
          const float4 v = *reinterpret_cast<const float4*>(as + row * (BK + 4) + kq);
          a[0][i] = v.x; a[1][i] = v.y; a[2][i] = v.z; a[3][i] = v.w;
        }
      } else {
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
#pragma unroll
          for (int r = 0; r < TM / 4; ++r) {
            const float4 v = *reinterpret_cast<const float4*>(as + (kq + kk) * (BM + 4) + r * 64 + ty * 4);
            a[kk][r * 4 + 0] = v.x; a[kk][r * 4 + 1] = v.y; a[kk][r * 4 + 2] = v.z; a[kk][r * 4 + 3] = v.w;
          }
      }
#pragma unroll
      for (int kk = 0; kk < 4; ++kk)
#pragma unroll
        for (int c = 0; c < TN / 4; ++c) {
          const float4 v = *reinterpret_cast<const float4*>(bs + (kq + kk) * (BN + 4) + c * 64 + tx * 4);
          b[kk][c * 4 + 0] = v.x; b[kk][c * 4 + 1] = v.y; b[kk][c * 4 + 2] = v.z; b[kk][c * 4 + 3] = v.w;
        }
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
        for (int i = 0; i < TM; ++i)
#pragma unroll
          for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[kk][i], b[kk][j], acc[i][j]);
        if (COLSUM) {
          if (do_cs) {
#pragma unroll
            for (int j = 0; j < TN; ++j) cs[j] += b[kk][j];
          }
        }
      }
    }
  }
  cp_async_wait<0>();

  if (COLSUM) {
    if (do_cs) {
#pragma unroll
      for (int j = 0; j < TN; ++j) {
        const int col = n0 + (j >> 2) * 64 + tx * 4 + (j & 3);
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

// ---- fp32-accurate tensor-core variant: 3xTF32 on mma.sync fragments -----------------------------
// The FFMA pipe issues one 3-register warp instruction per 2 cycles per SM sub-partition, so even a
// perfect FFMA kernel tops out near 36 TFLOP/s on this chip.  Same cp.async pipeline as above, but the
// inner product runs on the tensor cores with error-compensated TF32: every fp32 operand x is split in
// registers into hi = tf32(x) and lo = tf32(x - hi) and each 16x8x8 tile does  a_lo.b_hi + a_hi.b_lo +
// a_hi.b_hi  (the dropped a_lo.b_lo term is ~2^-22 relative), accumulating in fp32 -- within the 1e-5
// bound of the fp32 path.  Register fragments are reused across the three products and across the
// warp's 2x2 (4x4) MMA tiles, which is what makes this cheaper than an SS-operand tcgen05.mma.kind::tf32
// pipeline here: that one would re-read A and B from shared memory for each of the three products and is
// shared-memory-bandwidth bound at these tile shapes (128 B/clk for a 128x128x8 tf32 MMA per 64 clk).
__device__ __forceinline__ uint32_t to_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ void split_tf32(float x, uint32_t& hi, uint32_t& lo) {
  hi = to_tf32(x);
  lo = to_tf32(x - __uint_as_float(hi));
}
__device__ __forceinline__ void mma_tf32(float* c, const uint32_t* a, const uint32_t* b) {
  asm volatile(
      "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

template <int BM, int BN, bool A_KMAJ, typename Epi, bool COLSUM>
__global__ void __launch_bounds__(NTHREADS, 2)
tf32x3_gemm_kernel(const float* __restrict__ A, const float* __restrict__ Bm, int M, int N, int K, int64_t lda,
                   int64_t ldb, int klen, Epi epi, int64_t zstride, float* __restrict__ colsum_part, int np) {
  constexpr int ST = 4;
  constexpr int PA = A_KMAJ ? BK + 4 : BM + 8;  // pitches chosen so that every fragment load is conflict-free
  constexpr int PB = BN + 8;
  constexpr int A_FLOATS = A_KMAJ ? BM * PA : BK * PA;
  constexpr int B_FLOATS = BK * PB;
  constexpr int WM = BM / 2, WN = BN / 4;  // 8 warps as 2 (m) x 4 (n)
  constexpr int MT = WM / 16, NT = WN / 8;
  extern __shared__ __align__(16) float sm[];
  float* As = sm;
  float* Bs = sm + ST * A_FLOATS;

  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int g = lane >> 2, tig = lane & 3;
  const int wm0 = (warp >> 2) * WM, wn0 = (warp & 3) * WN;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN, z = blockIdx.z;
  const int kbeg = z * klen;
  const int kend = min(K, kbeg + klen);
  const int nslab = (kend - kbeg + BK - 1) / BK;

  auto load_stage = [&](int st, int k0) {
    float* as = As + st * A_FLOATS;
    float* bs = Bs + st * B_FLOATS;
#pragma unroll
    for (int i = 0; i < BM / 64; ++i) {
      const int f = t + i * NTHREADS;
      if (A_KMAJ) {
        const int m = f >> 2, kq = (f & 3) * 4;
        const int rem = (m0 + m < M) ? (kend - (k0 + kq)) * 4 : 0;
        const int nb = rem < 0 ? 0 : (rem > 16 ? 16 : rem);
        cp_async16(as + m * PA + kq, nb ? A + (int64_t)(m0 + m) * lda + k0 + kq : A, nb);
      } else {
        const int kk = f / (BM / 4), mq = (f % (BM / 4)) * 4;
        const int rem = (k0 + kk < kend) ? (M - (m0 + mq)) * 4 : 0;
        const int nb = rem < 0 ? 0 : (rem > 16 ? 16 : rem);
        cp_async16(as + kk * PA + mq, nb ? A + (int64_t)(k0 + kk) * lda + m0 + mq : A, nb);
      }
    }
#pragma unroll
    for (int i = 0; i < BN / 64; ++i) {
      const int f = t + i * NTHREADS;
      const int kk = f / (BN / 4), nq = (f % (BN / 4)) * 4;
      const int rem = (k0 + kk < kend) ? (N - (n0 + nq)) * 4 : 0;
      const int nb = rem < 0 ? 0 : (rem > 16 ? 16 : rem);
      cp_async16(bs + kk * PB + nq, nb ? Bm + (int64_t)(k0 + kk) * ldb + n0 + nq : Bm, nb);
    }
  };

  // two accumulators: the small cross terms (a_lo.b_hi + a_hi.b_lo, ~2^-11 of the big one) are summed
  // separately so that they are not rounded away against the running hi.hi sum; added once at the end
  float c[MT][NT][4], cl[MT][NT][4];
#pragma unroll
  for (int i = 0; i < MT; ++i)
#pragma unroll
    for (int j = 0; j < NT; ++j)
#pragma unroll
      for (int e = 0; e < 4; ++e) c[i][j][e] = cl[i][j][e] = 0.f;
  float cs[NT];
#pragma unroll
  for (int j = 0; j < NT; ++j) cs[j] = 0.f;
  const bool do_cs = COLSUM && blockIdx.y == 0 && (warp >> 2) == 0;

#pragma unroll
  for (int s = 0; s < ST - 1; ++s) {
    if (s < nslab) load_stage(s, kbeg + s * BK);
    cp_async_commit();
  }
  for (int s = 0; s < nslab; ++s) {
    cp_async_wait<ST - 2>();
    __syncthreads();
    if (s + ST - 1 < nslab) load_stage((s + ST - 1) % ST, kbeg + (s + ST - 1) * BK);
    cp_async_commit();
    const float* as = As + (s % ST) * A_FLOATS;
    const float* bs = Bs + (s % ST) * B_FLOATS;
#pragma unroll
    for (int k0 = 0; k0 < BK; k0 += 8) {
      uint32_t ah[MT][4], al[MT][4], bh[NT][2], bl[NT][2];
#pragma unroll
      for (int i = 0; i < MT; ++i) {
        const int r = wm0 + i * 16 + g;
        float x0, x1, x2, x3;  // (r, k0+tig) (r+8, k0+tig) (r, k0+tig+4) (r+8, k0+tig+4)
        if (A_KMAJ) {
          x0 = as[r * PA + k0 + tig];
          x1 = as[(r + 8) * PA + k0 + tig];
          x2 = as[r * PA + k0 + tig + 4];
          x3 = as[(r + 8) * PA + k0 + tig + 4];
        } else {
          x0 = as[(k0 + tig) * PA + r];
          x1 = as[(k0 + tig) * PA + r + 8];
          x2 = as[(k0 + tig + 4) * PA + r];
          x3 = as[(k0 + tig + 4) * PA + r + 8];
        }
        split_tf32(x0, ah[i][0], al[i][0]);
        split_tf32(x1, ah[i][1], al[i][1]);
        split_tf32(x2, ah[i][2], al[i][2]);
        split_tf32(x3, ah[i][3], al[i][3]);
      }
#pragma unroll
      for (int j = 0; j < NT; ++j) {
        const int n = wn0 + j * 8 + g;
        const float y0 = bs[(k0 + tig) * PB + n];      // (k0+tig, n)
        const float y1 = bs[(k0 + tig + 4) * PB + n];  // (k0+tig+4, n)
        if (COLSUM) {
          if (do_cs) cs[j] += y0 + y1;
        }
        split_tf32(y0, bh[j][0], bl[j][0]);
        split_tf32(y1, bh[j][1], bl[j][1]);
      }
#pragma unroll
      for (int i = 0; i < MT; ++i)
#pragma unroll
        for (int j = 0; j < NT; ++j) {
          mma_tf32(cl[i][j], al[i], bh[j]);
          mma_tf32(cl[i][j], ah[i], bl[j]);
          mma_tf32(c[i][j], ah[i], bh[j]);
        }
    }
  }
  cp_async_wait<0>();

  if (COLSUM) {
    if (do_cs) {
#pragma unroll
      for (int j = 0; j < NT; ++j) {
        float v = cs[j];
        v += __shfl_xor_sync(0xffffffffu, v, 1);
        v += __shfl_xor_sync(0xffffffffu, v, 2);
        const int col = n0 + wn0 + j * 8 + g;
        if (tig == 0 && col < N) colsum_part[(int64_t)z * np + col] = v;
      }
    }
  }
  // C fragment: c0 (g, 2tig) c1 (g, 2tig+1) c2 (g+8, 2tig) c3 (g+8, 2tig+1)
  const int64_t zoff = (int64_t)z * zstride;
#pragma unroll
  for (int i = 0; i < MT; ++i)
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int row = m0 + wm0 + i * 16 + g + h * 8;
      if (row >= M) continue;
#pragma unroll
      for (int j = 0; j < NT; ++j) {
        const int col = n0 + wn0 + j * 8 + 2 * tig;
        if (col >= N) continue;
        epi.apply4(row + zoff, col,
                   make_float4(c[i][j][2 * h] + cl[i][j][2 * h], c[i][j][2 * h + 1] + cl[i][j][2 * h + 1], 0.f, 0.f),
                   min(2, N - col), false);
      }
    }
}

template <int BM, int BN, bool A_KMAJ>
constexpr size_t tf32_smem_bytes() {
  return (size_t)4 * ((A_KMAJ ? BM * (BK + 4) : BK * (BM + 8)) + BK * (BN + 8)) * sizeof(float);
}

template <int BM, int BN, bool A_KMAJ>
constexpr size_t async_smem_bytes() {
  return (size_t)4 * ((A_KMAJ ? BM * (BK + 4) : BK * (BM + 4)) + BK * (BN + 4)) * sizeof(float);
}

template <typename T>
bool vec_ok(const T* p, int64_t ld) {
  return (ld % 4 == 0) && ((reinterpret_cast<uintptr_t>(p) % (4 * sizeof(T))) == 0);
}

template <int BM, int BN, bool A_KMAJ, typename Epi, bool COLSUM>
int launch_async(const float* A, const float* Bm, int M, int N, int K, int64_t lda, int64_t ldb, int splits, int klen,
                 bool veco, Epi epi, int64_t zstride, float* cs_part, int np, cudaStream_t st) {
  auto kern = simt_gemm_async_kernel<BM, BN, A_KMAJ, Epi, COLSUM>;
  constexpr size_t smem = async_smem_bytes<BM, BN, A_KMAJ>();
  if (smem > 48 * 1024)
    if (int rc = ensure_dynamic_smem((const void*)kern, smem)) return rc;
  dim3 grid(cdiv(N, BN), cdiv(M, BM), splits);
  kern<<<grid, NTHREADS, smem, st>>>(A, Bm, M, N, K, lda, ldb, klen, veco, epi, zstride, cs_part, np);
  count_launch();
  DDNN_LAUNCH_CHECK();
  return DDNN_OK;
}

template <int BM, int BN, bool A_KMAJ, typename Epi, bool COLSUM>
int launch_tf32(const float* A, const float* Bm, int M, int N, int K, int64_t lda, int64_t ldb, int splits, int klen,
                Epi epi, int64_t zstride, float* cs_part, int np, cudaStream_t st) {
  auto kern = tf32x3_gemm_kernel<BM, BN, A_KMAJ, Epi, COLSUM>;
  constexpr size_t smem = tf32_smem_bytes<BM, BN, A_KMAJ>();
  if (smem > 48 * 1024)
    if (int rc = ensure_dynamic_smem((const void*)kern, smem)) return rc;
  dim3 grid(cdiv(N, BN), cdiv(M, BM), splits);
  kern<<<grid, NTHREADS, smem, st>>>(A, Bm, M, N, K, lda, ldb, klen, epi, zstride, cs_part, np);
  count_launch();
  DDNN_LAUNCH_CHECK();
  return DDNN_OK;
}

std::atomic<int> g_f32_core{0};  // 0 = FFMA (default: exact fp32 FMAs), 1 = 3xTF32 tensor-core inner product

template <typename T> struct IsF32 { static constexpr bool value = false; };
template <> struct IsF32<float> { static constexpr bool value = true; };

template <bool A_KMAJ, bool B_KMAJ, typename TA, typename TB, typename Epi, bool COLSUM>
int launch(const TA* A, const TB* Bm, int M, int N, int K, int64_t lda, int64_t ldb, int splits, bool veco,
           Epi epi, int64_t zstride, float* cs_part, int np, cudaStream_t st) {
  const bool veca = vec_ok(A, lda), vecb = vec_ok(Bm, ldb);
  int klen = cdiv(cdiv(K, splits), BK) * BK;
  if (klen <= 0) klen = BK;
  const long big_tiles = (long)cdiv(M, 128) * cdiv(N, 128) * splits;
  if constexpr (IsF32<TA>::value && IsF32<TB>::value && !B_KMAJ) {
    if (veca && vecb && g_f32_core.load() == 1) {  // cp.async pipeline + 3xTF32 tensor-core inner product
      if (big_tiles >= 148 && N > 64)
        return launch_tf32<128, 128, A_KMAJ, Epi, COLSUM>((const float*)A, (const float*)Bm, M, N, K, lda, ldb, splits,
                                                          klen, epi, zstride, cs_part, np, st);
      return launch_tf32<64, 64, A_KMAJ, Epi, COLSUM>((const float*)A, (const float*)Bm, M, N, K, lda, ldb, splits,
                                                      klen, epi, zstride, cs_part, np, st);
    }
    if (veca && vecb) {  // 16-byte aligned rows: the cp.async pipeline, FFMA inner product
      if (big_tiles >= 148 && N > 64)
        return launch_async<128, 128, A_KMAJ, Epi, COLSUM>((const float*)A, (const float*)Bm, M, N, K, lda, ldb, splits,
                                                           klen, veco, epi, zstride, cs_part, np, st);
      return launch_async<64, 64, A_KMAJ, Epi, COLSUM>((const float*)A, (const float*)Bm, M, N, K, lda, ldb, splits,
                                                       klen, veco, epi, zstride, cs_part, np, st);
    }
  }
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

void simt_set_f32_core(int tensor) { g_f32_core.store(tensor ? 1 : 0); }

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
  // fill the 2 x 148 resident CTA slots (launch bounds: 2 CTAs / SM) in ONE wave: tiles x splits <= 296
  // (rounding up instead spilled 16 of 312 CTAs into a second wave on the MNIST layer: 2x the time),
  // at least 4 k-slabs per split
  long tiles = (long)cdiv(K, 64) * cdiv(N, 64);
  long want = (2 * 148) / tiles;
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
