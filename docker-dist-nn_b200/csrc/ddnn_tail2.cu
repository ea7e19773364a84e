// Fused training step of the last TWO layers of a small fp32 MLP (the MNIST class of the reference's
// configs: 784-128-64-10 and 784-32-16-10): one kernel takes the activations A1 (B x K1) that feed the
// hidden layer (K1 -> N2) and does
//   A2 = f2(A1.W2 + b2)            logits = A2.W3 + b3 -> softmax-CE -> row loss, dZ3
//   dZ2 = (dZ3.W3^T) * f2'(A2)     dZ1 = (dZ2.W2^T) * f1'(A1)                 (dZ1 -> HBM)
//   split partials of dW2 = A1^T.dZ2, db2, dW3 = A2^T.dZ3, db3               (-> the deterministic reducer)
// for a tile of <= 32 rows per CTA.  A2, dZ2, dZ3 and the logits never leave the SM; A1 is read once.  At
// 4096 rows these layers are launch-latency-bound as separate kernels (forward 8.5 us + fused output layer
// 17 us + dX 8 us on the critical path of an 81 us step, each about 7 us of fixed cost); fused they are one
// launch of ~0.2 GFLOP.  fp32 FFMA with register tiles: (2 rows x 4 cols), (2 rows x 8 k) and (4 k x 8 n) per
// thread, operands broadcast or 16-byte vectors from shared memory, W2 staged twice ([k][n] and [n][k]) so
// that both products read it conflict-free.
#include "ddnn_internal.h"

namespace ddnn {
namespace {

constexpr int T2_ROWS = 32;    // row capacity of a tile
constexpr int T2_MAXK = 128;   // K1 <= 128: 16 lanes x 8 k in the dZ1 product, 32 x 4 k in the dW2 product
constexpr int T2_MAXN = 64;    // N2 <= 64: 16 lanes x 4 cols in the forward product, 8 x 8 n in the dW2 product
constexpr int T2_A2P = T2_MAXN + 8;  // pitch of A2 rows: 4 rows of a warp land in distinct banks in the layer-3 phase

struct Tail2Smem {
  float w2s[T2_MAXK][T2_MAXN];   // W2[k][n]
  float w2t[T2_MAXN][T2_MAXK];   // W2^T[n][k]
  float w3t[16][T2_MAXN];        // W3^T[c][f], classes padded to 16 with zeros
  float a1s[T2_ROWS][T2_MAXK];
  float a2s[T2_ROWS][T2_A2P];
  float dz2s[T2_ROWS][T2_MAXN];
  float dz3s[T2_ROWS][16];
  float b2s[T2_MAXN];
  float b3s[16];
};

__global__ void __launch_bounds__(256)
tail2_kernel(const float* __restrict__ A1, const float* __restrict__ W2, const float* __restrict__ b2,
             const float* __restrict__ W3, const float* __restrict__ b3, const int32_t* __restrict__ labels,
             float* __restrict__ dZ1, float* __restrict__ row_loss, float* __restrict__ dw2_part,
             float* __restrict__ db2_part, float* __restrict__ dw3_part, float* __restrict__ db3_part, int64_t B, int K1,
             int N2, int N3, int np3, int act1, int act2, float inv_denom, int rows_per_tile, int num_tiles) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  Tail2Smem& s = *reinterpret_cast<Tail2Smem*>(smem_raw);
  const int t = threadIdx.x;
  pdl_launch_dependents();
  pdl_wait();

  // ---- stage the weights (zero padded): all of a thread's 16-byte loads are issued before the first store ----
  {
    constexpr int PER = T2_MAXK * (T2_MAXN / 4) / 256;  // 8
    float4 v[PER];
#pragma unroll
    for (int j = 0; j < PER; ++j) {
      const int i = t + j * 256, k = i / (T2_MAXN / 4), n = (i % (T2_MAXN / 4)) * 4;
      v[j] = (k < K1 && n < N2) ? __ldg(reinterpret_cast<const float4*>(W2 + (int64_t)k * N2 + n))
                                : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int j = 0; j < PER; ++j) {
      const int i = t + j * 256, k = i / (T2_MAXN / 4), n = (i % (T2_MAXN / 4)) * 4;
      *reinterpret_cast<float4*>(&s.w2s[k][n]) = v[j];
      s.w2t[n][k] = v[j].x; s.w2t[n + 1][k] = v[j].y; s.w2t[n + 2][k] = v[j].z; s.w2t[n + 3][k] = v[j].w;
    }
  }
  {
    float v[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {  // W3 is [N2][N3] with any N3: scalar loads, consecutive threads along a row
      const int i = t + j * 256, f = i >> 4, c = i & 15;
      v[j] = (f < N2 && c < N3) ? __ldg(W3 + (int64_t)f * N3 + c) : 0.f;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int i = t + j * 256, f = i >> 4, c = i & 15;
      s.w3t[c][f] = v[j];
    }
  }
  if (t < T2_MAXN) s.b2s[t] = t < N2 ? b2[t] : 0.f;
  if (t < 16) s.b3s[t] = t < N3 ? b3[t] : 0.f;

  // persistent accumulators of this CTA's split partials
  float dw2acc[4][8];  // rows k = 4 kq .. +3, cols n = 8 no .. +7
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) dw2acc[i][j] = 0.f;
  float dw3acc[4] = {0.f, 0.f, 0.f, 0.f};  // feature f = t & 63, classes 4 (t >> 6) .. +3
  float db2acc = 0.f, db3acc = 0.f;        // threads 0..63 / 0..15

  const int rp = t >> 4, cq = t & 15;   // forward / dZ2 mapping: rows 2rp, 2rp+1; cols 4cq..4cq+3
  const int ko = t & 15;                // dZ1 mapping: rows 2rp, 2rp+1; k = 8ko..8ko+7
  const int kq = t >> 3, no = t & 7;    // dW2 mapping: k = 4kq..4kq+3; n = 8no..8no+7
  const int row3 = t >> 3, sub = t & 7; // layer-3 mapping: 8 lanes per row, features sub + 8j

  for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
    const int64_t r0 = (int64_t)tile * rows_per_tile;
    const int nrows = (int)min((int64_t)rows_per_tile, B - r0);
    __syncthreads();  // weights staged / previous tile fully consumed
    // ---- A1 tile: coalesced 16-byte loads, rows past the end and columns past K1 are zero ----
    {
      float4 v[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int i = t + j * 256, r = i / (T2_MAXK / 4), c4 = (i % (T2_MAXK / 4)) * 4;
        v[j] = (r < nrows && c4 < K1) ? *reinterpret_cast<const float4*>(A1 + (r0 + r) * K1 + c4)
                                      : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int i = t + j * 256, r = i / (T2_MAXK / 4), c4 = (i % (T2_MAXK / 4)) * 4;
        *reinterpret_cast<float4*>(&s.a1s[r][c4]) = v[j];
      }
    }
    __syncthreads();

    // ---- phase 1: A2 = f2(A1.W2 + b2), thread = 2 rows x 4 cols ----
    {
      float acc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
      const float* a0p = s.a1s[2 * rp];
      const float* a1p = s.a1s[2 * rp + 1];
#pragma unroll 2
      for (int k = 0; k < K1; k += 4) {
        const float4 x0 = *reinterpret_cast<const float4*>(a0p + k);
        const float4 x1 = *reinterpret_cast<const float4*>(a1p + k);
        const float xa[4] = {x0.x, x0.y, x0.z, x0.w}, xb[4] = {x1.x, x1.y, x1.z, x1.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float4 w = *reinterpret_cast<const float4*>(&s.w2s[k + j][4 * cq]);
          acc[0][0] = fmaf(xa[j], w.x, acc[0][0]); acc[0][1] = fmaf(xa[j], w.y, acc[0][1]);
          acc[0][2] = fmaf(xa[j], w.z, acc[0][2]); acc[0][3] = fmaf(xa[j], w.w, acc[0][3]);
          acc[1][0] = fmaf(xb[j], w.x, acc[1][0]); acc[1][1] = fmaf(xb[j], w.y, acc[1][1]);
          acc[1][2] = fmaf(xb[j], w.z, acc[1][2]); acc[1][3] = fmaf(xb[j], w.w, acc[1][3]);
        }
      }
      const float4 bb = *reinterpret_cast<const float4*>(&s.b2s[4 * cq]);
#pragma unroll
      for (int h = 0; h < 2; ++h)
        *reinterpret_cast<float4*>(&s.a2s[2 * rp + h][4 * cq]) =
            make_float4(act_fwd(acc[h][0] + bb.x, act2), act_fwd(acc[h][1] + bb.y, act2),
                        act_fwd(acc[h][2] + bb.z, act2), act_fwd(acc[h][3] + bb.w, act2));
    }
    __syncthreads();

    // ---- phase 2: logits, softmax-CE, dZ3 (8 lanes per row) ----
    {
      float z[16];
#pragma unroll
      for (int c = 0; c < 16; ++c) z[c] = 0.f;
#pragma unroll
      for (int j = 0; j < T2_MAXN / 8; ++j) {
        const float a = s.a2s[row3][sub + 8 * j];  // columns >= N2 meet zero rows of W3^T
#pragma unroll
        for (int c = 0; c < 16; ++c) z[c] = fmaf(a, s.w3t[c][sub + 8 * j], z[c]);
      }
      float m = -INFINITY;
#pragma unroll
      for (int c = 0; c < 16; ++c) {
        z[c] += __shfl_xor_sync(0xffffffffu, z[c], 1);
        z[c] += __shfl_xor_sync(0xffffffffu, z[c], 2);
        z[c] += __shfl_xor_sync(0xffffffffu, z[c], 4);
        z[c] += s.b3s[c];
        if (c < N3) m = fmaxf(m, z[c]);
      }
      float sum = 0.f;
#pragma unroll
      for (int c = 0; c < 16; ++c)
        if (c < N3) sum += expf(z[c] - m);
      const bool ok = row3 < nrows;
      const int y = ok ? labels[r0 + row3] : -1;
      if (sub == 0) {
        const float inv = 1.0f / sum;
        float d[16], zy = 0.f;
#pragma unroll
        for (int c = 0; c < 16; ++c) {
          d[c] = (ok && c < N3) ? (expf(z[c] - m) * inv - (c == y ? 1.f : 0.f)) * inv_denom : 0.f;
          if (c == y) zy = z[c];
        }
#pragma unroll
        for (int q = 0; q < 4; ++q)
          *reinterpret_cast<float4*>(&s.dz3s[row3][4 * q]) = make_float4(d[4 * q], d[4 * q + 1], d[4 * q + 2], d[4 * q + 3]);
        if (ok) row_loss[r0 + row3] = (y >= 0 && y < N3) ? -((zy - m) - logf(sum)) : 0.f;
      }
    }
    __syncthreads();

    // ---- phase 3: dZ2 = (dZ3.W3^T) * f2'(A2), thread = 2 rows x 4 cols ----
    {
      float acc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
#pragma unroll
      for (int c4 = 0; c4 < 16; c4 += 4) {
        const float4 d0 = *reinterpret_cast<const float4*>(&s.dz3s[2 * rp][c4]);
        const float4 d1 = *reinterpret_cast<const float4*>(&s.dz3s[2 * rp + 1][c4]);
        const float da[4] = {d0.x, d0.y, d0.z, d0.w}, db[4] = {d1.x, d1.y, d1.z, d1.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float4 w = *reinterpret_cast<const float4*>(&s.w3t[c4 + j][4 * cq]);
          acc[0][0] = fmaf(da[j], w.x, acc[0][0]); acc[0][1] = fmaf(da[j], w.y, acc[0][1]);
          acc[0][2] = fmaf(da[j], w.z, acc[0][2]); acc[0][3] = fmaf(da[j], w.w, acc[0][3]);
          acc[1][0] = fmaf(db[j], w.x, acc[1][0]); acc[1][1] = fmaf(db[j], w.y, acc[1][1]);
          acc[1][2] = fmaf(db[j], w.z, acc[1][2]); acc[1][3] = fmaf(db[j], w.w, acc[1][3]);
        }
      }
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const float4 a = *reinterpret_cast<const float4*>(&s.a2s[2 * rp + h][4 * cq]);
        *reinterpret_cast<float4*>(&s.dz2s[2 * rp + h][4 * cq]) =
            make_float4(acc[h][0] * act_bwd(a.x, act2), acc[h][1] * act_bwd(a.y, act2),
                        acc[h][2] * act_bwd(a.z, act2), acc[h][3] * act_bwd(a.w, act2));
      }
    }
    // dW3 partial (thread = feature x 4 classes) and db3 need only A2 and dZ3, which are complete
    {
      const int f = t & 63, c4 = (t >> 6) * 4;
#pragma unroll 4
      for (int r = 0; r < T2_ROWS; ++r) {
        const float a = s.a2s[r][f];
        const float4 d = *reinterpret_cast<const float4*>(&s.dz3s[r][c4]);
        dw3acc[0] = fmaf(a, d.x, dw3acc[0]); dw3acc[1] = fmaf(a, d.y, dw3acc[1]);
        dw3acc[2] = fmaf(a, d.z, dw3acc[2]); dw3acc[3] = fmaf(a, d.w, dw3acc[3]);
      }
      if (t < 16) {
#pragma unroll 8
        for (int r = 0; r < T2_ROWS; ++r) db3acc += s.dz3s[r][t];
      }
    }
    __syncthreads();

    // ---- phase 4: dZ1 = (dZ2.W2^T) * f1'(A1), thread = 2 rows x 8 k ----
    if (dZ1) {
      float acc[2][8];
#pragma unroll
      for (int h = 0; h < 2; ++h)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[h][j] = 0.f;
#pragma unroll 2
      for (int n = 0; n < N2; n += 4) {
        const float4 d0 = *reinterpret_cast<const float4*>(&s.dz2s[2 * rp][n]);
        const float4 d1 = *reinterpret_cast<const float4*>(&s.dz2s[2 * rp + 1][n]);
        const float da[4] = {d0.x, d0.y, d0.z, d0.w}, db[4] = {d1.x, d1.y, d1.z, d1.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float4 w0 = *reinterpret_cast<const float4*>(&s.w2t[n + j][8 * ko]);
          const float4 w1 = *reinterpret_cast<const float4*>(&s.w2t[n + j][8 * ko + 4]);
          const float w[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            acc[0][e] = fmaf(da[j], w[e], acc[0][e]);
            acc[1][e] = fmaf(db[j], w[e], acc[1][e]);
          }
        }
      }
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int r = 2 * rp + h;
        if (r < nrows && 8 * ko < K1) {  // K1 % 4 == 0: the second half may be out
          const float4 a0 = *reinterpret_cast<const float4*>(&s.a1s[r][8 * ko]);
          const float4 a1 = *reinterpret_cast<const float4*>(&s.a1s[r][8 * ko + 4]);
          float* o = dZ1 + (r0 + r) * K1 + 8 * ko;
          *reinterpret_cast<float4*>(o) =
              make_float4(acc[h][0] * act_bwd(a0.x, act1), acc[h][1] * act_bwd(a0.y, act1),
                          acc[h][2] * act_bwd(a0.z, act1), acc[h][3] * act_bwd(a0.w, act1));
          if (8 * ko + 4 < K1)
            *reinterpret_cast<float4*>(o + 4) =
                make_float4(acc[h][4] * act_bwd(a1.x, act1), acc[h][5] * act_bwd(a1.y, act1),
                            acc[h][6] * act_bwd(a1.z, act1), acc[h][7] * act_bwd(a1.w, act1));
        }
      }
    }
    // ---- phase 5: dW2 += A1^T.dZ2, thread = 4 k x 8 n; db2 ----
#pragma unroll 2
    for (int r = 0; r < T2_ROWS; ++r) {
      const float4 a = *reinterpret_cast<const float4*>(&s.a1s[r][4 * kq]);
      const float4 d0 = *reinterpret_cast<const float4*>(&s.dz2s[r][8 * no]);
      const float4 d1 = *reinterpret_cast<const float4*>(&s.dz2s[r][8 * no + 4]);
      const float av[4] = {a.x, a.y, a.z, a.w};
      const float dv[8] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) dw2acc[i][j] = fmaf(av[i], dv[j], dw2acc[i][j]);
    }
    if (t < T2_MAXN) {
#pragma unroll 8
      for (int r = 0; r < T2_ROWS; ++r) db2acc += s.dz2s[r][t];
    }
  }

  // ---- split partials of this CTA ----
  const int zb = blockIdx.x;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int k = 4 * kq + i;
    if (k < K1 && 8 * no < N2) {
      float* o = dw2_part + ((int64_t)zb * K1 + k) * N2 + 8 * no;
      *reinterpret_cast<float4*>(o) = make_float4(dw2acc[i][0], dw2acc[i][1], dw2acc[i][2], dw2acc[i][3]);
      if (8 * no + 4 < N2)
        *reinterpret_cast<float4*>(o + 4) = make_float4(dw2acc[i][4], dw2acc[i][5], dw2acc[i][6], dw2acc[i][7]);
    }
  }
  if (t < N2) db2_part[(int64_t)zb * N2 + t] = db2acc;
  {
    const int f = t & 63, c4 = (t >> 6) * 4;
    if (f < N2 && c4 < np3)  // np3 % 4 == 0
      *reinterpret_cast<float4*>(dw3_part + ((int64_t)zb * N2 + f) * np3 + c4) =
          make_float4(dw3acc[0], dw3acc[1], dw3acc[2], dw3acc[3]);
  }
  if (t < np3) db3_part[(int64_t)zb * np3 + t] = db3acc;
}

// ======================= tensor-core variant =======================
// The FFMA kernel above issues ~6 200 warp instructions per warp for a 28-row tile (ncu: 7.3 M for 4096 rows,
// 23.5 us) with only 8 warps per SM: issue/latency-bound.  Here the three big products run on
// mma.sync.m16n8k8 (tf32 in, fp32 accumulate) as error-compensated 3xTF32: every operand lives in shared
// memory as a `hi` (tf32-rounded) and a `lo` (remainder) array, split ONCE when it is written (W2 at staging,
// A1 at the tile load, dZ2 in phase 3), so a k-step is a handful of fragment loads and 3 MMAs per output tile,
// correction terms in their own accumulators.  Fragments: ldmatrix on b16 pairs delivers fp32 elements when
// the reduction index is contiguous in shared memory (A of the forward and dZ1 products, W2 as B of the dZ1
// product); otherwise scalar loads with pitches == 8 (mod 32) words, which are conflict-free for the
// (lane/4, lane%4) fragment pattern.  The 10-class layer stays on FFMA (a few hundred instructions).
constexpr int TM_PA1 = T2_MAXK + 4;  // 132: A1 rows, 528 B (ldmatrix: odd multiple of 16 B mod 128)
constexpr int TM_PW = T2_MAXN + 8;   // 72 == 8 (mod 32): W2 [k1][n2] and dZ2 [row][n2]

struct Tail2MmaSmem {
  float a1h[T2_ROWS][TM_PA1], a1l[T2_ROWS][TM_PA1];
  float w2h[T2_MAXK][TM_PW], w2l[T2_MAXK][TM_PW];
  float dz2h[T2_ROWS][TM_PW], dz2l[T2_ROWS][TM_PW];
  float a2s[T2_ROWS][T2_A2P];
  float w3t[16][T2_MAXN];
  float dz3s[T2_ROWS][16];
  float b2s[T2_MAXN];
  float b3s[16];
};

__device__ __forceinline__ uint32_t tm_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ float tm_rna(float x) {  // tf32 rounding on the bit pattern (finite inputs)
  return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u);
}
// four 8x8 b16 matrices = four (8 rows x 4 fp32) blocks: thread (g = lane/4, t = lane%4) gets element (g, t) of each
__device__ __forceinline__ void tm_ldsm4(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr) : "memory");
}
// D += A[16x8] . B[8x8]: a = {(g,t), (g+8,t), (g,t+4), (g+8,t+4)}, b = {(k=t,n=g), (k=t+4,n=g)},
// d = {(g,2t), (g,2t+1), (g+8,2t), (g+8,2t+1)}
__device__ __forceinline__ void tm_mma(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, "
               "{%0, %1, %2, %3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// 3xTF32: corr += Al.Bh + Ah.Bl ; main += Ah.Bh
__device__ __forceinline__ void tm_mma3(float (&main)[4], float (&corr)[4], const uint32_t (&ah)[4],
                                        const uint32_t (&al)[4], uint32_t bh0, uint32_t bh1, uint32_t bl0,
                                        uint32_t bl1) {
  tm_mma(corr, al, bh0, bh1);
  tm_mma(corr, ah, bl0, bl1);
  tm_mma(main, ah, bh0, bh1);
}

__global__ void __launch_bounds__(256)
tail2_mma_kernel(const float* __restrict__ A1, const float* __restrict__ W2, const float* __restrict__ b2,
                 const float* __restrict__ W3, const float* __restrict__ b3, const int32_t* __restrict__ labels,
                 float* __restrict__ dZ1, float* __restrict__ row_loss, float* __restrict__ dw2_part,
                 float* __restrict__ db2_part, float* __restrict__ dw3_part, float* __restrict__ db3_part, int64_t B,
                 int K1, int N2, int N3, int np3, int act1, int act2, float inv_denom, int rows_per_tile,
                 int num_tiles) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  Tail2MmaSmem& s = *reinterpret_cast<Tail2MmaSmem*>(smem_raw);
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int g = lane >> 2, tq = lane & 3;
  pdl_launch_dependents();
  pdl_wait();

  // ---- stage W2 as hi / lo (zero padded to 128 x 64), W3^T, biases ----
  {
    constexpr int PER = T2_MAXK * (T2_MAXN / 4) / 256;  // 8
    float4 v[PER];
#pragma unroll
    for (int j = 0; j < PER; ++j) {
      const int i = t + j * 256, k = i / (T2_MAXN / 4), n = (i % (T2_MAXN / 4)) * 4;
      v[j] = (k < K1 && n < N2) ? __ldg(reinterpret_cast<const float4*>(W2 + (int64_t)k * N2 + n))
                                : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int j = 0; j < PER; ++j) {
      const int i = t + j * 256, k = i / (T2_MAXN / 4), n = (i % (T2_MAXN / 4)) * 4;
      const float4 h = make_float4(tm_rna(v[j].x), tm_rna(v[j].y), tm_rna(v[j].z), tm_rna(v[j].w));
      *reinterpret_cast<float4*>(&s.w2h[k][n]) = h;
      *reinterpret_cast<float4*>(&s.w2l[k][n]) = make_float4(v[j].x - h.x, v[j].y - h.y, v[j].z - h.z, v[j].w - h.w);
    }
  }
  {
    float v[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int i = t + j * 256, f = i >> 4, c = i & 15;
      v[j] = (f < N2 && c < N3) ? __ldg(W3 + (int64_t)f * N3 + c) : 0.f;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int i = t + j * 256, f = i >> 4, c = i & 15;
      s.w3t[c][f] = v[j];
    }
  }
  if (t < T2_MAXN) s.b2s[t] = t < N2 ? b2[t] : 0.f;
  if (t < 16) s.b3s[t] = t < N3 ? b3[t] : 0.f;

  // persistent accumulators: dW2 rows 16 warp + {g, g+8}, all 8 column tiles (main + correction); dW3; db
  float dwm[8][4], dwc[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) dwm[i][j] = dwc[i][j] = 0.f;
  float dw3acc[4] = {0.f, 0.f, 0.f, 0.f};
  float db2acc = 0.f, db3acc = 0.f;

  const int rp = t >> 4, cq = t & 15;    // phase 3 mapping: rows 2rp, 2rp+1; cols 4cq..4cq+3
  const int row3 = t >> 3, sub = t & 7;  // layer-3 mapping: 8 lanes per row
  const int k8 = (K1 + 7) >> 3, n8 = (N2 + 7) >> 3;  // k-steps of the products (zero padded operands)
  // ldmatrix row address pattern of an x4 load of a 16 x 8 fp32 block: matrix j = lane/8 -> rows (j&1)*8, cols (j>>1)*4
  const int lm_row = (lane & 7) + ((lane >> 3) & 1) * 8, lm_col = (lane >> 4) * 4;

  for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
    const int64_t r0 = (int64_t)tile * rows_per_tile;
    const int nrows = (int)min((int64_t)rows_per_tile, B - r0);
    __syncthreads();  // weights staged / previous tile fully consumed
    {
      float4 v[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int i = t + j * 256, r = i / (T2_MAXK / 4), c4 = (i % (T2_MAXK / 4)) * 4;
        v[j] = (r < nrows && c4 < K1) ? *reinterpret_cast<const float4*>(A1 + (r0 + r) * K1 + c4)
                                      : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int i = t + j * 256, r = i / (T2_MAXK / 4), c4 = (i % (T2_MAXK / 4)) * 4;
        const float4 h = make_float4(tm_rna(v[j].x), tm_rna(v[j].y), tm_rna(v[j].z), tm_rna(v[j].w));
        *reinterpret_cast<float4*>(&s.a1h[r][c4]) = h;
        *reinterpret_cast<float4*>(&s.a1l[r][c4]) = make_float4(v[j].x - h.x, v[j].y - h.y, v[j].z - h.z, v[j].w - h.w);
      }
    }
    __syncthreads();

    // ---- phase 1: A2 = f2(A1.W2 + b2); warp = column tile 8 warp .. +7, both 16-row tiles ----
    if (8 * warp < N2) {
      float cm[2][4], cc[2][4];
#pragma unroll
      for (int m = 0; m < 2; ++m)
#pragma unroll
        for (int e = 0; e < 4; ++e) cm[m][e] = cc[m][e] = 0.f;
      const uint32_t ah_addr = tm_u32(&s.a1h[lm_row][lm_col]), al_addr = tm_u32(&s.a1l[lm_row][lm_col]);
      const float* bh = &s.w2h[tq][8 * warp + g];
      const float* bl = &s.w2l[tq][8 * warp + g];
#pragma unroll 2
      for (int ks = 0; ks < k8; ++ks) {
        const uint32_t bh0 = __float_as_uint(bh[(8 * ks) * TM_PW]), bh1 = __float_as_uint(bh[(8 * ks + 4) * TM_PW]);
        const uint32_t bl0 = __float_as_uint(bl[(8 * ks) * TM_PW]), bl1 = __float_as_uint(bl[(8 * ks + 4) * TM_PW]);
#pragma unroll
        for (int m = 0; m < 2; ++m) {
          uint32_t ah[4], al[4];
          tm_ldsm4(ah_addr + (m * 16 * TM_PA1 + ks * 8) * 4, ah);
          tm_ldsm4(al_addr + (m * 16 * TM_PA1 + ks * 8) * 4, al);
          tm_mma3(cm[m], cc[m], ah, al, bh0, bh1, bl0, bl1);
        }
      }
      const int col = 8 * warp + 2 * tq;
      const float bb0 = s.b2s[col], bb1 = s.b2s[col + 1];
#pragma unroll
      for (int m = 0; m < 2; ++m) {
        *reinterpret_cast<float2*>(&s.a2s[16 * m + g][col]) =
            make_float2(act_fwd(cm[m][0] + cc[m][0] + bb0, act2), act_fwd(cm[m][1] + cc[m][1] + bb1, act2));
        *reinterpret_cast<float2*>(&s.a2s[16 * m + g + 8][col]) =
            make_float2(act_fwd(cm[m][2] + cc[m][2] + bb0, act2), act_fwd(cm[m][3] + cc[m][3] + bb1, act2));
      }
    } else {
      // column tiles past N2: A2 = 0 there (they meet zero rows of W3^T and are never written out)
      const int col = 8 * warp + 2 * tq;
#pragma unroll
      for (int m = 0; m < 2; ++m) {
        *reinterpret_cast<float2*>(&s.a2s[16 * m + g][col]) = make_float2(0.f, 0.f);
        *reinterpret_cast<float2*>(&s.a2s[16 * m + g + 8][col]) = make_float2(0.f, 0.f);
      }
    }
    __syncthreads();

    // ---- phase 2: logits, softmax-CE, dZ3 (8 lanes per row) ----
    {
      float z[16];
#pragma unroll
      for (int c = 0; c < 16; ++c) z[c] = 0.f;
#pragma unroll
      for (int j = 0; j < T2_MAXN / 8; ++j) {
        const float a = s.a2s[row3][sub + 8 * j];  // columns >= N2 meet zero rows of W3^T
#pragma unroll
        for (int c = 0; c < 16; ++c) z[c] = fmaf(a, s.w3t[c][sub + 8 * j], z[c]);
      }
      float m = -INFINITY;
#pragma unroll
      for (int c = 0; c < 16; ++c) {
        z[c] += __shfl_xor_sync(0xffffffffu, z[c], 1);
        z[c] += __shfl_xor_sync(0xffffffffu, z[c], 2);
        z[c] += __shfl_xor_sync(0xffffffffu, z[c], 4);
        z[c] += s.b3s[c];
        if (c < N3) m = fmaxf(m, z[c]);
      }
      float sum = 0.f;
#pragma unroll
      for (int c = 0; c < 16; ++c)
        if (c < N3) sum += expf(z[c] - m);
      const bool ok = row3 < nrows;
      const int y = ok ? labels[r0 + row3] : -1;
      if (sub == 0) {
        const float inv = 1.0f / sum;
        float d[16], zy = 0.f;
#pragma unroll
        for (int c = 0; c < 16; ++c) {
          d[c] = (ok && c < N3) ? (expf(z[c] - m) * inv - (c == y ? 1.f : 0.f)) * inv_denom : 0.f;
          if (c == y) zy = z[c];
        }
#pragma unroll
        for (int q = 0; q < 4; ++q)
          *reinterpret_cast<float4*>(&s.dz3s[row3][4 * q]) = make_float4(d[4 * q], d[4 * q + 1], d[4 * q + 2], d[4 * q + 3]);
        if (ok) row_loss[r0 + row3] = (y >= 0 && y < N3) ? -((zy - m) - logf(sum)) : 0.f;
      }
    }
    __syncthreads();


    // ---- phase 3: dZ2 = (dZ3.W3^T) * f2'(A2) (FFMA, thread = 2 rows x 4 cols), stored as hi / lo ----
    {
      float acc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
#pragma unroll
      for (int c4 = 0; c4 < 16; c4 += 4) {
        const float4 d0 = *reinterpret_cast<const float4*>(&s.dz3s[2 * rp][c4]);
        const float4 d1 = *reinterpret_cast<const float4*>(&s.dz3s[2 * rp + 1][c4]);
        const float da[4] = {d0.x, d0.y, d0.z, d0.w}, db[4] = {d1.x, d1.y, d1.z, d1.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float4 w = *reinterpret_cast<const float4*>(&s.w3t[c4 + j][4 * cq]);
          acc[0][0] = fmaf(da[j], w.x, acc[0][0]); acc[0][1] = fmaf(da[j], w.y, acc[0][1]);
          acc[0][2] = fmaf(da[j], w.z, acc[0][2]); acc[0][3] = fmaf(da[j], w.w, acc[0][3]);
          acc[1][0] = fmaf(db[j], w.x, acc[1][0]); acc[1][1] = fmaf(db[j], w.y, acc[1][1]);
          acc[1][2] = fmaf(db[j], w.z, acc[1][2]); acc[1][3] = fmaf(db[j], w.w, acc[1][3]);
        }
      }
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const float4 a = *reinterpret_cast<const float4*>(&s.a2s[2 * rp + h][4 * cq]);
        const float4 d = make_float4(acc[h][0] * act_bwd(a.x, act2), acc[h][1] * act_bwd(a.y, act2),
                                     acc[h][2] * act_bwd(a.z, act2), acc[h][3] * act_bwd(a.w, act2));
        const float4 hi = make_float4(tm_rna(d.x), tm_rna(d.y), tm_rna(d.z), tm_rna(d.w));
        *reinterpret_cast<float4*>(&s.dz2h[2 * rp + h][4 * cq]) = hi;
        *reinterpret_cast<float4*>(&s.dz2l[2 * rp + h][4 * cq]) = make_float4(d.x - hi.x, d.y - hi.y, d.z - hi.z, d.w - hi.w);
      }
    }
    // dW3 partial (thread = feature x 4 classes) and db3 need only A2 and dZ3, which are complete
    {
      const int f = t & 63, c4 = (t >> 6) * 4;
#pragma unroll 4
      for (int r = 0; r < T2_ROWS; ++r) {
        const float a = s.a2s[r][f];
        const float4 d = *reinterpret_cast<const float4*>(&s.dz3s[r][c4]);
        dw3acc[0] = fmaf(a, d.x, dw3acc[0]); dw3acc[1] = fmaf(a, d.y, dw3acc[1]);
        dw3acc[2] = fmaf(a, d.z, dw3acc[2]); dw3acc[3] = fmaf(a, d.w, dw3acc[3]);
      }
      if (t < 16) {
#pragma unroll 8
        for (int r = 0; r < T2_ROWS; ++r) db3acc += s.dz3s[r][t];
      }
    }
    __syncthreads();

    // ---- phase 4: dZ1 = (dZ2.W2^T) * f1'(A1); warp = column tiles 2 warp, 2 warp + 1 (k1), both row tiles ----
    if (dZ1 && 16 * warp < K1) {
      float cm[2][2][4], cc[2][2][4];
#pragma unroll
      for (int m = 0; m < 2; ++m)
#pragma unroll
        for (int n = 0; n < 2; ++n)
#pragma unroll
          for (int e = 0; e < 4; ++e) cm[m][n][e] = cc[m][n][e] = 0.f;
      const uint32_t ah_addr = tm_u32(&s.dz2h[lm_row][lm_col]), al_addr = tm_u32(&s.dz2l[lm_row][lm_col]);
      // B(k = n2, n = k1) = W2[k1][n2]: n2 contiguous -> ldmatrix; matrix j: column tile j/2, k half j%2
      const int brow = 16 * warp + (lane >> 4) * 8 + (lane & 7), bcol = ((lane >> 3) & 1) * 4;
      const uint32_t bh_addr = tm_u32(&s.w2h[brow][bcol]), bl_addr = tm_u32(&s.w2l[brow][bcol]);
#pragma unroll 2
      for (int ks = 0; ks < n8; ++ks) {
        uint32_t bh[4], bl[4];
        tm_ldsm4(bh_addr + ks * 32, bh);
        tm_ldsm4(bl_addr + ks * 32, bl);
#pragma unroll
        for (int m = 0; m < 2; ++m) {
          uint32_t ah[4], al[4];
          tm_ldsm4(ah_addr + (m * 16 * TM_PW + ks * 8) * 4, ah);
          tm_ldsm4(al_addr + (m * 16 * TM_PW + ks * 8) * 4, al);
          tm_mma3(cm[m][0], cc[m][0], ah, al, bh[0], bh[1], bl[0], bl[1]);
          tm_mma3(cm[m][1], cc[m][1], ah, al, bh[2], bh[3], bl[2], bl[3]);
        }
      }
#pragma unroll
      for (int m = 0; m < 2; ++m)
#pragma unroll
        for (int n = 0; n < 2; ++n) {
          const int k1 = 16 * warp + 8 * n + 2 * tq;
          if (k1 >= K1) continue;  // K1 % 4 == 0 and k1 is even: k1 + 1 < K1 as well
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const int r = 16 * m + g + 8 * h;
            if (r >= nrows) continue;
            const float x0 = s.a1h[r][k1] + s.a1l[r][k1], x1 = s.a1h[r][k1 + 1] + s.a1l[r][k1 + 1];
            *reinterpret_cast<float2*>(dZ1 + (r0 + r) * K1 + k1) =
                make_float2((cm[m][n][2 * h] + cc[m][n][2 * h]) * act_bwd(x0, act1),
                            (cm[m][n][2 * h + 1] + cc[m][n][2 * h + 1]) * act_bwd(x1, act1));
          }
        }
    }
    // ---- phase 5: dW2 += A1^T.dZ2; warp = rows k1 = 16 warp .. +15, all column tiles; reduction over the tile rows ----
    if (16 * warp < K1) {
#pragma unroll 1
      for (int ks = 0; ks < T2_ROWS / 8; ++ks) {
        // A(m = k1, k = row) = A1[row][k1]: scalar loads
        uint32_t ah[4], al[4];
        const int ra = 8 * ks + tq, ca = 16 * warp + g;
        ah[0] = __float_as_uint(s.a1h[ra][ca]);     ah[1] = __float_as_uint(s.a1h[ra][ca + 8]);
        ah[2] = __float_as_uint(s.a1h[ra + 4][ca]); ah[3] = __float_as_uint(s.a1h[ra + 4][ca + 8]);
        al[0] = __float_as_uint(s.a1l[ra][ca]);     al[1] = __float_as_uint(s.a1l[ra][ca + 8]);
        al[2] = __float_as_uint(s.a1l[ra + 4][ca]); al[3] = __float_as_uint(s.a1l[ra + 4][ca + 8]);
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
          if (nt < n8) {
            // B(k = row, n = n2) = dZ2[row][n2]: scalar loads, pitch 72 == 8 (mod 32): conflict-free
            const uint32_t bh0 = __float_as_uint(s.dz2h[ra][8 * nt + g]), bh1 = __float_as_uint(s.dz2h[ra + 4][8 * nt + g]);
            const uint32_t bl0 = __float_as_uint(s.dz2l[ra][8 * nt + g]), bl1 = __float_as_uint(s.dz2l[ra + 4][8 * nt + g]);
            tm_mma3(dwm[nt], dwc[nt], ah, al, bh0, bh1, bl0, bl1);
          }
        }
      }
    }
    if (t < T2_MAXN) {
#pragma unroll 8
      for (int r = 0; r < T2_ROWS; ++r) db2acc += s.dz2h[r][t] + s.dz2l[r][t];
    }
  }

  // ---- split partials of this CTA ----
  const int zb = blockIdx.x;
  if (16 * warp < K1) {
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      const int n = 8 * nt + 2 * tq;
      if (n >= N2) continue;  // N2 % 4 == 0 and n is even
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int k1 = 16 * warp + g + 8 * h;
        if (k1 < K1)
          *reinterpret_cast<float2*>(dw2_part + ((int64_t)zb * K1 + k1) * N2 + n) =
              make_float2(dwm[nt][2 * h] + dwc[nt][2 * h], dwm[nt][2 * h + 1] + dwc[nt][2 * h + 1]);
      }
    }
  }
  if (t < N2) db2_part[(int64_t)zb * N2 + t] = db2acc;
  {
    const int f = t & 63, c4 = (t >> 6) * 4;
    if (f < N2 && c4 < np3)
      *reinterpret_cast<float4*>(dw3_part + ((int64_t)zb * N2 + f) * np3 + c4) =
          make_float4(dw3acc[0], dw3acc[1], dw3acc[2], dw3acc[3]);
  }
  if (t < np3) db3_part[(int64_t)zb * np3 + t] = db3acc;
}

int sm_count() {
  int dev = 0, n = 0;
  if (cudaGetDevice(&dev) != cudaSuccess ||
      cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
    return 148;
  return n;
}

struct Plan {
  int rows_per_tile, tiles, grid;
};
Plan plan(int64_t B) {
  const int sms = sm_count();
  long rpt = (B + sms - 1) / sms;        // one tile per SM when the batch allows it
  rpt = (rpt + 3) / 4 * 4;
  if (rpt < 4) rpt = 4;
  if (rpt > T2_ROWS) rpt = T2_ROWS;
  Plan p;
  p.rows_per_tile = (int)rpt;
  p.tiles = (int)((B + rpt - 1) / rpt);
  p.grid = p.tiles < sms ? p.tiles : sms;
  return p;
}

}  // namespace

bool tail2_ok(int64_t K1, int64_t N2, int64_t N3, int dtype) {
  return dtype == DDNN_F32 && K1 >= 4 && K1 <= T2_MAXK && K1 % 4 == 0 && N2 >= 4 && N2 <= T2_MAXN && N2 % 4 == 0 &&
         N3 >= 1 && N3 <= 16;
}
int tail2_splits(int64_t B) { return plan(B).grid; }
// Bound on tail2_splits(B') for every B' <= B (workspace sizing): the grid is not monotone in the batch (147 CTAs
// at 4096 rows, 148 at 3540), but it never exceeds the SM count nor the number of 4-row tiles.
int tail2_max_splits(int64_t B) {
  const int64_t t = (B + 3) / 4;
  const int sms = sm_count();
  return (int)(t < sms ? t : sms);
}

int tail2_launch(const float* A1, const float* W2, const float* b2, const float* W3, const float* b3,
                 const int32_t* labels, float* dZ1, float* row_loss, float* dw2_part, float* db2_part, float* dw3_part,
                 float* db3_part, int64_t B, int64_t K1, int64_t N2, int64_t N3, int act1, int act2, float inv_denom,
                 int tensor_cores, cudaStream_t st) {
  const Plan p = plan(B);
  const int np3 = (int)((N3 + 3) & ~(int64_t)3);
  if (tensor_cores) {
    if (int rc = ensure_dynamic_smem((const void*)tail2_mma_kernel, sizeof(Tail2MmaSmem))) return rc;
    DDNN_CUDA_CHECK(launch_pdl(tail2_mma_kernel, dim3(p.grid), dim3(256), sizeof(Tail2MmaSmem), st, A1, W2, b2, W3, b3,
                               labels, dZ1, row_loss, dw2_part, db2_part, dw3_part, db3_part, B, (int)K1, (int)N2,
                               (int)N3, np3, act1, act2, inv_denom, p.rows_per_tile, p.tiles));
    count_launch();
    return DDNN_OK;
  }
  if (int rc = ensure_dynamic_smem((const void*)tail2_kernel, sizeof(Tail2Smem))) return rc;
  DDNN_CUDA_CHECK(launch_pdl(tail2_kernel, dim3(p.grid), dim3(256), sizeof(Tail2Smem), st, A1, W2, b2, W3, b3, labels,
                             dZ1, row_loss, dw2_part, db2_part, dw3_part, db3_part, B, (int)K1, (int)N2, (int)N3, np3,
                             act1, act2, inv_denom, p.rows_per_tile, p.tiles));
  count_launch();
  return DDNN_OK;
}

}  // namespace ddnn
