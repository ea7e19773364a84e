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
#include <atomic>

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
  pdl_launch_dependents();
  pdl_wait();  // A comes from the forward GEMM before us, W / bias from the previous step's reducer
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

// ======================= bf16 tensor-core variant =======================
// The FFMA kernel above is instruction-bound (ncu: 53 M warp instructions, 108 us for 65 536 x 256 -> 10
// against 11 us of HBM time).  For bf16 operands the three products of the step are small MMAs:
//   logits  Z[16 x 16]   = A[16 x K] . W[K x 16]          (per warp: its 16 rows, k = features)
//   dW      [K x 16]    += A^T[K x 128] . dZ[128 x 16]    (per warp: 32 features, k = the tile's 128 rows)
//   dZprev  [16 x K]     = dZ[16 x 16] . W^T[16 x K]      (per warp: its 16 rows, k = classes)
// on mma.sync.m16n8k16 (bf16 in, fp32 accumulate; N padded to 16 with zero columns).  The softmax runs in
// the accumulator fragment layout (a row's 16 logits live in 4 lanes), and that layout IS the A-operand
// layout of the dZprev product, so dZ never leaves registers for it.  A 128-row tile of A is staged once
// in shared memory with cp.async (double-buffered: tile i+1 streams in during tile i), read by ldmatrix for
// all three products, overwritten in place with dZprev = (dZ.W^T) * f'(A) and stored with 16-byte rows.
constexpr int MT_ROWS = 128;   // 8 warps x 16 rows
constexpr int WK_PITCH = 24;   // bf16 per staged W / dZ row: 48 B, an odd multiple of 16 B -> conflict-free ldmatrix

__device__ __forceinline__ uint32_t sm_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr) : "memory");
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr) : "memory");
}
// D (+)= A[16 x 16] . B[16 x 8], bf16 operands, fp32 accumulators in the m16n8 fragment layout:
// d[0], d[1] = (row g, cols 2t, 2t+1), d[2], d[3] = (row g+8, same cols) with g = lane/4, t = lane%4
__device__ __forceinline__ void mma_bf16(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                         uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, "
               "{%0, %1, %2, %3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, int src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  const __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&h);
}

template <int ACT>
__global__ void __launch_bounds__(256, 2)
outlayer_mma_kernel(const bf16* __restrict__ A, const bf16* __restrict__ W, const float* __restrict__ bias,
                    const int32_t* __restrict__ labels, bf16* __restrict__ dZ, float* __restrict__ row_loss,
                    bf16* __restrict__ dZprev, float* __restrict__ dw_part, float* __restrict__ db_part, int64_t B,
                    int K, int N, int np, float inv_denom, int num_tiles) {
  extern __shared__ __align__(16) uint8_t sm_raw[];
  const int PA = K + 8;  // bf16 per staged A row: (2K + 16) bytes, an odd multiple of 16 B
  bf16* at = reinterpret_cast<bf16*>(sm_raw);              // [MT_ROWS][PA]   the tile of A, later dZprev in place
  bf16* wk = at + MT_ROWS * PA;                            // [K][WK_PITCH]   W row-major, classes padded
  bf16* dzs = wk + K * WK_PITCH;                           // [MT_ROWS][WK_PITCH]  dZ rounded to bf16
  bf16* dzl = dzs + MT_ROWS * WK_PITCH;                    // [MT_ROWS][WK_PITCH]  the rounding remainder (dW only)
  float* red = reinterpret_cast<float*>(dzl + MT_ROWS * WK_PITCH);  // [8][16]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int cpr = K >> 3;  // 16-byte chunks per row
  pdl_launch_dependents();
  pdl_wait();  // A comes from the forward GEMM before us, W / bias from the previous step's reducer

  const int wr = warp * 16;  // this warp's first row inside a tile
  // Every warp stages ITS 16 rows (lane = 16-byte chunk of a row, K <= 256): phases 1 and 3 touch only those,
  // so the next tile's rows can stream in as soon as the warp has stored its part of this one.  Two CTAs per
  // SM cover each other's load latency.
  auto load_rows = [&](int tile) {
    const int64_t r0 = (int64_t)tile * MT_ROWS + wr;
    if (lane < cpr) {
#pragma unroll
      for (int r = 0; r < 16; ++r) {
        const bool ok = r0 + r < B;  // rows past the end are zero-filled (src-size 0)
        cp_async16(sm_u32(at + (wr + r) * PA + lane * 8), A + (ok ? (r0 + r) * K + lane * 8 : 0), ok ? 16 : 0);
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };

  int tile = blockIdx.x;
  if (tile < num_tiles) load_rows(tile);
  for (int i = tid; i < K * WK_PITCH; i += 256) {
    const int k = i / WK_PITCH, n = i - k * WK_PITCH;
    wk[i] = n < N ? W[(int64_t)k * N + n] : __float2bfloat16_rn(0.f);
  }
  float bv[2][2];
  bool cv[2][2];
#pragma unroll
  for (int nt = 0; nt < 2; ++nt)
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int c = nt * 8 + 2 * t + e;
      cv[nt][e] = c < N;
      bv[nt][e] = cv[nt][e] ? __ldg(bias + c) : 0.f;
    }
  float dwacc[2][2][4];
#pragma unroll
  for (int a = 0; a < 2; ++a)
#pragma unroll
    for (int b = 0; b < 2; ++b)
#pragma unroll
      for (int e = 0; e < 4; ++e) dwacc[a][b][e] = 0.f;
  float dbacc[2][2] = {{0.f, 0.f}, {0.f, 0.f}};
  // per-lane ldmatrix row addresses: matrix j = lane/8 of an x4 load, row i = lane%8
  const int li = lane & 7, j1 = (lane >> 3) & 1, j2 = lane >> 4;

  __syncthreads();  // W is staged
  for (; tile < num_tiles; tile += gridDim.x) {
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncwarp();  // this warp's 16 rows have landed
    const int64_t r0 = (int64_t)tile * MT_ROWS;

    // ---------------- phase 1: logits of 16 rows, softmax-CE in the accumulator layout ----------------
    float z[2][4];
#pragma unroll
    for (int nt = 0; nt < 2; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) z[nt][e] = 0.f;
    {
      // A operand (m = row, k = feature): matrices (rows 0-7 | 8-15) x (k 0-7 | 8-15)
      const uint32_t a_addr = sm_u32(at + (wr + li + j1 * 8) * PA + j2 * 8);
      // B operand (k = feature, n = class) from W[k][n]: transposed load, matrices (k 0-7 | 8-15) x (n 0-7 | 8-15)
      const uint32_t w_addr = sm_u32(wk + (li + j1 * 8) * WK_PITCH + j2 * 8);
#pragma unroll 4
      for (int ks = 0; ks < (K >> 4); ++ks) {
        uint32_t a0, a1, a2, a3, b0, b1, b2, b3;
        ldsm_x4(a_addr + ks * 32, a0, a1, a2, a3);
        ldsm_x4_t(w_addr + ks * (16 * WK_PITCH * 2), b0, b1, b2, b3);
        mma_bf16(z[0], a0, a1, a2, a3, b0, b1);
        mma_bf16(z[1], a0, a1, a2, a3, b2, b3);
      }
    }
    // dZ of this warp's 16 rows as the A operand (m = row, k = class) of the dZprev product: bf16 value and
    // its rounding remainder (the gradients keep fp32-level accuracy at the cost of a second, nearly free MMA)
    uint32_t dfrag[4], lfrag[4];
    {
      float m[2] = {-INFINITY, -INFINITY}, s[2] = {0.f, 0.f}, zy[2] = {0.f, 0.f};
      const int64_t row[2] = {r0 + wr + g, r0 + wr + g + 8};
      int y[2];
#pragma unroll
      for (int h = 0; h < 2; ++h) y[h] = row[h] < B ? __ldg(labels + row[h]) : -1;
#pragma unroll
      for (int nt = 0; nt < 2; ++nt)
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          z[nt][e] = cv[nt][e & 1] ? z[nt][e] + bv[nt][e & 1] : -INFINITY;
          m[e >> 1] = fmaxf(m[e >> 1], z[nt][e]);
        }
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        m[h] = fmaxf(m[h], __shfl_xor_sync(0xffffffffu, m[h], 1));
        m[h] = fmaxf(m[h], __shfl_xor_sync(0xffffffffu, m[h], 2));
      }
      float ex[2][4];
#pragma unroll
      for (int nt = 0; nt < 2; ++nt)
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          ex[nt][e] = __expf(z[nt][e] - m[e >> 1]);  // arguments <= 0; exp(-inf) = 0 for the padded classes
          s[e >> 1] += ex[nt][e];
          if (nt * 8 + 2 * t + (e & 1) == y[e >> 1]) zy[e >> 1] = z[nt][e];
        }
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        s[h] += __shfl_xor_sync(0xffffffffu, s[h], 1);
        s[h] += __shfl_xor_sync(0xffffffffu, s[h], 2);
        zy[h] += __shfl_xor_sync(0xffffffffu, zy[h], 1);
        zy[h] += __shfl_xor_sync(0xffffffffu, zy[h], 2);
        if (t == 0 && row[h] < B) row_loss[row[h]] = (y[h] >= 0 && y[h] < N) ? -((zy[h] - m[h]) - logf(s[h])) : 0.f;
      }
      float d[2][4];
      const float inv[2] = {1.0f / s[0], 1.0f / s[1]};
#pragma unroll
      for (int nt = 0; nt < 2; ++nt)
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int h = e >> 1, c = nt * 8 + 2 * t + (e & 1);
          d[nt][e] = row[h] < B ? (ex[nt][e] * inv[h] - (c == y[h] ? 1.f : 0.f)) * inv_denom : 0.f;
          if (!cv[nt][e & 1]) d[nt][e] = 0.f;
          if (dZ && cv[nt][e & 1] && row[h] < B) dZ[row[h] * N + c] = __float2bfloat16_rn(d[nt][e]);
        }
#pragma unroll
      for (int nt = 0; nt < 2; ++nt) {
        dbacc[nt][0] += d[nt][0] + d[nt][2];
        dbacc[nt][1] += d[nt][1] + d[nt][3];
      }
      dfrag[0] = pack_bf16(d[0][0], d[0][1]);  // row g,   classes 2t, 2t+1
      dfrag[1] = pack_bf16(d[0][2], d[0][3]);  // row g+8
      dfrag[2] = pack_bf16(d[1][0], d[1][1]);  // row g,   classes 8+2t, 9+2t
      dfrag[3] = pack_bf16(d[1][2], d[1][3]);  // row g+8
      // dW is an fp32 gradient (tolerance 1e-5): feed it dZ as bf16 hi + bf16 lo (two MMAs, error 2^-17)
#pragma unroll
      for (int nt = 0; nt < 2; ++nt)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const __nv_bfloat162 hi = *reinterpret_cast<const __nv_bfloat162*>(&dfrag[nt * 2 + h]);
          const float2 hf = __bfloat1622float2(hi);
          lfrag[nt * 2 + h] = pack_bf16(d[nt][2 * h] - hf.x, d[nt][2 * h + 1] - hf.y);
        }
      uint32_t* dz32 = reinterpret_cast<uint32_t*>(dzs);
      uint32_t* dl32 = reinterpret_cast<uint32_t*>(dzl);
#pragma unroll
      for (int nt = 0; nt < 2; ++nt)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int o = ((wr + g + 8 * h) * WK_PITCH + 8 * nt + 2 * t) >> 1;
          dz32[o] = dfrag[nt * 2 + h];
          dl32[o] = lfrag[nt * 2 + h];
        }
    }
    __syncthreads();  // the tile's dZ is complete

    // ---------------- phase 2: dW += A^T . dZ, this warp's 32 features over the tile's 128 rows ----------------
    {
      // B operand (k = row, n = class) from dZ[row][class]: transposed load
      const uint32_t d_addr = sm_u32(dzs + (li + j1 * 8) * WK_PITCH + j2 * 8);
      // A operand = A^T (m = feature, k = row) from A[row][feature]: transposed load,
      // matrices (m 0-7 | 8-15) x (k 0-7 | 8-15) = stored (cols f..f+7 | f+8..) x (rows 0-7 | 8-15)
      const uint32_t at_addr = sm_u32(at + (li + j2 * 8) * PA + warp * 32 + j1 * 8);
      const bool mt_ok[2] = {warp * 32 < K, warp * 32 + 16 < K};
#pragma unroll 2
      for (int ks = 0; ks < MT_ROWS / 16; ++ks) {
        uint32_t b0, b1, b2, b3, l0, l1, l2, l3;
        ldsm_x4_t(d_addr + ks * (16 * WK_PITCH * 2), b0, b1, b2, b3);
        ldsm_x4_t(d_addr + MT_ROWS * WK_PITCH * 2 + ks * (16 * WK_PITCH * 2), l0, l1, l2, l3);  // dzl follows dzs
#pragma unroll
        for (int mi = 0; mi < 2; ++mi) {
          if (!mt_ok[mi]) continue;  // warp-uniform
          uint32_t a0, a1, a2, a3;
          ldsm_x4_t(at_addr + ks * (16 * PA * 2) + mi * 32, a0, a1, a2, a3);
          mma_bf16(dwacc[mi][0], a0, a1, a2, a3, l0, l1);
          mma_bf16(dwacc[mi][1], a0, a1, a2, a3, l2, l3);
          mma_bf16(dwacc[mi][0], a0, a1, a2, a3, b0, b1);
          mma_bf16(dwacc[mi][1], a0, a1, a2, a3, b2, b3);
        }
      }
    }
    __syncthreads();  // every warp has read the whole tile: its rows may now be overwritten
    if (dZprev) {
      // ---------------- phase 3: dZprev = (dZ . W^T) * f'(A) for this warp's 16 rows ----------------
      // B operand (k = class, n = feature) from W[feature][class]: plain load, matrices
      // (n 0-7 | 8-15 of a 16-feature group) x (k 0-7 | 8-15)
      const uint32_t w_addr = sm_u32(wk + (li + j2 * 8) * WK_PITCH + j1 * 8);
      __nv_bfloat162* arow = reinterpret_cast<__nv_bfloat162*>(at + (wr + g) * PA + 2 * t);
      const int half_pitch8 = 8 * PA / 2;  // 8 rows further, in bf16 pairs
#pragma unroll 2
      for (int fg = 0; fg < (K >> 4); ++fg) {
        uint32_t b0, b1, b2, b3;
        ldsm_x4(w_addr + fg * (16 * WK_PITCH * 2), b0, b1, b2, b3);
        float c[2][4];
#pragma unroll
        for (int h = 0; h < 2; ++h)
#pragma unroll
          for (int e = 0; e < 4; ++e) c[h][e] = 0.f;
        mma_bf16(c[0], lfrag[0], lfrag[1], lfrag[2], lfrag[3], b0, b1);  // rounding remainder of dZ first
        mma_bf16(c[1], lfrag[0], lfrag[1], lfrag[2], lfrag[3], b2, b3);
        mma_bf16(c[0], dfrag[0], dfrag[1], dfrag[2], dfrag[3], b0, b1);  // features 16 fg + 0..7
        mma_bf16(c[1], dfrag[0], dfrag[1], dfrag[2], dfrag[3], b2, b3);  // features 16 fg + 8..15
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          __nv_bfloat162* pa = arow + fg * 8 + h * 4;  // (row g, features 16 fg + 8 h + 2t, +1)
          __nv_bfloat162* pb = pa + half_pitch8;       // row g + 8
          const float2 xa = __bfloat1622float2(*pa), xb = __bfloat1622float2(*pb);
          *pa = __floats2bfloat162_rn(c[h][0] * act_bwd(xa.x, ACT), c[h][1] * act_bwd(xa.y, ACT));
          *pb = __floats2bfloat162_rn(c[h][2] * act_bwd(xb.x, ACT), c[h][3] * act_bwd(xb.y, ACT));
        }
      }
      __syncwarp();
      if (lane < cpr) {
        const bf16* src = at + wr * PA + lane * 8;
        bf16* dst = dZprev + (r0 + wr) * K + lane * 8;
        const int nr = (int)min((int64_t)16, B - (r0 + wr));  // warp-uniform, may be <= 0 in the ragged last tile
#pragma unroll
        for (int r4 = 0; r4 < 16; r4 += 4) {
          uint4 v[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) v[u] = *reinterpret_cast<const uint4*>(src + (r4 + u) * PA);
#pragma unroll
          for (int u = 0; u < 4; ++u)
            if (r4 + u < nr) *reinterpret_cast<uint4*>(dst + (int64_t)(r4 + u) * K) = v[u];
        }
      }
    }
    // this warp's rows are free again: stream in its part of the CTA's next tile
    if (tile + (int)gridDim.x < num_tiles) load_rows(tile + gridDim.x);
  }

  // ---------------- split partials of this CTA: dw_part[z][K][np], db_part[z][np] ----------------
  const int zb = blockIdx.x;
#pragma unroll
  for (int mi = 0; mi < 2; ++mi) {
    const int f0 = warp * 32 + mi * 16;
    if (f0 >= K) continue;
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
      const int n = nt * 8 + 2 * t;
      if (n >= np) continue;  // np % 4 == 0 and n is even: n + 1 < np as well
      float* o = dw_part + ((int64_t)zb * K + f0 + g) * np + n;
      *reinterpret_cast<float2*>(o) = make_float2(dwacc[mi][nt][0], dwacc[mi][nt][1]);
      *reinterpret_cast<float2*>(o + (int64_t)8 * np) = make_float2(dwacc[mi][nt][2], dwacc[mi][nt][3]);
    }
  }
#pragma unroll
  for (int nt = 0; nt < 2; ++nt)
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      float v = dbacc[nt][e];
      v += __shfl_xor_sync(0xffffffffu, v, 4);
      v += __shfl_xor_sync(0xffffffffu, v, 8);
      v += __shfl_xor_sync(0xffffffffu, v, 16);
      if (g == 0) red[warp * 16 + nt * 8 + 2 * t + e] = v;
    }
  __syncthreads();
  if (tid < np) {
    float v = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) v += red[w * 16 + tid];
    db_part[(int64_t)zb * np + tid] = v;
  }
}

int sm_count() {
  int dev = 0, n = 0;
  if (cudaGetDevice(&dev) != cudaSuccess ||
      cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
    return 148;
  return n;
}
bool use_mma(int64_t K, int dtype) { return dtype == DDNN_BF16 && K >= 16 && K % 16 == 0; }
size_t mma_smem_bytes(int64_t K) {
  return (size_t)MT_ROWS * (K + 8) * 2 + (size_t)K * WK_PITCH * 2 + (size_t)2 * MT_ROWS * WK_PITCH * 2 + 8 * 16 * 4;
}
int mma_grid(int64_t B) {
  const int64_t tiles = cdiv(B, MT_ROWS);
  const int slots = 2 * sm_count();  // two resident CTAs per SM
  return (int)(tiles < slots ? tiles : slots);
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

static std::atomic<int> g_outlayer_mma{1};
void outlayer_set_tensor_cores(int on) { g_outlayer_mma.store(on ? 1 : 0); }

int outlayer_splits(int64_t B, int64_t K, int dtype) {
  if (g_outlayer_mma.load() && use_mma(K, dtype)) return mma_grid(B);
  return cdiv(B, rows_per_block(B));
}
// Bound on the split count of EVERY batch of at most B rows, over both variants: a caller sizes its workspace
// once for its largest batch and then runs smaller (ragged last) batches through the same buffer, and
// cdiv(B', rows_per_block(B')) is not monotone in B' (512 blocks at 65536 rows, 592 at 37888).  The FFMA variant
// never has more than 4 x 148 blocks nor more blocks than 64-row tiles; the MMA grid grows with B.
int outlayer_max_splits(int64_t B, int64_t K, int dtype) {
  const int64_t tiles = cdiv(B, (int64_t)TILE_ROWS);
  const int a = (int)(tiles < 4 * 148 ? tiles : 4 * 148);
  const int b = use_mma(K, dtype) ? mma_grid(B) : 0;
  return a > b ? a : b;
}

static int outlayer_launch_mma(const void* A, const void* W, const float* bias, const int32_t* labels, void* dZ,
                               float* row_loss, void* dZprev, float* dw_part, float* db_part, int64_t B, int64_t K,
                               int64_t N, int act_prev, float inv_denom, cudaStream_t st) {
  const int np = (int)((N + 3) & ~(int64_t)3);
  const size_t smem = mma_smem_bytes(K);
  const int tiles = (int)cdiv(B, MT_ROWS);
#define DDNN_OUT_MMA(ACT)                                                                                          \
  do {                                                                                                             \
    if (int rc_ = ensure_dynamic_smem((const void*)outlayer_mma_kernel<ACT>, smem)) return rc_;                    \
    DDNN_CUDA_CHECK(launch_pdl(outlayer_mma_kernel<ACT>, dim3(mma_grid(B)), dim3(256), smem, st, (const bf16*)A,   \
                               (const bf16*)W, bias, labels, (bf16*)dZ, row_loss, (bf16*)dZprev, dw_part, db_part, \
                               B, (int)K, (int)N, np, inv_denom, tiles));                                         \
  } while (0)
  switch (act_prev) {
    case DDNN_ACT_RELU: DDNN_OUT_MMA(DDNN_ACT_RELU); break;
    case DDNN_ACT_SIGMOID: DDNN_OUT_MMA(DDNN_ACT_SIGMOID); break;
    default: DDNN_OUT_MMA(DDNN_ACT_LINEAR); break;
  }
#undef DDNN_OUT_MMA
  count_launch();
  DDNN_LAUNCH_CHECK();
  return DDNN_OK;
}

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
  DDNN_CUDA_CHECK(launch_pdl(outlayer_fused_kernel<NV, MC, T>, dim3(blocks), dim3(256), smem, st, (const T*)A,  \
                             (const T*)W, bias, labels, (T*)dZ, row_loss, (T*)dZprev, dw_part, db_part, B, (int)K,  \
                             (int)N, np, act_prev, inv_denom, rpb))
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
  if (dtype == DDNN_BF16 && g_outlayer_mma.load() && use_mma(K, dtype))
    return outlayer_launch_mma(A, W, bias, labels, dZ, row_loss, dZprev, dw_part, db_part, B, K, N, act_prev, inv_denom,
                               st);
  if (dtype == DDNN_BF16)
    return outlayer_launch<bf16>(A, W, bias, labels, dZ, row_loss, dZprev, dw_part, db_part, B, K, N, act_prev,
                                 inv_denom, st);
  set_error("outlayer_fused: unsupported dtype %d", dtype);
  return DDNN_ERR_UNSUPPORTED;
}

}  // namespace ddnn
