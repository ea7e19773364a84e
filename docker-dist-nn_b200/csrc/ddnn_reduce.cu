// Deterministic split-K reduction fused with the dW/db epilogue (SGD step or fp32 store),
// plus the bias-gradient column reduction used beside the tcgen05 dW GEMM.
#include "ddnn_internal.h"

namespace ddnn {
namespace {

// part: [splits][rows][np] fp32.  Row index `rows_w` .. is the bias row (db partials live in
// a second array).  One thread = 4 consecutive columns, partials summed in z order.
template <typename EpiW, typename EpiB>
__global__ void __launch_bounds__(256)
reduce_dw_kernel(const float* __restrict__ dw_part, const float* __restrict__ db_part, int splits, int64_t rows,
                 int np, int N, EpiW ew, EpiB eb, bool vecw, bool vecb, int has_db) {
  const int q = np >> 2;
  const int64_t total = (rows + (has_db ? 1 : 0)) * q;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row = idx / q;
    const int col = (int)(idx - row * q) * 4;
    if (col >= N) continue;
    const int nv = min(4, N - col);
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    if (row < rows) {
      for (int z = 0; z < splits; ++z) {
        float4 v = ld4(dw_part + ((int64_t)z * rows + row) * np + col);
        s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
      }
      ew.apply4(row, col, s, nv, vecw);
    } else {
      for (int z = 0; z < splits; ++z) {
        float4 v = ld4(db_part + (int64_t)z * np + col);
        s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
      }
      eb.apply4(0, col, s, nv, vecb);
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
colsum_kernel(const T* __restrict__ dZ, float* __restrict__ part, int64_t B, int N, int np, int rlen, bool vec) {
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

template <typename T>
bool vec_ok(const T* p, int64_t ld) {
  return p && (ld % 4 == 0) && ((reinterpret_cast<uintptr_t>(p) % (4 * sizeof(T))) == 0);
}

}  // namespace

int reduce_dw(const float* dw_part, const float* db_part, int splits, int64_t K, int64_t N, float* W_master,
              void* W_shadow, float* bias, float* dW_out, float* db_out, float lr, int apply, cudaStream_t st) {
  const int np = (int)((N + 3) & ~(int64_t)3);
  const int64_t rows = dw_part ? K : 0;  // no dW partials: only the bias row is reduced
  const int has_db = db_part != nullptr;
  const int64_t total = (rows + has_db) * (np / 4);
  if (total == 0) return DDNN_OK;
  int blocks = (int)((total + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  if (blocks < 1) blocks = 1;
  if (apply) {
    EpiSgd ew{W_master, (bf16*)W_shadow, N, lr};
    EpiSgd eb{bias, nullptr, N, lr};
    const bool vw = vec_ok(W_master, N) && (!W_shadow || vec_ok((bf16*)W_shadow, N));
    reduce_dw_kernel<EpiSgd, EpiSgd><<<blocks, 256, 0, st>>>(dw_part, db_part, splits, rows, np, (int)N, ew, eb, vw,
                                                           vec_ok(bias, N), has_db);
  } else {
    EpiStoreF32 ew{dW_out, N};
    EpiStoreF32 eb{db_out, N};
    reduce_dw_kernel<EpiStoreF32, EpiStoreF32><<<blocks, 256, 0, st>>>(dw_part, db_part, splits, rows, np, (int)N, ew, eb,
                                                                     vec_ok(dW_out, N), vec_ok(db_out, N), has_db);
  }
  count_launch();
  DDNN_LAUNCH_CHECK();
  return DDNN_OK;
}

int colsum_splits(int64_t B, int64_t N) {
  long strips = cdiv(N, 128);
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
    colsum_kernel<float><<<grid, 256, 0, st>>>((const float*)dZ, db_part, B, (int)N, np, rlen,
                                               vec_ok((const float*)dZ, N));
  } else if (dtype == DDNN_BF16) {
    colsum_kernel<bf16><<<grid, 256, 0, st>>>((const bf16*)dZ, db_part, B, (int)N, np, rlen,
                                              vec_ok((const bf16*)dZ, N));
  } else {
    set_error("colsum: unsupported dtype %d", dtype);
    return DDNN_ERR_UNSUPPORTED;
  }
  count_launch();
  DDNN_LAUNCH_CHECK();
  return DDNN_OK;
}

}  // namespace ddnn
