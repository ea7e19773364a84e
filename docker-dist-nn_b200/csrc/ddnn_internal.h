// Internal launch interface between ddnn_api.cu and the kernel translation units.
#pragma once
#include "ddnn_common.cuh"

namespace ddnn {

// ---- SIMT fp32-accumulate GEMMs (ddnn_simt.cu) ---------------------------------------
int simt_forward(const void* X, const void* W, const float* bias, void* A, int64_t B, int64_t K,
                 int64_t N, int act, int dtype, int out_dtype, cudaStream_t st);
int simt_dx(const void* dZ, const void* W, const void* Aprev, void* dZprev, int64_t B, int64_t K,
            int64_t N, int act_prev, int dtype, cudaStream_t st);
// Writes split-K partials: dw_part[z][K][Np] and db_part[z][Np]  (Np = N rounded up to 4).
int simt_dw_partials(const void* X, const void* dZ, float* dw_part, float* db_part, int splits,
                     int64_t B, int64_t K, int64_t N, int dtype, cudaStream_t st);
int simt_dw_splits(int64_t B, int64_t K, int64_t N);
void simt_set_f32_core(int tensor);  // 1: 3xTF32 mma inner product (default), 0: FFMA

// ---- skinny output layers, N <= 16 (ddnn_skinny.cu) ------------------------------------
bool skinny_ok(int64_t K, int64_t N);
int skinny_forward(const void* X, const void* W, const float* bias, void* A, int64_t B, int64_t K,
                   int64_t N, int act, int dtype, int out_dtype, cudaStream_t st);
int skinny_dx(const void* dZ, const void* W, const void* Aprev, void* dZprev, int64_t B, int64_t K,
              int64_t N, int act_prev, int dtype, cudaStream_t st);
int skinny_dw_splits(int64_t B, int64_t K, int64_t N);
int skinny_dw_partials(const void* X, const void* dZ, float* dw_part, float* db_part, int splits,
                       int64_t B, int64_t K, int64_t N, int dtype, cudaStream_t st);

// ---- fused skinny output layer: logits + softmax-CE + dX + dW/db partials (ddnn_outlayer.cu) ----
bool outlayer_fused_ok(const void* A, int64_t K, int64_t N, int dtype);
int outlayer_splits(int64_t B, int64_t K, int dtype);      // of the variant the current engine setting selects
int outlayer_max_splits(int64_t B, int64_t K, int dtype);  // over both variants (workspace sizing)
void outlayer_set_tensor_cores(int on);  // bf16: 1 = mma.sync variant (default), 0 = FFMA variant
int outlayer_fused(const void* A, const void* W, const float* bias, const int32_t* labels, void* dZ,
                   float* row_loss, void* dZprev, float* dw_part, float* db_part, int64_t B, int64_t K,
                   int64_t N, int act_prev, float inv_denom, int dtype, cudaStream_t st);
int loss_reduce(const float* row_loss, float* loss, int64_t B, float scale, int accumulate, cudaStream_t st);

// ---- split-K reduction + dW epilogue (ddnn_reduce.cu) --------------------------------
// Sums partials in z order, then SGD (apply) or fp32 store.
int reduce_dw(const float* dw_part, const float* db_part, int splits, int64_t K, int64_t N,
              float* W_master, void* W_shadow, float* bias, float* dW_out, float* db_out, float lr,
              int apply, cudaStream_t st);
// same with separate split counts for the dW partials and the db partials (one launch instead of two)
int reduce_dw2(const float* dw_part, const float* db_part, int splits, int splits_b, int64_t K, int64_t N,
               float* W_master, void* W_shadow, float* bias, float* dW_out, float* db_out, float lr,
               int apply, cudaStream_t st);
// W -= lr * G, bias -= lr * gb, shadow = bf16(W) (any N): the update after chunked gradient accumulation
int sgd_apply(float* W, void* W_shadow, float* bias, const float* G, const float* gb, float lr, int64_t K, int64_t N,
              cudaStream_t st);
// db partials from dZ alone (used by the tcgen05 path): db_part[z][Np]
int colsum_partials(const void* dZ, float* db_part, int splits, int64_t B, int64_t N, int dtype,
                    cudaStream_t st);
int colsum_splits(int64_t B, int64_t N);

// ---- softmax family (ddnn_softmax.cu) -------------------------------------------------
int softmax_rows(const float* Z, float* P, int64_t B, int64_t N, cudaStream_t st);
int softmax_ce(const float* Z, const int32_t* labels, float* P, void* dZ, float* row_loss,
               float* loss, int64_t B, int64_t N, float inv_denom, int dz_dtype, int accumulate,
               cudaStream_t st);
int activate(const float* V, float* out, int64_t B, int64_t N, int act, cudaStream_t st);
int cast_f32_bf16(const float* in, bf16* out, int64_t n, cudaStream_t st);
int cast_bf16_f32(const bf16* in, float* out, int64_t n, cudaStream_t st);

// ---- tcgen05 / TMA / TMEM GEMMs (ddnn_umma.cu) ----------------------------------------
bool umma_shape_ok(int64_t B, int64_t K, int64_t N);
void umma_set_sm_reserve(int sms);
void umma_set_pair(int on);  // 1: CTA-pair (cta_group::2) kernel where it applies (default)
// sig (optional): pipeline signalling folded into the kernel, see ddnn_signals in include/ddnn.h
int umma_forward(const bf16* X, const bf16* W, const float* bias, void* A, int64_t B, int64_t K,
                 int64_t N, int act, int out_dtype, cudaStream_t st, const ddnn_signals* sig = nullptr);
int umma_dx(const bf16* dZ, const bf16* W, const bf16* Aprev, bf16* dZprev, int64_t B, int64_t K,
            int64_t N, int act_prev, cudaStream_t st, const ddnn_signals* sig = nullptr);
// Split-K plan for dW on tcgen05; splits == 1 means the SGD/store epilogue is fused.
int umma_dw_splits(int64_t B, int64_t K, int64_t N);
// db_part/db_fused (optional): when the split-K plan puts one tile on every CTA the kernel also writes the
// bias-gradient partials db_part[split][Np] (ones-operand MMA) and sets *db_fused = 1; otherwise the caller
// runs colsum_partials.
int umma_dw(const bf16* X, const bf16* dZ, float* dw_part, int splits, float* W_master,
            bf16* W_shadow, float* dW_out, float lr, int apply, int64_t B, int64_t K, int64_t N,
            float* db_part, int* db_fused, cudaStream_t st);

// ---- fused last two layers of a small fp32 MLP (ddnn_tail2.cu) ----------------------------
bool tail2_ok(int64_t K1, int64_t N2, int64_t N3, int dtype);
int tail2_splits(int64_t B);
int tail2_max_splits(int64_t B);  // bound over every batch of at most B rows
int tail2_launch(const float* A1, const float* W2, const float* b2, const float* W3, const float* b3,
                 const int32_t* labels, float* dZ1, float* row_loss, float* dw2_part, float* db2_part,
                 float* dw3_part, float* db3_part, int64_t B, int64_t K1, int64_t N2, int64_t N3, int act1,
                 int act2, float inv_denom, int tensor_cores /* 1: 3xTF32 mma.sync products, 0: FFMA */,
                 cudaStream_t st);

// ---- fp32 operands on tcgen05 as error-compensated TF32 (ddnn_tf32.cu) -----------------
constexpr int64_t DDNN_TF32_MAX_CHAIN = 2048;  // longest reduction accumulated in one TMEM accumulator
bool tf32_shape_ok(int64_t B, int64_t K, int64_t N);
int64_t tf32_max_reduction(int64_t M, int64_t N);  // forward / dX with an (M x N) output: longest reduction taken
int tf32_forward(const float* X, const float* W, const float* bias, float* A, int64_t B, int64_t K, int64_t N,
                 int act, cudaStream_t st);
int tf32_dx(const float* dZ, const float* W, const float* Aprev, float* dZprev, int64_t B, int64_t K, int64_t N,
            int act_prev, cudaStream_t st);
int tf32_dw_splits(int64_t B, int64_t K, int64_t N);
// split-K partials dw_part[split][K][Np]; like umma_dw, also writes db_part[split][Np] (*db_fused = 1) when every
// CTA has one tile, otherwise the caller runs colsum_partials; then reduce_dw2
int tf32_dw_partials(const float* X, const float* dZ, float* dw_part, int splits, int64_t B, int64_t K,
                     int64_t N, float* db_part, int* db_fused, cudaStream_t st);

// dW + db + SGD (or gradient store) in one launch: split-K over a thread-block cluster, reduced through distributed
// shared memory.  *taken = 0: no cluster plan for this shape, run tf32_dw_partials + reduce_dw2 instead.
int tf32_dw_cluster(const float* X, const float* dZ, float* W_master, float* bias, float* dW_out, float* db_out, float lr,
                    int apply, int64_t B, int64_t K, int64_t N, int* taken, cudaStream_t st);
void tf32_set_cluster_dw(int on);

}  // namespace ddnn
