// extern "C" boundary of libddnn.so (see include/ddnn.h).  Validates arguments, picks the
// engine (tcgen05 GEMMs for aligned bf16 layers, SIMT kernels otherwise) and launches.
#include <cstring>
#include <mutex>
#include <map>
#include <set>
#include <string>
#include <utility>

#include "ddnn_internal.h"

namespace ddnn {

static thread_local std::string g_err;
std::atomic<long long> g_launches{0};
std::atomic<int> g_pdl{1};
static std::atomic<int> g_engine{DDNN_ENGINE_AUTO};
static std::atomic<int> g_f32_tc{DDNN_F32_TC_DEFAULT};  // inner product of the fp32 GEMMs, see ddnn_set_f32_tensor_cores

void set_error(const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_err = buf;
}

int ensure_dynamic_smem(const void* kernel, size_t bytes) {
  // the attribute is a per-(kernel, device) high-water mark: raise it when a later shape needs more
  static std::mutex mu;
  static std::map<std::pair<const void*, int>, size_t> done;
  int dev = 0;
  DDNN_CUDA_CHECK(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> g(mu);
  auto it = done.find({kernel, dev});
  if (it != done.end() && it->second >= bytes) return DDNN_OK;
  DDNN_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  done[{kernel, dev}] = bytes;
  return DDNN_OK;
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

static int check_dims(const char* fn, int64_t B, int64_t K, int64_t N) {
  if (B < 0 || K <= 0 || N <= 0 || B > INT32_MAX || K > INT32_MAX || N > INT32_MAX) {
    set_error("%s: bad dimensions B=%lld K=%lld N=%lld", fn, (long long)B, (long long)K, (long long)N);
    return DDNN_ERR_DIM;
  }
  return DDNN_OK;
}

// tcgen05 path: bf16 operands, 16-byte aligned rows (widths % 8) and pointers.
static bool use_umma(int dtype, int64_t B, int64_t K, int64_t N, std::initializer_list<const void*> ptrs) {
  const int e = g_engine.load();
  if (e == DDNN_ENGINE_SIMT || dtype != DDNN_BF16) return false;
  if (!umma_shape_ok(B, K, N)) return false;
  for (const void* p : ptrs)
    if (p && !aligned16(p)) return false;
  if (e == DDNN_ENGINE_TCGEN05) return true;
  // AUTO: tiny layers are launch-latency bound either way; keep them on the SIMT kernels
  return K >= 64 && N >= 64;
}

// Skinny kernels (N <= 16): one coalesced pass over the big operand instead of a 64-wide tile.
// fp32 operands on tcgen05 (error-compensated TF32): 16-byte aligned rows and pointers, widths >= 32.
// `red` = length of an un-split reduction: the tensor core adds into its fp32 accumulator with truncation, so
// the error grows linearly with the chain (measured vs float64: 2.6e-6 at 1024, 1.2e-5 at 4096); chains longer
// than DDNN_TF32_MAX_CHAIN per accumulator set (the narrow tiles rotate their k-blocks over 4 or 2 sets) stay on
// FFMA (forward / dX) or get more splits (dW).
static bool use_tf32(int dtype, int64_t B, int64_t K, int64_t N, int64_t red, std::initializer_list<const void*> ptrs) {
  if (g_engine.load() == DDNN_ENGINE_SIMT || dtype != DDNN_F32 || g_f32_tc.load() != 2) return false;
  if (!tf32_shape_ok(B, K, N) || red > tf32_max_reduction(B, N)) return false;
  for (const void* p : ptrs)
    if (p && !aligned16(p)) return false;
  return true;
}

static bool use_skinny(int64_t K, int64_t N) {
  return g_engine.load() == DDNN_ENGINE_AUTO && skinny_ok(K, N);
}

}  // namespace ddnn

using namespace ddnn;

extern "C" {

int ddnn_version(void) { return 100; }

const char* ddnn_last_error(void) { return g_err.c_str(); }

int ddnn_activation_from_name(const char* name) {
  if (!name) return DDNN_ACT_LINEAR;
  if (!strcmp(name, "relu")) return DDNN_ACT_RELU;
  if (!strcmp(name, "sigmoid")) return DDNN_ACT_SIGMOID;
  if (!strcmp(name, "softmax")) return DDNN_ACT_SOFTMAX;
  return DDNN_ACT_LINEAR;  // grpc_node.py:72-73: everything else is the identity
}

int ddnn_set_engine(int engine) {
  if (engine < DDNN_ENGINE_AUTO || engine > DDNN_ENGINE_TCGEN05) {
    set_error("ddnn_set_engine: bad engine %d", engine);
    return DDNN_ERR_ARG;
  }
  g_engine.store(engine);
  outlayer_set_tensor_cores(engine != DDNN_ENGINE_SIMT);  // SIMT pins the fused output layer to its FFMA variant
  return DDNN_OK;
}

int ddnn_set_pdl(int on) {
  g_pdl.store(on ? 1 : 0);
  return DDNN_OK;
}

int ddnn_set_cta_pairs(int on) {
  umma_set_pair(on);
  return DDNN_OK;
}

int ddnn_set_f32_tensor_cores(int mode) {
  if (mode < 0 || mode > 2) {
    set_error("ddnn_set_f32_tensor_cores: mode %d is not 0 (FFMA), 1 (mma.sync 3xTF32) or 2 (tcgen05 3xTF32)", mode);
    return DDNN_ERR_ARG;
  }
  g_f32_tc.store(mode);
  simt_set_f32_core(mode == 1);
  return DDNN_OK;
}

int ddnn_set_cluster_split_k(int on) {
  tf32_set_cluster_dw(on);
  return DDNN_OK;
}

int ddnn_set_sm_reserve(int sms) {
  if (sms < 0 || sms > 64) {
    set_error("ddnn_set_sm_reserve: %d is out of range [0, 64]", sms);
    return DDNN_ERR_ARG;
  }
  umma_set_sm_reserve(sms);
  return DDNN_OK;
}

int64_t ddnn_launch_count(void) { return (int64_t)g_launches.load(); }

static int check_signals(const char* fn, const ddnn_signals* sig) {
  if (sig && sig->done_flag && !sig->counter) {
    set_error("%s: ddnn_signals.done_flag needs a counter", fn);
    return DDNN_ERR_ARG;
  }
  return DDNN_OK;
}

int ddnn_dense_forward(const void* X, const void* W, const float* bias, void* A, int64_t B, int64_t K, int64_t N,
                       int act, int dtype, int out_dtype, void* stream) {
  return ddnn_dense_forward_signalled(X, W, bias, A, B, K, N, act, dtype, out_dtype, nullptr, stream);
}

int ddnn_dense_forward_signalled(const void* X, const void* W, const float* bias, void* A, int64_t B, int64_t K,
                                 int64_t N, int act, int dtype, int out_dtype, const ddnn_signals* sig, void* stream) {
  int rc = check_dims("ddnn_dense_forward", B, K, N);
  if (rc) return rc;
  if ((rc = check_signals("ddnn_dense_forward_signalled", sig))) return rc;
  if (rc) return rc;
  if (B == 0) return DDNN_OK;
  if (!X || !W || !bias || !A) {
    set_error("ddnn_dense_forward: null pointer");
    return DDNN_ERR_ARG;
  }
  if (act < DDNN_ACT_LINEAR || act > DDNN_ACT_SOFTMAX) {
    set_error("ddnn_dense_forward: bad activation %d", act);
    return DDNN_ERR_ACTIVATION;
  }
  if (act == DDNN_ACT_SOFTMAX && out_dtype != DDNN_F32) {
    set_error("ddnn_dense_forward: softmax output must be fp32");
    return DDNN_ERR_UNSUPPORTED;
  }
  cudaStream_t st = (cudaStream_t)stream;
  const int gemm_act = act == DDNN_ACT_SOFTMAX ? DDNN_ACT_LINEAR : act;
  const bool umma = !use_skinny(K, N) && use_umma(dtype, B, K, N, {X, W, A});
  // the tcgen05 kernels take the signals themselves (a softmax pass after the GEMM would run past done_flag);
  // every other path brackets the plain launch with the stream-ordered flag operations
  const bool in_kernel = sig && umma && act != DDNN_ACT_SOFTMAX;
  if (sig && !in_kernel && sig->wait_flag && (rc = ddnn_flag_wait(sig->wait_flag, sig->wait_value, stream))) return rc;
  if (use_skinny(K, N))
    rc = skinny_forward(X, W, bias, A, B, K, N, gemm_act, dtype, out_dtype, st);
  else if (umma)
    rc = umma_forward((const bf16*)X, (const bf16*)W, bias, A, B, K, N, gemm_act, out_dtype, st,
                      in_kernel ? sig : nullptr);
  else if (out_dtype == DDNN_F32 && use_tf32(dtype, B, K, N, K, {X, W, A, bias}))
    rc = tf32_forward((const float*)X, (const float*)W, bias, (float*)A, B, K, N, gemm_act, st);
  else
    rc = simt_forward(X, W, bias, A, B, K, N, gemm_act, dtype, out_dtype, st);
  if (rc) return rc;
  if (act == DDNN_ACT_SOFTMAX) rc = softmax_rows((const float*)A, (float*)A, B, N, st);
  if (rc) return rc;
  if (sig && !in_kernel && sig->done_flag) rc = ddnn_flag_write(sig->done_flag, sig->done_value, stream);
  return rc;
}

int ddnn_softmax_rows(const float* Z, float* P, int64_t B, int64_t N, void* stream) {
  int rc = check_dims("ddnn_softmax_rows", B, 1, N);
  if (rc) return rc;
  if (B == 0) return DDNN_OK;
  if (!Z || !P) {
    set_error("ddnn_softmax_rows: null pointer");
    return DDNN_ERR_ARG;
  }
  return softmax_rows(Z, P, B, N, (cudaStream_t)stream);
}

int ddnn_activate(const float* V, float* out, int64_t B, int64_t N, int act, void* stream) {
  int rc = check_dims("ddnn_activate", B, 1, N);
  if (rc) return rc;
  if (B == 0) return DDNN_OK;
  if (!V || !out) {
    set_error("ddnn_activate: null pointer");
    return DDNN_ERR_ARG;
  }
  if (act < DDNN_ACT_LINEAR || act > DDNN_ACT_SOFTMAX) {
    set_error("ddnn_activate: bad activation %d", act);
    return DDNN_ERR_ACTIVATION;
  }
  return activate(V, out, B, N, act, (cudaStream_t)stream);
}

int ddnn_softmax_ce(const float* Z, const int32_t* labels, float* P, void* dZ, float* row_loss, float* loss,
                    int64_t B, int64_t N, float inv_denom, int dz_dtype, int accumulate, void* stream) {
  int rc = check_dims("ddnn_softmax_ce", B, 1, N);
  if (rc) return rc;
  if (B == 0) return DDNN_OK;
  if (!Z || !labels) {
    set_error("ddnn_softmax_ce: null pointer");
    return DDNN_ERR_ARG;
  }
  if (P == Z) {
    set_error("ddnn_softmax_ce: P must not alias Z");
    return DDNN_ERR_ARG;
  }
  return softmax_ce(Z, labels, P, dZ, row_loss, loss, B, N, inv_denom, dz_dtype, accumulate, (cudaStream_t)stream);
}

int ddnn_dense_backward_dx(const void* dZ, const void* W, const void* Aprev, void* dZprev, int64_t B, int64_t K,
                           int64_t N, int act_prev, int dtype, void* stream) {
  return ddnn_dense_backward_dx_signalled(dZ, W, Aprev, dZprev, B, K, N, act_prev, dtype, nullptr, stream);
}

int ddnn_dense_backward_dx_signalled(const void* dZ, const void* W, const void* Aprev, void* dZprev, int64_t B,
                                     int64_t K, int64_t N, int act_prev, int dtype, const ddnn_signals* sig,
                                     void* stream) {
  int rc = check_dims("ddnn_dense_backward_dx", B, K, N);
  if (rc) return rc;
  if ((rc = check_signals("ddnn_dense_backward_dx_signalled", sig))) return rc;
  if (B == 0) return DDNN_OK;
  if (!dZ || !W || !dZprev) {
    set_error("ddnn_dense_backward_dx: null pointer");
    return DDNN_ERR_ARG;
  }
  if (act_prev == DDNN_ACT_SOFTMAX) {
    set_error("ddnn_dense_backward_dx: hidden softmax layers are not supported (row Jacobian)");
    return DDNN_ERR_ACTIVATION;
  }
  if (act_prev < DDNN_ACT_LINEAR || act_prev > DDNN_ACT_SOFTMAX) {
    set_error("ddnn_dense_backward_dx: bad activation %d", act_prev);
    return DDNN_ERR_ACTIVATION;
  }
  if (act_prev != DDNN_ACT_LINEAR && !Aprev) {
    set_error("ddnn_dense_backward_dx: Aprev is required for activation %d", act_prev);
    return DDNN_ERR_ARG;
  }
  cudaStream_t st = (cudaStream_t)stream;
  const bool umma = !use_skinny(K, N) && use_umma(dtype, B, K, N, {dZ, W, Aprev, dZprev});
  if (umma)  // signals (if any) handled inside the kernel
    return umma_dx((const bf16*)dZ, (const bf16*)W, (const bf16*)Aprev, (bf16*)dZprev, B, K, N, act_prev, st, sig);
  if (sig && sig->wait_flag && (rc = ddnn_flag_wait(sig->wait_flag, sig->wait_value, stream))) return rc;
  if (use_skinny(K, N))
    rc = skinny_dx(dZ, W, Aprev, dZprev, B, K, N, act_prev, dtype, st);
  else if (use_tf32(dtype, B, N, K, N, {dZ, W, Aprev, dZprev}))  // output (B x K), reduction over N
    rc = tf32_dx((const float*)dZ, (const float*)W, (const float*)Aprev, (float*)dZprev, B, K, N, act_prev, st);
  else
    rc = simt_dx(dZ, W, Aprev, dZprev, B, K, N, act_prev, dtype, st);
  if (rc) return rc;
  if (sig && sig->done_flag) rc = ddnn_flag_write(sig->done_flag, sig->done_value, stream);
  return rc;
}

static size_t np_of(int64_t N) { return (size_t)((N + 3) & ~(int64_t)3); }

// Split count a deferred ddnn_output_layer_train wrote into a workspace, keyed by the workspace pointer:
// ddnn_output_layer_finish reads the partials with THAT layout even if the engine switch was flipped in
// between (the two calls are meant to be issued at different times on different streams).
static std::mutex g_defer_mu;
static std::map<const void*, int> g_defer_splits;
static void defer_record(const void* ws, int splits) {
  std::lock_guard<std::mutex> g(g_defer_mu);
  g_defer_splits[ws] = splits;
}
static int defer_lookup(const void* ws, int fallback) {
  std::lock_guard<std::mutex> g(g_defer_mu);
  auto it = g_defer_splits.find(ws);
  return it == g_defer_splits.end() ? fallback : it->second;
}

size_t ddnn_dw_workspace_bytes(int64_t B, int64_t K, int64_t N, int dtype) {
  if (B <= 0 || K <= 0 || N <= 0) return 0;
  const size_t np = np_of(N);
  // SIMT plan: dW partials [splits][K][np] + db partials [splits][np]
  const size_t s1 = (size_t)simt_dw_splits(B, K, N);
  size_t simt = s1 * ((size_t)K + 1) * np * sizeof(float);
  size_t umma = 0;
  if (dtype == DDNN_BF16 && umma_shape_ok(B, K, N)) {
    const size_t s2 = (size_t)umma_dw_splits(B, K, N);
    const size_t cs = (size_t)colsum_splits(B, N);
    umma = (s2 > 1 ? s2 * (size_t)K * np : 0) * sizeof(float) + (cs > s2 ? cs : s2) * np * sizeof(float);
  }
  size_t need = simt > umma ? simt : umma;
  if (dtype == DDNN_F32 && tf32_shape_ok(B, K, N)) {
    const size_t s3 = (size_t)tf32_dw_splits(B, K, N), cs = (size_t)colsum_splits(B, N);
    const size_t t = (s3 * (size_t)K * np + (cs > s3 ? cs : s3) * np) * sizeof(float);
    if (t > need) need = t;
  }
  if (skinny_ok(K, N)) {
    const size_t sk = (size_t)skinny_dw_splits(B, K, N) * ((size_t)K + 1) * np * sizeof(float);
    if (sk > need) need = sk;
  }
  return (need + 255) & ~(size_t)255;
}

int ddnn_dense_backward_dw_sgd(const void* X, const void* dZ, float* W_master, void* W_shadow, float* bias,
                               float* dW_out, float* db_out, float lr, int apply, int64_t B, int64_t K, int64_t N,
                               int dtype, void* workspace, size_t workspace_bytes, void* stream) {
  int rc = check_dims("ddnn_dense_backward_dw_sgd", B, K, N);
  if (rc) return rc;
  if (B == 0) return DDNN_OK;
  if (!X || !dZ) {
    set_error("ddnn_dense_backward_dw_sgd: null pointer");
    return DDNN_ERR_ARG;
  }
  if (apply ? (!W_master || !bias) : (!dW_out || !db_out)) {
    set_error("ddnn_dense_backward_dw_sgd: missing %s", apply ? "W_master/bias" : "dW_out/db_out");
    return DDNN_ERR_ARG;
  }
  const size_t need = ddnn_dw_workspace_bytes(B, K, N, dtype);
  if (!workspace || workspace_bytes < need || !aligned16(workspace)) {
    set_error("ddnn_dense_backward_dw_sgd: workspace %zu bytes < %zu required (or misaligned)", workspace_bytes, need);
    return DDNN_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  const size_t np = np_of(N);
  float* ws = (float*)workspace;
  if (use_skinny(K, N)) {
    const int splits = skinny_dw_splits(B, K, N);
    float* dw_part = ws;
    float* db_part = ws + (size_t)splits * K * np;
    rc = skinny_dw_partials(X, dZ, dw_part, db_part, splits, B, K, N, dtype, st);
    if (rc) return rc;
    return reduce_dw(dw_part, db_part, splits, K, N, W_master, W_shadow, bias, dW_out, db_out, lr, apply, st);
  }
  if (use_umma(dtype, B, K, N, {X, dZ, W_master, W_shadow, dW_out})) {
    const int splits = umma_dw_splits(B, K, N);
    float* dw_part = splits > 1 ? ws : nullptr;
    float* db_part = ws + (splits > 1 ? (size_t)splits * K * np : 0);
    int cs = colsum_splits(B, N), db_fused = 0;
    rc = umma_dw((const bf16*)X, (const bf16*)dZ, dw_part, splits, W_master, (bf16*)W_shadow, dW_out, lr, apply, B, K,
                 N, db_part, &db_fused, st);
    if (rc) return rc;
    if (db_fused) {
      cs = splits;  // the GEMM wrote one bias-gradient partial per split
    } else {
      rc = colsum_partials(dZ, db_part, cs, B, N, dtype, st);
      if (rc) return rc;
    }
    if (splits > 1)  // weights (splits) and bias row (cs) reduced in one launch
      return reduce_dw2(dw_part, db_part, splits, cs, K, N, W_master, W_shadow, bias, dW_out, db_out, lr, apply, st);
    return reduce_dw(dw_part, db_part, cs, K, N, W_master, W_shadow, bias, dW_out, db_out, lr, apply, st);
  }
  if (use_tf32(dtype, B, K, N, 0, {X, dZ, W_master, dW_out})) {
    int taken = 0;  // split-K reduced inside the kernel across a thread-block cluster: one launch, no partials in HBM
    rc = tf32_dw_cluster((const float*)X, (const float*)dZ, W_master, bias, dW_out, db_out, lr, apply, B, K, N, &taken, st);
    if (rc || taken) return rc;
    const int splits = tf32_dw_splits(B, K, N);
    int cs = colsum_splits(B, N), db_fused = 0;
    float* dw_part = ws;
    float* db_part = ws + (size_t)splits * K * np;
    rc = tf32_dw_partials((const float*)X, (const float*)dZ, dw_part, splits, B, K, N, db_part, &db_fused, st);
    if (rc) return rc;
    if (db_fused) {
      cs = splits;
    } else {
      rc = colsum_partials(dZ, db_part, cs, B, N, dtype, st);
      if (rc) return rc;
    }
    return reduce_dw2(dw_part, db_part, splits, cs, K, N, W_master, W_shadow, bias, dW_out, db_out, lr, apply, st);
  }
  const int splits = simt_dw_splits(B, K, N);
  float* dw_part = ws;
  float* db_part = ws + (size_t)splits * K * np;
  rc = simt_dw_partials(X, dZ, dw_part, db_part, splits, B, K, N, dtype, st);
  if (rc) return rc;
  return reduce_dw(dw_part, db_part, splits, K, N, W_master, W_shadow, bias, dW_out, db_out, lr, apply, st);
}

int ddnn_dense_backward_dw_accumulate(const void* X, const void* dZ, float* dW_acc, float* db_acc, int first, int64_t B,
                                      int64_t K, int64_t N, int dtype, void* workspace, size_t workspace_bytes,
                                      void* stream) {
  if (!dW_acc || !db_acc) {
    set_error("ddnn_dense_backward_dw_accumulate: null accumulator");
    return DDNN_ERR_ARG;
  }
  if (first)  // plain gradient store
    return ddnn_dense_backward_dw_sgd(X, dZ, nullptr, nullptr, nullptr, dW_acc, db_acc, 0.0f, 0, B, K, N, dtype,
                                      workspace, workspace_bytes, stream);
  // acc - (-1) * g: the update epilogue with the accumulator as the "weight" (an FMA with multiplier 1: exact sum,
  // rounded once)
  return ddnn_dense_backward_dw_sgd(X, dZ, dW_acc, nullptr, db_acc, nullptr, nullptr, -1.0f, 1, B, K, N, dtype,
                                    workspace, workspace_bytes, stream);
}

int ddnn_sgd_apply(float* W_master, void* W_shadow, float* bias, const float* dW, const float* db, float lr, int64_t K,
                   int64_t N, void* stream) {
  int rc = check_dims("ddnn_sgd_apply", 1, K, N);
  if (rc) return rc;
  if (!W_master || !bias || !dW || !db) {
    set_error("ddnn_sgd_apply: null pointer");
    return DDNN_ERR_ARG;
  }
  return sgd_apply(W_master, W_shadow, bias, dW, db, lr, K, N, (cudaStream_t)stream);
}

size_t ddnn_output_layer_workspace_bytes(int64_t B, int64_t K, int64_t N, int dtype) {
  if (B <= 0 || K <= 0 || N <= 0) return 0;
  if (!outlayer_fused_ok(nullptr, K, N, dtype)) return 0;  // shapes the entry point rejects need no scratch
  const size_t need = (size_t)outlayer_max_splits(B, K, dtype) * ((size_t)K + 1) * np_of(N) * sizeof(float);
  return (need + 255) & ~(size_t)255;
}

int ddnn_output_layer_train(const void* A, const void* W, const int32_t* labels, void* dZ, float* row_loss,
                            float* loss, void* dZprev, float* W_master, void* W_shadow, float* bias, float* dW_out,
                            float* db_out, float lr, int apply, int64_t B, int64_t K, int64_t N, int act_prev,
                            float inv_denom, int accumulate_loss, int dtype, void* workspace,
                            size_t workspace_bytes, void* stream) {
  int rc = check_dims("ddnn_output_layer_train", B, K, N);
  if (rc) return rc;
  if (B == 0) return DDNN_OK;
  if (!A || !W || !labels || !row_loss || !bias) {
    set_error("ddnn_output_layer_train: null pointer");
    return DDNN_ERR_ARG;
  }
  if (apply < 0 || apply > DDNN_OUTLAYER_DEFER) {
    set_error("ddnn_output_layer_train: apply = %d", apply);
    return DDNN_ERR_ARG;
  }
  if (apply != DDNN_OUTLAYER_DEFER && (apply ? !W_master : (!dW_out || !db_out))) {
    set_error("ddnn_output_layer_train: missing %s", apply ? "W_master" : "dW_out/db_out");
    return DDNN_ERR_ARG;
  }
  if (act_prev < DDNN_ACT_LINEAR || act_prev >= DDNN_ACT_SOFTMAX) {
    set_error("ddnn_output_layer_train: activation %d of the feeding layer is not supported", act_prev);
    return DDNN_ERR_ACTIVATION;
  }
  if (!outlayer_fused_ok(A, K, N, dtype)) {
    set_error("ddnn_output_layer_train: needs N <= 16, K <= 256, 16-byte aligned rows (got K=%lld N=%lld)",
              (long long)K, (long long)N);
    return DDNN_ERR_UNSUPPORTED;
  }
  const size_t need = ddnn_output_layer_workspace_bytes(B, K, N, dtype);
  if (!workspace || workspace_bytes < need || !aligned16(workspace)) {
    set_error("ddnn_output_layer_train: workspace %zu bytes < %zu required (or misaligned)", workspace_bytes, need);
    return DDNN_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  const int splits = outlayer_splits(B, K, dtype);
  const size_t np = np_of(N);
  float* dw_part = (float*)workspace;
  float* db_part = dw_part + (size_t)splits * K * np;
  rc = outlayer_fused(A, W, bias, labels, dZ, row_loss, dZprev, dw_part, db_part, B, K, N, act_prev, inv_denom, dtype,
                      st);
  if (rc) return rc;
  if (apply == DDNN_OUTLAYER_DEFER) {  // ddnn_output_layer_finish reduces, on any stream
    defer_record(workspace, splits);
    return DDNN_OK;
  }
  if (loss) {
    rc = loss_reduce(row_loss, loss, B, inv_denom, accumulate_loss, st);
    if (rc) return rc;
  }
  return reduce_dw(dw_part, db_part, splits, K, N, W_master, W_shadow, bias, dW_out, db_out, lr, apply, st);
}

int ddnn_output_layer_finish(const float* row_loss, float* loss, float* W_master, void* W_shadow, float* bias,
                             float* dW_out, float* db_out, float lr, int apply, int64_t B, int64_t K, int64_t N,
                             float inv_denom, int accumulate_loss, int dtype, void* workspace,
                             size_t workspace_bytes, void* stream) {
  int rc = check_dims("ddnn_output_layer_finish", B, K, N);
  if (rc) return rc;
  if (B == 0) return DDNN_OK;
  if (apply != 0 && apply != 1) {
    set_error("ddnn_output_layer_finish: apply = %d", apply);
    return DDNN_ERR_ARG;
  }
  if (!row_loss || !bias || (apply ? !W_master : (!dW_out || !db_out))) {
    set_error("ddnn_output_layer_finish: null pointer");
    return DDNN_ERR_ARG;
  }
  if (!outlayer_fused_ok(nullptr, K, N, dtype)) {
    set_error("ddnn_output_layer_finish: shape K=%lld N=%lld is not a fused output layer", (long long)K, (long long)N);
    return DDNN_ERR_UNSUPPORTED;
  }
  const size_t need = ddnn_output_layer_workspace_bytes(B, K, N, dtype);
  if (!workspace || workspace_bytes < need || !aligned16(workspace)) {
    set_error("ddnn_output_layer_finish: workspace %zu bytes < %zu required (or misaligned)", workspace_bytes, need);
    return DDNN_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  const int splits = defer_lookup(workspace, outlayer_splits(B, K, dtype));  // what the deferred call wrote
  if ((size_t)splits * ((size_t)K + 1) * np_of(N) * sizeof(float) > workspace_bytes) {
    set_error("ddnn_output_layer_finish: the deferred call on this workspace used %d splits, which do not fit %zu bytes",
              splits, workspace_bytes);
    return DDNN_ERR_WORKSPACE;
  }
  float* dw_part = (float*)workspace;
  float* db_part = dw_part + (size_t)splits * K * np_of(N);
  if (loss) {
    rc = loss_reduce(row_loss, loss, B, inv_denom, accumulate_loss, st);
    if (rc) return rc;
  }
  return reduce_dw(dw_part, db_part, splits, K, N, W_master, W_shadow, bias, dW_out, db_out, lr, apply, st);
}

// ---- fused last two layers (ddnn_tail2.cu) ----
// workspace: [dW2 partials S x K1 x N2][db2 S x N2][dW3 partials S x N2 x Np3][db3 S x Np3]
struct Tail2Ws {
  float *dw2, *db2, *dw3, *db3;
  size_t bytes;
};
static Tail2Ws tail2_ws(void* ws, int64_t B, int64_t K1, int64_t N2, int64_t N3, int splits = 0) {
  const size_t S = (size_t)(splits > 0 ? splits : tail2_splits(B)), np3 = np_of(N3);
  Tail2Ws w;
  w.dw2 = (float*)ws;
  w.db2 = w.dw2 + S * K1 * N2;
  w.dw3 = w.db2 + S * N2;
  w.db3 = w.dw3 + S * N2 * np3;
  w.bytes = ((S * ((size_t)K1 * N2 + N2 + (size_t)N2 * np3 + np3) * sizeof(float)) + 255) & ~(size_t)255;
  return w;
}

size_t ddnn_tail2_workspace_bytes(int64_t B, int64_t K1, int64_t N2, int64_t N3, int dtype) {
  if (B <= 0 || !tail2_ok(K1, N2, N3, dtype)) return 0;
  // sized for the largest split count of ANY batch of at most B rows: one buffer serves a ragged last batch too
  return tail2_ws(nullptr, B, K1, N2, N3, tail2_max_splits(B)).bytes;
}

int ddnn_tail2_train(const void* A1, const void* W2, const float* b2, const void* W3, const float* b3,
                     const int32_t* labels, void* dZ1, float* row_loss, int act1, int act2, int64_t B, int64_t K1,
                     int64_t N2, int64_t N3, float inv_denom, int dtype, void* workspace, size_t workspace_bytes,
                     void* stream) {
  int rc = check_dims("ddnn_tail2_train", B, K1, N2);
  if (rc) return rc;
  if (B == 0) return DDNN_OK;
  if (!tail2_ok(K1, N2, N3, dtype)) {
    set_error("ddnn_tail2_train: needs fp32, K1 <= 128, N2 <= 64 (both %% 4 == 0), N3 <= 16 (got %lld %lld %lld)",
              (long long)K1, (long long)N2, (long long)N3);
    return DDNN_ERR_UNSUPPORTED;
  }
  if (!A1 || !W2 || !b2 || !W3 || !b3 || !labels || !row_loss) {
    set_error("ddnn_tail2_train: null pointer");
    return DDNN_ERR_ARG;
  }
  if (!aligned16(A1) || (dZ1 && !aligned16(dZ1))) {
    set_error("ddnn_tail2_train: A1 / dZ1 must be 16-byte aligned");
    return DDNN_ERR_ARG;
  }
  for (int a : {act1, act2})
    if (a < DDNN_ACT_LINEAR || a >= DDNN_ACT_SOFTMAX) {
      set_error("ddnn_tail2_train: activation %d is not linear / relu / sigmoid", a);
      return DDNN_ERR_ACTIVATION;
    }
  const Tail2Ws w = tail2_ws(workspace, B, K1, N2, N3);
  if (!workspace || workspace_bytes < w.bytes || !aligned16(workspace)) {
    set_error("ddnn_tail2_train: workspace %zu bytes < %zu required (or misaligned)", workspace_bytes, w.bytes);
    return DDNN_ERR_WORKSPACE;
  }
  return tail2_launch((const float*)A1, (const float*)W2, b2, (const float*)W3, b3, labels, (float*)dZ1, row_loss,
                      w.dw2, w.db2, w.dw3, w.db3, B, K1, N2, N3, act1, act2, inv_denom,
                      (g_engine.load() != DDNN_ENGINE_SIMT && g_f32_tc.load() != 0) ? 1 : 0, (cudaStream_t)stream);
}

int ddnn_tail2_finish(const float* row_loss, float* loss, float* W2, float* b2, float* W3, float* b3, float* dW2_out,
                      float* db2_out, float* dW3_out, float* db3_out, float lr, int apply, int64_t B, int64_t K1,
                      int64_t N2, int64_t N3, float inv_denom, int accumulate_loss, int dtype, void* workspace,
                      size_t workspace_bytes, void* stream) {
  int rc = check_dims("ddnn_tail2_finish", B, K1, N2);
  if (rc) return rc;
  if (B == 0) return DDNN_OK;
  if (!tail2_ok(K1, N2, N3, dtype)) {
    set_error("ddnn_tail2_finish: unsupported shape / dtype");
    return DDNN_ERR_UNSUPPORTED;
  }
  if (!row_loss || (apply ? (!W2 || !b2 || !W3 || !b3) : (!dW2_out || !db2_out || !dW3_out || !db3_out))) {
    set_error("ddnn_tail2_finish: null pointer");
    return DDNN_ERR_ARG;
  }
  const Tail2Ws w = tail2_ws(workspace, B, K1, N2, N3);
  if (!workspace || workspace_bytes < w.bytes || !aligned16(workspace)) {
    set_error("ddnn_tail2_finish: workspace %zu bytes < %zu required (or misaligned)", workspace_bytes, w.bytes);
    return DDNN_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  const int S = tail2_splits(B);
  if (loss) {
    rc = loss_reduce(row_loss, loss, B, inv_denom, accumulate_loss, st);
    if (rc) return rc;
  }
  rc = reduce_dw(w.dw3, w.db3, S, N2, N3, W3, nullptr, b3, dW3_out, db3_out, lr, apply ? 1 : 0, st);
  if (rc) return rc;
  return reduce_dw(w.dw2, w.db2, S, K1, N2, W2, nullptr, b2, dW2_out, db2_out, lr, apply ? 1 : 0, st);
}

int ddnn_cast_f32_to_bf16(const float* in, void* out, int64_t n, void* stream) {
  if (n <= 0) return DDNN_OK;
  if (!in || !out || !aligned16(in) || (reinterpret_cast<uintptr_t>(out) & 7u)) {
    set_error("ddnn_cast_f32_to_bf16: null or misaligned pointer");
    return DDNN_ERR_ARG;
  }
  return cast_f32_bf16(in, (bf16*)out, n, (cudaStream_t)stream);
}
int ddnn_cast_bf16_to_f32(const void* in, float* out, int64_t n, void* stream) {
  if (n <= 0) return DDNN_OK;
  if (!in || !out || !aligned16(out) || (reinterpret_cast<uintptr_t>(in) & 7u)) {
    set_error("ddnn_cast_bf16_to_f32: null or misaligned pointer");
    return DDNN_ERR_ARG;
  }
  return cast_bf16_f32((const bf16*)in, out, n, (cudaStream_t)stream);
}

}  // extern "C"
