/*
 * ddnn.h -- C ABI of libddnn.so: the docker-dist-nn layer hot path on B200 (sm_100a).
 *
 * The reference has no FFI: its hot path is three Python methods of one class,
 * /root/reference/src/grpc_node.py:16-97 (Layer.__init__ / activate / process_matrix),
 * which call numpy.  Each entry point below names the numpy call site it replaces.
 * Backward, softmax-CE and SGD have no reference counterpart;
 * they are specified by BASELINE.json's north_star and derived from the forward.
 *
 * Conventions
 *  - every pointer is a DEVICE pointer owned by the caller (incl. workspace); the
 *    library allocates no device memory.  Its state: a mutex-guarded cache of TMA
 *    descriptors keyed by (pointer, shape), the per-device kernel attributes it has set,
 *    and the process-wide ddnn_set_* switches below -- test / profiling knobs that select
 *    between equivalent kernels; set them before serving, not while calls are in flight
 *    on other threads (the defaults never need changing);
 *  - matrices are row-major, (batch, features); weights are (K = in, N = out) like
 *    grpc_node.py:51 (`np.array([...]).T`);
 *  - `stream` is a cudaStream_t passed as void*; all work is stream-ordered, nothing
 *    synchronises;
 *  - return value: 0 = ok, otherwise a ddnn_status; ddnn_last_error() gives the text
 *    for the calling thread.  Nothing throws across the ABI.
 *  - B == 0 is legal everywhere and does nothing (grpc_node.py:108-111 returns an
 *    empty Matrix for an empty request).
 */
#ifndef DDNN_H_
#define DDNN_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  DDNN_OK = 0,
  DDNN_ERR_DIM = 1,          /* shape mismatch -> Python raises ValueError (grpc_node.py:83-84) */
  DDNN_ERR_ACTIVATION = 2,   /* activation not supported by this entry point */
  DDNN_ERR_CUDA = 3,         /* a CUDA call failed; text in ddnn_last_error() */
  DDNN_ERR_ARG = 4,          /* null / misaligned pointer, bad enum */
  DDNN_ERR_WORKSPACE = 5,    /* workspace smaller than ddnn_*_workspace_bytes() */
  DDNN_ERR_UNSUPPORTED = 6   /* dtype/shape combination has no kernel */
} ddnn_status;

/* grpc_node.py:62-73.  Any string that is not relu/sigmoid/softmax is the identity. */
typedef enum { DDNN_ACT_LINEAR = 0, DDNN_ACT_RELU = 1, DDNN_ACT_SIGMOID = 2, DDNN_ACT_SOFTMAX = 3 } ddnn_act;

/* Element type of activations / weight operands.  Accumulation is always fp32. */
typedef enum { DDNN_F32 = 0, DDNN_BF16 = 1 } ddnn_dtype;

/* Which GEMM engine to use.  AUTO: tcgen05 when dtype is bf16 and the shape meets the
 * TMA alignment rules (feature widths multiples of 8), else the SIMT kernels. */
typedef enum { DDNN_ENGINE_AUTO = 0, DDNN_ENGINE_SIMT = 1, DDNN_ENGINE_TCGEN05 = 2 } ddnn_engine;

int ddnn_version(void);
const char* ddnn_last_error(void);

/* Maps an activation string the way Layer.activate does (grpc_node.py:62-73). */
int ddnn_activation_from_name(const char* name);

/* Force an engine for subsequent calls of this process (tests / profiling only: process-wide, see above). */
int ddnn_set_engine(int engine);

/*
 * A[B,N] = act(X[B,K] . W[K,N] + bias[N])
 * Replaces `np.dot(current_matrix, params["weights"]) + params["biases"]`
 * (grpc_node.py:87) and `self.activate(raw_output, ...)` (grpc_node.py:90, :62-73).
 * X, W have element type `dtype`; A has `out_dtype`; bias is fp32.
 * act == SOFTMAX applies the row softmax of grpc_node.py:68-71 in a second launch.
 */
int ddnn_dense_forward(const void* X, const void* W, const float* bias, void* A,
                       int64_t B, int64_t K, int64_t N,
                       int act, int dtype, int out_dtype, void* stream);

/* P[B,N] = row softmax of Z (grpc_node.py:68-71); in place allowed.  fp32 in/out. */
int ddnn_softmax_rows(const float* Z, float* P, int64_t B, int64_t N, void* stream);

/* Elementwise activation, replaces Layer.activate for relu/sigmoid/linear/softmax on an
 * arbitrary fp32 matrix (grpc_node.py:62-73).  In place allowed. */
int ddnn_activate(const float* V, float* out, int64_t B, int64_t N, int act, void* stream);

/*
 * Softmax cross-entropy at the output node.  No reference counterpart; semantics of
 * nn.CrossEntropyLoss(mean) at /root/reference/scripts/generate_mnist_pytorch.py:37.
 *   p = softmax(Z);  row_loss[i] = -log p[i,y_i];  dZ = (p - onehot(y)) * inv_denom
 *   *loss (+)= inv_denom * sum_i row_loss[i]      (accumulate != 0 adds to *loss)
 * Z fp32 [B,N]; labels int32 [B]; P (fp32, nullable) receives the probabilities;
 * dZ has element type dz_dtype; row_loss is fp32 [B] scratch/output owned by the caller;
 * loss is one fp32 on the device.
 */
int ddnn_softmax_ce(const float* Z, const int32_t* labels, float* P, void* dZ,
                    float* row_loss, float* loss, int64_t B, int64_t N,
                    float inv_denom, int dz_dtype, int accumulate, void* stream);

/*
 * dZprev[B,K] = (dZ[B,N] . W[K,N]^T) * f'(Aprev[B,K])
 * f' from the activation of the PREVIOUS layer, evaluated on its stored output:
 * relu -> 1[A>0], sigmoid -> A(1-A), linear -> 1.  SOFTMAX is rejected (hidden softmax
 * needs the row Jacobian; DDNN_ERR_ACTIVATION).  No reference counterpart.
 */
int ddnn_dense_backward_dx(const void* dZ, const void* W, const void* Aprev, void* dZprev,
                           int64_t B, int64_t K, int64_t N,
                           int act_prev, int dtype, void* stream);

/*
 * dW[K,N] = X[B,K]^T . dZ[B,N];  db[N] = column sums of dZ;  then either
 *   apply != 0 :  W_master -= lr*dW ; bias -= lr*db ; W_shadow = bf16(W_master) if given
 *   apply == 0 :  dW_out = dW, db_out = db            (fp32; for gradient checks)
 * W_master is fp32 [K,N] (in f32 mode it IS the weight the forward reads); W_shadow is
 * the bf16 copy the bf16 forward/dX read (nullable).  Split-K partial sums go through
 * `workspace` (fp32, >= ddnn_dw_workspace_bytes) and are reduced in a fixed order, so
 * results are run-to-run deterministic.  No reference counterpart.
 */
int ddnn_dense_backward_dw_sgd(const void* X, const void* dZ,
                               float* W_master, void* W_shadow, float* bias,
                               float* dW_out, float* db_out,
                               float lr, int apply,
                               int64_t B, int64_t K, int64_t N, int dtype,
                               void* workspace, size_t workspace_bytes, void* stream);

size_t ddnn_dw_workspace_bytes(int64_t B, int64_t K, int64_t N, int dtype);

/*
 * The weight gradient of one ROW CHUNK of the batch, accumulated in caller-owned fp32 buffers:
 *   first != 0 :  dW_acc  = X^T . dZ ;  db_acc  = column sums of dZ
 *   first == 0 :  dW_acc += X^T . dZ ;  db_acc += column sums of dZ      (one fp32 rounding per element and call)
 * followed, after the last chunk, by ddnn_sgd_apply:  W_master -= lr * dW_acc ; bias -= lr * db_acc ; W_shadow =
 * bf16(W_master) if given.  Same kernels, workspace rule and determinism as ddnn_dense_backward_dw_sgd (the
 * accumulation is its update epilogue with the accumulator in W_master's place).  Two uses: a pipeline stage
 * computes the gradient of the micro-batches it has finished while it waits for its neighbours (zero-bubble
 * order), and a reduction over a very long batch is cut into chunks whose operands stay in L2 (the 16384 x 16384
 * x 65536 GEMM in one piece re-reads its operands from HBM 25 times over).  The chunking is part of the
 * arithmetic: results are bit-identical for equal chunk boundaries, and differ from the one-piece GEMM by fp32
 * rounding of the partial sums.  No reference counterpart (grpc_node.py has no backward pass).
 */
int ddnn_dense_backward_dw_accumulate(const void* X, const void* dZ, float* dW_acc, float* db_acc, int first,
                                      int64_t B, int64_t K, int64_t N, int dtype,
                                      void* workspace, size_t workspace_bytes, void* stream);
int ddnn_sgd_apply(float* W_master, void* W_shadow, float* bias, const float* dW, const float* db, float lr,
                   int64_t K, int64_t N, void* stream);

/*
 * Same computation and arguments as ddnn_dense_backward_dw_sgd, minus the workspace: the
 * split-K scratch is an array built into the library (DDNN_STATIC_SCRATCH_BYTES per GPU, held
 * from module load to process exit).  This entry point has no counterpart in the reference.
 * It exists for hosts that cannot give the library device memory of their own: a process
 * whose framework has already claimed the GPU's memory for its private pool (a plain
 * cudaMalloc then fails), an FFI stub that can only forward pointers it was handed, or a
 * request path on which cudaMalloc/cudaFree are banned because they stall the whole device.
 *  - Nothing is allocated or freed, here or in anything it calls.
 *  - Layers whose split-K plan needs more than DDNN_STATIC_SCRATCH_BYTES
 *    (ddnn_dw_workspace_bytes tells) are refused with DDNN_ERR_WORKSPACE; nothing is launched.
 *  - The first call on each device looks up the array's device address (one-time set-up, done
 *    with that device current).  Make that first call before starting any stream capture.
 *  - All calls share the one array and it keeps the last call's contents; no part of it needs
 *    to be zero on entry.  Calls that may execute concurrently on the same GPU (different
 *    streams) race on it: keep them on one stream or order them with events.
 */
#define DDNN_STATIC_SCRATCH_BYTES (8u * 1024u * 1024u)
int ddnn_dense_backward_dw_sgd_static(const void* X, const void* dZ,
                                      float* W_master, void* W_shadow, float* bias,
                                      float* dW_out, float* db_out,
                                      float lr, int apply,
                                      int64_t B, int64_t K, int64_t N, int dtype, void* stream);

/*
 * Whole training step of a skinny OUTPUT layer in one pass over its input A[B,K] (N <= 16 classes,
 * K <= 256, rows of A 16-byte aligned; otherwise DDNN_ERR_UNSUPPORTED -- use the separate entry points):
 *   logits = A.W + bias ; softmax-CE (row_loss, *loss, dZ = (p - onehot) * inv_denom)
 *   dZprev = (dZ.W^T) * f'(A)  with f' of the FEEDING layer's activation (nullable: no input gradient)
 *   dW = A^T.dZ, db = sum dZ   then SGD (apply) or fp32 store, like ddnn_dense_backward_dw_sgd
 * W is the operand the GEMMs read (bf16 shadow in bf16 mode, the fp32 weight in f32 mode); W_master/W_shadow
 * /bias are what the update writes.  dZ is optional.  No reference counterpart (grpc_node.py has no loss).
 */
int ddnn_output_layer_train(const void* A, const void* W, const int32_t* labels, void* dZ,
                            float* row_loss, float* loss, void* dZprev,
                            float* W_master, void* W_shadow, float* bias, float* dW_out, float* db_out,
                            float lr, int apply, int64_t B, int64_t K, int64_t N, int act_prev,
                            float inv_denom, int accumulate_loss, int dtype,
                            void* workspace, size_t workspace_bytes, void* stream);

/*
 * Two-call form: pass apply = DDNN_OUTLAYER_DEFER to ddnn_output_layer_train and it only runs the fused kernel
 * (row_loss, dZ, dZprev and the split partials in `workspace` are final; *loss and the weights are untouched),
 * then ddnn_output_layer_finish reduces the loss and the partials and applies SGD / stores the gradients -- on
 * any stream that is ordered after the first call.  Lets the host take the reduction off the critical path
 * (dX of the layer below only needs dZprev).  Same B, K, N, dtype and workspace in both calls; the workspace
 * must not be reused in between.  No reference counterpart.
 */
#define DDNN_OUTLAYER_DEFER 2
int ddnn_output_layer_finish(const float* row_loss, float* loss,
                             float* W_master, void* W_shadow, float* bias, float* dW_out, float* db_out,
                             float lr, int apply, int64_t B, int64_t K, int64_t N,
                             float inv_denom, int accumulate_loss, int dtype,
                             void* workspace, size_t workspace_bytes, void* stream);
size_t ddnn_output_layer_workspace_bytes(int64_t B, int64_t K, int64_t N, int dtype);

/*
 * Fused training step of the last TWO layers of a small fp32 MLP (hidden layer K1 -> N2 with activation act2,
 * output layer N2 -> N3 with softmax-CE): K1 <= 128 and N2 <= 64 (both multiples of 4), N3 <= 16, fp32, A1 and
 * dZ1 16-byte aligned; otherwise DDNN_ERR_UNSUPPORTED -- use the per-layer entry points.  One kernel reads
 * A1[B,K1] (the output of the layer whose activation is act1) once and produces row_loss, dZ1 = gradient w.r.t.
 * that layer's pre-activation (nullable) and the split partials of dW2, db2, dW3, db3 in `workspace`; A2, the
 * logits and dZ2 / dZ3 never reach HBM.  ddnn_tail2_finish (any stream ordered after the first call, same
 * workspace) reduces the loss and the partials and applies SGD (apply = 1: W2, b2, W3, b3 updated in place) or
 * stores the four gradients (apply = 0).  Arithmetic, loss and update rule are those of ddnn_dense_forward /
 * ddnn_softmax_ce / ddnn_dense_backward_* applied layer by layer.  No reference counterpart (grpc_node.py runs
 * one process_matrix per layer and has no backward pass).
 */
size_t ddnn_tail2_workspace_bytes(int64_t B, int64_t K1, int64_t N2, int64_t N3, int dtype);
int ddnn_tail2_train(const void* A1, const void* W2, const float* b2, const void* W3, const float* b3,
                     const int32_t* labels, void* dZ1, float* row_loss, int act1, int act2,
                     int64_t B, int64_t K1, int64_t N2, int64_t N3, float inv_denom, int dtype,
                     void* workspace, size_t workspace_bytes, void* stream);
int ddnn_tail2_finish(const float* row_loss, float* loss, float* W2, float* b2, float* W3, float* b3,
                      float* dW2_out, float* db2_out, float* dW3_out, float* db3_out,
                      float lr, int apply, int64_t B, int64_t K1, int64_t N2, int64_t N3,
                      float inv_denom, int accumulate_loss, int dtype,
                      void* workspace, size_t workspace_bytes, void* stream);

/* out[i] = bf16(in[i]) / fp32(in[i]) converters for staging host-provided fp32 data. */
int ddnn_cast_f32_to_bf16(const float* in, void* out, int64_t n, void* stream);
int ddnn_cast_bf16_to_f32(const void* in, float* out, int64_t n, void* stream);

/* 1 (default) = launch the kernels of a train step with programmatic dependent launch (the next kernel's launch
 * and prologue overlap the tail of the previous one; every such kernel waits for its predecessor's completion
 * before its first global-memory access), 0 = plain stream order.  No reference counterpart. */
int ddnn_set_pdl(int on);

/* bf16 tcgen05 GEMMs: 1 (default) = use the CTA-pair kernel (tcgen05.mma.cta_group::2, 256x256 tile
 * per 2-CTA cluster) when a problem has >= 2 waves of such tiles, 0 = always the single-CTA kernel. */
int ddnn_set_cta_pairs(int on);

/* fp32 path inner product for 16-byte-aligned forward / dX / dW GEMMs, all within the 1e-5 bound:
 *   0 = FFMA (cp.async SIMT kernels);
 *   1 = error-compensated 3xTF32 on mma.sync fragments inside the SIMT tiling (no faster than FFMA, kept for
 *       comparison);
 *   2 = error-compensated 3xTF32 on tcgen05.mma.kind::tf32 (TMA-fed, operands split hi/lo in shared memory by
 *       converter warps, main and correction accumulators in TMEM), widths >= 32 and % 4 == 0.
 * The default is DDNN_F32_TC_DEFAULT; other values return DDNN_ERR_ARG. */
#define DDNN_F32_TC_DEFAULT 2
int ddnn_set_f32_tensor_cores(int mode);

/* fp32 GEMMs on tcgen05: 1 (default) = when a one-wave plan exists, the reduction is split over a thread-block
 * cluster and the partial tiles are summed inside the kernel through distributed shared memory -- dW: the same kernel
 * applies SGD / stores the gradients (one launch, no partial sums in HBM); forward / dX with few output tiles and a
 * long reduction: each CTA walks 1 / S of the k-blocks, the epilogue runs on the sum.  0 = dW partials in the
 * workspace + reducer launch, forward / dX un-split.  Test /
 * profiling switch, process-wide.  No reference counterpart. */
int ddnn_set_cluster_split_k(int on);

/* Leaves `sms` SMs free of the persistent tcgen05 GEMM CTAs (default 0).  The pipeline sets this
 * so that NCCL's send/recv kernels, which cannot co-reside with a 210 KB-shared-memory GEMM CTA,
 * always find an SM and run beside the GEMM instead of between GEMMs.  No reference counterpart. */
int ddnn_set_sm_reserve(int sms);

/*
 * Peer-memory transport of the layer-per-GPU pipeline (replaces the gRPC hop of
 * /root/reference/src/grpc_node.py:120-135; no reference counterpart as an API).
 * ddnn_ipc_export : CUDA-IPC handle (64 bytes) of the allocation containing `ptr` + ptr's offset
 *                   in it; the bytes are sent to the neighbour process by any host channel.
 * ddnn_ipc_import : maps that allocation into this process (once per handle) and returns the
 *                   peer pointer.  It is a normal device pointer for every entry point above: give
 *                   it as the OUTPUT of ddnn_dense_forward / ddnn_dense_backward_dx and the GEMM
 *                   epilogue stores straight into the neighbour GPU's memory over NVLink.
 * ddnn_flag_write : stream-ordered 4-byte store (after a system-scope fence) to a local or peer flag.
 * ddnn_flag_wait  : makes `stream` wait until *flag >= value (cuStreamWaitValue32; flag is local).
 */
int ddnn_ipc_export(const void* ptr, void* handle64, int64_t* offset);
int ddnn_ipc_import(const void* handle64, int64_t offset, void** ptr);
int ddnn_flag_write(int32_t* flag, int32_t value, void* stream);
int ddnn_flag_wait(const int32_t* flag, int32_t value, void* stream);

/*
 * Pipeline signalling folded into a GEMM (no reference counterpart; replaces the per-hop RPC completion of
 * /root/reference/src/grpc_node.py:130-135).  With a separate ddnn_flag_wait before and ddnn_flag_write after every
 * micro-batch GEMM the kernels of a stage no longer follow each other directly in the stream, so programmatic
 * dependent launch cannot overlap them; the *_signalled entry points move both ends into the kernel:
 *   wait_flag  (nullable): the kernel's operand loader polls  *wait_flag >= wait_value  (acquire, system scope)
 *              before its first load of X / dZ -- the operand another GPU stores into this GPU's memory;
 *   done_flag  (nullable): the last CTA to finish stores done_value to *done_flag (local or peer-mapped) after a
 *              system-scope fence, i.e. after every tile of the output is visible to the flag's reader;
 *   counter    : 4 bytes of device memory owned by the caller, zero before the first use, shared only by kernels
 *              of ONE stream (each GEMM leaves it at zero); required when done_flag is given.
 * On shapes / dtypes that do not run on the tcgen05 kernels the same effect is produced with ddnn_flag_wait, the
 * plain kernel and ddnn_flag_write.  sig == NULL makes them identical to the plain entry points.  The polling
 * kernel occupies the GPU until the flag arrives: the producer must not depend on later work of this stream, and
 * a producer on ANOTHER STREAM OF THE SAME GPU must use kernels that are already loaded (the first launch of a
 * kernel loads its module lazily, which can wait for the device to go idle).
 */
typedef struct ddnn_signals {
  const int32_t* wait_flag;
  int32_t wait_value;
  int32_t* done_flag;
  int32_t done_value;
  uint32_t* counter;
} ddnn_signals;
int ddnn_dense_forward_signalled(const void* X, const void* W, const float* bias, void* A,
                                 int64_t B, int64_t K, int64_t N, int act, int dtype, int out_dtype,
                                 const ddnn_signals* sig, void* stream);
int ddnn_dense_backward_dx_signalled(const void* dZ, const void* W, const void* Aprev, void* dZprev,
                                     int64_t B, int64_t K, int64_t N, int act_prev, int dtype,
                                     const ddnn_signals* sig, void* stream);

/* Count of kernels this library has launched in this process (bench.py gpu_launches). */
int64_t ddnn_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* DDNN_H_ */
