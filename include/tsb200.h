/*
 * tsb200.h -- C ABI of the B200-native engine behind TorchSystem's training handlers.
 *
 * The reference (mapache-software/torch-system 1.4.0) has no FFI: its hot path is reached through
 * Python structural protocols (examples/tinysys/tinysys/domain.py:10-40). This header is the
 * boundary a binding for that path would target: plain pointers and sizes, no torch types.
 * Each entry point names the reference call it replaces (paths relative to the reference root).
 *
 * Conventions
 *   - every function returns 0 on success, a negative TS_E_* code otherwise; ts_last_error()
 *     returns a thread-local, human-readable message for the last failure on this thread;
 *   - all data pointers are DEVICE pointers on the device the handle was created for, unless
 *     marked HOST; `stream` is a cudaStream_t passed as void* (NULL = legacy default stream);
 *   - nothing here synchronises the stream except where stated; nothing falls back to the CPU;
 *   - the library never allocates device memory: the caller sizes a workspace with
 *     ts_mlp_workspace_bytes() and hands it over with ts_mlp_set_workspace().
 *
 * Parameter layout: all parameters live in ONE flat fp32 buffer in torch order
 * (layer0.weight [out,in] row-major, layer0.bias, layer1.weight, ...; examples/tinysys/modules/mlp.py:8-11),
 * each tensor starting at the element offset reported by ts_mlp_param_layout() (4 KiB aligned).
 * Gradients and the Adam moments use the same layout in buffers of the same size.
 */
#ifndef TSB200_H
#define TSB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TS_ABI_VERSION 2

/* status codes */
#define TS_OK 0
#define TS_E_INVALID (-1)   /* bad argument (NULL, size, enum) */
#define TS_E_CUDA (-2)      /* a CUDA runtime/driver call failed; message has the CUDA error */
#define TS_E_STATE (-3)     /* call order: parameters / optimizer / workspace not bound */
#define TS_E_UNSUPPORTED (-4)
#define TS_E_TOO_LARGE (-5) /* size beyond a stated maximum */
#define TS_E_TIMEOUT (-6)   /* data parallel: a peer rank did not arrive within the configured time */

/* activation of the hidden layers (examples/tinysys/main.py:60 uses ReLU; tests/test_storage.py:38 GELU) */
#define TS_ACT_RELU 0
#define TS_ACT_GELU 1

/* arithmetic mode of the dense layers */
#define TS_PREC_F32 0  /* fp32 SIMT everywhere: matches the CPU reference to 1e-5 */
#define TS_PREC_BF16 1 /* bf16 operands on tcgen05 tensor cores, fp32 accumulate in TMEM: 1e-2 */

/* element type of the `inputs` buffer */
#define TS_DT_F32 0
#define TS_DT_BF16 1
#define TS_DT_U8 2 /* raw pixels, mapped through the table given to ts_mlp_set_input_lut */

/* optimizer (examples/tinysys/main.py:62 ships Adam(lr=1e-3); SGD is the north-star addition) */
#define TS_OPT_SGD 0
#define TS_OPT_ADAM 1

typedef struct ts_mlp ts_mlp; /* opaque aggregate handle */

/* Hyper-parameters, torch.optim semantics (torch/optim/adam.py _single_tensor_adam, sgd.py _single_tensor_sgd);
 * `maximize` is not supported. Zero-initialise the struct: every field after `eps` defaults to "off". */
typedef struct ts_opt_hyper {
    double lr;
    double beta1, beta2;     /* Adam only */
    double eps;              /* Adam only */
    double weight_decay;     /* SGD and Adam: grad += weight_decay * param (or decoupled, see below) */
    double momentum;         /* SGD: momentum buffer lives in the `exp_avg` array given to ts_opt_bind */
    double dampening;        /* SGD */
    int32_t nesterov;        /* SGD */
    int32_t amsgrad;         /* Adam: needs the `max_exp_avg_sq` array */
    int32_t decoupled_weight_decay; /* Adam: param *= 1 - lr * weight_decay instead (AdamW) */
    int32_t reserved;
} ts_opt_hyper;

/* One record per fit/evaluate call, written by the fused head kernel (16 bytes, device memory).
 * `loss` is the batch-mean cross-entropy that Classifier.fit/evaluate return
 * (examples/tinysys/tinysys/classifier.py:35,39); `n_correct` is what
 * MulticlassAccuracy.update would count from (argmax, targets) (tinysys/metrics.py:15). */
typedef struct ts_batch_record {
    float loss;
    int32_t n_correct;
    int32_t n_rows;
    int32_t reserved;
} ts_batch_record;

/* Running metric state (device memory, 32 bytes): torcheval Mean + micro MulticlassAccuracy
 * (examples/tinysys/tinysys/metrics.py:8-27). */
typedef struct ts_metrics_state {
    double loss_sum;   /* sum of per-batch mean losses, fp64 like torcheval.Mean */
    int64_t n_batches; /* Mean's weight: one per update */
    int64_t n_correct;
    int64_t n_total;
} ts_metrics_state;

int32_t ts_abi_version(void);
const char *ts_last_error(void);

/* ---- aggregate construction: replaces Classifier.__init__ + .to(device)
 *      (examples/tinysys/tinysys/classifier.py:10-16, tinysys/services/compilation.py:23-34) ---- */

/* dims[n_dims] = {in, hidden..., classes}: (n_dims-2) x [Linear, Dropout(p), act] then Linear. */
int32_t ts_mlp_create(ts_mlp **out, const int32_t *dims /*HOST*/, int32_t n_dims, int32_t activation,
                      float dropout_p, int32_t precision, int32_t device);
int32_t ts_mlp_destroy(ts_mlp *m);

/* TS_OK, or TS_E_INVALID once a batch has held a target outside [0, classes) -- e.g. CrossEntropyLoss's default
 * ignore_index -100 (examples/tinysys/main.py:61), which this engine does not implement: such a batch's loss is NaN, the
 * head kernel records the row in a host-visible word, and every later ts_fit_step / ts_eval_step (and this call, which a
 * host can make after any synchronisation, as B200Metrics.compute does) fails until ts_mlp_clear_status. */
int32_t ts_mlp_status(const ts_mlp *m);
int32_t ts_mlp_clear_status(ts_mlp *m);

/* offsets[2*(n_dims-1)] (HOST, in elements) of each parameter tensor in the flat buffers; *total = buffer length. */
int32_t ts_mlp_param_layout(const ts_mlp *m, int64_t *offsets /*HOST*/, int64_t *total /*HOST*/);

/* uint8 inputs (TS_DT_U8): every byte is replaced by lut256[byte] on its way into the first layer -- the per-sample
 * ToTensor() + Normalize((0.1307,), (0.3081,)) of the reference's dataset (examples/tinysys/datasets/mnist.py:7) moved
 * onto the device, so a batch costs 1 byte per pixel of PCIe / HBM traffic instead of 4. lut256: HOST array of 256 floats
 * (fill it with the host-side arithmetic you want reproduced, e.g. (b / 255 - mean) / std in fp32). */
int32_t ts_mlp_set_input_lut(ts_mlp *m, const float *lut256 /*HOST*/);

/* Alias the caller's parameter and gradient storage (fp32 master weights stay where torch sees them). */
int32_t ts_mlp_bind_params(ts_mlp *m, float *params, float *grads);

/* replaces torch.optim.Adam/AdamW/SGD construction (examples/tinysys/main.py:62). State arrays are flat like the
 * parameters: `exp_avg` (Adam's first moment, or SGD's momentum buffer when momentum != 0), `exp_avg_sq` (Adam),
 * `max_exp_avg_sq` (Adam with amsgrad); the ones a configuration does not use may be NULL.
 * `step` is the number of optimizer steps already taken (resume; SGD with momentum copies the first gradient into
 * the momentum buffer when step == 0, as torch does for a missing buffer). */
int32_t ts_opt_bind(ts_mlp *m, int32_t kind, const ts_opt_hyper *hyper /*HOST*/, float *exp_avg, float *exp_avg_sq,
                    float *max_exp_avg_sq, int64_t step);
int64_t ts_opt_step_count(const ts_mlp *m);

int64_t ts_mlp_workspace_bytes(const ts_mlp *m, int64_t max_batch);
int32_t ts_mlp_set_workspace(ts_mlp *m, void *workspace, int64_t bytes, int64_t max_batch);

/* Re-derive the bf16 shadow weights from the fp32 masters (after load_state_dict / manual edits). */
int32_t ts_mlp_sync_weights(ts_mlp *m, void *stream);

/* ---- hot path ---- */

/* Classifier.fit (examples/tinysys/tinysys/classifier.py:29-35): zero_grad, forward with dropout,
 * mean cross-entropy, backward, optimizer step; returns the PRE-update argmax and loss.
 *   inputs       [B, dims[0]] row-major, dtype `inputs_dtype` (the handler's Flatten is a view)
 *   targets      [B] int64 in [0, classes)
 *   masks        HOST array of (n_dims-2) device pointers to uint8 keep-masks [B, hidden_l] (values 0/1), or
 *                NULL to draw the masks in-kernel from a counter-based hash keyed by (`seed`, optimizer step, layer)
 *   predictions  [B] int64 out
 *   record       one ts_batch_record out (its .loss is the returned loss tensor)
 */
/* Precondition on `inputs` in TS_PREC_BF16 mode: fp32 / uint8 batches are converted by a kernel that reads `inputs` BEFORE it
 * waits for the kernel in front of it on `stream` (it overlaps that kernel, normally this library's previous update). The
 * kernel that PRODUCED `inputs` must therefore not be the one immediately in front of this call on `stream`: have at least
 * one other kernel or an event / stream synchronisation in between (loaders that prepare an epoch up front, or copy on
 * another stream and hand over with an event, satisfy this). */
int32_t ts_fit_step(ts_mlp *m, const void *inputs, int32_t inputs_dtype, const int64_t *targets, int64_t batch,
                    const uint8_t *const *masks /*HOST*/, uint64_t seed, int64_t *predictions,
                    ts_batch_record *record, void *stream);

/* Data-parallel split of the same call: ts_fit_backward leaves the local mean gradient in `grads`
 * (caller all-reduces it), ts_opt_apply consumes `grads * grad_scale` (1/world) and updates. */
int32_t ts_fit_backward(ts_mlp *m, const void *inputs, int32_t inputs_dtype, const int64_t *targets, int64_t batch,
                        const uint8_t *const *masks /*HOST*/, uint64_t seed, int64_t *predictions,
                        ts_batch_record *record, void *stream);
int32_t ts_opt_apply(ts_mlp *m, float grad_scale, void *stream);
/* ts_opt_apply for the parameter tensors [first_tensor, first_tensor + n_tensors) only (torch order W0, b0, W1, ...), so
 * the update of a bucket whose all-reduce has finished can run while the next bucket is still being reduced. The first
 * range of a step passes advance_step = 1 (optimizer step counter, Adam bias correction), the others 0. */
int32_t ts_opt_apply_range(ts_mlp *m, float grad_scale, int32_t first_tensor, int32_t n_tensors, int32_t advance_step, void *stream);

/* ts_fit_backward for callers that overlap the gradient all-reduce with the rest of the backward pass (no counterpart
 * in the reference, which is single-process; BASELINE north star: "NCCL allreduces gradients ... overlapped with
 * backward"). `layer_events[l]` (HOST array of n_layers cudaEvent_t) is recorded on `stream` as soon as the kernels
 * producing layer l's gradients are enqueued -- last layer first. `tensor_final[t]` (HOST out, 2*n_layers: W0, b0, W1,
 * ...) says whether gradient tensor t is already complete in `grads` at that event (the caller may all-reduce it on
 * another stream that waits for the event) or only when the whole call has run (small tensors kept as partials).
 * ts_dp_reserve_sms makes the backward GEMMs of later calls leave `n_sms` SMs free for that collective's kernels. */
int32_t ts_fit_backward_overlap(ts_mlp *m, const void *inputs, int32_t inputs_dtype, const int64_t *targets, int64_t batch,
                                const uint8_t *const *masks /*HOST*/, uint64_t seed, int64_t *predictions,
                                ts_batch_record *record, void *stream, void *const *layer_events /*HOST*/,
                                int32_t *tensor_final /*HOST*/);
int32_t ts_dp_reserve_sms(ts_mlp *m, int32_t n_sms);

#define TS_DP_FLAG_WORDS 64
/* Data parallel without a separate collective (no counterpart in the reference, which is single-process): after
 * ts_dp_bind, ts_fit_step's optimizer kernel sums this rank's gradient partials into `grad_bufs[parity][rank]`, signals
 * its peers, reads the same chunk of every rank's buffer over the peer mappings (NVLink / NVSwitch), sums in rank order,
 * scales by 1/world and updates -- all ranks end a step with bit-identical weights.
 *   grad_bufs  HOST array of 2*world device pointers, buffer b of rank r at [b*world + r]; each as long as the flat
 *              parameter buffer (ts_mlp_param_layout), mapped on THIS device (e.g. torch symmetric memory, CUDA IPC)
 *   flag_bufs  HOST array of world device pointers to each rank's flag array: TS_DP_FLAG_WORDS uint32, zero-initialised
 *              (word q < world: last step rank q has published; words 8 and 9: arrival tickets; words 16..23: time trace)
 * Every rank must call ts_fit_step the same number of times. The kernel's CTAs wait for each other and for the peers, so
 * its whole grid (at most 2 CTAs per SM, 4 for ts_dp_bind_nvls) must be resident at once: do not run other work that
 * fills the GPU's SMs concurrently with a data-parallel ts_fit_step. A wait for a peer is bounded by wall-clock time
 * (ts_dp_set_timeout, default 600 s or TSB200_DP_TIMEOUT_S): when it expires the kernel skips its update and sets a
 * host-visible error word instead of trapping; the next ts_fit_step (or ts_dp_status) then returns TS_E_TIMEOUT. */
int32_t ts_dp_bind(ts_mlp *m, int32_t rank, int32_t world, float *const *grad_bufs /*HOST*/, uint32_t *const *flag_bufs /*HOST*/);
/* Optional low-latency form of that exchange for the one-hidden-layer path at batches of 2,048 rows or more (no counterpart in
 * the reference): the kernel that computes dW1 sends its share of the gradient as 16-byte packets {three floats, step number}
 * to the rank that owns the share, which adds the ranks in rank order and sends the sum to everybody; receivers poll their OWN
 * memory for the step number. No fence, flag or remote read on the critical path. Call after ts_dp_bind, on every rank.
 *   inbox_bufs   HOST array of world device pointers: rank r's inbox, TS_DP_LL_INBOX_BYTES bytes, zero-initialised, 16-byte
 *                aligned, mapped on THIS device (e.g. torch symmetric memory)
 * TSB200_DP_LL=0 (environment, read once) keeps the exchange of ts_dp_bind. */
#define TS_DP_LL_INBOX_BYTES 28311552
int32_t ts_dp_bind_ll(ts_mlp *m, void *const *inbox_bufs /*HOST*/, int64_t inbox_bytes);
/* Bound, in seconds, on how long a data-parallel kernel waits for a peer rank (no counterpart in the reference). */
int32_t ts_dp_set_timeout(ts_mlp *m, double seconds);
/* TS_OK, or TS_E_TIMEOUT once a data-parallel kernel has given up waiting for a peer (no counterpart in the reference). */
int32_t ts_dp_status(const ts_mlp *m);

/* The same for large models, with the reduction done inside the NVSwitch (no counterpart in the reference): gradient,
 * parameter and (bf16 mode) shadow buffers live in symmetric memory that also has a MULTICAST mapping (`*_mc`, e.g. torch
 * symmetric memory's multicast_ptr). ts_fit_step's optimizer kernel then, per step: sums this rank's partials into
 * `grads_local`; rank barrier; for its 1/world of the 1024-element chunks reads the cross-rank gradient sum with one
 * multimem.ld_reduce per float4, applies SGD / Adam, and stores the new parameters and shadows through the multicast
 * mappings, i.e. into every rank's buffers at once; rank barrier. `params_local` replaces the buffer given to
 * ts_mlp_bind_params (copy the weights into it first), the shadow buffer holds ts_mlp_shadow_elems() bf16 elements
 * (call ts_mlp_sync_weights after binding). flag_bufs as for ts_dp_bind. */
int64_t ts_mlp_shadow_elems(const ts_mlp *m);
int32_t ts_dp_bind_nvls(ts_mlp *m, int32_t rank, int32_t world, float *grads_local, const float *grads_mc, float *params_local,
                        float *params_mc, void *shadow_local, void *shadow_mc, uint32_t *const *flag_bufs /*HOST*/);
/* Optional second set of bf16 shadows (symmetric memory + its multicast mapping, ts_mlp_shadow_elems() elements; no
 * counterpart in the reference). Nets on the layer-by-layer path exchange every layer on a side stream while the backward pass
 * of the layers below is still running (TSB200_NVLS_OVERLAP_SMS when ts_dp_bind_nvls runs: -1, the default, = exchange CTAs
 * co-resident with the GEMM CTAs; n > 0 = on n SMs the backward GEMMs leave free; 0 = off, one kernel after the backward pass). With ONE shadow set a layer's exchange has to wait for the dX GEMM that still reads its weights; with
 * two, step s writes the set step s+1 reads and the exchange starts as soon as the layer's dW is enqueued. Call
 * ts_mlp_sync_weights afterwards. */
int32_t ts_dp_nvls_second_shadow(ts_mlp *m, void *shadow2_local, void *shadow2_mc);

/* Classifier.evaluate (examples/tinysys/tinysys/classifier.py:37-39): dropout off, no state change. */
int32_t ts_eval_step(ts_mlp *m, const void *inputs, int32_t inputs_dtype, const int64_t *targets, int64_t batch,
                     int64_t *predictions, ts_batch_record *record, void *stream);

/* Evaluate with no device allocation anywhere. No counterpart in the reference; added for hosts that are not
 * in a position to hand the library a workspace: a serving process whose GPU memory is wholly owned by another
 * framework's pool, a latency-sensitive loop that cannot afford the device-wide synchronisation cudaMalloc and
 * cudaFree imply, or a plug-in that is simply not given an allocator. Same inputs, outputs and arithmetic as
 * ts_eval_step for TS_PREC_F32 handles with fp32 inputs (other combinations: TS_E_UNSUPPORTED), and it needs
 * only ts_mlp_create + ts_mlp_bind_params, not ts_mlp_set_workspace.
 *   - All intermediates (hidden activations, the loss/accuracy partials, the completion ticket) live in a
 *     16 MiB array compiled into the library; every GPU that loads the library carries it until the process exits.
 *   - Limit: batch * (sum of hidden widths) * 4 bytes plus about 8 KiB must fit in those 16 MiB (batch 4096 with one
 *     256-wide hidden layer uses 4 MiB; GELU handles need twice the activation space). Larger requests return
 *     TS_E_TOO_LARGE and launch nothing.
 *   - The first call on each device looks the array's device address up (cudaGetSymbolAddress); make that first
 *     call before starting any stream capture on that device.
 *   - The array is shared by every handle and every call in the process and still holds the previous call's
 *     data: the function re-zeroes the one word that has to start at zero, but two calls that can overlap on the
 *     GPU (different streams, or host threads sharing a device) race on it. Serialise them.
 *   - It calls no cudaMalloc / cudaFree or any other allocating routine. */
int32_t ts_eval_step_static(ts_mlp *m, const void *inputs, int32_t inputs_dtype, const int64_t *targets, int64_t batch,
                            int64_t *predictions, ts_batch_record *record, void *stream);

/* ---- metrics: replaces Metrics.update/compute/reset (examples/tinysys/tinysys/metrics.py:13-27) ---- */

/* Fold `count` records of a ring of `ring_len` records, starting at slot `first`, into `state`
 * in slot order (fp64 adds, same order as successive Mean.update calls). */
int32_t ts_metrics_accumulate(ts_metrics_state *state, const ts_batch_record *ring, int64_t ring_len, int64_t first,
                              int64_t count, void *stream);
/* Metrics.update for tensors that did not come from this engine: counts predictions == targets on the GPU. */
int32_t ts_metrics_update_raw(ts_metrics_state *state, const float *loss, const int64_t *predictions,
                              const int64_t *targets, int64_t batch, void *stream);
int32_t ts_metrics_reset(ts_metrics_state *state, void *stream);
/* Copies the state to HOST memory; synchronises `stream`. */
int32_t ts_metrics_read(const ts_metrics_state *state, ts_metrics_state *out /*HOST*/, void *stream);

/* ---- measurement support (no counterpart in the reference, which ships no profiling hooks) ----
 * Between ts_profile_begin() and ts_profile_end() every kernel launch is bracketed by CUDA events recorded on
 * the launching stream; ts_profile_end() synchronises them and returns, per kernel kind, the summed device
 * time in milliseconds and the launch count (arrays of at least ts_profile_kind_count() entries, HOST). */
int32_t ts_profile_begin(void);
int32_t ts_profile_end(double *ms_per_kind /*HOST*/, int64_t *launches_per_kind /*HOST*/, int32_t n_kinds);
int32_t ts_profile_kind_count(void);
const char *ts_profile_kind_name(int32_t kind);

/* Test hook (no counterpart in the reference): the tcgen05 GEMM on its own, D[M,N] = A[M,K] * B[K,N] with bf16
 * operands and fp32 output `out[split][M][ldo]`. a_major/b_major 0: operand stored [MN][K]; 1: stored [K][MN]. */
int32_t ts_test_gemm_bf16(const void *A, int32_t a_major, int64_t lda, const void *B, int32_t b_major, int64_t ldb, int32_t M,
                          int32_t N, int32_t K, float *out, int64_t ldo, int32_t splits, int64_t split_stride,
                          uint64_t *debug_clocks /* 32 x u64 device buffer for an in-kernel clock trace, or NULL */, void *stream);

/* Test hook: 32 clock64 samples taken by CTA 0 of the most recent fused forward+head launch (HOST buffer); synchronises. */
int32_t ts_test_fused_trace(uint64_t *out /*HOST, 32 entries*/);
/* The same for the single-kernel small-batch step (ts_fit_step on at most 64 rows, one hidden layer, bf16 mode). */
int32_t ts_test_step_trace(uint64_t *out /*HOST, 32 entries*/);
/* The same for the dW1 + update kernel of the fused training path (16 samples). */
int32_t ts_test_wgrad_trace(uint64_t *out /*HOST, 16 entries*/);
/* The same for the streaming evaluate kernel (CTA 0, third slab; 32 samples). */
int32_t ts_test_eval_trace(uint64_t *out /*HOST, 32 entries*/);

/* number of kernels this library has launched on the calling process so far (bench.py's gpu_launches) */
int64_t ts_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* TSB200_H */
