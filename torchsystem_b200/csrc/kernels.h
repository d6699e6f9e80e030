// Internal launcher API shared by engine.cu, simt_kernels.cu and tc_gemm.cu.
#pragma once

#include "common.cuh"

namespace tsb {

constexpr int kMaxLayers = 8;          // dense layers per aggregate
constexpr int kMaxClasses = 32;        // width of the last layer handled by the fused head kernel
constexpr int kChunk = 1024;           // flat-buffer alignment of every parameter tensor (elements)
constexpr int kMaxSplits = 32;         // split-K partials per weight tensor
constexpr int kMaxRowBlocks = 1024;    // 128-row blocks whose small-gradient partials are kept apart (= kMaxHeadBlocks: batches up to 131,072 rows)
constexpr int kReduceAbove = 128;      // a tensor with more partials than this is pre-reduced in groups of kReduceFan before the optimizer
constexpr int kReduceFan = 32;
constexpr int kReduceSlots = kMaxRowBlocks / kReduceFan;   // extra partial slots behind the kMaxRowBlocks regular ones
constexpr int kMaxHeadBlocks = 1024;   // head-kernel grid cap (partials buffer size)

// ---------------------------------------------------------------- fp32 SIMT GEMM (precision mode F32)
enum { EPI_STORE = 0, EPI_FWD = 1, EPI_ACTGRAD = 2 };

struct SgemmArgs {
    const float *A, *B;
    float *C;
    int M, N, K;
    int64_t sam, sak, sbk, sbn;  // A(m,k) = A[m*sam + k*sak], B(k,n) = B[k*sbk + n*sbn]
    int ldc;
    int k_chunk;                 // K range per blockIdx.z (multiple of 16)
    int64_t c_split_stride;      // elements between split outputs
    // epilogue operands
    const float *bias;           // EPI_FWD
    float *U_out;                // EPI_FWD, GELU: post-dropout pre-activation
    const float *A_prev;         // EPI_ACTGRAD, ReLU: the layer's own output
    const float *U_prev;         // EPI_ACTGRAD, GELU
    DropoutSpec drop;
    int act;
    int training;
};
int launch_sgemm(int epi, const SgemmArgs &a, int splits, cudaStream_t st);

// ---------------------------------------------------------------- fused head: last Linear + CE + argmax + dZ + dA
struct HeadPartial {
    float loss_sum;
    int correct;
};

struct HeadArgs {
    const void *A;        // [B, K] last hidden activation (float or bf16)
    const float *W;       // [C, K] fp32 master
    const float *bias;    // [C]
    const int64_t *y;     // [B]
    int64_t B;
    int K, C;
    int a_is_bf16;        // element type of A and dH
    int round_w_bf16;     // bf16 mode: contract with bf16-rounded weights (what the shadow holds)
    int64_t *pred;        // [B]
    float *dz;            // [B, C] or nullptr (evaluate)
    void *dH;             // [B, K] or nullptr
    const float *U_prev;  // GELU: pre-activation of the last hidden layer
    DropoutSpec drop;     // dropout of the last hidden layer (for the activation gradient)
    int act;
    HeadPartial *partials;
    unsigned int *ticket;
    ts_batch_record *record;
    uint32_t *bad_target;  // optional host-visible word: set to row + 1 when a target lies outside [0, C) (e.g. ignore_index)
};
int launch_head(const HeadArgs &a, int sm_count, cudaStream_t st);

// Tensor-core head (bf16 mode, wide hidden layers): softmax / NLL / argmax / dZ from the logits GEMM's output, and the
// fold of the transposed dW_last partials + db_last into `grads` (the contractions themselves are tc_gemm launches).
int launch_logits_head(const float *Z, const float *bias, const int64_t *y, int64_t B, int C, int64_t *pred, float *dz,
                       __nv_bfloat16 *dzb, HeadPartial *partials, unsigned int *ticket, ts_batch_record *record, uint32_t *bad_target,
                       cudaStream_t st);
int launch_wlast_finish(const float *wt, int64_t split_stride, int splits, int K, int C, const float *dz, int64_t B, float *gw,
                        float *gb, cudaStream_t st);
// dH of the last hidden layer from dZ (bf16, [B,16]) and the zero-padded [16,K] last-layer shadow: an HBM-bound SIMT pass with the
// activation gradient, the dropout factor and the 32-row bias-gradient partials (same layout as the ACTGRAD GEMM epilogue's)
bool head_dh_supported(int K, int C, int64_t colsum_stride);
int launch_head_dh(const __nv_bfloat16 *dzb, const __nv_bfloat16 *Wp, const __nv_bfloat16 *prev, int64_t B, int K, int C, int act,
                   const DropoutSpec &drop, int training, __nv_bfloat16 *dH, float *colsum, int64_t colsum_stride, cudaStream_t st);

// dW_last[c,k] = sum_r dz[r,c] * A[r,k], db_last[c] = sum_r dz[r,c]; `splits` row ranges -> partials
int launch_head_wgrad(const void *A, int a_is_bf16, const float *dz, int64_t B, int K, int C, float *w_partial,
                      int64_t w_stride, float *b_partial, int64_t b_stride, int splits, cudaStream_t st);

// dW_last, db_last and the last hidden layer's db in one launch; one partial per 128-row block (ceil(B/128) of them)
int launch_small_grads(const void *A, const void *dH, int is_bf16, const float *dz, int64_t B, int K, int C, float *w_partial,
                       int64_t w_stride, float *bl_partial, int64_t bl_stride, float *bh_partial, int64_t bh_stride,
                       cudaStream_t st);

// column sums of X[B,H] (float or bf16) over `splits` row ranges -> partial[s*stride + h]
int launch_colsum(const void *X, int is_bf16, int64_t B, int H, float *partial, int64_t stride, int splits,
                  cudaStream_t st);

// ---------------------------------------------------------------- optimizer over the flat buffers
struct TensorDesc {
    int64_t offset;       // element offset in the flat buffers
    int64_t count;        // elements
    int64_t partial_off;  // element offset of this tensor's partials in the partial scratch
    int64_t partial_stride;
    int64_t shadow_off;   // element offset in the bf16 shadow buffer, -1 if none
    int32_t nsplit;       // 0: gradient already in `grads`; >0: sum this many partials
};

struct OptArgs {
    TensorDesc t[2 * kMaxLayers];
    int n_tensors;
    float *params, *grads, *exp_avg, *exp_avg_sq, *max_exp_avg_sq;
    const float *partials;
    __nv_bfloat16 *shadow;
    const int32_t *chunk_tensor;
    int kind;
    int update;          // 0: only reduce partials into grads (data-parallel backward)
    float *params_mc;           // NVLS data parallel: multicast mappings of the parameter and shadow buffers; updates are
    __nv_bfloat16 *shadow_mc;   //   stored through them (multimem.st) and so reach every rank. nullptr: ordinary stores
    int chunk_begin;     // first 1024-element chunk this launch covers (ts_opt_apply_range), 0 for the whole buffer
    int partial_depth;   // partial rows in flight per thread (4 KiB of shared memory each): 16 for small models, 4 for large ones
    float grad_scale;
    float lr, beta1, beta2, eps, one_minus_beta1, one_minus_beta2, step_size, bc2_sqrt;
    float weight_decay, wd_mul /* 1 - lr * weight_decay (decoupled) */, momentum, one_minus_dampening;
    int nesterov, amsgrad, decoupled, first_step /* SGD momentum: buffer := grad */;
};
int launch_optimizer(const OptArgs &a, int64_t n_chunks, cudaStream_t st);

// Two-level sum for tensors with many row-block partials: group g = partials [32 g, 32 g + 32) summed in index order into
// slot dst_off + g * stride; the optimizer then sums the groups. Fixed order, no atomics.
struct ReduceArgs {
    int n;
    int64_t off[4], stride[4], count[4], dst_off[4];
    int nsplit[4];
};
int launch_reduce_partials(float *partials, const ReduceArgs &r, cudaStream_t st);

// data parallel: peer-mapped gradient buffers and flag arrays of every rank (this step's parity)
constexpr int kMaxRanks = 8;
// Inbox of the low-latency exchange: per step parity, kMaxRanks regions for what the peers send to the owner of a CTA's share
// plus one region for the sums the owners send back; a region holds, per CTA of the kernel, kLlQuads x 256 sixteen-byte
// packets {three floats, step number}.   2 * kLlParityBytes = TS_DP_LL_INBOX_BYTES (include/tsb200.h)
constexpr int kLlCtas = 128, kLlQuads = 3;
constexpr int64_t kLlParityBytes = (int64_t)(kMaxRanks + 1) * kLlCtas * kLlQuads * 256 * 16;
__host__ __device__ inline int64_t ll_index(int region, int cta, int tid) { return ((int64_t)(region * kLlCtas + cta) * kLlQuads) * 256 + tid; }
struct DpArgs {
    const float *grads[kMaxRanks];   // rank r's reduced gradient (flat layout), readable from this GPU
    uint32_t *flags[kMaxRanks];      // rank r's flag array: entry q = last step whose gradient rank q has published; [8] arrival ticket
    // low-latency exchange of the dW1 + update kernel (ts_dp_bind_ll; ll_own == nullptr: the pull exchange above)
    float4 *ll_peer[kMaxRanks];      // rank r's inbox (this step's parity), writable from this GPU
    const float4 *ll_own;            // this rank's inbox (this step's parity)
    int ll_direct;                   // two ranks: each sends its share straight to the other and both add in rank order (one hop)
    int world, rank;
    uint32_t step;
    unsigned long long timeout_ns;   // how long a wait for a peer may last before the kernel gives up (no trap)
    uint32_t *err;                   // host-visible (mapped pinned) word: set to the step number on a timeout
};
int launch_dp_optimizer(const OptArgs &a, const DpArgs &dp, int64_t n_chunks, int sm_count, cudaStream_t st);
// NVLS variant (large models): every rank reduces 1/world of the chunks inside the NVSwitch (multimem.ld_reduce over the
// multicast mapping `grads_mc`), updates them and broadcasts parameters + bf16 shadows with multimem.st (a.params_mc / a.shadow_mc)
int launch_nvls_optimizer(const OptArgs &a, const DpArgs &dp, const float *grads_mc, int64_t n_chunks, int sm_count, cudaStream_t st);

// one layer's chunk range [a.chunk_begin, a.chunk_begin + n_chunks) of the same exchange, on `n_ctas` CTAs (side stream, overlap)
int launch_nvls_layer(const OptArgs &a, const DpArgs &dp, const float *grads_mc, int64_t n_chunks, int n_ctas, cudaStream_t st);
// the single-GPU counterpart: one layer's chunk range updated by CTAs that share SMs with the backward GEMMs (side stream)
int launch_layer_optimizer(const OptArgs &a, int64_t n_chunks, int n_ctas, cudaStream_t st);

int launch_cast_bf16(const float *src, __nv_bfloat16 *dst, int64_t n, cudaStream_t st, bool wait_at_end = false);
// uint8 -> lut[byte], written as bf16 (dst_bf16) or fp32 (dst_f32); exactly one destination is non-null
int launch_cast_u8(const uint8_t *src, __nv_bfloat16 *dst_bf16, float *dst_f32, int64_t n, const float *lut, cudaStream_t st,
                   bool wait_at_end = false);

// ---------------------------------------------------------------- metrics
int launch_metrics_accumulate(ts_metrics_state *state, const ts_batch_record *ring, int64_t ring_len, int64_t first,
                              int64_t count, cudaStream_t st);
int launch_metrics_update_raw(ts_metrics_state *state, const float *loss, const int64_t *pred, const int64_t *tgt,
                              int64_t B, cudaStream_t st);

// ---------------------------------------------------------------- dW1 = dH^T X + optimizer update in one kernel (wgrad_update.cu)
struct WgradUpdateArgs {
    const __nv_bfloat16 *dH;    // [B, H] bf16, written by the fused kernel
    const __nv_bfloat16 *X;     // [B, D] bf16 inputs
    int64_t B;
    int D, H;
    int64_t off_w1, sh_w1;      // element offsets of W1 in the flat buffers / the bf16 shadow buffer
    int small_chunk_begin, small_chunks;   // the 1024-element chunks of the remaining tensors (b1, W2, b2)
    OptArgs opt;                // flat buffers, tensor table (row-block partials of the remaining tensors), hyper-parameters
    DpArgs dp;                  // world > 1: data parallel over peer mappings ('p2p'), gradient sum inside this kernel
    int sm_count;
};
bool wgrad_update_supported(int D, int H, int64_t B);
int launch_wgrad_update(const WgradUpdateArgs &a, cudaStream_t st);
int read_wgrad_trace(unsigned long long *out_host);   // 16 clock64 samples of CTA 0 of the last launch

}  // namespace tsb

// ---------------------------------------------------------------- tcgen05 bf16 GEMM (precision mode BF16), tc_gemm.cu
namespace tsb {

enum { TC_EPI_STORE_F32 = 0, TC_EPI_FWD = 1, TC_EPI_ACTGRAD = 2 };
enum { TC_MAJOR_K = 0, TC_MAJOR_MN = 1 };

// D[M,N] = A[M,K] * B[K,N], bf16 operands fed by TMA, fp32 accumulation in TMEM.
//   a_major K : A stored [M][K] (row pitch lda)     a_major MN : A stored [K][M]
//   b_major K : B stored [N][K] (row pitch ldb)     b_major MN : B stored [K][N]
struct TcGemmArgs {
    const __nv_bfloat16 *A, *B;
    int a_major, b_major;
    int64_t lda, ldb;
    int M, N, K;
    int epi;
    void *out;             // STORE_F32: float [splits][M][ldo]; FWD / ACTGRAD: bf16 [M][ldo]
    int64_t ldo;
    int splits;            // split-K (STORE_F32 only)
    int k_chunk;           // K elements per split (multiple of 64); 0 = derive from splits
    int64_t split_stride;  // elements between split outputs
    const float *bias;             // FWD: [N]
    const __nv_bfloat16 *A_prev;   // ACTGRAD: the layer's own bf16 output [M][ldo] (ReLU), or its stored pre-activation (GELU)
    int act;                       // TS_ACT_RELU / TS_ACT_GELU (FWD and ACTGRAD)
    __nv_bfloat16 *U_out;          // FWD, GELU: post-dropout pre-activation [M][ldo], kept for the backward pass
    float *colsum_out;             // ACTGRAD (optional): column sums of the stored output per 32-row group -> row (4 * row_tile + quarter)
    int64_t colsum_stride;         //   of this array (the hidden-layer bias gradient partials; fixed order, no atomics)
    DropoutSpec drop;
    int training;
    unsigned long long *dbg;       // optional device buffer of 32 clock64 samples from CTA (0,0,0)
};
int launch_tc_gemm(const TcGemmArgs &a, int sm_count, cudaStream_t st);

// ---------------------------------------------------------------- fused one-hidden-layer forward + head (fused_mlp.cu)
struct FusedMlpArgs {
    const __nv_bfloat16 *X;     // [B, D] bf16 inputs, or nullptr when X_f32 is given
    const float *X_f32;         // [B, D] fp32 inputs: converted to bf16 on chip (no separate cast kernel)
    __nv_bfloat16 *x_bf16_out;  // with X_f32 in training: where the bf16 copy for the dW1 GEMM goes
    const __nv_bfloat16 *W1;    // [H, D] bf16 shadow of layer-0 weight
    const __nv_bfloat16 *W2p;   // [16, H] bf16 shadow of the last weight, rows >= C zero
    const float *b1, *b2;
    const int64_t *y;
    int64_t B;
    int D, H, C;
    int64_t *pred;
    float *dz;                  // training outputs (nullptr for evaluate)
    __nv_bfloat16 *A_out, *dH_out;
    // optional: per-CTA partials of dW_last [C][H], db_last [C], db_hidden [H] computed in-kernel (H >= 128, C <= 15)
    float *w2_partial, *b2_partial, *b1_partial;
    int64_t w2_stride, b2_stride, b1_stride;
    DropoutSpec drop;
    int training;
    HeadPartial *partials;
    unsigned int *ticket;
    ts_batch_record *record;
    uint32_t *bad_target;       // optional host-visible word, as in HeadArgs
};
// ---------------------------------------------------------------- single-kernel fit step for small batches (step_kernel.cu)
struct FitStepArgs {
    const __nv_bfloat16 *X;     // [B, D] bf16 inputs, or nullptr when X_f32 is given
    const float *X_f32;         // [B, D] fp32 inputs, converted on chip
    const __nv_bfloat16 *W1;    // [H, D] bf16 shadow of layer-0 weight (rewritten by the kernel)
    const __nv_bfloat16 *W2p;   // [16, H] bf16 shadow of the last weight (rewritten by the kernel)
    const int64_t *y;
    int64_t B;
    int D, H, C;
    int64_t *pred;
    DropoutSpec drop;
    int training;
    ts_batch_record *record;
    int64_t off_w1, off_b1, off_w2, off_b2, sh_w1, sh_w2;
    OptArgs opt;                // flat buffers, tensor table (nsplit 0: gradients final in `grads`) and this update's hyper-parameters
    uint32_t *bad_target;       // optional host-visible word, as in HeadArgs
    unsigned int *flag;         // device counter the compute cluster raises by `ns` per step; the updater CTAs wait for flag_target
    unsigned int flag_target;
    int n_chunks, sm_count;
};
bool fit_step_supported(int D, int H, int C, int64_t B);
int launch_fit_step(const FitStepArgs &a, cudaStream_t st);
int read_step_trace(unsigned long long *out_host);

bool fused_mlp_supported(int D, int H, int C);
inline bool fused_grads_supported(int H, int C, int64_t B) { return H >= 128 && C <= 15 && (B + 127) / 128 <= kMaxRowBlocks; }
int fused_cluster_size(int64_t B, int H, int sm_count);   // CTAs per 128-row slab the launcher will use
int launch_fused_mlp(const FusedMlpArgs &a, int sm_count, cudaStream_t st);
// persistent, pipelined evaluate for batches whose slabs are single CTAs (eval_stream.cu); same arguments, evaluate only
bool eval_stream_supported(int D, int H, int C);
int launch_eval_stream(const FusedMlpArgs &a, int sm_count, cudaStream_t st);
int read_eval_trace(unsigned long long *out_host);    // 32 clock64 samples of CTA 0's third slab
int read_fused_trace(unsigned long long *out_host);   // 32 clock64 samples of CTA 0 of the last launch

}  // namespace tsb
