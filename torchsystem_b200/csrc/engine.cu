// C ABI of the tsb200 engine (include/tsb200.h): handle management, workspace carving and the
// per-batch kernel sequences behind Classifier.fit / Classifier.evaluate / Metrics.
#include <cmath>
#include <climits>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <atomic>
#include <cstring>
#include <mutex>
#include <new>
#include <vector>

#include "kernels.h"

namespace tsb {

static thread_local char g_err[1024] = "";
long long g_launch_count = 0;

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

static inline int64_t round_up(int64_t v, int64_t a) { return (v + a - 1) / a * a; }

// ---- optional per-kernel timing: CUDA events recorded on the launching stream around each launch.
enum { K_CAST = 0, K_FWD_GEMM, K_HEAD, K_HEAD_WGRAD, K_COLSUM, K_DW_GEMM, K_DX_GEMM, K_OPT, K_FUSED, K_STEP, K_WGRAD_UPD, K_COUNT };
static const char *const kKindNames[K_COUNT] = {"cast_bf16", "fwd_gemm", "head", "head_wgrad", "colsum", "dw_gemm",
                                                "dx_gemm",   "optimizer", "fused_fwd_head", "fit_step_small", "dw_gemm_update"};
struct ProfPair { int kind; cudaEvent_t a, b; };
static bool g_prof_on = false;
static std::vector<ProfPair> g_prof;

struct ProfScope {
    int kind; cudaStream_t st; cudaEvent_t a = nullptr, b = nullptr;
    ProfScope(int k, cudaStream_t s) : kind(k), st(s) {
        if (!g_prof_on) return;
        if (cudaEventCreate(&a) != cudaSuccess || cudaEventCreate(&b) != cudaSuccess) { a = b = nullptr; return; }
        cudaEventRecord(a, st);
    }
    ~ProfScope() {
        if (!a) return;
        cudaEventRecord(b, st);
        g_prof.push_back({kind, a, b});
    }
};

}  // namespace tsb

using namespace tsb;

struct ts_mlp {
    int n_layers = 0;  // dense layers
    int dims[kMaxLayers + 1] = {0};
    int act = TS_ACT_RELU;
    float p = 0.0f, scale = 1.0f;
    int prec = TS_PREC_F32;
    int device = 0;
    int sm_count = 148;
    int bwd_sms = 148;            // SMs the backward GEMMs use: fewer than sm_count leaves room for a concurrently running collective
    cudaEvent_t *ov_events = nullptr;   // per-call (ts_fit_backward_overlap): event of layer l recorded once its gradients are enqueued
    int32_t *ov_final = nullptr;        // per-call out: gradient tensor t is final in `grads` when its layer's event fires
    unsigned int step_flag = 0;   // value of the single-kernel step's gradient-ready counter after the launches so far
    int eval_stream = 1;          // evaluate, large batches: persistent pipelined kernel (TSB200_EVAL_STREAM=0: the single-slab fused kernel)
    int wgrad_update = 1;         // fused path: dW1 GEMM, split-K reduce and optimizer update in one kernel (TSB200_WGRAD_UPDATE=0: off)
    int step_kernel = 1;          // batches of at most 64 rows: the whole fit step in one kernel (TSB200_STEP_KERNEL=0: off)
    int inkernel_cast = -1;       // fp32 batches converted inside the fused kernel: 1 always, 0 never (cast kernel in front), -1 by batch size

    // flat parameter layout (torch order: W0, b0, W1, b1, ...)
    int n_tensors = 0;
    int64_t off[2 * kMaxLayers] = {0}, cnt[2 * kMaxLayers] = {0};
    int max_split[2 * kMaxLayers] = {0};
    int64_t total = 0, n_chunks = 0;
    std::vector<int32_t> chunk_tensor_host;

    float *params = nullptr, *grads = nullptr;
    bool opt_bound = false;
    int opt_kind = TS_OPT_SGD;
    ts_opt_hyper hyper = {0, 0, 0, 0};
    float *exp_avg = nullptr, *exp_avg_sq = nullptr, *max_exp_avg_sq = nullptr;
    int64_t step = 0;

    // workspace (caller-owned memory)
    bool ws_set = false;
    int64_t max_batch = 0;
    int32_t *d_chunk_tensor = nullptr;
    unsigned int *ticket = nullptr;
    HeadPartial *head_partials = nullptr;
    void *acts[kMaxLayers] = {nullptr};
    float *upre[kMaxLayers] = {nullptr};
    float *dz = nullptr;
    void *dh[2] = {nullptr, nullptr};
    float *partials = nullptr;
    int64_t partial_off[2 * kMaxLayers] = {0}, partial_stride[2 * kMaxLayers] = {0};
    __nv_bfloat16 *x_bf16 = nullptr, *shadow = nullptr;
    __nv_bfloat16 *x_bf16_alt = nullptr;   // second input buffer of the fused path: lets the next batch's cast overlap this step's optimizer
    int x_parity = 0;
    float lut_host[256] = {0};    // uint8 inputs: table applied on the way in (ts_mlp_set_input_lut)
    bool lut_set = false;
    float *d_lut = nullptr;       // its device copy (workspace)
    float *x_f32 = nullptr;       // fp32 mode: uint8 batches are expanded here
    bool first_step = false;      // the step being applied is the first one (SGD momentum buffer := gradient)
    // tensor-core head (bf16 mode, models the fused kernel does not take, <= 16 classes): logits / dZ operand / dW_last^T partials
    bool tc_head = false;
    float *zbuf = nullptr, *wt_partial = nullptr;
    __nv_bfloat16 *dzb = nullptr;
    int64_t shadow_off[kMaxLayers] = {0};   // hidden-layer weights, dense [out][in]; last layer: [16][in], rows >= classes zero
    int64_t shadow_total = 0;
    bool fused = false;                     // one hidden layer of a width the fused kernel takes

    // data parallel over peer-mapped memory (ts_dp_bind)
    int dp_world = 1, dp_rank = 0;
    float *dp_grads[2][kMaxRanks] = {{nullptr}};
    uint32_t *dp_flags[kMaxRanks] = {nullptr};
    char *ll_inbox[kMaxRanks] = {nullptr};         // ts_dp_bind_ll: rank r's inbox of the low-latency exchange (two parities)
    bool ll_bound = false;
    // NVLS variant (ts_dp_bind_nvls): multicast mappings of the symmetric gradient / parameter / shadow buffers
    // nvls overlap: each layer's reduce + update runs on a side stream, on a few SMs, under the backward pass of the layers below
    cudaStream_t dp_stream = nullptr;
    cudaEvent_t dp_ev_layer[kMaxLayers] = {nullptr}, dp_ev_done = nullptr;
    int nvls_overlap_ctas = 0;            // 0: one whole-buffer kernel after the backward pass
    bool nvls_coresident = false;
    int local_overlap = 1;                // single GPU, layer-by-layer path: each layer's optimizer update on the side stream under the
                                          // rest of the backward pass (TSB200_OPT_OVERLAP=0: one optimizer launch at the end)         // the exchange CTAs share SMs with the backward GEMM CTAs instead of owning reserved SMs
    uint32_t dp_seq = 0;                  // launch counter of the nvls kernels: the rank barriers count in it
    uint32_t *status_host = nullptr, *status_dev = nullptr;   // mapped pinned words: [0] = row + 1 of a target outside [0, C)
    uint32_t *dp_err_host = nullptr, *dp_err_dev = nullptr;   // mapped pinned word the DP kernels set when a peer never arrives
    double dp_timeout_s = 600.0;                               // TSB200_DP_TIMEOUT_S / ts_dp_set_timeout
    bool nvls = false;
    const float *grads_mc = nullptr;
    float *params_mc = nullptr;
    __nv_bfloat16 *shadow_mc = nullptr, *shadow_ext = nullptr;
    __nv_bfloat16 *shadow_alt = nullptr, *shadow_mc_alt = nullptr;   // second set of bf16 shadows (ts_dp_nvls_second_shadow): the update of
                                                                      // step s writes the set step s+1 reads, so it never races with this step's dX

    int hidden_layers() const { return n_layers - 1; }
    int classes() const { return dims[n_layers]; }
    size_t elem() const { return prec == TS_PREC_BF16 ? 2 : 4; }
};

namespace {

struct Carver {
    char *base;
    int64_t used = 0;
    explicit Carver(void *b) : base(reinterpret_cast<char *>(b)) {}
    void *take(int64_t bytes) {
        void *p = base ? base + used : nullptr;
        used += round_up(bytes < 1 ? 1 : bytes, 256);
        return p;
    }
};

// Lays the workspace out; with base == nullptr only measures it.
int64_t carve(ts_mlp *m, void *base, int64_t Bm) {
    Carver c(base);
    const int L = m->n_layers;
    m->d_chunk_tensor = (int32_t *)c.take(m->n_chunks * 4);
    m->d_lut = (float *)c.take(256 * 4);
    m->x_f32 = m->prec == TS_PREC_F32 ? (float *)c.take(Bm * m->dims[0] * 4) : nullptr;
    m->ticket = (unsigned int *)c.take(256);
    m->head_partials = (HeadPartial *)c.take(sizeof(HeadPartial) * kMaxHeadBlocks);
    int maxH = 0;
    for (int l = 0; l < L - 1; ++l) {
        const int H = m->dims[l + 1];
        maxH = H > maxH ? H : maxH;
        m->acts[l] = c.take(Bm * H * (int64_t)m->elem());
        m->upre[l] = (m->act == TS_ACT_GELU) ? (float *)c.take(Bm * H * 4) : nullptr;
    }
    m->dz = (float *)c.take(Bm * m->classes() * 4);
    m->dh[0] = c.take(Bm * maxH * (int64_t)m->elem());
    m->dh[1] = (L > 2) ? c.take(Bm * maxH * (int64_t)m->elem()) : nullptr;
    int64_t pfloats = 0;
    for (int t = 0; t < m->n_tensors; ++t) {
        m->partial_stride[t] = round_up(m->cnt[t], 4) + 32;   // +128 B: successive partials of an element fall into different L2 slices
        m->partial_off[t] = pfloats;
        // tensors that can have more than kReduceAbove partials get kReduceSlots more slots for the pre-reduced groups
        if (m->max_split[t] > 1) pfloats += (int64_t)(m->max_split[t] + (m->max_split[t] > kReduceAbove ? kReduceSlots : 0)) * m->partial_stride[t];
    }
    m->partials = (float *)c.take(pfloats * 4);
    if (m->prec == TS_PREC_BF16) {
        m->x_bf16 = (__nv_bfloat16 *)c.take(Bm * m->dims[0] * 2);
        m->x_bf16_alt = m->fused ? (__nv_bfloat16 *)c.take(Bm * m->dims[0] * 2) : nullptr;
        __nv_bfloat16 *ws_shadow = (__nv_bfloat16 *)c.take(m->shadow_total * 2);
        // nets on the layer-by-layer path keep a second set: a layer's update (side stream) writes the set the NEXT step reads
        // while this step's dX GEMM still reads the current one
        __nv_bfloat16 *ws_shadow2 = (!m->fused && m->n_layers > 2) ? (__nv_bfloat16 *)c.take(m->shadow_total * 2) : nullptr;
        if (!m->shadow_ext) { m->shadow = ws_shadow; m->shadow_alt = ws_shadow2; }   // NVLS data parallel keeps its own sets in symmetric memory
    }
    if (m->tc_head) {
        m->zbuf = (float *)c.take(Bm * 16 * 4);
        m->dzb = (__nv_bfloat16 *)c.take(Bm * 16 * 2);
        m->wt_partial = (float *)c.take((int64_t)kMaxSplits * m->dims[L - 1] * 16 * 4);
    }
    return c.used;
}

int splits_for(int64_t B, int max_split, int granule, int *k_chunk) {
    int64_t s = (B + 63) / 64;
    if (s > max_split) s = max_split;
    if (s < 1) s = 1;
    int64_t kc = round_up((B + s - 1) / s, granule);
    *k_chunk = (int)kc;
    return (int)((B + kc - 1) / kc);
}

DropoutSpec drop_spec(const ts_mlp *m, const uint8_t *const *masks, uint64_t seed, int layer, bool training) {
    DropoutSpec d;
    d.mask = (masks && training) ? masks[layer] : nullptr;
    d.key = hash32(hash32((uint32_t)seed ^ 0x9E3779B9u) ^ hash32((uint32_t)(seed >> 32) + 0x7F4A7C15u) ^
                   hash32((uint32_t)m->step * 0x85EBCA6Bu + (uint32_t)layer + 1u));
    d.p = training ? m->p : 0.0f;
    d.thresh = (uint32_t)((double)d.p * 65536.0 + 0.5);
    d.scale = training ? m->scale : 1.0f;
    return d;
}

// An earlier batch held a target outside [0, classes): its loss is NaN; say so instead of training on.
int target_status(const ts_mlp *m) {
    const uint32_t bad = m->status_host ? *(volatile uint32_t *)m->status_host : 0u;
    TS_REQUIRE(bad == 0u, TS_E_INVALID,
               "a target outside [0, %d) was seen (row %u of its batch): CrossEntropyLoss's ignore_index (-100) is not implemented, "
               "targets must be class indices; the loss of that batch is NaN (ts_mlp_clear_status resets this)", m->classes(), bad - 1u);
    return TS_OK;
}

int check_ready(const ts_mlp *m, int64_t B, bool need_opt) {
    TS_REQUIRE(m != nullptr, TS_E_INVALID, "null handle");
    TS_REQUIRE(m->params && m->grads, TS_E_STATE, "parameters not bound: call ts_mlp_bind_params first");
    TS_REQUIRE(m->ws_set, TS_E_STATE, "workspace not set: call ts_mlp_set_workspace first");
    TS_REQUIRE(!need_opt || m->opt_bound, TS_E_STATE, "optimizer not bound: call ts_opt_bind first");
    TS_REQUIRE(B >= 1 && B <= m->max_batch, TS_E_INVALID, "batch %lld outside [1, %lld] (workspace capacity)",
               (long long)B, (long long)m->max_batch);
    return target_status(m);
}

// Forward through the hidden layers; *a_last receives the last hidden activation.
int forward_hidden(ts_mlp *m, const void *x, int x_dtype, int64_t B, const uint8_t *const *masks, uint64_t seed,
                   bool training, const void **a0_out, const void **a_last, cudaStream_t st) {
    const void *a_in = x;
    if (x_dtype == TS_DT_U8) {
        TS_REQUIRE(m->lut_set, TS_E_STATE, "uint8 inputs need ts_mlp_set_input_lut first");
        ProfScope ps(K_CAST, st);
        const bool f32 = m->prec == TS_PREC_F32;
        int rc = launch_cast_u8((const uint8_t *)x, f32 ? nullptr : m->x_bf16, f32 ? m->x_f32 : nullptr, B * m->dims[0], m->d_lut, st);
        if (rc) return rc;
        a_in = f32 ? (const void *)m->x_f32 : (const void *)m->x_bf16;
    } else if (m->prec == TS_PREC_F32) {
        TS_REQUIRE(x_dtype == TS_DT_F32, TS_E_UNSUPPORTED, "TS_PREC_F32 takes fp32 or uint8 inputs");
    } else if (x_dtype == TS_DT_F32) {
        ProfScope ps(K_CAST, st);
        int rc = launch_cast_bf16((const float *)x, m->x_bf16, B * m->dims[0], st);
        if (rc) return rc;
        a_in = m->x_bf16;
    } else {
        TS_REQUIRE(x_dtype == TS_DT_BF16, TS_E_INVALID, "unknown inputs dtype %d", x_dtype);
    }
    *a0_out = a_in;
    for (int l = 0; l < m->n_layers - 1; ++l) {
        const int in = m->dims[l], out = m->dims[l + 1];
        const DropoutSpec ds = drop_spec(m, masks, seed, l, training);
        int rc;
        ProfScope ps(K_FWD_GEMM, st);
        if (m->prec == TS_PREC_F32) {
            SgemmArgs g{};
            g.A = (const float *)a_in; g.sam = in; g.sak = 1;
            g.B = m->params + m->off[2 * l]; g.sbk = 1; g.sbn = in;
            g.C = (float *)m->acts[l]; g.ldc = out;
            g.M = (int)B; g.N = out; g.K = in;
            g.k_chunk = (int)round_up(in, 16); g.c_split_stride = 0;
            g.bias = m->params + m->off[2 * l + 1];
            g.U_out = m->upre[l];
            g.drop = ds; g.act = m->act; g.training = training && m->p > 0.0f;
            rc = launch_sgemm(EPI_FWD, g, 1, st);
        } else {
            TcGemmArgs g{};
            g.A = (const __nv_bfloat16 *)a_in; g.a_major = TC_MAJOR_K; g.lda = in;
            g.B = m->shadow + m->shadow_off[l]; g.b_major = TC_MAJOR_K; g.ldb = in;
            g.M = (int)B; g.N = out; g.K = in;
            g.epi = TC_EPI_FWD; g.out = m->acts[l]; g.ldo = out; g.splits = 1; g.split_stride = 0;
            g.act = m->act; g.U_out = (__nv_bfloat16 *)m->upre[l];      // GELU: pre-activations kept in bf16
            g.bias = m->params + m->off[2 * l + 1];
            g.drop = ds; g.training = training && m->p > 0.0f;
            rc = launch_tc_gemm(g, m->sm_count, st);
        }
        if (rc) return rc;
        a_in = m->acts[l];
    }
    *a_last = a_in;
    return TS_OK;
}

void fill_head(const ts_mlp *m, HeadArgs &h, const void *a_last, const int64_t *y, int64_t B, int64_t *pred,
               ts_batch_record *rec) {
    const int L = m->n_layers;
    h.A = a_last;
    h.W = m->params + m->off[2 * (L - 1)];
    h.bias = m->params + m->off[2 * (L - 1) + 1];
    h.y = y; h.B = B; h.K = m->dims[L - 1]; h.C = m->dims[L];
    h.a_is_bf16 = m->prec == TS_PREC_BF16;
    h.round_w_bf16 = m->prec == TS_PREC_BF16;
    h.pred = pred; h.dz = nullptr; h.dH = nullptr; h.U_prev = nullptr;
    h.act = m->act;
    h.partials = m->head_partials; h.ticket = m->ticket; h.record = rec;
    h.bad_target = m->status_dev;
}

// Tensor table of the optimizer kernels. Tensors with many row-block partials (batches beyond 16,384 rows) are summed in
// groups of kReduceFan by one extra launch first; the optimizer then sums the groups.
int fill_tensor_table(ts_mlp *m, const int *nsplit, OptArgs &o, bool *any_partial, cudaStream_t st, int t_begin = 0, int t_end = 1 << 30) {
    o.n_tensors = m->n_tensors;
    *any_partial = false;
    for (int t = 0; t < m->n_tensors; ++t) {
        o.t[t].offset = m->off[t]; o.t[t].count = m->cnt[t];
        o.t[t].partial_off = m->partial_off[t]; o.t[t].partial_stride = m->partial_stride[t];
        const int l = t / 2;
        o.t[t].shadow_off = (m->prec == TS_PREC_BF16 && t % 2 == 0) ? m->shadow_off[l] : -1;
        o.t[t].nsplit = (nsplit && t >= t_begin && t < t_end) ? nsplit[t] : 0;    // tensors outside the range: not this launch's business
        *any_partial |= o.t[t].nsplit > 0;
    }
    ReduceArgs r{};
    for (int t = 0; t < m->n_tensors; ++t) {
        if (o.t[t].nsplit <= kReduceAbove) continue;
        TS_REQUIRE(r.n < 4 && o.t[t].nsplit <= m->max_split[t] && m->max_split[t] == kMaxRowBlocks, TS_E_INVALID,
                   "tensor %d: %d partials cannot be pre-reduced", t, o.t[t].nsplit);
        const int i = r.n++;
        r.off[i] = m->partial_off[t]; r.stride[i] = m->partial_stride[t]; r.count[i] = m->cnt[t]; r.nsplit[i] = o.t[t].nsplit;
        r.dst_off[i] = m->partial_off[t] + (int64_t)kMaxRowBlocks * m->partial_stride[t];
        o.t[t].partial_off = r.dst_off[i];
        o.t[t].nsplit = (r.nsplit[i] + kReduceFan - 1) / kReduceFan;
    }
    if (r.n > 0) {
        ProfScope ps(K_COLSUM, st);
        int rc = launch_reduce_partials(m->partials, r, st);
        if (rc) return rc;
    }
    o.params = m->params; o.grads = m->grads; o.exp_avg = m->exp_avg; o.exp_avg_sq = m->exp_avg_sq;
    o.max_exp_avg_sq = m->max_exp_avg_sq;
    o.partials = m->partials; o.shadow = m->shadow; o.chunk_tensor = m->d_chunk_tensor;
    o.kind = m->opt_kind;
    o.partial_depth = m->n_chunks <= 3 * (int64_t)m->sm_count ? 16 : 4;
    return TS_OK;
}

// Advances the step counter and fills the hyper-parameters of one update.
void fill_update(ts_mlp *m, OptArgs &o, bool advance = true) {
    if (advance) {
        m->first_step = m->step == 0;
        m->step += 1;
    }
    o.first_step = m->first_step ? 1 : 0;
    o.weight_decay = (float)m->hyper.weight_decay;
    o.wd_mul = (float)(1.0 - m->hyper.lr * m->hyper.weight_decay);
    o.momentum = (float)m->hyper.momentum;
    o.one_minus_dampening = (float)(1.0 - m->hyper.dampening);
    o.nesterov = m->hyper.nesterov; o.amsgrad = m->hyper.amsgrad; o.decoupled = m->hyper.decoupled_weight_decay;
    const double b1 = m->hyper.beta1, b2 = m->hyper.beta2;
    o.lr = (float)m->hyper.lr;
    o.beta1 = (float)b1; o.beta2 = (float)b2; o.eps = (float)m->hyper.eps;
    o.one_minus_beta1 = (float)(1.0 - b1); o.one_minus_beta2 = (float)(1.0 - b2);
    if (m->opt_kind == TS_OPT_ADAM) {
        // host doubles exactly as torch computes them (adam.py: bias_correction*, step_size)
        const double bc1 = 1.0 - pow(b1, (double)m->step);
        const double bc2 = 1.0 - pow(b2, (double)m->step);
        o.step_size = (float)(m->hyper.lr / bc1);
        o.bc2_sqrt = (float)sqrt(bc2);
    }
}

int run_optimizer(ts_mlp *m, const int *nsplit, bool update, float grad_scale, cudaStream_t st) {
    OptArgs o{};
    bool any_partial = false;
    if (!update) {      // nothing to launch when no tensor has partials to fold
        for (int t = 0; t < m->n_tensors; ++t) any_partial |= nsplit && nsplit[t] > 0;
        if (!any_partial) return TS_OK;
    }
    int rc = fill_tensor_table(m, nsplit, o, &any_partial, st);
    if (rc) return rc;
    o.update = update ? 1 : 0; o.grad_scale = grad_scale;
    if (update) fill_update(m, o);
    ProfScope ps(K_OPT, st);
    return launch_optimizer(o, m->n_chunks, st);
}

// One-hidden-layer bf16 aggregates: cast (if needed) + ONE kernel for forward, loss, argmax and, when `train`,
// dZ / A / dH (fused_mlp.cu).
int fused_forward(ts_mlp *m, const void *x, int x_dtype, const int64_t *y, int64_t B, const uint8_t *const *masks, uint64_t seed,
                  bool train, int64_t *pred, ts_batch_record *rec, const void **a0_out, cudaStream_t st) {
    const void *a0 = x;
    TS_REQUIRE(x_dtype == TS_DT_F32 || x_dtype == TS_DT_BF16 || x_dtype == TS_DT_U8, TS_E_INVALID, "unknown inputs dtype %d", x_dtype);
    TS_REQUIRE(x_dtype != TS_DT_U8 || m->lut_set, TS_E_STATE, "uint8 inputs need ts_mlp_set_input_lut first");
    // fp32 batches: a cast kernel in front, or converted inside the fused kernel (which then also leaves the bf16 copy
    // the dW1 GEMM reads). Measured on B200: on-chip conversion wins when a slab is one CTA (large batches, HBM-bound:
    // 4 B/element read once instead of 4 + 2 + 2), and loses when the slab is a cluster, where every CTA of the cluster
    // would ingest and convert the whole fp32 tile. TSB200_INKERNEL_CAST=0/1 at ts_mlp_create overrides.
    const bool inkernel = m->inkernel_cast >= 0 ? m->inkernel_cast != 0 : fused_cluster_size(B, m->dims[1], m->sm_count) == 1;
    FusedMlpArgs f{};
    if (x_dtype != TS_DT_BF16) m->x_parity ^= 1;     // every converting route alternates between the two bf16 input buffers
    __nv_bfloat16 *xb = m->x_parity ? m->x_bf16_alt : m->x_bf16;
    if (x_dtype == TS_DT_F32 && inkernel) {
        f.X_f32 = (const float *)x;
        f.x_bf16_out = train ? xb : nullptr;
        a0 = xb;
    } else if (x_dtype == TS_DT_F32) {
        // The cast does not wait for the kernel in front (the previous step's optimizer) before converting, only before
        // exiting, so it overlaps it. Why its destination is free: the two buffers alternate per cast, so this one was
        // last read by the fused kernel / dW GEMM that followed the cast before the previous one; this cast can only
        // start once the most recent fused kernel has released its dependents, which that kernel does after its own
        // wait, i.e. after everything older -- those readers included -- has completed.
        ProfScope ps(K_CAST, st);
        int rc = launch_cast_bf16((const float *)x, xb, B * m->dims[0], st, true);
        if (rc) return rc;
        a0 = xb;
        f.X = xb;
    } else if (x_dtype == TS_DT_U8) {   // same early-running conversion as the fp32 cast, through the 256-entry table
        ProfScope ps(K_CAST, st);
        int rc = launch_cast_u8((const uint8_t *)x, xb, nullptr, B * m->dims[0], m->d_lut, st, true);
        if (rc) return rc;
        a0 = xb;
        f.X = xb;
    } else {
        f.X = (const __nv_bfloat16 *)x;
    }
    *a0_out = a0;
    f.W1 = m->shadow + m->shadow_off[0];
    f.W2p = m->shadow + m->shadow_off[1];
    f.b1 = m->params + m->off[1];
    f.b2 = m->params + m->off[3];
    f.y = y; f.B = B; f.D = m->dims[0]; f.H = m->dims[1]; f.C = m->dims[2];
    f.pred = pred;
    const bool fuse_grads = train && fused_grads_supported(m->dims[1], m->dims[2], B);
    f.dz = (train && !fuse_grads) ? m->dz : nullptr;      // fp32 dZ in global memory is only read by the SIMT gradient kernels
    f.A_out = (train && !fuse_grads) ? (__nv_bfloat16 *)m->acts[0] : nullptr;
    f.dH_out = train ? (__nv_bfloat16 *)m->dh[0] : nullptr;
    if (fuse_grads) {   // dW_last, db_last, db_hidden partials per 128-row CTA, straight from the tensor cores
        f.w2_partial = m->partials + m->partial_off[2]; f.w2_stride = m->partial_stride[2];
        f.b2_partial = m->partials + m->partial_off[3]; f.b2_stride = m->partial_stride[3];
        f.b1_partial = m->partials + m->partial_off[1]; f.b1_stride = m->partial_stride[1];
    }
    f.drop = drop_spec(m, masks, seed, 0, train);
    f.training = train && m->p > 0.0f;
    f.partials = m->head_partials; f.ticket = m->ticket; f.record = rec; f.bad_target = m->status_dev;
    ProfScope ps(K_FUSED, st);
    // evaluate on more slabs than half the SMs (a slab is then a single CTA): the persistent kernel overlaps a slab's epilogue
    // with the next slab's loads, conversion and main loop
    if (!train && m->eval_stream && fused_cluster_size(B, m->dims[1], m->sm_count) == 1 && (B + 127) / 128 >= m->sm_count &&
        eval_stream_supported(m->dims[0], m->dims[1], m->dims[2]))
        return launch_eval_stream(f, m->sm_count, st);
    return launch_fused_mlp(f, m->sm_count, st);
}

// Data-parallel update: the split-partial sum AND the cross-rank gradient sum happen inside the optimizer kernel, over
// peer mappings (dp_optimizer_kernel).
int run_dp_optimizer(ts_mlp *m, const int *nsplit, cudaStream_t st) {
    OptArgs o{};
    bool any_partial = false;
    int rc = fill_tensor_table(m, nsplit, o, &any_partial, st);
    if (rc) return rc;
    o.update = 1; o.grad_scale = 1.0f / (float)m->dp_world;
    const int parity = (int)(m->step & 1);
    fill_update(m, o);
    DpArgs dp{};
    dp.world = m->dp_world; dp.rank = m->dp_rank; dp.step = (uint32_t)m->step;
    dp.timeout_ns = (unsigned long long)(m->dp_timeout_s * 1e9);
    dp.err = m->dp_err_dev;
    for (int r = 0; r < m->dp_world; ++r) { dp.grads[r] = m->dp_grads[parity][r]; dp.flags[r] = m->dp_flags[r]; }
    ProfScope ps(K_OPT, st);
    if (m->nvls) {
        dp.step = ++m->dp_seq;
        o.params_mc = m->params_mc;
        o.shadow_mc = m->prec == TS_PREC_BF16 ? m->shadow_mc : nullptr;
        return launch_nvls_optimizer(o, dp, m->grads_mc, m->n_chunks, m->sm_count, st);
    }
    return launch_dp_optimizer(o, dp, m->n_chunks, m->sm_count, st);
}

// Last layer + loss on the tensor cores (bf16 mode, hidden layers wider than the fused kernel takes):
//   Z[B,16] = A_last W_last^T as a tcgen05 GEMM against the zero-padded [16, K] shadow, then one thread per row for
//   softmax / NLL / argmax / dZ (logits_head_kernel). The SIMT head_kernel spent 404 us on this at 8192 x 4096.
int tc_head_forward(ts_mlp *m, const void *a_last, const int64_t *y, int64_t B, bool train, int64_t *pred, ts_batch_record *rec,
                    cudaStream_t st) {
    const int L = m->n_layers, K = m->dims[L - 1];
    TcGemmArgs g{};
    g.A = (const __nv_bfloat16 *)a_last; g.a_major = TC_MAJOR_K; g.lda = K;
    g.B = m->shadow + m->shadow_off[L - 1]; g.b_major = TC_MAJOR_K; g.ldb = K;
    g.M = (int)B; g.N = 16; g.K = K;
    g.epi = TC_EPI_STORE_F32; g.out = m->zbuf; g.ldo = 16; g.splits = 1;
    ProfScope ps(K_HEAD, st);
    int rc = launch_tc_gemm(g, m->sm_count, st);
    if (rc) return rc;
    return launch_logits_head(m->zbuf, m->params + m->off[2 * (L - 1) + 1], y, B, m->dims[L], pred, train ? m->dz : nullptr, m->dzb,
                              m->head_partials, m->ticket, rec, m->status_dev, st);
}

// Backward of the last layer on the tensor cores: dH = (dZ W_last) * act'(A_last) / (1 - p) with the activation gradient
// in the GEMM epilogue (K = 16), dW_last^T = A_last^T dZ split over the batch, then the fold into `grads`.
int tc_head_backward(ts_mlp *m, const void *a_last, int64_t B, const uint8_t *const *masks, uint64_t seed, int *nsplit, cudaStream_t st) {
    const int L = m->n_layers, K = m->dims[L - 1], C = m->dims[L];
    int rc;
    const int tb_ = 2 * (L - 2) + 1;
    static const bool dh_gemm = [] { const char *e = getenv("TSB200_HEAD_DH_GEMM"); return e && atoi(e) != 0; }();
    if (!dh_gemm && head_dh_supported(K, C, m->partial_stride[tb_])) {
        const int rows32 = (int)((B + 31) / 32);
        float *cs = nullptr;
        if (rows32 <= kMaxRowBlocks) { cs = m->partials + m->partial_off[tb_]; nsplit[tb_] = rows32; }
        ProfScope ps(K_HEAD, st);
        rc = launch_head_dh(m->dzb, m->shadow + m->shadow_off[L - 1],
                            m->act == TS_ACT_RELU ? (const __nv_bfloat16 *)a_last : (const __nv_bfloat16 *)m->upre[L - 2], B, K, C, m->act,
                            drop_spec(m, masks, seed, L - 2, true), m->p > 0.0f, (__nv_bfloat16 *)m->dh[0], cs, m->partial_stride[tb_], st);
        if (rc) return rc;
    } else {
        TcGemmArgs g{};
        g.A = m->dzb; g.a_major = TC_MAJOR_K; g.lda = 16;
        g.B = m->shadow + m->shadow_off[L - 1]; g.b_major = TC_MAJOR_MN; g.ldb = K;
        g.M = (int)B; g.N = K; g.K = 16;
        g.epi = TC_EPI_ACTGRAD; g.out = m->dh[0]; g.ldo = K; g.splits = 1;
        g.act = m->act;
        g.A_prev = m->act == TS_ACT_RELU ? (const __nv_bfloat16 *)a_last : (const __nv_bfloat16 *)m->upre[L - 2];
        g.drop = drop_spec(m, masks, seed, L - 2, true); g.training = m->p > 0.0f;
        const int tb = 2 * (L - 2) + 1, rows4 = 4 * (int)((B + 127) / 128);
        if (rows4 <= kMaxRowBlocks) {    // db of the last hidden layer = column sums of this dH, taken in the epilogue
            g.colsum_out = m->partials + m->partial_off[tb]; g.colsum_stride = m->partial_stride[tb];
            nsplit[tb] = rows4;
        }
        ProfScope ps(K_HEAD, st);
        rc = launch_tc_gemm(g, m->bwd_sms, st);
        if (rc) return rc;
    }
    const int m_tiles = (K + 127) / 128;
    int S = m->sm_count / m_tiles;
    S = S < 1 ? 1 : (S > kMaxSplits ? kMaxSplits : S);
    int64_t kc = round_up((B + S - 1) / S, 128);
    S = (int)((B + kc - 1) / kc);
    {
        TcGemmArgs g{};
        g.A = (const __nv_bfloat16 *)a_last; g.a_major = TC_MAJOR_MN; g.lda = K;
        g.B = m->dzb; g.b_major = TC_MAJOR_MN; g.ldb = 16;
        g.M = K; g.N = 16; g.K = (int)B;
        g.epi = TC_EPI_STORE_F32; g.out = m->wt_partial; g.ldo = 16; g.splits = S; g.k_chunk = (int)kc;
        g.split_stride = (int64_t)K * 16;
        ProfScope ps(K_HEAD_WGRAD, st);
        rc = launch_tc_gemm(g, m->bwd_sms, st);
        if (rc) return rc;
        rc = launch_wlast_finish(m->wt_partial, (int64_t)K * 16, S, K, C, m->dz, B, m->grads + m->off[2 * (L - 1)],
                                 m->grads + m->off[2 * (L - 1) + 1], st);
    }
    return rc;
}

// ts_fit_backward_overlap: layer l's weight / bias gradients have been enqueued; tell the caller which of them are already
// final in `grads` (no partials left to fold) and mark the point in the stream.
int mark_layer_ready(ts_mlp *m, int l, const int *nsplit, cudaStream_t st) {
    if (m->ov_events == nullptr) return TS_OK;
    if (m->ov_final) { m->ov_final[2 * l] = nsplit[2 * l] == 0; m->ov_final[2 * l + 1] = nsplit[2 * l + 1] == 0; }
    TS_CUDA_CHECK(cudaEventRecord(m->ov_events[l], st));
    return TS_OK;
}

// The word the data-parallel kernels write when a wait for a peer times out: pinned host memory mapped into the device.
int ensure_dp_error_word(ts_mlp *m) {
    if (m->dp_err_host != nullptr) return TS_OK;
    TS_CUDA_CHECK(cudaSetDevice(m->device));
    void *h = nullptr, *d = nullptr;
    TS_CUDA_CHECK(cudaHostAlloc(&h, 64, cudaHostAllocMapped));
    memset(h, 0, 64);
    TS_CUDA_CHECK(cudaHostGetDevicePointer(&d, h, 0));
    m->dp_err_host = (uint32_t *)h;
    m->dp_err_dev = (uint32_t *)d;
    if (const char *e = getenv("TSB200_DP_TIMEOUT_S")) {
        const double v = atof(e);
        if (v > 0.0) m->dp_timeout_s = v;
    }
    return TS_OK;
}

int ensure_side_stream(ts_mlp *m) {
    if (m->dp_stream != nullptr) return TS_OK;
    TS_CUDA_CHECK(cudaSetDevice(m->device));
    int lo = 0, hi = 0;
    TS_CUDA_CHECK(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    TS_CUDA_CHECK(cudaStreamCreateWithPriority(&m->dp_stream, cudaStreamNonBlocking, hi));   // its small CTAs go first
    for (int l = 0; l < m->n_layers; ++l) TS_CUDA_CHECK(cudaEventCreateWithFlags(&m->dp_ev_layer[l], cudaEventDisableTiming));
    TS_CUDA_CHECK(cudaEventCreateWithFlags(&m->dp_ev_done, cudaEventDisableTiming));
    return TS_OK;
}

// Single GPU, layer-by-layer path: the optimizer update of layer l's tensors on the side stream, ordered behind everything
// enqueued on `st` so far, by CTAs small enough to share SMs with the backward GEMMs that are still running.
int update_layer(ts_mlp *m, int l, const int *nsplit, bool first_of_step, cudaStream_t st) {
    OptArgs o{};
    bool any_partial = false;
    int rc = fill_tensor_table(m, nsplit, o, &any_partial, st, 2 * l, 2 * l + 2);
    if (rc) return rc;
    o.update = 1; o.grad_scale = 1.0f;
    fill_update(m, o, first_of_step);
    o.shadow = m->shadow_alt;                            // the set the NEXT step reads
    o.partial_depth = 0;
    o.chunk_begin = (int)(m->off[2 * l] / kChunk);
    const int64_t chunk_end = (m->off[2 * l + 1] + round_up(m->cnt[2 * l + 1], kChunk)) / kChunk;
    TS_CUDA_CHECK(cudaEventRecord(m->dp_ev_layer[l], st));
    TS_CUDA_CHECK(cudaStreamWaitEvent(m->dp_stream, m->dp_ev_layer[l], 0));
    return launch_layer_optimizer(o, chunk_end - o.chunk_begin, m->sm_count - 16, m->dp_stream);
}

// nvls overlap: reduce + update of layer l's tensors (weight and bias) on the side stream, ordered behind everything
// enqueued on `st` so far. Called once the layer's gradients are enqueued AND nothing enqueued later reads its weights.
int exchange_layer(ts_mlp *m, int l, const int *nsplit, bool first_of_step, cudaStream_t st) {
    OptArgs o{};
    bool any_partial = false;
    int rc = fill_tensor_table(m, nsplit, o, &any_partial, st, 2 * l, 2 * l + 2);
    if (rc) return rc;
    o.update = 1; o.grad_scale = 1.0f / (float)m->dp_world;
    fill_update(m, o, first_of_step);
    o.params_mc = m->params_mc;
    o.shadow_mc = m->prec == TS_PREC_BF16 ? m->shadow_mc : nullptr;
    if (m->shadow_alt) { o.shadow = m->shadow_alt; o.shadow_mc = m->shadow_mc_alt; }   // the set the NEXT step reads
    o.partial_depth = m->nvls_coresident ? 0 : 4;
    o.chunk_begin = (int)(m->off[2 * l] / kChunk);
    const int64_t chunk_end = (m->off[2 * l + 1] + round_up(m->cnt[2 * l + 1], kChunk)) / kChunk;
    DpArgs dp{};
    dp.world = m->dp_world; dp.rank = m->dp_rank; dp.step = ++m->dp_seq;
    dp.timeout_ns = (unsigned long long)(m->dp_timeout_s * 1e9);
    dp.err = m->dp_err_dev;
    for (int r = 0; r < m->dp_world; ++r) dp.flags[r] = m->dp_flags[r];
    TS_CUDA_CHECK(cudaEventRecord(m->dp_ev_layer[l], st));
    TS_CUDA_CHECK(cudaStreamWaitEvent(m->dp_stream, m->dp_ev_layer[l], 0));
    return launch_nvls_layer(o, dp, m->grads_mc, chunk_end - o.chunk_begin, m->nvls_overlap_ctas, m->dp_stream);
}

int fit_impl(ts_mlp *m, const void *x, int x_dtype, const int64_t *y, int64_t B, const uint8_t *const *masks,
             uint64_t seed, int64_t *pred, ts_batch_record *rec, cudaStream_t st, bool update) {
    int rc = check_ready(m, B, update);
    if (rc) return rc;
    TS_REQUIRE(x && y && pred && rec, TS_E_INVALID, "null inputs/targets/predictions/record pointer");
    TS_CUDA_CHECK(cudaSetDevice(m->device));
    const int L = m->n_layers;
    const bool bf16 = m->prec == TS_PREC_BF16;
    const bool dp = m->dp_world > 1 && update;
    TS_REQUIRE(!dp || m->dp_err_host == nullptr || *(volatile uint32_t *)m->dp_err_host == 0u, TS_E_TIMEOUT,
               "data parallel: a peer rank did not arrive within %.0f s at step %u; that step's update was skipped and the "
               "replicas are no longer in step (ts_dp_set_timeout / TSB200_DP_TIMEOUT_S set the bound)",
               m->dp_timeout_s, *(volatile uint32_t *)m->dp_err_host);
    if (dp && !m->nvls) m->grads = m->dp_grads[m->step & 1][m->dp_rank];     // this step's peer-visible gradient buffer
    const void *a0 = nullptr, *a_last = nullptr;
    int nsplit[2 * kMaxLayers] = {0};
    if (m->fused && m->step_kernel && update && !dp && (x_dtype == TS_DT_F32 || x_dtype == TS_DT_BF16) &&
        fit_step_supported(m->dims[0], m->dims[1], m->dims[2], B)) {
        // small batch (the size the reference ships): forward, loss, backward and the optimizer update in ONE kernel
        FitStepArgs f{};
        if (x_dtype == TS_DT_F32) f.X_f32 = (const float *)x; else f.X = (const __nv_bfloat16 *)x;
        f.W1 = m->shadow + m->shadow_off[0]; f.W2p = m->shadow + m->shadow_off[1];
        f.y = y; f.B = B; f.D = m->dims[0]; f.H = m->dims[1]; f.C = m->dims[2];
        f.pred = pred; f.record = rec; f.bad_target = m->status_dev;
        f.drop = drop_spec(m, masks, seed, 0, true);
        f.training = m->p > 0.0f;
        f.off_w1 = m->off[0]; f.off_b1 = m->off[1]; f.off_w2 = m->off[2]; f.off_b2 = m->off[3];
        f.sh_w1 = m->shadow_off[0]; f.sh_w2 = m->shadow_off[1];
        bool any_partial = false;
        rc = fill_tensor_table(m, nullptr, f.opt, &any_partial, st);
        if (rc) return rc;
        f.opt.update = 1; f.opt.grad_scale = 1.0f;
        fill_update(m, f.opt);
        const unsigned int target = m->step_flag + (unsigned int)(m->dims[1] / 64);   // every CTA of the compute cluster adds one
        f.flag = m->ticket + 16; f.flag_target = target;
        f.n_chunks = (int)m->n_chunks; f.sm_count = m->sm_count;
        ProfScope ps(K_STEP, st);
        rc = launch_fit_step(f, st);
        if (rc == TS_OK) m->step_flag = target;                // a launch that did not happen must not move the counter's target
        return rc;
    }
    if (m->fused) {
        rc = fused_forward(m, x, x_dtype, y, B, masks, seed, true, pred, rec, &a0, st);
        if (rc) return rc;
        a_last = m->acts[0];
    } else {
        rc = forward_hidden(m, x, x_dtype, B, masks, seed, true, &a0, &a_last, st);
        if (rc) return rc;
        if (m->tc_head) {
            rc = tc_head_forward(m, a_last, y, B, true, pred, rec, st);
            if (rc) return rc;
            rc = tc_head_backward(m, a_last, B, masks, seed, nsplit, st);
        } else {
            HeadArgs h{};
            fill_head(m, h, a_last, y, B, pred, rec);
            h.dz = m->dz; h.dH = m->dh[0];
            h.U_prev = m->upre[L - 2];
            h.drop = drop_spec(m, masks, seed, L - 2, true);
            ProfScope ps(K_HEAD, st);
            rc = launch_head(h, m->sm_count, st);
        }
        if (rc) return rc;
    }

    const int64_t row_blocks = (B + 127) / 128;
    const bool tc_head = m->tc_head && !m->fused;      // last-layer gradients already in `grads` (tc_head_backward)
    const bool small = !tc_head && m->classes() <= 16 && row_blocks <= kMaxRowBlocks;   // one launch for dW_last, db_last, db of the last hidden layer
    const bool in_kernel = m->fused && fused_grads_supported(m->dims[1], m->dims[2], B);   // the fused kernel already left them
    if (in_kernel) {
        nsplit[2] = nsplit[3] = nsplit[1] = (int)row_blocks;
        if (m->wgrad_update && update && (!dp || !m->nvls) && m->ov_events == nullptr && row_blocks <= kReduceAbove &&
            wgrad_update_supported(m->dims[0], m->dims[1], B)) {
            // one kernel for the rest of the step: dW1 with the batch split over a cluster, reduced through distributed shared
            // memory and applied by the CTA holding the sum; its last cluster folds and updates the small tensors
            WgradUpdateArgs w{};
            w.dH = (const __nv_bfloat16 *)m->dh[0]; w.X = (const __nv_bfloat16 *)a0;
            w.B = B; w.D = m->dims[0]; w.H = m->dims[1];
            w.off_w1 = m->off[0]; w.sh_w1 = m->shadow_off[0];
            w.small_chunk_begin = (int)(m->off[1] / kChunk);
            w.small_chunks = (int)(m->n_chunks - m->off[1] / kChunk);
            bool any_partial = false;
            rc = fill_tensor_table(m, nsplit, w.opt, &any_partial, st);
            if (rc) return rc;
            w.opt.update = 1; w.opt.grad_scale = dp ? 1.0f / (float)m->dp_world : 1.0f;
            const int parity = (int)(m->step & 1);     // which of the two peer-visible gradient buffers this step uses
            fill_update(m, w.opt);
            w.sm_count = m->sm_count;
            w.dp.world = 1;
            if (dp) {   // 'p2p' data parallel: the cross-rank gradient sum happens inside this kernel
                w.dp.world = m->dp_world; w.dp.rank = m->dp_rank; w.dp.step = (uint32_t)m->step;
                w.dp.timeout_ns = (unsigned long long)(m->dp_timeout_s * 1e9);
                w.dp.err = m->dp_err_dev;
                for (int r = 0; r < m->dp_world; ++r) { w.dp.grads[r] = m->dp_grads[parity][r]; w.dp.flags[r] = m->dp_flags[r]; }
                if (m->ll_bound) {
                    for (int r = 0; r < m->dp_world; ++r) w.dp.ll_peer[r] = reinterpret_cast<float4 *>(m->ll_inbox[r] + parity * kLlParityBytes);
                    w.dp.ll_own = reinterpret_cast<const float4 *>(m->ll_inbox[m->dp_rank] + parity * kLlParityBytes);
                    static const bool two_hops = [] { const char *e = getenv("TSB200_DP_LL"); return e && atoi(e) == 2; }();   // A/B runs
                    w.dp.ll_direct = m->dp_world == 2 && !two_hops;
                }
            }
            ProfScope ps(K_WGRAD_UPD, st);
            return launch_wgrad_update(w, st);
        }
    } else if (tc_head) {
        // nothing to launch here
    } else if (small) {
        const int tw = 2 * (L - 1), tb = tw + 1, th = 2 * (L - 2) + 1;
        const bool direct = row_blocks == 1;
        float *wp = direct ? m->grads + m->off[tw] : m->partials + m->partial_off[tw];
        float *bp = direct ? m->grads + m->off[tb] : m->partials + m->partial_off[tb];
        float *hp = direct ? m->grads + m->off[th] : m->partials + m->partial_off[th];
        ProfScope ps(K_HEAD_WGRAD, st);
        rc = launch_small_grads(a_last, m->dh[0], bf16, m->dz, B, m->dims[L - 1], m->dims[L], wp, m->partial_stride[tw], bp,
                                m->partial_stride[tb], hp, m->partial_stride[th], st);
        if (rc) return rc;
        nsplit[tw] = nsplit[tb] = nsplit[th] = direct ? 0 : (int)row_blocks;
    } else {   // last layer: dW = dz^T a_last, db = colsum(dz)
        const int tw = 2 * (L - 1), tb = tw + 1;
        int kc;
        const int S = splits_for(B, kMaxSplits, 1, &kc);
        float *wp = S > 1 ? m->partials + m->partial_off[tw] : m->grads + m->off[tw];
        float *bp = S > 1 ? m->partials + m->partial_off[tb] : m->grads + m->off[tb];
        ProfScope ps(K_HEAD_WGRAD, st);
        rc = launch_head_wgrad(a_last, bf16, m->dz, B, m->dims[L - 1], m->dims[L], wp, m->partial_stride[tw], bp,
                               m->partial_stride[tb], S, st);
        if (rc) return rc;
        nsplit[tw] = nsplit[tb] = S > 1 ? S : 0;
    }
    rc = mark_layer_ready(m, L - 1, nsplit, st);
    if (rc) return rc;
    // nvls overlap (wide nets, tensor-core last layer): W_last was last read by tc_head_backward, already enqueued
    const bool ov = dp && m->nvls && m->nvls_overlap_ctas > 0 && tc_head;
    // single GPU: the same schedule with a plain per-layer optimizer kernel (needs the second shadow set of the workspace)
    const bool ov_local = !dp && update && tc_head && m->local_overlap && m->shadow_alt != nullptr && !m->shadow_ext &&
                          m->ov_events == nullptr;
    if (ov) {
        rc = exchange_layer(m, L - 1, nsplit, true, st);
        if (rc) return rc;
    }
    if (ov_local) {
        rc = ensure_side_stream(m);
        if (rc) return rc;
        rc = update_layer(m, L - 1, nsplit, true, st);
        if (rc) return rc;
    }
    int cur = 0;
    for (int l = L - 2; l >= 0; --l) {
        const int in = m->dims[l], out = m->dims[l + 1];
        const void *dHl = m->dh[cur];
        const void *a_prev = (l == 0) ? a0 : m->acts[l - 1];
        const int tw = 2 * l, tb = tw + 1;
        if (!((small || in_kernel) && l == L - 2) && nsplit[tb] == 0) {   // db_l = colsum(dH_l), unless it came with the kernel that produced dH_l
            int kc;
            const int S = splits_for(B, kMaxSplits, 1, &kc);
            float *bp = S > 1 ? m->partials + m->partial_off[tb] : m->grads + m->off[tb];
            ProfScope ps(K_COLSUM, st);
            rc = launch_colsum(dHl, bf16, B, out, bp, m->partial_stride[tb], S, st);
            if (rc) return rc;
            nsplit[tb] = S > 1 ? S : 0;
        }
        {   // dW_l[out,in] = dH_l^T a_prev, batch is the contraction (split across CTAs)
            int kc;
            const int S = splits_for(B, m->max_split[tw], bf16 ? 128 : 16, &kc);
            float *wp = S > 1 ? m->partials + m->partial_off[tw] : m->grads + m->off[tw];
            ProfScope ps(K_DW_GEMM, st);
            if (!bf16) {
                SgemmArgs g{};
                g.A = (const float *)dHl; g.sam = 1; g.sak = out;
                g.B = (const float *)a_prev; g.sbk = in; g.sbn = 1;
                g.C = wp; g.ldc = in; g.M = out; g.N = in; g.K = (int)B;
                g.k_chunk = kc; g.c_split_stride = m->partial_stride[tw];
                rc = launch_sgemm(EPI_STORE, g, S, st);
            } else {
                TcGemmArgs g{};
                g.A = (const __nv_bfloat16 *)dHl; g.a_major = TC_MAJOR_MN; g.lda = out;
                g.B = (const __nv_bfloat16 *)a_prev; g.b_major = TC_MAJOR_MN; g.ldb = in;
                g.M = out; g.N = in; g.K = (int)B;
                g.epi = TC_EPI_STORE_F32; g.out = wp; g.ldo = in; g.splits = S; g.k_chunk = kc; g.split_stride = m->partial_stride[tw];
                rc = launch_tc_gemm(g, m->bwd_sms, st);
            }
            if (rc) return rc;
            nsplit[tw] = S > 1 ? S : 0;
        }
        rc = mark_layer_ready(m, l, nsplit, st);
        if (rc) return rc;
        // nvls overlap with two shadow sets: layer l's exchange can start as soon as dW_l and db_l are enqueued -- the dX GEMM
        // below still reads W_l, but from this step's set, which the update does not write
        const bool ov_early = ov && m->shadow_alt != nullptr;
        if (ov_early) {
            rc = exchange_layer(m, l, nsplit, false, st);
            if (rc) return rc;
        }
        if (ov_local) {
            rc = update_layer(m, l, nsplit, false, st);
            if (rc) return rc;
        }
        if (l > 0) {   // dH_{l-1} = (dH_l W_l) * act'(layer l-1) * keep/(1-p)
            const DropoutSpec ds = drop_spec(m, masks, seed, l - 1, true);
            ProfScope ps(K_DX_GEMM, st);
            if (!bf16) {
                SgemmArgs g{};
                g.A = (const float *)dHl; g.sam = out; g.sak = 1;
                g.B = m->params + m->off[tw]; g.sbk = in; g.sbn = 1;
                g.C = (float *)m->dh[cur ^ 1]; g.ldc = in; g.M = (int)B; g.N = in; g.K = out;
                g.k_chunk = (int)round_up(out, 16); g.c_split_stride = 0;
                g.A_prev = (const float *)m->acts[l - 1]; g.U_prev = m->upre[l - 1];
                g.drop = ds; g.act = m->act; g.training = m->p > 0.0f;
                rc = launch_sgemm(EPI_ACTGRAD, g, 1, st);
            } else {
                TcGemmArgs g{};
                g.A = (const __nv_bfloat16 *)dHl; g.a_major = TC_MAJOR_K; g.lda = out;
                g.B = m->shadow + m->shadow_off[l]; g.b_major = TC_MAJOR_MN; g.ldb = in;
                g.M = (int)B; g.N = in; g.K = out;
                g.epi = TC_EPI_ACTGRAD; g.out = m->dh[cur ^ 1]; g.ldo = in; g.splits = 1;
                g.act = m->act;
                g.A_prev = m->act == TS_ACT_RELU ? (const __nv_bfloat16 *)m->acts[l - 1] : (const __nv_bfloat16 *)m->upre[l - 1];
                const int tbp = 2 * (l - 1) + 1, rows4 = 4 * (int)((B + 127) / 128);
                if (rows4 <= kMaxRowBlocks) {    // db of layer l-1 = column sums of the dH this GEMM stores
                    g.colsum_out = m->partials + m->partial_off[tbp]; g.colsum_stride = m->partial_stride[tbp];
                    nsplit[tbp] = rows4;
                }
                g.drop = ds; g.training = m->p > 0.0f;
                rc = launch_tc_gemm(g, m->bwd_sms, st);
            }
            if (rc) return rc;
            cur ^= 1;
        }
        if (ov && !ov_early) {   // one shadow set: wait until the dX GEMM above, the last reader of W_l in this step, is enqueued
            rc = exchange_layer(m, l, nsplit, false, st);
            if (rc) return rc;
        }
    }
    if (ov || ov_local) {       // the next forward pass reads every parameter: the compute stream waits for the last exchange / update
        TS_CUDA_CHECK(cudaEventRecord(m->dp_ev_done, m->dp_stream));
        TS_CUDA_CHECK(cudaStreamWaitEvent(st, m->dp_ev_done, 0));
        if (m->shadow_alt) {   // the set just written becomes the one the next step reads
            std::swap(m->shadow, m->shadow_alt);
            std::swap(m->shadow_mc, m->shadow_mc_alt);
        }
        return TS_OK;
    }
    if (!dp) return run_optimizer(m, nsplit, update, 1.0f, st);
    return run_dp_optimizer(m, nsplit, st);
}

}  // namespace

// ======================================================================================= C ABI
extern "C" {

int32_t ts_abi_version(void) { return TS_ABI_VERSION; }
const char *ts_last_error(void) { return g_err; }
int64_t ts_launch_count(void) { return g_launch_count; }

int32_t ts_mlp_create(ts_mlp **out, const int32_t *dims, int32_t n_dims, int32_t activation, float dropout_p,
                      int32_t precision, int32_t device) {
    TS_REQUIRE(out && dims, TS_E_INVALID, "null out/dims");
    *out = nullptr;
    TS_REQUIRE(n_dims >= 3 && n_dims <= kMaxLayers + 1, TS_E_INVALID,
               "n_dims = %d: need input, at least one hidden layer, classes (3..%d entries)", n_dims, kMaxLayers + 1);
    for (int i = 0; i < n_dims; ++i) TS_REQUIRE(dims[i] >= 1, TS_E_INVALID, "dims[%d] = %d", i, dims[i]);
    TS_REQUIRE(activation == TS_ACT_RELU || activation == TS_ACT_GELU, TS_E_INVALID, "unknown activation %d", activation);
    TS_REQUIRE(dropout_p >= 0.0f && dropout_p < 1.0f, TS_E_INVALID, "dropout %g outside [0,1)", dropout_p);
    TS_REQUIRE(precision == TS_PREC_F32 || precision == TS_PREC_BF16, TS_E_INVALID, "unknown precision %d", precision);
    TS_REQUIRE(dims[n_dims - 1] <= kMaxClasses, TS_E_UNSUPPORTED, "at most %d classes", kMaxClasses);
    for (int i = 1; i < n_dims - 1; ++i)
        TS_REQUIRE(dims[i] % 8 == 0, TS_E_UNSUPPORTED, "hidden width %d must be a multiple of 8", dims[i]);
    if (precision == TS_PREC_BF16) {
        TS_REQUIRE(activation == TS_ACT_RELU || dims[n_dims - 1] <= 16, TS_E_UNSUPPORTED,
                   "TS_PREC_BF16 with GELU needs at most 16 classes (tensor-core last layer); use TS_PREC_F32");
        TS_REQUIRE(dims[0] % 8 == 0, TS_E_UNSUPPORTED, "TS_PREC_BF16 needs an input width that is a multiple of 8 (TMA pitch)");
    }
    int n_dev = 0;
    TS_CUDA_CHECK(cudaGetDeviceCount(&n_dev));
    TS_REQUIRE(device >= 0 && device < n_dev, TS_E_INVALID, "device %d of %d", device, n_dev);
    int major = 0, sms = 0;
    TS_CUDA_CHECK(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
    TS_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    TS_REQUIRE(major == 10, TS_E_UNSUPPORTED, "this library is built for sm_100a only; device %d is sm_%dx", device, major);

    ts_mlp *m = new (std::nothrow) ts_mlp();
    TS_REQUIRE(m, TS_E_INVALID, "out of host memory");
    m->n_layers = n_dims - 1;
    for (int i = 0; i < n_dims; ++i) m->dims[i] = dims[i];
    m->act = activation; m->p = dropout_p; m->scale = 1.0f / (float)(1.0 - (double)dropout_p);
    m->prec = precision; m->device = device; m->sm_count = sms; m->bwd_sms = sms;
    if (const char *e = getenv("TSB200_INKERNEL_CAST")) m->inkernel_cast = atoi(e) != 0 ? 1 : 0;
    if (const char *e = getenv("TSB200_STEP_KERNEL")) m->step_kernel = atoi(e) != 0 ? 1 : 0;
    if (const char *e = getenv("TSB200_WGRAD_UPDATE")) m->wgrad_update = atoi(e) != 0 ? 1 : 0;
    if (const char *e = getenv("TSB200_EVAL_STREAM")) m->eval_stream = atoi(e) != 0 ? 1 : 0;
    if (const char *e = getenv("TSB200_OPT_OVERLAP")) m->local_overlap = atoi(e) != 0 ? 1 : 0;
    m->n_tensors = 2 * m->n_layers;
    int64_t o = 0, so = 0;
    for (int l = 0; l < m->n_layers; ++l) {
        const int64_t in = dims[l], outw = dims[l + 1];
        m->off[2 * l] = o; m->cnt[2 * l] = in * outw; o += round_up(in * outw, kChunk);
        m->off[2 * l + 1] = o; m->cnt[2 * l + 1] = outw; o += round_up(outw, kChunk);
        // split-K budget: enough CTAs for two waves, none for matrices that already fill the machine
        const int64_t tiles = ((outw + 127) / 128) * ((in + 127) / 128);
        int64_t s = sms / tiles;   // one wave of CTAs
        s = s < 1 ? 1 : (s > kMaxSplits ? kMaxSplits : s);
        m->max_split[2 * l] = (l == m->n_layers - 1) ? kMaxRowBlocks : (int)s;
        // bf16 mode: hidden bias gradients come out of the dH-producing GEMM epilogues, 4 partial rows per 128-row tile
        m->max_split[2 * l + 1] = (l >= m->n_layers - 2 || precision == TS_PREC_BF16) ? kMaxRowBlocks : kMaxSplits;
        m->shadow_off[l] = so;
        so += round_up(in * (l < m->n_layers - 1 ? outw : (outw > 16 ? outw : 16)), 128);   // last layer: at least 16 class rows
    }
    m->shadow_total = so;
    m->fused = precision == TS_PREC_BF16 && m->n_layers == 2 && activation == TS_ACT_RELU && fused_mlp_supported(dims[0], dims[1], dims[2]);
    m->tc_head = precision == TS_PREC_BF16 && !m->fused && dims[n_dims - 1] <= 16;
    if (const char *e = getenv("TSB200_SIMT_HEAD"))      // A/B switch for measurements (GELU always takes the tensor-core last layer)
        m->tc_head = m->tc_head && (atoi(e) == 0 || activation == TS_ACT_GELU);
    {   // status word the head kernels can reach: pinned host memory mapped into the device
        void *h = nullptr, *d = nullptr;
        if (cudaSetDevice(device) == cudaSuccess && cudaHostAlloc(&h, 64, cudaHostAllocMapped) == cudaSuccess) {
            memset(h, 0, 64);
            if (cudaHostGetDevicePointer(&d, h, 0) == cudaSuccess) { m->status_host = (uint32_t *)h; m->status_dev = (uint32_t *)d; }
            else cudaFreeHost(h);
        }
        (void)cudaGetLastError();
    }
    m->total = o;
    m->n_chunks = o / kChunk;
    m->chunk_tensor_host.assign((size_t)m->n_chunks, -1);
    for (int t = 0; t < m->n_tensors; ++t)
        for (int64_t c = m->off[t] / kChunk; c < (m->off[t] + round_up(m->cnt[t], kChunk)) / kChunk; ++c)
            m->chunk_tensor_host[(size_t)c] = t;
    *out = m;
    return TS_OK;
}

int32_t ts_mlp_destroy(ts_mlp *m) {
    if (m && m->status_host) cudaFreeHost(m->status_host);
    if (m && m->dp_err_host) cudaFreeHost(m->dp_err_host);
    if (m && m->dp_stream) {
        cudaStreamSynchronize(m->dp_stream);
        for (int l = 0; l < m->n_layers; ++l) cudaEventDestroy(m->dp_ev_layer[l]);
        cudaEventDestroy(m->dp_ev_done);
        cudaStreamDestroy(m->dp_stream);
    }
    delete m;
    return TS_OK;
}

int32_t ts_mlp_status(const ts_mlp *m) {
    TS_REQUIRE(m, TS_E_INVALID, "null handle");
    return target_status(m);
}

int32_t ts_mlp_clear_status(ts_mlp *m) {
    TS_REQUIRE(m, TS_E_INVALID, "null handle");
    if (m->status_host) *(volatile uint32_t *)m->status_host = 0u;
    return TS_OK;
}

int32_t ts_mlp_param_layout(const ts_mlp *m, int64_t *offsets, int64_t *total) {
    TS_REQUIRE(m && offsets && total, TS_E_INVALID, "null argument");
    for (int t = 0; t < m->n_tensors; ++t) offsets[t] = m->off[t];
    *total = m->total;
    return TS_OK;
}

int32_t ts_mlp_bind_params(ts_mlp *m, float *params, float *grads) {
    TS_REQUIRE(m && params && grads, TS_E_INVALID, "null argument");
    TS_REQUIRE(((uintptr_t)params % 256 == 0) && ((uintptr_t)grads % 256 == 0), TS_E_INVALID,
               "flat parameter/gradient buffers must be 256-byte aligned");
    m->params = params;
    m->grads = grads;
    return TS_OK;
}

int32_t ts_opt_bind(ts_mlp *m, int32_t kind, const ts_opt_hyper *hyper, float *exp_avg, float *exp_avg_sq,
                    float *max_exp_avg_sq, int64_t step) {
    TS_REQUIRE(m && hyper, TS_E_INVALID, "null argument");
    TS_REQUIRE(kind == TS_OPT_SGD || kind == TS_OPT_ADAM, TS_E_INVALID, "unknown optimizer kind %d", kind);
    TS_REQUIRE(kind == TS_OPT_SGD || (exp_avg && exp_avg_sq), TS_E_INVALID, "Adam needs exp_avg and exp_avg_sq buffers");
    TS_REQUIRE(kind != TS_OPT_ADAM || !hyper->amsgrad || max_exp_avg_sq, TS_E_INVALID, "amsgrad needs the max_exp_avg_sq buffer");
    TS_REQUIRE(kind != TS_OPT_SGD || hyper->momentum == 0.0 || exp_avg, TS_E_INVALID,
               "SGD with momentum needs the momentum buffer (exp_avg argument)");
    TS_REQUIRE(hyper->weight_decay >= 0.0 && hyper->momentum >= 0.0, TS_E_INVALID, "negative weight_decay or momentum");
    TS_REQUIRE(!hyper->nesterov || (hyper->momentum > 0.0 && hyper->dampening == 0.0), TS_E_INVALID,
               "Nesterov momentum requires a momentum and zero dampening");     // torch/optim/sgd.py __init__
    TS_REQUIRE(step >= 0, TS_E_INVALID, "negative step");
    m->opt_kind = kind; m->hyper = *hyper; m->exp_avg = exp_avg; m->exp_avg_sq = exp_avg_sq; m->max_exp_avg_sq = max_exp_avg_sq;
    m->step = step;
    m->opt_bound = true;
    return TS_OK;
}

int64_t ts_opt_step_count(const ts_mlp *m) { return m ? m->step : -1; }

int32_t ts_dp_bind(ts_mlp *m, int32_t rank, int32_t world, float *const *grad_bufs, uint32_t *const *flag_bufs) {
    TS_REQUIRE(m && grad_bufs && flag_bufs, TS_E_INVALID, "null argument");
    TS_REQUIRE(world >= 1 && world <= kMaxRanks && rank >= 0 && rank < world, TS_E_INVALID, "rank %d of %d (at most %d ranks)", rank,
               world, kMaxRanks);
    for (int b = 0; b < 2; ++b)
        for (int r = 0; r < world; ++r) {
            TS_REQUIRE(grad_bufs[b * world + r] != nullptr && ((uintptr_t)grad_bufs[b * world + r] % 16) == 0, TS_E_INVALID,
                       "gradient buffer %d of rank %d is null or not 16-byte aligned", b, r);
            m->dp_grads[b][r] = grad_bufs[b * world + r];
        }
    for (int r = 0; r < world; ++r) {
        TS_REQUIRE(flag_bufs[r] != nullptr, TS_E_INVALID, "flag array of rank %d is null", r);
        m->dp_flags[r] = flag_bufs[r];
    }
    m->dp_world = world;
    m->dp_rank = rank;
    return ensure_dp_error_word(m);
}

int32_t ts_dp_bind_ll(ts_mlp *m, void *const *inbox_bufs, int64_t inbox_bytes) {
    TS_REQUIRE(m && inbox_bufs, TS_E_INVALID, "null argument");
    TS_REQUIRE(m->dp_world >= 2 && !m->nvls, TS_E_STATE, "ts_dp_bind_ll follows ts_dp_bind on a world of at least 2 ranks");
    TS_REQUIRE(inbox_bytes >= 2 * kLlParityBytes, TS_E_INVALID, "inbox of %lld bytes, TS_DP_LL_INBOX_BYTES = %lld", (long long)inbox_bytes,
               (long long)(2 * kLlParityBytes));
    for (int r = 0; r < m->dp_world; ++r) {
        TS_REQUIRE(inbox_bufs[r] != nullptr && ((uintptr_t)inbox_bufs[r] % 16) == 0, TS_E_INVALID, "inbox of rank %d null or not 16-byte aligned", r);
        m->ll_inbox[r] = static_cast<char *>(inbox_bufs[r]);
    }
    static const bool off = [] { const char *e = getenv("TSB200_DP_LL"); return e && atoi(e) == 0; }();
    m->ll_bound = !off;
    return TS_OK;
}

int32_t ts_dp_set_timeout(ts_mlp *m, double seconds) {
    TS_REQUIRE(m, TS_E_INVALID, "null handle");
    TS_REQUIRE(seconds > 0.0 && seconds < 1e7, TS_E_INVALID, "timeout %g s", seconds);
    m->dp_timeout_s = seconds;
    return TS_OK;
}

int32_t ts_dp_status(const ts_mlp *m) {
    TS_REQUIRE(m, TS_E_INVALID, "null handle");
    TS_REQUIRE(m->dp_err_host == nullptr || *(volatile uint32_t *)m->dp_err_host == 0u, TS_E_TIMEOUT,
               "data parallel: a peer rank did not arrive within %.0f s at step %u", m->dp_timeout_s, *(volatile uint32_t *)m->dp_err_host);
    return TS_OK;
}

int64_t ts_mlp_shadow_elems(const ts_mlp *m) { return m ? (m->prec == TS_PREC_BF16 ? m->shadow_total : 0) : -1; }

int32_t ts_dp_bind_nvls(ts_mlp *m, int32_t rank, int32_t world, float *grads_local, const float *grads_mc, float *params_local,
                        float *params_mc, void *shadow_local, void *shadow_mc, uint32_t *const *flag_bufs) {
    TS_REQUIRE(m && grads_local && grads_mc && params_local && params_mc && flag_bufs, TS_E_INVALID, "null argument");
    TS_REQUIRE(world >= 1 && world <= kMaxRanks && rank >= 0 && rank < world, TS_E_INVALID, "rank %d of %d (at most %d ranks)", rank,
               world, kMaxRanks);
    TS_REQUIRE(m->prec != TS_PREC_BF16 || (shadow_local && shadow_mc), TS_E_INVALID, "bf16 mode needs the symmetric shadow buffer");
    for (const void *q : {(const void *)grads_local, (const void *)grads_mc, (const void *)params_local, (const void *)params_mc,
                          (const void *)shadow_local, (const void *)shadow_mc})
        TS_REQUIRE(((uintptr_t)q % 256) == 0, TS_E_INVALID, "buffers must be 256-byte aligned");
    for (int r = 0; r < world; ++r) {
        TS_REQUIRE(flag_bufs[r] != nullptr, TS_E_INVALID, "flag array of rank %d is null", r);
        m->dp_flags[r] = flag_bufs[r];
    }
    m->params = params_local; m->grads = grads_local;
    m->grads_mc = grads_mc; m->params_mc = params_mc;
    m->shadow_ext = (__nv_bfloat16 *)shadow_local; m->shadow_mc = (__nv_bfloat16 *)shadow_mc;
    if (m->shadow_ext) m->shadow = m->shadow_ext;
    m->shadow_alt = nullptr; m->shadow_mc_alt = nullptr;     // the workspace's second set is not symmetric memory (ts_dp_nvls_second_shadow)
    m->dp_world = world; m->dp_rank = rank; m->nvls = true;
    // Wide nets (layer-by-layer tensor-core path): exchange each layer on a side stream while the backward pass of the
    // layers below is still running, on TSB200_NVLS_OVERLAP_SMS SMs (default: 20 from 4 ranks up, 0 = off below) that the backward GEMMs leave free.
    // (128 = 148 - 20 GEMM CTAs walk the 1024 tiles of a 8192 x 4096 layer in exactly 8 waves, as many as 132 would)
    // TSB200_NVLS_OVERLAP_SMS = -1 (the default): co-resident -- the GEMM kernels are capped at 128 registers per thread, so a
    // 256-thread exchange CTA fits on the same SM next to a GEMM CTA and no SM is reserved (L model, 8,192 rows per GPU: 2.50 ms
    // per step on 8 B200s, 2.51 ms on 2); = n > 0: n fat exchange CTAs on n SMs taken from the backward GEMMs (2.79 ms on 8 GPUs,
    // 3.34 ms on 2); 0: off, one whole-buffer kernel after the backward pass (3.02 / 3.05 ms).
    int ov_sms = world >= 2 ? -1 : 0;
    if (const char *e = getenv("TSB200_NVLS_OVERLAP_SMS")) ov_sms = atoi(e);
    m->nvls_coresident = ov_sms < 0;
    if (ov_sms < 0) ov_sms = m->sm_count - 16;
    if (world > 1 && ov_sms > 0 && (m->nvls_coresident || ov_sms < m->sm_count / 2) && m->tc_head && !m->fused && m->n_layers > 2) {
        int rc = ensure_side_stream(m);
        if (rc) return rc;
        m->nvls_overlap_ctas = ov_sms;
        m->bwd_sms = m->nvls_coresident ? m->sm_count : m->sm_count - ov_sms;
    }
    return ensure_dp_error_word(m);
}

int32_t ts_dp_nvls_second_shadow(ts_mlp *m, void *shadow2_local, void *shadow2_mc) {
    TS_REQUIRE(m && shadow2_local && shadow2_mc, TS_E_INVALID, "null argument");
    TS_REQUIRE(m->nvls && m->shadow_ext, TS_E_STATE, "call ts_dp_bind_nvls (bf16 mode) first");
    TS_REQUIRE(((uintptr_t)shadow2_local % 256) == 0 && ((uintptr_t)shadow2_mc % 256) == 0, TS_E_INVALID, "buffers must be 256-byte aligned");
    m->shadow_alt = (__nv_bfloat16 *)shadow2_local;
    m->shadow_mc_alt = (__nv_bfloat16 *)shadow2_mc;
    return TS_OK;
}

int64_t ts_mlp_workspace_bytes(const ts_mlp *m, int64_t max_batch) {
    if (!m || max_batch < 1) { set_error("ts_mlp_workspace_bytes: bad argument"); return TS_E_INVALID; }
    ts_mlp tmp = *m;   // carve() writes pointers; measure on a copy
    return carve(&tmp, nullptr, max_batch);
}

int32_t ts_mlp_set_workspace(ts_mlp *m, void *workspace, int64_t bytes, int64_t max_batch) {
    TS_REQUIRE(m && workspace, TS_E_INVALID, "null argument");
    TS_REQUIRE((uintptr_t)workspace % 256 == 0, TS_E_INVALID, "workspace must be 256-byte aligned");
    TS_REQUIRE(max_batch >= 1, TS_E_INVALID, "max_batch %lld", (long long)max_batch);
    const int64_t need = ts_mlp_workspace_bytes(m, max_batch);
    TS_REQUIRE(bytes >= need, TS_E_INVALID, "workspace of %lld bytes < %lld needed for batch %lld", (long long)bytes,
               (long long)need, (long long)max_batch);
    TS_CUDA_CHECK(cudaSetDevice(m->device));
    carve(m, workspace, max_batch);
    m->max_batch = max_batch;
    // one-time set-up copies (synchronous): chunk map, ticket
    TS_CUDA_CHECK(cudaMemcpy(m->d_chunk_tensor, m->chunk_tensor_host.data(), (size_t)m->n_chunks * 4, cudaMemcpyHostToDevice));
    TS_CUDA_CHECK(cudaMemset(m->ticket, 0, 256));
    m->step_flag = 0;
    if (m->lut_set) TS_CUDA_CHECK(cudaMemcpy(m->d_lut, m->lut_host, sizeof(m->lut_host), cudaMemcpyHostToDevice));
    m->ws_set = true;
    return TS_OK;
}

int32_t ts_mlp_set_input_lut(ts_mlp *m, const float *lut256) {
    TS_REQUIRE(m && lut256, TS_E_INVALID, "null argument");
    memcpy(m->lut_host, lut256, sizeof(m->lut_host));
    m->lut_set = true;
    if (m->ws_set) {
        TS_CUDA_CHECK(cudaSetDevice(m->device));
        TS_CUDA_CHECK(cudaDeviceSynchronize());          // a batch converted with the old table may still be in flight
        TS_CUDA_CHECK(cudaMemcpy(m->d_lut, m->lut_host, sizeof(m->lut_host), cudaMemcpyHostToDevice));
    }
    return TS_OK;
}

int32_t ts_mlp_sync_weights(ts_mlp *m, void *stream) {
    TS_REQUIRE(m, TS_E_INVALID, "null handle");
    if (m->prec != TS_PREC_BF16) return TS_OK;
    TS_REQUIRE(m->params && m->ws_set, TS_E_STATE, "bind parameters and workspace before ts_mlp_sync_weights");
    TS_CUDA_CHECK(cudaSetDevice(m->device));
    for (__nv_bfloat16 *set : {m->shadow, m->shadow_alt}) {
        if (set == nullptr) continue;
        TS_CUDA_CHECK(cudaMemsetAsync(set, 0, (size_t)m->shadow_total * 2, (cudaStream_t)stream));   // padded class rows stay zero
        for (int l = 0; l < m->n_layers; ++l) {
            int rc = launch_cast_bf16(m->params + m->off[2 * l], set + m->shadow_off[l], m->cnt[2 * l], (cudaStream_t)stream);
            if (rc) return rc;
        }
    }
    return TS_OK;
}

int32_t ts_fit_step(ts_mlp *m, const void *inputs, int32_t inputs_dtype, const int64_t *targets, int64_t batch,
                    const uint8_t *const *masks, uint64_t seed, int64_t *predictions, ts_batch_record *record, void *stream) {
    return fit_impl(m, inputs, inputs_dtype, targets, batch, masks, seed, predictions, record, (cudaStream_t)stream, true);
}

int32_t ts_fit_backward(ts_mlp *m, const void *inputs, int32_t inputs_dtype, const int64_t *targets, int64_t batch,
                        const uint8_t *const *masks, uint64_t seed, int64_t *predictions, ts_batch_record *record, void *stream) {
    return fit_impl(m, inputs, inputs_dtype, targets, batch, masks, seed, predictions, record, (cudaStream_t)stream, false);
}

int32_t ts_fit_backward_overlap(ts_mlp *m, const void *inputs, int32_t inputs_dtype, const int64_t *targets, int64_t batch,
                                const uint8_t *const *masks, uint64_t seed, int64_t *predictions, ts_batch_record *record, void *stream,
                                void *const *layer_events, int32_t *tensor_final) {
    TS_REQUIRE(m && layer_events && tensor_final, TS_E_INVALID, "null handle / layer_events / tensor_final");
    for (int l = 0; l < m->n_layers; ++l) TS_REQUIRE(layer_events[l] != nullptr, TS_E_INVALID, "layer_events[%d] is null", l);
    for (int t = 0; t < m->n_tensors; ++t) tensor_final[t] = 0;
    m->ov_events = (cudaEvent_t *)layer_events;
    m->ov_final = tensor_final;
    const int rc = fit_impl(m, inputs, inputs_dtype, targets, batch, masks, seed, predictions, record, (cudaStream_t)stream, false);
    m->ov_events = nullptr;
    m->ov_final = nullptr;
    return rc;
}

int32_t ts_dp_reserve_sms(ts_mlp *m, int32_t n_sms) {
    TS_REQUIRE(m, TS_E_INVALID, "null handle");
    TS_REQUIRE(n_sms >= 0 && n_sms < m->sm_count, TS_E_INVALID, "cannot reserve %d of %d SMs", n_sms, m->sm_count);
    m->bwd_sms = m->sm_count - n_sms;
    return TS_OK;
}

int32_t ts_opt_apply(ts_mlp *m, float grad_scale, void *stream) {
    TS_REQUIRE(m, TS_E_INVALID, "null handle");
    TS_REQUIRE(m->params && m->grads && m->ws_set && m->opt_bound, TS_E_STATE, "bind parameters, optimizer and workspace first");
    TS_CUDA_CHECK(cudaSetDevice(m->device));
    return run_optimizer(m, nullptr, true, grad_scale, (cudaStream_t)stream);
}

int32_t ts_opt_apply_range(ts_mlp *m, float grad_scale, int32_t first_tensor, int32_t n_tensors, int32_t advance_step, void *stream) {
    TS_REQUIRE(m, TS_E_INVALID, "null handle");
    TS_REQUIRE(m->params && m->grads && m->ws_set && m->opt_bound, TS_E_STATE, "bind parameters, optimizer and workspace first");
    TS_REQUIRE(first_tensor >= 0 && n_tensors >= 1 && first_tensor + n_tensors <= m->n_tensors, TS_E_INVALID,
               "tensor range [%d, %d) outside [0, %d)", first_tensor, first_tensor + n_tensors, m->n_tensors);
    TS_REQUIRE(advance_step || m->step >= 1, TS_E_STATE, "the first range of a step must advance the step counter");
    TS_CUDA_CHECK(cudaSetDevice(m->device));
    OptArgs o{};
    bool any_partial = false;
    int rc = fill_tensor_table(m, nullptr, o, &any_partial, (cudaStream_t)stream);
    if (rc) return rc;
    o.update = 1; o.grad_scale = grad_scale;
    fill_update(m, o, advance_step != 0);
    const int last = first_tensor + n_tensors - 1;
    o.chunk_begin = (int)(m->off[first_tensor] / kChunk);
    const int64_t chunk_end = (m->off[last] + round_up(m->cnt[last], kChunk)) / kChunk;
    ProfScope ps(K_OPT, (cudaStream_t)stream);
    return launch_optimizer(o, chunk_end - o.chunk_begin, (cudaStream_t)stream);
}

int32_t ts_eval_step(ts_mlp *m, const void *inputs, int32_t inputs_dtype, const int64_t *targets, int64_t batch,
                     int64_t *predictions, ts_batch_record *record, void *stream) {
    int rc = check_ready(m, batch, false);
    if (rc) return rc;
    TS_REQUIRE(inputs && targets && predictions && record, TS_E_INVALID, "null inputs/targets/predictions/record pointer");
    TS_CUDA_CHECK(cudaSetDevice(m->device));
    const void *a0 = nullptr, *a_last = nullptr;
    if (m->fused)
        return fused_forward(m, inputs, inputs_dtype, targets, batch, nullptr, 0, false, predictions, record, &a0, (cudaStream_t)stream);
    rc = forward_hidden(m, inputs, inputs_dtype, batch, nullptr, 0, false, &a0, &a_last, (cudaStream_t)stream);
    if (rc) return rc;
    if (m->tc_head) return tc_head_forward(m, a_last, targets, batch, false, predictions, record, (cudaStream_t)stream);
    HeadArgs h{};
    fill_head(m, h, a_last, targets, batch, predictions, record);
    h.drop = drop_spec(m, nullptr, 0, m->n_layers - 2, false);
    ProfScope ps(K_HEAD, (cudaStream_t)stream);
    return launch_head(h, m->sm_count, (cudaStream_t)stream);
}

int32_t ts_metrics_accumulate(ts_metrics_state *state, const ts_batch_record *ring, int64_t ring_len, int64_t first,
                              int64_t count, void *stream) {
    TS_REQUIRE(state && ring && ring_len > 0 && first >= 0 && count >= 0 && count <= ring_len, TS_E_INVALID,
               "bad ring arguments (len %lld, first %lld, count %lld)", (long long)ring_len, (long long)first, (long long)count);
    return launch_metrics_accumulate(state, ring, ring_len, first % ring_len, count, (cudaStream_t)stream);
}

int32_t ts_metrics_update_raw(ts_metrics_state *state, const float *loss, const int64_t *predictions, const int64_t *targets,
                              int64_t batch, void *stream) {
    TS_REQUIRE(state && loss && predictions && targets && batch >= 0, TS_E_INVALID, "null argument");
    return launch_metrics_update_raw(state, loss, predictions, targets, batch, (cudaStream_t)stream);
}

int32_t ts_metrics_reset(ts_metrics_state *state, void *stream) {
    TS_REQUIRE(state, TS_E_INVALID, "null state");
    TS_CUDA_CHECK(cudaMemsetAsync(state, 0, sizeof(ts_metrics_state), (cudaStream_t)stream));
    return TS_OK;
}

int32_t ts_metrics_read(const ts_metrics_state *state, ts_metrics_state *out, void *stream) {
    TS_REQUIRE(state && out, TS_E_INVALID, "null argument");
    TS_CUDA_CHECK(cudaMemcpyAsync(out, state, sizeof(ts_metrics_state), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    TS_CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream));
    return TS_OK;
}

int32_t ts_test_gemm_bf16(const void *A, int32_t a_major, int64_t lda, const void *B, int32_t b_major, int64_t ldb, int32_t M,
                          int32_t N, int32_t K, float *out, int64_t ldo, int32_t splits, int64_t split_stride, uint64_t *debug_clocks,
                          void *stream) {
    TS_REQUIRE(A && B && out, TS_E_INVALID, "null argument");
    int dev = 0, sms = 0;
    TS_CUDA_CHECK(cudaGetDevice(&dev));
    TS_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    TcGemmArgs g{};
    g.A = (const __nv_bfloat16 *)A; g.a_major = a_major; g.lda = lda;
    g.B = (const __nv_bfloat16 *)B; g.b_major = b_major; g.ldb = ldb;
    g.M = M; g.N = N; g.K = K; g.epi = TC_EPI_STORE_F32; g.out = out; g.ldo = ldo;
    g.splits = splits; g.k_chunk = 0; g.split_stride = split_stride;
    g.dbg = (unsigned long long *)debug_clocks;
    return launch_tc_gemm(g, sms, (cudaStream_t)stream);
}

int32_t ts_test_fused_trace(uint64_t *out) {
    TS_REQUIRE(out, TS_E_INVALID, "null argument");
    TS_CUDA_CHECK(cudaDeviceSynchronize());
    return read_fused_trace((unsigned long long *)out);
}

int32_t ts_test_eval_trace(uint64_t *out) {
    TS_REQUIRE(out, TS_E_INVALID, "null argument");
    TS_CUDA_CHECK(cudaDeviceSynchronize());
    return read_eval_trace((unsigned long long *)out);
}

int32_t ts_test_wgrad_trace(uint64_t *out) {
    TS_REQUIRE(out, TS_E_INVALID, "null argument");
    TS_CUDA_CHECK(cudaDeviceSynchronize());
    return read_wgrad_trace((unsigned long long *)out);
}

int32_t ts_test_step_trace(uint64_t *out) {
    TS_REQUIRE(out, TS_E_INVALID, "null argument");
    TS_CUDA_CHECK(cudaDeviceSynchronize());
    return read_step_trace((unsigned long long *)out);
}

int32_t ts_profile_begin(void) {
    for (auto &p : g_prof) { cudaEventDestroy(p.a); cudaEventDestroy(p.b); }
    g_prof.clear();
    g_prof_on = true;
    return TS_OK;
}

int32_t ts_profile_end(double *ms_per_kind, int64_t *launches_per_kind, int32_t n_kinds) {
    g_prof_on = false;
    TS_REQUIRE(ms_per_kind && launches_per_kind && n_kinds >= K_COUNT, TS_E_INVALID, "need room for %d kinds", K_COUNT);
    for (int i = 0; i < n_kinds; ++i) { ms_per_kind[i] = 0.0; launches_per_kind[i] = 0; }
    for (auto &p : g_prof) {
        TS_CUDA_CHECK(cudaEventSynchronize(p.b));
        float ms = 0.0f;
        TS_CUDA_CHECK(cudaEventElapsedTime(&ms, p.a, p.b));
        ms_per_kind[p.kind] += ms;
        launches_per_kind[p.kind] += 1;
        cudaEventDestroy(p.a);
        cudaEventDestroy(p.b);
    }
    g_prof.clear();
    return TS_OK;
}

int32_t ts_profile_kind_count(void) { return K_COUNT; }
const char *ts_profile_kind_name(int32_t kind) { return (kind >= 0 && kind < K_COUNT) ? kKindNames[kind] : ""; }

}  // extern "C"

// ======================================================================================= evaluate without any allocation
// ts_eval_step_static: the same arithmetic as ts_eval_step (fp32 handles), but every intermediate lives in one
// fixed array that is part of the library image, so neither the caller nor the library has to obtain device memory.
namespace {
constexpr int64_t kStaticScratchBytes = 16ll << 20;   // 16 MiB, resident on every GPU that loads the library
constexpr int kStaticMaxDevices = 64;
std::atomic<void *> g_static_scratch_addr[kStaticMaxDevices];   // device address of the array, per device (null until resolved)
std::mutex g_static_scratch_mu;
}  // namespace

// In host code this name is only a handle for cudaGetSymbolAddress; kernels get the resolved device address.
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

extern "C" int32_t ts_eval_step_static(ts_mlp *m, const void *inputs, int32_t inputs_dtype, const int64_t *targets,
                                       int64_t batch, int64_t *predictions, ts_batch_record *record, void *stream) {
    TS_REQUIRE(m != nullptr, TS_E_INVALID, "null handle");
    TS_REQUIRE(m->params != nullptr, TS_E_STATE, "parameters not bound: call ts_mlp_bind_params first");
    TS_REQUIRE(inputs && targets && predictions && record, TS_E_INVALID, "null inputs/targets/predictions/record pointer");
    TS_REQUIRE(m->prec == TS_PREC_F32 && inputs_dtype == TS_DT_F32, TS_E_UNSUPPORTED,
               "ts_eval_step_static runs fp32 handles on fp32 inputs only; use ts_eval_step with a workspace otherwise");
    TS_REQUIRE(batch >= 1, TS_E_INVALID, "batch %lld", (long long)batch);
    TS_REQUIRE(m->device >= 0 && m->device < kStaticMaxDevices, TS_E_UNSUPPORTED, "device index %d beyond %d", m->device,
               kStaticMaxDevices);
    cudaStream_t st = (cudaStream_t)stream;
    const int L = m->n_layers;

    // how much of the array this call needs: ticket | head partials | per hidden layer: activation (+ GELU pre-activation)
    // (the batch is bounded first, so the products below cannot overflow)
    int64_t need = 256 + round_up((int64_t)sizeof(HeadPartial) * kMaxHeadBlocks, 256);
    const bool batch_sane = batch <= kStaticScratchBytes / 4;
    for (int l = 0; l < L - 1 && batch_sane; ++l)
        need += round_up(batch * m->dims[l + 1] * 4, 256) * (m->act == TS_ACT_GELU ? 2 : 1);
    if (!batch_sane) need = INT64_MAX;
    TS_REQUIRE(need <= kStaticScratchBytes, TS_E_TOO_LARGE,
               "ts_eval_step_static: batch %lld needs %lld bytes of scratch, the built-in array holds %lld (batch * sum of hidden "
               "widths * 4 must stay below about 16 MiB); use ts_eval_step with a caller workspace for larger batches",
               (long long)batch, (long long)need, (long long)kStaticScratchBytes);

    TS_CUDA_CHECK(cudaSetDevice(m->device));
    void *scratch = g_static_scratch_addr[m->device].load(std::memory_order_acquire);
    if (scratch == nullptr) {   // set-up, once per device, with that device current (host threads may race here: one lock)
        std::lock_guard<std::mutex> lock(g_static_scratch_mu);
        scratch = g_static_scratch_addr[m->device].load(std::memory_order_relaxed);
        if (scratch == nullptr) {
            TS_CUDA_CHECK(cudaGetSymbolAddress(&scratch, g_static_scratch));
            TS_REQUIRE(scratch != nullptr, TS_E_CUDA, "cudaGetSymbolAddress returned a null address for the static scratch");
            g_static_scratch_addr[m->device].store(scratch, std::memory_order_release);
        }
    }
    Carver c(scratch);
    unsigned int *ticket = (unsigned int *)c.take(256);
    HeadPartial *partials = (HeadPartial *)c.take(sizeof(HeadPartial) * kMaxHeadBlocks);
    // the array keeps whatever the previous call left: the completion ticket is the one thing that must start at zero
    TS_CUDA_CHECK(cudaMemsetAsync(ticket, 0, 256, st));

    const float *a_in = (const float *)inputs;
    for (int l = 0; l < L - 1; ++l) {
        const int in = m->dims[l], out = m->dims[l + 1];
        float *a_out = (float *)c.take(batch * out * 4);
        float *u_out = (m->act == TS_ACT_GELU) ? (float *)c.take(batch * out * 4) : nullptr;
        SgemmArgs g{};
        g.A = a_in; g.sam = in; g.sak = 1;
        g.B = m->params + m->off[2 * l]; g.sbk = 1; g.sbn = in;
        g.C = a_out; g.ldc = out; g.M = (int)batch; g.N = out; g.K = in;
        g.k_chunk = (int)round_up(in, 16); g.c_split_stride = 0;
        g.bias = m->params + m->off[2 * l + 1];
        g.U_out = u_out; g.act = m->act; g.training = 0;
        g.drop = drop_spec(m, nullptr, 0, l, false);
        ProfScope ps(K_FWD_GEMM, st);
        int rc = launch_sgemm(EPI_FWD, g, 1, st);
        if (rc) return rc;
        a_in = a_out;
    }
    HeadArgs h{};
    fill_head(m, h, a_in, targets, batch, predictions, record);
    h.a_is_bf16 = 0; h.round_w_bf16 = 0;
    h.partials = partials; h.ticket = ticket;
    h.drop = drop_spec(m, nullptr, 0, L - 2, false);
    ProfScope ps(K_HEAD, st);
    return launch_head(h, m->sm_count, st);
}
