// SIMT kernels of the tsb200 engine: the fp32 dense path (1e-5 parity mode), the fused head
// (last Linear + cross-entropy + argmax + dZ + dA, shared by both precision modes), gradient
// reductions, the flat optimizer and the metric accumulators. sm_100a only.
#include "opt_math.cuh"
#include "kernels.h"

namespace tsb {

// =====================================================================================
// fp32 GEMM, 64x64x16 tiles, 256 threads, 4x4 outputs per thread, optional split-K.
// Generic strides cover the three contractions of a Linear layer:
//   forward  Y = X W^T   (A K-contiguous, B K-contiguous)      mlp.py:14,17
//   dX = dY W            (A K-contiguous, B N-contiguous)      autograd of the same
//   dW = dY^T X          (A M-contiguous, B N-contiguous), K = batch, split across blockIdx.z
// =====================================================================================
template <int EPI>
__global__ void __launch_bounds__(256) sgemm_kernel(SgemmArgs p) {
    __shared__ __align__(16) float As[16][68];
    __shared__ __align__(16) float Bs[16][68];
    const int tid = threadIdx.x;
    const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
    const int kb = blockIdx.z * p.k_chunk;
    const int ke = min(p.K, kb + p.k_chunk);
    const int ty = tid >> 4, tx = tid & 15;
    const bool a_kc = (p.sak == 1), b_kc = (p.sbk == 1);

    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;

    for (int k0 = kb; k0 < ke; k0 += 16) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            int m, k;
            if (a_kc) { k = tid & 15; m = (tid >> 4) + i * 16; } else { m = tid & 63; k = (tid >> 6) + i * 4; }
            const int gm = m0 + m, gk = k0 + k;
            As[k][m] = (gm < p.M && gk < ke) ? p.A[(int64_t)gm * p.sam + (int64_t)gk * p.sak] : 0.0f;
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            int n, k;
            if (b_kc) { k = tid & 15; n = (tid >> 4) + i * 16; } else { n = tid & 63; k = (tid >> 6) + i * 4; }
            const int gn = n0 + n, gk = k0 + k;
            Bs[k][n] = (gn < p.N && gk < ke) ? p.B[(int64_t)gk * p.sbk + (int64_t)gn * p.sbn] : 0.0f;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            const float4 a = *reinterpret_cast<const float4 *>(&As[k][ty * 4]);
            const float4 b = *reinterpret_cast<const float4 *>(&Bs[k][tx * 4]);
            const float av[4] = {a.x, a.y, a.z, a.w};
            const float bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        __syncthreads();
    }

    float *C = p.C + (int64_t)blockIdx.z * p.c_split_stride;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int gm = m0 + ty * 4 + i;
        if (gm >= p.M) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int gn = n0 + tx * 4 + j;
            if (gn >= p.N) continue;
            const int64_t o = (int64_t)gm * p.ldc + gn;
            float v = acc[i][j];
            if (EPI == EPI_FWD) {
                // Linear bias, then Dropout, then the activation: mlp.py:14-16
                float u = v + p.bias[gn];
                if (p.training) u *= p.drop.factor(gm, gn, p.N);
                if (p.U_out) p.U_out[o] = u;
                v = (p.act == TS_ACT_RELU) ? fmaxf(u, 0.0f) : gelu_f(u);
            } else if (EPI == EPI_ACTGRAD) {
                // dH = dA * act'(.) * keep/(1-p); ReLU uses the sign of its output (threshold_backward)
                const float keep = p.training ? p.drop.factor(gm, gn, p.N) : 1.0f;
                if (p.act == TS_ACT_RELU) v = (p.A_prev[o] > 0.0f) ? v * keep : 0.0f;
                else v = v * gelu_grad_f(p.U_prev[o]) * keep;
            }
            C[o] = v;
        }
    }
}

int launch_sgemm(int epi, const SgemmArgs &a, int splits, cudaStream_t st) {
    dim3 grid((a.N + 63) / 64, (a.M + 63) / 64, splits);
    if (epi == EPI_STORE) sgemm_kernel<EPI_STORE><<<grid, 256, 0, st>>>(a);
    else if (epi == EPI_FWD) sgemm_kernel<EPI_FWD><<<grid, 256, 0, st>>>(a);
    else sgemm_kernel<EPI_ACTGRAD><<<grid, 256, 0, st>>>(a);
    TS_LAUNCH_CHECK();
    return TS_OK;
}

// =====================================================================================
// Fused head. One warp per row: logits (K-long dot products against W held in shared memory),
// log-softmax / NLL, first-max argmax, dZ = (softmax - onehot)/B, then dA = dZ W and the hidden
// activation gradient -> dH. Loss sum and correct count are reduced warp -> block -> grid (last
// block done finalises in a fixed order), so Metrics never re-reads logits or predictions.
//   classifier.py:23-27 (forward, loss), :35/:39 (argmax), metrics.py:14-15 (what is counted)
// =====================================================================================
template <typename T>
__device__ __forceinline__ void load4(const T *p, float (&v)[4]);
template <>
__device__ __forceinline__ void load4<float>(const float *p, float (&v)[4]) {
    const float4 t = *reinterpret_cast<const float4 *>(p);
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
}
template <>
__device__ __forceinline__ void load4<__nv_bfloat16>(const __nv_bfloat16 *p, float (&v)[4]) {
    const uint2 t = *reinterpret_cast<const uint2 *>(p);
    const __nv_bfloat162 lo = *reinterpret_cast<const __nv_bfloat162 *>(&t.x);
    const __nv_bfloat162 hi = *reinterpret_cast<const __nv_bfloat162 *>(&t.y);
    v[0] = __low2float(lo); v[1] = __high2float(lo); v[2] = __low2float(hi); v[3] = __high2float(hi);
}
__device__ __forceinline__ void store4(float *p, const float (&v)[4]) {
    *reinterpret_cast<float4 *>(p) = make_float4(v[0], v[1], v[2], v[3]);
}
__device__ __forceinline__ void store4(__nv_bfloat16 *p, const float (&v)[4]) {
    __nv_bfloat162 lo = __floats2bfloat162_rn(v[0], v[1]);
    __nv_bfloat162 hi = __floats2bfloat162_rn(v[2], v[3]);
    uint2 t;
    t.x = *reinterpret_cast<uint32_t *>(&lo);
    t.y = *reinterpret_cast<uint32_t *>(&hi);
    *reinterpret_cast<uint2 *>(p) = t;
}

template <typename T, int MAXC>
__global__ void __launch_bounds__(256) head_kernel(HeadArgs p) {
    extern __shared__ __align__(16) float Ws[];  // [C][K]
    __shared__ float s_loss[8];
    __shared__ int s_corr[8];
    __shared__ bool s_last;

    const int K = p.K, C = p.C;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < C * K; i += blockDim.x) {
        const float w = p.W[i];
        Ws[i] = p.round_w_bf16 ? round_bf16(w) : w;
    }
    __syncthreads();

    const T *A = reinterpret_cast<const T *>(p.A);
    T *dH = reinterpret_cast<T *>(p.dH);
    const float inv_B = 1.0f / (float)p.B;
    float loss_acc = 0.0f;
    int corr_acc = 0;

    for (int64_t r = (int64_t)blockIdx.x * 8 + warp; r < p.B; r += (int64_t)gridDim.x * 8) {
        const T *a_row = A + r * K;
        float z[MAXC];
#pragma unroll
        for (int c = 0; c < MAXC; ++c) z[c] = 0.0f;
        for (int k = lane * 4; k < K; k += 128) {
            float a4[4];
            load4<T>(a_row + k, a4);
#pragma unroll
            for (int c = 0; c < MAXC; ++c) {
                if (c < C) {
                    const float4 w = *reinterpret_cast<const float4 *>(&Ws[c * K + k]);
                    z[c] = fmaf(a4[0], w.x, z[c]);
                    z[c] = fmaf(a4[1], w.y, z[c]);
                    z[c] = fmaf(a4[2], w.z, z[c]);
                    z[c] = fmaf(a4[3], w.w, z[c]);
                }
            }
        }
#pragma unroll
        for (int c = 0; c < MAXC; ++c)
            if (c < C) z[c] = warp_sum(z[c]) + p.bias[c];

        // every lane now holds the full logit row
        float mx = z[0];
        int arg = 0;
#pragma unroll
        for (int c = 1; c < MAXC; ++c)
            if (c < C && z[c] > mx) { mx = z[c]; arg = c; }   // strict '>' : first maximum wins, like torch.argmax
        float s = 0.0f;
#pragma unroll
        for (int c = 0; c < MAXC; ++c)
            if (c < C) s += expf(z[c] - mx);
        const float lse = mx + logf(s);
        const int yy = (int)p.y[r];
        float zy = 0.0f;
#pragma unroll
        for (int c = 0; c < MAXC; ++c)
            if (c == yy) zy = z[c];
        float row_loss = lse - zy;
        if (yy < 0 || yy >= C) {                                       // target outside [0,C): poison the loss, tell the host
            row_loss = __int_as_float(0x7fc00000);
            if (lane == 0 && p.bad_target != nullptr) *reinterpret_cast<volatile uint32_t *>(p.bad_target) = (uint32_t)r + 1u;
        }
        loss_acc += row_loss;
        corr_acc += (arg == yy);
        if (lane == 0) p.pred[r] = arg;

        if (p.dz != nullptr) {
            float dzv[MAXC];
#pragma unroll
            for (int c = 0; c < MAXC; ++c)
                dzv[c] = (c < C) ? (expf(z[c] - lse) - (c == yy ? 1.0f : 0.0f)) * inv_B : 0.0f;
            if (lane == 0) {
#pragma unroll
                for (int c = 0; c < MAXC; ++c)
                    if (c < C) p.dz[r * C + c] = dzv[c];
            }
            if (dH != nullptr) {
                // bf16 mode: dA = dZ W contracts the bf16-rounded dZ, like the tensor-core path (fused_mlp.cu)
                if (p.round_w_bf16) {
#pragma unroll
                    for (int c = 0; c < MAXC; ++c) dzv[c] = round_bf16(dzv[c]);
                }
                for (int k = lane * 4; k < K; k += 128) {
                    float d[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
                    for (int c = 0; c < MAXC; ++c) {
                        if (c < C) {
                            const float4 w = *reinterpret_cast<const float4 *>(&Ws[c * K + k]);
                            d[0] = fmaf(dzv[c], w.x, d[0]);
                            d[1] = fmaf(dzv[c], w.y, d[1]);
                            d[2] = fmaf(dzv[c], w.z, d[2]);
                            d[3] = fmaf(dzv[c], w.w, d[3]);
                        }
                    }
                    if (p.act == TS_ACT_RELU) {
                        float a4[4];
                        load4<T>(a_row + k, a4);
                        const float sc = (p.drop.p > 0.0f) ? p.drop.scale : 1.0f;
#pragma unroll
                        for (int i = 0; i < 4; ++i) d[i] = (a4[i] > 0.0f) ? d[i] * sc : 0.0f;
                    } else {
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            d[i] = d[i] * gelu_grad_f(p.U_prev[r * K + k + i]) * p.drop.factor(r, k + i, K);
                    }
                    store4(dH + r * K + k, d);
                }
            }
        }
    }

    // warp -> block -> grid reduction of (loss sum, correct count)
    if (lane == 0) { s_loss[warp] = loss_acc; s_corr[warp] = corr_acc; }
    __syncthreads();
    if (threadIdx.x == 0) {
        float l = 0.0f;
        int c = 0;
        for (int w = 0; w < 8; ++w) { l += s_loss[w]; c += s_corr[w]; }
        p.partials[blockIdx.x].loss_sum = l;
        p.partials[blockIdx.x].correct = c;
        __threadfence();
        const unsigned int t = atomicAdd(p.ticket, 1u);
        s_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (s_last && warp == 0) {
        __threadfence();
        float l = 0.0f;
        int c = 0;
        for (int b = lane; b < (int)gridDim.x; b += 32) {
            const volatile HeadPartial *q = p.partials + b;
            l += q->loss_sum;
            c += q->correct;
        }
        l = warp_sum(l);
        c = warp_sum(c);
        if (lane == 0) {
            p.record->loss = l / (float)p.B;   // reduction='mean'
            p.record->n_correct = c;
            p.record->n_rows = (int32_t)p.B;
            p.record->reserved = 0;
            *p.ticket = 0u;                    // ready for the next launch on this stream
        }
    }
}

int launch_head(const HeadArgs &a, int sm_count, cudaStream_t st) {
    TS_REQUIRE(a.C >= 1 && a.C <= kMaxClasses, TS_E_UNSUPPORTED, "head kernel supports 1..%d classes, got %d",
               kMaxClasses, a.C);
    TS_REQUIRE(a.K % 8 == 0, TS_E_UNSUPPORTED, "last hidden width must be a multiple of 8, got %d", a.K);
    const size_t smem = (size_t)a.C * a.K * sizeof(float);
    TS_REQUIRE(smem <= 200 * 1024, TS_E_TOO_LARGE, "classes*hidden = %d exceeds the head kernel's 51200-element weight tile",
               a.C * a.K);
    int per_sm = (int)((220 * 1024) / (smem + 1024));
    if (per_sm > 4) per_sm = 4;
    if (per_sm < 1) per_sm = 1;
    int64_t want = (a.B + 31) / 32;   // >= 4 rows per warp
    int grid = (int)(want < 1 ? 1 : want);
    if (grid > sm_count * per_sm) grid = sm_count * per_sm;
    if (grid > kMaxHeadBlocks) grid = kMaxHeadBlocks;

#define TS_HEAD_LAUNCH(T, MAXC)                                                                                  \
    do {                                                                                                         \
        if (smem > 48 * 1024)                                                                                    \
            TS_CUDA_CHECK(cudaFuncSetAttribute(head_kernel<T, MAXC>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                               (int)smem));                                                      \
        head_kernel<T, MAXC><<<grid, 256, smem, st>>>(a);                                                        \
    } while (0)

    if (a.a_is_bf16) {
        if (a.C <= 16) TS_HEAD_LAUNCH(__nv_bfloat16, 16); else TS_HEAD_LAUNCH(__nv_bfloat16, 32);
    } else {
        if (a.C <= 16) TS_HEAD_LAUNCH(float, 16); else TS_HEAD_LAUNCH(float, 32);
    }
#undef TS_HEAD_LAUNCH
    TS_LAUNCH_CHECK();
    return TS_OK;
}

// =====================================================================================
// dW_last / db_last partials: thread per hidden column k, dz rows broadcast from shared memory.
// =====================================================================================
__device__ __forceinline__ uint32_t pack2_bf16(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t *>(&h);
}

// =====================================================================================
// Tensor-core head (bf16 mode, hidden layers too wide for the fused kernel): the two big contractions of the last
// layer run as tcgen05 GEMMs (engine.cu), and these two small kernels do what is left.
//
// logits_head_kernel: thread = batch row. Z[r, 0..15] (fp32, from the Z = A W^T GEMM, classes padded to 16) + bias
//   -> log-softmax / NLL / first-max argmax -> dZ (fp32 [B, C] for db_last, and bf16 [B, 16] zero-padded as the operand
//   of the dA and dW GEMMs); loss sum and correct count warp -> block -> grid like head_kernel.
// =====================================================================================
__global__ void __launch_bounds__(256) logits_head_kernel(const float *__restrict__ Z, const float *__restrict__ bias,
                                                          const int64_t *__restrict__ y, int64_t B, int C, int64_t *__restrict__ pred,
                                                          float *__restrict__ dz, __nv_bfloat16 *__restrict__ dzb,
                                                          HeadPartial *partials, unsigned int *ticket, ts_batch_record *record,
                                                          uint32_t *bad_target) {
    __shared__ float s_loss[8];
    __shared__ int s_corr[8];
    __shared__ bool s_last;
    pdl_launch_dependents();
    pdl_wait();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t r = (int64_t)blockIdx.x * 256 + threadIdx.x;
    float row_loss = 0.0f;
    int corr = 0;
    if (r < B) {
        float z[16];
        const float4 *zr = reinterpret_cast<const float4 *>(Z + r * 16);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const float4 v = zr[q];
            z[4 * q] = v.x; z[4 * q + 1] = v.y; z[4 * q + 2] = v.z; z[4 * q + 3] = v.w;
        }
#pragma unroll
        for (int c = 0; c < 16; ++c) z[c] = (c < C) ? z[c] + bias[c] : 0.0f;
        float mx = z[0];
        int arg = 0;
#pragma unroll
        for (int c = 1; c < 16; ++c)
            if (c < C && z[c] > mx) { mx = z[c]; arg = c; }      // strict '>': first maximum wins, like torch.argmax
        float sum = 0.0f;
#pragma unroll
        for (int c = 0; c < 16; ++c)
            if (c < C) sum += expf(z[c] - mx);
        const float lse = mx + logf(sum);
        const int yy = (int)y[r];
        float zy = 0.0f;
#pragma unroll
        for (int c = 0; c < 16; ++c)
            if (c == yy) zy = z[c];
        row_loss = lse - zy;
        if (yy < 0 || yy >= C) {                                        // target outside [0,C): poison the loss, tell the host
            row_loss = __int_as_float(0x7fc00000);
            if (bad_target != nullptr) *reinterpret_cast<volatile uint32_t *>(bad_target) = (uint32_t)r + 1u;
        }
        corr = (arg == yy);
        pred[r] = arg;
        if (dz != nullptr) {
            const float inv_B = 1.0f / (float)B;
            float d[16];
#pragma unroll
            for (int c = 0; c < 16; ++c) d[c] = (c < C) ? (expf(z[c] - lse) - (c == yy ? 1.0f : 0.0f)) * inv_B : 0.0f;
            for (int c = 0; c < C; ++c) dz[r * C + c] = d[c];
            uint4 lo, hi;
            lo.x = pack2_bf16(d[0], d[1]); lo.y = pack2_bf16(d[2], d[3]); lo.z = pack2_bf16(d[4], d[5]); lo.w = pack2_bf16(d[6], d[7]);
            hi.x = pack2_bf16(d[8], d[9]); hi.y = pack2_bf16(d[10], d[11]); hi.z = pack2_bf16(d[12], d[13]); hi.w = pack2_bf16(d[14], d[15]);
            uint4 *o = reinterpret_cast<uint4 *>(dzb + r * 16);
            o[0] = lo; o[1] = hi;
        }
    }
    row_loss = warp_sum(row_loss);
    corr = warp_sum(corr);
    if (lane == 0) { s_loss[warp] = row_loss; s_corr[warp] = corr; }
    __syncthreads();
    if (threadIdx.x == 0) {
        float l = 0.0f;
        int c = 0;
        for (int w = 0; w < 8; ++w) { l += s_loss[w]; c += s_corr[w]; }
        partials[blockIdx.x].loss_sum = l;
        partials[blockIdx.x].correct = c;
        __threadfence();
        s_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (s_last && warp == 0) {
        __threadfence();
        float l = 0.0f;
        int c = 0;
        for (int b = lane; b < (int)gridDim.x; b += 32) {
            const volatile HeadPartial *q = partials + b;
            l += q->loss_sum;
            c += q->correct;
        }
        l = warp_sum(l);
        c = warp_sum(c);
        if (lane == 0) {
            record->loss = l / (float)B;   // reduction='mean'
            record->n_correct = c;
            record->n_rows = (int32_t)B;
            record->reserved = 0;
            *ticket = 0u;
        }
    }
}

int launch_logits_head(const float *Z, const float *bias, const int64_t *y, int64_t B, int C, int64_t *pred, float *dz,
                       __nv_bfloat16 *dzb, HeadPartial *partials, unsigned int *ticket, ts_batch_record *record, uint32_t *bad_target,
                       cudaStream_t st) {
    const int64_t blocks = (B + 255) / 256;
    TS_REQUIRE(blocks <= kMaxHeadBlocks, TS_E_TOO_LARGE, "batch %lld needs %lld head blocks, more than %d", (long long)B,
               (long long)blocks, kMaxHeadBlocks);
    TS_CUDA_CHECK(launch_pdl(logits_head_kernel, dim3((unsigned)blocks), dim3(256), 0, st, 1, Z, bias, y, B, C, pred, dz, dzb, partials,
                             ticket, record, bad_target));
    TS_LAUNCH_CHECK();
    return TS_OK;
}

// wlast_finish_kernel: dW_last[c, k] = sum over split-K partials of the transposed product dWt[s][k][16] (what the
// dW^T = A^T dZ GEMM writes), and db_last[c] = sum_r dz[r, c] (fp32 dZ, fixed-order tree). Writes straight into `grads`.
__global__ void __launch_bounds__(256) wlast_finish_kernel(const float *__restrict__ wt, int64_t split_stride, int splits, int K, int C,
                                                           const float *__restrict__ dz, int64_t B, float *__restrict__ gw,
                                                           float *__restrict__ gb) {
    pdl_launch_dependents();
    pdl_wait();
    const int wblocks = (K + 255) / 256;
    if ((int)blockIdx.x < wblocks) {
        const int k = blockIdx.x * 256 + threadIdx.x;
        if (k >= K) return;
        float acc[16];
#pragma unroll
        for (int c = 0; c < 16; ++c) acc[c] = 0.0f;
        for (int s = 0; s < splits; ++s) {
            const float4 *q = reinterpret_cast<const float4 *>(wt + (int64_t)s * split_stride + (int64_t)k * 16);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float4 v = q[j];
                acc[4 * j] += v.x; acc[4 * j + 1] += v.y; acc[4 * j + 2] += v.z; acc[4 * j + 3] += v.w;
            }
        }
#pragma unroll
        for (int c = 0; c < 16; ++c)
            if (c < C) gw[(int64_t)c * K + k] = acc[c];         // coalesced over k for every class row
        return;
    }
    // one block per class: the bias gradient. Thread t sums rows t, t+256, ... (8 loads in flight), then a fixed-order tree.
    __shared__ float red[256];
    const int c = (int)blockIdx.x - wblocks;
    float a = 0.0f;
    int64_t r = threadIdx.x;
    for (; r + 7 * 256 < B; r += 8 * 256) {
        float v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) v[u] = dz[(r + u * 256) * C + c];
#pragma unroll
        for (int u = 0; u < 8; ++u) a += v[u];
    }
    for (; r < B; r += 256) a += dz[r * C + c];
    red[threadIdx.x] = a;
    __syncthreads();
    for (int w = 128; w > 0; w >>= 1) {
        if ((int)threadIdx.x < w) red[threadIdx.x] += red[threadIdx.x + w];
        __syncthreads();
    }
    if (threadIdx.x == 0) gb[c] = red[0];
}

int launch_wlast_finish(const float *wt, int64_t split_stride, int splits, int K, int C, const float *dz, int64_t B, float *gw,
                        float *gb, cudaStream_t st) {
    const unsigned blocks = (unsigned)((K + 255) / 256) + (unsigned)C;
    TS_CUDA_CHECK(launch_pdl(wlast_finish_kernel, dim3(blocks), dim3(256), 0, st, 1, wt, split_stride, splits, K, C, dz, B, gw, gb));
    TS_LAUNCH_CHECK();
    return TS_OK;
}

// dH of the last hidden layer (bf16 mode, wide nets): dH[r,k] = (sum_c dz[r,c] W_last[c,k]) * act'(.) * keep / (1 - p), with the
// bias gradient of that layer (column sums of the ROUNDED dH, one partial per 32 rows) taken on the way. The contraction is only
// C <= 16 long, so this is an HBM-bound elementwise pass (read A_prev, write dH: 4 B per element), not a GEMM: as a tcgen05 tile
// loop it ran with 8 epilogue warps per SM in flight and took 129 us at 8192 x 4096; here every SM holds 2048 threads.
// Thread = 4 neighbouring columns, its W_last column quad in registers; CTA = 32 rows x 1024 columns.
template <int CP>
__global__ void __launch_bounds__(256) head_dh_kernel(const __nv_bfloat16 *__restrict__ dzb, const __nv_bfloat16 *__restrict__ Wp,
                                                      const __nv_bfloat16 *__restrict__ prev, int64_t B, int K, int act, DropoutSpec drop,
                                                      int training, __nv_bfloat16 *__restrict__ dH, float *__restrict__ colsum,
                                                      int64_t colsum_stride) {
    __shared__ __align__(16) float sdz[32][CP];
    const int64_t r0 = (int64_t)blockIdx.y * 32;
    const int nr = (int)min((int64_t)32, B - r0);
    const int k = (blockIdx.x * 256 + threadIdx.x) * 4;
    pdl_wait();
    pdl_launch_dependents();
    float w[CP][4];
    if (k < K) {
#pragma unroll
        for (int c = 0; c < CP; ++c) {
            const uint2 pk = *reinterpret_cast<const uint2 *>(Wp + (int64_t)c * K + k);
            const __nv_bfloat162 lo = *reinterpret_cast<const __nv_bfloat162 *>(&pk.x), hi = *reinterpret_cast<const __nv_bfloat162 *>(&pk.y);
            w[c][0] = __low2float(lo); w[c][1] = __high2float(lo); w[c][2] = __low2float(hi); w[c][3] = __high2float(hi);
        }
    }
    for (int i = threadIdx.x; i < 32 * CP; i += 256) {
        const int rr = i / CP, c = i % CP;
        sdz[rr][c] = rr < nr ? __bfloat162float(dzb[(r0 + rr) * 16 + c]) : 0.0f;
    }
    __syncthreads();
    if (k >= K) return;
    const float sc = training ? drop.scale : 1.0f;
    float cs[4] = {0.0f, 0.0f, 0.0f, 0.0f};
    constexpr int U = 8;       // rows whose loads are in flight together
    for (int rb = 0; rb < nr; rb += U) {
        uint2 pv[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
            pv[u] = rb + u < nr ? __ldcs(reinterpret_cast<const uint2 *>(prev + (r0 + rb + u) * K + k)) : make_uint2(0u, 0u);
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (rb + u >= nr) break;
            const int rr = rb + u;
            float a[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
            for (int c4 = 0; c4 < CP; c4 += 4) {
                const float4 d = *reinterpret_cast<const float4 *>(&sdz[rr][c4]);
                const float dv[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
                for (int j = 0; j < 4; ++j)
#pragma unroll
                    for (int e = 0; e < 4; ++e) a[e] = fmaf(dv[j], w[c4 + j][e], a[e]);
            }
            const __nv_bfloat162 plo = *reinterpret_cast<const __nv_bfloat162 *>(&pv[u].x), phi = *reinterpret_cast<const __nv_bfloat162 *>(&pv[u].y);
            const float pf[4] = {__low2float(plo), __high2float(plo), __low2float(phi), __high2float(phi)};
            if (act == TS_ACT_RELU) {      // A_prev > 0 implies the unit was kept
#pragma unroll
                for (int e = 0; e < 4; ++e) a[e] = pf[e] > 0.0f ? a[e] * sc : 0.0f;
            } else {                       // prev = the stored pre-activation u; a dropped unit has u = 0, gelu'(0) = 1/2: keep factor explicit
                float keep[4] = {1.0f, 1.0f, 1.0f, 1.0f};
                if (training) drop.factor4(r0 + rr, k, K, keep);
#pragma unroll
                for (int e = 0; e < 4; ++e) a[e] = a[e] * gelu_grad_f(pf[e]) * keep[e];
            }
            const __nv_bfloat162 lo = __floats2bfloat162_rn(a[0], a[1]), hi = __floats2bfloat162_rn(a[2], a[3]);
            cs[0] += __low2float(lo); cs[1] += __high2float(lo); cs[2] += __low2float(hi); cs[3] += __high2float(hi);
            uint2 pk;
            pk.x = *reinterpret_cast<const uint32_t *>(&lo);
            pk.y = *reinterpret_cast<const uint32_t *>(&hi);
            *reinterpret_cast<uint2 *>(dH + (r0 + rr) * K + k) = pk;
        }
    }
    if (colsum != nullptr) *reinterpret_cast<float4 *>(colsum + (int64_t)blockIdx.y * colsum_stride + k) = make_float4(cs[0], cs[1], cs[2], cs[3]);
}

// The same pass for widths that are multiples of 64, with the 16-long contraction on warp-level MMAs (m16n8k16, one k-step):
// the FMA form above spends as long on its 12-16 FMAs per element as on the memory traffic. Warp = 64 rows x 64 columns in
// four 16-row blocks; the 8 column tiles of a block are PERMUTED columns of W_last, chosen so that the accumulator fragment
// of a thread is 2 x 8 contiguous columns of its row (tile j, fragment column 2q+e  <->  column 8q + 2(j&3) + e + 32(j>>2)):
// A_prev is read and dH written as 16-byte pieces, four neighbouring lanes covering 64 contiguous bytes.
__device__ __forceinline__ void mma_bf16_16x8x16(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__global__ void __launch_bounds__(256, 2) head_dh_mma_kernel(const __nv_bfloat16 *__restrict__ dzb, const __nv_bfloat16 *__restrict__ Wp,
                                                             const __nv_bfloat16 *__restrict__ prev, int64_t B, int K, int act,
                                                             DropoutSpec drop, int training, __nv_bfloat16 *__restrict__ dH,
                                                             float *__restrict__ colsum, int64_t colsum_stride) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, i = lane >> 2, q = lane & 3;
    const int cb = (blockIdx.x * 8 + warp) * 64;          // this warp's 64 columns
    const int64_t r0 = (int64_t)blockIdx.y * 64;
    pdl_wait();
    pdl_launch_dependents();
    if (cb >= K) return;
    // B fragments: tile j, fragment column i = this lane's n; k = 2q, 2q+1 (b0) and 2q+8, 2q+9 (b1)
    uint32_t wb[8][2];
    const unsigned short *W16 = reinterpret_cast<const unsigned short *>(Wp);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int n = cb + 8 * (i >> 1) + 2 * (j & 3) + (i & 1) + 32 * (j >> 2);
        wb[j][0] = (uint32_t)W16[(int64_t)(2 * q) * K + n] | ((uint32_t)W16[(int64_t)(2 * q + 1) * K + n] << 16);
        wb[j][1] = (uint32_t)W16[(int64_t)(2 * q + 8) * K + n] | ((uint32_t)W16[(int64_t)(2 * q + 9) * K + n] << 16);
    }
    const float sc = training ? drop.scale : 1.0f;
    const int col = cb + 8 * q;                            // this thread's columns: col..col+7 and col+32..col+39
    auto fetch = [&](int blk, uint4 (&pv)[4]) {            // rows i and i + 8 of the block, two 16-byte pieces each
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int64_t row = r0 + blk * 16 + i + 8 * h;
#pragma unroll
            for (int g = 0; g < 2; ++g)
                pv[h * 2 + g] = (blk < 4 && row < B) ? __ldcs(reinterpret_cast<const uint4 *>(prev + row * K + col + 32 * g)) : make_uint4(0u, 0u, 0u, 0u);
        }
    };
    uint4 pv_next[4];
    fetch(0, pv_next);
    float cs[16];
#pragma unroll
    for (int e = 0; e < 16; ++e) cs[e] = 0.0f;
#pragma unroll 1
    for (int blk = 0; blk < 4; ++blk) {
        const int64_t rb = r0 + blk * 16;
        if (rb < B) {
            uint4 pv[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) pv[e] = pv_next[e];
            fetch(blk + 1, pv_next);
            uint32_t a[4];
            const uint32_t *dz32 = reinterpret_cast<const uint32_t *>(dzb);
            const int64_t ra = rb + i, rc = rb + i + 8;
            a[0] = ra < B ? dz32[ra * 8 + q] : 0u;
            a[1] = rc < B ? dz32[rc * 8 + q] : 0u;
            a[2] = ra < B ? dz32[ra * 8 + q + 4] : 0u;
            a[3] = rc < B ? dz32[rc * 8 + q + 4] : 0u;
            float acc[8][4];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.0f;
                mma_bf16_16x8x16(acc[j], a, wb[j][0], wb[j][1]);
            }
#pragma unroll
            for (int h = 0; h < 2; ++h) {                 // row i (h = 0: c0, c1) and row i + 8 (h = 1: c2, c3)
                const int64_t row = rb + i + 8 * h;
#pragma unroll
                for (int g = 0; g < 2; ++g) {              // columns col + 32 g .. + 7: tiles j = 4g .. 4g + 3
                    const uint4 pq = pv[h * 2 + g];
                    const uint32_t pw[4] = {pq.x, pq.y, pq.z, pq.w};
                    float v[8];
#pragma unroll
                    for (int jj = 0; jj < 4; ++jj) { v[2 * jj] = acc[4 * g + jj][2 * h]; v[2 * jj + 1] = acc[4 * g + jj][2 * h + 1]; }
                    float keep[8] = {1.0f, 1.0f, 1.0f, 1.0f, 1.0f, 1.0f, 1.0f, 1.0f};
                    if (act != TS_ACT_RELU && training && row < B) {
                        drop.factor4(row, col + 32 * g, K, *reinterpret_cast<float(*)[4]>(&keep[0]));
                        drop.factor4(row, col + 32 * g + 4, K, *reinterpret_cast<float(*)[4]>(&keep[4]));
                    }
                    uint32_t ow[4];
#pragma unroll
                    for (int e2 = 0; e2 < 4; ++e2) {
                        const __nv_bfloat162 pp = *reinterpret_cast<const __nv_bfloat162 *>(&pw[e2]);
                        const float p0 = __low2float(pp), p1 = __high2float(pp);
                        float x0 = v[2 * e2], x1 = v[2 * e2 + 1];
                        if (act == TS_ACT_RELU) {          // A_prev > 0 implies the unit was kept
                            x0 = p0 > 0.0f ? x0 * sc : 0.0f;
                            x1 = p1 > 0.0f ? x1 * sc : 0.0f;
                        } else {                           // prev = the stored pre-activation; gelu'(0) = 1/2, so the keep factor is explicit
                            x0 = x0 * gelu_grad_f(p0) * keep[2 * e2];
                            x1 = x1 * gelu_grad_f(p1) * keep[2 * e2 + 1];
                        }
                        const __nv_bfloat162 o2 = __floats2bfloat162_rn(x0, x1);
                        if (row < B) { cs[g * 8 + 2 * e2] += __low2float(o2); cs[g * 8 + 2 * e2 + 1] += __high2float(o2); }
                        ow[e2] = *reinterpret_cast<const uint32_t *>(&o2);
                    }
                    if (row < B) *reinterpret_cast<uint4 *>(dH + row * K + col + 32 * g) = make_uint4(ow[0], ow[1], ow[2], ow[3]);
                }
            }
        }
        if ((blk & 1) && colsum != nullptr) {              // one bias-gradient partial per 32 rows: fold the 8 row lanes
            const int64_t prow = blockIdx.y * 2 + (blk >> 1);
#pragma unroll
            for (int e = 0; e < 16; ++e) {
                float c = cs[e];
                c += __shfl_xor_sync(0xffffffffu, c, 4);
                c += __shfl_xor_sync(0xffffffffu, c, 8);
                c += __shfl_xor_sync(0xffffffffu, c, 16);
                cs[e] = c;
            }
            if (i == 0 && prow * 32 < B) {
                float *o = colsum + prow * colsum_stride + col;
#pragma unroll
                for (int g = 0; g < 2; ++g) {
                    *reinterpret_cast<float4 *>(o + 32 * g) = make_float4(cs[g * 8], cs[g * 8 + 1], cs[g * 8 + 2], cs[g * 8 + 3]);
                    *reinterpret_cast<float4 *>(o + 32 * g + 4) = make_float4(cs[g * 8 + 4], cs[g * 8 + 5], cs[g * 8 + 6], cs[g * 8 + 7]);
                }
            }
#pragma unroll
            for (int e = 0; e < 16; ++e) cs[e] = 0.0f;
        }
    }
}

bool head_dh_supported(int K, int C, int64_t colsum_stride) { return K % 4 == 0 && C >= 1 && C <= 16 && colsum_stride % 4 == 0; }

int launch_head_dh(const __nv_bfloat16 *dzb, const __nv_bfloat16 *Wp, const __nv_bfloat16 *prev, int64_t B, int K, int C, int act,
                   const DropoutSpec &drop, int training, __nv_bfloat16 *dH, float *colsum, int64_t colsum_stride, cudaStream_t st) {
    TS_REQUIRE(head_dh_supported(K, C, colsum_stride), TS_E_INVALID, "head dH: width %d / classes %d not supported", K, C);
    if (K % 64 == 0) {
        const dim3 grid_mma((unsigned)((K / 64 + 7) / 8), (unsigned)((B + 63) / 64));
        TS_CUDA_CHECK(launch_pdl(head_dh_mma_kernel, grid_mma, dim3(256), 0, st, 1, dzb, Wp, prev, B, K, act, drop, training, dH, colsum,
                                 colsum_stride));
        TS_LAUNCH_CHECK();
        return TS_OK;
    }
    const dim3 grid((unsigned)((K / 4 + 255) / 256), (unsigned)((B + 31) / 32));
#define TS_HEAD_DH(CP_) \
    TS_CUDA_CHECK(launch_pdl(head_dh_kernel<CP_>, grid, dim3(256), 0, st, 1, dzb, Wp, prev, B, K, act, drop, training, dH, colsum, colsum_stride))
    if (C <= 4) TS_HEAD_DH(4);
    else if (C <= 8) TS_HEAD_DH(8);
    else if (C <= 12) TS_HEAD_DH(12);
    else TS_HEAD_DH(16);
#undef TS_HEAD_DH
    TS_LAUNCH_CHECK();
    return TS_OK;
}

template <typename T, int MAXC>
__global__ void __launch_bounds__(128) head_wgrad_kernel(const T *__restrict__ A, const float *__restrict__ dz, int64_t B,
                                                         int K, int C, float *__restrict__ w_partial, int64_t w_stride,
                                                         float *__restrict__ b_partial, int64_t b_stride, int rows_per_split) {
    __shared__ float sdz[64][MAXC];
    const int k = blockIdx.x * 128 + threadIdx.x;
    const int s = blockIdx.y;
    const int64_t r0 = (int64_t)s * rows_per_split;
    const int64_t r1 = min(B, r0 + rows_per_split);
    float acc[MAXC];
#pragma unroll
    for (int c = 0; c < MAXC; ++c) acc[c] = 0.0f;
    float bsum = 0.0f;   // thread c < C of block x == 0 also sums column c of dz
    for (int64_t rt = r0; rt < r1; rt += 64) {
        const int nr = (int)min((int64_t)64, r1 - rt);
        for (int i = threadIdx.x; i < 64 * MAXC; i += 128) {
            const int rr = i / MAXC, c = i % MAXC;
            sdz[rr][c] = (rr < nr && c < C) ? dz[(rt + rr) * C + c] : 0.0f;
        }
        __syncthreads();
        if (k < K) {
            for (int rr = 0; rr < nr; ++rr) {
                const float a = to_f32(A[(rt + rr) * K + k]);
#pragma unroll
                for (int c = 0; c < MAXC; ++c) acc[c] = fmaf(sizeof(T) == 2 ? round_bf16(sdz[rr][c]) : sdz[rr][c], a, acc[c]);
            }
        }
        if (blockIdx.x == 0 && threadIdx.x < C)
            for (int rr = 0; rr < nr; ++rr) bsum += sdz[rr][threadIdx.x];
        __syncthreads();
    }
    if (k < K) {
#pragma unroll
        for (int c = 0; c < MAXC; ++c)
            if (c < C) w_partial[(int64_t)s * w_stride + (int64_t)c * K + k] = acc[c];
    }
    if (blockIdx.x == 0 && threadIdx.x < C) b_partial[(int64_t)s * b_stride + threadIdx.x] = bsum;
}

int launch_head_wgrad(const void *A, int a_is_bf16, const float *dz, int64_t B, int K, int C, float *w_partial,
                      int64_t w_stride, float *b_partial, int64_t b_stride, int splits, cudaStream_t st) {
    const int rows = (int)((B + splits - 1) / splits);
    dim3 grid((K + 127) / 128, splits);
#define TS_WG(T, MAXC)                                                                                       \
    head_wgrad_kernel<T, MAXC><<<grid, 128, 0, st>>>(reinterpret_cast<const T *>(A), dz, B, K, C, w_partial, \
                                                     w_stride, b_partial, b_stride, rows)
    if (a_is_bf16) { if (C <= 16) TS_WG(__nv_bfloat16, 16); else TS_WG(__nv_bfloat16, 32); }
    else { if (C <= 16) TS_WG(float, 16); else TS_WG(float, 32); }
#undef TS_WG
    TS_LAUNCH_CHECK();
    return TS_OK;
}

// =====================================================================================
// Small gradients of the last two layers in one launch, from what the head left behind:
//   dW_last[c,k] = sum_r dz[r,c] * A[r,k]     db_last[c] = sum_r dz[r,c]     db_hidden[k] = sum_r dH[r,k]
// A CTA covers 128 rows x 128 hidden columns in two passes: thread = (row group g of 8, column quad), so every
// thread has 8 row loads of A and of dH in flight at once; the 8 row groups are then folded through shared memory in a fixed
// order and the CTA writes one partial per output (row block s = blockIdx.y).
// =====================================================================================
template <typename T, int MAXC>
__global__ void __launch_bounds__(256) small_grads_kernel(const T *__restrict__ A, const T *__restrict__ dH,
                                                          const float *__restrict__ dz, int64_t B, int K, int C,
                                                          float *__restrict__ w_partial, int64_t w_stride,
                                                          float *__restrict__ bl_partial, int64_t bl_stride,
                                                          float *__restrict__ bh_partial, int64_t bh_stride) {
    constexpr int ROWS = 128;                                  // rows per CTA = rows per partial
    __shared__ __align__(16) float sdz[ROWS][MAXC];
    extern __shared__ __align__(16) float red_dyn[];            // [8 row groups][MAXC + 1 outputs][128 columns]; output MAXC = db_hidden
    float(*red)[MAXC + 1][128] = reinterpret_cast<float(*)[MAXC + 1][128]>(red_dyn);
    const int g = threadIdx.x >> 5, quad = threadIdx.x & 31;
    const int k = blockIdx.x * 128 + quad * 4;
    const int64_t r0 = (int64_t)blockIdx.y * ROWS;
    for (int i = threadIdx.x; i < ROWS * MAXC; i += 256) {
        const int rr = i / MAXC, c = i % MAXC;
        sdz[rr][c] = (r0 + rr < B && c < C) ? dz[(r0 + rr) * C + c] : 0.0f;
    }
    __syncthreads();
    float acc[4][MAXC];
    float accb[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int c = 0; c < MAXC; ++c) acc[j][c] = 0.0f;
    if (k < K) {
#pragma unroll 1
        for (int pass = 0; pass < ROWS / 64; ++pass) {
            float a[8][4], h[8][4];
#pragma unroll
            for (int i = 0; i < 8; ++i) {                      // 16 independent 8-byte loads in flight per thread
                const int64_t r = r0 + pass * 64 + i * 8 + g;
                if (r < B) { load4<T>(A + r * K + k, a[i]); load4<T>(dH + r * K + k, h[i]); }
                else { a[i][0] = a[i][1] = a[i][2] = a[i][3] = 0.0f; h[i][0] = h[i][1] = h[i][2] = h[i][3] = 0.0f; }
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int rr = pass * 64 + i * 8 + g;
                float d[MAXC];
#pragma unroll
                for (int c = 0; c < MAXC; c += 4) {
                    const float4 t = *reinterpret_cast<const float4 *>(&sdz[rr][c]);
                    d[c] = t.x; d[c + 1] = t.y; d[c + 2] = t.z; d[c + 3] = t.w;
                }
                if (sizeof(T) == 2) {   // bf16 mode: the weight-gradient contraction takes the bf16-rounded dZ (as on the tensor cores)
#pragma unroll
                    for (int c = 0; c < MAXC; ++c) d[c] = round_bf16(d[c]);
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    accb[j] += h[i][j];
#pragma unroll
                    for (int c = 0; c < MAXC; ++c) acc[j][c] = fmaf(d[c], a[i][j], acc[j][c]);
                }
            }
        }
    }
    // every warp parks its accumulators, then all threads fold the 8 row groups in a fixed order (bit-reproducible)
#pragma unroll
    for (int c = 0; c <= MAXC; ++c) {
        float4 t;
        if (c < MAXC) t = make_float4(acc[0][c], acc[1][c], acc[2][c], acc[3][c]);
        else t = make_float4(accb[0], accb[1], accb[2], accb[3]);
        *reinterpret_cast<float4 *>(&red[g][c][quad * 4]) = t;
    }
    __syncthreads();
    for (int o = threadIdx.x; o < 128 * (MAXC + 1); o += 256) {
        const int col = o & 127, c = o >> 7;
        float v[8];
#pragma unroll
        for (int gg = 0; gg < 8; ++gg) v[gg] = red[gg][c][col];
        const float sum = ((v[0] + v[1]) + (v[2] + v[3])) + ((v[4] + v[5]) + (v[6] + v[7]));
        const int kk = blockIdx.x * 128 + col;
        if (kk >= K) continue;
        if (c < C) w_partial[(int64_t)blockIdx.y * w_stride + (int64_t)c * K + kk] = sum;
        else if (c == MAXC) bh_partial[(int64_t)blockIdx.y * bh_stride + kk] = sum;
    }
    if (blockIdx.x == 0 && g == 0) {
        // db_last: lanes split the 128 rows four ways per class, then shuffle-fold (fixed order)
        for (int c = 0; c < C; ++c) {
            float part = (sdz[quad][c] + sdz[quad + 32][c]) + (sdz[quad + 64][c] + sdz[quad + 96][c]);
            part = warp_sum(part);
            if (quad == 0) bl_partial[(int64_t)blockIdx.y * bl_stride + c] = part;
        }
    }
}

int launch_small_grads(const void *A, const void *dH, int is_bf16, const float *dz, int64_t B, int K, int C, float *w_partial,
                       int64_t w_stride, float *bl_partial, int64_t bl_stride, float *bh_partial, int64_t bh_stride,
                       cudaStream_t st) {
    TS_REQUIRE(C <= 16, TS_E_UNSUPPORTED, "small-gradient kernel handles at most 16 classes");
    dim3 grid((K + 127) / 128, (unsigned)((B + 127) / 128));
    const size_t smem = 8 * 17 * 128 * sizeof(float);
    static bool attr_set[2][64] = {{false}};
    int dev = 0;
    TS_CUDA_CHECK(cudaGetDevice(&dev));
    if (!attr_set[is_bf16 ? 1 : 0][dev & 63]) {
        if (is_bf16) TS_CUDA_CHECK(cudaFuncSetAttribute(small_grads_kernel<__nv_bfloat16, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        else TS_CUDA_CHECK(cudaFuncSetAttribute(small_grads_kernel<float, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_set[is_bf16 ? 1 : 0][dev & 63] = true;
    }
    if (is_bf16)
        small_grads_kernel<__nv_bfloat16, 16><<<grid, 256, smem, st>>>((const __nv_bfloat16 *)A, (const __nv_bfloat16 *)dH, dz, B, K, C, w_partial,
                                                                    w_stride, bl_partial, bl_stride, bh_partial, bh_stride);
    else
        small_grads_kernel<float, 16><<<grid, 256, smem, st>>>((const float *)A, (const float *)dH, dz, B, K, C, w_partial, w_stride,
                                                            bl_partial, bl_stride, bh_partial, bh_stride);
    TS_LAUNCH_CHECK();
    return TS_OK;
}

// =====================================================================================
// Column sums (hidden-layer bias gradients): thread per column, `splits` row ranges.
// =====================================================================================
template <typename T>
__global__ void __launch_bounds__(128) colsum_kernel(const T *__restrict__ X, int64_t B, int H, float *__restrict__ partial,
                                                     int64_t stride, int rows_per_split) {
    const int h = blockIdx.x * 128 + threadIdx.x;
    if (h >= H) return;
    const int64_t r0 = (int64_t)blockIdx.y * rows_per_split;
    const int64_t r1 = min(B, r0 + rows_per_split);
    float acc = 0.0f;
    for (int64_t r = r0; r < r1; ++r) acc += to_f32(X[r * H + h]);
    partial[(int64_t)blockIdx.y * stride + h] = acc;
}

int launch_colsum(const void *X, int is_bf16, int64_t B, int H, float *partial, int64_t stride, int splits,
                  cudaStream_t st) {
    const int rows = (int)((B + splits - 1) / splits);
    dim3 grid((H + 127) / 128, splits);
    if (is_bf16)
        colsum_kernel<__nv_bfloat16><<<grid, 128, 0, st>>>(reinterpret_cast<const __nv_bfloat16 *>(X), B, H, partial, stride, rows);
    else
        colsum_kernel<float><<<grid, 128, 0, st>>>(reinterpret_cast<const float *>(X), B, H, partial, stride, rows);
    TS_LAUNCH_CHECK();
    return TS_OK;
}

// =====================================================================================
// Flat optimizer: one block per 1024-element chunk of the flat parameter buffer; sums the
// split partials (fixed order), optionally applies SGD / Adam exactly as torch's single-tensor
// CPU path does (torch/optim/adam.py _single_tensor_adam, sgd.py _single_tensor_sgd), and
// refreshes the bf16 shadow of weight matrices.
// =====================================================================================
__global__ void __launch_bounds__(256) optimizer_kernel(OptArgs p) {
    extern __shared__ float4 slots[];
    pdl_launch_dependents();
    const int chunk = (int)blockIdx.x + p.chunk_begin;
    const int tix = p.chunk_tensor[chunk];          // static map: may be read before the predecessor has finished
    pdl_wait();
    if (tix < 0) return;
    const TensorDesc d = p.t[tix];
    const int64_t i0 = (int64_t)chunk * kChunk + threadIdx.x * 4;
    const int64_t j0 = i0 - d.offset;
    if (j0 >= d.count) return;
    const int n = (int)min((int64_t)4, d.count - j0);

    float g[4] = {0.0f, 0.0f, 0.0f, 0.0f};
    if (d.nsplit > 0) {
        const float *q = p.partials + d.partial_off + j0;
        sum_partials(q, d.partial_stride, d.nsplit, g, slots, p.partial_depth);
        if (n == 4) *reinterpret_cast<float4 *>(p.grads + i0) = make_float4(g[0], g[1], g[2], g[3]);
        else for (int e = 0; e < n; ++e) p.grads[i0 + e] = g[e];
    } else {
        const float4 t = *reinterpret_cast<const float4 *>(p.grads + i0);
        g[0] = t.x; g[1] = t.y; g[2] = t.z; g[3] = t.w;
    }
    if (!p.update) return;
    apply_update(p, d, i0, j0, n, g);
}

// =====================================================================================
// Data-parallel optimizer with the gradient all-reduce folded in (no separate collective, no separate partial-sum
// launch). Every rank keeps its local gradient in a peer-mapped buffer (NVLink / NVSwitch). This kernel
//   1. sums the split partials (fixed order) into this rank's peer-visible buffer,
//   2. publishes "my gradient of step s is complete" into every rank's flag array and waits until the local flag array
//      shows step s for every rank,
//   3. reads the same chunk of all `world` gradient buffers straight over the peer mappings, sums them in rank
//      order (identical order on every rank -> bit-identical replicas), scales by 1/world and applies SGD / Adam.
// One-shot (every rank reads everything) is the right shape here: the whole gradient is well under 1 MB, so latency,
// not bandwidth, is what the step pays. Buffers alternate with the step parity, so nobody overwrites a gradient a
// peer may still be reading (a rank passes step s+1's wait only after every peer finished the kernel of step s).
// Every CTA spins in part 2, so the grid must be co-resident: the launcher caps it and the CTAs loop over chunks.
// =====================================================================================
__device__ __forceinline__ uint32_t global_ns32() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return (uint32_t)t;
}

__global__ void __launch_bounds__(256) dp_optimizer_kernel(OptArgs p, DpArgs dp, int n_chunks) {
    extern __shared__ float4 slots[];
    pdl_launch_dependents();
    uint32_t *own = dp.flags[dp.rank];
    uint32_t *trace = own + 16;                        // spare words of this rank's flag array: a 5-point time trace (ns)
    const bool tracer = blockIdx.x == 1 && threadIdx.x == 0;
    if (tracer) trace[0] = global_ns32();
    pdl_wait();
    if (tracer) trace[1] = global_ns32();

    // 1. local gradient: sum the split partials of this CTA's chunks into the peer-visible buffer (p.grads is this
    //    rank's buffer of this step's parity)
    for (int chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x) {
        const int tix = p.chunk_tensor[chunk];
        if (tix < 0) continue;
        const TensorDesc d = p.t[tix];
        const int64_t i0 = (int64_t)chunk * kChunk + threadIdx.x * 4;
        const int64_t j0 = i0 - d.offset;
        if (j0 >= d.count || d.nsplit <= 0) continue;
        const int n = (int)min((int64_t)4, d.count - j0);
        const float *q = p.partials + d.partial_off + j0;
        float g[4] = {0.0f, 0.0f, 0.0f, 0.0f};
        sum_partials(q, d.partial_stride, d.nsplit, g, slots, p.partial_depth);
        // padding lanes (e >= n) carry zeros: the flat buffers are padded to whole chunks
        *reinterpret_cast<float4 *>(p.grads + i0) = make_float4(g[0], n > 1 ? g[1] : 0.0f, n > 2 ? g[2] : 0.0f, n > 3 ? g[3] : 0.0f);
    }

    // 2. one handshake per step: the last CTA of this rank to finish part 1 publishes "my gradient of step s is
    //    complete" into every rank's flag array (release, system scope). Per-CTA system-scope releases, or a full system
    //    fence in every CTA, cost several microseconds each while the next batch's cast streams through the same SMs.
    __shared__ bool s_last, s_timed_out;
    if (threadIdx.x == 0) s_timed_out = false;
    __syncthreads();
    if (tracer) trace[5] = global_ns32();
    if (threadIdx.x == 0) {
        __threadfence();                                           // this CTA's gradient stores before its ticket
        const uint32_t arrived = atomicAdd(own + 8, 1u);
        s_last = arrived == gridDim.x - 1;
        if (s_last) { own[8] = 0u; trace[6] = global_ns32(); }     // next step's kernel starts from zero
    }
    __syncthreads();
    if (s_last && threadIdx.x < dp.world) {
        // one thread per destination rank: the system-scope releases (a few microseconds each) run side by side
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(dp.flags[threadIdx.x] + dp.rank), "r"(dp.step) : "memory");
        if (threadIdx.x == 0) trace[7] = global_ns32();
    }
    if (threadIdx.x < dp.world && !wait_for_flag(own + threadIdx.x, dp.step, dp)) s_timed_out = true;
    __syncthreads();
    if (s_timed_out) return;                // a peer never arrived: nothing is updated, the host reports TS_E_TIMEOUT
    if (tracer) trace[2] = global_ns32();

    // 3. all ranks' gradient chunk in rank order (identical order on every rank), scaled update
    for (int chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x) {
        const int tix = p.chunk_tensor[chunk];
        if (tix < 0) continue;
        const TensorDesc d = p.t[tix];
        const int64_t i0 = (int64_t)chunk * kChunk + threadIdx.x * 4;
        const int64_t j0 = i0 - d.offset;
        if (j0 >= d.count) continue;
        const int n = (int)min((int64_t)4, d.count - j0);
        float4 v[8];
#pragma unroll
        for (int r = 0; r < 8; ++r)   // flat buffers are padded to whole chunks: the float4 is always in bounds
            v[r] = (r < dp.world) ? __ldcv(reinterpret_cast<const float4 *>(dp.grads[r] + i0)) : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        float g[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
        for (int r = 0; r < 8; ++r)
            if (r < dp.world) { g[0] += v[r].x; g[1] += v[r].y; g[2] += v[r].z; g[3] += v[r].w; }
        if (tracer && chunk == 1) trace[3] = global_ns32();
        apply_update(p, d, i0, j0, n, g);
        if (tracer && chunk == 1) trace[4] = global_ns32();
    }
}

// =====================================================================================
// Data-parallel optimizer for large models, reduction inside the NVSwitch (NVLS). All buffers involved live in symmetric
// memory with a multicast mapping. Per step, one kernel:
//   1. sum the split partials of every chunk into this rank's gradient buffer              (as dp_optimizer_kernel)
//   2. barrier A across ranks: every rank's gradient is complete
//   3. this rank owns the chunks c with c % world == rank: one multimem.ld_reduce per float4 returns the sum over all
//      ranks (added up by the switch), SGD / Adam on it (optimizer state is only ever touched for owned chunks), and
//      the new parameters + bf16 shadows go out through multimem.st: one store, delivered to every rank's buffer
//   4. barrier B across ranks: everybody's parameter stores have landed (and everybody is done reading the gradients, so
//      one gradient buffer suffices)
// Per GPU and step this moves ~1x the gradient out and ~1.4x in over NVLink and does 1/world of the optimizer work,
// against all-reduce (2x each way) followed by the full optimizer.
// =====================================================================================
__device__ __forceinline__ bool rank_barrier(uint32_t *own, const DpArgs &dp, int ticket_slot, uint32_t value, bool sys_fence) {
    __shared__ bool s_last_b, s_timed_out_b;
    if (threadIdx.x == 0) s_timed_out_b = false;
    __syncthreads();
    if (threadIdx.x == 0) {
        if (sys_fence) __threadfence_system(); else __threadfence();
        const uint32_t arrived = atomicAdd(own + ticket_slot, 1u);
        s_last_b = arrived == gridDim.x - 1;
        if (s_last_b) own[ticket_slot] = 0u;
    }
    __syncthreads();
    if (s_last_b && threadIdx.x < dp.world)
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(dp.flags[threadIdx.x] + dp.rank), "r"(value) : "memory");
    if (threadIdx.x < dp.world && !wait_for_flag(own + threadIdx.x, value, dp)) s_timed_out_b = true;
    __syncthreads();
    return !s_timed_out_b;                  // false: a peer never arrived (error word written); the caller returns
}

__global__ void __launch_bounds__(256) nvls_optimizer_kernel(OptArgs p, DpArgs dp, const float *grads_mc, int n_chunks) {
    extern __shared__ float4 slots[];
    pdl_launch_dependents();
    uint32_t *own = dp.flags[dp.rank];
    pdl_wait();
    for (int chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x) {
        const int tix = p.chunk_tensor[chunk];
        if (tix < 0) continue;
        const TensorDesc d = p.t[tix];
        const int64_t i0 = (int64_t)chunk * kChunk + threadIdx.x * 4;
        const int64_t j0 = i0 - d.offset;
        if (j0 >= d.count || d.nsplit <= 0) continue;
        const int n = (int)min((int64_t)4, d.count - j0);
        float g[4] = {0.0f, 0.0f, 0.0f, 0.0f};
        sum_partials(p.partials + d.partial_off + j0, d.partial_stride, d.nsplit, g, slots, p.partial_depth);
        *reinterpret_cast<float4 *>(p.grads + i0) = make_float4(g[0], n > 1 ? g[1] : 0.0f, n > 2 ? g[2] : 0.0f, n > 3 ? g[3] : 0.0f);
    }
    if (!rank_barrier(own, dp, 8, 2u * dp.step, false)) return;
    const int n_owned = (n_chunks - dp.rank + dp.world - 1) / dp.world;
    // four owned chunks per round: a reduction through the switch takes a few microseconds to come back, so one request per
    // thread at a time leaves the links mostly idle
    for (int j = blockIdx.x * 4; j < n_owned; j += gridDim.x * 4) {
        float g[4][4];
        int tix[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int chunk = (j + u) * dp.world + dp.rank;
            tix[u] = (j + u < n_owned) ? p.chunk_tensor[chunk] : -1;
            if (tix[u] < 0) continue;
            const int64_t i0 = (int64_t)chunk * kChunk + threadIdx.x * 4;
            if (i0 - p.t[tix[u]].offset >= p.t[tix[u]].count) { tix[u] = -1; continue; }
            asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                         : "=f"(g[u][0]), "=f"(g[u][1]), "=f"(g[u][2]), "=f"(g[u][3]) : "l"(grads_mc + i0) : "memory");
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (tix[u] < 0) continue;
            const int chunk = (j + u) * dp.world + dp.rank;
            const TensorDesc d = p.t[tix[u]];
            const int64_t i0 = (int64_t)chunk * kChunk + threadIdx.x * 4;
            const int64_t j0 = i0 - d.offset;
            apply_update(p, d, i0, j0, (int)min((int64_t)4, d.count - j0), g[u]);
        }
    }
    (void)rank_barrier(own, dp, 9, 2u * dp.step + 1u, true);
}

// =====================================================================================
// The same exchange for ONE LAYER's chunk range [chunk_begin, chunk_begin + n_chunks), meant to run on a side stream
// next to the backward GEMMs of the earlier layers (engine.cu, nvls overlap). The GEMMs are persistent CTAs that own a
// whole SM each, so this kernel must stay on the few SMs they leave free: its grid is one 1024-thread CTA per reserved
// SM (the link, not the SM count, bounds a multimem reduction), each CTA working as four 256-thread groups that take
// one 1024-element chunk each and keep two reductions in flight per thread.
// =====================================================================================
constexpr int kLayerThreads = 1024, kLayerGroups = kLayerThreads / 256, kLayerUnroll = 2;
__global__ void __launch_bounds__(kLayerThreads, 1) nvls_layer_kernel(OptArgs p, DpArgs dp, const float *grads_mc, int n_chunks) {
    extern __shared__ float4 slots[];
    uint32_t *own = dp.flags[dp.rank];
    const int tid = threadIdx.x & 255, grp = threadIdx.x >> 8;
    const int groups = (int)blockDim.x >> 8;                 // 4 (fat CTAs on reserved SMs) or 1 (co-resident with the GEMM CTAs)
    const int vblock = blockIdx.x * groups + grp, vgrid = gridDim.x * groups;
    // 1. this rank's split partials of the range (hidden-layer bias gradients in bf16 mode) into its gradient buffer
    for (int c = vblock; c < n_chunks; c += vgrid) {
        const int chunk = p.chunk_begin + c;
        const int tix = p.chunk_tensor[chunk];
        if (tix < 0) continue;
        const TensorDesc d = p.t[tix];
        const int64_t i0 = (int64_t)chunk * kChunk + tid * 4;
        const int64_t j0 = i0 - d.offset;
        if (j0 >= d.count || d.nsplit <= 0) continue;
        const int n = (int)min((int64_t)4, d.count - j0);
        float g[4] = {0.0f, 0.0f, 0.0f, 0.0f};
        if (p.partial_depth > 0) {
            sum_partials(p.partials + d.partial_off + j0, d.partial_stride, d.nsplit, g, slots, p.partial_depth, blockDim.x);
        } else {    // co-resident launch: no shared memory to stage through; these are the bias tensors, a few thousand elements
            const float *q = p.partials + d.partial_off + j0;
            for (int sp = 0; sp < d.nsplit; ++sp) {
                const float4 t = *reinterpret_cast<const float4 *>(q + (int64_t)sp * d.partial_stride);
                g[0] += t.x; g[1] += t.y; g[2] += t.z; g[3] += t.w;
            }
        }
        *reinterpret_cast<float4 *>(p.grads + i0) = make_float4(g[0], n > 1 ? g[1] : 0.0f, n > 2 ? g[2] : 0.0f, n > 3 ? g[3] : 0.0f);
    }
    if (!rank_barrier(own, dp, 8, 2u * dp.step, false)) return;
    // 2. owned chunks of the range: chunk % world == rank, as in the whole-buffer kernel (the optimizer state of a chunk
    //    always lives on the same rank, whichever kernel updates it)
    const int first = ((dp.rank - p.chunk_begin) % dp.world + dp.world) % dp.world;
    const int n_owned = n_chunks > first ? (n_chunks - first + dp.world - 1) / dp.world : 0;
    for (int j = vblock * kLayerUnroll; j < n_owned; j += vgrid * kLayerUnroll) {
        float g[kLayerUnroll][4];
        int tix[kLayerUnroll];
#pragma unroll
        for (int u = 0; u < kLayerUnroll; ++u) {
            const int chunk = p.chunk_begin + first + (j + u) * dp.world;
            tix[u] = (j + u < n_owned) ? p.chunk_tensor[chunk] : -1;
            if (tix[u] < 0) continue;
            const int64_t i0 = (int64_t)chunk * kChunk + tid * 4;
            if (i0 - p.t[tix[u]].offset >= p.t[tix[u]].count) { tix[u] = -1; continue; }
            asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                         : "=f"(g[u][0]), "=f"(g[u][1]), "=f"(g[u][2]), "=f"(g[u][3]) : "l"(grads_mc + i0) : "memory");
        }
#pragma unroll
        for (int u = 0; u < kLayerUnroll; ++u) {
            if (tix[u] < 0) continue;
            const int chunk = p.chunk_begin + first + (j + u) * dp.world;
            const TensorDesc d = p.t[tix[u]];
            const int64_t i0 = (int64_t)chunk * kChunk + tid * 4;
            const int64_t j0 = i0 - d.offset;
            apply_update(p, d, i0, j0, (int)min((int64_t)4, d.count - j0), g[u]);
        }
    }
    (void)rank_barrier(own, dp, 9, 2u * dp.step + 1u, true);
}

// =====================================================================================
// One layer's optimizer update on a single GPU, co-resident with the backward GEMMs of the layers below (side stream): the
// chunk range [chunk_begin, chunk_begin + n_chunks) of the flat buffers, 256 threads and at most 64 registers per thread, no
// dynamic shared memory, so a CTA fits on an SM next to a persistent GEMM CTA (which is capped at 128 registers).
// =====================================================================================
__global__ void __maxnreg__(64) layer_optimizer_kernel(OptArgs p, int n_chunks) {
    for (int c = blockIdx.x; c < n_chunks; c += gridDim.x) {
        const int chunk = p.chunk_begin + c;
        const int tix = p.chunk_tensor[chunk];
        if (tix < 0) continue;
        const TensorDesc d = p.t[tix];
        const int64_t i0 = (int64_t)chunk * kChunk + threadIdx.x * 4;
        const int64_t j0 = i0 - d.offset;
        if (j0 >= d.count) continue;
        const int n = (int)min((int64_t)4, d.count - j0);
        float g[4] = {0.0f, 0.0f, 0.0f, 0.0f};
        if (d.nsplit > 0) {      // split partials (bias tensors; small weight matrices): summed in order, straight from global memory
            const float *q = p.partials + d.partial_off + j0;
            for (int sp = 0; sp < d.nsplit; ++sp) {
                const float4 t = *reinterpret_cast<const float4 *>(q + (int64_t)sp * d.partial_stride);
                g[0] += t.x; g[1] += t.y; g[2] += t.z; g[3] += t.w;
            }
            if (n == 4) *reinterpret_cast<float4 *>(p.grads + i0) = make_float4(g[0], g[1], g[2], g[3]);
            else for (int e = 0; e < n; ++e) p.grads[i0 + e] = g[e];
        } else {
            const float4 t = *reinterpret_cast<const float4 *>(p.grads + i0);
            g[0] = t.x; g[1] = t.y; g[2] = t.z; g[3] = t.w;
        }
        apply_update(p, d, i0, j0, n, g);
    }
}

int launch_layer_optimizer(const OptArgs &a, int64_t n_chunks, int n_ctas, cudaStream_t st) {
    TS_REQUIRE(n_chunks >= 1 && n_ctas >= 1, TS_E_INVALID, "layer optimizer: %lld chunks, %d CTAs", (long long)n_chunks, n_ctas);
    const int64_t grid = n_chunks < n_ctas ? n_chunks : n_ctas;
    layer_optimizer_kernel<<<dim3((unsigned)grid), dim3(256), 0, st>>>(a, (int)n_chunks);
    TS_LAUNCH_CHECK();
    return TS_OK;
}

int launch_nvls_layer(const OptArgs &a, const DpArgs &dp, const float *grads_mc, int64_t n_chunks, int n_ctas, cudaStream_t st) {
    TS_REQUIRE(grads_mc != nullptr && a.params_mc != nullptr, TS_E_INVALID, "nvls optimizer: multicast mappings missing");
    TS_REQUIRE(n_ctas >= 1 && n_chunks >= 1, TS_E_INVALID, "nvls layer exchange: %d CTAs, %lld chunks", n_ctas, (long long)n_chunks);
    const size_t smem = (size_t)a.partial_depth * kLayerThreads * 16;
    static bool done[64] = {false};
    int dev = 0;
    TS_CUDA_CHECK(cudaGetDevice(&dev));
    if (!done[dev & 63]) {
        TS_CUDA_CHECK(cudaFuncSetAttribute(nvls_layer_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 * kLayerThreads * 16));
        done[dev & 63] = true;
    }
    TS_REQUIRE(a.partial_depth <= 4, TS_E_INVALID, "nvls layer exchange runs at partial depth <= 4");
    if (a.partial_depth == 0) {
        // co-resident form: 256 threads x 64 registers and no dynamic shared memory per CTA, so a CTA fits on an SM NEXT TO a
        // persistent GEMM CTA (352 threads x <= 128 registers, 225 KiB): no SM has to be taken from the backward pass
        int64_t grid = n_chunks < n_ctas ? n_chunks : n_ctas;
        nvls_layer_kernel<<<dim3((unsigned)grid), dim3(256), 0, st>>>(a, dp, grads_mc, (int)n_chunks);
        TS_LAUNCH_CHECK();
        return TS_OK;
    }
    int64_t grid = (n_chunks + kLayerGroups - 1) / kLayerGroups;
    if (grid > n_ctas) grid = n_ctas;
    nvls_layer_kernel<<<dim3((unsigned)grid), dim3(kLayerThreads), smem, st>>>(a, dp, grads_mc, (int)n_chunks);
    TS_LAUNCH_CHECK();
    return TS_OK;
}

// 64 KiB of dynamic shared memory needs the opt-in attribute, once per device.
static int ensure_optimizer_smem() {
    static bool done[64] = {false};
    int dev = 0;
    TS_CUDA_CHECK(cudaGetDevice(&dev));
    if (!done[dev & 63]) {
        TS_CUDA_CHECK(cudaFuncSetAttribute(dp_optimizer_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kPartialSmem));
        TS_CUDA_CHECK(cudaFuncSetAttribute(optimizer_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kPartialSmem));
        TS_CUDA_CHECK(cudaFuncSetAttribute(nvls_optimizer_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kPartialSmem));
        done[dev & 63] = true;
    }
    return TS_OK;
}

int launch_dp_optimizer(const OptArgs &a, const DpArgs &dp, int64_t n_chunks, int sm_count, cudaStream_t st) {
    // every CTA must be resident at once (they wait for each other and for the peer GPUs): with up to 64 KiB of shared
    // memory each 3 per SM fit; 2 leaves room for the early-running cast of the next batch
    int rc = ensure_optimizer_smem();
    if (rc) return rc;
    static int cached_dp[17] = {0};       // resident CTAs per SM by partial depth (same binary on every device of the process)
    int &per_sm_c = cached_dp[a.partial_depth & 15];
    if (per_sm_c == 0)
        TS_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_c, dp_optimizer_kernel, 256, (size_t)a.partial_depth * 4096));
    int per_sm = per_sm_c;
    TS_REQUIRE(per_sm >= 1, TS_E_CUDA, "data-parallel optimizer kernel does not fit on an SM");
    if (per_sm > 2) per_sm = 2;
    int64_t grid = (int64_t)sm_count * per_sm;
    if (grid > n_chunks) grid = n_chunks;
    TS_CUDA_CHECK(launch_pdl(dp_optimizer_kernel, dim3((unsigned)grid), dim3(256), (size_t)a.partial_depth * 4096, st, 1, a, dp, (int)n_chunks));
    TS_LAUNCH_CHECK();
    return TS_OK;
}

__global__ void __launch_bounds__(256) reduce_partials_kernel(float *partials, ReduceArgs r) {
    pdl_launch_dependents();
    pdl_wait();
    const int t = blockIdx.z, grp = blockIdx.y;
    const int s0 = grp * kReduceFan;
    if (s0 >= r.nsplit[t]) return;
    const int64_t j0 = ((int64_t)blockIdx.x * 256 + threadIdx.x) * 4;
    if (j0 >= r.count[t]) return;
    const int s1 = min(s0 + kReduceFan, r.nsplit[t]);
    const float *q = partials + r.off[t] + j0;          // partial rows are padded past `count`: whole float4 loads
    float g[4] = {0.0f, 0.0f, 0.0f, 0.0f};
    for (int s = s0; s < s1; s += 8) {
        float4 v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u)
            v[u] = (s + u < s1) ? *reinterpret_cast<const float4 *>(q + (int64_t)(s + u) * r.stride[t]) : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
#pragma unroll
        for (int u = 0; u < 8; ++u)
            if (s + u < s1) { g[0] += v[u].x; g[1] += v[u].y; g[2] += v[u].z; g[3] += v[u].w; }
    }
    *reinterpret_cast<float4 *>(partials + r.dst_off[t] + (int64_t)grp * r.stride[t] + j0) = make_float4(g[0], g[1], g[2], g[3]);
}

int launch_reduce_partials(float *partials, const ReduceArgs &r, cudaStream_t st) {
    int64_t max_count = 0;
    int max_groups = 0;
    for (int t = 0; t < r.n; ++t) {
        max_count = r.count[t] > max_count ? r.count[t] : max_count;
        const int g = (r.nsplit[t] + kReduceFan - 1) / kReduceFan;
        max_groups = g > max_groups ? g : max_groups;
    }
    const dim3 grid((unsigned)((max_count + 1023) / 1024), (unsigned)max_groups, (unsigned)r.n);
    TS_CUDA_CHECK(launch_pdl(reduce_partials_kernel, grid, dim3(256), 0, st, 1, partials, r));
    TS_LAUNCH_CHECK();
    return TS_OK;
}

int launch_nvls_optimizer(const OptArgs &a, const DpArgs &dp, const float *grads_mc, int64_t n_chunks, int sm_count, cudaStream_t st) {
    int rc = ensure_optimizer_smem();
    if (rc) return rc;
    TS_REQUIRE(grads_mc != nullptr && a.params_mc != nullptr, TS_E_INVALID, "nvls optimizer: multicast mappings missing");
    for (int t = 0; t < a.n_tensors; ++t)
        TS_REQUIRE(a.shadow_mc == nullptr || a.t[t].shadow_off < 0 || a.t[t].count % 4 == 0, TS_E_UNSUPPORTED,
                   "nvls optimizer: weight tensor %d has %lld elements, not a multiple of 4", t, (long long)a.t[t].count);
    // every CTA spins at the two barriers: the grid must be co-resident (64 KiB of shared memory per CTA for small models:
    // 3 per SM; 16 KiB for large ones: the 8 that the thread count allows)
    static int cached_nvls[17] = {0};
    int &per_sm_c = cached_nvls[a.partial_depth & 15];
    if (per_sm_c == 0)
        TS_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_c, nvls_optimizer_kernel, 256, (size_t)a.partial_depth * 4096));
    int per_sm = per_sm_c;
    TS_REQUIRE(per_sm >= 1, TS_E_CUDA, "nvls optimizer kernel does not fit on an SM");
    if (per_sm > 4) per_sm = 4;           // room for the early-running cast of the next batch
    int64_t grid = (int64_t)sm_count * per_sm;
    if (grid > n_chunks) grid = n_chunks;
    TS_CUDA_CHECK(launch_pdl(nvls_optimizer_kernel, dim3((unsigned)grid), dim3(256), (size_t)a.partial_depth * 4096, st, 1, a, dp, grads_mc,
                             (int)n_chunks));
    TS_LAUNCH_CHECK();
    return TS_OK;
}

int launch_optimizer(const OptArgs &a, int64_t n_chunks, cudaStream_t st) {
    int rc = ensure_optimizer_smem();
    if (rc) return rc;
    TS_CUDA_CHECK(launch_pdl(optimizer_kernel, dim3((unsigned)n_chunks), dim3(256), (size_t)a.partial_depth * 4096, st, 1, a));
    TS_LAUNCH_CHECK();
    return TS_OK;
}

// =====================================================================================
// wait_at_end = 0: ordinary stream order (dst may still be read by the kernel in front).
// wait_at_end = 1: the conversion runs concurrently with the kernel in front (the previous step's optimizer) and only
//   waits for it before exiting, so "this kernel is complete" still implies "everything before it is complete" for the
//   kernel behind. The caller guarantees nothing in flight touches dst (alternating buffers, see engine.cu).
__global__ void __launch_bounds__(256) cast_bf16_kernel(const float *__restrict__ src, __nv_bfloat16 *__restrict__ dst, int64_t n, int wait_at_end) {
    pdl_launch_dependents();
    if (!wait_at_end) pdl_wait();
    const int64_t stride = (int64_t)gridDim.x * blockDim.x * 8;
    for (int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 8; i < n; i += stride) {
        if (i + 8 <= n) {
            const float4 a = *reinterpret_cast<const float4 *>(src + i);
            const float4 b = *reinterpret_cast<const float4 *>(src + i + 4);
            __nv_bfloat162 o[4] = {__floats2bfloat162_rn(a.x, a.y), __floats2bfloat162_rn(a.z, a.w),
                                   __floats2bfloat162_rn(b.x, b.y), __floats2bfloat162_rn(b.z, b.w)};
            *reinterpret_cast<uint4 *>(dst + i) = *reinterpret_cast<uint4 *>(o);
        } else {
            for (int64_t e = i; e < n; ++e) dst[e] = __float2bfloat16_rn(src[e]);
        }
    }
    if (wait_at_end) pdl_wait();
}

// uint8 pixels -> lut[pixel] (ToTensor + Normalize folded into 256 values), 16 pixels per thread and iteration.
template <typename OutT>
__global__ void __launch_bounds__(256) cast_u8_kernel(const uint8_t *__restrict__ src, OutT *__restrict__ dst, int64_t n,
                                                      const float *__restrict__ lut, int wait_at_end) {
    __shared__ float s_lut[256];
    pdl_launch_dependents();
    if (!wait_at_end) pdl_wait();
    s_lut[threadIdx.x] = lut[threadIdx.x];
    __syncthreads();
    const int64_t stride = (int64_t)gridDim.x * blockDim.x * 16;
    for (int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 16; i < n; i += stride) {
        if (i + 16 <= n && ((reinterpret_cast<uintptr_t>(src + i) & 15) == 0)) {
            const uint4 raw = *reinterpret_cast<const uint4 *>(src + i);
            const uint32_t wv[4] = {raw.x, raw.y, raw.z, raw.w};
            float v[16];
#pragma unroll
            for (int k = 0; k < 16; ++k) v[k] = s_lut[(wv[k >> 2] >> ((k & 3) * 8)) & 0xFFu];
            if (sizeof(OutT) == 2) {
                uint32_t pk[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * k], v[2 * k + 1]);
                    pk[k] = *reinterpret_cast<const uint32_t *>(&h);
                }
                uint4 *o = reinterpret_cast<uint4 *>(dst + i);
                o[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                o[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
            } else {
                float4 *o = reinterpret_cast<float4 *>(dst + i);
#pragma unroll
                for (int k = 0; k < 4; ++k) o[k] = make_float4(v[4 * k], v[4 * k + 1], v[4 * k + 2], v[4 * k + 3]);
            }
        } else {
            for (int64_t e = i; e < n && e < i + 16; ++e) from_f32(dst[e], s_lut[src[e]]);
        }
    }
    if (wait_at_end) pdl_wait();
}

int launch_cast_u8(const uint8_t *src, __nv_bfloat16 *dst_bf16, float *dst_f32, int64_t n, const float *lut, cudaStream_t st,
                   bool wait_at_end) {
    if (n <= 0) return TS_OK;
    TS_REQUIRE((dst_bf16 != nullptr) != (dst_f32 != nullptr) && lut != nullptr, TS_E_INVALID, "cast_u8: one destination and a table");
    int64_t blocks = (n / 16 + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (blocks < 1) blocks = 1;
    if (dst_bf16)
        TS_CUDA_CHECK(launch_pdl(cast_u8_kernel<__nv_bfloat16>, dim3((unsigned)blocks), dim3(256), 0, st, 1, src, dst_bf16, n, lut,
                                 wait_at_end ? 1 : 0));
    else
        TS_CUDA_CHECK(launch_pdl(cast_u8_kernel<float>, dim3((unsigned)blocks), dim3(256), 0, st, 1, src, dst_f32, n, lut,
                                 wait_at_end ? 1 : 0));
    TS_LAUNCH_CHECK();
    return TS_OK;
}

int launch_cast_bf16(const float *src, __nv_bfloat16 *dst, int64_t n, cudaStream_t st, bool wait_at_end) {
    if (n <= 0) return TS_OK;
    int64_t blocks = (n / 8 + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (blocks < 1) blocks = 1;
    TS_CUDA_CHECK(launch_pdl(cast_bf16_kernel, dim3((unsigned)blocks), dim3(256), 0, st, 1, src, dst, n, wait_at_end ? 1 : 0));
    TS_LAUNCH_CHECK();
    return TS_OK;
}

// =====================================================================================
// Metrics (metrics.py:13-27): fold per-batch records into the running state in slot order.
// =====================================================================================
__global__ void metrics_accumulate_kernel(ts_metrics_state *state, const ts_batch_record *ring, int64_t ring_len,
                                          int64_t first, int64_t count) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    double ls = state->loss_sum;
    int64_t nc = state->n_correct, nt = state->n_total;
    for (int64_t i = 0; i < count; ++i) {
        const ts_batch_record r = ring[(first + i) % ring_len];
        ls += (double)r.loss;          // Mean.update: weighted_sum += loss (fp64), weights += 1
        nc += r.n_correct;             // MulticlassAccuracy.update: num_correct += (pred == tgt).sum()
        nt += r.n_rows;                //                            num_total += B
    }
    state->loss_sum = ls;
    state->n_batches += count;
    state->n_correct = nc;
    state->n_total = nt;
}

int launch_metrics_accumulate(ts_metrics_state *state, const ts_batch_record *ring, int64_t ring_len, int64_t first,
                              int64_t count, cudaStream_t st) {
    if (count <= 0) return TS_OK;
    metrics_accumulate_kernel<<<1, 32, 0, st>>>(state, ring, ring_len, first, count);
    TS_LAUNCH_CHECK();
    return TS_OK;
}

__global__ void __launch_bounds__(256) metrics_update_raw_kernel(ts_metrics_state *state, const float *loss,
                                                                 const int64_t *pred, const int64_t *tgt, int64_t B) {
    __shared__ int s_c[8];
    int c = 0;
    for (int64_t i = threadIdx.x; i < B; i += blockDim.x) c += (pred[i] == tgt[i]);
    c = warp_sum(c);
    if ((threadIdx.x & 31) == 0) s_c[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int w = 0; w < 8; ++w) t += s_c[w];
        state->loss_sum += (double)loss[0];
        state->n_batches += 1;
        state->n_correct += t;
        state->n_total += B;
    }
}

int launch_metrics_update_raw(ts_metrics_state *state, const float *loss, const int64_t *pred, const int64_t *tgt,
                              int64_t B, cudaStream_t st) {
    metrics_update_raw_kernel<<<1, 256, 0, st>>>(state, loss, pred, tgt, B);
    TS_LAUNCH_CHECK();
    return TS_OK;
}

}  // namespace tsb
