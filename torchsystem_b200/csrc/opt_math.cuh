// Device-side pieces of the fused optimizer.
// opt_update_element: one element: SGD / Adam exactly as torch's single-tensor CPU path orders the operations
// (torch/optim/sgd.py _single_tensor_sgd, adam.py _single_tensor_adam). Shared by the flat optimizer kernels
// (simt_kernels.cu) and the single-kernel small-batch step (step_kernel.cu), so both apply bit-identical updates.
#pragma once

#include "kernels.h"

namespace tsb {

// g: this element's gradient (already summed); w: parameter; m: exp_avg or SGD momentum buffer; v: exp_avg_sq;
// x: max_exp_avg_sq (amsgrad). Unused state is passed through untouched.
__device__ __forceinline__ void opt_update_element(const OptArgs &p, float g, float &w, float &m, float &v, float &x) {
    float grad = g * p.grad_scale;
    if (p.kind == TS_OPT_SGD) {                                        // sgd.py _single_tensor_sgd
        if (p.weight_decay != 0.0f) grad = grad + p.weight_decay * w;  // grad.add(param, alpha=weight_decay)
        if (p.momentum != 0.0f) {
            // buf = clone(grad) on the first step, else buf.mul_(momentum).add_(grad, alpha=1 - dampening)
            m = p.first_step ? grad : m * p.momentum + p.one_minus_dampening * grad;
            grad = p.nesterov ? grad + p.momentum * m : m;
        }
        w = w - p.lr * grad;                                           // param.add_(grad, alpha=-lr)
    } else {                                                           // adam.py _single_tensor_adam
        if (p.weight_decay != 0.0f) {
            if (p.decoupled) w = w * p.wd_mul;                         // param.mul_(1 - lr * weight_decay)   (AdamW)
            else grad = grad + p.weight_decay * w;
        }
        m = m + p.one_minus_beta1 * (grad - m);                        // exp_avg.lerp_(grad, 1 - beta1)
        v = v * p.beta2;                                               // exp_avg_sq.mul_(beta2)
        v = v + p.one_minus_beta2 * (grad * grad);                     //   .addcmul_(grad, grad, value=1 - beta2)
        float second = v;
        if (p.amsgrad) { x = fmaxf(x, v); second = x; }                // torch.maximum(max_exp_avg_sq, exp_avg_sq)
        const float denom = sqrtf(second) / p.bc2_sqrt + p.eps;        // (sqrt(.) / sqrt(bias_correction2)).add_(eps)
        w = w - p.step_size * (m / denom);                             // param.addcdiv_(exp_avg, denom, value=-step_size)
    }
}


// SGD / Adam on up to four consecutive elements, exactly as torch's single-tensor CPU path orders the operations.
__device__ __forceinline__ void apply_update(const OptArgs &p, const TensorDesc &d, int64_t i0, int64_t j0, int n, const float (&g)[4]) {
    // Whole float4 accesses (the flat buffers are padded to 1024-element chunks and i0 is a multiple of 4); all
    // loads are issued before the first use, the padding lanes (e >= n) are written back unchanged.
    const float4 w4 = *reinterpret_cast<const float4 *>(p.params + i0);
    float4 m4 = make_float4(0.0f, 0.0f, 0.0f, 0.0f), v4 = m4, x4 = m4;
    const bool adam = p.kind != TS_OPT_SGD;
    const bool use_m = adam || p.momentum != 0.0f;       // exp_avg doubles as SGD's momentum buffer
    const bool use_x = adam && p.amsgrad;
    if (use_m) m4 = *reinterpret_cast<const float4 *>(p.exp_avg + i0);
    if (adam) v4 = *reinterpret_cast<const float4 *>(p.exp_avg_sq + i0);
    if (use_x) x4 = *reinterpret_cast<const float4 *>(p.max_exp_avg_sq + i0);
    float w[4] = {w4.x, w4.y, w4.z, w4.w}, m[4] = {m4.x, m4.y, m4.z, m4.w}, v[4] = {v4.x, v4.y, v4.z, v4.w};
    float x[4] = {x4.x, x4.y, x4.z, x4.w};
#pragma unroll
    for (int e = 0; e < 4; ++e)
        if (e < n) opt_update_element(p, g[e], w[e], m[e], v[e], x[e]);
    if (p.params_mc != nullptr)      // one store, delivered to this address in every rank's parameter buffer by the NVSwitch
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p.params_mc + i0), "f"(w[0]), "f"(w[1]),
                     "f"(w[2]), "f"(w[3]) : "memory");
    else
        *reinterpret_cast<float4 *>(p.params + i0) = make_float4(w[0], w[1], w[2], w[3]);
    if (use_m) *reinterpret_cast<float4 *>(p.exp_avg + i0) = make_float4(m[0], m[1], m[2], m[3]);
    if (adam) *reinterpret_cast<float4 *>(p.exp_avg_sq + i0) = make_float4(v[0], v[1], v[2], v[3]);
    if (use_x) *reinterpret_cast<float4 *>(p.max_exp_avg_sq + i0) = make_float4(x[0], x[1], x[2], x[3]);
    if (d.shadow_off >= 0 && p.shadow != nullptr) {
        __nv_bfloat16 *sh = p.shadow + d.shadow_off + j0;
        if (p.shadow_mc != nullptr) {    // the launcher has checked that every shadowed tensor is a whole number of quads
            __nv_bfloat162 lo = __floats2bfloat162_rn(w[0], w[1]), hi = __floats2bfloat162_rn(w[2], w[3]);
            asm volatile("multimem.st.relaxed.sys.global.v2.bf16x2 [%0], {%1,%2};" ::"l"(p.shadow_mc + d.shadow_off + j0),
                         "r"(*reinterpret_cast<uint32_t *>(&lo)), "r"(*reinterpret_cast<uint32_t *>(&hi)) : "memory");
        } else if (n == 4 && (((uintptr_t)sh) & 7) == 0) {
            __nv_bfloat162 lo = __floats2bfloat162_rn(w[0], w[1]), hi = __floats2bfloat162_rn(w[2], w[3]);
            uint2 pk;
            pk.x = *reinterpret_cast<uint32_t *>(&lo);
            pk.y = *reinterpret_cast<uint32_t *>(&hi);
            *reinterpret_cast<uint2 *>(sh) = pk;
        } else {
            for (int e = 0; e < n; ++e) sh[e] = __float2bfloat16_rn(w[e]);
        }
    }
}

// Sum of `nsplit` partial rows at q (row stride `stride`), s ascending: a fixed order. Always whole float4 loads (partial
// rows are padded past `count`, so a tail thread may read and ignore up to three floats of padding). The loads go
// through cp.async into this thread's own shared-memory slots, kPartialDepth of them in flight per round, and are
// added in order afterwards: with plain loads ptxas interleaves the adds with the loads and the in-order issue then
// pays close to one memory round trip per partial -- and the tensors with one partial per 128-row block (32 at batch
// 4096) are the critical path of the whole launch.
constexpr int kPartialDepth = 16;                          // small models; large ones run 4 deep to keep 8 CTAs per SM
constexpr int kPartialSmem = kPartialDepth * 256 * 16;     // most dynamic shared memory of the optimizer kernels
__device__ __forceinline__ void sum_partials(const float *q, int64_t stride, int nsplit, float (&g)[4], float4 *slots, int depth,
                                             uint32_t nthreads = 256u) {
    const uint32_t base = (uint32_t)__cvta_generic_to_shared(slots + threadIdx.x);
    for (int s = 0; s < nsplit; s += depth) {
#pragma unroll
        for (int u = 0; u < kPartialDepth; ++u)
            if (u < depth && s + u < nsplit)
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(base + (uint32_t)u * nthreads * 16u), "l"(q + (int64_t)(s + u) * stride) : "memory");
        asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
#pragma unroll
        for (int u = 0; u < kPartialDepth; ++u)
            if (u < depth && s + u < nsplit) {
                const float4 v = slots[u * nthreads + threadIdx.x];
                g[0] += v.x; g[1] += v.y; g[2] += v.z; g[3] += v.w;
            }
    }
}



// Wait until *flag has reached `value` (release/acquire at system scope: the flag is written by a peer GPU). The wait is
// bounded by wall-clock time (dp.timeout_ns on %globaltimer, minutes by default like a collective library's watchdog, not
// by a poll count: a peer may legitimately be seconds late, e.g. while rank 0 writes a checkpoint). On expiry the step
// number goes into the host-visible error word and the caller skips its update; nothing traps, the context stays usable.
__device__ __forceinline__ bool wait_for_flag(const uint32_t *flag, uint32_t value, const DpArgs &dp) {
    unsigned long long t0 = 0;
    for (uint32_t spins = 0;; ++spins) {
        uint32_t seen;
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(flag) : "memory");
        if ((int32_t)(seen - value) >= 0) return true;
        __nanosleep(32);
        if ((spins & 1023u) == 1023u) {
            unsigned long long now;
            asm volatile("mov.u64 %0, %globaltimer;" : "=l"(now));
            if (t0 == 0) t0 = now;
            else if (now - t0 > dp.timeout_ns) {
                if (dp.err != nullptr) *reinterpret_cast<volatile uint32_t *>(dp.err) = dp.step ? dp.step : 1u;
                return false;
            }
        }
    }
}

}  // namespace tsb
