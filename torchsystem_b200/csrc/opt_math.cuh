// One element of the fused optimizer: SGD / Adam exactly as torch's single-tensor CPU path orders the operations
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

}  // namespace tsb
