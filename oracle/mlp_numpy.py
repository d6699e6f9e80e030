"""CPU oracle for the TorchSystem training hot path (TEST INFRASTRUCTURE ONLY).

This module restates, in plain numpy, the arithmetic that the reference's
`train` / `validate` handlers drive per batch. Only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s `cpu_baseline` leg may import it; the
product path (`torchsystem_b200/`) never does and fails loudly without its CUDA
library.

Reference call sites restated here (paths relative to /root/reference):

* `examples/tinysys/modules/mlp.py:13-18`         MLP.forward: Linear -> Dropout -> act -> Linear
* `examples/tinysys/tinysys/classifier.py:22-27`  Classifier.forward / loss (CrossEntropyLoss, mean)
* `examples/tinysys/tinysys/classifier.py:29-35`  Classifier.fit (zero_grad, fwd, loss, bwd, step, argmax)
* `examples/tinysys/tinysys/classifier.py:37-39`  Classifier.evaluate
* `examples/tinysys/main.py:62`                   Adam(lr=1e-3) with torch defaults
* `examples/tinysys/tinysys/metrics.py:8-27`      Metrics (torcheval Mean + MulticlassAccuracy)

The arithmetic itself lives in third-party packages that are not vendored in the
reference tree: torch (pinned 2.5.1 in poetry.lock:1372, 2.11.0 on this image)
and torcheval 0.0.7 (poetry.lock:1424, not installable here). Their published
algorithms are restated below:

* Linear:           y = x @ W.T + b                       (W is [out, in], row-major)
* Dropout (train):  y = x * (bernoulli(1-p) / (1-p))      (mask drawn by the caller)
* CrossEntropyLoss: mean_i(logsumexp(z_i) - z_i[y_i])
* argmax(dim=1):    first maximum wins on ties
* Adam:             torch/optim/adam.py `_single_tensor_adam` (non-capturable branch)
* SGD:              torch/optim/sgd.py `_single_tensor_sgd` (weight decay, momentum, dampening, Nesterov)
*                   Adam also with weight decay (L2 or decoupled = AdamW) and amsgrad
* Mean:             fp64 weighted_sum += loss, weights += 1 ; compute = sum / weights
* MulticlassAccuracy(micro): num_correct += (pred == tgt).sum(), num_total += B

Parity pinning: the reference's own tests hold no golden vector for this path
(SURVEY.md section 4), so the oracle is pinned against outputs of the reference itself:
`tests/golden/make_golden.py` imports the unmodified reference handlers,
Classifier and MLP from /root/reference, runs them on seeded inputs and commits
the results; `tests/test_oracle.py` checks this module against those files.

`emulate_bf16=True` reproduces the rounding points of the bf16 tensor-core mode
of the CUDA path (operands rounded to bf16, fp32 accumulation), so predictions
can be compared bit-exactly; see DESIGN.md "Numerics".
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

ACT_RELU = 0
ACT_GELU = 1
ACT_IDENTITY = 2

OPT_SGD = 0
OPT_ADAM = 1


def bf16_round(a: np.ndarray) -> np.ndarray:
    """Round fp32 values to the nearest bf16 (ties to even); result stays fp32."""
    u = np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)
    bias = ((u >> np.uint32(16)) & np.uint32(1)) + np.uint32(0x7FFF)
    return ((u + bias) & np.uint32(0xFFFF0000)).view(np.float32)


def _act(u: np.ndarray, kind: int) -> np.ndarray:
    if kind == ACT_RELU:
        return np.maximum(u, np.float32(0))
    if kind == ACT_GELU:  # exact (erf) GELU, torch.nn.GELU() default
        from math import sqrt
        from scipy.special import erf  # scipy ships on both boxes
        return (u * np.float32(0.5) * (np.float32(1) + erf(u / np.float32(sqrt(2.0))))).astype(np.float32)
    return u


def _act_grad(u: np.ndarray, a: np.ndarray, kind: int) -> np.ndarray:
    """d act(u) / du. ReLU uses the output sign like torch's threshold_backward."""
    if kind == ACT_RELU:
        return (a > 0).astype(np.float32)
    if kind == ACT_GELU:
        from math import sqrt, pi
        from scipy.special import erf
        cdf = np.float32(0.5) * (np.float32(1) + erf(u / np.float32(sqrt(2.0))))
        pdf = np.exp(np.float32(-0.5) * u * u) / np.float32(sqrt(2.0 * pi))
        return (cdf + u * pdf).astype(np.float32)
    return np.ones_like(u)


@dataclass
class OracleMLP:
    """Parameters + optimizer state of the aggregate, all fp32 numpy arrays.

    `weights[l]` is [out_l, in_l] (torch Linear layout), `biases[l]` is [out_l].
    Hidden layers apply dropout then the activation (mlp.py:14-16); the last
    layer is a plain Linear (mlp.py:17).
    """
    weights: list[np.ndarray]
    biases: list[np.ndarray]
    dropout: float = 0.0
    activation: int = ACT_RELU
    opt_kind: int = OPT_ADAM
    lr: float = 1e-3
    betas: tuple[float, float] = (0.9, 0.999)
    eps: float = 1e-8
    weight_decay: float = 0.0
    decoupled_weight_decay: bool = False   # AdamW (adam.py:417-419)
    amsgrad: bool = False
    momentum: float = 0.0                  # SGD (sgd.py:358-372)
    dampening: float = 0.0
    nesterov: bool = False
    emulate_bf16: bool = False
    step: int = 0
    exp_avg: list[np.ndarray] = field(default_factory=list)
    exp_avg_sq: list[np.ndarray] = field(default_factory=list)
    max_exp_avg_sq: list[np.ndarray] = field(default_factory=list)
    momentum_buffer: list = field(default_factory=list)      # None until the first step, like torch

    def __post_init__(self):
        self.weights = [np.array(w, dtype=np.float32) for w in self.weights]
        self.biases = [np.array(b, dtype=np.float32) for b in self.biases]
        if not self.exp_avg:
            self.exp_avg = [np.zeros_like(p) for p in self.parameters()]
            self.exp_avg_sq = [np.zeros_like(p) for p in self.parameters()]
        if not self.max_exp_avg_sq:
            self.max_exp_avg_sq = [np.zeros_like(p) for p in self.parameters()]
        if not self.momentum_buffer:
            self.momentum_buffer = [None for _ in self.parameters()]

    def parameters(self) -> list[np.ndarray]:
        """torch order: layer0.weight, layer0.bias, layer1.weight, ... (mlp.py:8-11)."""
        out = []
        for w, b in zip(self.weights, self.biases):
            out += [w, b]
        return out

    # ---------------------------------------------------------------- forward
    def _forward(self, x: np.ndarray, masks: list[np.ndarray] | None):
        rnd = bf16_round if self.emulate_bf16 else (lambda t: t)
        a = rnd(np.ascontiguousarray(x, dtype=np.float32).reshape(x.shape[0], -1))  # Flatten, classifier.py:23
        acts, pre = [a], []
        n = len(self.weights)
        scale = np.float32(1.0) / np.float32(1.0 - self.dropout)
        for l in range(n - 1):
            u = a @ rnd(self.weights[l]).T + self.biases[l]
            if masks is not None and self.dropout > 0.0:
                u = u * (masks[l].astype(np.float32) * scale)
            pre.append(u)
            a = rnd(_act(u, self.activation))
            acts.append(a)
        z = a @ rnd(self.weights[-1]).T + self.biases[-1]
        return z.astype(np.float32), acts, pre

    @staticmethod
    def _cross_entropy(z: np.ndarray, y: np.ndarray):
        zmax = z.max(axis=1, keepdims=True)
        e = np.exp(z - zmax)
        s = e.sum(axis=1, keepdims=True)
        lse = (np.log(s) + zmax)[:, 0]
        nll = lse - z[np.arange(z.shape[0]), y]
        loss = np.float32(nll.astype(np.float32).sum(dtype=np.float64) / z.shape[0])
        soft = (e / s).astype(np.float32)
        return loss, soft

    def evaluate(self, x: np.ndarray, y: np.ndarray):
        """classifier.py:37-39: dropout off, no state change."""
        z, _, _ = self._forward(x, None)
        loss, _ = self._cross_entropy(z, y)
        return z.argmax(axis=1).astype(np.int64), loss

    def fit(self, x: np.ndarray, y: np.ndarray, masks: list[np.ndarray] | None = None, grad_hook=None):
        """classifier.py:29-35. `masks[l]` is the 0/1 keep mask of hidden layer l
        (what `bernoulli_(1-p)` drew); returns pre-update predictions and loss.
        `grad_hook(grads)` may replace the gradient list (data-parallel averaging)."""
        rnd = bf16_round if self.emulate_bf16 else (lambda t: t)
        B = x.shape[0]
        z, acts, pre = self._forward(x, masks)
        loss, soft = self._cross_entropy(z, y)
        pred = z.argmax(axis=1).astype(np.int64)

        dz = soft.copy()
        dz[np.arange(B), y] -= np.float32(1)
        dz = (dz / np.float32(B)).astype(np.float32)

        n = len(self.weights)
        gw, gb = [None] * n, [None] * n
        scale = np.float32(1.0) / np.float32(1.0 - self.dropout)
        # last layer: every tensor-core contraction takes the bf16-rounded dz (dW = dZ^T A, dA = dZ W);
        # the bias gradient sums the fp32 dz
        gw[-1] = rnd(dz).T @ acts[-1]
        gb[-1] = dz.sum(axis=0)
        da = rnd(dz) @ rnd(self.weights[-1])
        for l in range(n - 2, -1, -1):
            g = _act_grad(rnd(pre[l]), acts[l + 1], self.activation)     # tensor-core mode stores the pre-activation in bf16
            if masks is not None and self.dropout > 0.0:
                g = g * (masks[l].astype(np.float32) * scale)
            dh = (da * g).astype(np.float32)
            dhr = rnd(dh)
            gb[l] = dhr.sum(axis=0)          # column sums of the stored (bf16 in tensor-core mode) dH
            gw[l] = dhr.T @ acts[l]
            if l > 0:
                da = dhr @ rnd(self.weights[l])
        grads = []
        for w, b in zip(gw, gb):
            grads += [w.astype(np.float32), b.astype(np.float32)]
        if grad_hook is not None:
            grads = grad_hook(grads)
        self._optimizer_step(grads)
        return pred, loss

    # -------------------------------------------------------------- optimizer
    def _optimizer_step(self, grads: list[np.ndarray]):
        self.step += 1
        params = self.parameters()
        f = np.float32
        wd = self.weight_decay
        if self.opt_kind == OPT_SGD:                                   # sgd.py _single_tensor_sgd
            for i, (p, g) in enumerate(zip(params, grads)):
                if wd != 0:
                    g = g + f(wd) * p                                  # grad.add(param, alpha=weight_decay)
                if self.momentum != 0:
                    buf = self.momentum_buffer[i]
                    if buf is None:
                        buf = g.copy()                                 # first step: clone(grad)
                    else:
                        buf *= f(self.momentum)
                        buf += f(1.0 - self.dampening) * g             # buf.mul_(momentum).add_(grad, alpha=1-dampening)
                    self.momentum_buffer[i] = buf
                    g = g + f(self.momentum) * buf if self.nesterov else buf
                p -= f(self.lr) * g                                    # param.add_(grad, alpha=-lr)
            return
        b1, b2 = self.betas
        bc1 = 1.0 - b1 ** self.step
        bc2 = 1.0 - b2 ** self.step
        step_size = self.lr / bc1
        bc2_sqrt = bc2 ** 0.5
        for p, g, m, v, mx in zip(params, grads, self.exp_avg, self.exp_avg_sq, self.max_exp_avg_sq):
            if wd != 0:
                if self.decoupled_weight_decay:
                    p *= f(1.0 - self.lr * wd)                         # param.mul_(1 - lr * weight_decay)
                else:
                    g = g + f(wd) * p
            m += (g - m) * f(1.0 - b1)                                 # exp_avg.lerp_(grad, 1-beta1)
            v *= f(b2)
            v += f(1.0 - b2) * g * g                                   # mul_(beta2).addcmul_(g, g, 1-beta2)
            if self.amsgrad:
                np.maximum(mx, v, out=mx)                              # torch.maximum(max_exp_avg_sq, exp_avg_sq)
                denom = np.sqrt(mx) / f(bc2_sqrt) + f(self.eps)
            else:
                denom = np.sqrt(v) / f(bc2_sqrt) + f(self.eps)
            p -= f(step_size) * (m / denom)                            # addcdiv_(m, denom, -step_size)


class OracleMetrics:
    """metrics.py:8-27 with torcheval 0.0.7's Mean / MulticlassAccuracy restated.

    `Mean.update(loss)` on a 0-d loss adds it with weight 1, so `compute()['loss']`
    is the unweighted mean of per-batch mean losses (a short last batch counts as
    much as a full one); the state is fp64. Accuracy is micro-averaged.
    """

    def __init__(self):
        self.reset()

    def update(self, batch: int, loss, predictions: np.ndarray, targets: np.ndarray) -> None:
        self.weighted_sum += float(np.float32(loss))
        self.weights += 1.0
        self.num_correct += int((np.asarray(predictions) == np.asarray(targets)).sum())
        self.num_total += int(np.asarray(targets).shape[0])

    def compute(self) -> dict:
        loss = self.weighted_sum / self.weights if self.weights else 0.0
        acc = self.num_correct / self.num_total if self.num_total else float('nan')
        return {'loss': loss, 'accuracy': acc}

    def reset(self) -> None:
        self.weighted_sum = 0.0
        self.weights = 0.0
        self.num_correct = 0
        self.num_total = 0
