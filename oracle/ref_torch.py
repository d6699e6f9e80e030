"""The reference's CPU path, restated on stock torch (TEST / BASELINE INFRASTRUCTURE ONLY).

The reference's per-batch arithmetic is a fixed sequence of stock eager torch
calls. This module issues the same sequence, on the
host cores, without importing anything from /root/reference, so that it can
travel to the GPU box and serve as

* the checker for the `-m gpu` parity tests (`CpuClassifier.fit/evaluate`), and
* the timed CPU arm of `bench.py` (`cpu_baseline`, `--impl reference`).

It is never imported by the product package. Call sites followed
(relative to /root/reference):

* examples/tinysys/tinysys/services/training.py:29-37   train handler loop
* examples/tinysys/tinysys/services/training.py:39-48   validate handler loop (inference_mode)
* examples/tinysys/tinysys/classifier.py:22-39          forward / loss / fit / evaluate
* examples/tinysys/modules/mlp.py:8-18                  Linear -> Dropout -> activation -> Linear
* examples/tinysys/tinysys/metrics.py:13-27             Mean + MulticlassAccuracy (torcheval 0.0.7, restated)
* examples/tinysys/main.py:60-62                        MLP(784,256,10,0.1,ReLU), CrossEntropyLoss(), Adam(lr=1e-3)

`tests/test_oracle.py` pins it (and `oracle/mlp_numpy.py`) against the golden
files that `tests/golden/make_golden.py` produced from the unmodified reference.
"""
from __future__ import annotations

import torch
from torch import Tensor, nn
import torch.nn.functional as F


def make_layers(dims: list[int], dropout: float, activation: str = 'relu') -> nn.Sequential:
    """[784, 256, 10] -> Linear(784,256), Dropout(p), ReLU, Linear(256,10): mlp.py:8-11, any depth."""
    mods: list[nn.Module] = []
    for i in range(len(dims) - 2):
        mods += [nn.Linear(dims[i], dims[i + 1]), nn.Dropout(dropout),
                 nn.ReLU() if activation == 'relu' else nn.GELU()]
    mods.append(nn.Linear(dims[-2], dims[-1]))
    return nn.Sequential(*mods)


class CpuClassifier:
    """Same observable behaviour as the reference Classifier (classifier.py:9-39)
    on whatever device `net` lives on (CPU for the baseline)."""

    def __init__(self, net: nn.Module, optimizer: torch.optim.Optimizer):
        self.nn = net
        self.optimizer = optimizer
        self.epoch = 0

    @property
    def phase(self) -> str:
        return 'train' if self.nn.training else 'evaluation'

    @phase.setter
    def phase(self, value: str) -> None:
        self.nn.train() if value == 'train' else self.nn.eval()

    def _logits(self, inputs: Tensor) -> Tensor:
        return self.nn(inputs.flatten(1))

    def fit(self, inputs: Tensor, targets: Tensor) -> tuple[Tensor, Tensor]:
        self.optimizer.zero_grad()
        outputs = self._logits(inputs)
        loss = F.cross_entropy(outputs, targets)
        loss.backward()
        self.optimizer.step()
        return torch.argmax(outputs, dim=1), loss

    def evaluate(self, inputs: Tensor, targets: Tensor) -> tuple[Tensor, Tensor]:
        outputs = self._logits(inputs)
        return torch.argmax(outputs, dim=1), F.cross_entropy(outputs, targets)


class CpuMetrics:
    """torcheval Mean (fp64 state) + micro MulticlassAccuracy, metrics.py:8-27."""

    def __init__(self):
        self.reset()

    def update(self, batch: int, loss: Tensor, predictions: Tensor, targets: Tensor) -> None:
        self.weighted_sum += loss.detach().to(torch.float64).sum()
        self.weights += float(loss.numel())
        self.num_correct += (predictions == targets).sum()
        self.num_total += targets.shape[0]

    def compute(self) -> dict[str, Tensor]:
        return {'loss': self.weighted_sum / self.weights,
                'accuracy': self.num_correct.to(torch.float32) / self.num_total}

    def reset(self) -> None:
        self.weighted_sum = torch.zeros((), dtype=torch.float64)
        self.weights = 0.0
        self.num_correct = torch.zeros((), dtype=torch.int64)
        self.num_total = 0


def train(model: CpuClassifier, loader, metrics: CpuMetrics, device: str = 'cpu') -> dict[str, Tensor]:
    """training.py:29-37 without the event dispatch."""
    model.phase = 'train'
    for batch, (inputs, targets) in enumerate(loader, start=1):
        inputs, targets = inputs.to(device), targets.to(device)
        predictions, loss = model.fit(inputs, targets)
        metrics.update(batch, loss, predictions, targets)
    return metrics.compute()


def validate(model: CpuClassifier, loader, metrics: CpuMetrics, device: str = 'cpu') -> dict[str, Tensor]:
    """training.py:39-48 without the event dispatch."""
    with torch.inference_mode():
        model.phase = 'evaluation'
        for batch, (inputs, targets) in enumerate(loader, start=1):
            inputs, targets = inputs.to(device), targets.to(device)
            predictions, loss = model.evaluate(inputs, targets)
            metrics.update(batch, loss, predictions, targets)
        return metrics.compute()


def draw_dropout_mask(shape: tuple[int, ...], p: float) -> Tensor:
    """The keep mask that `nn.Dropout(p)` is about to draw on CPU, without
    disturbing the global generator: torch's CPU dropout is
    `empty_like(x).bernoulli_(1 - p)` on the default generator."""
    state = torch.get_rng_state()
    mask = torch.empty(shape, dtype=torch.float32).bernoulli_(1.0 - p)
    torch.set_rng_state(state)
    return mask.to(torch.uint8)
