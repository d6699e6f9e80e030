"""Generate the golden vectors under tests/golden/ from the UNMODIFIED reference.

Run in the build container only (needs /root/reference; the GPU box has none):

    python tests/golden/make_golden.py

What runs: the reference's own `iterate` / `train` / `validate` handlers
(/root/reference/examples/tinysys/tinysys/services/training.py:21-48) through
`service.handle('iterate', ...)`, its `Classifier` (tinysys/classifier.py) and its
`MLP` (modules/mlp.py), on torch CPU (version recorded in each file). The only
stand-in is `Metrics`: tinysys/metrics.py needs torcheval 0.0.7, which is not
installable here, so `oracle/ref_torch.CpuMetrics` (a restatement of torcheval's
Mean / MulticlassAccuracy) is used and flagged as such in DESIGN.md.

Recorded per case: inputs, initial weights, the dropout keep-masks the reference
drew (re-drawn from the same generator state right before each `fit`), every
batch's (predictions, loss), each phase's `results` as carried by the
Trained / Validated events, and the final weights and optimizer state.
"""

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = ['/root/reference', '/root/reference/examples/tinysys', ROOT]

from torch.nn import ReLU, GELU  # noqa: E402
from torchsystem.registry import register, gethash  # noqa: E402
from torchsystem.services import Consumer  # noqa: E402
from tinysys.services import training  # noqa: E402
from tinysys.classifier import Classifier  # noqa: E402
from modules.mlp import MLP  # noqa: E402
from oracle.ref_torch import CpuMetrics, draw_dropout_mask, make_layers  # noqa: E402

register(MLP)
register(ReLU)
register(GELU)


def synthetic(n: int, seed: int = 1234):
    """unit-normal 28x28 inputs, labels from a fixed random teacher."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(n, 1, 28, 28, generator=g, dtype=torch.float32)
    teacher = torch.randn(784, 10, generator=g, dtype=torch.float32)
    y = torch.argmax(x.flatten(1) @ teacher, dim=1)
    return x, y


class ListLoader:
    """A Loader (tinysys/domain.py:26-29) over pre-cut batches; when `hidden` is set it
    records the dropout mask(s) the next `fit` call will draw."""

    def __init__(self, x, y, batch, hidden=None, p=0.0, masks=None):
        self.batches = [(x[i:i + batch], y[i:i + batch]) for i in range(0, x.shape[0], batch)]
        self.hidden, self.p, self.masks = hidden, p, masks

    def __iter__(self):
        for xb, yb in self.batches:
            if self.hidden is not None and self.p > 0.0:
                # the masks of successive Dropout layers are drawn in forward order
                state = torch.get_rng_state()
                drawn = [torch.empty(xb.shape[0], h).bernoulli_(1.0 - self.p).to(torch.uint8) for h in self.hidden]
                torch.set_rng_state(state)
                self.masks.append(drawn)
            yield xb, yb


class RecordingMetrics(CpuMetrics):
    def __init__(self):
        super().__init__()
        self.log = []

    def update(self, batch, loss, predictions, targets):
        super().update(batch, loss, predictions, targets)
        self.log.append((batch, float(loss.item()), predictions.clone().numpy()))


def run_case(name, dims, dropout, activation, opt, lr, batch, n_train, n_val, epochs, last_partial=0):
    x, y = synthetic(n_train + n_val + last_partial)
    xt, yt = x[:n_train + last_partial], y[:n_train + last_partial]
    xv, yv = x[n_train + last_partial:], y[n_train + last_partial:]

    torch.manual_seed(0)
    act = ReLU() if activation == 'relu' else GELU()
    if len(dims) == 3:
        net = MLP(dims[0], dims[1], dims[2], dropout=dropout, activation=act)
        model_hash = gethash(net)
    else:
        net = make_layers(dims, dropout, activation)   # deeper net through the same reference Classifier
        model_hash = ''
    init = [p.detach().clone().numpy() for p in net.parameters()]
    criterion = torch.nn.CrossEntropyLoss()
    optimizer = (torch.optim.Adam(net.parameters(), lr=lr) if opt == 'adam'
                 else torch.optim.SGD(net.parameters(), lr=lr))
    classifier = Classifier(net, criterion, optimizer)

    masks: list = []
    loaders = [('train', ListLoader(xt, yt, batch, hidden=dims[1:-1], p=dropout, masks=masks)),
               ('validation', ListLoader(xv, yv, batch))]
    metrics = RecordingMetrics()

    events = []
    consumer = Consumer()

    @consumer.handler
    def on_phase(event: training.Trained | training.Validated):
        events.append((type(event).__name__, event.model.epoch,
                       float(event.results['loss'].item()), float(event.results['accuracy'].item())))

    @consumer.handler
    def on_iterated(event: training.Iterated):
        events.append(('Iterated', event.model.epoch, 0.0, 0.0))

    training.producer.consumers.clear() if hasattr(training.producer, 'consumers') else None
    training.producer.register(consumer)
    training.provider.override(training.device, lambda: 'cpu')

    torch.manual_seed(42)
    for _ in range(epochs):
        training.service.handle('iterate', classifier, loaders, metrics)

    out = {
        'torch_version': np.array(torch.__version__),
        'dims': np.array(dims), 'dropout': np.array(dropout), 'activation': np.array(activation),
        'optimizer': np.array(opt), 'lr': np.array(lr), 'batch': np.array(batch), 'epochs': np.array(epochs),
        'model_hash': np.array(model_hash), 'final_epoch': np.array(classifier.epoch),
        'x_train': xt.numpy(), 'y_train': yt.numpy(), 'x_val': xv.numpy(), 'y_val': yv.numpy(),
        'batch_index': np.array([b for b, _, _ in metrics.log]),
        'batch_loss': np.array([l for _, l, _ in metrics.log], dtype=np.float32),
        'batch_pred': np.concatenate([p for _, _, p in metrics.log]),
        'event_name': np.array([e[0] for e in events]),
        'event_epoch': np.array([e[1] for e in events]),
        'event_loss': np.array([e[2] for e in events], dtype=np.float64),
        'event_accuracy': np.array([e[3] for e in events], dtype=np.float64),
    }
    for i, p in enumerate(init):
        out[f'init_{i}'] = p
    for i, p in enumerate(net.parameters()):
        out[f'final_{i}'] = p.detach().numpy()
        st = optimizer.state.get(p, {})
        if 'exp_avg' in st:
            out[f'exp_avg_{i}'] = st['exp_avg'].numpy()
            out[f'exp_avg_sq_{i}'] = st['exp_avg_sq'].numpy()
            out['opt_step'] = np.array(int(st['step'].item()))
    for s, per_layer in enumerate(masks):
        for l, m in enumerate(per_layer):
            out[f'mask_{s}_{l}'] = np.packbits(m.numpy(), axis=1)
    out['n_fit_calls'] = np.array(len(masks))
    path = os.path.join(HERE, f'{name}.npz')
    np.savez_compressed(path, **out)
    print(f'{name}: {os.path.getsize(path) / 1e6:.2f} MB, events={events}')


if __name__ == '__main__':
    # the shipped aggregate (main.py:60-65): MLP(784,256,10,0.1,ReLU), Adam(1e-3), batch 32
    run_case('tinysys_adam_b32', [784, 256, 10], 0.1, 'relu', 'adam', 1e-3, 32, 128, 64, 2)
    # SGD (north-star addition), ragged last train batch (16 extra samples)
    run_case('sgd_h64_ragged', [784, 64, 10], 0.1, 'relu', 'sgd', 0.01, 32, 96, 64, 2, last_partial=16)
    # GELU activation (examples/tinysys/tests/test_storage.py:38), no dropout
    run_case('gelu_h64', [784, 64, 10], 0.0, 'gelu', 'adam', 1e-3, 32, 96, 32, 1)
    # deeper net (config-4 shape class) through the same reference Classifier
    run_case('deep_h64_h48', [784, 64, 48, 10], 0.1, 'relu', 'adam', 1e-3, 32, 96, 32, 2)
    print('hash goldens:', gethash(MLP(784, 256, 10, dropout=0.1, activation=ReLU())),
          gethash(MLP(784, 256, 10, dropout=0.5, activation=ReLU())))
