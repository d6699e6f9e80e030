"""The wiring of the reference's examples/tinysys/main.py:49-71 on the B200 engine, runnable as is (python examples/main_b200.py).

Same steps as the reference driver: register the types whose constructor arguments define the aggregate's id, override the
`device` dependency, register consumers on the producer, build nn / criterion / optimizer, compile the aggregate, call
service.handle('iterate', classifier, loaders, metrics) once per epoch. What differs, and why:

* the aggregate is built by `lower_classifier` (a Compiler step returning B200Classifier) instead of `torch.compile`: the
  reference's pipeline ends in `bool(OptimizedModule)`, which raises on torch >= 2.6; this driver runs on
  the installed torch 2.11;
* data: MNIST cannot be downloaded here (no network), so the loaders hold a synthetic 28x28 10-class set of the same shapes
  and dtypes, resident on the device (DeviceLoader; batch 32 as shipped, main.py:65);
* consumers: a logging consumer with the reference's format (services/logging.py:15-21); the tinydb / tensorboard consumers
  need packages this image does not have (out of scope).

When `torchsystem` is importable its own Service / Producer / Consumer / register are used unchanged (torchsystem_b200.compat)."""
import argparse
import os
import sys
from logging import INFO, basicConfig, getLogger

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
from torch.nn import CrossEntropyLoss, ReLU
from torch.optim import Adam

from torchsystem_b200 import B200Metrics, DeviceLoader, MLP, lower_classifier, training
from torchsystem_b200.compat import Consumer, register

logger = getLogger('main_b200')


def synthetic_digits(n: int, seed: int):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(n, 1, 28, 28, generator=g)
    teacher = torch.randn(784, 10, generator=torch.Generator().manual_seed(99))
    return x, torch.argmax(x.flatten(1) @ teacher, dim=1)


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument('--epochs', type=int, default=10)               # main.py:70
    ap.add_argument('--batch-size', type=int, default=32)           # main.py:65
    ap.add_argument('--train-samples', type=int, default=60000)
    ap.add_argument('--val-samples', type=int, default=10000)
    ap.add_argument('--precision', default='bf16', choices=['bf16', 'f32'])
    args = ap.parse_args(argv)

    basicConfig(level=INFO)
    for cls in (MLP, ReLU, CrossEntropyLoss):                        # main.py:28-31: identity of the aggregate = md5 of these
        register(cls)
    register(Adam, excluded_args=[0])                                # main.py:32

    def device():
        return 'cuda'                                                # main.py:37-38; this engine has no CPU path

    history = []
    consumer = Consumer()

    @consumer.handler
    def on_results(event: training.Trained | training.Validated):    # services/logging.py:15-21
        name = 'train' if isinstance(event, training.Trained) else 'validation'
        loss, accuracy = event.results['loss'].item(), event.results['accuracy'].item()
        logger.info(f'--- {name} epoch {event.model.epoch}: loss={loss:.4f} accuracy={accuracy:.4f}')
        history.append((name, event.model.epoch, loss, accuracy))

    training.provider.override(training.device, device)              # main.py:49
    training.producer.consumers.clear()
    training.producer.register(consumer)                             # main.py:50-52

    torch.manual_seed(0)
    nn = MLP(784, 256, 10, dropout=0.1, activation=ReLU())           # main.py:60
    criterion = CrossEntropyLoss()                                   # main.py:61
    optimizer = Adam(nn.parameters(), lr=1e-3)                       # main.py:62
    metrics = B200Metrics(device=device())                           # main.py:63
    xt, yt = synthetic_digits(args.train_samples, 1)
    xv, yv = synthetic_digits(args.val_samples, 2)
    loaders = [('train', DeviceLoader(xt, yt, args.batch_size, device=device(), shuffle=True, seed=0)),        # main.py:64-67
               ('validation', DeviceLoader(xv, yv, args.batch_size, device=device()))]

    class Eager:                                                     # what the reference's build_classifier step returns
        pass
    eager = Eager()
    eager.nn, eager.criterion, eager.optimizer, eager.epoch = nn, criterion, optimizer, 0
    classifier = lower_classifier(eager, device=device(), precision=args.precision, max_batch=args.batch_size)   # main.py:68
    logger.info(f'aggregate id {classifier.id}')

    for _ in range(args.epochs):                                     # main.py:70-71
        training.service.handle('iterate', classifier, loaders, metrics)
    return classifier, history


if __name__ == '__main__':
    main()
