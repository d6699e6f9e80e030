"""The `iterate` / `train` / `validate` handlers and their events, for running where the reference is not
installed (the GPU box). Same names, arguments and event flow as
/root/reference/examples/tinysys/tinysys/services/training.py:15-63 -- which itself runs unmodified against this
package's protocol classes (tests/test_host_logic.py::test_handlers_trace_equals_unmodified_reference_handlers)."""
from __future__ import annotations

from typing import Any, Sequence

from torch import Tensor, inference_mode

from .compat import Depends, Provider, Service, Producer, event

provider = Provider()
service = Service(provider=provider)
producer = Producer()


def device() -> str:
    """Dependency stub: override with `provider.override(device, lambda: 'cuda')`."""
    ...


@event
class Trained:
    model: Any
    results: dict[str, Tensor]


@event
class Validated:
    model: Any
    results: dict[str, Tensor]


@event
class Iterated:
    model: Any
    loaders: Sequence[tuple[str, Any]]


def _run_phase(model, loader, metrics, device, step):
    for batch, (inputs, targets) in enumerate(loader, start=1):
        inputs, targets = inputs.to(device), targets.to(device)
        predictions, loss = step(inputs, targets)
        metrics.update(batch, loss, predictions, targets)
    return metrics.compute()


@service.handler
def train(model, loader, metrics, device: str = Depends(device)):
    model.phase = 'train'
    producer.dispatch(Trained(model, _run_phase(model, loader, metrics, device, model.fit)))


@service.handler
def validate(model, loader, metrics, device: str = Depends(device)):
    with inference_mode():
        model.phase = 'evaluation'
        results = _run_phase(model, loader, metrics, device, model.evaluate)
        producer.dispatch(Validated(model, results))


@service.handler
def iterate(model, loaders: Sequence[tuple[str, Any]], metrics):
    for phase, loader in loaders:
        (train if phase == 'train' else validate)(model, loader, metrics)
        metrics.reset()
    model.epoch += 1
    producer.dispatch(Iterated(model, loaders))
