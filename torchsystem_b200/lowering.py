"""The compiler step that swaps the eager aggregate for the B200 one.

In the reference pipeline (/root/reference/examples/tinysys/tinysys/services/compilation.py:23-64) step 3 is
`compile_classifier` (`torch.compile`); `lower_classifier` takes its place:

    @compiler.step
    def lower(classifier, device = Depends(device)):
        return lower_classifier(classifier, device=device)

It returns a plain truthy `Aggregate`, which also sidesteps the `bool(OptimizedModule)` failure the
reference's `Compiler.compile` hits on torch >= 2.6 without touching torchsystem."""
from __future__ import annotations

from .classifier import B200Classifier


def lower_classifier(classifier, device='cuda', precision: str = 'bf16', **kwargs) -> B200Classifier:
    lowered = B200Classifier(classifier.nn, classifier.criterion, classifier.optimizer, precision=precision,
                             device=device, **kwargs)
    epoch = getattr(classifier, 'epoch', 0)
    if epoch:
        object.__setattr__(lowered, 'epoch', epoch)   # carry the epoch over without firing onepoch()
    return lowered
