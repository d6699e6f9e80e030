"""Network definitions accepted by the lowering step.

`MLP` keeps the constructor of the reference's example network
(/root/reference/examples/tinysys/modules/mlp.py:6-11: Linear -> Dropout -> activation -> Linear, dropout BEFORE
the activation) so that `gethash(MLP(784, 256, 10, dropout=0.1, activation=ReLU()))` reproduces the identity the
reference's notebook shows (examples/tinysys/notebook.ipynb:34); `hidden_size` may also be a list for deeper
stacks (north-star config 4). It is an ordinary picklable torch Module: `torch.save(model.nn)` and
`load_state_dict` keep working (examples/tinysys/tinysys/repository.py:13-17)."""
from __future__ import annotations

from dataclasses import dataclass
from typing import Sequence

import torch
from torch import Tensor, nn


class MLP(nn.Module):
    def __init__(self, input_size: int, hidden_size, output_size: int, dropout: float, activation: nn.Module):
        super().__init__()
        widths = [hidden_size] if isinstance(hidden_size, int) else list(hidden_size)
        if len(widths) == 1:     # attribute names of the reference network, so state_dicts are interchangeable
            self.input_layer = nn.Linear(input_size, widths[0], bias=True)
            self.dropout = nn.Dropout(dropout)
            self.activation = activation
            self.output_layer = nn.Linear(widths[0], output_size)
        else:
            self.hidden = nn.ModuleList(nn.Linear(i, o) for i, o in zip([input_size] + widths[:-1], widths))
            self.dropout = nn.Dropout(dropout)
            self.activation = activation
            self.output_layer = nn.Linear(widths[-1], output_size)

    def linears(self) -> list[nn.Linear]:
        if hasattr(self, 'input_layer'):
            return [self.input_layer, self.output_layer]
        return list(self.hidden) + [self.output_layer]

    def forward(self, features: Tensor) -> Tensor:
        for layer in self.linears()[:-1]:
            features = self.activation(self.dropout(layer(features)))
        return self.output_layer(features)


@dataclass
class MLPSpec:
    """What the engine needs to know about a network: widths, dropout p, activation, and its Linear modules."""
    dims: list[int]
    dropout: float
    activation: str
    linears: list[nn.Linear]


def inspect_mlp(net: nn.Module) -> MLPSpec:
    """Recognise `(Linear [Dropout] activation)* Linear` in a module (the reference MLP, this package's MLP,
    or an nn.Sequential of that shape). Anything else raises TypeError: there is no generic fallback."""
    if hasattr(net, 'linears') and callable(net.linears):
        linears = list(net.linears())
        drops = [m for m in net.modules() if isinstance(m, nn.Dropout)]
        acts = [m for m in net.modules() if isinstance(m, (nn.ReLU, nn.GELU))]
    else:
        seq = [m for m in net.children()]
        linears = [m for m in seq if isinstance(m, nn.Linear)]
        drops = [m for m in seq if isinstance(m, nn.Dropout)]
        acts = [m for m in seq if isinstance(m, (nn.ReLU, nn.GELU))]
        other = [m for m in seq if not isinstance(m, (nn.Linear, nn.Dropout, nn.ReLU, nn.GELU))]
        if other:
            raise TypeError(f'unsupported layer(s) for the B200 engine: {[type(m).__name__ for m in other]}')
        # order check: ... Linear, [Dropout], act, Linear ...; the last child must be a Linear
        if not seq or not isinstance(seq[-1], nn.Linear):
            raise TypeError('the network must end with a Linear layer')
        kinds = ''.join('L' if isinstance(m, nn.Linear) else 'D' if isinstance(m, nn.Dropout) else 'A' for m in seq)
        import re
        if not re.fullmatch(r'(LD?A)+L', kinds):
            raise TypeError(f'layer order {kinds!r} is not (Linear [Dropout] activation)* Linear with dropout before activation')
    if len(linears) < 2:
        raise TypeError('the B200 engine needs at least one hidden Linear layer and an output Linear layer')
    for a, b in zip(linears[:-1], linears[1:]):
        if a.out_features != b.in_features:
            raise TypeError('Linear layers do not chain')
    if any(l.bias is None for l in linears):
        raise TypeError('Linear layers without bias are not supported')
    ps = {float(d.p) for d in drops}
    if len(ps) > 1:
        raise TypeError('all Dropout layers must share one probability')
    kinds = {type(a) for a in acts}
    if len(kinds) != 1:
        raise TypeError('exactly one activation type (ReLU or GELU) is supported')
    act = acts[0]
    if isinstance(act, nn.GELU) and getattr(act, 'approximate', 'none') != 'none':
        raise TypeError('only exact (erf) GELU is supported')
    dims = [linears[0].in_features] + [l.out_features for l in linears]
    return MLPSpec(dims=dims, dropout=ps.pop() if ps else 0.0,
                   activation='relu' if isinstance(act, nn.ReLU) else 'gelu', linears=linears)
