"""Loaders that satisfy the reference `Loader` protocol (/root/reference/examples/tinysys/tinysys/domain.py:26-29:
`__iter__` yielding `(inputs, targets)`), sized for a GPU that finishes a batch in microseconds.

The handlers call `inputs.to(device), targets.to(device)` on whatever is yielded
(/root/reference/examples/tinysys/tinysys/services/training.py:33,44); for tensors already on that device `.to`
returns the tensor itself, so both loaders make that line free.

DeviceLoader   the whole (rank-sharded) dataset resident in HBM; batches are views.
PinnedLoader   dataset in pinned host memory; batches are copied host->device on a side stream, two or more
               batches ahead, and handed over already resident (the end-to-end path of bench.py)."""
from __future__ import annotations

import torch
from torch import Tensor

from .parallel import shard_bounds


class DeviceLoader:
    def __init__(self, inputs: Tensor, targets: Tensor, batch_size: int, device: str | torch.device = 'cuda',
                 shuffle: bool = False, seed: int = 0, drop_last: bool = False, rank: int = 0, world: int = 1,
                 dtype: torch.dtype | None = None):
        lo, hi = shard_bounds(inputs.shape[0], rank, world, batch_size if world > 1 else 1) if world > 1 else (0, inputs.shape[0])
        self.inputs = inputs[lo:hi].to(device=device, dtype=dtype or inputs.dtype).contiguous()
        self.targets = targets[lo:hi].to(device=device, dtype=torch.int64).contiguous()
        self.batch_size, self.shuffle, self.drop_last = int(batch_size), shuffle, drop_last
        self._gen = torch.Generator(device='cpu').manual_seed(seed)
        self._views = None       # unshuffled: the batch views, made once (two slices cost the host ~4 us per batch)

    def __len__(self) -> int:
        n = self.inputs.shape[0]
        return n // self.batch_size if self.drop_last else (n + self.batch_size - 1) // self.batch_size

    def __iter__(self):
        n, b = self.inputs.shape[0], self.batch_size
        if self.shuffle:   # seeded permutation, gathered once per epoch on the device
            perm = torch.randperm(n, generator=self._gen).to(self.inputs.device)
            x, y = self.inputs[perm], self.targets[perm]
        else:
            x, y = self.inputs, self.targets
        stop = n - n % b if self.drop_last else n
        if self.shuffle:
            return ((x[i:i + b], y[i:i + b]) for i in range(0, stop, b))
        if self._views is None or self._views[0] is not x or self._views[1] is not y:
            self._views = (x, y, [(x[i:i + b], y[i:i + b]) for i in range(0, stop, b)])
        return iter(self._views[2])


class PinnedLoader:
    def __init__(self, inputs: Tensor, targets: Tensor, batch_size: int, device: str | torch.device = 'cuda',
                 prefetch: int = 3, rank: int = 0, world: int = 1, dtype: torch.dtype | None = None):
        lo, hi = shard_bounds(inputs.shape[0], rank, world, batch_size) if world > 1 else (0, inputs.shape[0])
        # dtype=torch.bfloat16: convert once on the host (what bf16 mode rounds the inputs to anyway): half the PCIe bytes
        self.inputs = inputs[lo:hi].to(dtype or inputs.dtype).contiguous().pin_memory()
        self.targets = targets[lo:hi].to(torch.int64).contiguous().pin_memory()
        self.batch_size, self.prefetch = int(batch_size), max(1, int(prefetch))
        self.device = torch.device(device)
        self.copy_stream = torch.cuda.Stream(device=self.device)
        self.h2d_bytes = 0

    def __len__(self) -> int:
        return (self.inputs.shape[0] + self.batch_size - 1) // self.batch_size

    def _issue(self, i: int):
        b = self.batch_size
        xs, ys = self.inputs[i:i + b], self.targets[i:i + b]
        with torch.cuda.stream(self.copy_stream):
            x = xs.to(self.device, non_blocking=True)
            y = ys.to(self.device, non_blocking=True)
            done = torch.cuda.Event()
            done.record(self.copy_stream)
        self.h2d_bytes += xs.numel() * xs.element_size() + ys.numel() * ys.element_size()
        return x, y, done

    def __iter__(self):
        n, b = self.inputs.shape[0], self.batch_size
        starts = list(range(0, n, b))
        inflight = [self._issue(s) for s in starts[:self.prefetch]]
        nxt = len(inflight)
        compute = torch.cuda.current_stream(self.device)
        while inflight:
            x, y, done = inflight.pop(0)
            compute.wait_event(done)
            # the copy stream allocated x, y: tell the allocator the compute stream uses them too
            x.record_stream(compute)
            y.record_stream(compute)
            if nxt < len(starts):
                inflight.append(self._issue(starts[nxt]))
                nxt += 1
            yield x, y
