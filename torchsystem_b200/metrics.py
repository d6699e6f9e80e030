"""B200Metrics: the `Metrics` protocol (/root/reference/examples/tinysys/tinysys/domain.py:31-40) with the state of
the reference implementation (/root/reference/examples/tinysys/tinysys/metrics.py:8-27: torcheval `Mean` of the
per-batch losses in fp64 + micro `MulticlassAccuracy`) kept on the GPU.

`B200Classifier.fit/evaluate` already leave a 16-byte record per batch (loss, correct count, rows) written by
the fused head kernel, so `update` never re-reads predictions or targets: it only notes which record to fold
in, and `compute` folds the noted records in order with one tiny kernel. Tensors that did not come from the
engine go through `ts_metrics_update_raw` (a GPU kernel, not a CPU path)."""
from __future__ import annotations

from logging import getLogger

import torch
from torch import Tensor

from . import _lib
from ._lib import lib, check
from .engine import RING, _raw_stream
from .parallel import DataParallel

logger = getLogger(__name__)


class B200Metrics:
    def __init__(self, device: str | torch.device = 'cuda', log_every: int = 100, process_group=None):
        dev = torch.device(device)
        if dev.type != 'cuda':
            raise RuntimeError('B200Metrics keeps its state on a CUDA device; there is no CPU path')
        if dev.index is None:
            dev = torch.device('cuda', torch.cuda.current_device())
        self.device = dev
        self.log_every = log_every
        self.state = torch.zeros(4, dtype=torch.int64, device=dev)    # ts_metrics_state (32 bytes)
        self._run = None        # (engine, first_slot, count) of records noted but not folded in yet
        self._engines = []      # engines whose records were seen (their status word is checked in compute)
        self._dp = DataParallel(process_group)

    def _stream(self):
        return _raw_stream(self.device.index)

    def _flush(self) -> None:
        if self._run is not None:
            engine, first, count = self._run
            check(lib.ts_metrics_accumulate(self.state.data_ptr(), engine.ring.data_ptr(), RING, first, count,
                                            self._stream()), 'ts_metrics_accumulate')
            self._run = None

    def update(self, batch: int, loss: Tensor, predictions: Tensor, targets: Tensor) -> None:
        tag = getattr(loss, '_ts_record', None)
        if tag is not None:
            engine, slot = tag
            if engine not in self._engines:
                self._engines.append(engine)
            run = self._run
            if run is not None and run[0] is engine and (run[1] + run[2]) % RING == slot and run[2] < RING // 2:
                self._run = (engine, run[1], run[2] + 1)
            else:
                self._flush()
                self._run = (engine, slot, 1)
        else:
            self._flush()
            check(lib.ts_metrics_update_raw(self.state.data_ptr(), loss.detach().to(self.device, torch.float32).data_ptr(),
                                            predictions.data_ptr(), targets.data_ptr(), targets.shape[0], self._stream()),
                  'ts_metrics_update_raw')
        if self.log_every and batch % self.log_every == 0:
            logger.info(f'--- Batch {batch}: loss={loss.item()}')          # metrics.py:16-17

    def counters(self) -> tuple[float, int, int, int]:
        """(loss_sum, n_batches, n_correct, n_total), summed over data-parallel ranks."""
        self._flush()
        out = _lib.MetricsState()
        check(lib.ts_metrics_read(self.state.data_ptr(), out, self._stream()), 'ts_metrics_read')
        for engine in self._engines:       # the read above synchronised: a target outside [0, classes) is known by now
            check(lib.ts_mlp_status(engine.handle), 'ts_mlp_status')
        return self._dp.reduce_metrics(out.loss_sum, out.n_batches, out.n_correct, out.n_total, self.device)

    def compute(self) -> dict[str, Tensor]:
        loss_sum, n_batches, n_correct, n_total = self.counters()
        loss = loss_sum / n_batches if n_batches else 0.0
        acc = n_correct / n_total if n_total else float('nan')
        return {'loss': torch.tensor(loss, dtype=torch.float64), 'accuracy': torch.tensor(acc, dtype=torch.float32)}

    def reset(self) -> None:
        self._run = None
        check(lib.ts_metrics_reset(self.state.data_ptr(), self._stream()), 'ts_metrics_reset')
