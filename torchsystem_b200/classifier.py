"""B200Classifier: the aggregate the `train` / `validate` / `iterate` handlers drive, with its per-batch
work done by hand-written sm_100a kernels instead of eager PyTorch.

It honours the reference's `Model` protocol (/root/reference/examples/tinysys/tinysys/domain.py:10-24) and
mirrors the reference aggregate (/root/reference/examples/tinysys/tinysys/classifier.py:9-39):

    id, epoch, phase, nn, criterion, optimizer          attributes the handlers and consumers read
    fit(inputs, targets)      -> (predictions, loss)    classifier.py:29-35
    evaluate(inputs, targets) -> (predictions, loss)    classifier.py:37-39

`nn` stays an ordinary torch Module whose Parameters alias the engine's fp32 master weights, so
`torch.save(model.nn)`, `load_state_dict` and `gethash(model.nn)` behave as before."""
from __future__ import annotations

import os

import torch
from torch import Tensor, nn

from .compat import Aggregate, gethash
from .engine import Engine
from .modules import inspect_mlp
from . import parallel


def _check_criterion(criterion: nn.Module) -> None:
    if not isinstance(criterion, nn.CrossEntropyLoss):
        raise TypeError(f'B200Classifier fuses CrossEntropyLoss; got {type(criterion).__name__}')
    if (criterion.weight is not None or criterion.reduction != 'mean' or criterion.label_smoothing != 0.0):
        raise TypeError('only CrossEntropyLoss() with default weight/reduction/label_smoothing is supported')


def _optimizer_config(optimizer: torch.optim.Optimizer):
    """(kind, hyper-parameters) of a torch.optim.SGD / Adam / AdamW with one param group, as the fused optimizer takes
    them (torch/optim/sgd.py _single_tensor_sgd, adam.py _single_tensor_adam)."""
    if len(optimizer.param_groups) != 1:
        raise TypeError('exactly one optimizer param group is supported')
    g = optimizer.param_groups[0]
    if g.get('maximize', False):
        raise NotImplementedError('maximize=True is not supported by the fused optimizer')
    if type(optimizer) in (torch.optim.Adam, torch.optim.AdamW):
        decoupled = bool(g.get('decoupled_weight_decay', type(optimizer) is torch.optim.AdamW))
        return 'adam', dict(lr=float(g['lr']), betas=tuple(float(b) for b in g['betas']), eps=float(g['eps']),
                            weight_decay=float(g.get('weight_decay', 0.0)), amsgrad=bool(g.get('amsgrad', False)),
                            decoupled_weight_decay=decoupled)
    if type(optimizer) is torch.optim.SGD:
        return 'sgd', dict(lr=float(g['lr']), weight_decay=float(g.get('weight_decay', 0.0)),
                           momentum=float(g.get('momentum', 0.0)), dampening=float(g.get('dampening', 0.0)),
                           nesterov=bool(g.get('nesterov', False)))
    raise TypeError(f'unsupported optimizer {type(optimizer).__name__}: the fused step implements SGD, Adam and AdamW')


class B200Classifier(Aggregate):
    """Drop-in for the reference `Classifier(nn, criterion, optimizer)`.

    precision   'bf16' (tcgen05 tensor cores, fp32 accumulate; parity 1e-2) or 'f32' (SIMT; parity 1e-5)
    dropout_seed  key of the in-kernel counter-based dropout masks (pass `masks=` to `fit` to supply them instead)
    process_group  torch.distributed group for data-parallel training (default: the world, if initialised)
    uint8_norm  (mean, std): `fit` / `evaluate` then also take uint8 image batches and apply the dataset's
                ToTensor() + Normalize((mean,), (std,)) (examples/tinysys/datasets/mnist.py:7) on the device, 1 byte per pixel
                over PCIe / from HBM instead of 4
    dp_mode     'nvls': like 'p2p' one kernel per step and no separate collective, but for large models: every rank reduces
                1/world of the gradient inside the NVSwitch (multimem.ld_reduce), updates that share of the parameters and
                broadcasts them (multimem.st); needs symmetric memory with a multicast mapping.
                'p2p': every rank's optimizer kernel reads all ranks' gradients over NVLink peer mappings (no separate
                collective; right for gradients of about a megabyte); 'nccl': ts_fit_backward -> NCCL all-reduce of the flat gradient
                -> ts_opt_apply -- right for large models. Opt-in variants (both measured slower on 8 B200s, see
                _fit_bucketed): TSB200_DP_BUCKETS=1 pipelines bucketed all-reduces with per-bucket optimizer updates
                (ts_opt_apply_range); TSB200_DP_OVERLAP=1 starts each layer's all-reduce during the backward pass
                (ts_fit_backward_overlap)
    """

    def __init__(self, nn: nn.Module, criterion: nn.Module, optimizer: torch.optim.Optimizer, *, precision: str = 'bf16',
                 device: str | torch.device = 'cuda', max_batch: int = 4096, dropout_seed: int = 42, process_group=None,
                 dp_mode: str = 'p2p', uint8_norm: tuple[float, float] | None = None):
        super().__init__()
        if dp_mode not in ('p2p', 'nccl', 'nvls'):
            raise ValueError(f"dp_mode must be 'p2p', 'nccl' or 'nvls', got {dp_mode!r}")
        _check_criterion(criterion)
        spec = inspect_mlp(nn)
        kind, hyper = _optimizer_config(optimizer)
        dev = torch.device(device)
        if dev.type == 'cuda' and dev.index is None:
            dev = torch.device('cuda', torch.cuda.current_device())
        self.nn = nn
        self.epoch = 0
        self.criterion = criterion
        self.optimizer = optimizer
        self.precision = precision
        self.dropout_seed = int(dropout_seed)
        self._params = [p for lin in spec.linears for p in (lin.weight, lin.bias)]
        owned = {id(p) for p in self._params}
        extra = [p for p in nn.parameters() if id(p) not in owned]
        if extra:
            raise TypeError('the network has parameters outside its Linear layers; unsupported')
        nn.to(dev)
        self.engine = Engine(spec, precision, dev, max_batch=max_batch)
        self.engine.adopt(self._params)
        if uint8_norm is not None:        # accept raw uint8 images: ToTensor + Normalize(mean, std) happen on the device
            self.engine.set_input_normalization(float(uint8_norm[0]), float(uint8_norm[1]))
        step = self._adopt_optimizer_state(kind, hyper)
        self.engine.bind_optimizer(kind, step=step, **hyper)
        self._hyper_key = self._param_group_key()
        # optimizer.state_dict() must show the live step count (the moments alias the engine's buffers already)
        optimizer.register_state_dict_pre_hook(self._before_state_dict)
        self._state_gathered_at = -1
        self._last_phase = None
        self._dp = parallel.DataParallel(process_group)
        self._p2p = None
        self._overlap = False
        self._nvls = dp_mode == 'nvls' and self._dp.world > 1
        if self._dp.world > 1:
            self._dp.broadcast_(self.engine.params)          # same initial weights on every rank
            self.engine.sync_weights()
            if dp_mode == 'p2p':
                self._p2p = parallel.bind_peer_gradients(self.engine, self._params, self._dp)
            elif dp_mode == 'nvls':
                self._p2p = parallel.bind_nvls(self.engine, self._params, self._dp)
            elif dp_mode == 'nccl':
                # buckets of whole layers, at least 16 M elements each (the last one absorbs what is left)
                self._comm = torch.cuda.Stream(device=dev)
                self._buckets = parallel.layer_buckets(self.engine.offsets, self.engine.total)
                self._bucket_events = [torch.cuda.Event() for _ in self._buckets]
                self._bucketed = os.environ.get('TSB200_DP_BUCKETS', '0') != '0'
            if dp_mode == 'nccl' and os.environ.get('TSB200_DP_OVERLAP', '0') != '0':
                # opt-in: all-reduce per layer on a side stream, under the rest of the backward pass; the backward GEMMs
                # leave `reserve` SMs to the collective's kernels (they cannot share an SM with a 225 KB persistent GEMM
                # CTA). Measured slower than the bucketed scheme on 8 B200s (NCCL needs more SMs than it pays to give up).
                self._overlap = True
                self._layer_events = [torch.cuda.Event() for _ in spec.linears]
                for e in self._layer_events:
                    e.record()                               # creates the underlying cudaEvent_t
                self.engine.reserve_sms(int(os.environ.get('TSB200_DP_RESERVE_SMS', '16')))

    # ------------------------------------------------------------------ identity / hooks
    @property
    def id(self) -> str:
        return gethash(self.nn)          # classifier.py:18-20

    def onphase(self) -> None:
        # weights may have been replaced between phases (repository.restore -> load_state_dict)
        self.engine.sync_weights()
        # data parallel: line the ranks up on the host whenever the phase CHANGES (train -> evaluation -> train ...). Rank 0 may
        # come out of an I/O consumer (storage.py:84-86 saves on every Iterated) seconds after the others; they then wait
        # here, in the process group's barrier, not inside a spinning optimizer kernel. Setting the same phase again (a handler
        # called twice in a row) costs nothing.
        if self.phase != self._last_phase:
            self._last_phase = self.phase
            self._dp.barrier()

    def sync(self) -> None:
        """Call after editing `nn`'s weights by hand; re-derives the bf16 operand copies."""
        self.engine.sync_weights()

    def _adopt_optimizer_state(self, kind: str, hyper: dict) -> int:
        """Resume: take over the state the torch optimizer already holds (Adam: exp_avg / exp_avg_sq / max_exp_avg_sq /
        step; SGD: momentum_buffer) and make the optimizer's state dict alias the engine's flat buffers from now on."""
        eng = self.engine
        step = 0
        if kind == 'adam':
            names = ['exp_avg', 'exp_avg_sq'] + (['max_exp_avg_sq'] if hyper['amsgrad'] else [])
        elif hyper['momentum'] != 0:
            names = ['momentum_buffer']
        else:
            return step
        flat = {n: torch.zeros_like(eng.params) for n in names}
        for i, p in enumerate(self._params):
            o, n = eng.offsets[i], p.numel()
            st = self.optimizer.state.get(p, {})
            have = all(st.get(k) is not None for k in names)
            if have:
                for k in names:
                    flat[k][o:o + n].copy_(st[k].reshape(-1))
                if kind == 'adam':
                    step = int(st['step']) if not isinstance(st['step'], Tensor) else int(st['step'].item())
                else:
                    step = max(step, 1)          # a momentum buffer exists: not the first step any more
            new = {k: flat[k][o:o + n].view(p.shape) for k in names}
            if kind == 'adam':
                new['step'] = torch.tensor(float(step))
            self.optimizer.state[p] = new
        eng.exp_avg = flat.get('exp_avg', flat.get('momentum_buffer'))
        eng.exp_avg_sq = flat.get('exp_avg_sq')
        eng.max_exp_avg_sq = flat.get('max_exp_avg_sq')
        return step

    def sync_optimizer_state(self) -> None:
        """Refresh the `step` entries torch's optimizer.state_dict() reports (the moments alias live buffers). Runs
        automatically before every optimizer.state_dict()."""
        step = self.engine.step_count
        for p in self._params:
            st = self.optimizer.state.get(p)
            if st is not None and 'step' in st:
                st['step'] = torch.tensor(float(step))

    def gather_optimizer_state(self) -> None:
        """dp_mode='nvls' only, COLLECTIVE (every rank must call it): each rank updates the optimizer moments of its own
        1/world of the 1024-element chunks, so before a checkpoint the owned chunks are summed into every rank's buffers."""
        eng = self.engine
        if self._dp.world > 1 and self._nvls:
            own = (torch.arange(eng.total // 1024, device=eng.device) % self._dp.world == self._dp.rank)
            own = own.repeat_interleave(1024).to(torch.float32)
            for buf in (eng.exp_avg, eng.exp_avg_sq, eng.max_exp_avg_sq):
                if buf is not None:
                    buf.mul_(own)
                    self._dp.all_reduce_sum_(buf)
        self._state_gathered_at = eng.step_count

    def _before_state_dict(self, optimizer) -> None:
        if self._dp.world > 1 and self._nvls and self.engine.opt_kind is not None and self.engine.exp_avg is not None \
                and self._state_gathered_at != self.engine.step_count:
            raise RuntimeError("dp_mode='nvls' shards the optimizer moments over the ranks: call model.gather_optimizer_state() "
                               'on EVERY rank (a collective) after the last step, then optimizer.state_dict()')
        self.sync_optimizer_state()

    def _param_group_key(self) -> tuple:
        g = self.optimizer.param_groups[0]
        return (g['lr'], g.get('weight_decay'), g.get('momentum'), g.get('dampening'), g.get('nesterov'), g.get('betas'),
                g.get('eps'), g.get('amsgrad'), g.get('maximize'), g.get('decoupled_weight_decay'))

    def _refresh_hyper(self) -> None:
        kind, hyper = _optimizer_config(self.optimizer)      # follows lr schedulers / manual edits of the param group
        self.engine.bind_optimizer(kind, step=self.engine.step_count, **hyper)
        self._hyper_key = self._param_group_key()

    # ------------------------------------------------------------------ the hot path
    def forward(self, input: Tensor) -> Tensor:
        raise RuntimeError('B200Classifier fuses forward, loss and argmax; call fit() or evaluate()')

    def fit(self, inputs: Tensor, targets: Tensor, *, masks=None) -> tuple[Tensor, Tensor]:
        """One optimisation step on a batch; returns (argmax of the pre-update logits, mean CE loss).
        `masks`: optional list of [B, hidden] 0/1 keep-masks, one per hidden layer (parity runs)."""
        if self._param_group_key() != self._hyper_key:       # lr schedulers, manual edits of any hyper-parameter
            self._refresh_hyper()
        if self._dp.world == 1:
            return self.engine.fit(inputs, targets, masks, self.dropout_seed)
        if self._p2p is not None:      # gradient all-reduce folded into the optimizer kernel over NVLink peer mappings
            return self.engine.fit(inputs, targets, masks, self.dropout_seed + self._dp.rank)
        if self._overlap:
            return self._fit_overlapped(inputs, targets, masks)
        return self._fit_bucketed(inputs, targets, masks)

    def _fit_bucketed(self, inputs: Tensor, targets: Tensor, masks) -> tuple[Tensor, Tensor]:
        """'nccl' mode: backward pass, one all-reduce of the flat gradient, optimizer. With TSB200_DP_BUCKETS=1 the gradient
        is instead all-reduced bucket by bucket (about 64 MB each, whole layers) on a side stream while the optimizer
        already updates the buckets that have arrived (ts_opt_apply_range). Measured on 8 B200s, L model, 8,192 rows per
        GPU: 3.10 ms per step plain, 3.40 ms bucketed, 3.31 ms with the all-reduce started during the backward pass
        (TSB200_DP_OVERLAP=1): NCCL wants the whole GPU, so the plain sequence is the default."""
        eng, dp = self.engine, self._dp
        out = eng.fit(inputs, targets, masks, self.dropout_seed + dp.rank, update=False)
        if len(self._buckets) == 1 or not self._bucketed:
            dp.all_reduce_sum_(eng.grads)
            eng.apply(1.0 / dp.world)
            return out
        main = torch.cuda.current_stream(eng.index)
        self._comm.wait_stream(main)
        for i, (t0, nt, o, n) in enumerate(self._buckets):
            with torch.cuda.stream(self._comm):
                dp.all_reduce_sum_(eng.grads[o:o + n])
                self._bucket_events[i].record(self._comm)
            main.wait_event(self._bucket_events[i])
            eng.apply_range(1.0 / dp.world, t0, nt, advance=i == 0)
        return out

    def _fit_overlapped(self, inputs: Tensor, targets: Tensor, masks) -> tuple[Tensor, Tensor]:
        """'nccl' mode with the all-reduce of each layer's gradient issued on a side stream as soon as that layer's
        backward kernels are enqueued (last layer first), so it runs under the remaining dX / dW GEMMs; the small
        tensors that are still partial sums at that point (biases) go in one packed all-reduce at the end."""
        eng, dp = self.engine, self._dp
        main = torch.cuda.current_stream(eng.index)
        pred, loss, final = eng.fit_backward_overlap(inputs, targets, masks, self.dropout_seed + dp.rank, self._layer_events)
        n_layers = len(self._layer_events)
        late = []
        with torch.cuda.stream(self._comm):
            for l in range(n_layers - 1, -1, -1):
                self._comm.wait_event(self._layer_events[l])
                for t in (2 * l, 2 * l + 1):
                    o, n = eng.offsets[t], self._params[t].numel()
                    if final[t]:
                        dp.all_reduce_sum_(eng.grads[o:o + n])
                    else:
                        late.append((o, n))
            self._comm.wait_stream(main)                 # the fold of the remaining partials is the last kernel of the call
            if late:
                packed = torch.cat([eng.grads[o:o + n] for o, n in late])
                dp.all_reduce_sum_(packed)
                for (o, n), piece in zip(late, packed.split([n for _, n in late])):
                    eng.grads[o:o + n].copy_(piece)
        main.wait_stream(self._comm)
        eng.apply(1.0 / dp.world)
        return pred, loss

    def evaluate(self, inputs: Tensor, targets: Tensor) -> tuple[Tensor, Tensor]:
        return self.engine.evaluate(inputs, targets)
