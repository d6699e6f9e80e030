"""Host-side owner of one `ts_mlp` handle: flat parameter / gradient / Adam buffers (torch tensors that
the network's Parameters alias), the workspace, the per-batch record ring, and the calls into the C ABI.

PyTorch is used here for device memory and streams only; all arithmetic runs in libtsb200.so."""
from __future__ import annotations

import ctypes as C

import torch
from torch import Tensor, nn

from . import _lib
from ._lib import lib, check
from .modules import MLPSpec

RING = 1024  # per-batch records kept alive; a returned loss tensor is overwritten RING batches later

# handle of the current stream of a device: torch's raw accessor when this build has it (what its own inductor uses),
# else through torch.cuda.Stream. Per-batch host time is what bounds the handlers once the kernels take ~30 us.
_get_raw = getattr(torch._C, '_cuda_getCurrentRawStream', None)


def _raw_stream(index: int) -> int:
    if _get_raw is not None:
        return _get_raw(index)
    return torch.cuda.current_stream(index).cuda_stream


class Engine:
    def __init__(self, spec: MLPSpec, precision: str, device: torch.device, max_batch: int = 4096):
        if device.type != 'cuda':
            raise RuntimeError('torchsystem_b200 runs on CUDA devices (sm_100a) only; there is no CPU path')
        self.spec, self.device = spec, device
        self.precision = {'f32': _lib.TS_PREC_F32, 'fp32': _lib.TS_PREC_F32, 'bf16': _lib.TS_PREC_BF16}[precision]
        self.index = device.index if device.index is not None else torch.cuda.current_device()
        dims = (C.c_int32 * len(spec.dims))(*spec.dims)
        act = _lib.TS_ACT_RELU if spec.activation == 'relu' else _lib.TS_ACT_GELU
        handle = C.c_void_p()
        check(lib.ts_mlp_create(C.byref(handle), dims, len(spec.dims), act, float(spec.dropout), self.precision,
                                self.index), 'ts_mlp_create')
        self.handle = handle
        n = 2 * (len(spec.dims) - 1)
        offs, total = (C.c_int64 * n)(), C.c_int64()
        check(lib.ts_mlp_param_layout(self.handle, offs, C.byref(total)), 'ts_mlp_param_layout')
        self.offsets, self.total = list(offs), total.value
        with torch.cuda.device(self.index):
            self.params = torch.zeros(self.total, dtype=torch.float32, device=device)
            self.grads = torch.zeros(self.total, dtype=torch.float32, device=device)
            self.exp_avg = self.exp_avg_sq = self.max_exp_avg_sq = None
            self.ring = torch.zeros(RING * 4, dtype=torch.float32, device=device)
        check(lib.ts_mlp_bind_params(self.handle, self.params.data_ptr(), self.grads.data_ptr()), 'ts_mlp_bind_params')
        # pre-built 0-d views of every record's loss, tagged so B200Metrics can find the record
        self.loss_views: list[Tensor] = []
        for s in range(RING):
            v = self.ring[4 * s]
            v._ts_record = (self, s)
            self.loss_views.append(v)
        self._ring_ptr = self.ring.data_ptr()
        self._slot = 0
        self.max_batch = 0
        self.workspace = None
        self.reserve(max_batch)
        self._hyper = None
        self.opt_kind = None

    def __del__(self):
        h, self.handle = getattr(self, 'handle', None), None
        if h:
            lib.ts_mlp_destroy(h)

    # ------------------------------------------------------------------ memory
    def reserve(self, max_batch: int) -> None:
        if max_batch <= self.max_batch:
            return
        need = lib.ts_mlp_workspace_bytes(self.handle, int(max_batch))
        if need < 0:
            check(int(need), 'ts_mlp_workspace_bytes')
        with torch.cuda.device(self.index):
            torch.cuda.current_stream().synchronize()    # old workspace may still be in use
            self.workspace = torch.empty(need, dtype=torch.uint8, device=self.device)
        check(lib.ts_mlp_set_workspace(self.handle, self.workspace.data_ptr(), need, int(max_batch)), 'ts_mlp_set_workspace')
        self.max_batch = int(max_batch)
        self.sync_weights()

    def set_input_normalization(self, mean: float, std: float) -> None:
        """uint8 batches are mapped through ToTensor() + Normalize((mean,), (std,)) on the device
        (/root/reference/examples/tinysys/datasets/mnist.py:7): the 256 possible results are computed here with torch's
        own fp32 arithmetic, so the device path reproduces the host transform bit for bit."""
        lut = torch.arange(256, dtype=torch.uint8).to(torch.float32).div(255).sub_(mean).div_(std)
        arr = (C.c_float * 256)(*lut.tolist())
        check(lib.ts_mlp_set_input_lut(self.handle, arr), 'ts_mlp_set_input_lut')

    def view(self, index: int, shape) -> Tensor:
        o = self.offsets[index]
        n = 1
        for s in shape:
            n *= s
        return self.params[o:o + n].view(shape)

    def adopt(self, parameters: list[nn.Parameter]) -> None:
        """Move the network's parameters into the flat buffer and re-point them at it, so `model.nn`
        keeps showing the live fp32 master weights (storage.py:84-86, repository.py:13-17 rely on that)."""
        with torch.no_grad():
            for i, p in enumerate(parameters):
                o, n = self.offsets[i], p.numel()
                self.params[o:o + n].copy_(p.detach().reshape(-1).to(self.device, torch.float32))
                p.data = self.params[o:o + n].view(p.shape)
                p.grad = self.grads[o:o + n].view(p.shape)
        self.sync_weights()

    def bind_optimizer(self, kind: str, lr: float, betas=(0.9, 0.999), eps: float = 1e-8, step: int = 0, *,
                       weight_decay: float = 0.0, momentum: float = 0.0, dampening: float = 0.0, nesterov: bool = False,
                       amsgrad: bool = False, decoupled_weight_decay: bool = False) -> None:
        """torch.optim.SGD / Adam / AdamW hyper-parameters (one param group). State lives in flat buffers laid out
        like the parameters: exp_avg (Adam's first moment or SGD's momentum buffer), exp_avg_sq, max_exp_avg_sq."""
        self.opt_kind = kind
        adam = kind == 'adam'
        if (adam or momentum != 0) and self.exp_avg is None:
            self.exp_avg = torch.zeros_like(self.params)
        if adam and self.exp_avg_sq is None:
            self.exp_avg_sq = torch.zeros_like(self.params)
        if adam and amsgrad and self.max_exp_avg_sq is None:
            self.max_exp_avg_sq = torch.zeros_like(self.params)
        h = _lib.OptHyper(float(lr), float(betas[0]), float(betas[1]), float(eps), float(weight_decay), float(momentum),
                          float(dampening), int(bool(nesterov)), int(bool(amsgrad)), int(bool(decoupled_weight_decay)), 0)
        ptr = lambda t: t.data_ptr() if t is not None else None
        check(lib.ts_opt_bind(self.handle, _lib.TS_OPT_ADAM if adam else _lib.TS_OPT_SGD, C.byref(h), ptr(self.exp_avg),
                              ptr(self.exp_avg_sq) if adam else None, ptr(self.max_exp_avg_sq) if adam and amsgrad else None,
                              int(step)), 'ts_opt_bind')
        self._hyper = (kind, float(lr), tuple(betas), float(eps), float(weight_decay), float(momentum), float(dampening),
                       bool(nesterov), bool(amsgrad), bool(decoupled_weight_decay))

    @property
    def step_count(self) -> int:
        return int(lib.ts_opt_step_count(self.handle))

    def sync_weights(self) -> None:
        if self.workspace is not None:
            check(lib.ts_mlp_sync_weights(self.handle, self._stream()), 'ts_mlp_sync_weights')

    # ------------------------------------------------------------------ hot path
    def _stream(self):
        return _raw_stream(self.index)       # the current stream's handle, without building a torch.cuda.Stream per call

    def _next_slot(self) -> int:
        s = self._slot
        self._slot = (s + 1) % RING
        return s

    @staticmethod
    def _dtype_code(x: Tensor) -> int:
        if x.dtype == torch.float32:
            return _lib.TS_DT_F32
        if x.dtype == torch.bfloat16:
            return _lib.TS_DT_BF16
        if x.dtype == torch.uint8:
            return _lib.TS_DT_U8          # needs set_input_normalization (the library checks)
        raise TypeError(f'inputs must be float32, bfloat16 or uint8, got {x.dtype}')

    def _check_batch(self, inputs: Tensor, targets: Tensor) -> int:
        B = inputs.shape[0]
        if inputs.device != self.device or targets.device != self.device:
            raise RuntimeError(f'inputs/targets must live on {self.device} (got {inputs.device}, {targets.device})')
        if inputs.numel() != B * self.spec.dims[0]:
            raise ValueError(f'inputs of shape {tuple(inputs.shape)} do not flatten to [B, {self.spec.dims[0]}]')
        if targets.dtype != torch.int64 or targets.shape != (B,):
            raise TypeError('targets must be an int64 vector of length B')
        if not inputs.is_contiguous() or not targets.is_contiguous():
            raise ValueError('inputs and targets must be contiguous')
        if B > self.max_batch:
            self.reserve(B)
        return B

    def _masks(self, masks, B: int):
        if masks is None:
            return None, None
        hidden = self.spec.dims[1:-1]
        if len(masks) != len(hidden):
            raise ValueError(f'need {len(hidden)} dropout masks')
        keep = []
        for m, h in zip(masks, hidden):
            m = m.to(self.device, torch.uint8).contiguous()
            if m.shape != (B, h):
                raise ValueError(f'mask shape {tuple(m.shape)} != {(B, h)}')
            keep.append(m)
        return (C.c_void_p * len(keep))(*[m.data_ptr() for m in keep]), keep

    def fit(self, inputs: Tensor, targets: Tensor, masks=None, seed: int = 0, update: bool = True):
        B = self._check_batch(inputs, targets)
        pred = torch.empty(B, dtype=torch.int64, device=self.device)
        slot = self._next_slot()
        mptr, keep = self._masks(masks, B)
        fn = lib.ts_fit_step if update else lib.ts_fit_backward
        rc = fn(self.handle, inputs.data_ptr(), self._dtype_code(inputs), targets.data_ptr(), B, mptr, seed,
                pred.data_ptr(), self._ring_ptr + _lib.RECORD_BYTES * slot, self._stream())
        if rc:
            check(rc, 'ts_fit_step')
        return pred, self.loss_views[slot]

    def fit_backward_overlap(self, inputs: Tensor, targets: Tensor, masks, seed: int, events):
        """ts_fit_backward_overlap: like fit(update=False), plus one CUDA event per layer recorded as soon as that layer's
        gradients are enqueued. Returns (pred, loss, final) with final[t] = gradient tensor t is complete at its event."""
        B = self._check_batch(inputs, targets)
        pred = torch.empty(B, dtype=torch.int64, device=self.device)
        slot = self._next_slot()
        mptr, keep = self._masks(masks, B)
        n = len(events)
        evs = (C.c_void_p * n)(*[e.cuda_event for e in events])
        final = (C.c_int32 * (2 * n))()
        rc = lib.ts_fit_backward_overlap(self.handle, inputs.data_ptr(), self._dtype_code(inputs), targets.data_ptr(), B, mptr, seed,
                                         pred.data_ptr(), self._ring_ptr + _lib.RECORD_BYTES * slot, self._stream(), evs, final)
        if rc:
            check(rc, 'ts_fit_backward_overlap')
        return pred, self.loss_views[slot], list(final)

    def reserve_sms(self, n: int) -> None:
        check(lib.ts_dp_reserve_sms(self.handle, int(n)), 'ts_dp_reserve_sms')

    def apply(self, grad_scale: float = 1.0) -> None:
        check(lib.ts_opt_apply(self.handle, float(grad_scale), self._stream()), 'ts_opt_apply')

    def apply_range(self, grad_scale: float, first_tensor: int, n_tensors: int, advance: bool) -> None:
        check(lib.ts_opt_apply_range(self.handle, float(grad_scale), int(first_tensor), int(n_tensors), int(bool(advance)),
                                     self._stream()), 'ts_opt_apply_range')

    def evaluate(self, inputs: Tensor, targets: Tensor):
        B = self._check_batch(inputs, targets)
        pred = torch.empty(B, dtype=torch.int64, device=self.device)
        slot = self._next_slot()
        rc = lib.ts_eval_step(self.handle, inputs.data_ptr(), self._dtype_code(inputs), targets.data_ptr(), B,
                              pred.data_ptr(), self._ring_ptr + _lib.RECORD_BYTES * slot, self._stream())
        if rc:
            check(rc, 'ts_eval_step')
        return pred, self.loss_views[slot]
