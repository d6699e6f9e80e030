"""Data-parallel plumbing: one process per GPU, `torch.distributed` (NCCL over NVLink 5 / NVSwitch on the
box, gloo in the CPU tests). The reference is single-process; this is the north star's
addition: every rank runs the same handlers on its shard of the loader, gradients are averaged once per
step, metric counters are summed once per `compute()`.

Everything here is device-agnostic on purpose, so the world_size-2 gloo tests exercise the same code."""
from __future__ import annotations

import os

import torch
import torch.distributed as dist
from torch import Tensor

LL_INBOX_BYTES = 28311552              # TS_DP_LL_INBOX_BYTES (include/tsb200.h)
LL_MAX_MODEL_BYTES = 16 << 20          # flat gradient size up to which the low-latency inbox is allocated (one-hidden-layer models)


def shard_bounds(n: int, rank: int, world: int, granule: int = 1) -> tuple[int, int]:
    """Contiguous, equal shards of `n` samples (the remainder is dropped so every rank sees the same number of
    equally sized batches: mean of local means == global mean, as CrossEntropyLoss(reduction='mean') needs)."""
    per = (n // world) // granule * granule
    return rank * per, rank * per + per


class DataParallel:
    def __init__(self, group=None):
        self.group = group
        if dist.is_available() and dist.is_initialized():
            self.world = dist.get_world_size(group)
            self.rank = dist.get_rank(group)
        else:
            self.world, self.rank = 1, 0

    def broadcast_(self, flat: Tensor, src: int = 0) -> None:
        if self.world > 1:
            dist.broadcast(flat, src=src, group=self.group)

    def barrier(self) -> None:
        if self.world > 1:
            dist.barrier(group=self.group)

    def all_reduce_sum_(self, flat: Tensor) -> None:
        if self.world > 1:
            dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group)

    def reduce_metrics(self, loss_sum: float, n_batches: int, n_correct: int, n_total: int, device) -> tuple:
        """Sum the four metric accumulators over ranks (fp64 carries the integer counts exactly below 2**53)."""
        if self.world == 1:
            return loss_sum, n_batches, n_correct, n_total
        t = torch.tensor([loss_sum, float(n_batches), float(n_correct), float(n_total)], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        v = t.tolist()
        return v[0], int(v[1]), int(v[2]), int(v[3])


def layer_buckets(offsets: list[int], total: int, min_elems: int = 16 << 20, small_tail: int = 4 << 20):
    """All-reduce buckets over the flat gradient buffer: whole layers (weight + bias, torch order W0, b0, W1, ...), merged
    until a bucket holds at least `min_elems` elements; a last bucket smaller than `small_tail` joins its neighbour.
    Returns [(first_tensor, n_tensors, first_element, n_elements)]; the buckets tile [0, total) exactly."""
    n_t = len(offsets)
    buckets, t0, o0 = [], 0, offsets[0]
    for l in range(n_t // 2):
        end_t = 2 * l + 2
        end_o = offsets[end_t] if end_t < n_t else total
        if end_o - o0 >= min_elems or end_t == n_t:
            buckets.append((t0, end_t - t0, o0, end_o - o0))
            t0, o0 = end_t, end_o
    if len(buckets) > 1 and buckets[-1][3] < small_tail:
        a, b = buckets[-2], buckets.pop()
        buckets[-1] = (a[0], a[1] + b[1], a[2], a[3] + b[3])
    return buckets


def bind_peer_gradients(engine, params, dp: DataParallel):
    """Put the flat gradient buffers (two, alternating with the step parity) and a small flag array into torch
    symmetric memory, exchange the peer mappings and hand them to the library (ts_dp_bind). Returns the objects
    that must stay alive, or raises if the GPUs cannot map each other's memory."""
    import ctypes as C
    import torch.distributed._symmetric_memory as symm
    from ._lib import lib, check
    group = dp.group if dp.group is not None else dist.group.WORLD
    dev, total = engine.device, engine.total
    with torch.cuda.device(engine.index):
        gbuf = symm.empty(2 * total, dtype=torch.float32, device=dev)
        flags = symm.empty(64, dtype=torch.int32, device=dev)      # TS_DP_FLAG_WORDS, see include/tsb200.h ts_dp_bind
        gbuf.zero_()
        flags.zero_()
        hg = symm.rendezvous(gbuf, group)
        hf = symm.rendezvous(flags, group)
        torch.cuda.synchronize(engine.index)
    dist.barrier(group=dp.group)
    world = dp.world
    gptrs = (C.c_void_p * (2 * world))(*[hg.buffer_ptrs[r] + b * total * 4 for b in (0, 1) for r in range(world)])
    fptrs = (C.c_void_p * world)(*[hf.buffer_ptrs[r] for r in range(world)])
    check(lib.ts_dp_bind(engine.handle, dp.rank, world, gptrs, fptrs), 'ts_dp_bind')
    # param.grad keeps showing this rank's local mean gradient of even steps (buffer 0)
    engine.grads = gbuf[:total]
    for i, p in enumerate(params):
        o, n = engine.offsets[i], p.numel()
        p.grad = engine.grads[o:o + n].view(p.shape)
    keep = [gbuf, flags, hg, hf]
    # small models (the one-hidden-layer path): inbox of the low-latency exchange inside the dW1 kernel (ts_dp_bind_ll)
    if total * 4 <= LL_MAX_MODEL_BYTES and os.environ.get('TSB200_DP_LL', '1') != '0':
        with torch.cuda.device(engine.index):
            inbox = symm.empty(LL_INBOX_BYTES // 4, dtype=torch.float32, device=dev)
            inbox.zero_()
            hi = symm.rendezvous(inbox, group)
            torch.cuda.synchronize(engine.index)
        dist.barrier(group=dp.group)
        iptrs = (C.c_void_p * world)(*[hi.buffer_ptrs[r] for r in range(world)])
        check(lib.ts_dp_bind_ll(engine.handle, iptrs, LL_INBOX_BYTES), 'ts_dp_bind_ll')
        keep += [inbox, hi]
    return tuple(keep)


def bind_nvls(engine, params, dp: DataParallel):
    """Data parallel for large models with the gradient reduced inside the NVSwitch: move the flat parameter, gradient and
    bf16 shadow buffers into torch symmetric memory, take their multicast mappings and hand everything to the library
    (ts_dp_bind_nvls). Returns the objects that must stay alive; raises if the fabric offers no multicast mapping."""
    import ctypes as C
    import torch.distributed._symmetric_memory as symm
    from ._lib import lib, check
    group = dp.group if dp.group is not None else dist.group.WORLD
    dev, total = engine.device, engine.total
    n_shadow = int(lib.ts_mlp_shadow_elems(engine.handle))
    with torch.cuda.device(engine.index):
        gbuf = symm.empty(total, dtype=torch.float32, device=dev)
        pbuf = symm.empty(total, dtype=torch.float32, device=dev)
        sbuf = symm.empty(max(n_shadow, 128), dtype=torch.bfloat16, device=dev)
        flags = symm.empty(64, dtype=torch.int32, device=dev)
        gbuf.zero_(); sbuf.zero_(); flags.zero_()
        pbuf.copy_(engine.params)
        handles = [symm.rendezvous(t, group) for t in (gbuf, pbuf, sbuf, flags)]
        torch.cuda.synchronize(engine.index)
    hg, hp, hs, hf = handles
    if not (hg.multicast_ptr and hp.multicast_ptr and hs.multicast_ptr):
        raise RuntimeError('symmetric memory on this system has no multicast mapping (NVLS): use dp_mode="nccl"')
    dist.barrier(group=dp.group)
    world, rank = dp.world, dp.rank
    fptrs = (C.c_void_p * world)(*[hf.buffer_ptrs[r] for r in range(world)])
    check(lib.ts_dp_bind_nvls(engine.handle, rank, world, gbuf.data_ptr(), hg.multicast_ptr, pbuf.data_ptr(), hp.multicast_ptr,
                              sbuf.data_ptr() if n_shadow else None, hs.multicast_ptr if n_shadow else None, fptrs), 'ts_dp_bind_nvls')
    keep = []
    if n_shadow and len(engine.spec.dims) > 3:
        # nets on the layer-by-layer path: a second shadow set lets each layer's exchange start under its own dX GEMM
        # (ts_dp_nvls_second_shadow; used when TSB200_NVLS_OVERLAP_SMS is set)
        with torch.cuda.device(engine.index):
            sbuf2 = symm.empty(n_shadow, dtype=torch.bfloat16, device=dev)
            sbuf2.zero_()
            hs2 = symm.rendezvous(sbuf2, group)
            torch.cuda.synchronize(engine.index)
        if hs2.multicast_ptr:
            check(lib.ts_dp_nvls_second_shadow(engine.handle, sbuf2.data_ptr(), hs2.multicast_ptr), 'ts_dp_nvls_second_shadow')
            keep = [sbuf2, hs2]
    engine.params, engine.grads = pbuf, gbuf
    for i, p in enumerate(params):
        o, n = engine.offsets[i], p.numel()
        p.data = pbuf[o:o + n].view(p.shape)
        p.grad = gbuf[o:o + n].view(p.shape)
    engine.sync_weights()
    torch.cuda.synchronize(engine.index)
    dist.barrier(group=dp.group)
    return gbuf, pbuf, sbuf, flags, handles, keep
