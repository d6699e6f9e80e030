"""world_size-2 gloo tests (CPU) of the data-parallel host logic in torchsystem_b200/parallel.py: averaging the
local mean gradients of equal shards reproduces the single-process step, metric counters sum exactly,
parameter broadcast makes replicas identical. The same DataParallel object drives NCCL on the GPU box."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import mlp_numpy as onp


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def _data(n=64, dims=(784, 32, 10)):
    g = torch.Generator().manual_seed(11)
    x = torch.randn(n, dims[0], generator=g).numpy()
    y = torch.randint(0, dims[-1], (n,), generator=g).numpy()
    torch.manual_seed(0)
    init = []
    for i, o in zip(dims[:-1], dims[1:]):
        lin = torch.nn.Linear(i, o)
        init += [lin.weight.detach().numpy().copy(), lin.bias.detach().numpy().copy()]
    return x, y, init


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from torchsystem_b200.parallel import DataParallel, shard_bounds
    dp = DataParallel()
    x, y, init = _data()
    if rank == 1:   # replicas start different; broadcast from rank 0 must fix that
        init = [w + 1.0 for w in init]
    flat = torch.from_numpy(np.concatenate([w.ravel() for w in init]))
    dp.broadcast_(flat)
    shapes = [w.shape for w in init]
    sizes = [int(np.prod(s)) for s in shapes]
    init = [c.reshape(s).copy() for c, s in zip(np.split(flat.numpy(), np.cumsum(sizes)[:-1]), shapes)]
    model = onp.OracleMLP(weights=init[0::2], biases=init[1::2], opt_kind=onp.OPT_ADAM, lr=1e-3)
    lo, hi = shard_bounds(x.shape[0], rank, world)

    def hook(grads):
        t = torch.from_numpy(np.concatenate([g.ravel() for g in grads]))
        dp.all_reduce_sum_(t)
        t /= dp.world
        return [c.reshape(s) for c, s in zip(np.split(t.numpy(), np.cumsum(sizes)[:-1]), shapes)]

    metrics = onp.OracleMetrics()
    for step in range(3):
        pred, loss = model.fit(x[lo:hi], y[lo:hi], None, grad_hook=hook)
        metrics.update(step + 1, loss, pred, y[lo:hi])
    tot = dp.reduce_metrics(metrics.weighted_sum, int(metrics.weights), metrics.num_correct, metrics.num_total, 'cpu')
    if rank == 0:
        np.savez(out, *model.parameters(), counters=np.array(tot, dtype=np.float64))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gradient_average_equals_full_batch(tmp_path):
    out = str(tmp_path / 'rank0.npz')
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    x, y, init = _data()
    single = onp.OracleMLP(weights=init[0::2], biases=init[1::2], opt_kind=onp.OPT_ADAM, lr=1e-3)
    correct = 0
    losses = 0.0
    halves = [(0, 32), (32, 64)]
    for _ in range(3):
        # what each rank saw before the step: predictions/loss of its own half under the shared weights
        for lo, hi in halves:
            p, l = single.evaluate(x[lo:hi], y[lo:hi])
            correct += int((p == y[lo:hi]).sum()); losses += float(l)
        single.fit(x, y, None)
    got = np.load(out)
    for i, w in enumerate(single.parameters()):
        np.testing.assert_allclose(got[f'arr_{i}'], w, rtol=2e-4, atol=2e-6)
    loss_sum, n_batches, n_correct, n_total = got['counters']
    assert (n_batches, n_correct, n_total) == (6, correct, 192)
    assert loss_sum == pytest.approx(losses, rel=1e-5)


def test_single_process_is_a_no_op():
    from torchsystem_b200.parallel import DataParallel
    dp = DataParallel()
    assert (dp.world, dp.rank) == (1, 0)
    t = torch.ones(4)
    dp.all_reduce_sum_(t); dp.broadcast_(t)
    assert t.tolist() == [1, 1, 1, 1]
    assert dp.reduce_metrics(1.5, 2, 3, 4, 'cpu') == (1.5, 2, 3, 4)
