"""Data-parallel path on real GPUs (needs >= 2 devices; skipped otherwise): two ranks over NCCL, each fitting its half
of every global batch through B200Classifier (dp_mode 'p2p': gradient sum inside the optimizer kernel over peer
mappings; 'nccl': ts_fit_backward -> all-reduce -> ts_opt_apply; 'nvls': reduction inside the NVSwitch, parameters
broadcast through multicast stores), must reproduce the single-process oracle fed the full batch;
B200Metrics.compute() sums the counters of both ranks."""
import os
import socket

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def _worker(rank, world, port, precision, dp_mode, out, dims=(784, 256, 10), B=1024, optimizer='sgd'):
    import torch.distributed as dist
    from torch import nn
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group('nccl', rank=rank, world_size=world, device_id=torch.device('cuda', rank))
    from torchsystem_b200 import B200Classifier, B200Metrics, MLP
    from tests.test_gpu_parity import synthetic, seeded_init
    dims = list(dims)
    x, y = synthetic(B * 3)
    if precision == 'bf16':
        x = x.bfloat16().float()
    init = seeded_init(dims, seed=rank, bf16=precision == 'bf16')      # different per rank: the broadcast must fix it
    net = MLP(dims[0], dims[1] if len(dims) == 3 else dims[1:-1], dims[-1], dropout=0.0, activation=nn.ReLU())
    with torch.no_grad():
        for p, w in zip([q for lin in net.linears() for q in (lin.weight, lin.bias)], init):
            p.copy_(torch.from_numpy(w))
    opt = torch.optim.SGD(net.parameters(), lr=0.05) if optimizer == 'sgd' else torch.optim.Adam(net.parameters(), lr=1e-3)
    model = B200Classifier(net, nn.CrossEntropyLoss(), opt, precision=precision,
                           device=f'cuda:{rank}', max_batch=B, dp_mode=dp_mode)
    assert (model._p2p is not None) == (dp_mode in ('p2p', 'nvls'))
    assert model._overlap == (dp_mode == 'nccl' and os.environ.get('TSB200_DP_OVERLAP', '0') != '0')
    metrics = B200Metrics(device=f'cuda:{rank}', log_every=0)
    half = B // world
    trail = {}
    for s in range(3):
        lo = s * B + rank * half
        xb, yb = x[lo:lo + half].cuda(), y[lo:lo + half].cuda()
        for i, q in enumerate(q for lin in net.linears() for q in (lin.weight, lin.bias)):
            trail[f'pre_{s}_{i}'] = q.detach().cpu().numpy().copy()        # this replica's weights BEFORE step s
        p, l = model.fit(xb, yb)
        trail[f'pred_{s}'] = p.cpu().numpy()
        metrics.update(s + 1, l, p, yb)
    counters = metrics.counters()
    if optimizer == 'adam':
        assert model.engine.step_count == 3
        if dp_mode == 'nvls':       # the moments are sharded over the ranks: state_dict refuses until they are gathered
            with pytest.raises(RuntimeError):
                model.optimizer.state_dict()
            model.gather_optimizer_state()
        sd = model.optimizer.state_dict()
        assert all(int(st['step']) == 3 for st in sd['state'].values())
        trail['exp_avg_0'] = sd['state'][0]['exp_avg'].cpu().numpy()
    np.savez(out.replace('rank0', f'rank{rank}'), *[q.detach().cpu().numpy() for lin in net.linears() for q in (lin.weight, lin.bias)],
             counters=np.array(counters, dtype=np.float64), **trail)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason='needs two GPUs')
@pytest.mark.parametrize('dp_mode', ['p2p', 'nccl', 'nvls'])
@pytest.mark.parametrize('precision,tol', [('f32', 1e-5), ('bf16', 1e-2)])
def test_two_gpu_data_parallel_matches_full_batch_oracle(tmp_path, precision, tol, dp_mode):
    import torch.multiprocessing as mp
    from oracle import mlp_numpy as onp
    from tests.test_gpu_parity import synthetic, seeded_init
    from tests.golden_util import assert_params_close
    out = str(tmp_path / 'rank0.npz')
    mp.spawn(_worker, args=(2, _free_port(), precision, dp_mode, out), nprocs=2, join=True)
    dims, B = [784, 256, 10], 1024
    x, y = synthetic(B * 3)
    if precision == 'bf16':
        x = x.bfloat16().float()
    init = seeded_init(dims, seed=0, bf16=precision == 'bf16')
    oracle = onp.OracleMLP(weights=init[0::2], biases=init[1::2], opt_kind=onp.OPT_SGD, lr=0.05, emulate_bf16=precision == 'bf16')
    from tests.parity_util import assert_predictions_explained
    got = [np.load(out.replace('rank0', f'rank{r}')) for r in range(2)]
    correct, total, half = 0, 0, B // 2
    for s in range(3):
        xb, yb = x[s * B:(s + 1) * B].numpy(), y[s * B:(s + 1) * B].numpy()
        pre = [got[0][f'pre_{s}_{i}'] for i in range(4)]
        for i in range(4):      # replicas are bit-identical before every step (same reduction order on every rank)
            assert np.array_equal(pre[i], got[1][f'pre_{s}_{i}']), f'step {s}: replicas differ in tensor {i}'
        pred = np.concatenate([got[r][f'pred_{s}'] for r in range(2)])       # rank r fitted rows [r * half, (r+1) * half)
        if precision == 'f32':
            p, _ = oracle.evaluate(xb, yb)
            assert np.array_equal(pred, p)
        else:       # exact under the replicas' own weights; differences from the oracle trajectory explained row by row
            assert_predictions_explained(pred, oracle, pre, xb, None, f'{dp_mode} step {s}')
        correct += int((pred == yb).sum()); total += B
        oracle.fit(xb, yb, None)
    for i in range(4):
        assert np.array_equal(got[0][f'arr_{i}'], got[1][f'arr_{i}']), 'final replicas differ'
    assert_params_close([got[0][f'arr_{i}'] for i in range(4)], oracle.parameters(), tol, adam=False)
    loss_sum, n_batches, n_correct, n_total = got[0]['counters']
    assert (n_batches, n_total) == (6, total)
    assert n_correct == correct         # the all-reduced fused counters are exactly the count of both ranks' predictions


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason='needs two GPUs')
@pytest.mark.parametrize('overlap', ['buckets', 'backward', 'nvls', 'nvls_reserved_sms', 'nvls_one_kernel'])
def test_two_gpu_nccl_wide_layers(tmp_path, monkeypatch, overlap):
    """'nccl' mode on a net with 2048- and 4096-wide layers: several all-reduce buckets with the optimizer applied bucket by
    bucket (TSB200_DP_BUCKETS=1), or (TSB200_DP_OVERLAP=1) each layer's gradient all-reduced on the side stream while the
    backward pass is still running, biases in a packed all-reduce at the end. Must match the full-batch oracle."""
    monkeypatch.setenv('TSB200_DP_OVERLAP', '1' if overlap == 'backward' else '0')
    monkeypatch.setenv('TSB200_DP_BUCKETS', '1' if overlap == 'buckets' else '0')
    # 'nvls': each layer's in-switch reduce + update on a side stream under the rest of the backward pass (TSB200_NVLS_OVERLAP_SMS,
    # for nets on the layer-by-layer path); 'nvls_one_kernel': the whole buffer in one kernel after the backward pass
    # ('nvls': the exchange CTAs co-resident with the GEMM CTAs; 'nvls_reserved_sms': on 20 SMs the GEMMs leave free)
    monkeypatch.setenv('TSB200_NVLS_OVERLAP_SMS', {'nvls_one_kernel': '0', 'nvls_reserved_sms': '20'}.get(overlap, '-1'))
    import torch.multiprocessing as mp
    from oracle import mlp_numpy as onp
    from tests.test_gpu_parity import synthetic, seeded_init
    from tests.golden_util import assert_params_close
    dims, B = (784, 4096, 4096, 2048, 10), 1024
    out = str(tmp_path / 'rank0.npz')
    mp.spawn(_worker, args=(2, _free_port(), 'bf16', 'nvls' if overlap.startswith('nvls') else 'nccl', out, dims, B), nprocs=2, join=True)
    x, y = synthetic(B * 3)
    x = x.bfloat16().float()
    init = seeded_init(list(dims), seed=0, bf16=True)
    oracle = onp.OracleMLP(weights=init[0::2], biases=init[1::2], opt_kind=onp.OPT_SGD, lr=0.05, emulate_bf16=True)
    for s in range(3):
        oracle.fit(x[s * B:(s + 1) * B].numpy(), y[s * B:(s + 1) * B].numpy(), None)
    got = np.load(out)
    assert_params_close([got[f'arr_{i}'] for i in range(8)], oracle.parameters(), 1e-2, adam=False)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason='needs two GPUs')
def test_two_gpu_nvls_overlap_is_bit_identical_to_one_kernel_adam(tmp_path, monkeypatch):
    """The per-layer overlapped exchange does the same arithmetic as the single whole-buffer kernel (same in-switch sums, same
    chunk ownership, Adam bias correction advanced once per step): after three Adam steps on a 4-layer net the replicas and
    the gathered first moments must agree bit for bit between the two schedules."""
    import torch.multiprocessing as mp
    dims, B = (784, 2048, 2048, 1024, 10), 1024
    outs = []
    for sms in ('-1', '20', '0'):
        monkeypatch.setenv('TSB200_NVLS_OVERLAP_SMS', sms)
        out = str(tmp_path / f'sms{sms.replace("-", "m")}_rank0.npz')
        mp.spawn(_worker, args=(2, _free_port(), 'bf16', 'nvls', out, dims, B, 'adam'), nprocs=2, join=True)
        outs.append([np.load(out.replace('rank0', f'rank{r}')) for r in range(2)])
    for r in range(2):
        for i in range(8):
            for k in (1, 2):
                assert np.array_equal(outs[0][r][f'arr_{i}'], outs[k][r][f'arr_{i}']), f'rank {r} tensor {i}: schedule {k} differs'
            assert np.array_equal(outs[0][0][f'arr_{i}'], outs[0][1][f'arr_{i}']), f'tensor {i}: replicas differ'
        assert np.array_equal(outs[0][r]['exp_avg_0'], outs[1][r]['exp_avg_0']) and np.array_equal(outs[0][r]['exp_avg_0'], outs[2][r]['exp_avg_0'])
    assert np.array_equal(outs[0][0]['exp_avg_0'], outs[0][1]['exp_avg_0']), 'gathered moments differ between ranks'


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason='needs two GPUs')
@pytest.mark.parametrize('optimizer', ['sgd', 'adam'])
def test_two_gpu_low_latency_exchange_is_bit_identical_to_publish_and_read(tmp_path, monkeypatch, optimizer):
    """2,048 rows per rank: the dW1 + update kernel exchanges gradients as 16-byte packets {three floats, step} that the receivers
    poll in their own memory (ts_dp_bind_ll). Same terms, same order as the publish + remote-read exchange: the weights after three
    steps must be bit-identical between the two exchanges and between the replicas, and within 1e-2 of the full-batch oracle."""
    import torch.multiprocessing as mp
    from oracle import mlp_numpy as onp
    from tests.test_gpu_parity import synthetic, seeded_init
    from tests.golden_util import assert_params_close
    dims, B = [784, 256, 10], 4096
    runs = {}
    for name, env in (('packets', '1'), ('two_hops', '2'), ('publish_read', '0')) if optimizer == 'sgd' else (('packets', '1'), ('publish_read', '0')):
        monkeypatch.setenv('TSB200_DP_LL', env)
        out = str(tmp_path / f'{name}_rank0.npz')
        mp.spawn(_worker, args=(2, _free_port(), 'bf16', 'p2p', out, tuple(dims), B, optimizer), nprocs=2, join=True)
        runs[name] = [np.load(out.replace('rank0', f'rank{r}')) for r in range(2)]
    for name, got in runs.items():
        for i in range(4):
            assert np.array_equal(got[0][f'arr_{i}'], got[1][f'arr_{i}']), f'{name}: replicas differ in tensor {i}'
            assert np.array_equal(got[0][f'arr_{i}'], runs['publish_read'][0][f'arr_{i}']), f'{name} differs from publish + read in tensor {i}'
    if optimizer == 'adam':
        assert np.array_equal(runs['packets'][0]['exp_avg_0'], runs['publish_read'][0]['exp_avg_0'])
    x, y = synthetic(B * 3)
    x = x.bfloat16().float()
    init = seeded_init(dims, seed=0, bf16=True)
    oracle = onp.OracleMLP(weights=init[0::2], biases=init[1::2], opt_kind=onp.OPT_SGD if optimizer == 'sgd' else onp.OPT_ADAM,
                           lr=0.05 if optimizer == 'sgd' else 1e-3, emulate_bf16=True)
    for s in range(3):
        oracle.fit(x[s * B:(s + 1) * B].numpy(), y[s * B:(s + 1) * B].numpy(), None)
    assert_params_close([runs['packets'][0][f'arr_{i}'] for i in range(4)], oracle.parameters(), 1e-2, lr=1e-3, steps=3, adam=optimizer == 'adam')


def _late_peer_worker(rank, world, port, precision, out, B=1024):
    import torch.distributed as dist
    from torch import nn
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), TSB200_DP_TIMEOUT_S='1.5')
    torch.cuda.set_device(rank)
    dist.init_process_group('nccl', rank=rank, world_size=world, device_id=torch.device('cuda', rank))
    from torchsystem_b200 import B200Classifier, MLP, TsError
    from tests.test_gpu_parity import synthetic
    x, y = synthetic(B)
    torch.manual_seed(0)
    net = MLP(784, 256, 10, dropout=0.0, activation=nn.ReLU())
    model = B200Classifier(net, nn.CrossEntropyLoss(), torch.optim.SGD(net.parameters(), lr=0.05), precision=precision,
                           device=f'cuda:{rank}', max_batch=B, dp_mode='p2p')
    xb, yb = x.cuda(), y.cuda()
    model.fit(xb, yb)                                   # step 1: both ranks
    torch.cuda.synchronize()
    w1 = net.linears()[0].weight.detach().clone()
    result = 'none'
    if rank == 0:
        model.fit(xb, yb)                               # step 2: rank 1 never comes
        torch.cuda.synchronize()                        # returns: the kernel gave up after 1.5 s, it did not trap
        unchanged = torch.equal(w1, net.linears()[0].weight.detach())
        try:
            model.fit(xb, yb)
            result = 'no error'
        except TsError as e:
            result = f'{"timeout" if "did not arrive" in str(e) and "-6" in str(e) else "other"}|unchanged={unchanged}|{str(e)[:160]}'
        torch.zeros(1, device='cuda').sum().item()      # the context is still usable
        with open(out, 'w') as f:
            f.write(result)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason='needs two GPUs')
@pytest.mark.parametrize('precision,rows', [('bf16', 1024), ('bf16', 2048), ('f32', 1024)])
def test_a_peer_that_never_arrives_is_a_status_not_a_trap(tmp_path, precision, rows):
    """A rank whose peer does not reach the step within the configured time (TSB200_DP_TIMEOUT_S) skips that step's update
    and reports TS_E_TIMEOUT at its next call; nothing traps, the CUDA context survives (bf16: the dW1 + update kernel's
    in-kernel exchange -- publish + remote reads at 1,024 rows, low-latency packets at 2,048; f32: dp_optimizer_kernel)."""
    import torch.multiprocessing as mp
    out = str(tmp_path / 'result.txt')
    mp.spawn(_late_peer_worker, args=(2, _free_port(), precision, out, rows), nprocs=2, join=True)
    result = open(out).read()
    assert result.startswith('timeout|unchanged=True'), result
