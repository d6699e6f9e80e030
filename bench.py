#!/usr/bin/env python
"""bench.py -- train() samples/sec of the TorchSystem hot path on B200 (BASELINE.json `metric`).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (BASELINE.json configs[1]): MLP(784,256,10), dropout 0.1 (in-kernel Philox),
ReLU, CrossEntropyLoss, SGD(lr=0.01), batch 4096, 1,048,576 synthetic samples per GPU. One STEP is one pass of the
`train` handler (model.fit + metrics.update per batch, metrics.compute, Trained event) over those 256 batches.
Weak scaling: every rank owns its own 1,048,576-sample shard, gradients are all-reduced every batch.

* value   whole-job samples/s, inputs resident in HBM (DeviceLoader); device-timed, max over ranks.
* e2e     same handler, inputs in pinned HOST memory copied host->device every step (PinnedLoader), the
          phase result read back device->host.
* roofline / cpu_baseline / clocks / gpu_launches: see DESIGN.md "Measurement".
* --impl reference  times the reference's CPU path (oracle/ref_torch.py: the same stock-torch call sequence as
          the reference's Classifier/handlers) on the host cores, on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

DIMS = [784, 256, 10]
BATCH = 4096
BATCHES_PER_STEP = 256
DROPOUT = 0.1
LR = 0.01
BYTES_PER_SAMPLE = 784 * 4 + 8 + 8          # x f32 + target i64 + prediction i64


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--precision', default=os.environ.get('TSB_BENCH_PRECISION', 'bf16'), choices=['bf16', 'f32'])
    ap.add_argument('--optimizer', default='sgd', choices=['sgd', 'adam'])
    ap.add_argument('--batches', type=int, default=BATCHES_PER_STEP, help='batches of 4096 per step')
    ap.add_argument('--dp-mode', default='auto', choices=['auto', 'p2p', 'nvls', 'nccl'], help='gradient exchange at N > 1')
    ap.add_argument('--e2e-steps', type=int, default=5)
    ap.add_argument('--cpu-seconds', type=float, default=12.0, help='budget of the cpu_baseline leg')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--skip-e2e', action='store_true', help='profiling runs only')
    ap.add_argument('--skip-extras', action='store_true', help='skip the non-headline extra measurements')
    return ap.parse_args()


# ----------------------------------------------------------------------------- CPU arm (reference path)
def cpu_reference_run(steps: int, warmup: int, batches_per_step: int, seconds_cap: float | None, optimizer: str, batch: int = BATCH,
                      min_warm_seconds: float = 3.0):
    """The reference's stock-torch CPU path over `batches_per_step` batches of `batch` per step. Warm-up steps are FULL steps
    (the first pass over a buffer pays page faults and thread-pool start-up) and last at least `min_warm_seconds`: the same code
    measured 0.36 M samples/s after 3 short warm-up steps and 0.45 M after ten seconds; `value` is taken from the MEDIAN step."""
    import torch
    from oracle import ref_torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    torch.manual_seed(0)
    net = ref_torch.make_layers(DIMS, DROPOUT, 'relu')
    opt = torch.optim.SGD(net.parameters(), lr=LR) if optimizer == 'sgd' else torch.optim.Adam(net.parameters(), lr=1e-3)
    model, metrics = ref_torch.CpuClassifier(net, opt), ref_torch.CpuMetrics()
    g = torch.Generator().manual_seed(1234)
    n = batches_per_step * batch
    x = torch.randn(n, 1, 28, 28, generator=g)
    y = torch.randint(0, 10, (n,), generator=g)
    loader = [(x[i:i + batch], y[i:i + batch]) for i in range(0, n, batch)]
    tw, done_w = time.perf_counter(), 0
    while done_w < max(warmup, 1) or time.perf_counter() - tw < min_warm_seconds:
        ref_torch.train(model, loader, metrics); metrics.reset()     # the CPU path needs seconds, not steps, to reach its rate
        done_w += 1
    times, t0 = [], time.perf_counter()
    for _ in range(steps):
        t1 = time.perf_counter()
        ref_torch.train(model, loader, metrics); metrics.reset()
        times.append(time.perf_counter() - t1)
        if seconds_cap is not None and time.perf_counter() - t0 > seconds_cap:
            break
    dt = time.perf_counter() - t0
    med = sorted(times)[len(times) // 2]
    return {'value': n / med, 'mean_value': len(times) * n / dt, 'seconds': dt, 'steps_done': len(times), 'cores': cores,
            'median_step_s': med,
            'sample': f'{len(times)} x {batches_per_step} batches of {batch} (S model, {optimizer}, dropout {DROPOUT}), median step, '
                      f'torch {torch.__version__} CPU'}


def reference_arm(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    per_step = 16                       # bounded sample: 16 batches of 4096 = 65,536 samples per step
    r = cpu_reference_run(args.steps, min(args.warmup, 3), per_step, None, args.optimizer)
    ms = r['median_step_s'] * 1e3
    line = {
        'impl': 'reference', 'metric': 'train_samples_per_sec', 'value': r['value'], 'unit': 'samples/s',
        'n_gpus': args.gpus, 'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': ms, 'higher_is_better': True,
        'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
        'config': {'workload': f'tinysys train handler, MLP(784,256,10) {args.optimizer} batch {BATCH}; CPU sample of '
                               f'{per_step} batches per step', 'batch': BATCH, 'samples_per_step': per_step * BATCH},
        'cpu_baseline': {'value': r['value'], 'unit': 'samples/s', 'cores': r['cores'], 'kind': 'port', 'sample': r['sample']},
        'e2e': {'value': r['value'], 'unit': 'samples/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region. The region is tens of milliseconds long, shorter than
    one `nvidia-smi -lms` period, so the samples come from NVML in a thread (every ~8 ms: an NVML query takes driver locks the
    kernel launches of the same process also need, so it is kept rare; always at least one sample taken while the GPU is busy); nvidia-smi is only the fallback when the NVML binding is missing."""
    Q = ('clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,'
         'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, index: int):
        self.sm, self.mx, self.reasons, self.p, self.t = [], [], set(), None, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            phys = index
            vis = os.environ.get('CUDA_VISIBLE_DEVICES')
            if vis:
                ids = [v.strip() for v in vis.split(',') if v.strip()]
                if index < len(ids) and ids[index].isdigit():
                    phys = int(ids[index])
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_sm = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            import threading
            self.stop_flag = threading.Event()
            self.t = threading.Thread(target=self._loop, daemon=True)
            self.t.start()
        except Exception:
            self.nv = None
            self.f = tempfile.NamedTemporaryFile('w+', suffix='.csv', delete=False)
            try:
                self.p = subprocess.Popen(['nvidia-smi', f'--id={index}', f'--query-gpu={self.Q}', '--format=csv,noheader,nounits',
                                           '-lms', '20'], stdout=self.f, stderr=subprocess.DEVNULL)
            except OSError:
                self.p = None

    def _sample(self):
        nv = self.nv
        self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
        self.mx.append(self.max_sm)
        r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
        for name, bit in (('hw_slowdown', nv.nvmlClocksThrottleReasonHwSlowdown),
                          ('hw_thermal_slowdown', nv.nvmlClocksThrottleReasonHwThermalSlowdown),
                          ('sw_thermal_slowdown', nv.nvmlClocksThrottleReasonSwThermalSlowdown),
                          ('sw_power_cap', nv.nvmlClocksThrottleReasonSwPowerCap)):
            if r & bit:
                self.reasons.add(name)

    def _loop(self):
        while not self.stop_flag.is_set():
            try:
                self._sample()
            except Exception:
                break
            time.sleep(0.008)

    def sample_now(self):
        """One sample from the calling thread (called right after the timed launches are enqueued: the GPU is busy)."""
        if self.nv is not None:
            try:
                self._sample()
            except Exception:
                pass

    def stop(self) -> dict:
        if self.nv is not None:
            self.stop_flag.set()
            self.t.join(timeout=2)
            sm = sorted(self.sm)
            return {'sm_mhz': sm[len(sm) // 2] if sm else None, 'sm_max_mhz': max(self.mx) if self.mx else None,
                    'reasons': sorted(self.reasons), 'samples': len(sm), 'source': 'nvml'}
        if self.p is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush(); self.f.seek(0)
        sm, mx, reasons = [], [], set()
        for row in self.f.read().splitlines():
            c = [t.strip() for t in row.split(',')]
            if len(c) < 8:
                continue
            try:
                sm.append(float(c[0])); mx.append(float(c[1]))
            except ValueError:
                continue
            for name, v in zip(('hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap'), c[4:8]):
                if v.lower().startswith('active'):
                    reasons.add(name)
        os.unlink(self.f.name)
        sm.sort()
        return {'sm_mhz': sm[len(sm) // 2] if sm else None, 'sm_max_mhz': max(mx) if mx else None,
                'reasons': sorted(reasons), 'samples': len(sm), 'source': 'nvidia-smi'}



L_DIMS = [784, 4096, 4096, 4096, 4096, 10]
L_FLOP_PER_SAMPLE = 315080704          # 6 * sum(in * out) of the dense layers, forward + backward
GLOBAL_BATCH_CFG4 = 65536


def config4_dp(torch, dist, dev, world, rank, x, y, precision, peaks, steps=6, warmup=2):
    """BASELINE.json configs[3]: MLP 784-4096x4-10, Adam, GLOBAL batch 65,536 split over the ranks (65,536 rows on one GPU at
    N=1), one fit per step. Device-timed, max over ranks. At N>1 every gradient-exchange mode is timed."""
    from torch import nn
    from torchsystem_b200 import B200Classifier, MLP
    per = GLOBAL_BATCH_CFG4 // world
    out = {'model': 'MLP 784-4096x4-10, Adam lr 1e-3, dropout 0.1', 'global_batch': GLOBAL_BATCH_CFG4, 'rows_per_gpu': per,
           'unit': 'samples/s', 'modes': {}}
    n_sets = max(1, min(4, x.shape[0] // per))
    for mode in (['single'] if world == 1 else ['nvls', 'nvls_overlap', 'nvls_overlap_20sm', 'nccl']):
        try:
            # 'nvls_overlap': each layer's in-switch reduce + update on a side stream under the rest of the backward pass, its CTAs
            # co-resident with the GEMM CTAs; '_20sm': on 20 SMs taken from the GEMMs instead; 'nvls': one kernel after the backward pass
            os.environ['TSB200_NVLS_OVERLAP_SMS'] = {'nvls_overlap': '-1', 'nvls_overlap_20sm': '20'}.get(mode, '0')
            torch.manual_seed(0)
            net = MLP(L_DIMS[0], L_DIMS[1:-1], L_DIMS[-1], dropout=DROPOUT, activation=nn.ReLU())
            model = B200Classifier(net, nn.CrossEntropyLoss(), torch.optim.Adam(net.parameters(), lr=1e-3), precision=precision,
                                   device=dev, max_batch=per, **({} if world == 1 else {'dp_mode': mode.split('_')[0]}))
            def step(i):
                j = i % n_sets
                return model.fit(x[j * per:(j + 1) * per], y[j * per:(j + 1) * per])
            for i in range(warmup):
                step(i)
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(steps):
                _, loss = step(i)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / steps
            if world > 1:
                t = torch.tensor([ms], dtype=torch.float64, device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                ms = t.item()
            tfl = per * L_FLOP_PER_SAMPLE / (ms * 1e-3) / 1e12
            out['modes'][mode] = {'ms_per_step': ms, 'value': GLOBAL_BATCH_CFG4 / (ms * 1e-3), 'tflops_per_gpu': tfl,
                                  'frac_of_bf16_sustained_peak': tfl / peaks.get('bf16_tflops_sustained', 1402.0),
                                  'last_loss': float(loss.item())}
            del model, net
            torch.cuda.empty_cache()
        except Exception as e:          # recorded, never fatal for the headline line
            out['modes'][mode] = {'error': repr(e)[:300]}
    ok = {k: v for k, v in out['modes'].items() if 'value' in v}
    if ok:
        best = max(ok, key=lambda k: ok[k]['value'])
        out.update(dp_mode=best, ms_per_step=ok[best]['ms_per_step'], value=ok[best]['value'], tflops_per_gpu=ok[best]['tflops_per_gpu'])
    out['note'] = ('same GLOBAL batch at every N: value(N) / value(1) from the N=1 record is the strong-scaling speed-up the '
                   'north star asks for (>= 7 at 8 GPUs)')
    return out


def dp_parity(torch, dist, dev, world, rank, precision, dp_mode):
    """Cross-rank parity evidence in the bench line (N > 1): three steps of the S model (dropout off, SGD) data-parallel over
    all ranks against the same three steps on rank 0 alone with the full global batch, and a bit-comparison of the replicas."""
    from torch import nn
    from torchsystem_b200 import B200Classifier, MLP
    tol = 1e-2 if precision == 'bf16' else 1e-5
    solo_group = dist.new_group([0])
    per = 2048          # rows per rank: the size class of the headline (dW1 + update kernel on clusters of 8)
    g = torch.Generator(device=dev).manual_seed(4242)            # same data on every rank
    xg = torch.randn(3 * per * world, 1, 28, 28, generator=g, device=dev).bfloat16().float()
    teacher = torch.randn(784, 10, generator=g, device=dev)
    yg = torch.argmax(xg.flatten(1) @ teacher, dim=1)

    def build(group, mode):
        torch.manual_seed(7)
        net = MLP(DIMS[0], DIMS[1], DIMS[2], dropout=0.0, activation=nn.ReLU())
        return B200Classifier(net, nn.CrossEntropyLoss(), torch.optim.SGD(net.parameters(), lr=0.05), precision=precision, device=dev,
                              max_batch=per * world, process_group=group, dp_mode=mode), net
    model, net = build(None, dp_mode)
    losses, correct = [], []
    for s in range(3):
        lo = (s * world + rank) * per
        p, l = model.fit(xg[lo:lo + per], yg[lo:lo + per])
        t = torch.stack([l.double(), (p == yg[lo:lo + per]).sum().double()])
        dist.all_reduce(t)
        losses.append(t[0].item() / world); correct.append(int(t[1].item()))
    flat = model.engine.params
    digest = torch.stack([flat.view(torch.int32).to(torch.int64).sum(), (flat.view(torch.int32).to(torch.int64) * 31 % 1000003).sum()])
    all_d = [torch.zeros_like(digest) for _ in range(world)]
    dist.all_gather(all_d, digest)
    identical = all(torch.equal(all_d[0], d) for d in all_d)
    res = {'dp_mode': dp_mode, 'replicas_bit_identical': bool(identical), 'dp_losses': losses, 'dp_correct': correct, 'tolerance': tol}
    if rank == 0:
        solo, net1 = build(solo_group, 'p2p')
        s_loss, s_corr = [], []
        for s in range(3):
            lo = s * world * per
            p, l = solo.fit(xg[lo:lo + per * world], yg[lo:lo + per * world])
            s_loss.append(l.item()); s_corr.append(int((p == yg[lo:lo + per * world]).sum().item()))
        wa = torch.cat([q.detach().flatten() for q in net.parameters()])
        wb = torch.cat([q.detach().flatten() for q in net1.parameters()])
        rel = ((wa - wb).norm() / wb.norm()).item()
        res.update(full_batch_losses=s_loss, full_batch_correct=s_corr, weights_rel_err=rel,
                   loss_rel_err=max(abs(a - b) / abs(b) for a, b in zip(losses, s_loss)),
                   ok=bool(identical and rel <= tol and all(abs(a - b) <= tol * abs(b) for a, b in zip(losses, s_loss))))
        del solo
    dist.barrier()
    del model
    return res


def h2d_ceiling(torch, dist, dev, world):
    """Pinned host -> device copy rate of this rank while every rank copies at once (what bounds `e2e`)."""
    buf = torch.empty(256 << 20, dtype=torch.uint8).pin_memory()
    dst = torch.empty_like(buf, device=dev)
    dst.copy_(buf, non_blocking=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(8):
        dst.copy_(buf, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    gbs = 8 * buf.numel() / (e0.elapsed_time(e1) * 1e-3) / 1e9
    if world > 1:
        t = torch.tensor([gbs], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        gbs = t.item()
    return gbs


# ----------------------------------------------------------------------------- B200 arm
def main():
    args = parse()
    if args.impl == 'reference':
        return reference_arm(args)

    import torch
    import torch.distributed as dist
    from torch import nn
    import __graft_entry__ as entry
    entry.build()
    from torchsystem_b200 import B200Classifier, B200Metrics, DeviceLoader, PinnedLoader, MLP, training, _lib

    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        # keep stdout to the single JSON line: NCCL prints its version banner there on first use
        os.environ.setdefault('NCCL_DEBUG', 'WARN')
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group('nccl', device_id=dev)
            dist.all_reduce(torch.zeros(1, device=dev))
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)

    n = args.batches * BATCH
    g = torch.Generator(device=dev).manual_seed(1234 + rank)
    x = torch.randn(n, 1, 28, 28, generator=g, device=dev)
    teacher = torch.randn(784, 10, generator=torch.Generator(device=dev).manual_seed(99), device=dev)
    y = torch.argmax(x.flatten(1) @ teacher, dim=1)

    def make_model(dp_mode):
        torch.manual_seed(0)
        net = MLP(DIMS[0], DIMS[1], DIMS[2], dropout=DROPOUT, activation=nn.ReLU())
        opt = torch.optim.SGD(net.parameters(), lr=LR) if args.optimizer == 'sgd' else torch.optim.Adam(net.parameters(), lr=1e-3)
        return B200Classifier(net, nn.CrossEntropyLoss(), opt, precision=args.precision, device=dev, max_batch=BATCH, dp_mode=dp_mode)

    # N > 1: which gradient exchange the S model uses is decided by a short trial of both one-kernel modes on this box
    # ('p2p': every rank reads all gradients over peer mappings; 'nvls': reduce-scatter / all-gather through the NVSwitch)
    dp_trial = None
    dp_mode = args.dp_mode
    if world > 1 and dp_mode == 'auto':
        dp_trial = {}
        for mode in ('p2p', 'nvls'):
            try:
                cand = make_model(mode)
                def run(k):
                    for i in range(k):
                        cand.fit(x[(i % args.batches) * BATCH:(i % args.batches + 1) * BATCH], y[(i % args.batches) * BATCH:(i % args.batches + 1) * BATCH])
                run(32)
                dist.barrier(); torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); run(128); e1.record()
                torch.cuda.synchronize()
                t = torch.tensor([e0.elapsed_time(e1) / 128 * 1e3], dtype=torch.float64, device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                dp_trial[mode] = t.item()
                del cand
            except Exception as e:
                dp_trial[mode] = None
                dp_trial[mode + '_error'] = repr(e)[:200]
        ok = {k: v for k, v in dp_trial.items() if isinstance(v, float)}
        # every rank saw the same all-reduced numbers, so every rank picks the same mode
        dp_mode = min(ok, key=ok.get) if ok else 'p2p'
    elif dp_mode == 'auto':
        dp_mode = 'p2p'
    model = make_model(dp_mode)
    metrics = B200Metrics(device=dev, log_every=0)
    training.provider.override(training.device, lambda: dev)
    results = []
    consumer = __import__('torchsystem_b200.compat', fromlist=['Consumer']).Consumer()

    def on_trained(e: training.Trained):
        results.append((e.results['loss'].item(), e.results['accuracy'].item()))     # device->host read of the result
    on_trained.__annotations__ = {'e': training.Trained}
    consumer.handler(on_trained)
    training.producer.consumers.clear()
    training.producer.register(consumer)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(loader, steps, warmup, sample_clocks=False):
        for _ in range(warmup):
            training.service.handle('train', model, loader, metrics); metrics.reset()
        barrier()
        clocks = ClockSampler(local) if sample_clocks and rank == 0 else None
        launches0 = _lib.lib.ts_launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            training.service.handle('train', model, loader, metrics); metrics.reset()
        e1.record()
        if clocks:
            clocks.sample_now()
        barrier()
        ms = e0.elapsed_time(e1)
        launches = _lib.lib.ts_launch_count() - launches0
        t = torch.tensor([ms, float(launches)], dtype=torch.float64, device=dev)
        if world > 1:
            tm = t.clone(); dist.all_reduce(tm, op=dist.ReduceOp.MAX)
            ts = t.clone(); dist.all_reduce(ts, op=dist.ReduceOp.SUM)
            ms, launches = tm[0].item(), int(ts[1].item())
        return ms, int(launches), (clocks.stop() if clocks else None)

    # ---- value: inputs resident in HBM
    dev_loader = DeviceLoader(x, y, BATCH, device=dev)
    ms, launches, clocks = timed(dev_loader, args.steps, args.warmup, sample_clocks=True)
    value = world * n * args.steps / (ms * 1e-3)

    # ---- roofline: one more identical step with CUDA events around every kernel (on the launching stream)
    _lib.profile_begin()
    training.service.handle('train', model, dev_loader, metrics); metrics.reset()
    prof = _lib.profile_end()
    total_ms = sum(v[0] for v in prof.values())
    top = max(prof, key=lambda k: prof[k][0])
    esz = 2 if args.precision == 'bf16' else 4
    H, D = DIMS[1], DIMS[0]
    algo_bytes = {   # per launch: what the kernel must read once + write once (DESIGN.md "Kernels")
        'cast_bf16': BATCH * D * 6,
        'fwd_gemm': BATCH * D * esz + H * D * esz + H * 4 + BATCH * H * esz,
        'head': BATCH * H * esz * 2 + BATCH * (8 + 8 + DIMS[2] * 4) + DIMS[2] * H * 4,
        'head_wgrad': BATCH * H * esz + BATCH * DIMS[2] * 4,
        'colsum': BATCH * H * esz,
        'dw_gemm': BATCH * H * esz + BATCH * D * esz + H * D * 4,
        'optimizer': sum(DIMS[i] * DIMS[i + 1] + DIMS[i + 1] for i in range(2)) * (12 if args.optimizer == 'sgd' else 28),
        # dW1 GEMM + in-cluster reduce + update: dH + x (bf16) read once, W1 master read + written, grads + bf16 shadow written,
        # optimizer moments read + written (Adam), the per-slab partials of the small tensors
        'dw_gemm_update': BATCH * H * 2 + BATCH * D * 2 + H * D * ((8 if args.optimizer == 'sgd' else 24) + 4 + 2)
                          + (BATCH // 128) * (DIMS[2] * H + H + DIMS[2]) * 4,
        # fused forward + head + dH: x (bf16 copy) + W1 + dH out + pred/target + per-slab gradient partials
        'fused_fwd_head': BATCH * D * 2 + H * D * 2 + BATCH * H * 2 + BATCH * 16 + (BATCH // 128) * (DIMS[2] * H + H + DIMS[2]) * 4,
    }
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
    except OSError:
        pass
    peak = float(peaks.get('hbm_gbs', 6650.0))
    top_ms, top_n = prof[top]
    achieved = algo_bytes.get(top, 0) / (top_ms / top_n * 1e-3) / 1e9 if top_n else 0.0
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, 'profiles', 'traffic.json'))).get(f'{top}:{args.precision}')
    except OSError:
        pass
    step_bytes = BYTES_PER_SAMPLE * n
    roofline = {
        'bound': 'hbm', 'kernel': top, 'achieved': achieved, 'peak': peak, 'unit': 'GB/s', 'frac': achieved / peak,
        'traffic': traffic, 'peak_source': 'MEASURED_PEAKS.json hbm_gbs' if peaks else 'fallback 6650 GB/s',
        'algorithmic_bytes_per_launch': algo_bytes.get(top), 'avg_launch_us': top_ms / max(top_n, 1) * 1e3,
        'kernel_share_of_step': {k: round(v[0] / total_ms, 4) for k, v in prof.items()},
        'kernel_avg_us': {k: round(v[0] / v[1] * 1e3, 3) for k, v in prof.items()},
        'step_streaming_gbs': step_bytes * world / (ms / args.steps * 1e-3) / 1e9 / world,
        'step_streaming_frac': step_bytes / (ms / args.steps * 1e-3) / 1e9 / peak,
    }

    # ---- extras (not the headline): the same train step with the loader holding bf16 (what bf16 mode rounds the inputs
    #      to anyway), and the validate handler at batch 65,536 (BASELINE.json configs[4]: HBM-bound evaluation)
    extras = {}
    if not args.skip_extras:
        bf_loader = DeviceLoader(x, y, BATCH, device=dev, dtype=torch.bfloat16)
        b_ms, _, _ = timed(bf_loader, args.steps, 2)
        extras['train_bf16_resident_inputs'] = {'value': world * n * args.steps / (b_ms * 1e-3), 'unit': 'samples/s',
                                                'note': 'DeviceLoader(dtype=bfloat16): no per-batch cast kernel, 1,584 B/sample streamed'}
        del bf_loader
        vb = 65536
        val_loader = DeviceLoader(x, y, vb, device=dev)
        results_v = []
        def run_val():
            training.service.handle('validate', model, val_loader, metrics); metrics.reset()
        for _ in range(2):
            run_val()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            run_val()
        e1.record()
        barrier()
        v_ms = e0.elapsed_time(e1)
        v_rate = n * args.steps / (v_ms * 1e-3)
        extras['validate_batch65536'] = {'value': world * v_rate, 'unit': 'samples/s', 'per_gpu_streaming_gbs': v_rate * 3152 / 1e9,
                                         'frac_of_hbm_peak': v_rate * 3152 / 1e9 / peak,
                                         'note': 'validate handler (inference_mode, evaluate + fused metrics), fp32 inputs resident in HBM; persistent pipelined '
                                                 'evaluate kernel (eval_stream.cu) from 148 slabs up'}

    # ---- the GPU incumbent: stock PyTorch eager on the same B200, same model / batch / optimizer,
    #      fp32 and under autocast(bf16); plain library code, not part of the product
    if not args.skip_extras and world == 1:
        def eager(autocast):
            torch.manual_seed(0)
            net_e = MLP(DIMS[0], DIMS[1], DIMS[2], dropout=DROPOUT, activation=nn.ReLU()).to(dev)
            opt_e = torch.optim.SGD(net_e.parameters(), lr=LR)
            crit = nn.CrossEntropyLoss()
            def step(i):
                xb, yb = x[i * BATCH:(i + 1) * BATCH].flatten(1), y[i * BATCH:(i + 1) * BATCH]
                opt_e.zero_grad()
                with torch.autocast('cuda', dtype=torch.bfloat16, enabled=autocast):
                    out = net_e(xb)
                    loss = crit(out, yb)
                loss.backward()
                opt_e.step()
                return out.argmax(1), loss
            for i in range(16):
                step(i)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(args.batches):
                step(i)
            e1.record()
            torch.cuda.synchronize()
            return n / (e0.elapsed_time(e1) * 1e-3)
        extras['torch_eager_same_gpu'] = {'fp32': eager(False), 'autocast_bf16': eager(True), 'unit': 'samples/s',
                                          'note': 'stock PyTorch eager Classifier.fit loop (zero_grad, forward, CrossEntropyLoss, backward, '
                                                  'SGD.step, argmax) on this GPU, batch 4096, inputs resident: the GPU incumbent'}

    # ---- extras for BASELINE.json configs[2] and configs[3] (one GPU only; not the headline)
    if not args.skip_extras and world == 1:
        import time
        # configs[2]: the tinysys aggregate as shipped (Adam, batch 32) through a full `iterate` (train + validation
        # phases, events dispatched to a consumer that reads the results like the logging consumer does); wall clock
        torch.manual_seed(0)
        net3 = MLP(DIMS[0], DIMS[1], DIMS[2], dropout=DROPOUT, activation=nn.ReLU())
        model3 = B200Classifier(net3, nn.CrossEntropyLoss(), torch.optim.Adam(net3.parameters(), lr=1e-3), precision=args.precision,
                                device=dev, max_batch=32)
        n_tr, n_va = 32 * 1024, 32 * 256
        loaders = [('train', DeviceLoader(x[:n_tr], y[:n_tr], 32, device=dev)),
                   ('validation', DeviceLoader(x[n_tr:n_tr + n_va], y[n_tr:n_tr + n_va], 32, device=dev))]
        training.service.handle('iterate', model3, loaders, metrics)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(3):
            training.service.handle('iterate', model3, loaders, metrics)
        torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) / 3
        extras['config3_iterate_batch32'] = {
            'value': (n_tr + n_va) / wall, 'unit': 'samples/s', 'wall_ms_per_epoch': wall * 1e3,
            'us_per_batch': wall / ((n_tr + n_va) / 32) * 1e6,
            'note': 'shipped aggregate (Adam lr 1e-3, dropout 0.1), iterate = 1024 train + 256 validation batches of 32, '
                    'events dispatched and read by a consumer; host wall clock (launch-bound at this batch size)'}
        del model3, loaders
        # configs[3]'s model on one GPU: 784 -> 4096 x 4 -> 10, the 8-GPU share (8,192) of the global batch 65,536;
        # 315,080,704 FLOP per sample against the measured cuBLAS bf16 peaks
        BL = 8192
        torch.manual_seed(0)
        netL = MLP(DIMS[0], [4096] * 4, DIMS[2], dropout=DROPOUT, activation=nn.ReLU())
        modelL = B200Classifier(netL, nn.CrossEntropyLoss(), torch.optim.Adam(netL.parameters(), lr=1e-3), precision=args.precision,
                                device=dev, max_batch=BL)
        xl, yl = x[:BL * 4], y[:BL * 4]
        for i in range(3):
            modelL.fit(xl[(i % 4) * BL:(i % 4 + 1) * BL], yl[(i % 4) * BL:(i % 4 + 1) * BL])
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for i in range(12):
            modelL.fit(xl[(i % 4) * BL:(i % 4 + 1) * BL], yl[(i % 4) * BL:(i % 4 + 1) * BL])
        e1.record()
        torch.cuda.synchronize()
        l_ms = e0.elapsed_time(e1) / 12
        tfl = BL * 315080704 / (l_ms * 1e-3) / 1e12
        extras['config4_L_model_one_gpu'] = {
            'value': BL / (l_ms * 1e-3), 'unit': 'samples/s', 'ms_per_step': l_ms, 'tflops': tfl,
            'frac_of_bf16_sustained_peak': tfl / (peaks or {}).get('bf16_tflops_sustained', 1363.5),
            'frac_of_bf16_burst_peak': tfl / (peaks or {}).get('bf16_tflops', 1650.7),
            'note': 'MLP 784-4096x4-10, Adam, batch 8192 per step, fp32 inputs resident; dense-layer FLOPs only'}
        del modelL, netL

    # ---- BASELINE.json configs[3] at its real global batch, every N (SURVEY.md section 8d "Config 4", section 8e)
    parity = None
    if not args.skip_extras:
        extras['config4_dp'] = config4_dp(torch, dist, dev, world, rank, x, y, args.precision, peaks or {})
        if world > 1:
            try:
                parity = dp_parity(torch, dist, dev, world, rank, args.precision, dp_mode)
            except Exception as e:
                parity = {'error': repr(e)[:300]}
    if not args.skip_extras and world == 1:
        # the GPU incumbent for the L model: stock PyTorch eager (cuBLASLt GEMMs), same model / batch / optimizer
        def eager_L(autocast):
            BL = 8192
            torch.manual_seed(0)
            net_e = MLP(L_DIMS[0], L_DIMS[1:-1], L_DIMS[-1], dropout=DROPOUT, activation=nn.ReLU()).to(dev)
            opt_e = torch.optim.Adam(net_e.parameters(), lr=1e-3)
            crit = nn.CrossEntropyLoss()
            def step(i):
                xb, yb = x[(i % 4) * BL:(i % 4 + 1) * BL].flatten(1), y[(i % 4) * BL:(i % 4 + 1) * BL]
                opt_e.zero_grad()
                with torch.autocast('cuda', dtype=torch.bfloat16, enabled=autocast):
                    out = net_e(xb)
                    loss = crit(out, yb)
                loss.backward()
                opt_e.step()
                return out.argmax(1), loss
            for i in range(3):
                step(i)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(8):
                step(i)
            e1.record()
            torch.cuda.synchronize()
            ms_e = e0.elapsed_time(e1) / 8
            del net_e, opt_e
            return {'ms_per_step': ms_e, 'value': BL / (ms_e * 1e-3), 'tflops': BL * L_FLOP_PER_SAMPLE / (ms_e * 1e-3) / 1e12}
        try:
            inc = {'fp32': eager_L(False), 'autocast_bf16': eager_L(True), 'unit': 'samples/s',
                   'note': 'stock PyTorch eager fit loop of MLP 784-4096x4-10 (Adam, batch 8192) on this GPU: the L-model incumbent'}
            mine = extras.get('config4_L_model_one_gpu', {}).get('value')
            if mine:
                inc['this_repo_over_autocast_bf16'] = mine / inc['autocast_bf16']['value']
                inc['this_repo_over_fp32'] = mine / inc['fp32']['value']
            extras['torch_eager_L_model_same_gpu'] = inc
        except Exception as e:
            extras['torch_eager_L_model_same_gpu'] = {'error': repr(e)[:300]}
        torch.cuda.empty_cache()

    # ---- e2e: host buffers, H2D every step, result read back
    if args.skip_e2e:
        e2e = None
    else:
      pin_loader = PinnedLoader(x.cpu(), y.cpu(), BATCH, device=dev, prefetch=3)
      pin_loader.h2d_bytes = 0
      e_ms, _, _ = timed(pin_loader, args.e2e_steps, 1)
      h2d_per_step = pin_loader.h2d_bytes // (args.e2e_steps + 1)
      if not args.skip_extras:
          pin16 = PinnedLoader(x.cpu(), y.cpu(), BATCH, device=dev, prefetch=3, dtype=torch.bfloat16)
          p_ms, _, _ = timed(pin16, args.e2e_steps, 1)
          extras['e2e_bf16_host_buffers'] = {'value': world * n * args.e2e_steps / (p_ms * 1e-3), 'unit': 'samples/s',
                                             'note': 'PinnedLoader(dtype=bfloat16): dataset converted once on the host, 1,576 B/sample over PCIe'}
          del pin16
          # raw uint8 images (what MNIST is on disk): ToTensor + Normalize run on the device, 784 B/sample over PCIe
          model.engine.set_input_normalization(0.1307, 0.3081)
          xu8 = torch.randint(0, 256, (n, 1, 28, 28), dtype=torch.uint8, generator=torch.Generator().manual_seed(5 + rank))
          pin8 = PinnedLoader(xu8, y.cpu(), BATCH, device=dev, prefetch=3)
          u_ms, _, _ = timed(pin8, args.e2e_steps, 1)
          extras['e2e_uint8_host_buffers'] = {'value': world * n * args.e2e_steps / (u_ms * 1e-3), 'unit': 'samples/s',
                                              'note': 'PinnedLoader over uint8 images, B200Classifier(uint8_norm=(0.1307, 0.3081)): the '
                                                      "dataset's ToTensor + Normalize applied on the device, 792 B/sample over PCIe"}
          del pin8, xu8
      ceiling = h2d_ceiling(torch, dist if world > 1 else None, dev, world)
      e2e_rate = world * n * args.e2e_steps / (e_ms * 1e-3)
      e2e = {'value': e2e_rate, 'unit': 'samples/s', 'h2d_bytes_per_step': h2d_per_step,
           'd2h_bytes_per_step': 32, 'steps': args.e2e_steps, 'ms_per_step': e_ms / args.e2e_steps,
           'h2d_gbs_per_gpu': h2d_per_step / (e_ms / args.e2e_steps * 1e-3) / 1e9,
           'h2d_ceiling_gbs_per_gpu': ceiling, 'frac_of_h2d_ceiling': h2d_per_step / (e_ms / args.e2e_steps * 1e-3) / 1e9 / ceiling,
           'ceiling_note': f'pinned host -> device copies of 256 MiB, all {world} rank(s) at once, slowest rank',
           'path': 'training.service.handle("train") over PinnedLoader (pinned host -> device on a copy stream, 3 batches ahead)'}

    # ---- cpu_baseline (rank 0, N=1 only)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        r = cpu_reference_run(1000, 1, 16, args.cpu_seconds, args.optimizer)
        cpu = {'value': r['value'], 'unit': 'samples/s', 'cores': r['cores'], 'kind': 'port', 'sample': r['sample']}

    if rank == 0 and world == 1 and not args.no_cpu_baseline and not args.skip_extras:
        # the reference's CPU path at the batch sizes of configs[0] (256) and configs[2] (32, as shipped)
        for name, b, nb in (('cpu_reference_batch256_adam', 256, 64), ('cpu_reference_batch32_adam', 32, 256)):
            try:
                r = cpu_reference_run(1000, 1, nb, 4.0, 'adam', batch=b, min_warm_seconds=1.5)
                extras[name] = {'value': r['value'], 'unit': 'samples/s', 'cores': r['cores'], 'us_per_batch': b / r['value'] * 1e6,
                                'sample': r['sample']}
            except Exception as e:
                extras[name] = {'error': repr(e)[:300]}
    if rank == 0:
        line = {
            'metric': 'train_samples_per_sec', 'value': value, 'unit': 'samples/s', 'n_gpus': world, 'steps': args.steps,
            'warmup': args.warmup, 'ms_per_step': ms / args.steps, 'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': None, 'dtype': args.precision, 'data': 'synthetic',
            'config': {'workload': f'tinysys train handler: MLP(784,256,10) dropout 0.1 ReLU, CrossEntropyLoss, {args.optimizer}, '
                                   f'batch {BATCH}, {args.batches} batches per step per GPU (BASELINE.json configs[1])',
                       'batch': BATCH, 'samples_per_step_per_gpu': n, 'optimizer': args.optimizer, 'precision': args.precision,
                       'parallelism': f'dp{world}', 'dp_mode': dp_mode if world > 1 else None, 'dp_mode_trial_us_per_batch': dp_trial, 'l2': f'inputs of one step ({n * 3136 / 1e9:.2f} GB per GPU) exceed the 126 MB L2',
                       'last_train_result': results[-1] if results else None},
            'clocks': clocks, 'e2e': e2e, 'gpu_launches': launches, 'roofline': roofline, 'cpu_baseline': cpu,
            'parity': parity, 'extras': extras,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
