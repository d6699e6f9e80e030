import os, sys, torch, torch.distributed as dist, collections
sys.path.insert(0, '/root/repo')
from torch import nn
from torch.profiler import profile, ProfilerActivity
rank = int(os.environ['RANK']); world = int(os.environ['WORLD_SIZE'])
torch.cuda.set_device(rank)
dist.init_process_group('nccl', device_id=torch.device('cuda', rank))
from torchsystem_b200 import B200Classifier, MLP
B = 4096
mode = sys.argv[1] if len(sys.argv) > 1 else 'p2p'
torch.manual_seed(0)
net = MLP(784, 256, 10, dropout=0.1, activation=nn.ReLU())
model = B200Classifier(net, nn.CrossEntropyLoss(), torch.optim.SGD(net.parameters(), lr=0.01), precision='bf16', device=f'cuda:{rank}', max_batch=B, dp_mode=mode)
x = torch.randn(B * 16, 1, 28, 28, device='cuda'); y = torch.randint(0, 10, (B * 16,), device='cuda')
def step(i): model.fit(x[(i % 16) * B:(i % 16 + 1) * B], y[(i % 16) * B:(i % 16 + 1) * B])
for i in range(64): step(i)
torch.cuda.synchronize(); dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(256): step(i)
e1.record(); torch.cuda.synchronize()
if rank == 0: print(f'{mode}: {e0.elapsed_time(e1) / 256 * 1e3:.1f} us per fit (world {world})', flush=True)
if model._p2p is not None:
    tr = model._p2p[1][16:24].cpu().tolist(); tr = [(v - tr[0]) & 0xffffffff for v in tr]
    print(f'rank {rank} dp kernel trace ns: pdl_wait_done {tr[1]}, flags_seen {tr[2]}, peers_read {tr[3]}, updated {tr[4]} | part1_done {tr[5]} last_cta_ticket {tr[6]} flags_sent {tr[7]}', flush=True)
import ctypes
from torchsystem_b200 import _lib
tr = (ctypes.c_uint64 * 16)()
if _lib.lib.ts_test_wgrad_trace(tr) == 0:
    t = [int(v) for v in tr]
    print(f'rank {rank} dW1+update CTA 0 (cycles after griddepcontrol.wait): last MMA issued {t[3]-t[1]}, tile staged {t[5]-t[1]}, cluster barrier {t[6]-t[1]}, '
          f'remote smem loads issued {t[7]-t[1]}, own gradient summed {t[10]-t[1]}, exchanged+updated {t[8]-t[1]}, end {t[9]-t[1]}', flush=True)
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for i in range(64): step(i)
    torch.cuda.synchronize()
if rank == 0:
    evs = [e for e in prof.events() if e.device_type.name == 'CUDA']
    agg = collections.defaultdict(list)
    for e in evs: agg[e.name[:60]].append(e.device_time)
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        print(f'{k:60s} n={len(v):4d} avg={sum(v)/len(v):7.2f} us')
    ts = sorted((e.time_range.start, e.time_range.end, e.name[:28]) for e in evs)
    n = len(ts) // 64
    mid = ts[n * 30:n * 30 + n + 1]; t0 = mid[0][0]
    print('timeline:', [(nm, round(s - t0, 1), round(e - t0, 1)) for s, e, nm in mid], flush=True)
dist.barrier(); dist.destroy_process_group()
