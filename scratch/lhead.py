import sys, os, torch
sys.path.insert(0, '/root/repo')
from torch import nn
from torch.profiler import profile, ProfilerActivity
from torchsystem_b200 import B200Classifier, MLP
B = 8192
torch.manual_seed(0)
net = MLP(784, [4096] * 4, 10, dropout=0.1, activation=nn.ReLU())
model = B200Classifier(net, nn.CrossEntropyLoss(), torch.optim.Adam(net.parameters(), lr=1e-3), precision='bf16', device='cuda:0', max_batch=B)
x = torch.randn(B, 784, device='cuda'); y = torch.randint(0, 10, (B,), device='cuda')
for _ in range(3): model.fit(x, y)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10): p, l = model.fit(x, y)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
print(f'L model fit (SIMT_HEAD={os.environ.get("TSB200_SIMT_HEAD")}): {ms:.3f} ms/step, {B*315080704/ms/1e9:.1f} TFLOP/s, loss {l.item():.4f}')
from torchsystem_b200 import _lib
_lib.profile_begin()
for _ in range(4): model.fit(x, y)
torch.cuda.synchronize()
pk = _lib.profile_end()
print('event-bracketed kinds (us per launch, launches per step):', {k: (round(v[0] / v[1] * 1e3, 1), v[1] // 4) for k, v in pk.items()})
print('sum per step: %.1f us' % (sum(v[0] for v in pk.values()) / 4 * 1e3))
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(3): model.fit(x, y)
    torch.cuda.synchronize()
import collections
agg = collections.defaultdict(list)
for e in prof.events():
    if e.device_type.name == 'CUDA': agg[e.name[:60]].append(e.device_time)
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1]))[:4]:
    print(f'{k:60s} n={len(v):3d} avg={sum(v)/len(v):9.1f} us total/step={sum(v)/3:9.1f}')
evs = sorted([e for e in prof.events() if e.device_type.name == 'CUDA'], key=lambda e: e.time_range.start)
n = len(evs) // 3
print('one step, in launch order:')
for e in evs[n:2 * n]:
    print(f'   {e.name[:70]:70s} {e.device_time:8.1f} us')
