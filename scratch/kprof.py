import sys, time, torch
sys.path.insert(0, '/root/repo')
from torch import nn
from torch.profiler import profile, ProfilerActivity
from torchsystem_b200 import B200Classifier, MLP
B = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
torch.manual_seed(0)
opt_kind = sys.argv[1] if len(sys.argv) > 1 else 'sgd'
net = MLP(784, 256, 10, dropout=0.1, activation=nn.ReLU())
opt = torch.optim.SGD(net.parameters(), lr=0.01) if opt_kind == 'sgd' else torch.optim.Adam(net.parameters(), lr=1e-3)
model = B200Classifier(net, nn.CrossEntropyLoss(), opt, precision='bf16', device='cuda:0', max_batch=B)
x = torch.randn(B * 16, 1, 28, 28, device='cuda'); y = torch.randint(0, 10, (B * 16,), device='cuda')
for i in range(32): model.fit(x[(i % 16) * B:(i % 16 + 1) * B], y[(i % 16) * B:(i % 16 + 1) * B])
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for i in range(64): model.fit(x[(i % 16) * B:(i % 16 + 1) * B], y[(i % 16) * B:(i % 16 + 1) * B])
    for i in range(64): model.evaluate(x[(i % 16) * B:(i % 16 + 1) * B], y[(i % 16) * B:(i % 16 + 1) * B])
    torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type.name == 'CUDA']
import collections
agg = collections.defaultdict(list)
for e in evs: agg[e.name[:60]].append(e.device_time if hasattr(e, 'device_time') else e.cuda_time)
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
    print(f'{k:60s} n={len(v):4d} avg={sum(v)/len(v):7.2f} us')
ts = sorted((e.time_range.start, e.time_range.end, e.name[:30]) for e in evs)
# steady-state timeline of one fit (kernels 5 per fit): print gaps for one step in the middle
mid = [t for t in ts][5*30:5*30+6]
t0 = mid[0][0]
print('timeline (us since first kernel of a fit):', [(n, round(s - t0, 1), round(e - t0, 1)) for s, e, n in mid])
