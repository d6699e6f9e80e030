"""L model (784 -> 4096 x4 -> 10, Adam, bf16) at 8,192 rows on one GPU: step time, then per-kernel-kind device time through the
library's own event brackets (serialises the kernels: shares, not overlap). python scratch/lprof.py [rows]"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from torch import nn
from torchsystem_b200 import B200Classifier, MLP, _lib
B = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
torch.manual_seed(0)
DROP = float(os.environ.get("LPROF_DROPOUT", "0.1"))
net = MLP(784, [4096] * 4, 10, dropout=DROP, activation=nn.ReLU())
model = B200Classifier(net, nn.CrossEntropyLoss(), torch.optim.Adam(net.parameters(), lr=1e-3), precision='bf16', device='cuda:0', max_batch=B)
x = torch.randn(B, 784, device='cuda'); y = torch.randint(0, 10, (B,), device='cuda')
for _ in range(3): model.fit(x, y)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10): p, l = model.fit(x, y)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
print(f'L model fit: B={B} {ms:.3f} ms/step, {B*315080704/ms/1e9:.1f} TFLOP/s, loss {l.item():.4f}  env: '
      + ' '.join(f'{k}={v}' for k, v in os.environ.items() if k.startswith('TSB')))
_lib.profile_begin()
for _ in range(5): model.fit(x, y)
prof = _lib.profile_end()
tot = 0.0
for k, (t, n) in sorted(prof.items(), key=lambda kv: -kv[1][0]):
    print(f'  {k:24s} {n/5:5.1f} launches/step  {t/5*1e3:8.1f} us/step  {t/n*1e3:8.1f} us each'); tot += t / 5
print(f'  sum {tot*1e3:.1f} us/step')
