"""S model (784-256-10, SGD, dropout 0.1, bf16) on one GPU: device time per fit of a plain fit loop over 256 resident fp32 batches.
python scratch/sstep.py [batch]"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from torch import nn
from torchsystem_b200 import B200Classifier, MLP
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
torch.manual_seed(0)
net = MLP(784, 256, 10, dropout=0.1, activation=nn.ReLU())
model = B200Classifier(net, nn.CrossEntropyLoss(), torch.optim.SGD(net.parameters(), lr=0.01), precision='bf16', device='cuda:0', max_batch=B)
n = 256
x = torch.randn(B * n, 1, 28, 28, device='cuda'); y = torch.randint(0, 10, (B * n,), device='cuda')
xs = [(x[i * B:(i + 1) * B], y[i * B:(i + 1) * B]) for i in range(n)]
for a, b in xs[:64]: model.fit(a, b)
torch.cuda.synchronize()
best = 1e9
for rep in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for a, b in xs: p, l = model.fit(a, b)
    e1.record(); torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1) / n * 1e3)
print(f'S step, batch {B}: {best:.2f} us per fit ({B / best:.1f} M samples/s), loss {l.item():.4f}  env: '
      + ' '.join(f'{k}={v}' for k, v in os.environ.items() if k.startswith('TSB')), flush=True)
