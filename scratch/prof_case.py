"""Small deterministic case for ncu / compute-sanitizer: every kernel family of the library once or a few times.
    python scratch/prof_case.py train      S model, batch 4096: cast -> fused -> dW1+update, 12 fits, 2 evaluates
    python scratch/prof_case.py validate   S model, evaluate at batch 65,536 (fp32 inputs converted on chip), 6 calls
    python scratch/prof_case.py small      S model, batch 32 (single-kernel step), 8 fits; L-shaped net 784-512x2-10 at 512 rows, 3 fits
"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from torch import nn
from torchsystem_b200 import B200Classifier, MLP
what = sys.argv[1] if len(sys.argv) > 1 else 'train'
torch.manual_seed(0)


def model(dims, opt, B):
    net = MLP(dims[0], dims[1] if len(dims) == 3 else dims[1:-1], dims[-1], dropout=0.1, activation=nn.ReLU())
    o = torch.optim.SGD(net.parameters(), lr=0.01) if opt == 'sgd' else torch.optim.Adam(net.parameters(), lr=1e-3)
    return B200Classifier(net, nn.CrossEntropyLoss(), o, precision='bf16', device='cuda:0', max_batch=B)


def data(B, n):
    g = torch.Generator(device='cuda').manual_seed(1)
    return torch.randn(B * n, 1, 28, 28, device='cuda', generator=g), torch.randint(0, 10, (B * n,), device='cuda', generator=g)


if what == 'train':
    B = 4096
    m = model([784, 256, 10], 'sgd', B)
    x, y = data(B, 4)
    for i in range(12):
        p, l = m.fit(x[(i % 4) * B:(i % 4 + 1) * B], y[(i % 4) * B:(i % 4 + 1) * B])
    for i in range(2):
        p, l = m.evaluate(x[i * B:(i + 1) * B], y[i * B:(i + 1) * B])
elif what == 'validate':
    B = 65536
    m = model([784, 256, 10], 'sgd', B)
    x, y = data(B, 2)
    for i in range(6):
        p, l = m.evaluate(x[(i % 2) * B:(i % 2 + 1) * B], y[(i % 2) * B:(i % 2 + 1) * B])
else:
    m = model([784, 256, 10], 'adam', 64)
    x, y = data(32, 8)
    for i in range(8):
        p, l = m.fit(x[i * 32:(i + 1) * 32], y[i * 32:(i + 1) * 32])
    m2 = model([784, 512, 512, 10], 'adam', 512)
    x2, y2 = data(512, 1)
    for i in range(3):
        p, l = m2.fit(x2, y2)
torch.cuda.synchronize()
print(what, 'ok: loss', float(l.item()))
