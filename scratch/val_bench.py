import sys, torch
sys.path.insert(0, '/root/repo')
from torch import nn
from torchsystem_b200 import B200Classifier, MLP
B = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
torch.manual_seed(0)
net = MLP(784, 256, 10, dropout=0.1, activation=nn.ReLU())
model = B200Classifier(net, nn.CrossEntropyLoss(), torch.optim.SGD(net.parameters(), lr=0.01), precision='bf16', device='cuda:0', max_batch=B)
nb = 8
x = torch.randn(B * nb, 784, device='cuda'); y = torch.randint(0, 10, (B * nb,), device='cuda')
def run(fn, iters):
    for i in range(4): fn(x[(i % nb) * B:(i % nb + 1) * B], y[(i % nb) * B:(i % nb + 1) * B])
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for i in range(iters): fn(x[(i % nb) * B:(i % nb + 1) * B], y[(i % nb) * B:(i % nb + 1) * B])
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3
for name, fn in (('evaluate', model.evaluate), ('fit', model.fit)):
    us = run(fn, 40)
    print(f'{name} B={B}: {us:.1f} us/batch, {B / us:.1f} M samples/s, fp32-stream {B * 3152 / us / 1e3:.0f} GB/s')
