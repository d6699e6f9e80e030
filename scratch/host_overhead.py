import sys, time, torch
sys.path.insert(0, '/root/repo')
from torch import nn
from torchsystem_b200 import B200Classifier, B200Metrics, MLP, DeviceLoader, training, _lib
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
torch.manual_seed(0)
net = MLP(784, 256, 10, dropout=0.1, activation=nn.ReLU())
model = B200Classifier(net, nn.CrossEntropyLoss(), torch.optim.SGD(net.parameters(), lr=0.01), precision='bf16', device='cuda:0', max_batch=B)
metrics = B200Metrics(device='cuda:0', log_every=0)
x = torch.randn(B * 64, 1, 28, 28, device='cuda'); y = torch.randint(0, 10, (B * 64,), device='cuda')
loader = DeviceLoader(x, y, B, device='cuda:0')
training.provider.override(training.device, lambda: torch.device('cuda', 0))
for _ in range(3):
    training.service.handle('train', model, loader, metrics); metrics.reset()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(4):
    training.service.handle('train', model, loader, metrics); metrics.reset()
t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
print(f'handler loop: host {1e6*(t1-t0)/256:.1f} us/batch, with sync {1e6*(t2-t0)/256:.1f} us/batch')
# raw fit calls only
xb, yb = x[:B], y[:B]
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(256): model.fit(xb, yb)
t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
print(f'fit only: host {1e6*(t1-t0)/256:.1f} us/call, with sync {1e6*(t2-t0)/256:.1f} us/call')
import cProfile, pstats
pr = cProfile.Profile(); pr.enable()
for _ in range(256): model.fit(xb, yb)
pr.disable(); torch.cuda.synchronize()
pstats.Stats(pr).sort_stats('cumulative').print_stats(12)
