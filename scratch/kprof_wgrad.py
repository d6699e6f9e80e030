"""Fused training path at batch 4096: device time per fit and the in-kernel clock traces of the fused kernel and of the dW1 + update kernel."""
import ctypes as C, os, sys, time, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from torch import nn
from torchsystem_b200 import B200Classifier, MLP, _lib
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
opt_kind = sys.argv[2] if len(sys.argv) > 2 else 'sgd'
torch.manual_seed(0)
net = MLP(784, 256, 10, dropout=0.1, activation=nn.ReLU())
opt = torch.optim.SGD(net.parameters(), lr=0.01) if opt_kind == 'sgd' else torch.optim.Adam(net.parameters(), lr=1e-3)
model = B200Classifier(net, nn.CrossEntropyLoss(), opt, precision='bf16', device='cuda:0', max_batch=B)
x = torch.randn(B * 32, 1, 28, 28, device='cuda'); y = torch.randint(0, 10, (B * 32,), device='cuda')
xs = [x[i * B:(i + 1) * B] for i in range(32)]; ys = [y[i * B:(i + 1) * B] for i in range(32)]
for i in range(64): model.fit(xs[i & 31], ys[i & 31])
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
N = 1024
t0 = time.perf_counter(); e0.record()
for i in range(N): model.fit(xs[i & 31], ys[i & 31])
e1.record(); t1 = time.perf_counter(); torch.cuda.synchronize()
print(f'B={B} {opt_kind}: device {e0.elapsed_time(e1) / N * 1e3:.2f} us/fit, host enqueue {(t1 - t0) / N * 1e6:.2f} us/fit')
out = (C.c_uint64 * 32)()
_lib.check(_lib.lib.ts_test_fused_trace(out), 'trace'); t = list(out)
print('fused trace (cycles):', [t[i] - t[0] for i in range(16)])
_lib.check(_lib.lib.ts_test_wgrad_trace(out), 'trace'); t = list(out)
names = ['start', 'after pdl_wait', 'first MMA', 'last MMA issued', 'acc ready', 'staged', 'cluster barrier', 'remote loads', 'updated', 'end']
print('wgrad trace (cycles): ' + ' | '.join(f'{n} {t[i] - t[0]}' for i, n in enumerate(names)))
