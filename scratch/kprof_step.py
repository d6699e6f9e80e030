"""Single-kernel small-batch step: device time per fit (CUDA events), host time per fit, and the in-kernel clock trace."""
import ctypes as C, os, sys, time, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from torch import nn
from torchsystem_b200 import B200Classifier, MLP, _lib
B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
opt_kind = sys.argv[2] if len(sys.argv) > 2 else 'adam'
torch.manual_seed(0)
net = MLP(784, 256, 10, dropout=0.1, activation=nn.ReLU())
opt = torch.optim.SGD(net.parameters(), lr=0.01) if opt_kind == 'sgd' else torch.optim.Adam(net.parameters(), lr=1e-3)
model = B200Classifier(net, nn.CrossEntropyLoss(), opt, precision='bf16', device='cuda:0', max_batch=64)
x = torch.randn(B * 64, 1, 28, 28, device='cuda'); y = torch.randint(0, 10, (B * 64,), device='cuda')
xs = [x[i * B:(i + 1) * B] for i in range(64)]; ys = [y[i * B:(i + 1) * B] for i in range(64)]
for i in range(64): model.fit(xs[i], ys[i])
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
N = 2000
t0 = time.perf_counter(); e0.record()
for i in range(N): model.fit(xs[i & 63], ys[i & 63])
e1.record(); t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
print(f'B={B} {opt_kind} step_kernel={os.environ.get("TSB200_STEP_KERNEL", "1")}: device {e0.elapsed_time(e1) / N * 1e3:.2f} us/fit, host enqueue {(t1 - t0) / N * 1e6:.2f} us/fit, '
      f'wall {(t2 - t0) / N * 1e6:.2f} us/fit')
if os.environ.get('TSB200_STEP_KERNEL', '1') != '0':
    out = (C.c_uint64 * 32)()
    _lib.check(_lib.lib.ts_test_step_trace(out), 'trace')
    t = list(out)
    names = {0: 'start', 1: 'after pdl_wait', 2: 'first MMA issued', 3: 'main loop issued', 4: 'X converted', 5: 'H ready', 6: 'logits exchanged',
             7: 'dA ready', 10: 'all MMAs issued', 11: 'gradients ready', 12: 'update done', 13: 'end'}
    print('trace (cycles since start): ' + ' | '.join(f'{names[i]} {t[i] - t[0]}' for i in sorted(names)))
