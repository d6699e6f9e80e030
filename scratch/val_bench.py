import sys, torch
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
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

import ctypes as C
from torchsystem_b200 import _lib
model.evaluate(x[:B], y[:B])
out = (C.c_uint64 * 32)()
_lib.check(_lib.lib.ts_test_eval_trace(out), 'trace'); t = list(out)
t0 = t[0]
print('MMA issuer (slab 3): acc_empty wait', t[1] - t0, '| kb0 x_full', t[2] - t0, 'w_full', t[3] - t0, '| kb1', t[4] - t0, t[5] - t0, '| kb2', t[6] - t0, t[7] - t0,
      '| kb3', t[8] - t0, t[9] - t0, '| last kb committed', t[10] - t0)
print('epilogue: start', t[11] - t0, 'acc_full', t[12] - t0, 'atoms done', t[13] - t0, 'z_full', t[14] - t0, 'slab done', t[15] - t0)
print('converter kb0..3 (start, xf_full, x_empty):', [(t[16 + 3 * i] - t0, t[17 + 3 * i] - t0, t[18 + 3 * i] - t0) for i in range(4)])
