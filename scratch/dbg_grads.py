import sys, numpy as np, torch
sys.path.insert(0, '/root/repo')
from torch import nn
from oracle import mlp_numpy as onp
from tests.test_gpu_parity import build_model, seeded_init, synthetic
dims, p, B = [784, 256, 10], 0.1, 4096
x, y = synthetic(B); x = x.bfloat16().float()
init = seeded_init(dims, bf16=True)
model, net = build_model(dims, p, 'relu', 'sgd', 0.0, init, 'bf16', max_batch=B)
g = torch.Generator().manual_seed(42)
mask = (torch.rand(B, dims[1], generator=g) >= p).to(torch.uint8)
model.fit(x.cuda(), y.cuda(), masks=[mask.cuda()])
oracle = onp.OracleMLP(weights=init[0::2], biases=init[1::2], dropout=p, opt_kind=onp.OPT_SGD, lr=1.0, emulate_bf16=True)
ref = [w.copy() for w in oracle.parameters()]
oracle.fit(x.numpy(), y.numpy(), [mask.numpy()])
grads = [r - w for r, w in zip(ref, oracle.parameters())]
got = [q.grad.cpu().numpy() for lin in net.linears() for q in (lin.weight, lin.bias)]
for name, a, b in zip(['dW1', 'db1', 'dW2', 'db2'], got, grads):
    err = np.abs(a - b); print(name, 'max abs err', err.max(), 'scale', np.abs(b).max(), 'rel L2', np.linalg.norm(a - b) / np.linalg.norm(b))
    if name in ('db1', 'db2'): print('   got', a[:6], '\n   ref', b[:6])
    if name == 'dW2': print('   got', a[0, :6], '\n   ref', b[0, :6]); print('   got', a[9, 250:], '\n   ref', b[9, 250:])
