"""L model at 256 rows: per-tensor relative error of the engine's gradients against the bf16-emulating oracle (SGD, lr 0)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import mlp_numpy as onp
from tests.test_gpu_parity import build_model, seeded_init, synthetic

dims = [784] + [int(a) for a in (sys.argv[1].split(',') if len(sys.argv) > 1 else ['4096'] * 4)] + [10]
p, B = 0.1, int(sys.argv[2]) if len(sys.argv) > 2 else 256
x, y = synthetic(B, seed=31)
x = x.bfloat16().float()
init = seeded_init(dims, bf16=True)
model, net = build_model(dims, p, 'relu', 'sgd', 0.0, init, 'bf16', max_batch=B)
oracle = onp.OracleMLP(weights=init[0::2], biases=init[1::2], dropout=p, opt_kind=onp.OPT_SGD, lr=0.0, emulate_bf16=True)
g = torch.Generator().manual_seed(9)
masks = [(torch.rand(B, h, generator=g) >= p).to(torch.uint8) for h in dims[1:-1]]
pr, l = model.fit(x.cuda(), y.cuda(), masks=[m.cuda() for m in masks])
cap = {}
oracle.fit(x.numpy(), y.numpy(), [m.numpy() for m in masks], grad_hook=lambda gs: cap.setdefault('g', [q.copy() for q in gs]) or gs)
got = [q.grad.detach().cpu().numpy() for lin in net.linears() for q in (lin.weight, lin.bias)]
print('loss', l.item(), 'pred equal', np.array_equal(pr.cpu().numpy(), oracle._forward(x.numpy(), [m.numpy() for m in masks])[0].argmax(1)))
for i, (a, b) in enumerate(zip(got, cap['g'])):
    d = a.astype(np.float64) - b
    rms = np.sqrt((b.astype(np.float64) ** 2).mean())
    print(f'tensor {i} shape {a.shape}: rms(g)={rms:.3e} rms(err)={np.sqrt((d**2).mean()):.3e} rel={np.sqrt((d**2).mean())/rms:.3e} max|err|={np.abs(d).max():.3e} '
          f'sign flips {np.mean(np.sign(a) != np.sign(b)):.4f} |g|<1e-7 frac {np.mean(np.abs(b) < 1e-7):.4f}')
