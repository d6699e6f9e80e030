import sys, os, numpy as np, torch
sys.path.insert(0, '/root/repo')
from torch import nn
from tests.test_gpu_parity import build_model, seeded_init, synthetic
dims, B = [784, 256, 10], int(sys.argv[1]) if len(sys.argv) > 1 else 4096
x, y = synthetic(B); x = x.bfloat16().float()
init = seeded_init(dims, bf16=True)
for p in (0.0, 0.1):
    res = {}
    for tag, xin in (('f32', x.flatten(1).cuda()), ('bf16', x.flatten(1).bfloat16().cuda())):
        model, net = build_model(dims, p, 'relu', 'sgd', 0.0, init, 'bf16', max_batch=B)
        g = torch.Generator().manual_seed(42)
        mask = (torch.rand(B, dims[1], generator=g) >= 0.1).to(torch.uint8).cuda()
        pe, le = model.evaluate(xin, y.cuda())
        pf, lf = model.fit(xin, y.cuda(), masks=[mask] if p > 0 else None)
        grads = [q.grad.clone() for lin in net.linears() for q in (lin.weight, lin.bias)]
        res[tag] = (pe.cpu(), le.item(), pf.cpu(), lf.item(), [gg.cpu() for gg in grads])
    a, b = res['f32'], res['bf16']
    print(f'NS={os.environ.get("TSB_FUSED_NS")} p={p}: eval pred mismatch {(a[0]!=b[0]).sum().item()} loss {a[1]:.7f} vs {b[1]:.7f} | fit pred mismatch {(a[2]!=b[2]).sum().item()} loss {a[3]:.7f} vs {b[3]:.7f} | grad rel diffs',
          [f'{((ga-gb).norm()/gb.norm()).item():.2e}' for ga, gb in zip(a[4], b[4])])
