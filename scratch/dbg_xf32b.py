import sys, os, numpy as np, torch
sys.path.insert(0, '/root/repo')
from tests.test_gpu_parity import build_model, seeded_init, synthetic
dims, B = [784, 256, 10], 4096
x, y = synthetic(B); x = x.bfloat16().float()
init = seeded_init(dims, bf16=True)
model, net = build_model(dims, 0.0, 'relu', 'sgd', 0.0, init, 'bf16', max_batch=B)
xin = x.flatten(1).cuda()
model.fit(xin, y.cuda())
torch.cuda.synchronize()
eng = model.engine
# x_bf16 lives in the workspace; find it by scanning for the converted copy: compare through a second conversion
want = xin.bfloat16()
ws = eng.workspace
# locate: try all 256-byte aligned offsets cheaply via first-row match
wv = ws.view(torch.bfloat16)
first = want[0, :16]
cand = (wv.unfold(0, 16, 128)[:, :] == first).all(dim=1).nonzero().flatten()
print('candidates', cand[:4].tolist())
off = cand[0].item() * 128
got = wv[off:off + B * 784].view(B, 784)
bad = (got != want)
print('x_bf16 mismatching elements', bad.sum().item(), 'rows with mismatch', bad.any(dim=1).sum().item())
if bad.any():
    r, c = bad.nonzero()[0].tolist(); print('first bad at', r, c, got[r, c].item(), want[r, c].item(), 'bad cols hist (by 64):', torch.bincount(bad.nonzero()[:, 1] // 64, minlength=13).tolist())
idx = bad.nonzero()[:400]
import collections
shift = collections.Counter()
for r, c in idx.tolist():
    v = got[r, c]
    m = (want[r] == v).nonzero().flatten().tolist()
    m2 = (want[:, c] == v).nonzero().flatten().tolist()
    shift[('samerow_dcol', tuple(x_ - c for x_ in m[:2]))] += 1 if m else 0
    if not m: shift[('samecol_drow', tuple(x_ - r for x_ in m2[:2]))] += 1
print(shift.most_common(8))
rows_bad = bad.any(dim=1).nonzero().flatten().tolist(); print('bad rows (first 30):', rows_bad[:30])
