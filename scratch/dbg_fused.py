import sys, torch, ctypes as C
sys.path.insert(0, '/root/repo')
from torch import nn
from torchsystem_b200 import B200Classifier, MLP, _lib
def trace(tag):
    buf = (C.c_uint64 * 32)()
    _lib.check(_lib.lib.ts_test_fused_trace(buf))
    d = list(buf); t0 = d[0]
    print(tag, 'CTA1:', {k: d[16+k]-d[16] for k in (1,2,3,4,5,8,14,15,9,12,13) if d[16+k]}, 'start skew', d[16]-d[0])
    names = {14:'z_sent',15:'cluster_waited',1:'setup',2:'first_mma',3:'last_mma_issued',4:'h_full_seen',5:'A_written',7:'A_copied',8:'z_full_seen',6:'z_mma_issued',9:'dz_written',10:'da_issued',11:'da_seen',12:'dH_done',13:'end'}
    print(tag, {names[i]: d[i]-t0 for i in sorted(names) if d[i] >= t0 and d[i]})
for p, B in ((0.1, 4096), (0.0, 4096)):
    torch.manual_seed(0)
    net = MLP(784, 256, 10, dropout=p, activation=nn.ReLU())
    model = B200Classifier(net, nn.CrossEntropyLoss(), torch.optim.SGD(net.parameters(), lr=0.01), precision='bf16', device='cuda:0', max_batch=B)
    x = torch.randn(B, 784, device='cuda'); y = torch.randint(0, 10, (B,), device='cuda')
    for _ in range(3): model.fit(x, y)
    trace(f'fit p={p}')
    for _ in range(3): model.evaluate(x, y)
    trace(f'eval p={p}')
