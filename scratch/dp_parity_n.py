"""bench.py's dp_parity (three data-parallel steps of the S model, 2,048 rows per rank, against rank 0 alone on the full batch;
replicas bit-compared) at N GPUs, with the low-latency packet exchange and with publish + remote reads, and a bit-comparison of
the two results. torchrun --nproc-per-node N scratch/dp_parity_n.py"""
import os, sys, json, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local)
dev = torch.device('cuda', local)
dist.init_process_group('nccl', device_id=dev)
from torch import nn
from torchsystem_b200 import B200Classifier, MLP

def weights_after(steps=4, per=2048):
    g = torch.Generator(device=dev).manual_seed(99)
    x = torch.randn(steps * per * world, 784, generator=g, device=dev); y = torch.randint(0, 10, (steps * per * world,), generator=g, device=dev)
    torch.manual_seed(3)
    net = MLP(784, 256, 10, dropout=0.1, activation=nn.ReLU())
    m = B200Classifier(net, nn.CrossEntropyLoss(), torch.optim.Adam(net.parameters(), lr=1e-3), precision='bf16', device=dev, max_batch=per, dp_mode='p2p')
    for s in range(steps):
        lo = (s * world + rank) * per
        m.fit(x[lo:lo + per], y[lo:lo + per])
    torch.cuda.synchronize()
    return m.engine.params.clone()

out = {}
for name, env in (('packets', '1'), ('publish_read', '0')):
    os.environ['TSB200_DP_LL'] = env
    res = bench.dp_parity(torch, dist, dev, world, rank, 'bf16', 'p2p')
    w = weights_after()
    out[name] = (res, w)
same = torch.equal(out['packets'][1], out['publish_read'][1])
t = torch.tensor([int(same)], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MIN)
if rank == 0:
    for name in out:
        r = out[name][0]
        print(name, json.dumps({k: r[k] for k in ('ok', 'replicas_bit_identical', 'weights_rel_err', 'loss_rel_err', 'dp_correct', 'full_batch_correct')}), flush=True)
    print(f'Adam + dropout, 4 steps: weights bit-identical between the two exchanges on all {world} ranks: {bool(t.item())}', flush=True)
dist.destroy_process_group()
