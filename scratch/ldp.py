"""Config 4: L model data-parallel, per-GPU batch = 65536 / world (or argv[2]). torchrun --nproc-per-node N scratch/ldp.py [mode] [rows]"""
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from torch import nn
from torchsystem_b200 import B200Classifier, MLP
mode = sys.argv[1] if len(sys.argv) > 1 else 'nccl'
rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local)
dist.init_process_group('nccl', device_id=torch.device('cuda', local))
B = int(sys.argv[2]) if len(sys.argv) > 2 else 65536 // world
torch.manual_seed(0)
net = MLP(784, [4096] * 4, 10, dropout=0.1, activation=nn.ReLU())
model = B200Classifier(net, nn.CrossEntropyLoss(), torch.optim.Adam(net.parameters(), lr=1e-3), precision='bf16', device=f'cuda:{local}',
                       max_batch=B, dp_mode=mode)
g = torch.Generator(device='cuda').manual_seed(rank)
x = torch.randn(B, 784, device='cuda', generator=g); y = torch.randint(0, 10, (B,), device='cuda', generator=g)
for _ in range(3): model.fit(x, y)
torch.cuda.synchronize(); dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10): p, l = model.fit(x, y)
e1.record(); torch.cuda.synchronize()
ms = torch.tensor([e0.elapsed_time(e1) / 10], device='cuda'); dist.all_reduce(ms, op=dist.ReduceOp.MAX)
if rank == 0:
    print(f'L model dp{world} ({mode}): per-GPU batch {B}, {ms.item():.3f} ms/step, {B * world / ms.item() * 1e3 / 1e6:.2f} M samples/s global, '
          f'{B * 315080704 / ms.item() / 1e9:.1f} TFLOP/s per GPU, overlap SMs {os.environ.get("TSB200_NVLS_OVERLAP_SMS", "16")}, loss {l.item():.4f}', flush=True)
dist.destroy_process_group()
