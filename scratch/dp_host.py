"""S model data parallel: device time per fit against the host's enqueue time per fit (is the step host-bound?).
torchrun --nproc-per-node N scratch/dp_host.py [mode]"""
import os, sys, time, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from torch import nn
from torchsystem_b200 import B200Classifier, MLP
mode = sys.argv[1] if len(sys.argv) > 1 else 'p2p'
rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local)
dist.init_process_group('nccl', device_id=torch.device('cuda', local))
B = 4096
torch.manual_seed(0)
net = MLP(784, 256, 10, dropout=0.1, activation=nn.ReLU())
model = B200Classifier(net, nn.CrossEntropyLoss(), torch.optim.SGD(net.parameters(), lr=0.01), precision='bf16', device=f'cuda:{local}',
                       max_batch=B, dp_mode=mode)
g = torch.Generator(device='cuda').manual_seed(rank)
x = torch.randn(B * 32, 1, 28, 28, device='cuda', generator=g); y = torch.randint(0, 10, (B * 32,), device='cuda', generator=g)
xs = [x[i * B:(i + 1) * B] for i in range(32)]; ys = [y[i * B:(i + 1) * B] for i in range(32)]
for i in range(64): model.fit(xs[i & 31], ys[i & 31])
torch.cuda.synchronize(); dist.barrier()
N = 512
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter(); e0.record()
for i in range(N): model.fit(xs[i & 31], ys[i & 31])
e1.record(); t1 = time.perf_counter(); torch.cuda.synchronize()
print(f'rank {rank}/{world} {mode}: device {e0.elapsed_time(e1) / N * 1e3:.2f} us/fit, host enqueue {(t1 - t0) / N * 1e6:.2f} us/fit, cpus {os.cpu_count()}', flush=True)
dist.barrier(); dist.destroy_process_group()
