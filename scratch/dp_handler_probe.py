"""Where a data-parallel train epoch spends its time on N GPUs (S model, batch 4096 per GPU): plain fit loop (host enqueue time and
time with the final sync), the train handler per batch at two epoch lengths (a per-epoch cost shows as a difference), and the
per-epoch tail alone. torchrun --nproc-per-node N scratch/dp_handler_probe.py"""
import os, sys, time, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from torch import nn
from torchsystem_b200 import B200Classifier, B200Metrics, MLP, DeviceLoader, training
rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local)
dev = torch.device('cuda', local)
dist.init_process_group('nccl', device_id=dev)
B = 4096
torch.manual_seed(0)
net = MLP(784, 256, 10, dropout=0.1, activation=nn.ReLU())
model = B200Classifier(net, nn.CrossEntropyLoss(), torch.optim.SGD(net.parameters(), lr=0.01), precision='bf16', device=dev, max_batch=B,
                       dp_mode='p2p' if world > 1 else None) if world > 1 else \
    B200Classifier(net, nn.CrossEntropyLoss(), torch.optim.SGD(net.parameters(), lr=0.01), precision='bf16', device=dev, max_batch=B)
metrics = B200Metrics(device=dev, log_every=0)
training.provider.override(training.device, lambda: dev)
g = torch.Generator(device=dev).manual_seed(rank)
x = torch.randn(B * 256, 1, 28, 28, device=dev, generator=g); y = torch.randint(0, 10, (B * 256,), device=dev, generator=g)
def report(name, host_s, total_s, n):
    t = torch.tensor([host_s, total_s], dtype=torch.float64, device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0: print(f'{name}: host {t[0].item() / n * 1e6:.1f} us/batch, with sync {t[1].item() / n * 1e6:.1f} us/batch (max over {world} ranks)', flush=True)
# (a) plain fit loop
for i in range(64): model.fit(x[i * B:(i + 1) * B], y[i * B:(i + 1) * B])
torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
t0 = time.perf_counter()
for i in range(256): model.fit(x[i * B:(i + 1) * B], y[i * B:(i + 1) * B])
t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
report('fit loop', t1 - t0, t2 - t0, 256)
# (b) handler, 256 and 64 batches per epoch
for nb in (256, 64):
    loader = DeviceLoader(x[:nb * B], y[:nb * B], B, device=dev)
    for _ in range(2): training.service.handle('train', model, loader, metrics); metrics.reset()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    reps = 4 * 256 // nb
    t0 = time.perf_counter()
    for _ in range(reps): training.service.handle('train', model, loader, metrics); metrics.reset()
    t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
    report(f'train handler, {nb} batches per epoch', t1 - t0, t2 - t0, reps * nb)
# (c) the epoch tail alone: compute (counter all-reduce + read) + reset
torch.cuda.synchronize(); dist.barrier()
t0 = time.perf_counter()
for _ in range(20): metrics.compute(); metrics.reset()
torch.cuda.synchronize(); t1 = time.perf_counter()
if rank == 0: print(f'metrics.compute + reset: {(t1 - t0) / 20 * 1e6:.0f} us per epoch', flush=True)
dist.destroy_process_group()
