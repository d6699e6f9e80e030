import os, torch, torch.distributed as dist
rank = int(os.environ['RANK']); world = int(os.environ['WORLD_SIZE'])
torch.cuda.set_device(rank)
dist.init_process_group('nccl', device_id=torch.device('cuda', rank))
import torch.distributed._symmetric_memory as symm
t = symm.empty(1 << 20, dtype=torch.float32, device=f'cuda:{rank}')
h = symm.rendezvous(t, dist.group.WORLD)
print(rank, 'multicast_ptr', hex(getattr(h, 'multicast_ptr', 0) or 0), 'has multimem ops', hasattr(torch.ops.symm_mem, 'multimem_all_reduce_'), flush=True)
try:
    t.fill_(rank + 1.0); dist.barrier(); torch.cuda.synchronize()
    torch.ops.symm_mem.multimem_all_reduce_(t, 'sum', dist.group.WORLD.group_name)
    torch.cuda.synchronize()
    print(rank, 'multimem_all_reduce_ ->', t[0].item(), flush=True)
except Exception as e:
    print(rank, 'multimem_all_reduce_ failed:', repr(e)[:200], flush=True)
dist.destroy_process_group()
