import os, torch, torch.distributed as dist
rank = int(os.environ['RANK']); world = int(os.environ['WORLD_SIZE'])
torch.cuda.set_device(rank)
dist.init_process_group('nccl', device_id=torch.device('cuda', rank))
try:
    import torch.distributed._symmetric_memory as symm
    t = symm.empty(1024, dtype=torch.float32, device=f'cuda:{rank}')
    h = symm.rendezvous(t, dist.group.WORLD)
    print(rank, 'symm ok', type(h).__name__, [hex(p) for p in h.buffer_ptrs], 'signal', [hex(p) for p in h.signal_pad_ptrs], 'pad size', h.signal_pad_size, flush=True)
    t.fill_(rank + 1); dist.barrier(); torch.cuda.synchronize()
    peer = h.get_buffer((rank + 1) % world, (1024,), torch.float32)
    print(rank, 'peer value', peer[0].item(), flush=True)
except Exception as e:
    import traceback; traceback.print_exc()
dist.destroy_process_group()
