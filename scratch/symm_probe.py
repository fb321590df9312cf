import os, torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem
rank=int(os.environ['RANK']); world=int(os.environ['WORLD_SIZE']); lr=int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(lr)
dist.init_process_group('nccl', device_id=torch.device('cuda', lr))
try:
    t = symm_mem.empty(1024, dtype=torch.int64, device=torch.device('cuda', lr))
    t.fill_(rank)
    hdl = symm_mem.rendezvous(t, dist.group.WORLD)
    print(rank, 'ptrs', [hex(p) for p in hdl.buffer_ptrs], 'sig', [hex(p) for p in hdl.signal_pad_ptrs][:2], type(hdl))
    print([n for n in dir(hdl) if not n.startswith('_')])
    hdl.barrier(channel=0)
    peer = hdl.get_buffer((rank+1)%world, (1024,), torch.int64)
    print(rank, 'peer value', int(peer[0]))
    hdl.barrier(channel=0)
except Exception as e:
    import traceback; traceback.print_exc()
dist.destroy_process_group()
