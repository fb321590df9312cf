import os, sys, time, datetime as dt
sys.path.insert(0, os.getcwd())
import torch, torch.distributed as dist
import ctypes as C, numpy as np
from smsgate_b200 import synth, _lib as L
from smsgate_b200.batch import Batch
from smsgate_b200.replay import ReplayEngine
from smsgate_b200.dist import GlobalDedupP2P
rank=int(os.environ['RANK']); world=int(os.environ['WORLD_SIZE']); lr=int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(lr)
dist.init_process_group('nccl', device_id=torch.device('cuda', lr))
NOW=dt.datetime(2026,1,1); n=1_000_000
backup=synth.generate(n, seed=100+rank)
b=Batch.from_strings(backup.bodies, backup.senders, backup.date_ms)
eng=ReplayEngine(lr); eng.load_cache(backup.cache)
d_bytes=torch.from_numpy(b.body_bytes.copy()).cuda(); d_off=torch.from_numpy(b.body_offsets.copy()).cuda()
cols=eng.device_columns(n, d_bytes.numel()); eng.reserve(n, d_bytes.numel())
ex=GlobalDedupP2P(eng, dist, n)
lib=L.lib()
def ev():
    e=torch.cuda.Event(enable_timing=True); e.record(); return e
for it in range(8):
    eng.replay_device(d_bytes, d_off, NOW, cols, base_index=rank*n)
    out=eng._dev_outputs(cols); parity=ex.step&1
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    ea=ev()
    L.check(eng._ctx, lib.smsgate_p2p_filter_push(eng._ctx, n, C.byref(out), C.byref(ex.desc)))
    eb=ev()
    ex._h[1].barrier(channel=0)
    ec=ev()
    L.check(eng._ctx, lib.smsgate_p2p_filter_reduce(eng._ctx, C.byref(ex.desc)))
    ed=ev()
    ex._h[1].barrier(channel=1)
    e0=ev()
    L.check(eng._ctx, lib.smsgate_p2p_scatter(eng._ctx, n, rank*n, parity, C.byref(out), C.byref(ex.desc)))
    e1=ev()
    ex._h[1].barrier(channel=0)
    e2=ev()
    L.check(eng._ctx, lib.smsgate_p2p_merge(eng._ctx, parity, C.byref(ex.desc)))
    e3=ev()
    ex._h[1].barrier(channel=1)
    e4=ev()
    L.check(eng._ctx, lib.smsgate_p2p_resolve(eng._ctx, n, rank*n, C.byref(out), C.byref(ex.desc), None))
    e5=ev()
    ex.step+=1
    torch.cuda.synchronize()
    if rank==0 and it>=4: print('push %.0f bar %.0f reduce %.0f bar %.0f scatter %.0f bar %.0f merge %.0f bar %.0f resolve %.0f | total %.0f us'%tuple(a.elapsed_time(b_)*1e3 for a,b_ in [(ea,eb),(eb,ec),(ec,ed),(ed,e0),(e0,e1),(e1,e2),(e2,e3),(e3,e4),(e4,e5),(ea,e5)]))
if rank==0: print('inbox count', int(ex._count[0]), int(ex._count[1]))
dist.destroy_process_group()
