import os, sys, time, datetime as dt
sys.path.insert(0, os.getcwd())
import torch, torch.distributed as dist
import ctypes as C, numpy as np
from smsgate_b200 import synth, _lib as L
from smsgate_b200.batch import Batch
from smsgate_b200.replay import ReplayEngine
from smsgate_b200.dist import GlobalDedup
rank=int(os.environ['RANK']); world=int(os.environ['WORLD_SIZE']); lr=int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(lr)
dist.init_process_group('nccl', device_id=torch.device('cuda', lr))
NOW=dt.datetime(2026,1,1); n=1_000_000
backup=synth.generate(n, seed=100+rank)
b=Batch.from_strings(backup.bodies, backup.senders, backup.date_ms)
eng=ReplayEngine(lr); eng.load_cache(backup.cache)
d_bytes=torch.from_numpy(b.body_bytes.copy()).cuda(); d_off=torch.from_numpy(b.body_offsets.copy()).cuda()
cols=eng.device_columns(n, d_bytes.numel()); eng.reserve(n, d_bytes.numel())
ex=GlobalDedup(eng, dist, n)
lib=L.lib()
def T():
    torch.cuda.synchronize(); return time.perf_counter()
for it in range(6):
    eng.replay_device(d_bytes, d_off, NOW, cols, base_index=rank*n)
    out=eng._dev_outputs(cols)
    t0=T()
    L.check(eng._ctx, lib.smsgate_dedup_export(eng._ctx, n, rank*n, world, C.byref(out), ex.records.data_ptr(), ex.rec_slot.data_ptr(), ex.h_counts.ctypes.data_as(C.c_void_p)))
    t1=T()
    sc=[int(x) for x in ex.h_counts]; rc=ex.plan.exchange_counts(sc)
    t2=T()
    recv=ex.plan.rows(ex.records, sc, rc, out=ex._buf('_recv', sum(rc), 32, torch.uint8))
    t3=T()
    L.check(eng._ctx, lib.smsgate_reserve_merge(eng._ctx, sum(rc)))
    ver=ex._buf('_verdicts', sum(rc), 2, torch.int64)
    L.check(eng._ctx, lib.smsgate_dedup_merge(eng._ctx, sum(rc), recv.data_ptr(), ver.data_ptr()))
    t4=T()
    back=ex.plan.rows(ver, rc, sc, out=ex._buf('_back', sum(sc), 2, torch.int64))
    t5=T()
    L.check(eng._ctx, lib.smsgate_dedup_apply(eng._ctx, n, rank*n, sum(sc), ex.rec_slot.data_ptr(), back.data_ptr(), C.byref(out)))
    t6=T()
    if rank==0 and it>=3: print('export %.0f counts %.0f a2a_rec %.0f merge %.0f a2a_ver %.0f apply %.0f us'%tuple((b_-a_)*1e6 for a_,b_ in [(t0,t1),(t1,t2),(t2,t3),(t3,t4),(t4,t5),(t5,t6)]), sc, rc)
dist.destroy_process_group()
