"""How much host<->device bandwidth a rank gets when 1 / all ranks of the box copy at once: the e2e figure of bench.py
moves 124 MB in and 65 MB out per rank and step, concurrently on all ranks.  Pinned host buffers, two streams (H2D and
D2H at the same time, like smsgate_replay_host*), CUDA events.  Launch with torchrun like bench.py; rank 0 prints one
JSON line.  usage: torchrun --nproc-per-node N scripts/pcie_probe.py"""
import json
import os

import torch
import torch.distributed as dist


def main():
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    nin, nout = 124 << 20, 65 << 20
    h_in = torch.empty(nin, dtype=torch.uint8).pin_memory()
    h_out = torch.empty(nout, dtype=torch.uint8).pin_memory()
    d_in = torch.empty(nin, dtype=torch.uint8, device="cuda")
    d_out = torch.empty(nout, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def once(both):
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        torch.cuda.synchronize()
        e0.record()
        s1.wait_event(e0)
        s2.wait_event(e0)
        with torch.cuda.stream(s1):
            d_in.copy_(h_in, non_blocking=True)
            e1.record(s1)
        if both:
            with torch.cuda.stream(s2):
                h_out.copy_(d_out, non_blocking=True)
                e2.record(s2)
        torch.cuda.synchronize()
        return e0.elapsed_time(e1), (e0.elapsed_time(e2) if both else 0.0)

    def measure(active, both):
        if world > 1:
            dist.barrier()
        ts = []
        for _ in range(8):
            if world > 1:
                dist.barrier()
            if active:
                ts.append(once(both))
        if not ts:
            return 0.0, 0.0
        ts = ts[2:]
        return sum(t[0] for t in ts) / len(ts), sum(t[1] for t in ts) / len(ts)

    res = {}
    for name, active, both in (("h2d_alone_rank0", rank == 0, False), ("h2d_d2h_alone_rank0", rank == 0, True),
                               ("h2d_all_ranks", True, False), ("h2d_d2h_all_ranks", True, True)):
        tin, tout = measure(active, both)
        t = torch.tensor([tin, tout], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        tin, tout = (float(x) for x in t.cpu())
        res[name] = {"h2d_ms": tin, "h2d_gbs_per_rank": nin / tin / 1e6 if tin else None,
                     "d2h_ms": tout, "d2h_gbs_per_rank": (nout / tout / 1e6 if tout else None)}
    if rank == 0:
        print(json.dumps({"n_gpus": world, "h2d_bytes": nin, "d2h_bytes": nout, "slowest_rank": res}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
