#!/usr/bin/env python
"""Bus bandwidth of dgs_comm_allreduce_async (NCCL over NVLink) by message size, to choose the bucket size.
    python -m torch.distributed.run --nproc-per-node N tools/nccl_probe.py"""
import json
import os
import pathlib
import sys

import torch
import torch.distributed as dist

ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from determined_clone_b200 import pytorch  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))


def bcast(obj):
    box = [obj]
    dist.broadcast_object_list(box, src=0)
    return box[0]


comm = pytorch.Comm(rank, world, torch.device("cuda", local), bcast)
st = torch.cuda.current_stream()
for mb in (1, 4, 8, 16, 25, 32, 64, 128, 256, 512, 1024):
    n = mb * (1 << 20) // 4
    x = torch.ones(n, device="cuda")
    for chunks in (1,) if mb < 256 else (1, 4):
        views = x.chunk(chunks)
        for _ in range(3):
            for v in views:
                comm.allreduce(v)
            comm.wait()
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        iters = 10
        e0.record(st)
        for _ in range(iters):
            for v in views:
                comm.allreduce(v)
            comm.wait()
        e1.record(st)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t)
        if rank == 0:
            alg = n * 4 / (ms * 1e-3) / 1e9
            print(json.dumps({"n_gpus": world, "mib": mb, "calls": chunks, "us": round(ms * 1e3, 1), "algbw_gbs": round(alg, 1),
                              "busbw_gbs": round(alg * 2 * (world - 1) / world, 1)}), flush=True)
comm.close()
dist.destroy_process_group()
