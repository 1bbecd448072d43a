#!/usr/bin/env python
"""DirectSort sharded over the ranks of a torchrun job (one GPU each): torchrun --nproc-per-node G tools/sort_multi.py N"""
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, ".")
from fhe_sorting_b200.sort import DirectSortSession, uniform_gap_input, shard_batches  # noqa: E402

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
for n in [int(a) for a in sys.argv[1:]] or [1024]:
    s = DirectSortSession(n, 17, seed=4242, device=local)
    s.set_shard(rank, world)
    x = uniform_gap_input(n)
    times = []
    for rep in range(2):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        out, ms, level = s.sort(x)
        torch.cuda.synchronize()
        t = torch.tensor([time.perf_counter() - t0], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        times.append(float(t.item()))
    err = float(np.abs(out - np.sort(x)).max())
    if rank == 0:
        print(json.dumps({"N": n, "gpus": world, "sharded": shard_batches(n, rank, world)[1], "sort_s": times, "max_err": err,
                          "level": level}), flush=True)
    s.close()
if world > 1:
    dist.destroy_process_group()
