"""Multi-GPU paths on real devices (needs >= 2 GPUs; the world_size-2 host logic also runs on CPU/gloo in
tests/test_multi_rank.py): the engine's own NCCL all-reduce + modular fix-up kernel, limb for limb, and sharded sorts
(batches for N = 128 on ring 2^13, giant-step groups for the single-batch N = 64) against the unsharded result."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)     # plumbing only: the data path is the engine's NCCL
    import ctypes as C
    import fhe_sorting_b200 as F
    from fhe_sorting_b200.sort import DirectSortSession, uniform_gap_input
    res = {}
    try:
        # ---- engine level: all-reduce of per-rank ciphertexts == the sum computed locally, every limb
        lib = F.load()
        ctx = F.Context(12, 6, 40, 60, batch=8, device=rank)
        keys = ctx.keygen(4242)                                      # test mode: identical keys on every rank
        ctx.eval_mult_keygen(keys)
        ident = [None]
        if rank == 0:
            buf = (C.c_ubyte * 128)()
            assert lib.fheb200_comm_unique_id(buf) == 0
            ident = [bytes(buf)]
        dist.broadcast_object_list(ident, src=0)
        assert lib.fheb200_ctx_comm_init(ctx.h, (C.c_ubyte * 128)(*ident[0]), rank, world) == 0, lib.fheb200_last_error()
        cts = [ctx.encrypt(keys, ctx.make_plaintext(np.arange(8) * 0.1 + r, 1, 0, 8)) for r in range(world)]
        want = cts[0]
        for c in cts[1:]:
            want = ctx.add(want, c)
        mine = ctx.clone(cts[rank])
        assert lib.fheb200_ct_all_reduce_sum(ctx.h, mine.h) == 0, lib.fheb200_last_error()
        g, w = ctx.ct_export(mine), ctx.ct_export(want)
        res["allreduce_limbs_equal"] = bool(np.array_equal(g[0], w[0]) and np.array_equal(g[1], w[1]))
        # a degree-2 ciphertext after a multiplication, at a lower level (fewer limbs, scaling primes only + q0)
        deep = [ctx.mult(c, c) for c in cts]
        want = deep[0]
        for c in deep[1:]:
            want = ctx.add(want, c)
        mine = ctx.clone(deep[rank])
        assert lib.fheb200_ct_all_reduce_sum(ctx.h, mine.h) == 0
        g, w = ctx.ct_export(mine), ctx.ct_export(want)
        res["allreduce_deep_equal"] = bool(np.array_equal(g[0], w[0]) and np.array_equal(g[1], w[1]))
        lib.fheb200_ctx_comm_destroy(ctx.h)
        # ---- sorts: sharded over batches (N = 128, 4 batches on ring 2^13) and over giant-step groups (N = 64, one batch)
        for n in (128, 64):
            x = uniform_gap_input(n, seed=3)
            s = DirectSortSession(n, 13, seed=500 + n, device=rank)
            base, _, level0 = s.sort(x)
            ranks0, _, _ = s.rank(x)
            s.set_shard(rank, world)
            out, _, level = s.sort(x)
            ranks, _, _ = s.rank(x)
            s.close()
            res["sort%d" % n] = (float(np.abs(out - np.sort(x)).max()), float(np.abs(out - base).max()),
                                 float(np.abs(ranks - ranks0).max()), level == level0)
        # ---- MEHP24 over several ciphertexts: 64 values as two 32 x 32 matrices on ring 2^12, the 3 comparison
        # blocks and 4 indicator products dealt over the ranks, partial sums all-reduced
        from fhe_sorting_b200.sort import Mehp24SortSession
        x = uniform_gap_input(64, seed=9)
        m = Mehp24SortSession(64, 12, 48, seed=777, device=rank)
        base, _, level0 = m.sort(x)
        m.set_shard(rank, world)
        out, _, level = m.sort(x)
        m.close()
        res["mehp24"] = (float(np.abs(out - np.sort(x)).max()), float(np.abs(out - base).max()), level == level0)
    except Exception as e:  # noqa: BLE001
        res["error"] = repr(e)
    dist.barrier()
    dist.destroy_process_group()
    q.put((rank, res))


def test_native_allreduce_and_sharded_sorts_two_gpus():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=300) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    for rank in (0, 1):
        r = got[rank]
        assert "error" not in r, r
        assert r["allreduce_limbs_equal"] and r["allreduce_deep_equal"]
        for n in (128, 64):
            err, vs_unsharded, rank_diff, same_level = r["sort%d" % n]
            assert err < 0.01 and vs_unsharded < 1e-3 and rank_diff < 1e-3 and same_level, (rank, n, r["sort%d" % n])
        err, vs_unsharded, same_level = r["mehp24"]
        assert err < 0.01 and vs_unsharded < 1e-4 and same_level, (rank, r["mehp24"])
