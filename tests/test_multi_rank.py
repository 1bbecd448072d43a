"""world_size-2 gloo tests (CPU) of the multi-GPU plan: batch sharding of DirectSort and the integer all-reduce +
mod-q fix-up that combines partial rank / output ciphertexts."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from fhe_sorting_b200 import sort as S


def test_shard_partition_covers_every_batch_once():
    for n, world in [(1024, 2), (1024, 4), (1024, 8), (1024, 16), (512, 4)]:
        B = n // min(n, 65536 // n)
        seen = []
        for r in range(world):
            b, sharded = S.shard_batches(n, r, world)
            assert sharded and len(b) == B // world
            seen += b
        assert sorted(seen) == list(range(B))
    for n, world in [(256, 8), (16, 2), (1024, 3)]:     # fewer batches than ranks (or not divisible): replicated
        B = n // min(n, 65536 // n)
        b, sharded = S.shard_batches(n, 1, world)
        assert not sharded and b == list(range(B))


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    moduli = [1152921504606584833, 1099512938497, 1099510054913]
    rng = np.random.default_rng(7)                       # same stream on both ranks: rank r takes slice r
    parts = np.stack([np.stack([rng.integers(0, m, 64, dtype=np.uint64) for m in moduli]) for _ in range(world)])
    got = S.combine_partial_residues(parts[rank], moduli, dist)
    want = np.zeros_like(parts[0])
    for i, m in enumerate(moduli):
        acc = np.zeros(64, dtype=object)
        for r in range(world):
            acc = acc + parts[r][i].astype(object)
        want[i] = np.array([int(x) % m for x in acc], dtype=np.uint64)
    ok = bool(np.array_equal(got, want))
    dist.barrier()
    dist.destroy_process_group()
    q.put((rank, ok))


def test_partial_ciphertext_combine_world2_gloo():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(0, True), (1, True)]


def _sort_worker(rank, world, port, q):
    """A whole sharded DirectSort on two ranks with no GPU: this repository's drivers over the CPU scheme layer
    (oracle/libfhesort_cpu.so), the partial ciphertexts combined by a gloo all-reduce of their words + the mod-q fix-up."""
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), FHEB200_CPU_THREADS="4")
    dist.init_process_group("gloo", rank=rank, world_size=world)
    res = {}
    try:
        import refdrivers
        L = refdrivers.cpu_library("own")
        n = 16                                             # one batch, 4 giant-step groups: sharded over the groups
        x = S.uniform_gap_input(n, seed=2)
        s = S.DirectSortSession(n, 12, seed=91, lib=L)
        base, _, level0 = s.sort(x)
        s.set_shard(rank, world, native=False, host=True)
        out, _, level = s.sort(x)
        ranks, _, _ = s.rank(x)
        s.close()
        res = {"err": float(np.abs(out - np.sort(x)).max()), "vs_unsharded": float(np.abs(out - base).max()),
               "rank_err": float(np.abs(ranks - np.argsort(np.argsort(x))).max()), "level": (level, level0)}
        # MEHP24 over several ciphertexts: 32 values as two 16 x 16 matrices on ring 2^10, the 3 comparison blocks and
        # 4 indicator products dealt over the two ranks, partial sums combined the same way
        x = S.uniform_gap_input(32, seed=9)
        m = S.Mehp24SortSession(32, 10, 48, seed=777, lib=L)
        base, _, l0 = m.sort(x)
        m.set_shard(rank, world, native=False, host=True)
        out, _, l1 = m.sort(x)
        m.close()
        res["mehp24"] = (float(np.abs(out - np.sort(x)).max()), float(np.abs(out - base).max()), l0 == l1)
    except Exception as e:  # noqa: BLE001
        res = {"error": repr(e)}
    dist.barrier()
    dist.destroy_process_group()
    q.put((rank, res))


def test_sharded_direct_sort_world2_gloo_on_the_cpu_scheme_layer():
    if not os.path.exists(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "libfhesort_cpu.so")):
        import pytest
        pytest.skip("oracle/libfhesort_cpu.so not built")
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_sort_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=600) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    for rank in (0, 1):
        r = got[rank]
        assert "error" not in r, r
        assert r["err"] < 1e-3 and r["vs_unsharded"] < 1e-5 and r["rank_err"] < 1e-3 and r["level"][0] == r["level"][1], (rank, r)
        assert r["mehp24"][0] < 1e-3 and r["mehp24"][1] < 1e-5 and r["mehp24"][2], (rank, r)
