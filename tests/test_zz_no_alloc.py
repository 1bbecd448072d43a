"""fheb200_ntt_forward_noalloc (allocation-free forward NTT, scratch in a static device array) against the existing
fheb200_raw_ntt, through the C ABI, on two calls in a row (the second call finds the first call's scratch contents).
Integer work: bit-exact.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_ntt_forward_noalloc_matches_raw_ntt_twice():
    import torch
    import fhe_sorting_b200 as F
    from fhe_sorting_b200.lib import _ck
    lib = F.load()
    ctx = F.Context(17, 4, 40, 60)
    n, nmod = ctx.n, ctx.L + ctx.K
    assert lib.fheb200_ntt_forward_noalloc_max_limbs() == 256
    rng = np.random.default_rng(2025)
    for call, qps in enumerate(([0, 1, 2, nmod - 1, 1, 3], [nmod - 1, 0, 2])):   # 60-bit, 40-bit and special primes; repeats
        host = np.stack([rng.integers(0, ctx.moduli[q], size=n, dtype=np.uint64) for q in qps])
        want = np.stack([ctx.raw_ntt(q, host[i]) for i, q in enumerate(qps)])
        d_in = torch.from_numpy(host.view(np.int64)).cuda()
        d_out = torch.zeros_like(d_in)
        torch.cuda.synchronize()
        arr = (C.c_int * len(qps))(*qps)
        _ck(lib.fheb200_ntt_forward_noalloc(ctx.h, arr, len(qps), C.c_void_p(d_in.data_ptr()), C.c_void_p(d_out.data_ptr())))
        ctx.sync()
        got = d_out.cpu().numpy().view(np.uint64)
        assert np.array_equal(got, want), "call %d" % call
        # in place as well
        _ck(lib.fheb200_ntt_forward_noalloc(ctx.h, arr, len(qps), C.c_void_p(d_in.data_ptr()), C.c_void_p(d_in.data_ptr())))
        ctx.sync()
        assert np.array_equal(d_in.cpu().numpy().view(np.uint64), want), "call %d in place" % call
    # beyond the static array: a clear error, no launch
    arr = (C.c_int * 257)(*([0] * 257))
    rc = lib.fheb200_ntt_forward_noalloc(ctx.h, arr, 257, C.c_void_p(d_in.data_ptr()), C.c_void_p(d_out.data_ptr()))
    assert rc != 0 and b"256" in lib.fheb200_last_error()
    small = F.Context(12, 4, 40, 60)
    rc = lib.fheb200_ntt_forward_noalloc(small.h, arr, 1, C.c_void_p(d_in.data_ptr()), C.c_void_p(d_out.data_ptr()))
    assert rc != 0 and b"2^17" in lib.fheb200_last_error()
