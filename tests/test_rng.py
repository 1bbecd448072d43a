"""Randomness of key generation and encryption: ChaCha20 keyed from OS entropy by default, the 64-bit
seeded generator only behind an explicit test flag."""
import ctypes as C

import numpy as np
import pytest

import fhe_sorting_b200 as F


def test_chacha20_block_known_answer():
    """RFC 8439 section 2.3.2: key 00..1f, counter 1, nonce 00:00:00:09:00:00:00:4a:00:00:00:00 -> first block words
    e4e7f110 15593bd1 1fdd0f50 c47120a3 (counter and nonce words map onto (index, stream) of the samplers)."""
    lib = F.load()
    key = (C.c_uint32 * 8)(*[int.from_bytes(bytes(range(4 * i, 4 * i + 4)), "little") for i in range(8)])
    out = (C.c_uint64 * 2)()
    index = 1 | (0x09000000 << 32)
    stream = 0x4A000000
    assert lib.fheb200_rng_block(key, C.c_uint64(stream), C.c_uint64(index), out) == 0
    assert out[0] == (0x15593BD1 << 32) | 0xE4E7F110
    assert out[1] == (0xC47120A3 << 32) | 0x1FDD0F50


@pytest.mark.gpu
def test_default_keys_come_from_os_entropy_and_seeds_need_the_test_flag():
    lib = F.load()
    ctx = F.Context(11, 3, 40, 60, batch=8)
    # a non-zero seed is refused until the test flag is set
    h = C.c_void_p()
    assert lib.fheb200_keygen(ctx.h, C.c_uint64(5), C.byref(h)) != 0
    assert b"deterministic" in lib.fheb200_last_error()
    k1, k2 = ctx.keygen(), ctx.keygen()
    s1, s2 = ctx.keys_export(k1, 0), ctx.keys_export(k2, 0)
    assert not np.array_equal(s1, s2)                       # every key pair draws fresh key material
    other = F.Context(11, 3, 40, 60, batch=8)
    assert not np.array_equal(other.keys_export(other.keygen(), 0), s1)
    # encryption under OS-entropy keys decrypts, and two encryptions of one plaintext differ
    v = np.array([0.5, -0.25, 0.125, 1, -1, 0, 0.75, -0.6])
    pt = ctx.make_plaintext(v, 1, 0, 8)
    a, b = ctx.encrypt(k2, pt), ctx.encrypt(k2, pt)
    assert np.abs(ctx.decrypt(k2, a, 8) - v).max() < 1e-6
    assert not np.array_equal(ctx.ct_export(a)[0], ctx.ct_export(b)[0])
    # the secret is ternary in the coefficient domain, with about n/3 of each value
    q0 = ctx.moduli[0]
    coef = ctx.raw_ntt(0, s2[0], inverse=True)
    vals, counts = np.unique(coef, return_counts=True)
    assert set(int(x) for x in vals) <= {0, 1, q0 - 1}
    assert counts.min() > ctx.n / 3 * 0.8
    # the public uniform polynomial is not drawn from the secret's generator state: a context that is handed the same 64
    # bytes reproduces the same keys (replication across ranks), any other bytes do not
    km = ctx.rng_keys()
    twin = F.Context(11, 3, 40, 60, batch=8)
    ctx.set_rng_keys(km), twin.set_rng_keys(km)
    assert np.array_equal(ctx.keys_export(ctx.keygen(), 0), twin.keys_export(twin.keygen(), 0))
    assert np.array_equal(ctx.keys_export(ctx.keygen(), 2), twin.keys_export(twin.keygen(), 2))
    twin.set_rng_keys(bytes(64 - 1) + b"\x01")
    assert not np.array_equal(ctx.keys_export(ctx.keygen(), 0), twin.keys_export(twin.keygen(), 0))


@pytest.mark.gpu
def test_seeded_keys_are_reproducible_in_test_mode():
    a, b = F.Context(11, 3, 40, 60, batch=8), F.Context(11, 3, 40, 60, batch=8)
    assert np.array_equal(a.keys_export(a.keygen(99), 0), b.keys_export(b.keygen(99), 0))
