"""ctypes binding of the CPU oracle (oracle/ckks_oracle.cpp).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "libckks_oracle.so")
_U64P = np.ctypeslib.ndpointer(dtype=np.uint64, flags="C_CONTIGUOUS")
_I64P = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")
_F64P = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")


def build():
    """(Re)build libckks_oracle.so when missing or older than its sources."""
    src = [os.path.join(_HERE, f) for f in ("ckks_oracle.cpp", "ckks_oracle.h")]
    if (not os.path.exists(_LIB)) or any(os.path.getmtime(s) > os.path.getmtime(_LIB) for s in src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _LIB


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    build()
    L = C.CDLL(_LIB)
    ctx = C.c_void_p
    i, u64, u32, dbl = C.c_int, C.c_uint64, C.c_uint32, C.c_double
    sig = {
        "orc_ctx_create": (ctx, [i, i, i, i, i]),
        "orc_ctx_free": (None, [ctx]),
        "orc_n": (i, [ctx]), "orc_L": (i, [ctx]), "orc_K": (i, [ctx]), "orc_alpha": (i, [ctx]),
        "orc_modulus": (u64, [ctx, i]), "orc_psi": (u64, [ctx, i]),
        "orc_scale_real": (dbl, [ctx, i]), "orc_scale_big": (dbl, [ctx, i]),
        "orc_set_threads": (None, [i]),
        "orc_ntt_fwd": (None, [ctx, i, _U64P]), "orc_ntt_inv": (None, [ctx, i, _U64P]),
        "orc_automorphism_eval": (None, [ctx, u32, i, _U64P, _U64P]),
        "orc_automorphism_coeff": (None, [ctx, u32, i, _U64P, _U64P]),
        "orc_rot_to_galois": (u32, [ctx, i]),
        "orc_add": (None, [ctx, i, _U64P, _U64P, _U64P]),
        "orc_sub": (None, [ctx, i, _U64P, _U64P, _U64P]),
        "orc_mul": (None, [ctx, i, _U64P, _U64P, _U64P]),
        "orc_mul_scalar": (None, [ctx, i, _U64P, _U64P, _U64P]),
        "orc_add_scalar": (None, [ctx, i, _U64P, _U64P, _U64P]),
        "orc_tensor": (None, [ctx, i] + [_U64P] * 7),
        "orc_rescale": (None, [ctx, i, _U64P, _U64P]),
        "orc_evk_words": (u64, [ctx]),
        "orc_beta": (i, [ctx, i]),
        "orc_modup": (None, [ctx, i, _U64P, _U64P]),
        "orc_inner_product": (None, [ctx, i, _U64P, _U64P, _U64P, _U64P]),
        "orc_moddown": (None, [ctx, i, _U64P, _U64P]),
        "orc_keyswitch": (None, [ctx, i, _U64P, _U64P, _U64P, _U64P]),
        "orc_rotate": (None, [ctx, i, u32, _U64P, _U64P, _U64P, _U64P, _U64P]),
        "orc_rotate_hoisted": (None, [ctx, i, u32, _U64P, _U64P, _U64P, _U64P, _U64P]),
        "orc_mult_relin": (None, [ctx, i] + [_U64P] * 7),
        "orc_rng_u64": (u64, [u64, u64, u64, u64]),
        "orc_sample_uniform": (None, [ctx, u64, u64, i, _U64P]),
        "orc_sample_gauss": (None, [ctx, u64, u64, _I64P]),
        "orc_sample_ternary": (None, [ctx, u64, u64, _I64P]),
        "orc_keygen_secret": (None, [ctx, u64, _U64P]),
        "orc_keygen_public": (None, [ctx, u64, _U64P, _U64P, _U64P]),
        "orc_keygen_switch": (None, [ctx, u64, u64, _U64P, _U64P, _U64P]),
        "orc_keygen_mult": (None, [ctx, u64, _U64P, _U64P]),
        "orc_keygen_rot": (None, [ctx, u64, u32, _U64P, _U64P]),
        "orc_encode": (None, [ctx, _F64P, i, i, dbl, i, _U64P]),
        "orc_encrypt": (None, [ctx, u64, u64, i, _U64P, _U64P, _U64P, _U64P, _U64P]),
        "orc_decrypt": (None, [ctx, i, _U64P, _U64P, _U64P, i, dbl, _F64P, C.c_void_p]),
    }
    for name, (res, args) in sig.items():
        f = getattr(L, name)
        f.restype, f.argtypes = res, args
    _lib = L
    return L


class Oracle:
    """Thin object wrapper: polynomials are numpy uint64 arrays of shape [limbs, n]."""

    def __init__(self, logn, depth, scale_bits=40, first_bits=60, dnum=3, threads=1):
        self.l = lib()
        self.l.orc_set_threads(threads)
        self.h = self.l.orc_ctx_create(logn, depth, scale_bits, first_bits, dnum)
        if not self.h:
            raise ValueError("bad oracle parameters")
        self.logn, self.depth = logn, depth
        self.n, self.L, self.K = self.l.orc_n(self.h), self.l.orc_L(self.h), self.l.orc_K(self.h)
        self.alpha = self.l.orc_alpha(self.h)
        self.moduli = [self.l.orc_modulus(self.h, i) for i in range(self.L + self.K)]

    def __del__(self):
        try:
            self.l.orc_ctx_free(self.h)
        except Exception:
            pass

    # -- parameters
    def psi(self, i):
        return self.l.orc_psi(self.h, i)

    def sf(self, level):
        return self.l.orc_scale_real(self.h, level)

    def sf_big(self, level):
        return self.l.orc_scale_big(self.h, level)

    def beta(self, ell):
        return self.l.orc_beta(self.h, ell)

    def galois(self, rot):
        return self.l.orc_rot_to_galois(self.h, rot)

    def zeros(self, limbs):
        return np.zeros((limbs, self.n), dtype=np.uint64)

    def random_poly(self, rng, ell, qp_offset=0):
        out = self.zeros(ell)
        for i in range(ell):
            out[i] = rng.integers(0, self.moduli[qp_offset + i], size=self.n, dtype=np.uint64)
        return out

    # -- primitives
    def ntt_fwd(self, qp, a):
        a = np.ascontiguousarray(a.copy())
        self.l.orc_ntt_fwd(self.h, qp, a)
        return a

    def ntt_inv(self, qp, a):
        a = np.ascontiguousarray(a.copy())
        self.l.orc_ntt_inv(self.h, qp, a)
        return a

    def automorphism_eval(self, k, a):
        out = np.empty_like(a)
        self.l.orc_automorphism_eval(self.h, k, a.shape[0], np.ascontiguousarray(a), out)
        return out

    def automorphism_coeff(self, k, qp, a):
        out = np.empty_like(a)
        self.l.orc_automorphism_coeff(self.h, k, qp, np.ascontiguousarray(a), out)
        return out

    def _bin(self, fn, a, b):
        out = np.empty_like(a)
        fn(self.h, a.shape[0], np.ascontiguousarray(a), np.ascontiguousarray(b), out)
        return out

    def add(self, a, b):
        return self._bin(self.l.orc_add, a, b)

    def sub(self, a, b):
        return self._bin(self.l.orc_sub, a, b)

    def mul(self, a, b):
        return self._bin(self.l.orc_mul, a, b)

    def mul_scalar(self, a, s):
        return self._bin(self.l.orc_mul_scalar, a, np.asarray(s, dtype=np.uint64))

    def add_scalar(self, a, s):
        return self._bin(self.l.orc_add_scalar, a, np.asarray(s, dtype=np.uint64))

    def tensor(self, a0, a1, b0, b1):
        d = [np.empty_like(a0) for _ in range(3)]
        self.l.orc_tensor(self.h, a0.shape[0], a0, a1, b0, b1, *d)
        return d

    def rescale(self, a):
        out = self.zeros(a.shape[0] - 1)
        self.l.orc_rescale(self.h, a.shape[0], np.ascontiguousarray(a), out)
        return out

    def modup(self, c):
        ell = c.shape[0]
        out = np.zeros((self.beta(ell), ell + self.K, self.n), dtype=np.uint64)
        self.l.orc_modup(self.h, ell, np.ascontiguousarray(c), out)
        return out

    def inner_product(self, ell, digits, evk):
        a0, a1 = self.zeros(ell + self.K), self.zeros(ell + self.K)
        self.l.orc_inner_product(self.h, ell, digits, evk, a0, a1)
        return a0, a1

    def moddown(self, ell, acc):
        out = self.zeros(ell)
        self.l.orc_moddown(self.h, ell, np.ascontiguousarray(acc), out)
        return out

    def keyswitch(self, c, evk):
        ell = c.shape[0]
        k0, k1 = self.zeros(ell), self.zeros(ell)
        self.l.orc_keyswitch(self.h, ell, np.ascontiguousarray(c), evk, k0, k1)
        return k0, k1

    def rotate(self, k, c0, c1, evk):
        ell = c0.shape[0]
        o0, o1 = self.zeros(ell), self.zeros(ell)
        self.l.orc_rotate(self.h, ell, k, np.ascontiguousarray(c0), np.ascontiguousarray(c1), evk, o0, o1)
        return o0, o1

    def rotate_hoisted(self, k, c0, digits, evk):
        ell = c0.shape[0]
        o0, o1 = self.zeros(ell), self.zeros(ell)
        self.l.orc_rotate_hoisted(self.h, ell, k, np.ascontiguousarray(c0), digits, evk, o0, o1)
        return o0, o1

    def mult_relin(self, a0, a1, b0, b1, evk):
        ell = a0.shape[0]
        o0, o1 = self.zeros(ell), self.zeros(ell)
        self.l.orc_mult_relin(self.h, ell, a0, a1, b0, b1, evk, o0, o1)
        return o0, o1

    # -- keys
    def evk_shape(self):
        return (self.beta(self.L), 2, self.L + self.K, self.n)

    def keygen_secret(self, seed):
        s = self.zeros(self.L + self.K)
        self.l.orc_keygen_secret(self.h, seed, s)
        return s

    def keygen_public(self, seed, s):
        b, a = self.zeros(self.L), self.zeros(self.L)
        self.l.orc_keygen_public(self.h, seed, s, b, a)
        return b, a

    def keygen_mult(self, seed, s):
        evk = np.zeros(self.evk_shape(), dtype=np.uint64)
        self.l.orc_keygen_mult(self.h, seed, s, evk)
        return evk

    def keygen_rot(self, seed, k, s):
        evk = np.zeros(self.evk_shape(), dtype=np.uint64)
        self.l.orc_keygen_rot(self.h, seed, k, s, evk)
        return evk

    # -- encode / encrypt / decrypt
    def encode(self, vals, slots, scale, ell):
        vals = np.ascontiguousarray(np.asarray(vals, dtype=np.float64))
        pt = self.zeros(ell)
        self.l.orc_encode(self.h, vals, len(vals), slots, scale, ell, pt)
        return pt

    def encrypt(self, seed, nonce, pk, pt):
        ell = pt.shape[0]
        c0, c1 = self.zeros(ell), self.zeros(ell)
        self.l.orc_encrypt(self.h, seed, nonce, ell, pk[0], pk[1], pt, c0, c1)
        return c0, c1

    def decrypt(self, c0, c1, s, slots, scale):
        out = np.zeros(slots, dtype=np.float64)
        self.l.orc_decrypt(self.h, c0.shape[0], np.ascontiguousarray(c0), np.ascontiguousarray(c1), s, slots,
                           scale, out, None)
        return out
