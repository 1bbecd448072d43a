"""ctypes binding of libfheb200.so (include/fhe_b200.h).

The library is the product: CUDA kernels for sm_100a behind a C ABI.  There is no CPU fallback — loading
works anywhere (so symbols can be checked), but creating a context raises when no CUDA device is present.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libfheb200.so")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "fhe_b200.h")

_U64P = np.ctypeslib.ndpointer(dtype=np.uint64, flags="C_CONTIGUOUS")
_F64P = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")


class FheError(RuntimeError):
    pass


class Params(C.Structure):
    _fields_ = [("log_ring_dim", C.c_int), ("mult_depth", C.c_int), ("scaling_mod_size", C.c_int),
                ("first_mod_size", C.c_int), ("batch_size", C.c_int), ("num_large_digits", C.c_int),
                ("device", C.c_int)]


_lib = None


def load():
    """dlopen the in-tree library; raises FheError when it has not been built (run __graft_entry__.build())."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FheError("libfheb200.so is missing: build it with `make -C fhe_sorting_b200` "
                       "(python -c 'import __graft_entry__ as g; g.build()'); there is no CPU fallback")
    L = C.CDLL(LIB_PATH)
    L.fheb200_last_error.restype = C.c_char_p
    L.fheb200_version.restype = C.c_char_p
    _lib = L
    return L


def _ck(rc):
    if rc != 0:
        raise FheError(load().fheb200_last_error().decode())


def launch_count():
    """CUDA kernels launched by libfheb200.so in this process so far."""
    L = load()
    L.fheb200_launch_count.restype = C.c_ulonglong
    return int(L.fheb200_launch_count())


def algo_bytes():
    """Algorithmic bytes (SURVEY.md 8d formulas) of the scheme operations issued by this process so far."""
    L = load()
    L.fheb200_algo_bytes.restype = C.c_ulonglong
    return int(L.fheb200_algo_bytes())


def op_counts():
    """(rescales, relinearisations, merged relinearise-and-rescale, rotations) since the library was loaded"""
    L = load()
    L.fheb200_op_count.restype = C.c_ulonglong
    return tuple(int(L.fheb200_op_count(i)) for i in range(4))


def params_chain(logn, depth, scale_bits=40, first_bits=60, dnum=3):
    """Host-only: modulus chain / roots / (L, K, alpha) for a parameter set (no GPU needed)."""
    L = load()
    p = Params(logn, depth, scale_bits, first_bits, 0, dnum, 0)
    mod = np.zeros(128, dtype=np.uint64)
    psi = np.zeros(128, dtype=np.uint64)
    nq, npp, al = C.c_int(), C.c_int(), C.c_int()
    _ck(L.fheb200_params_chain(C.byref(p), mod.ctypes.data_as(C.c_void_p), psi.ctypes.data_as(C.c_void_p), 128,
                               C.byref(nq), C.byref(npp), C.byref(al)))
    F = nq.value + npp.value
    return [int(x) for x in mod[:F]], [int(x) for x in psi[:F]], nq.value, npp.value, al.value


def bootstrap_plan_info(ring_dim, slots, budget_enc, budget_dec, q0, raise_scale, correction=12):
    """Host-only: depth, K, double-angle count, rotation indices and Chebyshev series of the bootstrapping plan."""
    L = load()
    depth, K, r, nrot, ncheb = (C.c_int() for _ in range(5))
    shift = C.c_double()
    rots = (C.c_int * 4096)()
    cheb = (C.c_double * 1024)()
    _ck(L.fheb200_bootstrap_plan_info(ring_dim, slots, budget_enc, budget_dec, C.c_double(q0), C.c_double(raise_scale),
                                      correction, C.byref(depth), C.byref(K), C.byref(r), rots, 4096, C.byref(nrot), cheb,
                                      1024, C.byref(ncheb), C.byref(shift)))
    return dict(depth=depth.value, K=K.value, r=r.value, shift=shift.value, rotations=list(rots[:nrot.value]),
                cheb=np.array(cheb[:ncheb.value]))


def bootstrap_plan_apply(ring_dim, slots, budget_enc, budget_dec, q0, raise_scale, which, values, correction=12):
    """Host-only: the plan's CoeffsToSlots (which=0) / SlotsToCoeffs (which=1) on 2*slots complex values, in the clear."""
    L = load()
    v = np.ascontiguousarray(np.asarray(values, dtype=np.complex128))
    assert v.shape == (2 * slots,)
    out = np.zeros_like(v)
    _ck(L.fheb200_bootstrap_plan_apply(ring_dim, slots, budget_enc, budget_dec, C.c_double(q0), C.c_double(raise_scale),
                                       correction, which, v.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p)))
    return out


class _Handle:
    _release = None

    def __init__(self, h):
        self.h = h

    def __del__(self):
        try:
            if self.h and _lib is not None:
                getattr(_lib, self._release)(self.h)
                self.h = None
        except Exception:
            pass


class Keys(_Handle):
    _release = "fheb200_keys_release"


class Precomp(_Handle):
    _release = "fheb200_pre_release"


class Plaintext(_Handle):
    _release = "fheb200_pt_release"

    def values(self, count=None):
        n = C.c_size_t()
        _ck(_lib.fheb200_pt_values(self.h, None, C.c_size_t(0), C.byref(n)))
        out = np.zeros(n.value, dtype=np.float64)
        _ck(_lib.fheb200_pt_values(self.h, out.ctypes.data_as(C.c_void_p), C.c_size_t(n.value), C.byref(n)))
        return out if count is None else out[:count]

    def set_length(self, n):
        _ck(_lib.fheb200_pt_set_length(self.h, C.c_size_t(n)))


class Ciphertext(_Handle):
    _release = "fheb200_ct_release"

    def info(self):
        lv, nsd, sl, nl, sc = C.c_int(), C.c_int(), C.c_int(), C.c_int(), C.c_double()
        _ck(_lib.fheb200_ct_info(self.h, C.byref(lv), C.byref(nsd), C.byref(sl), C.byref(nl), C.byref(sc)))
        return dict(level=lv.value, nsd=nsd.value, slots=sl.value, limbs=nl.value, scale=sc.value)

    @property
    def level(self):
        return self.info()["level"]

    @property
    def limbs(self):
        return self.info()["limbs"]

    def set_slots(self, s):
        _ck(_lib.fheb200_ct_set_slots(self.h, s))


class Context(_Handle):
    """Mirror of lbcrypto::CryptoContext<DCRTPoly> for the calls the reference makes."""
    _release = "fheb200_ctx_release"

    def __init__(self, logn, depth, scale_bits=40, first_bits=60, batch=0, dnum=3, device=0):
        L = load()
        p = Params(logn, depth, scale_bits, first_bits, batch, dnum, device)
        h = C.c_void_p()
        _ck(L.fheb200_ctx_create(C.byref(p), C.byref(h)))
        super().__init__(h)
        n, nq, npp, al, b = (C.c_int() for _ in range(5))
        _ck(L.fheb200_ctx_info(h, C.byref(n), C.byref(nq), C.byref(npp), C.byref(al), C.byref(b)))
        self.n, self.L, self.K, self.alpha, self.batch = n.value, nq.value, npp.value, al.value, b.value
        self.logn, self.depth = logn, depth
        self.moduli = [self.modulus(i)[0] for i in range(self.L + self.K)]

    # -- context
    def modulus(self, i):
        q, psi = C.c_uint64(), C.c_uint64()
        _ck(_lib.fheb200_ctx_modulus(self.h, i, C.byref(q), C.byref(psi)))
        return q.value, psi.value

    def sf(self, level):
        a, b = C.c_double(), C.c_double()
        _ck(_lib.fheb200_ctx_scaling_factor(self.h, level, C.byref(a), C.byref(b)))
        return a.value

    def sf_big(self, level):
        a, b = C.c_double(), C.c_double()
        _ck(_lib.fheb200_ctx_scaling_factor(self.h, level, C.byref(a), C.byref(b)))
        return b.value

    def set_stream(self, cuda_stream_ptr):
        _ck(_lib.fheb200_ctx_set_stream(self.h, C.c_void_p(cuda_stream_ptr)))

    def sync(self):
        _ck(_lib.fheb200_ctx_sync(self.h))

    def key_bytes(self):
        b = C.c_uint64()
        _ck(_lib.fheb200_ctx_key_bytes(self.h, C.byref(b)))
        return b.value

    def beta(self, ell):
        return (ell + self.alpha - 1) // self.alpha

    def galois(self, rot):
        g = C.c_uint32()
        _ck(_lib.fheb200_galois_for_rotation(self.h, rot, C.byref(g)))
        return g.value

    # -- keys
    def keygen(self, seed=0):
        """seed = 0: ChaCha20 keyed from OS entropy (the default).  seed != 0: TEST-ONLY deterministic key material from the
        64-bit seeded generator the CPU oracle shares (enables fheb200_ctx_allow_deterministic_keys on this context)."""
        if seed:
            _ck(_lib.fheb200_ctx_allow_deterministic_keys(self.h, 1))
        h = C.c_void_p()
        _ck(_lib.fheb200_keygen(self.h, C.c_uint64(seed), C.byref(h)))
        return Keys(h)

    def rng_keys(self):
        """64 bytes of generator key material (treat like the secret key): hand them to set_rng_keys on the other ranks."""
        buf = (C.c_ubyte * 64)()
        _ck(_lib.fheb200_ctx_get_rng_keys(self.h, buf))
        return bytes(buf)

    def set_rng_keys(self, keys64):
        assert len(keys64) == 64
        _ck(_lib.fheb200_ctx_set_rng_keys(self.h, (C.c_ubyte * 64)(*keys64)))

    def eval_mult_keygen(self, keys):
        _ck(_lib.fheb200_eval_mult_keygen(self.h, keys.h))

    def eval_rotate_keygen(self, keys, indices):
        arr = (C.c_int * len(indices))(*indices)
        _ck(_lib.fheb200_eval_rotate_keygen(self.h, keys.h, arr, C.c_size_t(len(indices))))

    # -- encode / encrypt / decrypt
    def make_plaintext(self, values, nsd=1, level=0, slots=0):
        v = np.ascontiguousarray(np.asarray(values, dtype=np.float64))
        h = C.c_void_p()
        _ck(_lib.fheb200_make_packed_plaintext(self.h, v.ctypes.data_as(C.c_void_p), C.c_size_t(len(v)), nsd, level,
                                               slots, C.byref(h)))
        return Plaintext(h)

    def rotate_plaintext(self, pt, index):
        """pt with its slots rotated left by index; encoded as a Galois permutation of pt's encoding."""
        h = C.c_void_p()
        _ck(_lib.fheb200_pt_rotate(self.h, pt.h, int(index), C.byref(h)))
        return Plaintext(h)

    def encrypt(self, keys, pt):
        h = C.c_void_p()
        _ck(_lib.fheb200_encrypt(self.h, keys.h, pt.h, C.byref(h)))
        return Ciphertext(h)

    def decrypt(self, keys, ct, count=None):
        h = C.c_void_p()
        _ck(_lib.fheb200_decrypt(self.h, keys.h, ct.h, C.byref(h)))
        return Plaintext(h).values(count)

    # -- arithmetic
    def _un(self, fn, a, *args):
        h = C.c_void_p()
        _ck(getattr(_lib, fn)(self.h, a.h, *args, C.byref(h)))
        return Ciphertext(h)

    def clone(self, a):
        return self._un("fheb200_ct_clone", a)

    def add(self, a, b):
        if isinstance(b, Ciphertext):
            return self._un("fheb200_eval_add", a, b.h)
        if isinstance(b, Plaintext):
            return self._un("fheb200_eval_add_plain", a, b.h)
        return self._un("fheb200_eval_add_const", a, C.c_double(float(b)))

    def sub(self, a, b):
        if isinstance(b, Ciphertext):
            return self._un("fheb200_eval_sub", a, b.h)
        if isinstance(b, Plaintext):
            return self._un("fheb200_eval_sub_plain", a, b.h)
        return self._un("fheb200_eval_add_const", a, C.c_double(-float(b)))

    def negate(self, a):
        return self._un("fheb200_eval_negate", a)

    def mult(self, a, b):
        if isinstance(b, Ciphertext):
            return self._un("fheb200_eval_mult", a, b.h)
        if isinstance(b, Plaintext):
            return self._un("fheb200_eval_mult_plain", a, b.h)
        return self._un("fheb200_eval_mult_const", a, C.c_double(float(b)))

    def mult_rescaled(self, a, b):
        """EvalMult + the lazy rescale of the product as one ModDown by P * q_last (engine_arith.cu: relinearize_reduced)."""
        return self._un("fheb200_eval_mult_rescaled", a, b.h)

    def square(self, a):
        return self._un("fheb200_eval_mult", a, a.h)

    def rescale(self, a):
        return self._un("fheb200_eval_rescale", a)

    def add_many(self, cts):
        arr = (C.c_void_p * len(cts))(*[c.h for c in cts])
        h = C.c_void_p()
        _ck(_lib.fheb200_eval_add_many(self.h, arr, C.c_size_t(len(cts)), C.byref(h)))
        return Ciphertext(h)

    def rotate(self, a, index):
        return self._un("fheb200_eval_rotate", a, C.c_int(index))

    def fast_rotation_precompute(self, a):
        h = C.c_void_p()
        _ck(_lib.fheb200_fast_rotation_precompute(self.h, a.h, C.byref(h)))
        return Precomp(h)

    def fast_rotation(self, a, index, pre):
        return self._un("fheb200_eval_fast_rotation", a, C.c_int(index), pre.h)

    def chebyshev(self, x, coeffs, a=-1.0, b=1.0):
        v = np.ascontiguousarray(np.asarray(coeffs, dtype=np.float64))
        return self._un("fheb200_eval_chebyshev_series", x, v.ctypes.data_as(C.c_void_p), C.c_size_t(len(v)),
                        C.c_double(a), C.c_double(b))

    def poly(self, x, coeffs):
        v = np.ascontiguousarray(np.asarray(coeffs, dtype=np.float64))
        return self._un("fheb200_eval_poly", x, v.ctypes.data_as(C.c_void_p), C.c_size_t(len(v)))

    # -- bootstrapping
    def bootstrap_setup(self, budget=(4, 4), slots=0):
        _ck(_lib.fheb200_eval_bootstrap_setup(self.h, int(budget[0]), int(budget[1]), int(slots)))

    def bootstrap_keygen(self, keys, slots=0):
        _ck(_lib.fheb200_eval_bootstrap_keygen(self.h, keys.h, int(slots)))

    def bootstrap(self, a, stop_after=0, iterations=1, precision=0):
        if iterations > 1:
            return self._un("fheb200_eval_bootstrap_iterated", a, C.c_int(iterations), C.c_int(precision))
        if stop_after:
            return self._un("fheb200_eval_bootstrap_partial", a, C.c_int(stop_after))
        return self._un("fheb200_eval_bootstrap", a)

    def bootstrap_depth(self, slots=0):
        d = C.c_int()
        _ck(_lib.fheb200_bootstrap_depth(self.h, int(slots), C.byref(d)))
        return d.value

    def conjugate(self, a):
        return self._un("fheb200_eval_conjugate", a)

    # -- raw RNS access (parity tests)
    def ct_export(self, ct, out=None):
        """Raw limbs of a ciphertext; `out` = (c0, c1) preallocated (e.g. pinned) uint64 [limbs][n] host arrays."""
        ell = ct.limbs
        if out is not None:
            c0, c1 = out
            assert c0.dtype == np.uint64 and c0.shape == (ell, self.n) and c0.flags.c_contiguous
            assert c1.dtype == np.uint64 and c1.shape == (ell, self.n) and c1.flags.c_contiguous
        else:
            c0 = np.zeros((ell, self.n), dtype=np.uint64)
            c1 = np.zeros((ell, self.n), dtype=np.uint64)
        _ck(_lib.fheb200_ct_export(self.h, ct.h, c0.ctypes.data_as(C.c_void_p), c1.ctypes.data_as(C.c_void_p),
                                   C.c_size_t(c0.size)))
        return c0, c1

    def ct_import(self, c0, c1, nsd, scale, slots):
        c0 = np.ascontiguousarray(c0, dtype=np.uint64)
        c1 = np.ascontiguousarray(c1, dtype=np.uint64)
        h = C.c_void_p()
        _ck(_lib.fheb200_ct_import(self.h, c0.ctypes.data_as(C.c_void_p), c1.ctypes.data_as(C.c_void_p), c0.shape[0],
                                   nsd, C.c_double(scale), slots, C.byref(h)))
        return Ciphertext(h)

    def ct_import_async(self, c0, c1, nsd, scale, slots):
        """H2D on the copy stream from PINNED uint64 [limbs][n] arrays; call ct_ready(ct) before the first use of the
        result, and leave c0/c1 untouched until the next sync()."""
        assert c0.dtype == np.uint64 and c0.flags.c_contiguous and c1.dtype == np.uint64 and c1.flags.c_contiguous
        h = C.c_void_p()
        _ck(_lib.fheb200_ct_import_async(self.h, c0.ctypes.data_as(C.c_void_p), c1.ctypes.data_as(C.c_void_p), c0.shape[0],
                                         nsd, C.c_double(scale), slots, C.byref(h)))
        return Ciphertext(h)

    def ct_ready(self, ct):
        _ck(_lib.fheb200_ct_ready(self.h, ct.h))

    def ct_export_async(self, ct, out):
        """D2H on the copy stream into PINNED (c0, c1); the arrays are valid after the next sync()."""
        c0, c1 = out
        ell = ct.limbs
        assert c0.dtype == np.uint64 and c0.shape == (ell, self.n) and c0.flags.c_contiguous
        assert c1.dtype == np.uint64 and c1.shape == (ell, self.n) and c1.flags.c_contiguous
        _ck(_lib.fheb200_ct_export_async(self.h, ct.h, c0.ctypes.data_as(C.c_void_p), c1.ctypes.data_as(C.c_void_p),
                                         C.c_size_t(c0.size)))

    def pt_export(self, pt, ell):
        out = np.zeros((ell, self.n), dtype=np.uint64)
        _ck(_lib.fheb200_pt_export(self.h, pt.h, out.ctypes.data_as(C.c_void_p), C.c_size_t(out.size)))
        return out

    def keys_export(self, keys, which):
        limbs = self.L + self.K if which == 0 else self.L
        out = np.zeros((limbs, self.n), dtype=np.uint64)
        _ck(_lib.fheb200_keys_export(self.h, keys.h, which, out.ctypes.data_as(C.c_void_p), C.c_size_t(out.size)))
        return out

    def evk_export(self, galois):
        out = np.zeros((self.beta(self.L), 2, self.L + self.K, self.n), dtype=np.uint64)
        _ck(_lib.fheb200_evk_export(self.h, C.c_uint32(galois), out.ctypes.data_as(C.c_void_p), C.c_size_t(out.size)))
        return out

    def raw_ntt(self, qp, poly, inverse=False):
        a = np.ascontiguousarray(poly.copy(), dtype=np.uint64)
        _ck(_lib.fheb200_raw_ntt(self.h, 1 if inverse else 0, qp, a.ctypes.data_as(C.c_void_p)))
        return a

    def raw_automorphism(self, galois, a):
        a = np.ascontiguousarray(a, dtype=np.uint64)
        out = np.empty_like(a)
        _ck(_lib.fheb200_raw_automorphism(self.h, C.c_uint32(galois), a.shape[0], a.ctypes.data_as(C.c_void_p),
                                          out.ctypes.data_as(C.c_void_p)))
        return out

    def raw_modup(self, c):
        c = np.ascontiguousarray(c, dtype=np.uint64)
        ell = c.shape[0]
        out = np.zeros((self.beta(ell), ell + self.K, self.n), dtype=np.uint64)
        _ck(_lib.fheb200_raw_modup(self.h, ell, c.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p),
                                   C.c_size_t(out.size)))
        return out

    def raw_rescale(self, a):
        a = np.ascontiguousarray(a, dtype=np.uint64)
        out = np.zeros((a.shape[0] - 1, self.n), dtype=np.uint64)
        _ck(_lib.fheb200_raw_rescale(self.h, a.shape[0], a.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p)))
        return out

    def raw_lincomb(self, inputs, weights, variant=0):
        """inputs [count][limbs2][n], weights [G][count][ell] residues -> [G][limbs2][n] (PS leaf sums, one kernel)."""
        a = np.ascontiguousarray(inputs, dtype=np.uint64)
        w = np.ascontiguousarray(weights, dtype=np.uint64)
        count, limbs2 = a.shape[0], a.shape[1]
        G, ell = w.shape[0], w.shape[2]
        out = np.zeros((G, limbs2, self.n), dtype=np.uint64)
        _ck(_lib.fheb200_raw_lincomb(self.h, ell, limbs2, count, G, a.ctypes.data_as(C.c_void_p), w.ctypes.data_as(C.c_void_p),
                                     out.ctypes.data_as(C.c_void_p), variant))
        return out

    def raw_mul_pt(self, a0, a1, pt, acc=None):
        a0, a1, pt = (np.ascontiguousarray(x, dtype=np.uint64) for x in (a0, a1, pt))
        o0 = np.ascontiguousarray(acc[0], dtype=np.uint64).copy() if acc is not None else np.zeros_like(a0)
        o1 = np.ascontiguousarray(acc[1], dtype=np.uint64).copy() if acc is not None else np.zeros_like(a0)
        vp = lambda x: x.ctypes.data_as(C.c_void_p)  # noqa: E731
        _ck(_lib.fheb200_raw_mul_pt(self.h, a0.shape[0], vp(a0), vp(a1), vp(pt), vp(o0), vp(o1), 1 if acc is not None else 0))
        return o0, o1

    def raw_tensor(self, a0, a1, b0=None, b1=None):
        vp = lambda x: x.ctypes.data_as(C.c_void_p) if x is not None else None  # noqa: E731
        a0, a1 = np.ascontiguousarray(a0, dtype=np.uint64), np.ascontiguousarray(a1, dtype=np.uint64)
        if b0 is not None:
            b0, b1 = np.ascontiguousarray(b0, dtype=np.uint64), np.ascontiguousarray(b1, dtype=np.uint64)
        d = [np.zeros_like(a0) for _ in range(3)]
        _ck(_lib.fheb200_raw_tensor(self.h, a0.shape[0], vp(a0), vp(a1), vp(b0), vp(b1), vp(d[0]), vp(d[1]), vp(d[2])))
        return d

    def raw_dh_accumulate(self, ell, acc, cP, pt, S, galois):
        vp = lambda x: x.ctypes.data_as(C.c_void_p) if x is not None else None  # noqa: E731
        acc = np.ascontiguousarray(acc, dtype=np.uint64) if acc is not None else None
        cP, pt = np.ascontiguousarray(cP, dtype=np.uint64), np.ascontiguousarray(pt, dtype=np.uint64)
        S = np.ascontiguousarray(S, dtype=np.uint64).copy()
        _ck(_lib.fheb200_raw_dh_accumulate(self.h, ell, vp(acc), vp(cP), vp(pt), vp(S), C.c_uint32(galois)))
        return S

    def raw_inner_product(self, ell, galois, digits):
        d = np.ascontiguousarray(digits, dtype=np.uint64)
        acc = np.zeros((2, ell + self.K, self.n), dtype=np.uint64)
        _ck(_lib.fheb200_raw_inner_product(self.h, ell, C.c_uint32(galois), d.ctypes.data_as(C.c_void_p),
                                           acc.ctypes.data_as(C.c_void_p)))
        return acc

    def raw_moddown(self, ell, acc):
        a = np.ascontiguousarray(acc, dtype=np.uint64)
        o0, o1 = np.zeros((ell, self.n), dtype=np.uint64), np.zeros((ell, self.n), dtype=np.uint64)
        _ck(_lib.fheb200_raw_moddown(self.h, ell, a.ctypes.data_as(C.c_void_p), o0.ctypes.data_as(C.c_void_p),
                                     o1.ctypes.data_as(C.c_void_p)))
        return o0, o1

    # -- device timing loops
    def bench_rotate(self, ct, index, reps):
        ms = C.c_float()
        _ck(_lib.fheb200_bench_rotate(self.h, ct.h, index, reps, C.byref(ms)))
        return ms.value

    def bench_rotations(self, ct, indices, reps, hoisted=False):
        ms = C.c_float()
        arr = (C.c_int * len(indices))(*indices)
        _ck(_lib.fheb200_bench_rotations(self.h, ct.h, arr, C.c_size_t(len(indices)), reps, 1 if hoisted else 0,
                                         C.byref(ms)))
        return ms.value

    def bench_kernel(self, name, ell, reps):
        ms, by = C.c_float(), C.c_double()
        _ck(_lib.fheb200_bench_kernel(self.h, name.encode(), ell, reps, C.byref(ms), C.byref(by)))
        return ms.value, by.value
