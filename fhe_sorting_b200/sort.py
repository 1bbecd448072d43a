"""ctypes binding of libfhesort.so (include/fhe_sort.h): the C++ sort drivers (host/sort_algo.h ...) over the CUDA engine."""
import ctypes as C
import os

import numpy as np

from .lib import FheError

_HERE = os.path.dirname(os.path.abspath(__file__))
SORT_LIB_PATH = os.path.join(_HERE, "libfhesort.so")
_lib = None


def load_sort():
    """dlopen libfhesort.so (this repository's host drivers); raises when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(SORT_LIB_PATH):
            raise FheError("%s is missing: build it with `make -C fhe_sorting_b200`" % SORT_LIB_PATH)
        L = C.CDLL(SORT_LIB_PATH)
        L.fhesort_last_error.restype = C.c_char_p
        _lib = L
    return _lib


def _ck(rc, L=None):
    if rc != 0:
        raise FheError((L or load_sort()).fhesort_last_error().decode())


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def uniform_gap_input(n, seed=None):
    """tests/utils.h:28-51 getVectorWithMinDiff(N, 0, 1, 1/N): the multiples of 1/N, shuffled (seeded here)."""
    rng = np.random.default_rng(n if seed is None else seed)
    return rng.permutation(n).astype(np.float64) / n


class DirectSortSession:
    """What tests/DirectSortTest.cpp's fixture sets up: context, keys in HBM, DirectSort<N>."""

    _create = "fhesort_session_create"

    def __init__(self, n, log_ring_dim=17, mult_depth=0, seed=0, device=0, lib=None):
        """seed = 0: keys from OS entropy (ChaCha20); seed != 0: TEST-ONLY deterministic keys (needed for replicated keys
        across ranks and for comparisons with the oracle).  lib: another library with the same C ABI (tests pass the
        reference's own drivers built under oracle/_ref)."""
        L = self.L = lib or load_sort()
        h = C.c_void_p()
        _ck(getattr(L, self._create)(n, log_ring_dim, mult_depth, C.c_uint64(seed), device, C.byref(h)), L)
        self.h, self.n = h, n
        d, k, ms = C.c_int(), C.c_int(), C.c_double()
        _ck(L.fhesort_session_info(h, C.byref(d), C.byref(k), C.byref(ms)), L)
        self.mult_depth, self.num_rotation_keys, self.setup_ms = d.value, k.value, ms.value

    def close(self):
        if self.h:
            self.L.fhesort_session_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_streams(self, k):
        """Concurrent CUDA streams over the independent batches of a phase (default 4)."""
        _ck(self.L.fhesort_session_set_streams(self.h, int(k)), self.L)

    def set_shard(self, rank, world, native=True, host=False):
        """Shard the sort's batches over `world` ranks of an initialised torch.distributed group (one GPU each).
        native=True: the engine opens its own NCCL communicator (the 128-byte id of rank 0 travels through
        torch.distributed.broadcast_object_list) and all-reduces on its own stream; native=False: torch's all_reduce
        through a callback (round-1 path, kept for comparison)."""
        if world > 1 and native and not host:
            import torch.distributed as dist
            from .lib import load
            ident = [None]
            if rank == 0:
                buf = (C.c_ubyte * 128)()
                E = load()
                if E.fheb200_comm_unique_id(buf) != 0:
                    raise FheError(E.fheb200_last_error().decode())
                ident = [bytes(buf)]
            dist.broadcast_object_list(ident, src=0)
            _ck(self.L.fhesort_session_comm_init(self.h, (C.c_ubyte * 128)(*ident[0]), rank, world), self.L)
            self._cb = None
            return
        if world <= 1:
            _ck(self.L.fhesort_session_set_shard(self.h, 0, 1, None, None), self.L)
            self._cb = None
            return
        import torch
        import torch.distributed as dist

        class _View:  # a borrowed device buffer seen through __cuda_array_interface__
            def __init__(self, ptr, words):
                self.__cuda_array_interface__ = {"shape": (words,), "typestr": "<i8", "data": (ptr, False), "version": 3}

        def allreduce(user, ptr, words):
            try:
                if host:  # a library whose ciphertexts live in host memory (the CPU scheme layer under tests/): gloo on the words
                    a = np.ctypeslib.as_array((C.c_int64 * words).from_address(ptr))
                    dist.all_reduce(torch.from_numpy(a), op=dist.ReduceOp.SUM)
                    return 0
                t = torch.as_tensor(_View(ptr, words), device="cuda")
                dist.all_reduce(t, op=dist.ReduceOp.SUM)
                torch.cuda.current_stream().synchronize()
                return 0
            except Exception as e:  # noqa: BLE001 - reported through the C ABI status
                print("all-reduce failed:", e)
                return 1

        self._cb = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_size_t)(allreduce)
        _ck(self.L.fhesort_session_set_shard(self.h, rank, world, self._cb, None), self.L)

    def _run(self, stage, values, aux=None, sign=(0, 0, 0)):
        v = np.ascontiguousarray(values, dtype=np.float64)
        assert v.shape == (self.n,)
        out = np.zeros(self.n)
        a = np.ascontiguousarray(aux, dtype=np.float64) if aux is not None else None
        ms, lv = C.c_double(), C.c_int()
        _ck(self.L.fhesort_run(self.h, stage, _dp(v), _dp(a) if a is not None else None, sign[0], sign[1], sign[2], _dp(out),
                               C.byref(ms), C.byref(lv)), self.L)
        return out, ms.value, lv.value

    def sort(self, values, sign=(0, 0, 0)):
        """DirectSort<N>::sort; returns (sorted values, elapsed ms incl. encrypt/decrypt, final level)."""
        return self._run(0, values, None, sign)

    def sort_hybrid(self, values, sign=(0, 0, 0)):
        """DirectSort<N>::sort_hybrid (rank + MEHP24-style placement); needs a HybridSortSession."""
        return self._run(3, values, None, sign)

    def rank(self, values, sign=(0, 0, 0)):
        return self._run(1, values, None, sign)

    def place(self, values, ranks):
        return self._run(2, values, ranks)

    def composed_rotate(self, values, rotation):
        v = np.ascontiguousarray(values, dtype=np.float64)
        out = np.zeros(len(v))
        _ck(self.L.fhesort_composed_rotate(self.h, _dp(v), len(v), rotation, _dp(out)), self.L)
        return out


def rotation_indices(n, log_ring_dim=17):
    L = load_sort()
    out = (C.c_int * 4096)()
    cnt = C.c_int()
    _ck(L.fhesort_rotation_indices(n, log_ring_dim, out, 4096, C.byref(cnt)))
    return list(out[:cnt.value])


def decompose(keys, rotation, wrap, algo="BINARY", lib=None):
    """Decomposer<1024>::decompose: the signed rotation amounts, largest first (src/rotation.h:53-102)."""
    L = lib or load_sort()
    k = (C.c_int * len(keys))(*keys)
    out = (C.c_int * 64)()
    cnt = C.c_int()
    _ck(L.fhesort_decompose(k, len(keys), rotation, wrap, {"NAF": 0, "BNAF": 1, "BINARY": 2}[algo], out, 64, C.byref(cnt)), L)
    return list(out[:cnt.value])


def doubled_sinc_coefficients(n):
    L = load_sort()
    buf = np.zeros(16384)
    cnt = C.c_int()
    _ck(L.fhesort_doubled_sinc_coefficients(n, _dp(buf), len(buf), C.byref(cnt)))
    return buf[:cnt.value].copy()


def sinc_coefficients(n):
    L = load_sort()
    buf = np.zeros(16384)
    cnt = C.c_int()
    _ck(L.fhesort_sinc_coefficients(n, _dp(buf), len(buf), C.byref(cnt)))
    return buf[:cnt.value].copy()


def shard_batches(n, rank, world, log_ring_dim=17):
    """(batches evaluated by `rank`, whether the sort is sharded at all) — host-only mirror of DirectSort::setShard."""
    L = load_sort()
    out = (C.c_int * 4096)()
    cnt, sh = C.c_int(), C.c_int()
    _ck(L.fhesort_shard_batches(n, log_ring_dim, rank, world, out, 4096, C.byref(cnt), C.byref(sh)))
    return list(out[:cnt.value]), bool(sh.value)


def combine_partial_residues(partial, moduli, dist=None):
    """Sum per-rank partial residue arrays [limbs][n] (uint64 < q) over a torch.distributed group and bring the sums
    back to [0, q): the host-side statement of AllReduceSum + fheb200_ct_reduce_partial_sums (used by the gloo tests)."""
    import torch
    t = torch.from_numpy(np.ascontiguousarray(partial).view(np.int64).copy())
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    s = t.numpy().view(np.uint64)
    q = np.asarray(moduli, dtype=np.uint64).reshape(-1, 1)
    return s % q


class Mehp24SortSession(DirectSortSession):
    """tests/mehp24/Mehp24SortTest.cpp's fixture: N values in the first row of an N x N slot matrix, mehp24::sortFG
    (sortLargeArrayFG above 256 values).  Only sort() applies."""
    _create = "fhesort_mehp24_session_create"


class KWaySortSession(DirectSortSession):
    """tests/k-way/KWaySort2Test.cpp's fixture: the HKC+21 k-way sorting network on N = k^M values with bootstrapping
    (depth 40, first modulus 60, scaling 59 bits, EvalBootstrapSetup + KeyGen).  Only sort() applies."""

    def __init__(self, k, m, log_ring_dim=17, mult_depth=0, seed=0, device=0, lib=None):
        L = self.L = lib or load_sort()
        h = C.c_void_p()
        _ck(L.fhesort_kway_session_create(k, m, log_ring_dim, mult_depth, C.c_uint64(seed), device, C.byref(h)), L)
        self.h, self.n, self.k, self.m = h, k ** m, k, m
        d, nk, ms = C.c_int(), C.c_int(), C.c_double()
        _ck(L.fhesort_session_info(h, C.byref(d), C.byref(nk), C.byref(ms)), L)
        self.mult_depth, self.num_rotation_keys, self.setup_ms = d.value, nk.value, ms.value

    def bootstraps(self):
        c = C.c_int()
        _ck(self.L.fhesort_session_bootstraps(self.h, C.byref(c)), self.L)
        return c.value


class HybridSortSession(DirectSortSession):
    """tests/DirectSortHTest.cpp's fixture: DirectSort<N> with the depth and rotation keys sort_hybrid() needs."""
    _create = "fhesort_hybrid_session_create"


class BitonicSortSession(DirectSortSession):
    """tests/BitonicSortTest.cpp's fixture: BitonicSort<N> on values in [0, 255], depth 58, bootstrapping above level 29."""
    _create = "fhesort_bitonic_session_create"

    def bootstraps(self):
        c = C.c_int()
        _ck(self.L.fhesort_session_bootstraps(self.h, C.byref(c)), self.L)
        return c.value


def kway_stage(k, m_total, stage, num_slots, lib=None):
    """(m, logDist, slope, rotate distance), indices[2][num_slots] of one stage of the k-way network (host only)."""
    L = lib or load_sort()
    geo = (C.c_int * 4)()
    i0 = np.zeros(num_slots, dtype=np.int32)
    i1 = np.zeros(num_slots, dtype=np.int32)
    _ck(L.fhesort_kway_stage(k, m_total, stage, num_slots, geo, i0.ctypes.data_as(C.c_void_p), i1.ctypes.data_as(C.c_void_p)), L)
    return tuple(geo), np.stack([i0, i1])


def kway_num_stages(k, m_total):
    return m_total + m_total * (m_total - 1) // 2 * ((k + 1) // 2)


SIGN_FUNCS = {"CompositeSign": 0, "SignumPolycircuit": 1, "Tanh": 2, "NaiveDiscrete": 3}
_ALGOS = {"NAF": 0, "BNAF": 1, "BINARY": 2}


class ToolkitSession(DirectSortSession):
    """Fixture of the reference's primitive-level tests (tests/RotationTest.cpp, SignTest.cpp, tests/k-way/SorterTest.cpp,
    utils/sign_characterizer.cpp): a context with the given depth / scaling size / batch / rotation keys and, optionally,
    bootstrapping with level budget {b, b}; no sorter."""

    def __init__(self, log_ring_dim, mult_depth, scaling_mod_size, batch, rotation_keys=(), bootstrap_budget=0, seed=0, device=0,
                 lib=None):
        L = self.L = lib or load_sort()
        h = C.c_void_p()
        keys = (C.c_int * max(1, len(rotation_keys)))(*rotation_keys)
        _ck(L.fhesort_toolkit_session_create(log_ring_dim, mult_depth, scaling_mod_size, batch, keys, len(rotation_keys),
                                             bootstrap_budget, C.c_uint64(seed), device, C.byref(h)), L)
        self.h, self.n, self.mult_depth = h, batch, mult_depth

    def sign(self, func, values, n=0, dg=0, df=0):
        """sign(x, cc, func, cfg) of src/sign.cpp:635-651 -> (values, final level, milliseconds)"""
        v = np.ascontiguousarray(values, dtype=np.float64)
        out = np.zeros(len(v))
        lv, ms = C.c_int(), C.c_double()
        _ck(self.L.fhesort_sign_eval(self.h, SIGN_FUNCS[func], n, dg, df, _dp(v), len(v), _dp(out), C.byref(lv), C.byref(ms)), self.L)
        return out, lv.value, ms.value

    def tree_rotate(self, n_template, values, start, end, rotations, algo="NAF"):
        """RotationTree<N>::buildTree(start, end), then treeRotate for every requested rotation -> [len(rotations)][count]"""
        v = np.ascontiguousarray(values, dtype=np.float64)
        rots = (C.c_int * len(rotations))(*rotations)
        out = np.zeros((len(rotations), len(v)))
        _ck(self.L.fhesort_tree_rotate(self.h, n_template, _ALGOS[algo], start, end, _dp(v), len(v), rots, len(rotations), _dp(out)), self.L)
        return out

    def kway_unit(self, op, k, m, in0, in1=None, in2=None, indices=None, param=0, aux=0):
        """One kwaySort method on fresh encryptions (op codes: host/sort_cabi.cpp fhesort_kway_unit) -> (out, level before, after)"""
        a = [None if x is None else np.ascontiguousarray(x, dtype=np.float64) for x in (in0, in1, in2)]
        count = len(a[0])
        out = np.zeros(2 * count)
        i0 = i1 = None
        if indices is not None:
            i0 = np.ascontiguousarray(indices[0], dtype=np.int32)
            i1 = np.ascontiguousarray(indices[1], dtype=np.int32)
        lb, la = C.c_int(), C.c_int()
        ip = lambda x: None if x is None else x.ctypes.data_as(C.c_void_p)  # noqa: E731
        _ck(self.L.fhesort_kway_unit(self.h, op, C.c_long(k), C.c_long(m), C.c_long(param), aux, _dp(a[0]),
                                     None if a[1] is None else _dp(a[1]), None if a[2] is None else _dp(a[2]), ip(i0), ip(i1),
                                     count, _dp(out), C.byref(lb), C.byref(la)), self.L)
        return (out if op == 11 else out[:count]), lb.value, la.value
