"""The C-ABI library loads on a CPU-only box, exports every symbol include/fhe_b200.h declares, and refuses to
compute without a GPU (no CPU fallback)."""
import ctypes
import os
import re

import pytest

import fhe_sorting_b200 as F

HEADER = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "fhe_b200.h")


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(fheb200_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported():
    lib = F.load()
    names = declared_symbols()
    assert len(names) >= 50
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, "declared in fhe_b200.h but not exported: %s" % missing


def test_every_declared_sort_symbol_is_exported():
    from fhe_sorting_b200.sort import load_sort
    text = re.sub(r"/\*.*?\*/", "", open(HEADER.replace("fhe_b200.h", "fhe_sort.h")).read(), flags=re.S)
    names = sorted(set(re.findall(r"\b(fhesort_[a-z0-9_]+)\s*\(", text)) - {"fhesort_allreduce_fn"})
    lib = load_sort()
    assert len(names) >= 10
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, "declared in fhe_sort.h but not exported: %s" % missing


def test_version_and_error_strings():
    lib = F.load()
    assert lib.fheb200_version().decode().startswith("fheb200")
    assert isinstance(lib.fheb200_last_error(), bytes)


def test_null_arguments_are_errors_not_crashes():
    lib = F.load()
    assert lib.fheb200_ctx_create(None, None) != 0
    assert lib.fheb200_ct_release(None) in (0, 1)
    assert b"null" in lib.fheb200_last_error().lower() or lib.fheb200_last_error() is not None


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(F.FheError) as e:
        F.Context(10, 2, 40, 60)
    assert "no CUDA device" in str(e.value) or "CUDA" in str(e.value)


def test_params_chain_is_host_only_and_matches_oracle():
    from oracle.oracle import Oracle
    for logn, depth, sb in [(10, 6, 40), (12, 9, 50), (13, 5, 59), (17, 25, 40)]:
        mod, psi, nq, npp, alpha = F.params_chain(logn, depth, sb, 60, 3)
        orc = Oracle(logn, depth, sb, 60, 3)
        assert (nq, npp, alpha) == (orc.L, orc.K, orc.alpha)
        assert mod == orc.moduli
        assert psi == [orc.psi(i) for i in range(nq + npp)]
        n2 = 2 << logn
        for q, w in zip(mod, psi):
            assert q % n2 == 1 and pow(w, n2 // 2, q) == q - 1
        assert len(set(mod)) == len(mod)
