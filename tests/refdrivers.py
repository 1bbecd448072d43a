"""Loader of oracle/_ref/libref_sort.so: the reference's own src/*.cpp, src/k-way/*.cpp and src/mehp24/*.cpp compiled
unmodified against the lbcrypto shim (oracle/Makefile `ref`; exists only where /root/reference was present at build time).
Test infrastructure: the product package does not know this library."""
import ctypes as C
import os

import pytest

PATH = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "libref_sort.so")
_lib = None


def available():
    return os.path.exists(PATH)


def library():
    """The reference drivers' C ABI (same entry points as libfhesort.so); skips the calling test when not built."""
    global _lib
    if not available():
        pytest.skip("oracle/_ref/libref_sort.so not built (no reference checkout on the build machine)")
    if _lib is None:
        L = C.CDLL(PATH)
        L.fhesort_last_error.restype = C.c_char_p
        _lib = L
    return _lib


_OR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle")
CPU_PATHS = {"reference": os.path.join(_OR, "_ref", "libref_sort_cpu.so"), "own": os.path.join(_OR, "libfhesort_cpu.so")}
_cpu = {}


def cpu_library(which="own"):
    """The sort drivers linked against oracle/libfheb200_cpu.so (oracle/cpu_scheme.cpp: the C ABI of include/fhe_b200.h on
    the host cores, no CUDA anywhere): which = "own" -> this repository's host/*.cpp, "reference" -> the reference's
    unmodified src/*.cpp.  An encrypted sort that shares no kernel with the product."""
    path = CPU_PATHS[which]
    if not os.path.exists(path):
        pytest.skip("%s not built" % os.path.relpath(path, os.path.dirname(_OR)))
    if which not in _cpu:
        L = C.CDLL(path)
        L.fhesort_last_error.restype = C.c_char_p
        _cpu[which] = L
    return _cpu[which]
