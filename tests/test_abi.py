"""CPU tests of the boundary: the library loads, exports every symbol that
include/dafs_cuda.h declares, and refuses to compute without a GPU (no CPU
fallback).  No compute calls here."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "dafs_cuda.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(dafs_cuda_\w+)\s*\(", src)))


def test_header_declares_something():
    assert len(declared_symbols()) >= 10


def test_library_exports_every_declared_symbol():
    import dafs_b200
    L = dafs_b200.lib()
    for sym in declared_symbols():
        assert hasattr(L, sym), "libdafs_cuda.so does not export " + sym


def test_no_cpu_fallback_without_gpu():
    import dafs_b200
    if dafs_b200.device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(dafs_b200.DafsCudaError) as e:
        dafs_b200.pairhmm_all([b"ACGU", b"ACGA"])
    assert e.value.code == -2      # DAFS_CUDA_ENODEV
    with pytest.raises(dafs_b200.DafsCudaError):
        dafs_b200.init(0)


def test_product_does_not_reference_oracle():
    """Nothing under dafs_b200/ or include/ may name the oracle."""
    bad = []
    for base in ("dafs_b200", "include"):
        for dp, _, fns in os.walk(os.path.join(ROOT, base)):
            for fn in fns:
                if fn.endswith((".so", ".o", ".log", ".pyc")):
                    continue
                txt = open(os.path.join(dp, fn), errors="replace").read()
                if re.search(r"oracle/|libdafs_oracle|libdafs_ref|oracle_util", txt):
                    bad.append(os.path.join(dp, fn))
    assert not bad, bad
