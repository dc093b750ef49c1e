"""GPU: the C++ adapters of dafs_b200/host LINKED and RUN against the reference's own classes
(tests/cxx/adapter_driver.cpp; the binary is built in the development container by oracle/Makefile
into oracle/_ref/, where the reference's headers and sources are, and travels to the GPU box like
the other built files).  CudaProbCons == ProbCons, CudaNussinov == SparseNussinov,
CudaNeedlemanWunsch == SparseNeedlemanWunsch, dafs_cuda::solve_by_dd == DAFS::solve_by_dd on
RF00005, bit for bit."""
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DRIVER = os.path.join(ROOT, "oracle", "_ref", "adapter_driver")


@pytest.mark.skipif(not os.path.exists(DRIVER), reason="oracle/_ref/adapter_driver not built (needs /root/reference at build time)")
@pytest.mark.parametrize("fasta", ["RF00005_0.fa", "RF00017_4.fa"])
def test_linked_adapters_equal_reference_classes(fasta):
    r = subprocess.run([DRIVER, os.path.join(ROOT, "tests", "golden", fasta)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.startswith("adapters ok:"), r.stdout
