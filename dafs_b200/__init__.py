"""dafs_b200 -- B200-native (sm_100a) engines for the DAFS hot path.

The product is ``libdafs_cuda.so`` (C ABI in ``include/dafs_cuda.h``) plus the
C++ adapters in ``dafs_b200/host`` that plug it behind the reference's
``Align::Model`` / ``Fold::Model`` interfaces.  This Python package is only the
thin ctypes binding the tests and ``bench.py`` drive it through; there is no
CPU fallback anywhere: if the library or a GPU is missing, calls raise.
"""
import os as _os

# Must be in the environment before the CUDA context exists (torch creates it lazily): the
# progressive stage keeps many launches in flight on separate streams, see csrc/common.cu.
_os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

from .capi import (  # noqa: F401
    DafsCudaError,
    CsrBatch,
    DeviceCsrBatch,
    PairHmmPlan,
    all_pairs,
    device_count,
    fp32_probe,
    fp64_probe,
    init,
    lib,
    library_path,
    pairhmm_all,
    pairhmm_pairs,
    pairhmm_tables,
)
from .capi_comm import Comm, HostCsr, comm_unique_id, pairhmm_pairs_into, plan_fetch_into  # noqa: F401
from .capi_mc import McCaskillPlan, mccaskill_all, mccaskill_constrained  # noqa: F401
from .capi_pct import pct, pct_last_ms  # noqa: F401
from .capi_prog import guide_tree, progressive  # noqa: F401
from .capi_dd import DDPlan, dd_solve_batch, nussinov_decode, nw_decode, nw_decode_static  # noqa: F401
from .synth import fasta_read, homolog_family, random_rna, splitmix64, synthetic_bpp  # noqa: F401
