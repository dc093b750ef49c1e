"""GPU test of the N > 1 route bench.py uses, with two real ranks (torchrun, NCCL): the gathered
CSR (library NCCL gather-v to rank 0, and per-rank delivery into shared host memory) equals the
one-rank CSR byte for byte.  tools/multirank_check.py does the work; skipped with < 2 devices."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_two_ranks_gathered_csr_equals_one_rank_csr():
    import dafs_b200
    if dafs_b200.device_count() < 2:
        pytest.skip("needs two visible devices")
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", str(port),
                        os.path.join(ROOT, "tools", "multirank_check.py")],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "multirank ok: 2 ranks" in r.stdout
