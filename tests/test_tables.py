"""CPU tests that pin the constant tables the kernels embed against the reference's own source:
(a1) the log-space pair-HMM tables against the reference's ProbabilisticModel constructor
(src/probconsRNA/ProbabilisticModel.h:55-88, run inside oracle/_ref), and (a7) the BL* energy
blob against a fresh extraction from src/boltzmann_param.c."""
import ctypes as C
import filecmp
import os
import subprocess
import sys

import numpy as np
import pytest

import oracle_util as ou

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_PARAM = "/root/reference/src/boltzmann_param.c"


def ref_tables():
    L = C.CDLL(ou.REF_SO)
    out = np.zeros(3 + 9 + 256 + 256 * 256, np.float32)
    L.ref_probcons_tables.argtypes = [C.c_void_p]
    L.ref_probcons_tables(out.ctypes.data)
    return out[:3], out[3:12], out[12:268], out[268:].reshape(256, 256)


@pytest.mark.skipif(not ou.have_ref(), reason="oracle/_ref not built")
def test_pairhmm_tables_equal_reference_constructor():
    import dafs_b200
    init, trans, ins, match = ref_tables()
    t = dafs_b200.pairhmm_tables()                      # host tables, no device needed
    bits = lambda a: np.asarray(a, np.float32).view(np.uint32)
    assert np.array_equal(bits(t[:3]), bits(init))
    # the recurrences read T[k][0], T[0][s], T[s][s]; T[1][2] = T[2][1] = log 0
    assert np.array_equal(bits(t[3:12]), bits(trans))
    # six symbol classes stand for the 256 bytes: A C G U/T N other
    reps = [b"Aa", b"Cc", b"Gg", b"UuTt", b"Nn"]
    cls = np.full(256, 5)
    for k, chars in enumerate(reps):
        for ch in chars:
            cls[ch] = k
    lib_ins, lib_match = t[12:18], t[18:54].reshape(6, 6)
    assert np.array_equal(bits(lib_ins[cls]), bits(ins))
    assert np.array_equal(bits(lib_match[cls][:, cls]), bits(match))


def test_oracle_restatement_tables_equal_library_tables():
    """The C restatement's tables (7 classes: U and T apart) carry the same values."""
    import dafs_b200

    class T(C.Structure):
        _fields_ = [("init", C.c_float * 3), ("trans", C.c_float * 9), ("ins", C.c_float * 7),
                    ("match", C.c_float * 49)]
    o = T()
    ou.oracle().oracle_pairhmm_tables(C.byref(o))
    t = dafs_b200.pairhmm_tables()
    m = {0: 0, 1: 1, 2: 2, 3: 3, 4: 3, 5: 4, 6: 5}      # oracle class -> library class
    assert np.array_equal(np.array(o.init, np.float32).view(np.uint32), t[:3].view(np.uint32))
    assert np.array_equal(np.array(o.trans, np.float32).view(np.uint32), t[3:12].view(np.uint32))
    om = np.array(o.match, np.float32).reshape(7, 7)
    for a in range(7):
        assert np.float32(o.ins[a]).view(np.uint32) == t[12 + m[a]].view(np.uint32)
        for b in range(7):
            assert om[a, b].view(np.uint32) == t[18 + 6 * m[a] + m[b]].view(np.uint32)


@pytest.mark.skipif(not os.path.exists(REF_PARAM), reason="reference tree absent")
def test_blstar_blob_is_a_fresh_extraction_of_the_reference_tables(tmp_path):
    out = str(tmp_path / "blstar37.i32")
    subprocess.check_call([sys.executable, os.path.join(ROOT, "tools", "extract_blstar.py"), REF_PARAM, out],
                          stdout=subprocess.DEVNULL)
    assert filecmp.cmp(out, os.path.join(ROOT, "dafs_b200", "data", "blstar37.i32"), shallow=False)
