"""GPU: the device-side set-up of DD problems (dd_pack.cuh: CBP enumeration, projections, Nussinov
entry lists, NW envelope, band, CBP quadruples -- row f2 of SURVEY.md 8(f)) builds constant blocks
that are BYTE-IDENTICAL to the host packer's (DAFS_DD_PACK_VERIFY=1 makes the library compare them)
and therefore the same results as the oracle.  DAFS_DD_DEVICE_PACK=1 routes the host-pointer entry
point through the device path; dafs_cuda_progressive uses it natively."""
import os

import numpy as np
import pytest

import oracle_util as ou
import dafs_b200 as db

pytestmark = pytest.mark.gpu


@pytest.fixture()
def device_pack(monkeypatch):
    monkeypatch.setenv("DAFS_DD_DEVICE_PACK", "1")
    monkeypatch.setenv("DAFS_DD_PACK_VERIFY", "1")


def solve(probs, trace=False):
    plan = db.DDPlan(probs, trace=trace)
    plan.run(); plan.sync()
    res, info = plan.fetch(), plan.info()
    plan.close()
    return res, info


@pytest.mark.parametrize("length,family,mutation", [(1, 0, 0.2), (5, 1, 0.2), (40, 1, 0.3), (73, 2, 0.75), (150, 3, 0.5), (300, 4, 0.4)])
def test_device_pack_equals_host_pack_and_oracle(device_pack, length, family, mutation):
    probs = []
    for k in range(4):
        p_x, p_y, p_z = ou.dd_problem(length, family=family + k, mutation=mutation, mix=True, l2=max(1, length - 3 * k))
        probs.append(dict(p_x=p_x, p_y=p_y, p_z=p_z, t_max=80, n1=1 + k, n2=2))
    got, info = solve(probs, trace=True)
    assert info["packed_on_device"] == 1
    for g, p in zip(got, probs):
        o = ou.oracle_dd(p["p_x"], p["p_y"], p["p_z"], n1=p["n1"], n2=p["n2"], t_max=80)
        assert g["iters"] == o["iters"] and g["score"].view(np.uint32) == o["score"].view(np.uint32)
        assert all(np.array_equal(g[k], o[k]) for k in "xyz")


def test_host_path_is_the_default():
    p_x, p_y, p_z = ou.dd_problem(60, family=2)
    _, info = solve([dict(p_x=p_x, p_y=p_y, p_z=p_z, t_max=10)])
    assert info["packed_on_device"] == 0


def test_dense_and_empty_matrices(device_pack):
    r = np.random.default_rng(5)
    L1, L2 = 48, 40
    dense = dict(p_x=np.triu((0.02 + 0.5 * r.random((L1, L1))).astype(np.float32), 1),
                 p_y=np.triu((0.02 + 0.5 * r.random((L2, L2))).astype(np.float32), 1),
                 p_z=(0.02 + 0.3 * r.random((L1, L2))).astype(np.float32), t_max=5)
    empty = dict(p_x=np.zeros((L1, L1), np.float32), p_y=np.zeros((L2, L2), np.float32), p_z=np.zeros((L1, L2), np.float32), t_max=5)
    got, info = solve([dense, empty])
    assert info["packed_on_device"] == 1 and got[0]["ncbp"] > 1000 and got[1]["ncbp"] == 0
    for g, p in zip(got, (dense, empty)):
        o = ou.oracle_dd(p["p_x"], p["p_y"], p["p_z"], t_max=5)
        assert g["iters"] == o["iters"] and all(np.array_equal(g[k], o[k]) for k in "xyz")
