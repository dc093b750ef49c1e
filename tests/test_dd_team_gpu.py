"""GPU: the latency configuration of the DD kernel (a team of 2 or 3 warps per Nussinov
structure, pipelined over row blocks) gives the same bits as one warp per structure and as
the oracle.  The team size is normally chosen from the batch size; DAFS_DD_TEAM forces it."""
import os

import numpy as np
import pytest

import oracle_util as ou
import dafs_b200 as db

pytestmark = pytest.mark.gpu


def solve(probs, team):
    old = os.environ.get("DAFS_DD_TEAM")
    os.environ["DAFS_DD_TEAM"] = str(team)
    try:
        plan = db.DDPlan(probs, trace=True)
        plan.run(); plan.sync()
        res = plan.fetch()
        info = plan.info()
        plan.close()
    finally:
        if old is None:
            del os.environ["DAFS_DD_TEAM"]
        else:
            os.environ["DAFS_DD_TEAM"] = old
    return res, info


@pytest.mark.parametrize("length,family,mutation", [(40, 1, 0.3), (73, 1, 0.75), (120, 2, 0.4), (150, 3, 0.5), (210, 4, 0.6)])
def test_team_equals_single_warp_and_oracle(length, family, mutation):
    probs = []
    for k in range(3):
        p_x, p_y, p_z = ou.dd_problem(length, family=family + k, mutation=mutation, mix=True)
        probs.append(dict(p_x=p_x, p_y=p_y, p_z=p_z, t_max=120))
    base, info1 = solve(probs, 1)
    assert info1["block"] == 96
    for team in (2, 3):
        got, info = solve(probs, team)
        assert info["block"] == 32 * (2 * team + 1)
        for g, b in zip(got, base):
            assert g["iters"] == b["iters"] and g["violated"] == b["violated"]
            assert g["score"].view(np.uint32) == b["score"].view(np.uint32)
            assert all(np.array_equal(g[k], b[k]) for k in "xyz")
            assert np.array_equal(g["trace"].view(np.uint32), b["trace"].view(np.uint32))
    o = ou.oracle_dd(probs[0]["p_x"], probs[0]["p_y"], probs[0]["p_z"], t_max=120)
    assert base[0]["iters"] == o["iters"] and base[0]["score"].view(np.uint32) == o["score"].view(np.uint32)
    assert all(np.array_equal(base[0][k], o[k]) for k in "xyz")


def test_team_mode_is_repeatable():
    """compute-sanitizer is closed on the B200 pool (profiles/r02_sanitizer.txt), so the race check
    of the team pipeline (progress flags in shared memory) is run-to-run identity: 40 runs per team
    size of the same batch, every bit of every result equal to the first run and to one warp."""
    probs = []
    for k in range(6):
        p_x, p_y, p_z = ou.dd_problem(100 + 20 * k, family=5 + k, mutation=0.5, mix=True)
        probs.append(dict(p_x=p_x, p_y=p_y, p_z=p_z, t_max=60))
    base, _ = solve(probs, 1)
    for team in (2, 3):
        old = os.environ.get("DAFS_DD_TEAM")
        os.environ["DAFS_DD_TEAM"] = str(team)
        try:
            plan = db.DDPlan(probs, trace=True)
            for rep in range(40):
                plan.run(); plan.sync()
                got = plan.fetch()
                for g, b in zip(got, base):
                    assert g["iters"] == b["iters"] and g["violated"] == b["violated"], (team, rep)
                    assert g["score"].view(np.uint32) == b["score"].view(np.uint32), (team, rep)
                    assert all(np.array_equal(g[k], b[k]) for k in "xyz"), (team, rep)
                    assert np.array_equal(g["trace"].view(np.uint32), b["trace"].view(np.uint32)), (team, rep)
            plan.close()
        finally:
            if old is None:
                del os.environ["DAFS_DD_TEAM"]
            else:
                os.environ["DAFS_DD_TEAM"] = old
