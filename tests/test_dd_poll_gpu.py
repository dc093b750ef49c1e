"""GPU: dafs_cuda_dd_plan_poll hands out each problem of a running DD batch as soon as its
completion flag is set, and what it hands out is what dafs_cuda_dd_plan_fetch returns after
the batch has finished.  dafs_cuda_progressive schedules guide-tree nodes on this call."""
import time

import numpy as np
import pytest

import oracle_util as ou
import dafs_b200 as db

pytestmark = pytest.mark.gpu


def same(a, b):
    return (a["iters"] == b["iters"] and a["violated"] == b["violated"] and a["status"] == b["status"]
            and a["score"].view(np.uint32) == b["score"].view(np.uint32)
            and all(np.array_equal(a[k], b[k]) for k in "xyz"))


def problems(n, lengths, t_maxes):
    probs = []
    for k in range(n):
        p_x, p_y, p_z = ou.dd_problem(lengths[k % len(lengths)], family=1 + k, mutation=0.3 + 0.1 * (k % 4), mix=True)
        probs.append(dict(p_x=p_x, p_y=p_y, p_z=p_z, t_max=t_maxes[k % len(t_maxes)]))
    return probs


@pytest.mark.parametrize("n", [1, 5, 40])
def test_poll_equals_fetch(n):
    probs = problems(n, [40, 90, 150], [5, 60, 200])
    with db.DDPlan(probs) as plan:
        plan.run()
        state, seen, polls = None, {}, 0
        deadline = time.time() + 60
        while len(seen) < n:
            assert time.time() < deadline, "poll never reported %d of %d problems" % (n - len(seen), n)
            state, fresh = plan.poll(state)
            assert not (set(fresh) & set(seen)), "a problem was reported twice"
            seen.update({k: {a: (b.copy() if isinstance(b, np.ndarray) else b) for a, b in r.items()} for k, r in fresh.items()})
            polls += 1
        state, fresh = plan.poll(state)                     # nothing is left
        assert fresh == {}
        plan.sync()
        full = plan.fetch()
    assert sorted(seen) == list(range(n))
    for k in range(n):
        assert same(seen[k], full[k]), k
    o = ou.oracle_dd(probs[0]["p_x"], probs[0]["p_y"], probs[0]["p_z"], t_max=probs[0]["t_max"])
    assert seen[0]["iters"] == o["iters"] and seen[0]["score"].view(np.uint32) == o["score"].view(np.uint32)
    assert all(np.array_equal(seen[0][k], o[k]) for k in "xyz")


def test_poll_after_rerun_starts_from_scratch():
    probs = problems(3, [60], [30])
    with db.DDPlan(probs) as plan:
        for _ in range(2):                                   # the completion flags are cleared by every run
            plan.run()
            state, seen = None, {}
            deadline = time.time() + 60
            while len(seen) < 3:
                assert time.time() < deadline
                state, fresh = plan.poll(state)
                seen.update(fresh)
            plan.sync()
            full = plan.fetch()
            assert all(same(seen[k], full[k]) for k in range(3))


def test_poll_before_run_is_an_error():
    probs = problems(1, [40], [5])
    with db.DDPlan(probs) as plan:
        with pytest.raises(Exception):
            plan.poll()
