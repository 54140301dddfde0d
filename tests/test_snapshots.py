"""GPU: batched snapshot history (mevo_batch_save_snapshot / mevo_batch_revert / mevo_env_read_snapshot) against the
reference's Environment::save_snapshot / revert_to_snapshot / Sim::revert (Environment.cpp:239-263, Sim.cpp:126-133), run live
through oracle/_ref: same state after the revert (pos, vel, pos_last = pos, vel_true = vel, time), actuator goals NOT reverted,
same trajectory afterwards."""
import numpy as np
import pytest

from helpers import rel_err
from modularevogym_b200 import capi, workloads
from oracle import ref as oref

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not oref.available(), reason="oracle/_ref/libmevo_ref.so not built")]

CASES = [("walker5x5", 3), ("walker5x5", 4), ("pusher", 5), ("obstacle10", 6), ("bridge", 7)]


def test_snapshot_slots_follow_reference_revert(capfd):
    worlds = [workloads.world_for(n, i) for n, i in CASES]
    refs = [oref.RefSim(w) for w in worlds]
    B = capi.Batch(0)
    for w in worlds:
        B.add_world(w)
    B.finalize()
    rngs = [np.random.default_rng(50 + i) for i in range(len(CASES))]
    idx = [r.actuator_indices("robot") for r in refs]

    def step_all(tight=True):
        for i, r in enumerate(refs):
            rows = np.stack((idx[i].astype(np.float64), workloads.clipped_actions(rngs[i], idx[i].size)), axis=1)
            r.set_action("robot", rows)
            assert B.set_action(i, "robot", rows)
        _, flags = B.step(1, apply_actions=False)
        for i, r in enumerate(refs):
            assert bool(flags[i]) == r.step()
            assert rel_err(B.get_state(i)[0], r.get_state()[0]) <= (1e-9 if tight else 1e-6), (CASES[i], "trajectory")

    assert B.read_snapshot(0, 3) is None                       # empty slot
    for _ in range(3):
        step_all()
    B.save_snapshot(1)                                           # the reference saved history[3] by itself (Sim.cpp:112-113)
    snap = [r.get_state() for r in refs]
    for _ in range(2):
        step_all()
    B.save_snapshot(2)                                           # t = 5
    goals_before = [B.get_goal(i) for i in range(len(CASES))]
    # ---- read-back: Sim::pos_at_time(3) / vel_at_time(3)
    for i, r in enumerate(refs):
        pos, vel, t = B.read_snapshot(i, 1)
        assert t == 3
        assert rel_err(pos, r.pos_at_time(3)) <= 1e-9 and rel_err(vel, r.pos_at_time(3, vel=True)) <= 1e-7
        assert B.read_snapshot(i, 2)[2] == 5
    # ---- revert a subset to t = 3; the others keep going
    subset = [0, 2, 3]
    B.revert(1, subset)
    for i in subset:
        refs[i].revert(3)
    for i, r in enumerate(refs):
        pos, vel, last, vt = B.get_state(i)
        pr, vr, lr, vtr = r.get_state()
        assert B.get_time(i) == r.get_time() == (3 if i in subset else 5)
        assert rel_err(pos, pr) <= 1e-9 and rel_err(vel, vr) <= 1e-7
        if i in subset:
            assert np.array_equal(last, pos) and np.array_equal(vt, vel)            # pos_last = pos, vel_true = vel
            assert np.array_equal(lr, pr) and np.array_equal(vtr, vr)
            assert np.array_equal(B.get_goal(i), goals_before[i])                   # goals are not reverted
            assert rel_err(B.get_goal(i), r.edges()["goal"]) <= 1e-12
    for _ in range(2):
        step_all(tight=False)
    # ---- revert everything to slot 2 (t = 5), then to the reset snapshot (slot 0)
    B.revert(2)
    for i, r in enumerate(refs):
        r.revert(5) if i not in subset else None
    # environments of `subset` no longer have t = 5 in the reference (revert erased the later history, Environment.cpp:262);
    # the batch keeps slot 2 until it is overwritten -- compare the others with the reference, the subset with the stored copy
    for i, r in enumerate(refs):
        assert B.get_time(i) == 5
        if i not in subset:
            assert rel_err(B.get_state(i)[0], r.get_state()[0]) <= 1e-9
        else:
            assert np.array_equal(B.get_state(i)[0], B.read_snapshot(i, 2)[0])
    B.revert(0)
    for i, r in enumerate(refs):
        r.revert(0)
        assert B.get_time(i) == 0
        assert np.array_equal(B.get_state(i)[0], r.get_state()[0])
    # ---- empty slot: the reference's message, nothing changes
    capfd.readouterr()
    before = B.get_state(1)[0]
    B.revert(5)
    assert "no save found" in capfd.readouterr().out
    assert np.array_equal(B.get_state(1)[0], before)
    with pytest.raises(capi.MevoError):
        B.save_snapshot(8)
    B.close()
