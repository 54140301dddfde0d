"""CPU: the committed fixtures still come out of the compiled reference, bit for bit.

``tests/golden/*.npz`` were written by ``tests/golden/make_golden.py`` from ``oracle/_ref/libmevo_ref.so`` (the unmodified
reference sources, ``oracle/Makefile``).  This test re-runs that library on every fixture's stored world, actions and initial
state and demands IDENTICAL arrays -- a fixture regenerated from a modified driver, a changed compiler flag or an edited
world builder shows up here.  Skipped where ``oracle/_ref`` was not built (it needs ``/root/reference``)."""
import numpy as np
import pytest

from helpers import Golden, golden_names
from oracle import ref as oref

pytestmark = pytest.mark.skipif(not oref.available(), reason="oracle/_ref/libmevo_ref.so not built")

NAMES = golden_names()


@pytest.mark.parametrize("name", NAMES)
def test_fixture_reproduces_from_reference(name):
    g = Golden(name)
    r = oref.RefSim(g.world)
    # topology
    assert r.num_points == g.P and r.num_edges == g.E
    ed = r.edges()
    mass, fixed = r.point_consts()
    for mine, key in ((r.get_state()[0], "pos0"), (mass, "mass"), (fixed, "fixed"), (ed["a"], "ea"), (ed["b"], "eb"), (ed["k"], "ek"),
                      (ed["l0"], "el0"), (ed["on_surface"], "on_surface")):
        assert np.array_equal(mine, g[key]), key
    for o in range(r.num_objects):
        assert np.array_equal(r.object_boxels(o), g[f"boxels_{o}"])
        assert np.array_equal(np.sort(r.object_surface_boxels(o)), g[f"surface_{o}"])
    assert np.array_equal(r.actuator_indices("robot"), g["act_idx"])
    # substep by substep through Environment::step
    p, v = g["init_pos"], g["init_vel"]
    r.set_state(p, v, p, v)
    cand_at = set(g["cand_at"].tolist())
    for k in range(g.n_sub):
        if k % 30 == 0:
            r.set_action("robot", g.action_rows(k // 30))
        if k in cand_at:
            assert np.array_equal(r.candidates(), g[f"cand_{k}"]), ("candidates", k)
        assert r.substep() == bool(g["sub_unstable"][k]), k
        pos, vel, _, vt = r.get_state()
        assert np.array_equal(pos, g["sub_pos"][k]), ("pos", k)
        assert np.array_equal(vel, g["sub_vel"][k]), ("vel", k)
        assert np.array_equal(vt, g["sub_vt"][k]), ("vel_true", k)
        assert np.array_equal(r.collision_forces(), g["sub_cf"][k]), ("collision_forces", k)
    # step by step through Sim::step on a fresh Sim
    r2 = oref.RefSim(g.world)
    if not np.array_equal(p, g["pos0"]) or np.any(v != 0):
        r2.set_state(p, v, p, v)
    for s in range(g.n_steps):
        r2.set_action("robot", g.action_rows(s))
        assert r2.step() == bool(g["step_unstable"][s]), s
        assert r2.get_time() == int(g["step_time"][s]), s
        pos, vel, _, _ = r2.get_state()
        assert np.array_equal(pos, g["step_pos"][s + 1]) and np.array_equal(vel, g["step_vel"][s + 1]), s
        assert r2.object_orientation_at_time(r2.get_time(), "robot") == g["step_orient"][s], s
