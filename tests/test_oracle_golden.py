"""CPU: the oracle (oracle/mevo_oracle.c + oracle/topology.py) against the committed fixtures that the
UNMODIFIED reference engine produced (tests/golden/make_golden.py).  This is what pins the oracle."""
import numpy as np
import pytest

from helpers import Golden, golden_names, pair_set, rel_err
from oracle.oracle import OracleSim
from oracle.topology import build_topology

NAMES = golden_names()

# Teacher-forced (state reset to the reference's before every substep).  Without contacts the
# restatement is bit-identical; with contacts the only difference is the order in which contact
# forces are summed into collision_forces (tree-descent order vs ascending leaf order) -> a few ulp
# of the force, times dt^2 on the position.
TF_POS_TOL = 1e-14
TF_VEL_TOL = 1e-11
# Free-running 60 substeps: the same ulp-level differences, amplified by stiff contact dynamics.
FREE_POS_TOL = 1e-7


def test_fixtures_present():
    assert len(NAMES) >= 16


@pytest.mark.parametrize("name", NAMES)
def test_topology_matches_reference(name):
    g = Golden(name)
    t = build_topology(g.world)
    assert t.P == g.P and t.E == g.E
    assert np.array_equal(t.pos0, g["pos0"])
    assert np.array_equal(t.mass, g["mass"])
    assert np.array_equal(t.fixed, g["fixed"])
    for mine, ref in ((t.ea, "ea"), (t.eb, "eb"), (t.ek, "ek"), (t.el0, "el0"), (t.on_surface, "on_surface")):
        assert np.array_equal(mine, g[ref]), ref
    assert len(t.objects) == g.n_obj
    for o, ob in enumerate(t.objects):
        info = g.obj_info[o]
        assert (int(ob.is_robot), ob.min_point, ob.max_point, ob.boxels.shape[0], ob.grid_w, ob.grid_h,
                len(ob.surface_boxels)) == tuple(int(v) for v in info)
        assert np.array_equal(ob.boxels, g[f"boxels_{o}"])
        assert np.array_equal(ob.surface_boxels, g[f"surface_{o}"])
    assert np.array_equal(t.objects[t.object_index("robot")].actuator_grid_indices, g["act_idx"])


@pytest.mark.parametrize("name", NAMES)
def test_teacher_forced_substeps(name):
    g = Golden(name)
    o = OracleSim(g.world)
    n_exact = 0
    for k in range(g.n_sub):
        if k % 30 == 0:
            o.set_action("robot", g.action_rows(k // 30))
        o.set_state(*g.state_before_substep(k))
        unstable = o.substep()
        pos, vel, _, vt = o.get_state()
        assert unstable == bool(g["sub_unstable"][k]), k
        assert rel_err(pos, g["sub_pos"][k]) <= TF_POS_TOL, k
        assert rel_err(vel, g["sub_vel"][k]) <= TF_VEL_TOL, k
        assert rel_err(vt, g["sub_vt"][k]) <= 1e-9, k
        assert rel_err(o.collision_forces(), g["sub_cf"][k]) <= 1e-12, k
        n_exact += int(np.array_equal(pos, g["sub_pos"][k]) and np.array_equal(vel, g["sub_vel"][k]))
    # substeps without contact must be bit-identical
    no_contact = int((np.abs(g["sub_cf"]).sum(axis=(1, 2)) == 0).sum())
    assert n_exact >= no_contact


@pytest.mark.parametrize("name", NAMES)
def test_candidate_sets(name):
    g = Golden(name)
    o = OracleSim(g.world)
    for k in g["cand_at"].tolist():
        o.set_state(*g.state_before_substep(k))
        mine = o.candidates_as_object_boxel(o.phase_collisions(True))
        assert pair_set(mine) == pair_set(g[f"cand_{k}"]), k


@pytest.mark.parametrize("name", [n for n in NAMES if "20x20" not in n])
def test_free_running_and_step_semantics(name):
    g = Golden(name)
    o = OracleSim(g.world)
    p, v = g["init_pos"], g["init_vel"]
    o.set_state(p, v, p, v)
    chaotic = name.startswith("collapse")        # exploding robot: only flags and time are compared
    for s in range(g.n_steps):
        o.set_action("robot", g.action_rows(s))
        unstable = o.step()
        assert unstable == bool(g["step_unstable"][s]), s
        assert o.get_time() == int(g["step_time"][s]), s
        if not chaotic and s < 2:
            assert rel_err(o.get_state()[0], g["step_pos"][s + 1]) <= FREE_POS_TOL, s


def test_revert_keeps_goals():
    """Environment::revert_to_snapshot (Environment.cpp:247-263) restores pos/vel only."""
    g = Golden("walker_s0")
    o = OracleSim(g.world)
    o.set_action("robot", g.action_rows(0))
    o.step()
    goal = o.edge_state()[1].copy()
    o.revert(0)
    assert o.get_time() == 0
    assert np.array_equal(o.get_state()[0], g["pos0"])
    assert np.array_equal(o.edge_state()[1], goal)
    assert not np.array_equal(goal, g["el0"])
