"""GPU: the CUDA engine, called through the C ABI (ctypes over libmevo_b200.so), against
 (a) the committed fixtures the unmodified reference engine produced, and
 (b) the oracle on seeded inputs.
All fixture worlds live in ONE heterogeneous batch (5x5 ... 20x20 robots, 1-3 objects, static and
dynamic terrain), so every size class of the step kernel is exercised in the same launches."""
import random

import numpy as np
import pytest

from helpers import Golden, golden_names, pair_set, rel_err
from modularevogym_b200 import capi
from modularevogym_b200.world import sample_robot, walker_world

pytestmark = pytest.mark.gpu

NAMES = golden_names()

# FP64 tolerances (relative to max(1, |x|max), see helpers.rel_err).  Teacher-forced = the state is
# reset to the reference's before every substep.  Sources of difference: (1) the order contact
# forces are summed in (candidate-list order here, tree-descent order in the reference), (2) tanh
# (libdevice vs glibc, <= 2 ulp) in the friction terms.  Everything else is bit-identical.
TF_POS_TOL = 1e-14
TF_VEL_TOL = 1e-11
TF_VT_TOL = 1e-9          # vel_true = dpos / 1e-4 amplifies position ulps by 1e4
TF_CF_TOL = 1e-11         # collision_forces, relative to max(1, |cf| max): O(1e4) forces summed in candidate-list order
FREE_POS_TOL = 1e-7       # 60 free-running substeps through stiff contacts


@pytest.fixture(scope="module")
def goldens():
    return [Golden(n) for n in NAMES]


@pytest.fixture(params=["warp", "group", "group+1", "group-1"])
def kernel_choice(request, monkeypatch):
    """Small environments run in the warp-per-environment kernel (wkernel.cuh) by default; MEVO_NO_WKERNEL=1
    (read when a batch is created) sends them through the group-per-environment kernel (kernels.cuh) like
    the large ones.  Both must reproduce the reference."""
    monkeypatch.setenv("MEVO_NO_WKERNEL", "0" if request.param == "warp" else "1")
    # group+1 / group-1: every fixture one size class up / down (MEVO_GROUP_SHIFT), so that each group size -- including the
    # 256-thread class that reads its tables from global memory -- sees 5x5, 10x10 and 20x20 worlds
    monkeypatch.setenv("MEVO_GROUP_SHIFT", {"group+1": "1", "group-1": "-1"}.get(request.param, "0"))
    return request.param


def make_batch(goldens, cand_dump=0):
    B = capi.Batch(0)
    for g in goldens:
        B.add_world(g.world)
    if cand_dump:
        B.enable_candidate_dump(cand_dump)
        B.enable_force_dump()
    B.finalize()
    return B


def test_topology_through_batch(goldens):
    B = make_batch(goldens)
    for i, g in enumerate(goldens):
        t = B.topology(i)
        for mine, ref in (("pos0", "pos0"), ("a", "ea"), ("b", "eb"), ("k", "ek"), ("l0", "el0"), ("on_surface", "on_surface")):
            assert np.array_equal(t[mine], g[ref]), (g.name, ref)
        assert np.array_equal(B.actuator_indices(i, "robot"), g["act_idx"])
        static = g.static_objects()
        for o in range(g.n_obj):
            assert bool(B.object_info(i, o)["is_static"]) == static[o]
    B.close()


def test_teacher_forced_substeps_and_candidates(goldens, kernel_choice):
    """Environment::step, one substep at a time, from the reference's own state."""
    B = make_batch(goldens, cand_dump=8192)
    n_sub = max(g.n_sub for g in goldens)
    worst = {"pos": 0.0, "vel": 0.0, "vt": 0.0, "cf": 0.0}
    exact = total = 0
    dyn = []
    for i, g in enumerate(goldens):
        m = np.ones(g.P, dtype=bool)
        for o, st in enumerate(g.static_objects()):
            if st:
                m[g.obj_info[o, 1]:g.obj_info[o, 2] + 1] = False
        dyn.append(m)
    for k in range(n_sub):
        live = [i for i, g in enumerate(goldens) if k < g.n_sub]
        for i in live:
            g = goldens[i]
            if k % 30 == 0:
                assert B.set_action(i, "robot", g.action_rows(k // 30))
            B.set_state(i, *g.state_before_substep(k))
        _, flags = B.step_substeps(1, 1, apply_actions=False)
        for i in live:
            g = goldens[i]
            pos, vel, last, vt = B.get_state(i)
            assert B.error_flags(i) == 0, (g.name, k)
            assert bool(flags[i]) == bool(g["sub_unstable"][k]), (g.name, k)
            e = (rel_err(pos, g["sub_pos"][k]), rel_err(vel, g["sub_vel"][k]), rel_err(vt, g["sub_vt"][k]))
            assert e[0] <= TF_POS_TOL and e[1] <= TF_VEL_TOL and e[2] <= TF_VT_TOL, (g.name, k, e)
            assert np.array_equal(last, pos), (g.name, k)
            # PhysicsEngine::resolve_collisions' own output (a-6), on the points the engine simulates
            ecf = rel_err(B.read_collision_forces(i)[dyn[i]], g["sub_cf"][k][dyn[i]])
            assert ecf <= TF_CF_TOL, (g.name, k, ecf)
            worst = {"pos": max(worst["pos"], e[0]), "vel": max(worst["vel"], e[1]), "vt": max(worst["vt"], e[2]), "cf": max(worst["cf"], ecf)}
            total += 1
            exact += int(np.array_equal(pos, g["sub_pos"][k]) and np.array_equal(vel, g["sub_vel"][k]))
            if k in g["cand_at"].tolist():
                static = g.static_objects()
                want = pair_set(g[f"cand_{k}"], static, drop_self=True)
                got = pair_set(B.read_candidates(i), static, drop_self=True)
                assert got == want, (g.name, k, sorted(got - want)[:4], sorted(want - got)[:4])
    print(f"\n[{kernel_choice}] teacher-forced: {exact}/{total} substeps bit-identical to the reference; worst rel err {worst}")
    no_contact = sum(int((np.abs(g['sub_cf']).sum(axis=(1, 2)) == 0).sum()) for g in goldens)
    assert exact >= no_contact       # without contacts the result is bit-identical
    B.close()


def test_free_running_steps(goldens, kernel_choice):
    """Sim::step for every fixture at once: 30 substeps per launch, actions through the per-env API;
    unstable flag and sim_time semantics (Sim.cpp:105-115)."""
    B = make_batch(goldens)
    for i, g in enumerate(goldens):
        p, v = g["init_pos"], g["init_vel"]
        B.set_state(i, p, v, p, v)
    n_steps = max(g.n_steps for g in goldens)
    for s in range(n_steps):
        live = [i for i, g in enumerate(goldens) if s < g.n_steps]
        for i in live:
            assert B.set_action(i, "robot", goldens[i].action_rows(s))
        _, flags = B.step(1, apply_actions=False)
        for i in live:
            g = goldens[i]
            assert bool(flags[i]) == bool(g["step_unstable"][s]), (g.name, s)
            assert B.get_time(i) == int(g["step_time"][s]), (g.name, s)
            if s < 2 and not g.name.startswith("collapse"):
                e = rel_err(B.get_state(i)[0], g["step_pos"][s + 1])
                assert e <= FREE_POS_TOL, (g.name, s, e)
    B.close()
