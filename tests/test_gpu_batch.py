"""GPU: batch-level behaviour of the engine through the C ABI -- the batched action path, reset
semantics, determinism, independence of environments, and size-independent properties at
BASELINE.json's full size (4096 Walker-v0 5x5 environments)."""
import random

import numpy as np
import pytest

from helpers import Golden, rel_err
from modularevogym_b200 import capi
from modularevogym_b200.world import sample_robot, walker_world
from oracle.oracle import OracleSim

pytestmark = pytest.mark.gpu


def walker(seed):
    random.seed(seed)
    r, c = sample_robot((5, 5))
    return walker_world(r, c)


def clipped_actions(rng, shape):
    return np.clip(rng.uniform(-1, 1, shape) + 1.0, 0.6, 1.6)      # evogym/envs/base.py:42,389


def test_batched_actions_vs_oracle_step_forced():
    """30-substep launches with actions taken from the dense device buffer (K1), state re-synchronised
    with the oracle at every step boundary."""
    seeds = list(range(100, 112))
    worlds = [walker(s) for s in seeds]
    oracles = [OracleSim(w) for w in worlds]
    B = capi.Batch(0)
    for w in worlds:
        B.add_world(w)
    B.finalize()
    stride = B.action_stride
    rng = np.random.default_rng(7)
    worst = 0.0
    for step in range(12):
        acts = np.ones((len(worlds), stride))
        for i, o in enumerate(oracles):
            n = len(o.actuator_indices("robot"))
            acts[i, :n] = clipped_actions(rng, n)
            st = o.get_state()
            B.set_state(i, *st)
            B.set_goal(i, o.edge_state()[1])
            o.set_action_values("robot", acts[i, :n])
        B.set_actions(acts)
        _, flags = B.step(1, apply_actions=True)
        for i, o in enumerate(oracles):
            assert bool(flags[i]) == o.step()
            so, sg = o.get_state(), B.get_state(i)
            assert np.array_equal(B.get_goal(i), o.edge_state()[1])         # K1 is exact
            e = rel_err(sg[0], so[0])
            worst = max(worst, e)
            assert e <= 1e-9, (seeds[i], step, e)
            assert B.get_time(i) == o.get_time()
    print(f"\nstep-forced vs oracle: worst rel pos err over 12 steps x 12 envs = {worst:.3e}")
    B.close()


def test_reset_restores_state_but_not_goals():
    """Sim::revert(0) (Environment.cpp:247-263): pos, vel <- snapshot 0; pos_last = pos; vel_true = vel;
    length_eq_goal untouched; time = 0."""
    B = capi.Batch(0)
    for s in (0, 1, 2):
        B.add_world(walker(s))
    B.finalize()
    before = [B.get_state(i) for i in range(3)]
    B.set_actions(np.full((3, B.action_stride), 1.4))
    B.step(3, apply_actions=True)
    goals = [B.get_goal(i) for i in range(3)]
    assert B.get_time(1) == 3
    B.reset([0, 2])
    for i in (0, 2):
        st = B.get_state(i)
        assert np.array_equal(st[0], before[i][0]) and np.array_equal(st[1], before[i][1])
        assert np.array_equal(st[2], st[0]) and np.array_equal(st[3], st[1])
        assert np.array_equal(B.get_goal(i), goals[i])
        assert B.get_time(i) == 0
    assert B.get_time(1) == 3 and not np.array_equal(B.get_state(1)[0], before[1][0])
    B.close()


def test_full_size_properties():
    """4096 Walker-v0 5x5 environments (BASELINE.json configs[1]): bit-reproducible, environments are
    independent of their neighbours in the batch, and a copy of an environment placed anywhere in
    the batch evolves bit-identically."""
    n = 4096
    n_designs = 64
    worlds = [walker(s) for s in range(n_designs)]

    def build(order):
        B = capi.Batch(0)
        for k in order:
            B.add_world(worlds[k])
        B.finalize()
        return B

    order = [i % n_designs for i in range(n)]
    rng = np.random.default_rng(11)
    per_design = clipped_actions(rng, (5, n_designs, 32))

    def run(B, order):
        stride = B.action_stride
        for s in range(5):
            acts = np.stack([per_design[s, k, :stride] for k in order])
            B.set_actions(acts)
            B.step(1, apply_actions=True)
        return B.object_com(0)

    B1 = build(order)
    com1 = run(B1, order)
    s1 = [B1.get_state(i) for i in (0, 64, 4095, 777)]
    assert B1.num_envs == n and B1.last_launches >= 1
    # copies of one design agree bit for bit wherever they sit
    c = com1.reshape(n // n_designs, n_designs, 2)
    assert np.array_equal(c, np.broadcast_to(c[0], c.shape))
    assert np.array_equal(s1[0][0], s1[1][0])
    assert np.isfinite(com1).all()
    B1.close()
    # a different neighbourhood (reversed order) leaves every environment's trajectory unchanged
    rev = order[::-1]
    B2 = build(rev)
    com2 = run(B2, rev)
    assert np.array_equal(com2[::-1], com1)
    B2.close()
    # against the oracle for a few of them (free-running 5 steps)
    for k in (0, 5, 63):
        o = OracleSim(worlds[k])
        stride = len(o.actuator_indices("robot"))
        for s in range(5):
            o.set_action_values("robot", per_design[s, k, :stride])
            o.step()
        a, b = o.topo.objects[0].min_point, o.topo.objects[0].max_point + 1
        assert rel_err(com1[k], o.get_state()[0][a:b].mean(axis=0)) <= 1e-8


def test_walker_observation_rows():
    """On-device Walker-v0 observation (evogym/envs/walk.py:58-61: get_vel_com_obs + get_relative_pos_obs,
    evogym/envs/base.py:214-240) against the same numpy expressions on the downloaded state."""
    gs = [Golden(n) for n in ("walker_s0", "walker_s1", "walker_dense", "pusher_s3")]
    B = capi.Batch(0)
    for g in gs:
        B.add_world(g.world)
    B.finalize()
    rng = np.random.default_rng(5)
    B.set_actions(np.clip(rng.uniform(-1, 1, (len(gs), B.action_stride)) + 1.0, 0.6, 1.6))
    B.step(2)
    obs = B.object_obs(0)
    width = obs.shape[1]
    for i, g in enumerate(gs):
        a, b = int(g.obj_info[0, 1]), int(g.obj_info[0, 2])
        pos, vel = B.get_state(i)[:2]
        p = pos[a:b + 1].T                      # (2, n) like Sim::object_pos_at_time
        v = vel[a:b + 1].T
        want = np.concatenate((np.mean(v, axis=1), (p - np.array([np.mean(p, axis=1)]).T).flatten()))
        assert width >= want.size
        assert np.max(np.abs(obs[i, :want.size] - want)) <= 1e-12
        assert not obs[i, want.size:].any()
    com = B.object_com(0)
    for i, g in enumerate(gs):
        a, b = int(g.obj_info[0, 1]), int(g.obj_info[0, 2])
        assert np.max(np.abs(com[i] - B.get_state(i)[0][a:b + 1].mean(axis=0))) <= 1e-12
    B.close()


def test_fetch_step_results_equals_separate_reads():
    """flags + centre of mass of a step in one round trip (mevo_batch_fetch_step_results) = step flags + object_com"""
    worlds = [walker(s) for s in range(300, 306)]
    A, B = capi.Batch(0), capi.Batch(0)
    for w in worlds:
        A.add_world(w); B.add_world(w)
    A.finalize(); B.finalize()
    acts = clipped_actions(np.random.default_rng(11), (len(worlds), A.action_stride))
    A.set_actions(acts); B.set_actions(acts)
    _, flags_a = A.step(1, apply_actions=True)
    com_a = A.object_com(0)
    B.step_async(1, apply_actions=True)
    flags_b, com_b = np.zeros(len(worlds), dtype=np.uint8), np.zeros((len(worlds), 2))
    n_unstable = B.fetch_step_results_ptr(flags_b.ctypes.data, 0, com_b.ctypes.data)
    assert n_unstable == int(flags_a.sum()) and np.array_equal(flags_a, flags_b) and np.array_equal(com_a, com_b)
    A.close(); B.close()


@pytest.mark.gpu
@pytest.mark.parametrize("env_id", ["Walker-v0", "Pusher-v0", "Climber-v0", "ObstacleTraverser-v1", "BridgeWalker-v0", "CaveCrawler-v0"])
def test_add_robot_envs_equals_add_world(env_id):
    """mevo_batch_add_robot_envs (one call for a whole design-loop rebuild: shared terrain + one robot per environment) builds the
    environments add_world builds from the per-environment WorldSpec: same topology arrays, same state after a step"""
    import random
    from modularevogym_b200 import capi
    from modularevogym_b200.vec_env import TASKS
    from modularevogym_b200.world import sample_large_robot, sample_robot, terrain_for
    random.seed(31)
    robots = [sample_robot((5, 5)) if k % 3 else sample_large_robot((4 + k % 4, 3 + k % 5)) for k in range(12)]
    builder = TASKS[env_id][1]
    A, B = capi.Batch(0), capi.Batch(0)
    for body, conn in robots:
        A.add_world(builder(body, conn))
    i = 0
    while i < len(robots):                               # grouped by terrain like the vec-env front (Climber-v0: pipe width)
        terrain, rx, ry = terrain_for(env_id, robots[i][0])
        j = i + 1
        while j < len(robots) and terrain_for(env_id, robots[j][0])[0] is terrain:
            j += 1
        first = B.add_robot_worlds(terrain, [r for r, _ in robots[i:j]], [c for _, c in robots[i:j]], "robot", rx, ry)
        assert first == i
        i = j
    A.finalize(); B.finalize()
    assert A.num_envs == B.num_envs == len(robots)
    acts = np.random.default_rng(3).uniform(0.6, 1.6, (len(robots), A.action_stride))
    for X in (A, B):
        X.set_actions(acts)
        X.step(1, True)
    for e in range(len(robots)):
        ta, tb = A.topology(e), B.topology(e)
        assert ta.keys() == tb.keys()
        for k in ta:
            assert np.array_equal(ta[k], tb[k]), (e, k)
        assert A.num_objects(e) == B.num_objects(e)
        for o in range(A.num_objects(e)):
            assert A.object_info(e, o) == B.object_info(e, o)
        for x, y in zip(A.get_state(e), B.get_state(e)):
            assert np.array_equal(x, y), e
    A.close(); B.close()
