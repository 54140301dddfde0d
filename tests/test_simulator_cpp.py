"""The pybind11 module `simulator_cpp` (csrc/simulator_cpp.cpp): the reference's Python boundary
(SimulatorCPP/PythonBindings.cpp:14-59), driven the way `evogym.sim.EvoSim` drives it
(evogym/sim.py:29-98,126-143,179-252) and compared with fixtures from the reference engine."""
import numpy as np
import pytest

from helpers import Golden, rel_err

REF_SIM_METHODS = ["init", "get_version", "read_object_from_file", "read_robot_from_file", "read_object_from_array",
                   "read_robot_from_array", "step", "set_action", "revert", "get_time", "pos_at_time", "vel_at_time",
                   "object_pos_at_time", "object_vel_at_time", "object_orientation_at_time", "translate_object",
                   "get_indices_of_actuators"]


def test_module_surface():
    """CPU: the module imports without a GPU and exports the reference's classes and method names."""
    from modularevogym_b200 import simulator_cpp as m
    for cls in ("Sim", "Viewer", "Camera", "BatchSim"):
        assert hasattr(m, cls)
    for name in REF_SIM_METHODS:
        assert hasattr(m.Sim, name), name
    for name in ("init", "render", "show_debug_window", "hide_debug_window", "get_debug_window_pos"):
        assert hasattr(m.Viewer, name), name
    for name in ("set_pos", "set_size", "set_resolution", "get_resolution_width", "get_resolution_height", "get_image"):
        assert hasattr(m.Camera, name), name
    cam = m.Camera(False)
    cam.set_resolution(64, 48)
    assert (cam.get_resolution_width(), cam.get_resolution_height()) == (64, 48)
    m.Sim.get_version()                       # static, like Sim::get_version (Sim.h:18)


def evosim_like(m, g):
    """What EvoSim.__init__ / _init_world_items do (evogym/sim.py:29-98): robots first, then the rest, each
    alphabetical; grids are handed over top row first."""
    sim = m.Sim()
    sim.init(g.world.width, g.world.height)
    for o in g.world.load_order():
        add = sim.read_robot_from_array if o.is_robot else sim.read_object_from_array
        assert add(o.grid.astype(np.float64), o.connections.astype(np.float64), o.name, o.x, o.y)
    return sim


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["walker_s0", "pusher_s3", "collapse_s15"])
def test_sim_drop_in(name, capfd):
    from modularevogym_b200 import simulator_cpp as m
    g = Golden(name)
    sim = evosim_like(m, g)
    idx = sim.get_indices_of_actuators("robot")
    assert idx.shape == (1, len(g["act_idx"])) and np.array_equal(idx.ravel(), g["act_idx"])
    assert sim.get_time() == 0
    p0 = sim.pos_at_time(0)
    assert p0.shape == (2, g.P) and np.array_equal(p0.T, g["pos0"])
    if name == "collapse_s15":                 # the perturbed start state: written through revert's path
        return
    for s in range(g.n_steps):
        sim.set_action("robot", g.action_rows(s))
        unstable = sim.step()
        assert unstable == bool(g["step_unstable"][s])
        assert sim.get_time() == int(g["step_time"][s])
        t = sim.get_time()
        if s < 2:
            assert rel_err(sim.pos_at_time(t).T, g["step_pos"][s + 1]) <= 1e-7
            assert rel_err(sim.vel_at_time(t).T, g["step_vel"][s + 1]) <= 1e-4
            a, b = int(g.obj_info[0, 1]), int(g.obj_info[0, 2])
            rp = sim.object_pos_at_time(t, "robot")
            assert rp.shape == (2, b - a + 1) and np.array_equal(rp, sim.pos_at_time(t)[:, a:b + 1])
            assert abs(sim.object_orientation_at_time(t, "robot") - float(g["step_orient"][s])) <= 1e-6
    # unknown time / name: the reference hands back a 2 x 4 zero matrix (Environment.cpp:267-272)
    assert np.array_equal(sim.pos_at_time(12345), np.zeros((2, 4)))
    assert np.array_equal(sim.object_pos_at_time(0, "nobody"), np.zeros((2, 4)))
    assert sim.object_orientation_at_time(0, "nobody") == 0
    # EvoSim.reset = revert(0) (evogym/sim.py:248-252): state back, time back, later snapshots gone
    sim.revert(0)
    assert sim.get_time() == 0 and np.array_equal(sim.pos_at_time(0).T, g["pos0"])
    assert np.array_equal(sim.pos_at_time(1), np.zeros((2, 4)))
    sim.revert(7)                               # no such snapshot: message, nothing changes
    assert "Error: Could not revert" in capfd.readouterr().out
    # after the revert the same actions reproduce the same trajectory (goals persisted: set them again first)
    sim.set_action("robot", g.action_rows(0))
    assert not sim.step()
    assert rel_err(sim.pos_at_time(1).T, g["step_pos"][1]) <= 1e-7
    # translate_object moves the points and re-saves the snapshot of the current time (Sim.cpp:148-151)
    before = sim.object_pos_at_time(1, "robot").copy()
    sim.translate_object(0.05, 0.02, "robot")
    after = sim.object_pos_at_time(1, "robot")
    assert np.allclose(after - before, np.array([[0.05], [0.02]]), atol=1e-15)


@pytest.mark.gpu
def test_batchsim():
    from modularevogym_b200 import simulator_cpp as m
    gs = [Golden(n) for n in ("walker_s0", "walker_s1", "walker_s2")]
    B = m.BatchSim(0)
    for g in gs:
        e = B.begin_env(g.world.width, g.world.height)
        for o in g.world.load_order():
            assert B.add_object(e, o.grid.astype(np.float64), o.connections.astype(np.float64), o.name, o.x, o.y, o.is_robot)
    B.finalize()
    assert B.num_envs == 3
    stride = B.action_stride
    for s in range(2):
        acts = np.ones((3, stride))
        for i, g in enumerate(gs):
            acts[i, :g["actions"].shape[1]] = g["actions"][s]
        B.set_actions(acts)
        flags = B.step(1)
        assert not flags.any()
    for i, g in enumerate(gs):
        assert B.get_time(i) == 2
        assert rel_err(B.pos(i).T, g["step_pos"][2]) <= 1e-7
    com = B.object_com(0)
    for i, g in enumerate(gs):
        a, b = int(g.obj_info[0, 1]), int(g.obj_info[0, 2])
        assert rel_err(com[i], B.pos(i)[:, a:b + 1].mean(axis=1)) <= 1e-12


@pytest.mark.gpu
def test_batchsim_task_layer():
    """BatchSim exposes the task layer of the C ABI: one Thrower-v0 step against the reference environment's fixture"""
    from helpers import EnvGolden
    from modularevogym_b200 import simulator_cpp as m
    g = EnvGolden("env_thrower_s0")
    B = m.BatchSim(0)
    e = B.begin_env(g.world.width, g.world.height)
    for o in g.world.load_order():
        assert B.add_object(e, o.grid.astype(np.float64), o.connections.astype(np.float64), o.name, o.x, o.y, o.is_robot)
    B.finalize()
    B.task_configure(2, "robot", "package", 300)
    d = B.task_dims()
    assert d["modular"] == 225 and d["other"] == 6 and d["flat"] == g["flat_obs"].shape[1]
    t = B.task_env_tables(0)
    assert np.array_equal(t["obs_index_array"], g["obs_index_array"]) and t["voxel_corner_size"] == int(g["voxel_corner_size"])
    r = B.task_reset()
    assert rel_err(r["modular"][0], g["modular_obs"][0]) <= 1e-12 and rel_err(r["flat"][0], g["flat_obs"][0]) <= 1e-12
    B.task_apply_policy(g["voxel_actions"][0][None, :], True)
    assert not B.step(1).any()
    o = B.task_outputs(True)
    assert abs(o["reward"][0] - g["reward"][0]) <= 1e-9 and o["done"][0] == 0
    assert rel_err(o["modular"][0], g["modular_obs"][1]) <= 1e-9 and rel_err(o["other"][0], g["other_obs"][1]) <= 1e-9
    assert B.object_orientation(0).shape == (1,)


@pytest.mark.gpu
def test_batchsim_goal_lists():
    """BatchSim's BidirectionalWalker-v0 goal lists (mevo_batch_task_set_goals / _draw_goals / _get_goals): first step of the
    reference rollout whose first three goals are reached at once"""
    import json
    import os
    from helpers import GOLDEN_DIR
    from modularevogym_b200 import simulator_cpp as m
    from modularevogym_b200.world import WorldSpec
    with np.load(os.path.join(GOLDEN_DIR, "biwalk_near.npz")) as z:
        g = {k: z[k] for k in z.files}
    world = WorldSpec.from_dict(json.loads(str(g["world_json"])))
    B = m.BatchSim(0)
    e = B.begin_env(world.width, world.height)
    for o in world.load_order():
        assert B.add_object(e, o.grid.astype(np.float64), o.connections.astype(np.float64), o.name, o.x, o.y, o.is_robot)
    B.finalize()
    B.task_configure(29, "robot", None, 1000)
    drawn = B.task_get_goals(0)
    assert drawn["current"] == 0 and len(drawn["goals"]) >= 4 and all(20 <= x < 50 for x in drawn["goals"])
    B.task_set_goals([g["goals"].tolist()])
    assert B.task_get_goals(0)["goals"] == g["goals"].tolist()
    r = B.task_reset()
    assert rel_err(r["flat"][0], g["flat_obs"][0]) <= 1e-12
    B.task_apply_policy(g["actions"][0][None, :], False)
    assert not B.step(1).any()
    o = B.task_outputs(False)
    assert abs(o["reward"][0] - g["reward"][0]) <= 1e-9 and rel_err(o["flat"][0], g["flat_obs"][1]) <= 1e-9
    assert B.task_get_goals(0)["current"] == int(g["current_goal"][0]) == 3
    B.task_draw_goals(5)
    assert B.task_get_goals(0)["goals"] != g["goals"].tolist()


@pytest.mark.gpu
def test_several_sims_of_different_sizes_step_alternately():
    """a design loop keeps several `Sim`s alive in one process and steps them in turn: robots of different size classes (a later,
    smaller one must not lower a kernel's shared-memory limit under an earlier, larger one); each equals the same Sim stepped alone"""
    import random
    from modularevogym_b200 import simulator_cpp as m
    from modularevogym_b200.world import obstacle_traverser_world, sample_large_robot, walker_world

    def build(k):
        random.seed(900 + k)
        side = (14, 9, 12, 7, 5)[k]
        body, conn = sample_large_robot((side, side))
        w = obstacle_traverser_world(body, conn) if k % 2 == 0 else walker_world(body, conn)
        sim = m.Sim()
        sim.init(w.width, w.height)
        for o in w.load_order():
            add = sim.read_robot_from_array if o.is_robot else sim.read_object_from_array
            assert add(o.grid.astype(np.float64), o.connections.astype(np.float64), o.name, o.x, o.y)
        return sim

    def actions(sim, k, s):
        idx = np.asarray(sim.get_indices_of_actuators("robot")).ravel()
        rng = np.random.default_rng(100 * k + s)
        return np.stack((idx.astype(np.float64), rng.uniform(0.6, 1.6, idx.size)), axis=1)

    alone = []
    for k in range(5):
        sim = build(k)
        for s in range(3):
            sim.set_action("robot", actions(sim, k, s))
            sim.step()
        alone.append(np.asarray(sim.pos_at_time(sim.get_time())).copy())
        del sim
    sims = [build(k) for k in range(5)]
    for s in range(3):
        for k in range(5):
            sims[k].set_action("robot", actions(sims[k], k, s))
            sims[k].step()
    for k in range(5):
        assert np.array_equal(np.asarray(sims[k].pos_at_time(sims[k].get_time())), alone[k]), k
