"""World / robot I/O: the JSON world format (``evogym/world.py:25-72``, the 33 files under
``evogym/envs/sim_files/``), the built-in world builders of ``modularevogym_b200/world.py``, and the ``.sob``-style text loader
``Sim.read_object_from_file`` / ``read_robot_from_file`` (``SimulatorCPP/Sim.cpp:30-48`` -> ``ObjectCreator.cpp:118-171``).

``tests/golden/worlds_topology.npz`` (``tests/golden/make_world_golden.py``) holds, per file, the arrays the REFERENCE built
from it through its own Python loader and load order, plus this repo's parse of the same file.
"""
import glob
import json
import os
import random

import numpy as np
import pytest

from helpers import GOLDEN_DIR
from modularevogym_b200 import capi
from modularevogym_b200 import world as W

SIM_FILES = "/root/reference/evogym/envs/sim_files"
have_reference = pytest.mark.skipif(not os.path.isdir(SIM_FILES), reason="reference checkout not present")

with np.load(os.path.join(GOLDEN_DIR, "worlds_topology.npz")) as _z:
    WORLDS = {k: _z[k] for k in _z.files}
NAMES = json.loads(str(WORLDS["names"]))


def test_all_33_reference_worlds_are_in_the_fixture():
    assert len(NAMES) == 33


@pytest.mark.parametrize("name", NAMES)
def test_world_spec_builds_the_reference_topology(name):
    """WorldSpec (this repo's parse of the JSON file, stored in the fixture) -> host world builder == what the reference's
    EvoWorld.from_json + EvoSim._init_world_items + ObjectCreator built, array for array."""
    spec = W.WorldSpec.from_dict(json.loads(str(WORLDS[name + "/world_json"])))
    hw = capi.HostWorld(spec)
    t = hw.topology()
    assert np.array_equal(t["pos0"], WORLDS[name + "/pos0"])
    assert np.array_equal(t["fixed"], WORLDS[name + "/fixed"])
    assert np.array_equal(t["a"], WORLDS[name + "/ea"]) and np.array_equal(t["b"], WORLDS[name + "/eb"])
    assert np.array_equal(t["k"], WORLDS[name + "/ek"]) and np.array_equal(t["l0"], WORLDS[name + "/el0"])
    assert np.array_equal(t["on_surface"], WORLDS[name + "/on_surface"])
    assert (t["mass"] == 1.0).all()
    infos = WORLDS[name + "/obj_info"]
    assert hw.num_objects == infos.shape[0]
    assert [o.name for o in spec.load_order()] == json.loads(str(WORLDS[name + "/obj_names"]))
    boxels = []
    for o in range(hw.num_objects):
        i = hw.object_info(o)
        assert [i[k] for k in ("is_robot", "min_point", "max_point", "n_boxels", "grid_w", "grid_h", "n_surface")] == infos[o].tolist()
        boxels.append(hw.object_boxels(o))
    assert np.array_equal(np.concatenate(boxels, axis=0), WORLDS[name + "/boxels"])


@have_reference
@pytest.mark.parametrize("name", NAMES)
def test_from_json_parses_the_reference_file(name):
    """WorldSpec.from_json on the reference's file itself reproduces the parse stored in the fixture"""
    spec = W.WorldSpec.from_json(os.path.join(SIM_FILES, name + ".json"))
    assert spec.to_dict() == json.loads(str(WORLDS[name + "/world_json"]))


def _same_world(a: W.WorldSpec, b: W.WorldSpec) -> bool:
    if set(a.objects) != set(b.objects) or (a.width, a.height) != (b.width, b.height):
        return False
    for n in a.objects:
        oa, ob = a.objects[n], b.objects[n]
        ca = set(map(tuple, np.sort(oa.connections.T, axis=1).tolist()))
        cb = set(map(tuple, np.sort(ob.connections.T, axis=1).tolist()))
        if not (np.array_equal(oa.grid, ob.grid) and (oa.x, oa.y) == (ob.x, ob.y) and ca == cb):
            return False
    return True


BUILDERS = [("Walker-v0", W.walker_world, (1, 1)), ("Pusher-v0", W.pusher_world, (1, 1)), ("Thrower-v0", W.thrower_world, (1, 1)),
            ("Climber-v0", W.climber_world, (1, 1)), ("ObstacleTraverser-v1", W.obstacle_traverser_world, (2, 4)),
            ("BridgeWalker-v0", W.bridge_walker_world, (2, 5)), ("ObstacleTraverser-v0", W.obstacle_traverser_v0_world, (2, 1)),
            ("UpStepper-v0", W.up_stepper_world, (1, 1)), ("DownStepper-v0", W.down_stepper_world, (1, 11)),
            ("Jumper-v0", W.jumper_world, (32, 1)), ("Hurdler-v0", W.hurdler_world, (2, 4)), ("Climber-v1", W.climber_v1_world, (1, 1)),
            ("CaveCrawler-v0", W.cave_crawler_world, (1, 2)), ("Climber-v2", W.climber_v2_world, (1, 1)),
            ("Carrier-v0", W.carrier_world, (1, 1)), ("Pusher-v1", W.pusher_v1_world, (13, 1)),
            ("Balancer-v0", W.balancer_world, (15, 3)), ("Balancer-v1", W.balancer_v1_world, (10, 1)),
            ("GapJumper-v0", W.gap_jumper_world, (1, 6)), ("PlatformJumper-v0", W.platform_jumper_world, (1, 6)),
            ("Carrier-v1", W.carrier_v1_world, (1, 4)), ("Lifter-v0", W.lifter_world, (2, 3)),
            ("Flipper-v0", W.flipper_world, (60, 1)), ("BeamToppler-v0", W.beam_toppler_world, (1, 1)),
            ("BeamSlider-v0", W.beam_slider_world, (1, 1)), ("ShapeChange", W.shape_change_world, (7, 1)),
            ("BidirectionalWalker-v0", W.bidirectional_walker_world, (33, 1))]
# (Traverser-v0 and Catcher-v0 add objects in their environments' code: pinned against the environments' own worlds in test_task_layer.py)


@pytest.mark.parametrize("env_id,builder,at", BUILDERS)
def test_builtin_world_builders_equal_the_world_files(env_id, builder, at):
    """the hard-coded terrain of the built-in builders == the reference's world file (connections as sets: the JSON lists each
    pair in both directions) + the robot where the reference's environment puts it"""
    random.seed(3)
    r, c = W.sample_robot((5, 5))
    ref = W.WorldSpec.from_dict(json.loads(str(WORLDS[env_id + "/world_json"])))
    ref.add_from_array("robot", r, at[0], at[1], c)
    assert _same_world(builder(r, c), ref)


def test_world_spec_json_round_trip(tmp_path):
    """to_dict / from_dict and the reference's JSON layout (grid_width, grid_height, objects{indices, types, neighbors})"""
    random.seed(5)
    r, c = W.sample_robot((5, 5))
    w = W.pusher_world(r, c)
    assert _same_world(W.WorldSpec.from_dict(w.to_dict()), w)
    # write the reference's file format by hand and read it back through from_json
    objs = {}
    for o in w.objects.values():
        h, wd = o.grid.shape
        idx, types, nb = [], [], {}
        loc = {}
        for rr in range(h):
            for cc in range(wd):
                if o.grid[rr, cc]:
                    gi = (o.y + (h - 1 - rr)) * w.width + (o.x + cc)
                    loc[rr * wd + cc] = gi
                    idx.append(gi); types.append(int(o.grid[rr, cc])); nb[str(gi)] = []
        for a, b in o.connections.T.tolist():
            nb[str(loc[a])].append(loc[b]); nb[str(loc[b])].append(loc[a])
        objs[o.name] = {"indices": idx, "types": types, "neighbors": nb}
    p = tmp_path / "world.json"
    p.write_text(json.dumps({"grid_width": w.width, "grid_height": w.height, "objects": objs}))
    assert _same_world(W.WorldSpec.from_json(str(p)), w)


# ------------------------------------------------------------------------------------------------ text loader (GPU: pybind Sim)

def _sob_text(grid: np.ndarray, conn: np.ndarray, trailing_newline: bool) -> str:
    h, w = grid.shape
    lines = [f"{w} {h}", " ".join(str(int(v)) for v in grid.ravel())]
    lines += [f"{a} {b}" for a, b in conn.T.tolist()]
    return "\n".join(lines) + ("\n" if trailing_newline else "")


@pytest.mark.gpu
@pytest.mark.parametrize("trailing_newline", [False, True])
def test_text_file_loader_matches_reference(tmp_path, trailing_newline):
    """Sim.read_robot_from_file / read_object_from_file of the pybind module against the reference's own loader (through
    oracle/_ref when it is there, else against the array loader): same points, same edges, same first env-step.  With a
    trailing newline the reference's `while (!file.eof())` loop reads the last pair twice (ObjectCreator.cpp:157-162) -- a
    duplicate connection, which changes nothing."""
    from modularevogym_b200 import simulator_cpp
    from oracle import ref as oref
    random.seed(11)
    robot, rconn = W.sample_robot((5, 5))
    ground = np.full((1, 30), 5)
    gconn = W.full_connectivity(ground)
    rf, gf = tmp_path / "robot.sob", tmp_path / "ground.sob"
    rf.write_text(_sob_text(robot, rconn, trailing_newline))
    gf.write_text(_sob_text(ground, gconn, trailing_newline))

    s = simulator_cpp.Sim()
    s.init(30, 12)
    assert s.read_robot_from_file(str(rf), "robot", 2, 1)
    assert s.read_object_from_file(str(gf), "ground", 0, 0)
    assert not s.read_object_from_file(str(tmp_path / "missing.sob"), "nothing", 0, 5)       # "Error: ... not found", False
    assert not s.read_object_from_file(str(gf), "ground2", 0, 0)                             # overlaps: refused like the reference
    pos0 = np.asarray(s.pos_at_time(0)).T.copy()
    idx = np.asarray(s.get_indices_of_actuators("robot")).ravel()

    a = simulator_cpp.Sim()                                    # same world through the array loader
    a.init(30, 12)
    assert a.read_robot_from_array(robot.astype(np.float64), rconn.astype(np.float64), "robot", 2, 1)
    assert a.read_object_from_array(ground.astype(np.float64), gconn.astype(np.float64), "ground", 0, 0)
    assert np.array_equal(pos0, np.asarray(a.pos_at_time(0)).T)
    assert np.array_equal(idx, np.asarray(a.get_indices_of_actuators("robot")).ravel())

    act = np.stack((idx.astype(np.float64), np.linspace(0.7, 1.5, idx.size)), axis=1)
    s.set_action("robot", act)
    assert not s.step()
    pos1 = np.asarray(s.pos_at_time(1)).T
    if oref.available():
        r = oref.RefSim(world_w=30, world_h=12)
        assert r.add_object_from_file(str(rf), "robot", 2, 1, True)
        assert r.add_object_from_file(str(gf), "ground", 0, 0, False)
        assert np.array_equal(pos0, r.get_state()[0])
        assert np.array_equal(idx, r.actuator_indices("robot"))
        r.set_action("robot", act)
        assert not r.step()
        assert np.abs(pos1 - r.get_state()[0]).max() <= 1e-9
    else:
        a.set_action("robot", act)
        assert not a.step()
        assert np.array_equal(pos1, np.asarray(a.pos_at_time(1)).T)
