"""CPU: the host world builder (csrc/topology.cpp through the C ABI's mevo_world_*) against the LIVE compiled reference
(oracle/_ref: ObjectCreator::read_object_from_array + SimObject::compute_surface + edge orientation + Robot::init) on seeded
random worlds: random shapes and voxel types, PARTIAL connectivity (adjacent voxels that are not connected keep separate points
and edges), several objects, objects the reference refuses (disconnected, overlapping, out of the grid).  Array-for-array
equality, and the same accept / refuse decisions.  Skipped where oracle/_ref was not built."""
import numpy as np
import pytest

from modularevogym_b200 import capi
from oracle import ref as oref

pytestmark = pytest.mark.skipif(not oref.available(), reason="oracle/_ref/libmevo_ref.so not built")


def random_object(rng, h, w, p_empty, p_drop):
    grid = rng.choice([0, 1, 2, 3, 4, 5], size=(h, w), p=[p_empty] + [(1 - p_empty) / 5] * 5)
    conn = []
    for r in range(h):
        for c in range(w):
            if grid[r, c] == 0:
                continue
            if c + 1 < w and grid[r, c + 1] and rng.random() >= p_drop:
                conn.append((r * w + c, r * w + c + 1))
            if r + 1 < h and grid[r + 1, c] and rng.random() >= p_drop:
                conn.append((r * w + c, (r + 1) * w + c))
    cn = np.asarray(conn, dtype=np.float64).T if conn else np.zeros((2, 0))
    return grid.astype(np.float64), cn


def same_world(hw: capi.HostWorld, r: "oref.RefSim", names):
    assert hw.num_points == r.num_points and hw.num_edges == r.num_edges and hw.num_objects == r.num_objects
    t, ed = hw.topology(), r.edges()
    mass, fixed = r.point_consts()
    assert np.array_equal(t["pos0"], r.get_state()[0])
    assert np.array_equal(t["mass"], mass) and np.array_equal(t["fixed"], fixed)
    for mine, ref in (("a", "a"), ("b", "b"), ("k", "k"), ("l0", "l0"), ("on_surface", "on_surface")):
        assert np.array_equal(t[mine], ed[ref]), ref
    for o in range(hw.num_objects):
        i, j = hw.object_info(o), r.object_info(o)
        assert all(i[k] == j[k] for k in j), (i, j)
        assert np.array_equal(hw.object_boxels(o), r.object_boxels(o))
        assert np.array_equal(hw.object_surface_boxels(o), np.sort(r.object_surface_boxels(o)))
        if j["is_robot"]:
            name = names[o]
            assert np.array_equal(hw.actuator_indices(name), r.actuator_indices(name))


@pytest.mark.parametrize("seed", range(40))
def test_random_world_matches_reference(seed, capfd):
    rng = np.random.default_rng(1000 + seed)
    W, H = 24, 16
    hw, r = capi.HostWorld(world_w=W, world_h=H), oref.RefSim(world_w=W, world_h=H)
    accepted = refused = 0
    names = []                                  # names of the objects that really exist (an all-empty grid "succeeds" without adding one)
    for k in range(int(rng.integers(2, 6))):
        h, w = int(rng.integers(1, 8)), int(rng.integers(1, 8))
        grid, cn = random_object(rng, h, w, p_empty=float(rng.choice([0.0, 0.2, 0.5])), p_drop=float(rng.choice([0.0, 0.0, 0.15, 0.5])))
        x, y = int(rng.integers(-1, W - 2)), int(rng.integers(-1, H - 2))           # sometimes outside the grid / overlapping
        is_robot = bool(((grid == 3) | (grid == 4)).any()) and bool(rng.random() < 0.7)
        name = f"obj{k}"
        n_before = r.num_objects
        ok_ref = r.add_object(grid, cn, name, x, y, is_robot)
        ok_mine = hw.add_object(grid, cn, name, x, y, is_robot)
        if r.num_objects > n_before:
            names.append(name)
        assert ok_mine == ok_ref, (seed, k, "accept / refuse differs", grid, x, y)
        accepted += ok_ref
        refused += not ok_ref
        same_world(hw, r, names)                # also after a refusal: the world must be untouched
    capfd.readouterr()                          # both print the reference's "Error: ..." lines on refusals


def test_repeated_terrain_objects_match_reference(capfd):
    """the builder keeps built objects of >= 32 cells that are not the first of their world and re-instantiates them with shifted
    ids (csrc/topology.cpp, "built-object cache"): the same terrain and free objects after robots of different sizes -- different
    point / edge offsets every time --, a duplicate name and an overlapping copy, all against the live reference"""
    rng = np.random.default_rng(77)
    W, H = 40, 16
    ground = np.full((2, 40), 5.0)
    ground[0, 5:9] = 0
    ground_cn = []
    for r_ in range(2):
        for c in range(40):
            if ground[r_, c] == 0:
                continue
            if c + 1 < 40 and ground[r_, c + 1]:
                ground_cn.append((r_ * 40 + c, r_ * 40 + c + 1))
            if r_ + 1 < 2 and ground[r_ + 1, c]:
                ground_cn.append((r_ * 40 + c, (r_ + 1) * 40 + c))
    ground_cn = np.asarray(ground_cn, dtype=np.float64).T
    block, block_cn = random_object(np.random.default_rng(5), 6, 7, 0.1, 0.1)          # 42 cells, partial connectivity
    for trial in range(8):
        hw, r = capi.HostWorld(world_w=W, world_h=H), oref.RefSim(world_w=W, world_h=H)
        names = []
        h, w = int(rng.integers(2, 6)), int(rng.integers(2, 6))
        robot, robot_cn = random_object(rng, h, w, 0.0, 0.0)
        robot[0, 0] = 3.0
        for grid, cn, name, x, y, is_robot in ((robot, robot_cn, "robot", 1, 2, True), (ground, ground_cn, "ground", 0, 0, False),
                                               (block, block_cn, "block", 20, 3, False), (block, block_cn, "block", 30, 3, False),
                                               (block, block_cn, "block2", 22, 4, False), (ground, ground_cn, "roof", 0, 12, False)):
            n_before = r.num_objects
            ok_ref = r.add_object(grid, cn, name, x, y, is_robot)
            assert hw.add_object(grid, cn, name, x, y, is_robot) == ok_ref, (trial, name)
            if r.num_objects > n_before:
                names.append(name)
            same_world(hw, r, names)
        assert names == ["robot", "ground", "block", "roof"]
    capfd.readouterr()
