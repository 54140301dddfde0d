"""GPU: teacher-forced per-substep parity against the LIVE compiled reference on seeded RANDOM worlds -- beyond the benchmarked
workloads: several robots per world (action slots across robots), free soft / rigid objects, all-FIXED pieces hanging in the air
(pruned, never integrated), partial connectivity (unconnected neighbours keep separate points and collide with each other), all
dropped onto a FIXED floor so that they fall, touch and stack.  Same protocol and tolerances as tests/test_gpu_parity.py, both
kernels.  Skipped where oracle/_ref was not built."""
import numpy as np
import pytest

from helpers import pair_set, rel_err
from modularevogym_b200 import capi
from modularevogym_b200.world import WorldSpec
from oracle import ref as oref
from test_random_worlds import random_object

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not oref.available(), reason="oracle/_ref/libmevo_ref.so not built")]

TF_POS_TOL, TF_VEL_TOL, TF_VT_TOL, TF_CF_TOL = 1e-14, 1e-11, 1e-9, 1e-11
N_WORLDS, N_SUB = 32, 90


def random_world(seed: int):
    rng = np.random.default_rng(7000 + seed)
    W, H = 28, 18
    probe = oref.RefSim(world_w=W, world_h=H)                  # the reference decides what is a valid object / placement
    spec = WorldSpec(width=W, height=H)
    floor = np.full((1, W), 5.0)
    cn = np.stack((np.arange(W - 1), np.arange(1, W))).astype(np.float64)
    assert probe.add_object(floor, cn, "floor", 0, 0, False)
    spec.add_from_array("floor", floor.astype(np.int64), 0, 0, cn.astype(np.int64))
    robots = []
    # every fourth world: 9 ... 14 small objects -- more simulated objects than the warp kernels' eight (group kernels, 6-bit object
    # field of the pair codes); at most four of them robots (robots own the eight self-collision counters)
    many = seed % 4 == 3
    for k in range(int(rng.integers(9, 15)) if many else int(rng.integers(2, 6))):
        for _try in range(20):
            h, w = (int(rng.integers(1, 4)), int(rng.integers(1, 4))) if many else (int(rng.integers(1, 7)), int(rng.integers(1, 7)))
            grid, c = random_object(rng, h, w, p_empty=float(rng.choice([0.0, 0.2, 0.4])), p_drop=float(rng.choice([0.0, 0.0, 0.2])))
            if many and k >= 4:
                grid = np.where((grid == 3) | (grid == 4), 2, grid)
            if not grid.any():
                continue
            x, y = int(rng.integers(0, W - w + 1)), int(rng.integers(1, 12 if many else 5))
            is_robot = bool(((grid == 3) | (grid == 4)).any())
            name = f"{'robot' if is_robot else 'thing'}{k}"
            n0 = probe.num_objects
            if probe.add_object(grid, c, name, x, y, is_robot) and probe.num_objects > n0:
                spec.add_from_array(name, grid.astype(np.int64), x, y, c.astype(np.int64))
                if is_robot:
                    robots.append(name)
                break
    return spec, robots


@pytest.mark.parametrize("kernel", ["warp", "group"])
def test_random_worlds_teacher_forced(kernel, monkeypatch):
    monkeypatch.setenv("MEVO_NO_WKERNEL", "0" if kernel == "warp" else "1")
    worlds = [random_world(s) for s in range(N_WORLDS)]
    refs = [oref.RefSim(w) for w, _ in worlds]
    B = capi.Batch(0)
    for w, _ in worlds:
        B.add_world(w)
    B.enable_candidate_dump(16384)
    B.enable_force_dump()
    B.finalize()
    n = len(worlds)
    static_objs = [[bool(B.object_info(i, o)["is_static"]) for o in range(B.num_objects(i))] for i in range(n)]
    dyn = []
    for i in range(n):
        m = np.ones(B.num_points(i), dtype=bool)
        for o in range(B.num_objects(i)):
            info = B.object_info(i, o)
            if info["is_static"]:
                m[info["min_point"]:info["max_point"] + 1] = False
        dyn.append(m)
        assert B.num_points(i) == refs[i].num_points and B.num_edges(i) == refs[i].num_edges
    # unconnected neighbouring voxels keep one side edge each: some points carry more than 8 springs (the group kernel's second
    # gather record, DesignHeader::o_csrx)
    deg = [np.bincount(np.concatenate((r.edges()["a"], r.edges()["b"]))).max() for r in refs]
    assert max(deg) > 8, deg
    assert max(sum(1 for o in range(B.num_objects(i)) if not static_objs[i][o]) for i in range(n)) > 8
    rngs = [np.random.default_rng(90 + i) for i in range(n)]
    live = [True] * n
    exact = total = contacts = 0
    worst = [0.0] * 4
    for k in range(N_SUB):
        if k % 30 == 0:
            for i, (w, robots) in enumerate(worlds):
                for name in robots:
                    idx = refs[i].actuator_indices(name)
                    rows = np.stack((idx.astype(np.float64), np.clip(rngs[i].uniform(-1, 1, idx.size) + 1.0, 0.6, 1.6)), axis=1)
                    refs[i].set_action(name, rows)
                    assert B.set_action(i, name, rows)
        want = {}
        for i, r in enumerate(refs):
            if live[i]:
                B.set_state(i, *r.get_state())
                if k in (0, 29, 60, 89):
                    want[i] = pair_set(r.candidates(), static_objs[i], drop_self=True)
        _, flags = B.step_substeps(1, 1, apply_actions=False)
        for i, r in enumerate(refs):
            if not live[i]:
                continue
            unstable = r.substep()
            assert bool(flags[i]) == unstable, (i, k)
            pr, vr, _, vtr = r.get_state()
            cfr = r.collision_forces()
            pos, vel, last, vt = B.get_state(i)
            cf = B.read_collision_forces(i)
            e = (rel_err(pos, pr), rel_err(vel, vr), rel_err(vt, vtr), rel_err(cf[dyn[i]], cfr[dyn[i]]))
            assert e[0] <= TF_POS_TOL and e[1] <= TF_VEL_TOL and e[2] <= TF_VT_TOL and e[3] <= TF_CF_TOL, (kernel, i, k, e)
            worst = [max(a, b) for a, b in zip(worst, e)]
            total += 1
            hit = bool(np.abs(cfr[dyn[i]]).sum() > 0)
            contacts += hit
            same = bool(np.array_equal(pos, pr) and np.array_equal(vel, vr))
            exact += same
            assert same or hit, (kernel, i, k, "not bit-identical although no contact acted")
            if i in want:
                got = pair_set(B.read_candidates(i), static_objs[i], drop_self=True)
                assert got == want[i], (kernel, i, k, sorted(got - want[i])[:3], sorted(want[i] - got)[:3])
            if unstable:
                live[i] = False
    print(f"\n[random worlds, {kernel}] {sum(d > 8 for d in deg)} of {n} worlds with a point of more than 8 springs (max {max(deg)}); {total} substeps, {exact} bit-identical, {contacts} with contacts, worst rel err pos/vel/vel_true/cf {worst}")
    B.close()


def test_object_limits_are_refused_with_a_message():
    """nine robots (eight self-collision counters) and 65 simulated objects (6-bit object field) are refused when the batch is packed"""
    def world(n_objects, n_robots):
        w = WorldSpec(width=3 * n_objects + 2, height=6)
        w.add_from_array("floor", np.full((1, 3 * n_objects + 2), 5, dtype=np.int64), 0, 0)
        for k in range(n_objects):
            w.add_from_array(f"{'robot' if k < n_robots else 'thing'}{k:02d}", np.array([[3 if k < n_robots else 2]], dtype=np.int64), 1 + 3 * k, 2)
        return w
    for n_objects, n_robots, msg in ((9, 9, "robots"), (65, 2, "dynamic objects")):
        B = capi.Batch(0)
        B.add_world(world(n_objects, n_robots))
        with pytest.raises(capi.MevoError, match=msg):
            B.finalize()
        B.close()
    B = capi.Batch(0)                       # 8 robots + 56 things: accepted, and steps
    B.add_world(world(64, 8))
    B.finalize()
    n_unstable, flags = B.step(2, apply_actions=False)
    assert n_unstable == 0 and not flags.any()
    B.close()
