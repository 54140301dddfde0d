"""GPU: seeded parity sweep of the BENCHMARKED workloads against the LIVE compiled reference.

``oracle/_ref/libmevo_ref.so`` -- the unmodified reference engine, built by ``oracle/Makefile`` where
``/root/reference`` exists -- travels to the GPU box with the repo (it is git-ignored, not gpurun-ignored), so this test
drives it there, side by side with the CUDA engine through the C ABI, on the robots ``bench.py`` times:

  (a) Walker-v0, bench robots ``random.seed(i); sample_robot((5,5))``, i = 0..255           (BASELINE configs[1])
  (b) 64 random 5x5 robots each on Pusher-v0 and Thrower-v0                                  (configs[2])
  (c) ``sample_large_robot`` of every side 7..20, two seeds each, on ObstacleTraverser-v1 and in a Climber-v0-style
      pipe (the reference's pipe is 5 voxels wide inside, so for these robots it is widened to side + 2)   (configs[3], [4])
  (d) 5x5 robots on BridgeWalker-v0's mixed FIXED + SOFT terrain (simulated as a whole)

All of them live in ONE heterogeneous batch, so every size class of both kernels runs in the same launches.  The
reference is stepped one ``Environment::step`` at a time with the bench's action law (U(-1,1) + 1 clipped to [0.6, 1.6]
from ``default_rng(1234 + i)``) for 3 env-steps; before every substep the CUDA state is reset to the reference's
(teacher forcing) and after it pos / vel / vel_true / collision_forces / unstable flag are compared, candidate pair sets
at six substeps; then both free-run 2 env-steps.  Skipped when ``oracle/_ref`` is absent.
"""
import json
import os
import random

import numpy as np
import pytest

from helpers import GOLDEN_DIR, Golden, pair_set, rel_err
from modularevogym_b200 import capi
from modularevogym_b200.world import (WorldSpec, climber_world, obstacle_traverser_world, pusher_world, sample_large_robot,
                                      sample_robot, thrower_world, walker_world)
from oracle import ref as oref

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not oref.available(), reason="oracle/_ref/libmevo_ref.so not built")]

TF_POS_TOL = 1e-14
TF_VEL_TOL = 1e-11
TF_VT_TOL = 1e-9
TF_CF_TOL = 1e-11         # relative to max(1, |collision_forces| max): forces are O(1e4), summed in a different order
FREE_POS_TOL = 1e-7
N_ENV_STEPS = 3
CAND_AT = (0, 1, 29, 30, 60, 89)


def sweep_cases(scale: int = 1):
    """[(class name, seed, WorldSpec)]; scale < 1 shrinks the sweep (development)."""
    out = []
    for i in range(256 // scale):
        random.seed(i)
        r, c = sample_robot((5, 5))
        out.append(("walker5x5", i, walker_world(r, c)))
    for i in range(64 // scale):
        random.seed(1000 + i)
        r, c = sample_robot((5, 5))
        out.append(("pusher5x5", 1000 + i, pusher_world(r, c)))
        random.seed(2000 + i)
        r, c = sample_robot((5, 5))
        out.append(("thrower5x5", 2000 + i, thrower_world(r, c)))
    for side in range(7, 21):
        for k in range(2 if scale == 1 else 1):
            seed = 3000 + 10 * side + k
            random.seed(seed)
            r, c = sample_large_robot((side, side))
            out.append((f"obstacle{side}", seed, obstacle_traverser_world(r, c)))
            random.seed(seed + 5)
            r, c = sample_large_robot((side, side))
            out.append((f"climber{side}", seed + 5, climber_world(r, c)))
    bridge = Golden("bridge_s6").world
    for k in range(4 // min(scale, 4)):
        random.seed(4000 + k)
        r, c = sample_robot((5, 5))
        w = WorldSpec.from_dict(bridge.to_dict())
        del w.objects["robot"]
        w.add_from_array("robot", r, 2, 5, c)
        out.append(("bridge5x5", 4000 + k, w))
    return out


def test_sweep_against_live_reference():
    scale = int(os.environ.get("MEVO_SWEEP_SCALE", "1"))
    cases = sweep_cases(scale)
    refs = [oref.RefSim(w) for _, _, w in cases]
    B = capi.Batch(0)
    for _, _, w in cases:
        B.add_world(w)
    B.enable_candidate_dump(16384)
    B.enable_force_dump()
    B.finalize()
    n = len(cases)
    rngs = [np.random.default_rng(1234 + seed) for _, seed, _ in cases]
    act_idx = [r.actuator_indices("robot") for r in refs]
    static_objs = [[bool(B.object_info(i, o)["is_static"]) for o in range(B.num_objects(i))] for i in range(n)]
    dyn_mask = []
    for i in range(n):
        m = np.ones(B.num_points(i), dtype=bool)
        for o in range(B.num_objects(i)):
            info = B.object_info(i, o)
            if info["is_static"]:
                m[info["min_point"]:info["max_point"] + 1] = False
        dyn_mask.append(m)

    stats = {}

    def stat(cls):
        return stats.setdefault(cls, {"envs": 0, "substeps": 0, "bit_identical": 0, "contact_substeps": 0, "pos": 0.0, "vel": 0.0, "vt": 0.0,
                                      "cf": 0.0, "free_pos": 0.0, "unstable": 0})

    for cls, _, _ in cases:
        stat(cls)["envs"] += 1

    def set_actions():
        for i, r in enumerate(refs):
            a = np.clip(rngs[i].uniform(-1.0, 1.0, len(act_idx[i])) + 1.0, 0.6, 1.6)
            rows = np.stack((act_idx[i].astype(np.float64), a), axis=1)
            r.set_action("robot", rows)
            assert B.set_action(i, "robot", rows)

    live = [True] * n       # an environment the reference reports unstable leaves the sweep (its state is garbage from there)
    for k in range(30 * N_ENV_STEPS):
        if k % 30 == 0:
            set_actions()
        want_cands = {}
        for i, r in enumerate(refs):
            if not live[i]:
                continue
            B.set_state(i, *r.get_state())
            if k in CAND_AT:
                want_cands[i] = pair_set(r.candidates(), static_objs[i], drop_self=True)
        _, flags = B.step_substeps(1, 1, apply_actions=False)
        for i, r in enumerate(refs):
            if not live[i]:
                continue
            cls = cases[i][0]
            st = stat(cls)
            unstable = r.substep()
            assert bool(flags[i]) == unstable, (cls, cases[i][1], k, "unstable flag")
            pos_r, vel_r, _, vt_r = r.get_state()
            cf_r = r.collision_forces()
            pos, vel, last, vt = B.get_state(i)
            cf = B.read_collision_forces(i)
            m = dyn_mask[i]
            e = (rel_err(pos, pos_r), rel_err(vel, vel_r), rel_err(vt, vt_r), rel_err(cf[m], cf_r[m]))
            assert e[0] <= TF_POS_TOL and e[1] <= TF_VEL_TOL and e[2] <= TF_VT_TOL and e[3] <= TF_CF_TOL, (cls, cases[i][1], k, e)
            assert np.array_equal(last, pos)
            st["substeps"] += 1
            st["bit_identical"] += int(np.array_equal(pos, pos_r) and np.array_equal(vel, vel_r))
            st["contact_substeps"] += int(np.abs(cf_r[m]).sum() > 0)
            for key, v in zip(("pos", "vel", "vt", "cf"), e):
                st[key] = max(st[key], v)
            if i in want_cands:
                got = pair_set(B.read_candidates(i), static_objs[i], drop_self=True)
                assert got == want_cands[i], (cls, cases[i][1], k, sorted(got - want_cands[i])[:4], sorted(want_cands[i] - got)[:4])
            if unstable:
                live[i] = False
                st["unstable"] += 1

    # free-running: 2 env-steps of Sim::step from the reference's state
    for i, r in enumerate(refs):
        if live[i]:
            B.set_state(i, *r.get_state())
    for s in range(2):
        set_actions()
        _, flags = B.step(1, apply_actions=False)
        for i, r in enumerate(refs):
            if not live[i]:
                continue
            unstable = r.step()
            assert bool(flags[i]) == unstable, (cases[i][0], cases[i][1], "free step", s)
            if unstable:
                live[i] = False
                continue
            e = rel_err(B.get_state(i)[0], r.get_state()[0])
            assert e <= FREE_POS_TOL, (cases[i][0], cases[i][1], "free step", s, e)
            stat(cases[i][0])["free_pos"] = max(stat(cases[i][0])["free_pos"], e)
    B.close()

    def group(name):
        return "".join(ch for ch in name if not ch.isdigit()) if name[-1].isdigit() and name not in ("walker5x5", "pusher5x5", "thrower5x5", "bridge5x5") else name
    merged = {}
    for cls, st in stats.items():
        g = merged.setdefault(group(cls), {k: 0 for k in st})
        for k, v in st.items():
            g[k] = g[k] + v if k in ("envs", "substeps", "bit_identical", "contact_substeps", "unstable") else max(g[k], v)
    print()
    for cls, st in merged.items():
        print(f"[sweep] {cls:12s} envs {st['envs']:4d}  substeps {st['substeps']:6d}  bit-identical {st['bit_identical']:6d}  with contacts "
              f"{st['contact_substeps']:6d}  worst rel err pos {st['pos']:.1e} vel {st['vel']:.1e} vel_true {st['vt']:.1e} cf {st['cf']:.1e}  "
              f"free-run pos {st['free_pos']:.1e}  unstable {st['unstable']}")
    out_dir = os.path.join(os.path.dirname(GOLDEN_DIR), "..", "gpurun_out")
    try:
        os.makedirs(out_dir, exist_ok=True)
        with open(os.path.join(out_dir, "parity_sweep.json"), "w") as f:
            json.dump({"per_class": stats, "merged": merged, "tolerances": {"pos": TF_POS_TOL, "vel": TF_VEL_TOL, "vel_true": TF_VT_TOL,
                                                                            "collision_forces": TF_CF_TOL, "free_pos": FREE_POS_TOL}}, f, indent=1)
    except OSError:
        pass
    # substeps without contacts must be bit-identical
    for cls, st in merged.items():
        assert st["bit_identical"] >= st["substeps"] - st["contact_substeps"], (cls, st)
