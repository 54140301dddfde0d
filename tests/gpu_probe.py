"""Diagnostic driver for the first GPU runs (prints instead of asserting).  Not a pytest file.

  python tests/gpu_probe.py [n_envs_for_timing]
"""
import os
import random
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from modularevogym_b200 import capi  # noqa: E402
from modularevogym_b200.world import sample_robot, walker_world  # noqa: E402
from oracle.oracle import OracleSim  # noqa: E402


def norm_pairs_oracle(orc, cand):
    """oracle candidates -> set of unordered (obj,boxel) pairs without (a,a) and static-static."""
    out = set()
    ob = orc.candidates_as_object_boxel(cand)
    static = [(not o.is_robot) and bool((o.boxels[:, 0] == 5).all()) for o in orc.topo.objects]
    for oa, ba, o2, b2 in ob.tolist():
        if (oa, ba) == (o2, b2):
            continue
        if static[oa] and static[o2]:
            continue
        a, b = (oa, ba), (o2, b2)
        out.add((a, b) if a <= b else (b, a))
    return out


def norm_pairs_gpu(c):
    out = set()
    for oa, ba, o2, b2 in c.tolist():
        a, b = (oa, ba), (o2, b2)
        out.add((a, b) if a <= b else (b, a))
    return out


def parity(n_robots=4, steps=3):
    worlds = []
    for sd in range(n_robots):
        random.seed(sd)
        r, c = sample_robot((5, 5))
        worlds.append(walker_world(r, c))
    B = capi.Batch(0)
    for w in worlds:
        B.add_world(w)
    B.enable_candidate_dump(2048)
    B.finalize()
    oracles = [OracleSim(w) for w in worlds]
    rngs = [np.random.default_rng(1234 + i) for i in range(n_robots)]
    tot = exact = 0
    maxerr = 0.0
    cand_bad = 0
    for s in range(steps):
        for i, o in enumerate(oracles):
            idx = o.actuator_indices("robot")
            a = np.clip(rngs[i].uniform(-1, 1, len(idx)) + 1, 0.6, 1.6)
            o.set_action_values("robot", a)
            B.set_action_values(i, "robot", a)
        for sub in range(30):
            # teacher forcing: GPU state := oracle state
            for i, o in enumerate(oracles):
                st = o.get_state()
                B.set_state(i, *st)
                B.set_goal(i, o.edge_state()[1])
            B.step_substeps(1, 1, apply_actions=False)
            for i, o in enumerate(oracles):
                co = o.phase_collisions(True)
                o.phase_actuators(); o.phase_rk4(); o.phase_constraints(); o.phase_true_vel()
                uo = o.phase_self_collide()
                so = o.get_state()
                sg = B.get_state(i)
                ex = all((x == y).all() for x, y in zip(so, sg))
                tot += 1
                exact += ex
                for x, y in zip(so[:2], sg[:2]):
                    maxerr = max(maxerr, float(np.abs(x - y).max() / max(1.0, np.abs(x).max())))
                if not ex and o.contact_count() == 0 and tot < 200:
                    d = [float(np.abs(x - y).max()) for x, y in zip(so, sg)]
                    print(f"  inexact w/o contact env{i} step{s} sub{sub} maxabs={d}")
                pg, po = norm_pairs_gpu(B.read_candidates(i)), norm_pairs_oracle(o, co)
                if pg != po:
                    cand_bad += 1
                    if cand_bad < 5:
                        print(f"  cand mismatch env{i} step{s} sub{sub}: gpu-only {sorted(pg - po)[:5]} oracle-only {sorted(po - pg)[:5]}")
                if B.error_flags(i):
                    print("  error flag", i, B.error_flags(i))
    print(f"teacher-forced substeps: {tot}, bit-exact: {exact}, max rel err: {maxerr:.3e}, cand mismatches: {cand_bad}")

    # free running: 5 more steps through the batched action path
    for i, o in enumerate(oracles):
        st = o.get_state()
        B.set_state(i, *st)
        B.set_goal(i, o.edge_state()[1])
    stride = B.action_stride
    for s in range(5):
        acts = np.ones((len(oracles), stride))
        for i, o in enumerate(oracles):
            idx = o.actuator_indices("robot")
            a = np.clip(rngs[i].uniform(-1, 1, len(idx)) + 1, 0.6, 1.6)
            o.set_action_values("robot", a)
            acts[i, :len(idx)] = a
            o.step()
        B.set_actions(acts)
        n, flags = B.step(1, apply_actions=True)
        err = 0.0
        for i, o in enumerate(oracles):
            so, sg = o.get_state(), B.get_state(i)
            err = max(err, float(np.abs(so[0] - sg[0]).max()))
        print(f"free-run step {s}: unstable={n} max |dpos| = {err:.3e}  times={[B.get_time(i) for i in range(len(oracles))]}")


def timing(n_envs=4096, steps=20):
    t0 = time.time()
    B = capi.Batch(0)
    for i in range(n_envs):
        random.seed(i)
        r, c = sample_robot((5, 5))
        B.add_world(walker_world(r, c))
    B.finalize()
    print(f"built {n_envs} envs in {time.time() - t0:.1f}s, device bytes {B.device_bytes / 1e6:.1f} MB, stride {B.action_stride}")
    rng = np.random.default_rng(0)
    acts = np.clip(rng.uniform(-1, 1, (n_envs, B.action_stride)) + 1, 0.6, 1.6)
    B.set_actions(acts)
    B.step(1)
    B.sync()
    t0 = time.time()
    nun = 0
    for s in range(steps):
        acts = np.clip(rng.uniform(-1, 1, (n_envs, B.action_stride)) + 1, 0.6, 1.6)
        B.set_actions(acts)
        n, flags = B.step(1)
        nun += n
    dt = time.time() - t0
    print(f"{n_envs} envs x {steps} steps: {dt * 1e3 / steps:.2f} ms/step -> {n_envs * steps / dt:.0f} env-steps/s (unstable events {nun})")
    # kernel only
    t0 = time.time()
    B.step_async(steps, True)
    B.sync()
    dt = time.time() - t0
    print(f"kernel-only ({steps} steps in one launch): {dt * 1e3 / steps:.2f} ms/step -> {n_envs * steps / dt:.0f} env-steps/s")


if __name__ == "__main__":
    print("devices:", capi.load().mevo_device_count())
    parity()
    timing(int(sys.argv[1]) if len(sys.argv) > 1 else 4096)
