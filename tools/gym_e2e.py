#!/usr/bin/env python
"""SURVEY 8f-3 end to end: the reference's OWN Python layer -- `gym.make(env_id, body=..., mode=...)` -> `evogym/envs/*.py`
-> `evogym.EvoSim` (`evogym/sim.py:18-252`) -- driving THIS repo's `simulator_cpp.Sim` (the CUDA engine) through real steps.

    python tools/gym_e2e.py [--pyref DIR] [--out FILE]

`modularevogym_b200.gym_shim.install()` registers this repo's pybind11 module as `evogym.simulator_cpp` (plus `np.float` and a
minimal `gym` where missing); the reference's unmodified Python files are imported from --pyref (default: oracle/_ref/pyref,
staged by `make -C oracle pyref`; /root/reference where that exists).  For every committed environment fixture
(tests/golden/env_*.npz: rollouts of the same Python environments over the reference's own engine) the script makes the
environment through gym's registry, replays the stored policy outputs in flat and in modular mode, and compares reward, done
and observations step by step (free running, no teacher forcing).  Then the reference's smoke script
(`examples/modularevogym_test.py:1-37`: make, `env.update` with a second robot, random rollout until done) on every fixture
environment.  Prints one JSON line per environment and a summary; exit code 1 on a mismatch.
"""
import argparse
import glob
import json
import os
import random
import sys
import warnings

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

TOL_REWARD = 1e-5        # free-running 12 env-steps (360 substeps) through contacts; teacher-forced parity is test_task_layer.py's
TOL_OBS = 1e-5


def rel_err(x, y):
    x, y = np.asarray(x, dtype=np.float64), np.asarray(y, dtype=np.float64)
    return float(np.abs(x - y).max() / max(1.0, float(np.abs(y).max()))) if x.size else 0.0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pyref", default="")
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    warnings.filterwarnings("ignore")
    pyref = args.pyref or next((p for p in (os.path.join(ROOT, "oracle", "_ref", "pyref"), "/root/reference") if os.path.isdir(os.path.join(p, "evogym"))), "")
    if not pyref:
        print(json.dumps({"skipped": "the reference's Python layer is not here (make -C oracle pyref)"}))
        return 0
    from modularevogym_b200 import gym_shim, simulator_cpp
    installed = gym_shim.install()
    sys.path.insert(0, pyref)
    import gym
    import evogym.envs  # noqa: F401  (registers the 32 environment ids)
    from evogym import EvoSim, sample_robot
    assert issubclass(EvoSim, simulator_cpp.Sim), "EvoSim must subclass this repo's Sim"
    log = [{"installed": installed, "pyref": pyref, "evogym": os.path.dirname(evogym.envs.__file__)}]
    print(json.dumps(log[0]), flush=True)
    ok = True

    for path in sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "env_*.npz"))):
        with np.load(path) as z:
            g = {k: z[k] for k in z.files}
        env_id = str(g["env_id"])
        body, conn = g["body"], g["connections"]
        rec = {"fixture": os.path.basename(path), "env_id": env_id}
        # fixtures that turn the robot over in the reference ENGINE before step `flip_step` (make_env_golden.py: the orientation rule
        # of DownStepper-v0 / ObstacleTraverser-v0) are replayed up to that step: the pybind surface has no state setter
        n_replay = int(g["flip_step"]) if "flip_step" in g else g["actions"].shape[0]
        rec["steps"] = n_replay
        # ---- flat mode: gym.make through the registry (TimeLimit wrapper included), stored policy outputs replayed
        env = gym.make(env_id, body=body, connections=conn)
        assert isinstance(env.unwrapped._sim, simulator_cpp.Sim)
        if "catch_offset" in g:                 # Catcher-v0 re-draws the package's launch position in reset(): the fixture's seed (make_env_golden.py)
            random.seed(1234)
        o = env.reset()
        if "catch_offset" in g:
            assert [env.unwrapped.offsetx, env.unwrapped.offsety] == g["catch_offset"].tolist()
        worst_r, worst_o = 0.0, rel_err(o, g["flat_obs"][0])
        dones = []
        for t in range(n_replay):
            o, r, d, _ = env.step(g["actions"][t].copy())
            worst_r = max(worst_r, abs(float(r) - float(g["reward"][t])))
            worst_o = max(worst_o, rel_err(o, g["flat_obs"][t + 1]))
            dones.append(bool(d))
        rec.update(flat_reward_err=worst_r, flat_obs_err=worst_o, flat_done_equal=dones == g["done"].astype(bool).tolist()[:n_replay])
        env.close()
        # ---- modular mode (ModularEvoGym's per-voxel observation / action layout)
        env = gym.make(env_id, body=body, connections=conn, mode="modular", env_id=env_id)
        if "catch_offset" in g:
            random.seed(1234)
        o = env.reset()
        worst_r, worst_o = 0.0, max(rel_err(o["modular"], g["modular_obs"][0]), rel_err(o["other"], g["other_obs"][0]))
        dones = []
        for t in range(n_replay):
            o, r, d, _ = env.step(g["voxel_actions"][t].copy())
            worst_r = max(worst_r, abs(float(r) - float(g["modular_reward"][t])))
            worst_o = max(worst_o, rel_err(o["modular"], g["modular_obs"][t + 1]), rel_err(o["other"], g["other_obs"][t + 1]))
            dones.append(bool(d))
        rec.update(modular_reward_err=worst_r, modular_obs_err=worst_o, modular_done_equal=dones == g["modular_done"].astype(bool).tolist()[:n_replay],
                   masks_equal=bool(np.array_equal(env.unwrapped.act_mask, g["act_mask"]) and np.array_equal(env.unwrapped.obs_mask, g["obs_mask"])))
        env.close()
        good = (rec["flat_reward_err"] <= TOL_REWARD and rec["flat_obs_err"] <= TOL_OBS and rec["modular_reward_err"] <= TOL_REWARD and
                rec["modular_obs_err"] <= TOL_OBS and rec["flat_done_equal"] and rec["modular_done_equal"] and rec["masks_equal"])
        rec["ok"] = bool(good)
        ok &= good
        log.append(rec)
        print(json.dumps(rec), flush=True)

    # ---- BidirectionalWalker-v0 (flat mode only in the reference): the fixture's goal list is put in place of the one reset() drew
    from evogym.envs.multi_goal import WalkToX
    for path in sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "biwalk_*.npz"))):
        with np.load(path) as z:
            g = {k: z[k] for k in z.files}
        env = gym.make("BidirectionalWalker-v0", body=g["body"], connections=g["connections"])
        assert isinstance(env.unwrapped._sim, simulator_cpp.Sim)
        o = env.reset()
        env.unwrapped.init_reward_goals([WalkToX(int(x)) for x in g["goals"]])
        env.unwrapped.current_goal = 0
        worst_r, worst_o, dones, currents = 0.0, rel_err(o[:-3], g["flat_obs"][0][:-3]), [], []
        for t in range(g["reward"].shape[0]):
            o, r, d, _ = env.step(g["actions"][t].copy())
            worst_r = max(worst_r, abs(float(r) - float(g["reward"][t])))
            worst_o = max(worst_o, rel_err(o, g["flat_obs"][t + 1]))
            dones.append(bool(d)); currents.append(int(env.unwrapped.current_goal))
        env.close()
        rec = {"fixture": os.path.basename(path), "env_id": "BidirectionalWalker-v0", "steps": len(dones), "flat_reward_err": worst_r,
               "flat_obs_err": worst_o, "flat_done_equal": dones == g["done"].astype(bool).tolist(),
               "goal_index_equal": currents == g["current_goal"].tolist()}
        rec["ok"] = bool(worst_r <= TOL_REWARD and worst_o <= TOL_OBS and rec["flat_done_equal"] and rec["goal_index_equal"])
        ok &= rec["ok"]
        log.append(rec)
        print(json.dumps(rec), flush=True)

    # ---- the reference's smoke script (examples/modularevogym_test.py:1-37) on the fixture environments
    random.seed(0)
    np.random.seed(0)
    for env_id in ("Walker-v0", "Pusher-v0", "Thrower-v0", "Climber-v0", "ObstacleTraverser-v1", "BridgeWalker-v0", "ObstacleTraverser-v0", "UpStepper-v0",
                   "DownStepper-v0", "Jumper-v0", "Hurdler-v0", "Climber-v1", "CaveCrawler-v0",
                   "Climber-v2", "Carrier-v0", "Pusher-v1", "Balancer-v0", "Balancer-v1", "GapJumper-v0",
                   "PlatformJumper-v0", "Carrier-v1", "Lifter-v0", "Flipper-v0",
                   "BeamToppler-v0", "BeamSlider-v0", "WingspanMazimizer-v0", "HeightMaximizer-v0",
                   "AreaMaximizer-v0", "AreaMinimizer-v0", "Traverser-v0", "Catcher-v0"):
        body, connections = sample_robot((5, 5))
        env = gym.make(env_id, body=body)
        env = gym.make(env_id, body=body, mode="modular", env_id=env_id)
        env.reset()
        new_body, new_connections = sample_robot((5, 5))
        env.update(body=new_body, connections=new_connections)          # design-loop rebuild (evogym/envs/walk.py:105-120)
        env.reset()
        steps, ret = 0, 0.0
        while True:
            ob, reward, done, info = env.step(np.random.uniform(low=-1.0, high=1.0, size=25))
            steps += 1
            ret += float(reward)
            if done:
                break
        env.close()
        rec = {"smoke": env_id, "steps_until_done": steps, "return": ret, "finite": bool(np.isfinite(ret) and np.isfinite(ob["modular"]).all()),
               "truncated": bool(info.get("TimeLimit.truncated", False))}
        ok &= rec["finite"]
        log.append(rec)
        print(json.dumps(rec), flush=True)

    # BidirectionalWalker-v0 has neither a modular mode nor update(): the same rollout in flat mode
    body, connections = sample_robot((5, 5))
    env = gym.make("BidirectionalWalker-v0", body=body, connections=connections)
    ob = env.reset()
    steps, ret = 0, 0.0
    while True:
        ob, reward, done, info = env.step(np.random.uniform(low=-1.0, high=1.0, size=env.action_space.shape[0]))
        steps += 1
        ret += float(reward)
        if done:
            break
    env.close()
    rec = {"smoke": "BidirectionalWalker-v0", "steps_until_done": steps, "return": ret, "finite": bool(np.isfinite(ret) and np.isfinite(ob).all()),
           "truncated": bool(info.get("TimeLimit.truncated", False)), "mode": "flat"}
    ok &= rec["finite"]
    log.append(rec)
    print(json.dumps(rec), flush=True)

    summary = {"summary": "ok" if ok else "MISMATCH", "fixtures": sum(1 for r in log if "fixture" in r), "tolerance": {"reward": TOL_REWARD, "obs": TOL_OBS}}
    print(json.dumps(summary), flush=True)
    if args.out:
        with open(args.out, "w") as f:
            for r in log + [summary]:
                f.write(json.dumps(r) + "\n")
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
