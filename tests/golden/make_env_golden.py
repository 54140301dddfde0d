"""Generates tests/golden/env_*.npz: rollouts of the reference's OWN gym environments (Python, unmodified,
imported from /root/reference) over the reference's OWN engine (oracle/_ref, compiled from the unmodified C++
sources), for the task layer either side of the hot path:

    evogym/envs/walk.py:13-103        WalkingFlat      (Walker-v0)
    evogym/envs/manipulate.py:318-417 PushSmallRect    (Pusher-v0)
    evogym/envs/manipulate.py:519-628 ThrowSmallRect   (Thrower-v0)

each in the reference's flat-observation mode and in ModularEvoGym's "modular" mode
(evogym/envs/base.py:427-506,746-753).  Run HERE only (needs /root/reference); the fixtures travel.

What is stubbed, and why it does not touch the numbers:
  * `gym` is not installed: modularevogym_b200/gym_shim.py provides Env / spaces.Box / spaces.Dict / register; the
    envs use them for space declarations only.
  * numpy 2.x has no `np.float`: aliased to `float` (it was an alias of `float`).
  * `evogym.simulator_cpp` (the pybind11 module) is replaced by a class with the same methods that forwards every
    call to the compiled reference `Sim` through oracle/ref.py -- the C++ that runs is the reference's.
  * `Viewer` / `Camera` (OpenGL) are inert.
"""
from __future__ import annotations

import json
import os
import random
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path.insert(0, ROOT)


def install_stubs():
    from oracle.ref import RefSim

    # ---- evogym.simulator_cpp over the compiled reference
    class Sim:
        def __init__(self):
            self._r = None

        @staticmethod
        def get_version():
            pass

        def init(self, w, h):
            self._r = RefSim(world_w=int(w), world_h=int(h))
            self._range = {}

        def _add(self, grid, conn, name, x, y, robot):
            # evogym/sim.py:83-84 hands np.empty((2, 1)) -- uninitialised memory -- over for an object without connections (the 1 x 1
            # block of Traverser-v0); whatever it holds, the reference then reads cell indices out of it.  Pass no connection instead.
            conn = np.asarray(conn, dtype=np.float64).reshape(2, -1)
            n_cells = np.asarray(grid).size
            valid = np.isfinite(conn).all(axis=0) & (conn >= 0).all(axis=0) & (conn < n_cells).all(axis=0)
            conn = np.ascontiguousarray(conn[:, valid])
            ok = self._r.add_object(np.asarray(grid, dtype=np.float64), conn, name, x, y, robot)
            if ok:
                info = self._r.object_info(self._r.num_objects - 1)
                self._range[name] = (info["min_point"], info["max_point"])
            return ok

        def read_robot_from_array(self, grid, conn, name, x, y):
            return self._add(grid, conn, name, x, y, True)

        def read_object_from_array(self, grid, conn, name, x, y):
            return self._add(grid, conn, name, x, y, False)

        def step(self):
            return self._r.step()

        def set_action(self, name, action):
            self._r.set_action(name, action)

        def revert(self, t):
            self._r.revert(t)

        def get_time(self):
            return self._r.get_time()

        def pos_at_time(self, t):
            return np.asfortranarray(self._r.pos_at_time(t).T)

        def vel_at_time(self, t):
            return np.asfortranarray(self._r.pos_at_time(t, vel=True).T)

        # NB: explicit Sim.* calls -- EvoSim overrides pos_at_time (divides by VOXEL_SIZE) and a C++ member of the real
        # pybind11 class never dispatches to a Python override
        def object_pos_at_time(self, t, name):
            a, b = self._range[name]
            return Sim.pos_at_time(self, t)[:, a:b + 1]

        def object_vel_at_time(self, t, name):
            a, b = self._range[name]
            return Sim.vel_at_time(self, t)[:, a:b + 1]

        def object_orientation_at_time(self, t, name):
            return self._r.object_orientation_at_time(t, name)

        def translate_object(self, x, y, name):
            self._r.translate_object(x, y, name)

        def get_indices_of_actuators(self, name):
            return self._r.actuator_indices(name).reshape(1, -1)

    class Viewer:
        def __init__(self, sim): pass
        def init(self): pass
        def render(self, *a, **k): pass
        def show_debug_window(self): pass
        def hide_debug_window(self): pass
        def get_debug_window_pos(self): return (0, 0)

    class Camera:
        def __init__(self, b): pass
        def set_pos(self, x, y): pass
        def set_size(self, x, y): pass
        def set_resolution(self, x, y): pass
        def get_resolution_width(self): return 0
        def get_resolution_height(self): return 0
        def get_image(self): return None

    mod = types.ModuleType("evogym.simulator_cpp")
    mod.Sim, mod.Viewer, mod.Camera = Sim, Viewer, Camera
    # np.float alias, minimal gym, and `evogym.simulator_cpp` = the compiled REFERENCE engine (not this repo's):
    # registered before `import evogym` because evogym/sim.py:14 imports it
    from modularevogym_b200 import gym_shim
    gym_shim.install(simulator_module=mod)
    sys.path.insert(0, REF)


# DownStepper-v0 / ObstacleTraverser-v0 end the episode when the robot lies on its back (traverse.py:226-228, 325-327); random
# rollouts of 12 steps never get there, so before step FLIP_STEP the robot is turned by pi about its centre of mass and lifted clear
# of the ground (engine state only: the step's reward still starts from the last snapshot's centre of mass, which has the same x)
FLIP_TASKS, FLIP_STEP = ("obstaclev0", "downstepper", "hurdler", "platformjumper"), 6


# GapJumper-v0 ends the episode when the robot has fallen between the platforms (centre of mass below 4 voxels, traverse.py:734-738):
# before step FLIP_STEP the robot is set down left of the first platform, just above the ground plane
DROP_TASKS = ("gapjumper",)


# Catcher-v0 rewards the package's fall only below 5 voxels (manipulate.py:750-751), which the tumbling package reaches long after
# 12 steps: before step FLIP_STEP the package is set down beside the robot, just above the ground
LOWER_TASKS = ("catcher",)


def lower_package(env):
    r = env._sim._r
    pos, vel, last, vt = r.get_state()
    names = list(r.object_names)
    ri, pi = r.object_info(names.index("robot")), r.object_info(names.index("package"))
    rs, ps = slice(ri["min_point"], ri["max_point"] + 1), slice(pi["min_point"], pi["max_point"] + 1)
    target = np.array([pos[rs][:, 0].max() + 0.3, 0.12])
    pos[ps] += target - pos[ps].min(axis=0)
    vel[ps] = 0.0
    vt[ps] = 0.0
    r.set_state(pos, vel, pos.copy(), vt)


# An unstable step saves no snapshot (Sim.cpp:105-115): the environment then reads the snapshot of the step before -- for the
# observation it returns and for the "before" values of the steps that follow.  Before step FLIP_STEP the robot is crushed to 4 % of
# its size about its centre of mass (engine state only), which the engine reports as unstable; the rollout goes on without a reset.
CRUSH_TASKS = ("walkercrush",)


def crush_robot(env):
    r = env._sim._r
    pos, vel, last, vt = r.get_state()
    info = r.object_info(list(r.object_names).index("robot"))
    sl = slice(info["min_point"], info["max_point"] + 1)
    c = pos[sl].mean(axis=0)
    pos[sl] = c + 0.04 * (pos[sl] - c)
    r.set_state(pos, vel, pos.copy(), vt)


def drop_robot(env):
    r = env._sim._r
    pos, vel, last, vt = r.get_state()
    info = r.object_info(list(r.object_names).index("robot"))
    sl = slice(info["min_point"], info["max_point"] + 1)
    low = pos[sl].min(axis=0)
    pos[sl] += np.array([-1.0, 0.02]) - low
    r.set_state(pos, np.zeros_like(vel), pos.copy(), np.zeros_like(vt))


# Flipper-v0 counts whole turns when the orientation wraps between 0 and 2 pi (flip.py:78-85): before step FLIP_STEP the robot is
# turned by +0.3 rad about its centre of mass (it sits at about -0.01 rad = 6.27 by then) and lifted clear of the floor
# UpStepper-v0 ends the episode of a tilted robot in FLAT mode only, and in a band pi / 12 wider than the other environments'
# (traverse.py:124-129): its robot is turned by 1.4 rad -- inside (pi/2 - pi/12, pi/2) -- so the flat rollout ends there, the modular one goes on
TWIST_TASKS = ("flipper", "upstepper")
TWIST_ANGLE = {"flipper": 0.3, "upstepper": 1.4}


def twist_robot(env, angle=0.3):
    r = env._sim._r
    pos, vel, last, vt = r.get_state()
    info = r.object_info(list(r.object_names).index("robot"))
    sl = slice(info["min_point"], info["max_point"] + 1)
    c = pos[sl].mean(axis=0)
    d = pos[sl] - c
    ca, sa = np.cos(angle), np.sin(angle)
    pos[sl] = c + np.stack((ca * d[:, 0] - sa * d[:, 1], sa * d[:, 0] + ca * d[:, 1]), axis=1) + np.array([0.0, 0.12])
    r.set_state(pos, np.zeros_like(vel), pos.copy(), np.zeros_like(vt))


def flip_robot(env):
    r = env._sim._r
    pos, vel, last, vt = r.get_state()
    info = r.object_info(list(r.object_names).index("robot"))
    sl = slice(info["min_point"], info["max_point"] + 1)
    c = pos[sl].mean(axis=0)
    pos[sl] = 2.0 * c - pos[sl] + np.array([0.0, 0.6])
    r.set_state(pos, np.zeros_like(vel), pos.copy(), np.zeros_like(vt))


def main():
    import warnings
    warnings.filterwarnings("ignore")
    install_stubs()

    from evogym.envs.walk import Duck, SoftBridge, WalkingFlat
    from evogym.envs.manipulate import CarrySmallRect, CarrySmallRectToTable, CatchSmallRect, LiftSmallRect, PushSmallRect, SlideBeam, ToppleBeam, PushSmallRectOnOppositeSide, ThrowSmallRect
    from evogym.envs.climb import Climb0, Climb1, Climb2
    from evogym.envs.traverse import BlockSoup, FloatingPlatform, Gaps, StepsDown, StepsUp, VerticalBarrier, WalkingBumpy, WalkingBumpy2
    from evogym.envs.jump import StationaryJump
    from evogym.envs.balance import Balance, BalanceJump
    from evogym.envs.flip import Flipping
    from evogym.envs.change_shape import MaximizeShape, MaximizeXShape, MaximizeYShape, MinimizeShape
    from evogym import sample_robot

    tasks = {"walker": (WalkingFlat, "Walker-v0"), "pusher": (PushSmallRect, "Pusher-v0"), "thrower": (ThrowSmallRect, "Thrower-v0"),
             "climber": (Climb0, "Climber-v0"), "obstacle": (WalkingBumpy2, "ObstacleTraverser-v1"),
             "bridgewalker": (SoftBridge, "BridgeWalker-v0"), "obstaclev0": (WalkingBumpy, "ObstacleTraverser-v0"),
             "upstepper": (StepsUp, "UpStepper-v0"), "downstepper": (StepsDown, "DownStepper-v0"),
             "jumper": (StationaryJump, "Jumper-v0"), "hurdler": (VerticalBarrier, "Hurdler-v0"), "climberv1": (Climb1, "Climber-v1"),
             "cavecrawler": (Duck, "CaveCrawler-v0"), "climberv2": (Climb2, "Climber-v2"),
             "carrier": (CarrySmallRect, "Carrier-v0"), "pusherv1": (PushSmallRectOnOppositeSide, "Pusher-v1"),
             "balancer": (Balance, "Balancer-v0"), "balancerv1": (BalanceJump, "Balancer-v1"),
             "gapjumper": (Gaps, "GapJumper-v0"), "platformjumper": (FloatingPlatform, "PlatformJumper-v0"),
             "carrierv1": (CarrySmallRectToTable, "Carrier-v1"), "lifter": (LiftSmallRect, "Lifter-v0"),
             "flipper": (Flipping, "Flipper-v0"), "beamtoppler": (ToppleBeam, "BeamToppler-v0"), "beamslider": (SlideBeam, "BeamSlider-v0"),
             "wingspan": (MaximizeXShape, "WingspanMazimizer-v0"), "height": (MaximizeYShape, "HeightMaximizer-v0"),
             "areamax": (MaximizeShape, "AreaMaximizer-v0"), "areamin": (MinimizeShape, "AreaMinimizer-v0"),
             "traverser": (BlockSoup, "Traverser-v0"), "catcher": (CatchSmallRect, "Catcher-v0"),
             "walkercrush": (WalkingFlat, "Walker-v0")}
    only = [a for a in sys.argv[1:] if not a.startswith("-")]          # e.g. `make_env_golden.py bridgewalker`: only these tasks
    n_steps = 12
    for task, (cls, env_id) in tasks.items():
        if only and task not in only:
            continue
        for seed in ((0, 1, 2) if task in ("walker", "pusher", "thrower") else (0,)):
            random.seed(100 + seed)
            np.random.seed(100 + seed)
            body, conn = sample_robot((5, 5))
            rng = np.random.default_rng(500 + seed)
            out = {"body": body.astype(np.int32), "connections": np.asarray(conn, dtype=np.int32), "env_id": env_id}
            if task in FLIP_TASKS or task in DROP_TASKS or task in TWIST_TASKS or task in LOWER_TASKS or task in CRUSH_TASKS:
                out["flip_step"] = np.int32(FLIP_STEP)
            # ---- flat mode
            env = cls(body, conn, mode=None, env_id=env_id)
            n_act = env.get_actuator_indices("robot").size
            acts = rng.uniform(-1.0, 1.0, size=(n_steps, n_act))       # policy outputs; BenchmarkBase.step adds 1, base clips
            if task == "catcher":                       # CatchSmallRect.reset re-draws the package's launch position (manipulate.py:815):
                random.seed(1234)                       # the same draw for the flat and the modular environment
            obs0 = env.reset()
            flat_obs, rew, done = [np.asarray(obs0, dtype=np.float64)], [], []
            pre = {k: [] for k in ("pos", "vel", "pos_last", "vel_true", "goal")}
            for t in range(n_steps):
                if task in FLIP_TASKS and t == FLIP_STEP:
                    flip_robot(env)
                if task in DROP_TASKS and t == FLIP_STEP:
                    drop_robot(env)
                if task in TWIST_TASKS and t == FLIP_STEP:
                    twist_robot(env, TWIST_ANGLE[task])
                if task in LOWER_TASKS and t == FLIP_STEP:
                    lower_package(env)
                if task in CRUSH_TASKS and t == FLIP_STEP:
                    crush_robot(env)
                # engine state the step starts from (teacher forcing in the GPU test)
                st = env._sim._r.get_state()
                for k, v in zip(("pos", "vel", "pos_last", "vel_true"), st):
                    pre[k].append(v.copy())
                pre["goal"].append(env._sim._r.edges()["goal"].copy())
                o, r, d, _ = env.step(acts[t].copy())
                flat_obs.append(np.asarray(o, dtype=np.float64)); rew.append(float(r)); done.append(bool(d))
            from modularevogym_b200.world import WorldSpec
            out["world_json"] = json.dumps(WorldSpec.from_evoworld(env.world).to_dict())
            if task == "catcher":
                out["catch_offset"] = np.asarray([env.offsetx, env.offsety], dtype=np.int32)
            out.update({"pre_" + k: np.stack(v) for k, v in pre.items()})
            out.update(actions=acts, flat_obs=np.stack(flat_obs), reward=np.asarray(rew), done=np.asarray(done),
                       actuator_indices=env.get_actuator_indices("robot").astype(np.int32))
            # ---- modular mode, same policy outputs scattered over the voxel grid
            envm = cls(body, conn, mode="modular", env_id=env_id)
            act_mask = envm.act_mask.astype(bool)
            if task == "catcher":
                random.seed(1234)
            m0 = envm.reset()
            if task == "catcher":
                assert (envm.offsetx, envm.offsety) == (env.offsetx, env.offsety)
            mod_obs, other_obs, mrew, mdone = [np.asarray(m0["modular"], dtype=np.float64)], [np.asarray(m0["other"], dtype=np.float64)], [], []
            vox_acts = np.zeros((n_steps, body.size))
            for t in range(n_steps):
                vox_acts[t, ~act_mask] = acts[t]
                if task in FLIP_TASKS and t == FLIP_STEP:
                    flip_robot(envm)
                if task in DROP_TASKS and t == FLIP_STEP:
                    drop_robot(envm)
                if task in TWIST_TASKS and t == FLIP_STEP:
                    twist_robot(envm, TWIST_ANGLE[task])
                if task in LOWER_TASKS and t == FLIP_STEP:
                    lower_package(envm)
                if task in CRUSH_TASKS and t == FLIP_STEP:
                    crush_robot(envm)
                o, r, d, _ = envm.step(vox_acts[t].copy())
                mod_obs.append(np.asarray(o["modular"], dtype=np.float64)); other_obs.append(np.asarray(o["other"], dtype=np.float64))
                mrew.append(float(r)); mdone.append(bool(d))
            out.update(voxel_actions=vox_acts, modular_obs=np.stack(mod_obs), other_obs=np.stack(other_obs),
                       modular_reward=np.asarray(mrew), modular_done=np.asarray(mdone),
                       act_mask=envm.act_mask.astype(np.int32), obs_mask=np.asarray(envm.obs_mask).astype(np.int32),
                       obs_index_array=np.asarray(envm.obs_index_array, dtype=np.int32),
                       voxel_corner_size=np.int32(envm.voxel_corner_size))
            path = os.path.join(HERE, f"env_{task}_s{seed}.npz")
            np.savez_compressed(path, **out)
            print(path, "reward sum", float(np.sum(rew)), "modular width", out["modular_obs"].shape[1],
                  "flat == modular rewards:", np.array_equal(out["reward"], out["modular_reward"]))

    # ---- BidirectionalWalker-v0 (multi_goal.py:91-205): flat mode only (the reference has no modular mode for it).  Three goal lists:
    # the environment's own random draw; a list whose first goals lie within two voxels of the start, so that get_reward walks
    # through several goals in one step (multi_goal.py:50-53); and a list the robot exhausts on the first step (done, +1)
    if not only or "biwalk" in only:
        from evogym.envs.multi_goal import BiWalk, WalkToX
        for tag, forced in (("random", None), ("near", [36, 35, 34, 22, 48]), ("all", [35, 36])):
            random.seed(100)
            np.random.seed(100)
            body, conn = sample_robot((5, 5))
            rng = np.random.default_rng(500)
            env = BiWalk(body, conn)
            random.seed(4242)
            obs0 = env.reset()                          # draws a new list
            if forced is not None:
                env.init_reward_goals([WalkToX(g) for g in forced])
                env.current_goal = 0
                obs0 = np.concatenate((obs0[:-3], np.array([0]), env.goals[0].get_obs({"robot_com_pos_final": env.object_pos_at_time(env.get_time(), "robot")})))
            goals = [g.x_goal for g in env.goals]
            n_act = env.get_actuator_indices("robot").size
            acts = rng.uniform(-1.0, 1.0, size=(n_steps, n_act))
            flat_obs, rew, done, cur = [np.asarray(obs0, dtype=np.float64)], [], [], []
            pre = {k: [] for k in ("pos", "vel", "pos_last", "vel_true", "goal")}
            for t in range(n_steps):
                st = env._sim._r.get_state()
                for k, v in zip(("pos", "vel", "pos_last", "vel_true"), st):
                    pre[k].append(v.copy())
                pre["goal"].append(env._sim._r.edges()["goal"].copy())
                o, r, d, _ = env.step(acts[t].copy())
                flat_obs.append(np.asarray(o, dtype=np.float64)); rew.append(float(r)); done.append(bool(d)); cur.append(env.current_goal)
                if d:
                    break
            from modularevogym_b200.world import WorldSpec
            out = {"body": body.astype(np.int32), "connections": np.asarray(conn, dtype=np.int32), "env_id": "BidirectionalWalker-v0",
                   "world_json": json.dumps(WorldSpec.from_evoworld(env.world).to_dict()), "goals": np.asarray(goals, dtype=np.int32),
                   "current_goal": np.asarray(cur, dtype=np.int32), "actions": acts[:len(rew)], "flat_obs": np.stack(flat_obs),
                   "reward": np.asarray(rew), "done": np.asarray(done),
                   "actuator_indices": env.get_actuator_indices("robot").astype(np.int32)}
            out.update({"pre_" + k: np.stack(v) for k, v in pre.items()})
            path = os.path.join(HERE, f"biwalk_{tag}.npz")
            np.savez_compressed(path, **out)
            print(path, "goals", goals, "current", cur, "reward", rew[:3], "done", done[-1])
    if only:
        return
    # ---- terrain sight + orientation observations (base.py:258-372): ObstacleTraverser-v1 (floor, "ground"), CaveCrawler-v0
    # (floor + ceil, "terrain"), Climber-v2 (ceil, "pipe"); flat mode, the envs' own sight_dist
    from evogym.envs.traverse import WalkingBumpy2
    from evogym.envs.walk import Duck
    from evogym.envs.climb import Climb2
    from modularevogym_b200.world import WorldSpec
    for tag, cls, env_id, terrain in (("obstacle", WalkingBumpy2, "ObstacleTraverser-v1", "ground"), ("cave", Duck, "CaveCrawler-v0", "terrain"),
                                      ("climber2", Climb2, "Climber-v2", "pipe")):
        random.seed(200)
        np.random.seed(200)
        body, conn = sample_robot((5, 5))
        env = cls(body, conn, mode=None, env_id=env_id)
        rng = np.random.default_rng(900)
        n_act = env.get_actuator_indices("robot").size
        acts = rng.uniform(-1.0, 1.0, size=(n_steps, n_act))
        env.reset()
        rec = {k: [] for k in ("pos", "vel", "pos_last", "vel_true", "floor", "ceil", "orient", "vel_com")}

        def snap():
            st = env._sim._r.get_state()
            for k, v in zip(("pos", "vel", "pos_last", "vel_true"), st):
                rec[k].append(v.copy())
            rec["floor"].append(np.asarray(env.get_floor_obs("robot", [terrain], env.sight_dist), dtype=np.float64))
            rec["ceil"].append(np.asarray(env.get_ceil_obs("robot", [terrain], env.sight_dist), dtype=np.float64))
            rec["orient"].append(float(env.object_orientation_at_time(env.get_time(), "robot")))
            rec["vel_com"].append(np.asarray(env.get_vel_com_obs("robot"), dtype=np.float64))
        snap()
        for t in range(n_steps):
            env.step(acts[t].copy())
            snap()
        out = {"world_json": json.dumps(WorldSpec.from_evoworld(env.world).to_dict()), "env_id": env_id, "terrain": terrain,
               "sight_dist": np.int32(env.sight_dist), "body": body.astype(np.int32)}
        out.update({k: np.stack(v) if k != "orient" else np.asarray(v) for k, v in rec.items()})
        path = os.path.join(HERE, f"sight_{tag}.npz")
        np.savez_compressed(path, **out)
        print(path, "floor[0]", out["floor"][0], "ceil[0]", out["ceil"][0])


if __name__ == "__main__":
    main()
