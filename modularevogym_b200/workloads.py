"""Synthetic workloads of BASELINE.json's configurations, shared by ``bench.py``, the tools and the tests.

Every workload is a rule ``(name, i) -> WorldSpec`` for environment ``i`` (a global index: rank * envs_per_gpu + local index),
seeded so that the GPU arm, the CPU reference arm and the parity sweep build the same robots:

==============  ====================================================================================================
``walker5x5``   configs[1]: Walker-v0 ground + ``random.seed(i); sample_robot((5, 5))``
``pusher``      configs[2]: Pusher-v0 (ground + 3x2 package) + ``random.seed(i); sample_robot((5, 5))``
``thrower``     configs[2]: Thrower-v0 (ground with posts + package) + the same robots
``obstacleS``   configs[3]: ObstacleTraverser-v1 + ``random.seed(i); sample_large_robot((S, S))``  (S = 7..20, e.g. obstacle10)
``climberS``    configs[3]: Climber-v0-style pipe (widened to S + 2 for S > 5) + the same robots
``bridge``      mixed FIXED + SOFT terrain of BridgeWalker-v0 (simulated as a whole) + 5x5 robots
``hetero``      configs[4]: Walker-v0 ground + robot of side s_i ~ U{5..20} (``default_rng(i)``)
==============  ====================================================================================================

Action law of all of them: ``a ~ U(-1, 1) + 1`` clipped to ``[0.6, 1.6]`` per actuator per step (``evogym/envs/base.py:42,389``).
"""
from __future__ import annotations

import random
import re
from typing import Tuple

import numpy as np

from .world import (WorldSpec, bridge_walker_world, climber_world, obstacle_traverser_world, pusher_world, sample_large_robot, sample_robot, thrower_world,
                    walker_world)

def robot_for(name: str, i: int) -> Tuple[np.ndarray, np.ndarray]:
    random.seed(i)
    m = re.fullmatch(r"(obstacle|climber)(\d+)", name)
    if m:
        s = int(m.group(2))
        return sample_robot((s, s)) if s <= 6 else sample_large_robot((s, s))
    if name == "hetero":
        s = int(np.random.default_rng(i).integers(5, 21))
        return sample_robot((s, s)) if s <= 6 else sample_large_robot((s, s))
    return sample_robot((5, 5))


def world_for(name: str, i: int) -> WorldSpec:
    r, c = robot_for(name, i)
    if name in ("walker5x5", "hetero"):
        return walker_world(r, c)
    if name == "pusher":
        return pusher_world(r, c)
    if name == "thrower":
        return thrower_world(r, c)
    if name.startswith("obstacle"):
        return obstacle_traverser_world(r, c)
    if name.startswith("climber"):
        return climber_world(r, c)
    if name == "bridge":
        return bridge_walker_world(r, c)
    raise ValueError(f"unknown workload {name!r}")


def describe(name: str) -> str:
    law = "; actions U(-1,1)+1 clipped to [0.6,1.6] per actuator per step"
    m = re.fullmatch(r"(obstacle|climber)(\d+)", name)
    if m:
        s = m.group(2)
        terr = ("ObstacleTraverser-v1 (60x10 bumpy FIXED ground)" if m.group(1) == "obstacle"
                else f"Climber-v0-style FIXED pipe ({int(s) + 2} wide inside walls: the reference's 7-wide pipe cannot hold a {s}x{s} robot)")
        return f"{terr} + robot i = sample_large_robot(({s},{s})) under random.seed(i)" + law
    return {
        "walker5x5": "Walker-v0 (100x1 FIXED ground) + random 5x5 robot per env, robot i = sample_robot((5,5)) under random.seed(i)",
        "pusher": "Pusher-v0 (100x1 FIXED ground + 3x2 SOFT package) + robot i = sample_robot((5,5)) under random.seed(i)",
        "thrower": "Thrower-v0 (70x10 FIXED ground with posts + 3x2 SOFT package) + robot i = sample_robot((5,5)) under random.seed(i)",
        "bridge": "BridgeWalker-v0 terrain (FIXED banks + SOFT deck, simulated as a whole) + robot i = sample_robot((5,5)) under random.seed(i)",
        "hetero": ("BASELINE.json configs[4] per GPU: Walker-v0 ground + robot i of side s_i ~ U{5..20} (default_rng(i)), "
                   "sample_robot((s,s)) for s <= 6 else sample_large_robot((s,s)) under random.seed(i)"),
    }[name] + law


def clipped_actions(rng, shape) -> np.ndarray:
    return np.clip(rng.uniform(-1.0, 1.0, shape) + 1.0, 0.6, 1.6)     # evogym/envs/base.py:42,389
