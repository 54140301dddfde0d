"""Shared helpers of the test-suite: fixture loading and candidate-set normalisation."""
from __future__ import annotations

import glob
import json
import os
from typing import Dict, List, Set, Tuple

import numpy as np

from modularevogym_b200.world import WorldSpec

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden_names() -> List[str]:
    """engine fixtures (make_golden.py); env_* / sight_* / biwalk_* are the task-layer fixtures (make_env_golden.py), worlds_* the world-file
    topologies (make_world_golden.py)"""
    names = sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN_DIR, "*.npz")))
    return [n for n in names if not n.startswith(("env_", "sight_", "worlds_", "adv_", "biwalk_"))]


def env_golden_names() -> List[str]:
    return sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN_DIR, "env_*.npz")))


class EnvGolden:
    """One rollout of a reference gym environment (layout: tests/golden/make_env_golden.py)."""

    def __init__(self, name: str):
        self.name = name
        with np.load(os.path.join(GOLDEN_DIR, name + ".npz")) as z:
            self.z = {k: z[k] for k in z.files}
        self.world = WorldSpec.from_dict(json.loads(str(self.z["world_json"])))
        self.env_id = str(self.z["env_id"])
        self.task = {"Walker-v0": 0, "Pusher-v0": 1, "Thrower-v0": 2, "Climber-v0": 3, "ObstacleTraverser-v1": 4,
                     "BridgeWalker-v0": 5, "ObstacleTraverser-v0": 6, "UpStepper-v0": 7, "DownStepper-v0": 8,
                     "Jumper-v0": 9, "Hurdler-v0": 10, "Climber-v1": 11, "CaveCrawler-v0": 12,
                     "Climber-v2": 13, "Carrier-v0": 14, "Pusher-v1": 15,
                     "Balancer-v0": 16, "Balancer-v1": 17, "GapJumper-v0": 18, "PlatformJumper-v0": 19,
                     "Carrier-v1": 20, "Lifter-v0": 21, "Flipper-v0": 22,
                     "BeamToppler-v0": 23, "BeamSlider-v0": 24, "WingspanMazimizer-v0": 25, "HeightMaximizer-v0": 26,
                     "AreaMaximizer-v0": 27, "AreaMinimizer-v0": 28, "Traverser-v0": 30, "Catcher-v0": 31}[self.env_id]
        self.package = "package" if self.task in (1, 2, 14, 15, 20, 21, 31) else "beam" if self.task in (23, 24) else None
        self.n_steps = int(self.z["reward"].shape[0])

    def __getitem__(self, k):
        return self.z[k]


class Golden:
    """One committed fixture (layout: tests/golden/make_golden.py::record)."""

    def __init__(self, name: str):
        self.name = name
        with np.load(os.path.join(GOLDEN_DIR, name + ".npz")) as z:
            self.z = {k: z[k] for k in z.files}       # decompress once
        self.world = WorldSpec.from_dict(json.loads(str(self.z["world_json"])))
        self.obj_info = self.z["obj_info"]
        self.n_obj = int(self.obj_info.shape[0])
        self.P = int(self.z["pos0"].shape[0])
        self.E = int(self.z["ea"].shape[0])
        self.n_sub = int(self.z["sub_pos"].shape[0])
        self.n_steps = int(self.z["step_unstable"].shape[0])

    def __getitem__(self, k):
        return self.z[k]

    def static_objects(self) -> List[bool]:
        """objects made of FIXED voxels only that are not robots (never move)"""
        out = []
        for o in range(self.n_obj):
            bx = self.z[f"boxels_{o}"]
            out.append(bool((bx[:, 0] == 5).all()) and not bool(self.obj_info[o, 0]))
        return out

    def state_before_substep(self, k: int):
        """(pos, vel, pos_last, vel_true) the reference had when substep k started"""
        if k == 0:
            p, v = self.z["init_pos"], self.z["init_vel"]
            return p, v, p, v
        p = self.z["sub_pos"][k - 1]
        return p, self.z["sub_vel"][k - 1], p, self.z["sub_vt"][k - 1]

    def action_rows(self, step: int) -> np.ndarray:
        """(A, 2) rows (grid_index, value) as EvoSim.set_action stacks them (evogym/sim.py:141)"""
        return np.stack((self.z["act_idx"].astype(np.float64), self.z["actions"][step]), axis=1)


Pair = Tuple[Tuple[int, int], Tuple[int, int]]


def pair_set(cands: np.ndarray, static: List[bool] = None, drop_self: bool = False) -> Set[Pair]:
    """(n,4) rows (obj_a, boxel_a, obj_b, boxel_b) -> set of unordered pairs.  With ``static`` given,
    pairs of two static objects are dropped (the engine never generates them: they cannot move
    anything); ``drop_self`` drops (a, a)."""
    out: Set[Pair] = set()
    for oa, ba, ob, bb in np.asarray(cands).reshape(-1, 4).tolist():
        a, b = (oa, ba), (ob, bb)
        if drop_self and a == b:
            continue
        if static is not None and static[oa] and static[ob]:
            continue
        out.add((a, b) if a <= b else (b, a))
    return out


def rel_err(x: np.ndarray, y: np.ndarray) -> float:
    """max |x - y| / max(1, max |y|): relative to the array's scale, absolute below 1"""
    x, y = np.asarray(x), np.asarray(y)
    if x.size == 0:
        return 0.0
    return float(np.abs(x - y).max() / max(1.0, float(np.abs(y).max())))
