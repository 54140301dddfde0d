"""TEST INFRASTRUCTURE ONLY: ctypes view of ``oracle/libmevo_oracle.so`` (the plain-C restatement
in ``mevo_oracle.c``) plus the Python-side bookkeeping of ``Sim`` (sim_time, snapshots).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline leg may import this.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Dict, Optional, Tuple

import numpy as np

from .topology import Topology, build_topology

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmevo_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "mevo_oracle.c")
    hdr = os.path.join(_HERE, "mevo_oracle.h")
    if (force or not os.path.exists(LIB_PATH)
            or os.path.getmtime(LIB_PATH) < max(os.path.getmtime(src), os.path.getmtime(hdr))):
        subprocess.check_call(["make", "-C", _HERE, "oracle"], stdout=subprocess.DEVNULL)
    return LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB_PATH)
        vp, i, d = C.c_void_p, C.c_int, C.c_double
        dp, ip, bp = C.POINTER(C.c_double), C.POINTER(C.c_int), C.POINTER(C.c_ubyte)
        sig = {
            "mo_create": (vp, [i, i, i, i, dp, dp, bp, ip, ip, dp, dp, bp, ip, ip]),
            "mo_destroy": (None, [vp]),
            "mo_get_state": (None, [vp, dp, dp, dp, dp]),
            "mo_set_state": (None, [vp, dp, dp, dp, dp]),
            "mo_get_edge_state": (None, [vp, dp, dp]),
            "mo_set_edge_state": (None, [vp, dp, dp]),
            "mo_get_collision_forces": (None, [vp, dp]),
            "mo_set_action": (None, [vp, ip, dp, i]),
            "mo_resolve_collisions": (i, [vp, ip, i]),
            "mo_update_actuators": (None, [vp]),
            "mo_step_rk4": (None, [vp, d]),
            "mo_resolve_edge_constraints": (None, [vp]),
            "mo_update_true_vel": (None, [vp, d]),
            "mo_is_any_robot_self_colliding": (i, [vp]),
            "mo_substep": (i, [vp]),
            "mo_step": (i, [vp, i]),
            "mo_contact_count": (C.c_long, [vp]),
        }
        for name, (res, args) in sig.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


def _dp(a):
    return None if a is None else a.ctypes.data_as(C.POINTER(C.c_double))


def _ip(a):
    return None if a is None else a.ctypes.data_as(C.POINTER(C.c_int))


def _bp(a):
    return a.ctypes.data_as(C.POINTER(C.c_ubyte))


class OracleSim:
    """Restated ``Sim`` for one world.  Point arrays are ``(P, 2)`` float64."""

    def __init__(self, world=None, topology: Optional[Topology] = None):
        self.L = lib()
        self.topo = topology if topology is not None else build_topology(world)
        t = self.topo
        rows, info = [], []
        first = 0
        for oi, o in enumerate(t.objects):
            surf = set(int(s) for s in o.surface_boxels)
            for j in range(o.boxels.shape[0]):
                rows.append(list(o.boxels[j, :10]) + [oi, 1 if j in surf else 0])
            info.append([int(o.is_robot), first, o.boxels.shape[0], o.grid_w])
            first += o.boxels.shape[0]
        self.boxels = np.ascontiguousarray(np.asarray(rows, dtype=np.int32).reshape(-1, 12))
        self.obj_info = np.ascontiguousarray(np.asarray(info, dtype=np.int32).reshape(-1, 4))
        self.h = self.L.mo_create(t.P, t.E, len(t.objects), self.boxels.shape[0],
                                  _dp(np.ascontiguousarray(t.pos0)), _dp(np.ascontiguousarray(t.mass)),
                                  _bp(np.ascontiguousarray(t.fixed)), _ip(np.ascontiguousarray(t.ea)),
                                  _ip(np.ascontiguousarray(t.eb)), _dp(np.ascontiguousarray(t.ek)),
                                  _dp(np.ascontiguousarray(t.el0)), _bp(np.ascontiguousarray(t.on_surface)),
                                  _ip(self.boxels), _ip(self.obj_info))
        self.sim_time = 0
        self.history: Dict[int, Tuple[np.ndarray, np.ndarray]] = {}
        self._save(0)
        # actuator grid index -> global boxel row, per robot
        self._act: Dict[str, Dict[int, int]] = {}
        for oi, o in enumerate(t.objects):
            if not o.is_robot:
                continue
            m = {}
            for j in range(o.boxels.shape[0]):
                if o.boxels[j, 0] in (3, 4):
                    m[int(o.boxels[j, 1])] = int(self.obj_info[oi, 1]) + j
            self._act[o.name] = m

    def __del__(self):
        try:
            if getattr(self, "h", None):
                self.L.mo_destroy(self.h)
                self.h = None
        except Exception:
            pass

    # -- state
    def get_state(self):
        out = [np.zeros((self.topo.P, 2)) for _ in range(4)]
        self.L.mo_get_state(self.h, *[_dp(a) for a in out])
        return tuple(out)

    def set_state(self, pos=None, vel=None, pos_last=None, vel_true=None):
        arrs = [None if a is None else np.ascontiguousarray(a, dtype=np.float64) for a in (pos, vel, pos_last, vel_true)]
        self.L.mo_set_state(self.h, *[_dp(a) for a in arrs])

    def edge_state(self):
        leq, goal = np.zeros(self.topo.E), np.zeros(self.topo.E)
        self.L.mo_get_edge_state(self.h, _dp(leq), _dp(goal))
        return leq, goal

    def set_edge_state(self, leq=None, goal=None):
        arrs = [None if a is None else np.ascontiguousarray(a, dtype=np.float64) for a in (leq, goal)]
        self.L.mo_set_edge_state(self.h, *[_dp(a) for a in arrs])

    def collision_forces(self):
        out = np.zeros((self.topo.P, 2))
        self.L.mo_get_collision_forces(self.h, _dp(out))
        return out

    # -- Sim API
    def actuator_indices(self, robot: str) -> np.ndarray:
        return self.topo.objects[self.topo.object_index(robot)].actuator_grid_indices.copy()

    def set_action(self, robot: str, action) -> None:
        a = np.asarray(action, dtype=np.float64).reshape(-1, 2)
        ids = np.ascontiguousarray([self._act[robot][int(g)] for g in a[:, 0]], dtype=np.int32)
        vals = np.ascontiguousarray(a[:, 1])
        self.L.mo_set_action(self.h, _ip(ids), _dp(vals), len(ids))

    def set_action_values(self, robot: str, values) -> None:
        idx = self.actuator_indices(robot)
        self.set_action(robot, np.stack((idx.astype(np.float64), np.asarray(values, dtype=np.float64).ravel()), axis=1))

    def _save(self, t: int) -> None:
        pos, vel, _, _ = self.get_state()
        self.history[t] = (pos, vel)

    def substep(self) -> bool:
        return bool(self.L.mo_substep(self.h))

    def step(self, n_substeps: int = 30) -> bool:
        """``Sim::step`` (Sim.cpp:105-115): unstable aborts without advancing time or saving."""
        if self.L.mo_step(self.h, n_substeps):
            return True
        self.sim_time += 1
        self._save(self.sim_time)
        return False

    def get_time(self) -> int:
        return self.sim_time

    def revert(self, t: int) -> None:
        """``Environment::revert_to_snapshot`` (Environment.cpp:247-263): goals are NOT reset."""
        if t not in self.history:
            return
        pos, vel = self.history[t]
        self.set_state(pos, vel, pos, vel)
        for k in [k for k in self.history if k > t]:
            del self.history[k]
        self.sim_time = t

    def pos_at_time(self, t: int, vel: bool = False):
        if t not in self.history:
            return None
        return self.history[t][1 if vel else 0].copy()

    # -- phases
    def phase_collisions(self, want_candidates: bool = False):
        if not want_candidates:
            return self.L.mo_resolve_collisions(self.h, None, 0)
        n = self.L.mo_resolve_collisions(self.h, None, 0)
        out = np.zeros((max(n, 1), 2), dtype=np.int32)
        self.L.mo_resolve_collisions(self.h, _ip(out), n)
        return out[:n]

    def phase_actuators(self): self.L.mo_update_actuators(self.h)
    def phase_rk4(self, dt: float = 0.0001): self.L.mo_step_rk4(self.h, dt)
    def phase_constraints(self): self.L.mo_resolve_edge_constraints(self.h)
    def phase_true_vel(self, dt: float = 0.0001): self.L.mo_update_true_vel(self.h, dt)
    def phase_self_collide(self) -> bool: return bool(self.L.mo_is_any_robot_self_colliding(self.h))
    def contact_count(self) -> int: return int(self.L.mo_contact_count(self.h))

    def candidates_as_object_boxel(self, cand: np.ndarray) -> np.ndarray:
        """Map (global boxel a, global boxel b) rows to the reference's (obj_a, boxel_a, obj_b, boxel_b)."""
        out = np.zeros((len(cand), 4), dtype=np.int32)
        for r, (ga, gb) in enumerate(cand):
            oa, ob = self.boxels[ga, 10], self.boxels[gb, 10]
            out[r] = (oa, ga - self.obj_info[oa, 1], ob, gb - self.obj_info[ob, 1])
        return out
