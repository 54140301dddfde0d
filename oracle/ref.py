"""TEST INFRASTRUCTURE ONLY: ctypes view of ``oracle/_ref/libmevo_ref.so`` -- the unmodified
reference engine compiled from ``/root/reference`` by ``oracle/Makefile`` (see ``ref_driver.cpp``).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / ``--impl reference``
legs may import this module.  The product package never does.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional, Tuple

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libmevo_ref.so")                # g++ -O2: the reference's Release default; parity uses this one
LIB_PATH_O3 = os.path.join(_HERE, "_ref", "libmevo_ref_o3.so")          # g++ -O3 -march=x86-64-v3: CPU-baseline timing only

_lib = None


def available(opt: str = "O2") -> bool:
    return os.path.exists(LIB_PATH_O3 if opt == "O3" else LIB_PATH)


def use_build(opt: str) -> None:
    """Select the build this PROCESS loads (call before the first RefSim): "O2" (default) or "O3"."""
    global _lib, LIB_PATH
    if _lib is not None:
        raise RuntimeError("the reference library is already loaded")
    if opt == "O3":
        LIB_PATH = LIB_PATH_O3


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise FileNotFoundError(f"{LIB_PATH} missing: run `make -C oracle ref` where /root/reference exists")
        L = C.CDLL(LIB_PATH)
        vp, i, d, cp = C.c_void_p, C.c_int, C.c_double, C.c_char_p
        dp = C.POINTER(C.c_double)
        ip = C.POINTER(C.c_int)
        bp = C.POINTER(C.c_ubyte)
        sig = {
            "mref_create": (vp, [i, i]),
            "mref_destroy": (None, [vp]),
            "mref_add_object": (i, [vp, dp, i, i, dp, i, cp, d, d, i, i]),
            "mref_add_object_from_file": (i, [vp, cp, cp, d, d, i, i]),
            "mref_num_points": (i, [vp]),
            "mref_num_edges": (i, [vp]),
            "mref_num_objects": (i, [vp]),
            "mref_get_state": (None, [vp, dp, dp, dp, dp]),
            "mref_set_state": (None, [vp, dp, dp, dp, dp]),
            "mref_get_point_consts": (None, [vp, dp, bp]),
            "mref_get_edges": (None, [vp, ip, ip, dp, dp, dp, dp, bp]),
            "mref_set_edge_state": (None, [vp, dp, dp]),
            "mref_object_info": (None, [vp, i, ip]),
            "mref_object_boxels": (None, [vp, i, ip]),
            "mref_object_surface_boxels": (None, [vp, i, ip]),
            "mref_object_name": (cp, [vp, i]),
            "mref_num_actuators": (i, [vp, cp]),
            "mref_actuator_indices": (None, [vp, cp, ip]),
            "mref_set_action": (None, [vp, cp, dp, i]),
            "mref_substep": (i, [vp]),
            "mref_step": (i, [vp]),
            "mref_steps": (i, [vp, i]),
            "mref_get_time": (i, [vp]),
            "mref_revert": (None, [vp, C.c_long]),
            "mref_translate_object": (None, [vp, d, d, cp]),
            "mref_object_orientation_at_time": (d, [vp, C.c_long, cp]),
            "mref_pos_at_time": (i, [vp, C.c_long, dp, i, i]),
            "mref_phase_collisions": (None, [vp]),
            "mref_phase_actuators": (None, [vp]),
            "mref_phase_rk4": (None, [vp]),
            "mref_phase_constraints": (None, [vp]),
            "mref_phase_true_vel": (None, [vp]),
            "mref_phase_self_collide": (i, [vp]),
            "mref_get_collision_forces": (None, [vp, dp]),
            "mref_candidates": (i, [vp, ip, i]),
        }
        for name, (res, args) in sig.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def _dp(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(C.POINTER(C.c_double))


def _ip(a: np.ndarray):
    return a.ctypes.data_as(C.POINTER(C.c_int))


def _bp(a: np.ndarray):
    return a.ctypes.data_as(C.POINTER(C.c_ubyte))


class RefSim:
    """One reference ``Sim``.  Point arrays are ``(P, 2)`` float64 (x, y per row) -- the memory
    order of the reference's column-major ``2 x P`` Eigen matrices."""

    def __init__(self, world=None, world_w: int = 1, world_h: int = 1, quiet: bool = True):
        self.L = lib()
        if world is not None:
            world_w, world_h = world.width, world.height
        self.h = self.L.mref_create(int(world_w), int(world_h))
        self.quiet = quiet
        self.object_names: List[str] = []
        if world is not None:
            for o in world.load_order():
                ok = self.add_object(o.grid, o.connections, o.name, o.x, o.y, o.is_robot)
                if not ok:
                    raise ValueError(f"reference refused object {o.name}")

    def __del__(self):
        try:
            if getattr(self, "h", None):
                self.L.mref_destroy(self.h)
                self.h = None
        except Exception:
            pass

    # -- construction
    def add_object(self, grid, connections, name: str, x: float, y: float, is_robot: bool) -> bool:
        g = np.ascontiguousarray(grid, dtype=np.float64)
        cn = np.ascontiguousarray(connections, dtype=np.float64).reshape(2, -1)
        ok = self.L.mref_add_object(self.h, _dp(g), g.shape[1], g.shape[0], _dp(cn), cn.shape[1],
                                    name.encode(), float(x), float(y), int(bool(is_robot)), int(self.quiet))
        if ok:
            self.object_names.append(name)
        return bool(ok)

    def add_object_from_file(self, path: str, name: str, x: float, y: float, is_robot: bool) -> bool:
        """``Sim::read_robot_from_file`` / ``read_object_from_file``: the reference's own text loader (ObjectCreator.cpp:118-171)"""
        ok = self.L.mref_add_object_from_file(self.h, str(path).encode(), name.encode(), float(x), float(y), int(bool(is_robot)), int(self.quiet))
        if ok:
            self.object_names.append(name)
        return bool(ok)

    # -- sizes
    @property
    def num_points(self) -> int:
        return self.L.mref_num_points(self.h)

    @property
    def num_edges(self) -> int:
        return self.L.mref_num_edges(self.h)

    @property
    def num_objects(self) -> int:
        return self.L.mref_num_objects(self.h)

    # -- state
    def get_state(self) -> Tuple[np.ndarray, np.ndarray, np.ndarray, np.ndarray]:
        n = self.num_points
        out = [np.zeros((n, 2)) for _ in range(4)]
        self.L.mref_get_state(self.h, *[_dp(a) for a in out])
        return tuple(out)

    def set_state(self, pos=None, vel=None, pos_last=None, vel_true=None) -> None:
        arrs = [None if a is None else np.ascontiguousarray(a, dtype=np.float64) for a in (pos, vel, pos_last, vel_true)]
        self.L.mref_set_state(self.h, *[_dp(a) for a in arrs])

    def point_consts(self) -> Tuple[np.ndarray, np.ndarray]:
        n = self.num_points
        mass = np.zeros(n)
        fixed = np.zeros(n, dtype=np.uint8)
        self.L.mref_get_point_consts(self.h, _dp(mass), _bp(fixed))
        return mass, fixed

    def edges(self) -> dict:
        n = self.num_edges
        a = np.zeros(n, dtype=np.int32); b = np.zeros(n, dtype=np.int32)
        k = np.zeros(n); l0 = np.zeros(n); leq = np.zeros(n); goal = np.zeros(n)
        surf = np.zeros(n, dtype=np.uint8)
        self.L.mref_get_edges(self.h, _ip(a), _ip(b), _dp(k), _dp(l0), _dp(leq), _dp(goal), _bp(surf))
        return dict(a=a, b=b, k=k, l0=l0, leq=leq, goal=goal, on_surface=surf)

    def set_edge_state(self, leq=None, goal=None) -> None:
        arrs = [None if a is None else np.ascontiguousarray(a, dtype=np.float64) for a in (leq, goal)]
        self.L.mref_set_edge_state(self.h, *[_dp(a) for a in arrs])

    def object_info(self, obj: int) -> dict:
        info = np.zeros(7, dtype=np.int32)
        self.L.mref_object_info(self.h, obj, _ip(info))
        keys = ("is_robot", "min_point", "max_point", "n_boxels", "grid_w", "grid_h", "n_surface")
        return dict(zip(keys, (int(v) for v in info)))

    def object_boxels(self, obj: int) -> np.ndarray:
        """(n_boxels, 18) ints: cell_type, grid_index, TL,TR,BR,BL, top,right,bot,left, 8 neighbour flags."""
        n = self.object_info(obj)["n_boxels"]
        out = np.zeros((n, 18), dtype=np.int32)
        self.L.mref_object_boxels(self.h, obj, _ip(out))
        return out

    def object_surface_boxels(self, obj: int) -> np.ndarray:
        n = self.object_info(obj)["n_surface"]
        out = np.zeros(n, dtype=np.int32)
        self.L.mref_object_surface_boxels(self.h, obj, _ip(out))
        return out

    def actuator_indices(self, robot: str) -> np.ndarray:
        n = self.L.mref_num_actuators(self.h, robot.encode())
        if n < 0:
            raise ValueError(f"{robot} is not a robot")
        out = np.zeros(n, dtype=np.int32)
        self.L.mref_actuator_indices(self.h, robot.encode(), _ip(out))
        return out

    # -- stepping
    def set_action(self, robot: str, action: np.ndarray) -> None:
        """``action`` is (n, 2): (grid_index, value) rows (``evogym/sim.py:141``)."""
        a = np.ascontiguousarray(action, dtype=np.float64).reshape(-1, 2)
        self.L.mref_set_action(self.h, robot.encode(), _dp(a), a.shape[0])

    def set_action_values(self, robot: str, values: np.ndarray) -> None:
        idx = self.actuator_indices(robot)
        self.set_action(robot, np.stack((idx.astype(np.float64), np.asarray(values, dtype=np.float64).ravel()), axis=1))

    def substep(self) -> bool:
        return bool(self.L.mref_substep(self.h))

    def step(self) -> bool:
        return bool(self.L.mref_step(self.h))

    def steps(self, n: int) -> int:
        return int(self.L.mref_steps(self.h, int(n)))

    def get_time(self) -> int:
        return self.L.mref_get_time(self.h)

    def revert(self, t: int) -> None:
        self.L.mref_revert(self.h, int(t))

    def translate_object(self, x: float, y: float, name: str) -> None:
        self.L.mref_translate_object(self.h, float(x), float(y), name.encode())

    def object_orientation_at_time(self, t: int, name: str) -> float:
        return float(self.L.mref_object_orientation_at_time(self.h, int(t), name.encode()))

    def pos_at_time(self, t: int, vel: bool = False) -> Optional[np.ndarray]:
        n = self.num_points
        out = np.zeros((n, 2))
        got = self.L.mref_pos_at_time(self.h, int(t), _dp(out), n, int(vel))
        return out if got == n else None

    # -- phases (Environment::step split up)
    def phase_collisions(self): self.L.mref_phase_collisions(self.h)
    def phase_actuators(self): self.L.mref_phase_actuators(self.h)
    def phase_rk4(self): self.L.mref_phase_rk4(self.h)
    def phase_constraints(self): self.L.mref_phase_constraints(self.h)
    def phase_true_vel(self): self.L.mref_phase_true_vel(self.h)
    def phase_self_collide(self) -> bool: return bool(self.L.mref_phase_self_collide(self.h))

    def collision_forces(self) -> np.ndarray:
        out = np.zeros((self.num_points, 2))
        self.L.mref_get_collision_forces(self.h, _dp(out))
        return out

    def candidates(self) -> np.ndarray:
        """(n, 4) int array (object_a, boxel_a, object_b, boxel_b) in reference descent order."""
        n = self.L.mref_candidates(self.h, None, 0)
        out = np.zeros((max(n, 1), 4), dtype=np.int32)
        self.L.mref_candidates(self.h, _ip(out), n)
        return out[:n]
