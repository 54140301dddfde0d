"""ctypes binding of ``libmevo_b200.so`` -- exactly the functions ``include/mevo_b200.h`` declares.

The product path: if the shared library is missing this module raises at import time of the first
handle (``load()``); there is no Python or CPU fallback for the physics.
"""
from __future__ import annotations

import ctypes as C
import os
import re
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MEVO_LIB") or os.path.join(PKG_DIR, "libmevo_b200.so")     # MEVO_LIB: development builds (e.g. -DMEVO_PROF)
HEADER_PATH = os.path.join(PKG_DIR, "..", "include", "mevo_b200.h")

_vp, _i, _d, _cp = C.c_void_p, C.c_int, C.c_double, C.c_char_p
_dp, _ip, _bp = C.POINTER(C.c_double), C.POINTER(C.c_int), C.POINTER(C.c_ubyte)

SIGNATURES: Dict[str, tuple] = {
    "mevo_abi_version": (_i, []),
    "mevo_last_error": (_cp, []),
    "mevo_device_count": (_i, []),
    "mevo_batch_create": (_vp, [_i]),
    "mevo_batch_destroy": (None, [_vp]),
    "mevo_batch_begin_env": (_i, [_vp, _i, _i]),
    "mevo_batch_add_object": (_i, [_vp, _i, _dp, _i, _i, _dp, _i, _cp, _d, _d, _i]),
    "mevo_batch_add_robot_envs": (_i, [_vp, _i, _i, _i, _i, _dp, _ip, _dp, _ip, _cp, _dp, _dp, _ip, _dp, _ip, _cp, _d, _d]),
    "mevo_batch_finalize": (_i, [_vp]),
    "mevo_batch_finalize_epoch": (_i, [_vp]),
    "mevo_batch_num_envs": (_i, [_vp]),
    "mevo_env_num_points": (_i, [_vp, _i]),
    "mevo_env_num_edges": (_i, [_vp, _i]),
    "mevo_env_num_objects": (_i, [_vp, _i]),
    "mevo_env_object_index": (_i, [_vp, _i, _cp]),
    "mevo_env_object_info": (_i, [_vp, _i, _i, _ip]),
    "mevo_env_object_boxels": (_i, [_vp, _i, _i, _ip]),
    "mevo_env_object_surface_boxels": (_i, [_vp, _i, _i, _ip]),
    "mevo_env_actuator_indices": (_i, [_vp, _i, _cp, _ip, _i]),
    "mevo_env_get_topology": (_i, [_vp, _i, _dp, _dp, _bp, _ip, _ip, _dp, _dp, _bp]),
    "mevo_world_create": (_vp, [_i, _i]),
    "mevo_world_destroy": (None, [_vp]),
    "mevo_world_add_object": (_i, [_vp, _dp, _i, _i, _dp, _i, _cp, _d, _d, _i]),
    "mevo_world_num_points": (_i, [_vp]),
    "mevo_world_num_edges": (_i, [_vp]),
    "mevo_world_num_objects": (_i, [_vp]),
    "mevo_world_object_info": (_i, [_vp, _i, _ip]),
    "mevo_world_object_boxels": (_i, [_vp, _i, _ip]),
    "mevo_world_object_surface_boxels": (_i, [_vp, _i, _ip]),
    "mevo_world_actuator_indices": (_i, [_vp, _cp, _ip, _i]),
    "mevo_world_get_topology": (_i, [_vp, _dp, _dp, _bp, _ip, _ip, _dp, _dp, _bp]),
    "mevo_env_set_action": (_i, [_vp, _i, _cp, _dp, _i]),
    "mevo_batch_action_stride": (_i, [_vp]),
    "mevo_batch_set_actions": (_i, [_vp, _dp, _i]),
    "mevo_batch_set_actions_device": (_i, [_vp, _vp, _i]),
    "mevo_batch_bind_actions_device": (_i, [_vp, _vp, _i]),
    "mevo_batch_set_auto_reset": (_i, [_vp, _i]),
    "mevo_batch_unstable_events": (C.c_longlong, [_vp]),
    "mevo_batch_object_com_device": (_i, [_vp, _i, _vp]),
    "mevo_batch_fetch_step_results": (_i, [_vp, _vp, _i, _vp]),
    "mevo_batch_step": (_i, [_vp, _i, _i, _bp]),
    "mevo_batch_step_substeps": (_i, [_vp, _i, _i, _i, _bp]),
    "mevo_batch_step_async": (_i, [_vp, _i, _i]),
    "mevo_batch_step_static_scratch": (_i, [_vp, _i, _vp, _i, _vp]),
    "mevo_batch_sync": (_i, [_vp]),
    "mevo_batch_last_launches": (_i, [_vp]),
    "mevo_env_get_time": (C.c_long, [_vp, _i]),
    "mevo_env_set_time": (_i, [_vp, _i, C.c_long]),
    "mevo_env_get_state": (_i, [_vp, _i, _dp, _dp, _dp, _dp]),
    "mevo_env_set_state": (_i, [_vp, _i, _dp, _dp, _dp, _dp]),
    "mevo_env_get_goal": (_i, [_vp, _i, _dp]),
    "mevo_env_set_goal": (_i, [_vp, _i, _dp]),
    "mevo_env_translate_object": (_i, [_vp, _i, _cp, _d, _d]),
    "mevo_batch_save_reset_state": (_i, [_vp]),
    "mevo_batch_reset": (_i, [_vp, _ip, _i]),
    "mevo_batch_save_snapshot": (_i, [_vp, _i]),
    "mevo_batch_revert": (_i, [_vp, _ip, _i, _i]),
    "mevo_env_read_snapshot": (C.c_long, [_vp, _i, _i, _dp, _dp]),
    "mevo_batch_object_com": (_i, [_vp, _i, _dp]),
    "mevo_batch_object_orientation": (_i, [_vp, _i, _dp]),
    "mevo_batch_object_orientation_device": (_i, [_vp, _i, _vp]),
    "mevo_batch_sight_obs": (_i, [_vp, _i, _cp, _i, _d, _i, _dp, _i]),
    "mevo_batch_sight_obs_device": (_i, [_vp, _i, _cp, _i, _d, _i, _vp, _i, _i]),
    "mevo_batch_object_obs_width": (_i, [_vp, _i]),
    "mevo_batch_object_obs": (_i, [_vp, _i, _dp, _i]),
    "mevo_batch_object_obs_device": (_i, [_vp, _i, _vp, _i]),
    "mevo_batch_task_configure": (_i, [_vp, _i, _cp, _cp, _i]),
    "mevo_batch_task_dims": (_i, [_vp, _ip]),
    "mevo_batch_task_env_tables": (_i, [_vp, _i, _ip, _ip, _ip, _i]),
    "mevo_modular_corner_index": (_i, [_ip, _i, _i, _ip]),
    "mevo_batch_task_apply_policy": (_i, [_vp, _dp, _i, _i]),
    "mevo_batch_task_apply_policy_device": (_i, [_vp, _vp, _i, _i, _i]),
    "mevo_batch_task_outputs": (_i, [_vp, _dp, _bp, _dp, _i, _dp, _i, _dp, _i, _i]),
    "mevo_batch_task_outputs_device": (_i, [_vp, _vp, _vp, _vp, _i, _vp, _i, _vp, _i, _i, _i]),
    "mevo_batch_task_reset": (_i, [_vp, _dp, _i, _dp, _i, _dp, _i]),
    "mevo_batch_task_bind_flag_planes": (_i, [_vp, _vp]),
    "mevo_batch_task_launch_sources": (_i, [_vp, _ip]),
    "mevo_batch_task_set_goals": (_i, [_vp, _i, _i, _ip, _ip, _i]),
    "mevo_batch_task_draw_goals": (_i, [_vp, C.c_ulonglong]),
    "mevo_batch_task_get_goals": (_i, [_vp, _i, _ip, _ip, _ip, _i]),
    "mevo_batch_task_reset_device": (_i, [_vp, _vp, _i, _vp, _i, _vp, _i, _i]),
    "mevo_env_render": (_i, [_vp, _i, _d, _d, _d, _d, _i, _i, _i, _bp]),
    "mevo_env_render_device": (_i, [_vp, _i, _d, _d, _d, _d, _i, _i, _i, _vp]),
    "mevo_batch_capture_begin": (_i, [_vp]),
    "mevo_batch_capture_end": (_i, [_vp]),
    "mevo_batch_graph_launch": (_i, [_vp, _i]),
    "mevo_batch_graph_destroy": (_i, [_vp, _i]),
    "mevo_batch_actions_device_ptr": (_vp, [_vp]),
    "mevo_batch_unstable_device_ptr": (_vp, [_vp]),
    "mevo_batch_cuda_stream": (_vp, [_vp]),
    "mevo_batch_device_bytes": (C.c_longlong, [_vp]),
    "mevo_batch_enable_candidate_dump": (_i, [_vp, _i]),
    "mevo_env_read_candidates": (_i, [_vp, _i, _ip, _i]),
    "mevo_batch_enable_force_dump": (_i, [_vp]),
    "mevo_env_read_collision_forces": (_i, [_vp, _i, _dp]),
    "mevo_env_error_flags": (_i, [_vp, _i]),
}

MEVO_TASK_WALKER, MEVO_TASK_PUSHER, MEVO_TASK_THROWER, MEVO_TASK_CLIMBER, MEVO_TASK_OBSTACLE_TRAVERSER, MEVO_TASK_BRIDGE_WALKER = 0, 1, 2, 3, 4, 5
MEVO_TASK_OBSTACLE_TRAVERSER_V0, MEVO_TASK_UP_STEPPER, MEVO_TASK_DOWN_STEPPER, MEVO_TASK_JUMPER = 6, 7, 8, 9
MEVO_TASK_HURDLER, MEVO_TASK_CLIMBER_V1, MEVO_TASK_CAVE_CRAWLER, MEVO_TASK_CLIMBER_V2 = 10, 11, 12, 13
MEVO_TASK_CARRIER, MEVO_TASK_PUSHER_V1, MEVO_TASK_BALANCER, MEVO_TASK_BALANCER_V1 = 14, 15, 16, 17
MEVO_TASK_GAP_JUMPER, MEVO_TASK_PLATFORM_JUMPER, MEVO_TASK_CARRIER_V1, MEVO_TASK_LIFTER, MEVO_TASK_FLIPPER = 18, 19, 20, 21, 22
MEVO_TASK_BEAM_TOPPLER, MEVO_TASK_BEAM_SLIDER, MEVO_TASK_WINGSPAN, MEVO_TASK_HEIGHT = 23, 24, 25, 26
MEVO_TASK_AREA_MAX, MEVO_TASK_AREA_MIN, MEVO_TASK_BIDIRECTIONAL_WALKER, MEVO_TASK_TRAVERSER, MEVO_TASK_CATCHER, MEVO_TASK_UP_STEPPER_FLAT = 27, 28, 29, 30, 31, 32
MEVO_DONE, MEVO_DONE_TRUNCATED, MEVO_DONE_UNSTABLE = 1, 2, 4      # done byte bits (include/mevo_b200.h)

_lib = None


class MevoError(RuntimeError):
    pass


def header_symbols(path: str = HEADER_PATH) -> List[str]:
    """Function names declared in ``include/mevo_b200.h`` (used by the export test)."""
    with open(path) as f:
        text = re.sub(r"/\*.*?\*/", "", f.read(), flags=re.S)
    return sorted(set(re.findall(r"\b(mevo_[a-z0-9_]+)\s*\(", text)))


def load():
    """Load the C-ABI library (fails loudly if it was not built)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise MevoError(f"{LIB_PATH} is missing: build it with `python -m modularevogym_b200._build` "
                            "(needs nvcc). The engine has no CPU fallback.")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        if L.mevo_abi_version() != 1:
            raise MevoError("libmevo_b200.so ABI version mismatch")
        _lib = L
    return _lib


def _dptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(_dp)


def _iptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(_ip)


def _bptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(_bp)


def _check(rc: int) -> int:
    if rc < 0:
        raise MevoError(load().mevo_last_error().decode())
    return rc


OBJ_INFO_KEYS = ("is_robot", "min_point", "max_point", "n_boxels", "grid_w", "grid_h", "n_surface", "is_static")


class HostWorld:
    """Host-only world builder handle (no GPU): ``mevo_world_*``."""

    def __init__(self, world=None, world_w: int = 1, world_h: int = 1):
        self.L = load()
        if world is not None:
            world_w, world_h = world.width, world.height
        self.h = self.L.mevo_world_create(int(world_w), int(world_h))
        if not self.h:
            raise MevoError(self.L.mevo_last_error().decode())
        if world is not None:
            for o in world.load_order():
                if not self.add_object(o.grid, o.connections, o.name, o.x, o.y, o.is_robot):
                    raise ValueError(f"object {o.name} refused")

    def __del__(self):
        try:
            if getattr(self, "h", None):
                self.L.mevo_world_destroy(self.h)
                self.h = None
        except Exception:
            pass

    def add_object(self, grid, connections, name: str, x: float, y: float, is_robot: bool) -> bool:
        g = np.ascontiguousarray(grid, dtype=np.float64)
        cn = np.ascontiguousarray(connections, dtype=np.float64).reshape(2, -1)
        return bool(_check(self.L.mevo_world_add_object(self.h, _dptr(g), g.shape[1], g.shape[0], _dptr(cn), cn.shape[1],
                                                        name.encode(), float(x), float(y), int(bool(is_robot)))))

    @property
    def num_points(self): return self.L.mevo_world_num_points(self.h)
    @property
    def num_edges(self): return self.L.mevo_world_num_edges(self.h)
    @property
    def num_objects(self): return self.L.mevo_world_num_objects(self.h)

    def object_info(self, obj: int) -> dict:
        info = np.zeros(8, dtype=np.int32)
        _check(self.L.mevo_world_object_info(self.h, obj, _iptr(info)))
        return dict(zip(OBJ_INFO_KEYS, (int(v) for v in info)))

    def object_boxels(self, obj: int) -> np.ndarray:
        out = np.zeros((self.object_info(obj)["n_boxels"], 18), dtype=np.int32)
        _check(self.L.mevo_world_object_boxels(self.h, obj, _iptr(out)))
        return out

    def object_surface_boxels(self, obj: int) -> np.ndarray:
        out = np.zeros(self.object_info(obj)["n_surface"], dtype=np.int32)
        _check(self.L.mevo_world_object_surface_boxels(self.h, obj, _iptr(out)))
        return out

    def actuator_indices(self, robot: str) -> np.ndarray:
        out = np.zeros(4096, dtype=np.int32)
        n = _check(self.L.mevo_world_actuator_indices(self.h, robot.encode(), _iptr(out), out.size))
        return out[:n].copy()

    def topology(self) -> dict:
        P, E = self.num_points, self.num_edges
        t = dict(pos0=np.zeros((P, 2)), mass=np.zeros(P), fixed=np.zeros(P, dtype=np.uint8),
                 a=np.zeros(E, dtype=np.int32), b=np.zeros(E, dtype=np.int32), k=np.zeros(E), l0=np.zeros(E),
                 on_surface=np.zeros(E, dtype=np.uint8))
        _check(self.L.mevo_world_get_topology(self.h, _dptr(t["pos0"]), _dptr(t["mass"]), _bptr(t["fixed"]), _iptr(t["a"]),
                                              _iptr(t["b"]), _dptr(t["k"]), _dptr(t["l0"]), _bptr(t["on_surface"])))
        return t


class Batch:
    """N environments on one GPU: ``mevo_batch_*`` / ``mevo_env_*``."""

    def __init__(self, device: int = 0):
        self.L = load()
        self.h = self.L.mevo_batch_create(int(device))
        if not self.h:
            raise MevoError(self.L.mevo_last_error().decode())
        self.device = int(device)

    def close(self):
        if getattr(self, "h", None):
            self.L.mevo_batch_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- construction
    def begin_env(self, world_w: int, world_h: int) -> int:
        return _check(self.L.mevo_batch_begin_env(self.h, int(world_w), int(world_h)))

    def add_object(self, env: int, grid, connections, name: str, x: float, y: float, is_robot: bool) -> bool:
        g = np.ascontiguousarray(grid, dtype=np.float64)
        cn = np.ascontiguousarray(connections, dtype=np.float64).reshape(2, -1)
        return bool(_check(self.L.mevo_batch_add_object(self.h, env, _dptr(g), g.shape[1], g.shape[0], _dptr(cn), cn.shape[1],
                                                        name.encode(), float(x), float(y), int(bool(is_robot)))))

    def add_world(self, world) -> int:
        """Add one environment from a ``WorldSpec`` (objects in the reference's load order)."""
        env = self.begin_env(world.width, world.height)
        for o in world.load_order():
            if not self.add_object(env, o.grid, o.connections, o.name, o.x, o.y, o.is_robot):
                raise ValueError(f"object {o.name} refused")
        return env

    def add_robot_worlds(self, terrain, bodies, connections, robot_name: str, rx: int, ry: int) -> int:
        """One call for ``[add_world(terrain + robot_i at (rx, ry)) for i ...]`` (``mevo_batch_add_robot_envs``): ``terrain`` is a
        ``WorldSpec`` without robots, ``bodies`` / ``connections`` the robots' grids (row 0 = top) and (2, k) connection lists."""
        objs = terrain.load_order()
        if any(o.is_robot for o in objs):
            raise ValueError("add_robot_worlds: the terrain must not contain robots")
        f64 = lambda arrs: (np.concatenate([np.asarray(a, dtype=np.float64).ravel() for a in arrs]) if arrs else np.zeros(0))
        i32 = lambda rows: np.ascontiguousarray(np.asarray(rows, dtype=np.int32).reshape(-1))
        t_grids, t_conn = f64([o.grid for o in objs]), f64([np.asarray(o.connections).reshape(2, -1) for o in objs])
        t_wh, t_k = i32([o.size for o in objs]), i32([np.asarray(o.connections).reshape(2, -1).shape[1] for o in objs])
        t_xy = np.ascontiguousarray(np.asarray([(o.x, o.y) for o in objs], dtype=np.float64).reshape(-1))
        bodies = [np.asarray(b) for b in bodies]
        conns = [np.asarray(c).reshape(2, -1) for c in connections]
        r_grids, r_conn = f64(bodies), f64(conns)
        r_wh, r_k = i32([(b.shape[1], b.shape[0]) for b in bodies]), i32([c.shape[1] for c in conns])
        names = "\n".join(o.name for o in objs).encode()
        return _check(self.L.mevo_batch_add_robot_envs(self.h, len(bodies), int(terrain.width), int(terrain.height), len(objs), _dptr(t_grids),
                                                       _iptr(t_wh), _dptr(t_conn), _iptr(t_k), names, _dptr(t_xy), _dptr(r_grids), _iptr(r_wh),
                                                       _dptr(r_conn), _iptr(r_k), robot_name.encode(), float(rx), float(ry)))

    def finalize(self) -> None:
        _check(self.L.mevo_batch_finalize(self.h))

    @property
    def finalize_epoch(self) -> int:
        return self.L.mevo_batch_finalize_epoch(self.h)

    # -- queries
    @property
    def num_envs(self) -> int:
        return self.L.mevo_batch_num_envs(self.h)

    def num_points(self, env: int) -> int: return _check(self.L.mevo_env_num_points(self.h, env))
    def num_edges(self, env: int) -> int: return _check(self.L.mevo_env_num_edges(self.h, env))
    def num_objects(self, env: int) -> int: return _check(self.L.mevo_env_num_objects(self.h, env))
    def object_index(self, env: int, name: str) -> int: return _check(self.L.mevo_env_object_index(self.h, env, name.encode()))

    def object_info(self, env: int, obj: int) -> dict:
        info = np.zeros(8, dtype=np.int32)
        _check(self.L.mevo_env_object_info(self.h, env, obj, _iptr(info)))
        return dict(zip(OBJ_INFO_KEYS, (int(v) for v in info)))

    def object_boxels(self, env: int, obj: int) -> np.ndarray:
        out = np.zeros((self.object_info(env, obj)["n_boxels"], 18), dtype=np.int32)
        _check(self.L.mevo_env_object_boxels(self.h, env, obj, _iptr(out)))
        return out

    def object_surface_boxels(self, env: int, obj: int) -> np.ndarray:
        out = np.zeros(self.object_info(env, obj)["n_surface"], dtype=np.int32)
        _check(self.L.mevo_env_object_surface_boxels(self.h, env, obj, _iptr(out)))
        return out

    def actuator_indices(self, env: int, robot: str) -> np.ndarray:
        out = np.zeros(4096, dtype=np.int32)
        n = _check(self.L.mevo_env_actuator_indices(self.h, env, robot.encode(), _iptr(out), out.size))
        return out[:n].copy()

    def topology(self, env: int) -> dict:
        P, E = self.num_points(env), self.num_edges(env)
        t = dict(pos0=np.zeros((P, 2)), mass=np.zeros(P), fixed=np.zeros(P, dtype=np.uint8),
                 a=np.zeros(E, dtype=np.int32), b=np.zeros(E, dtype=np.int32), k=np.zeros(E), l0=np.zeros(E),
                 on_surface=np.zeros(E, dtype=np.uint8))
        _check(self.L.mevo_env_get_topology(self.h, env, _dptr(t["pos0"]), _dptr(t["mass"]), _bptr(t["fixed"]), _iptr(t["a"]),
                                            _iptr(t["b"]), _dptr(t["k"]), _dptr(t["l0"]), _bptr(t["on_surface"])))
        return t

    # -- actions
    def set_action(self, env: int, robot: str, action) -> bool:
        a = np.ascontiguousarray(action, dtype=np.float64).reshape(-1, 2)
        return bool(_check(self.L.mevo_env_set_action(self.h, env, robot.encode(), _dptr(a), a.shape[0])))

    def set_action_values(self, env: int, robot: str, values) -> bool:
        idx = self.actuator_indices(env, robot)
        return self.set_action(env, robot, np.stack((idx.astype(np.float64), np.asarray(values, dtype=np.float64).ravel()), axis=1))

    @property
    def action_stride(self) -> int:
        return self.L.mevo_batch_action_stride(self.h)

    def set_actions(self, values: np.ndarray) -> None:
        """values: (n_envs, stride) float64 host array (pinned or pageable)."""
        v = np.ascontiguousarray(values, dtype=np.float64)
        assert v.ndim == 2 and v.shape[0] == self.num_envs
        _check(self.L.mevo_batch_set_actions(self.h, _dptr(v), v.shape[1]))

    def fetch_step_results_ptr(self, flags_host_ptr: int, obj: int, com_host_ptr: int) -> int:
        """flags of the last step + COM of ``obj`` into caller-owned (pinned) host buffers, one synchronisation"""
        return _check(self.L.mevo_batch_fetch_step_results(self.h, C.c_void_p(flags_host_ptr) if flags_host_ptr else None, int(obj),
                                                           C.c_void_p(com_host_ptr) if com_host_ptr else None))

    def set_actions_ptr(self, host_ptr: int, stride: int) -> None:
        _check(self.L.mevo_batch_set_actions(self.h, C.cast(host_ptr, _dp), int(stride)))

    def set_actions_device(self, device_ptr: int, stride: int) -> None:
        _check(self.L.mevo_batch_set_actions_device(self.h, C.c_void_p(device_ptr), int(stride)))

    def bind_actions_device(self, device_ptr: int, stride: int) -> None:
        _check(self.L.mevo_batch_bind_actions_device(self.h, C.c_void_p(device_ptr or None), int(stride)))

    def set_auto_reset(self, on: bool) -> None:
        _check(self.L.mevo_batch_set_auto_reset(self.h, int(bool(on))))

    @property
    def unstable_events(self) -> int:
        return int(self.L.mevo_batch_unstable_events(self.h))

    def object_com_device(self, obj: int, device_ptr: int) -> None:
        _check(self.L.mevo_batch_object_com_device(self.h, int(obj), C.c_void_p(device_ptr)))

    # -- stepping
    def step(self, n_steps: int = 1, apply_actions: bool = True, want_flags: bool = True):
        flags = np.zeros(self.num_envs, dtype=np.uint8) if want_flags else None
        n = _check(self.L.mevo_batch_step(self.h, int(n_steps), int(apply_actions), _bptr(flags)))
        return (n, flags) if want_flags else n

    def step_substeps(self, n_steps: int, n_substeps: int, apply_actions: bool = False):
        flags = np.zeros(self.num_envs, dtype=np.uint8)
        n = _check(self.L.mevo_batch_step_substeps(self.h, int(n_steps), int(n_substeps), int(apply_actions), _bptr(flags)))
        return n, flags

    def step_async(self, n_steps: int = 1, apply_actions: bool = True) -> None:
        _check(self.L.mevo_batch_step_async(self.h, int(n_steps), int(apply_actions)))

    def step_static_scratch(self, n_steps: int, actions_device_ptr: int, stride: int, unstable_device_ptr: int = 0) -> None:
        """allocation-free variant of ``step_async`` (``mevo_batch_step_static_scratch``): device action / flag buffers of the caller"""
        _check(self.L.mevo_batch_step_static_scratch(self.h, int(n_steps), C.c_void_p(actions_device_ptr or None), int(stride),
                                                     C.c_void_p(unstable_device_ptr or None)))

    def sync(self) -> None:
        _check(self.L.mevo_batch_sync(self.h))

    @property
    def last_launches(self) -> int:
        return self.L.mevo_batch_last_launches(self.h)

    # -- state
    def get_time(self, env: int) -> int:
        return int(self.L.mevo_env_get_time(self.h, env))

    def set_time(self, env: int, t: int) -> None:
        _check(self.L.mevo_env_set_time(self.h, env, int(t)))

    def get_state(self, env: int) -> Tuple[np.ndarray, np.ndarray, np.ndarray, np.ndarray]:
        P = self.num_points(env)
        out = [np.zeros((P, 2)) for _ in range(4)]
        _check(self.L.mevo_env_get_state(self.h, env, *[_dptr(a) for a in out]))
        return tuple(out)

    def set_state(self, env: int, pos=None, vel=None, pos_last=None, vel_true=None) -> None:
        arrs = [None if a is None else np.ascontiguousarray(a, dtype=np.float64) for a in (pos, vel, pos_last, vel_true)]
        _check(self.L.mevo_env_set_state(self.h, env, *[_dptr(a) for a in arrs]))

    def get_goal(self, env: int) -> np.ndarray:
        out = np.zeros(self.num_edges(env))
        _check(self.L.mevo_env_get_goal(self.h, env, _dptr(out)))
        return out

    def set_goal(self, env: int, goal) -> None:
        g = np.ascontiguousarray(goal, dtype=np.float64)
        _check(self.L.mevo_env_set_goal(self.h, env, _dptr(g)))

    def translate_object(self, env: int, name: str, dx: float, dy: float) -> None:
        _check(self.L.mevo_env_translate_object(self.h, env, name.encode(), float(dx), float(dy)))

    def save_reset_state(self) -> None:
        _check(self.L.mevo_batch_save_reset_state(self.h))

    def reset(self, env_ids: Optional[Sequence[int]] = None) -> None:
        if env_ids is None:
            _check(self.L.mevo_batch_reset(self.h, None, 0))
        else:
            ids = np.ascontiguousarray(env_ids, dtype=np.int32)
            _check(self.L.mevo_batch_reset(self.h, _iptr(ids), ids.size))

    def save_snapshot(self, slot: int) -> None:
        _check(self.L.mevo_batch_save_snapshot(self.h, int(slot)))

    def revert(self, slot: int, env_ids: Optional[Sequence[int]] = None) -> None:
        if env_ids is None:
            _check(self.L.mevo_batch_revert(self.h, None, 0, int(slot)))
        else:
            ids = np.ascontiguousarray(env_ids, dtype=np.int32)
            _check(self.L.mevo_batch_revert(self.h, _iptr(ids), ids.size, int(slot)))

    def read_snapshot(self, env: int, slot: int):
        """(pos, vel, time) of environment ``env`` in snapshot ``slot``; ``None`` for an empty slot"""
        P = self.num_points(env)
        pos, vel = np.zeros((P, 2)), np.zeros((P, 2))
        t = int(self.L.mevo_env_read_snapshot(self.h, int(env), int(slot), _dptr(pos), _dptr(vel)))
        if t < -1:
            raise MevoError(self.L.mevo_last_error().decode())
        return None if t == -1 else (pos, vel, t)

    def object_com(self, obj: int, out: Optional[np.ndarray] = None) -> np.ndarray:
        if out is None:
            out = np.zeros((self.num_envs, 2))
        _check(self.L.mevo_batch_object_com(self.h, int(obj), _dptr(out)))
        return out

    def object_orientation(self, obj: int) -> np.ndarray:
        out = np.zeros(self.num_envs)
        _check(self.L.mevo_batch_object_orientation(self.h, int(obj), _dptr(out)))
        return out

    def sight_obs(self, obj: int, terrain: Sequence[str], sight_dist: int, sight_range: float = 5.0, ceil: bool = False) -> np.ndarray:
        """``get_floor_obs`` / ``get_ceil_obs`` (reference ``evogym/envs/base.py:269-372``) of object ``obj`` in every env"""
        out = np.zeros((self.num_envs, 2 * sight_dist + 1))
        _check(self.L.mevo_batch_sight_obs(self.h, int(obj), ",".join(terrain).encode(), int(sight_dist), float(sight_range), int(ceil),
                                           _dptr(out), out.shape[1]))
        return out

    def object_obs(self, obj: int, stride: int = 0) -> np.ndarray:
        """(n_envs, stride) Walker-style observation rows of object `obj` (evogym/envs/walk.py:58-61)."""
        w = int(self.L.mevo_batch_object_obs_width(self.h, int(obj)))
        if w < 0:
            raise MevoError(self.L.mevo_last_error().decode())
        stride = max(int(stride), w)
        out = np.zeros((self.num_envs, stride))
        _check(self.L.mevo_batch_object_obs(self.h, int(obj), _dptr(out), stride))
        return out

    def object_obs_device(self, obj: int, device_ptr: int, stride: int) -> None:
        _check(self.L.mevo_batch_object_obs_device(self.h, int(obj), C.c_void_p(device_ptr), int(stride)))

    # -- task layer (gym environments either side of Sim::step; include/mevo_b200.h "task layer")
    def task_configure(self, task: int, robot: str = "robot", package: Optional[str] = None, max_episode_steps: int = 0) -> None:
        _check(self.L.mevo_batch_task_configure(self.h, int(task), robot.encode(), package.encode() if package else None,
                                                int(max_episode_steps)))

    def task_dims(self) -> dict:
        d = np.zeros(5, dtype=np.int32)
        _check(self.L.mevo_batch_task_dims(self.h, _iptr(d)))
        return dict(zip(("flat", "modular", "other", "max_voxels", "action_stride"), (int(v) for v in d)))

    def task_env_tables(self, env: int) -> dict:
        nv = self.task_dims()["max_voxels"]
        idx, am, om = np.full((nv, 4), -1, dtype=np.int32), np.zeros(nv, dtype=np.int32), np.zeros(nv, dtype=np.int32)
        corner = _check(self.L.mevo_batch_task_env_tables(self.h, int(env), _iptr(idx), _iptr(am), _iptr(om), nv))
        return {"obs_index_array": idx, "act_mask": am, "obs_mask": om, "voxel_corner_size": corner}

    def task_apply_policy(self, actions: np.ndarray, per_voxel: bool = False) -> None:
        a = np.ascontiguousarray(actions, dtype=np.float64)
        _check(self.L.mevo_batch_task_apply_policy(self.h, _dptr(a), a.shape[1], int(per_voxel)))

    def task_apply_policy_device(self, device_ptr: int, stride: int, per_voxel: bool, is_f32: bool) -> None:
        _check(self.L.mevo_batch_task_apply_policy_device(self.h, C.c_void_p(device_ptr), int(stride), int(per_voxel), int(is_f32)))

    def _task_bufs(self):
        d, n = self.task_dims(), self.num_envs
        return np.zeros((n, d["flat"])), np.zeros((n, d["modular"])), np.zeros((n, d["other"]))

    def task_outputs(self, reset_done: bool = True) -> dict:
        flat, mod, oth = self._task_bufs()
        rew, done = np.zeros(self.num_envs), np.zeros(self.num_envs, dtype=np.uint8)
        _check(self.L.mevo_batch_task_outputs(self.h, _dptr(rew), _bptr(done), _dptr(flat), flat.shape[1], _dptr(mod), mod.shape[1],
                                              _dptr(oth), oth.shape[1], int(reset_done)))
        return {"reward": rew, "done": done, "flat": flat, "modular": mod, "other": oth}

    def task_outputs_device(self, reward_ptr: int, done_ptr: int, flat_ptr: int, flat_stride: int, modular_ptr: int, modular_stride: int,
                            other_ptr: int, other_stride: int, is_f32: bool, reset_done: bool = True) -> None:
        vp = lambda p: C.c_void_p(p) if p else None
        _check(self.L.mevo_batch_task_outputs_device(self.h, vp(reward_ptr), vp(done_ptr), vp(flat_ptr), int(flat_stride), vp(modular_ptr),
                                                     int(modular_stride), vp(other_ptr), int(other_stride), int(is_f32), int(reset_done)))

    def task_reset(self) -> dict:
        flat, mod, oth = self._task_bufs()
        _check(self.L.mevo_batch_task_reset(self.h, _dptr(flat), flat.shape[1], _dptr(mod), mod.shape[1], _dptr(oth), oth.shape[1]))
        return {"flat": flat, "modular": mod, "other": oth}

    def task_reset_device(self, flat_ptr: int, flat_stride: int, modular_ptr: int, modular_stride: int, other_ptr: int, other_stride: int,
                          is_f32: bool) -> None:
        vp = lambda p: C.c_void_p(p) if p else None
        _check(self.L.mevo_batch_task_reset_device(self.h, vp(flat_ptr), int(flat_stride), vp(modular_ptr), int(modular_stride),
                                                   vp(other_ptr), int(other_stride), int(is_f32)))

    def task_launch_sources(self) -> np.ndarray:
        """Catcher-v0: per environment, the environment whose launch position (pegs, package start) it runs now"""
        out = np.zeros(self.num_envs, dtype=np.int32)
        _check(self.L.mevo_batch_task_launch_sources(self.h, _iptr(out)))
        return out

    def task_bind_flag_planes(self, device_ptr: int) -> None:
        """``device_ptr``: 3 * n bytes (a ``torch.bool`` tensor of shape (3, n)) that every later task_outputs call fills with done /
        truncated / unstable as 0 / 1; 0 switches it off"""
        _check(self.L.mevo_batch_task_bind_flag_planes(self.h, C.c_void_p(device_ptr) if device_ptr else None))

    def task_set_goals(self, goals, env_first: int = 0) -> None:
        """BidirectionalWalker-v0: ``goals`` is one sequence of x positions (voxels) per environment, starting at ``env_first``
        (the arguments of the reference's ``WalkToX`` goals, multi_goal.py:62-71); the current goal of those environments is the first."""
        counts = np.asarray([len(g) for g in goals], dtype=np.int32)
        stride = max(1, int(counts.max()) if counts.size else 1)
        table = np.zeros((len(goals), stride), dtype=np.int32)
        for i, g in enumerate(goals):
            table[i, :len(g)] = np.asarray(g, dtype=np.int32)
        _check(self.L.mevo_batch_task_set_goals(self.h, int(env_first), len(goals), _iptr(counts), _iptr(table), stride))

    def task_draw_goals(self, seed: int) -> None:
        """new goal lists for all environments, drawn on the device (set_random_goals(20, 50, 100), multi_goal.py:115-129)"""
        _check(self.L.mevo_batch_task_draw_goals(self.h, int(seed) & 0xFFFFFFFFFFFFFFFF))

    def task_get_goals(self, env: int) -> dict:
        count, current = C.c_int(0), C.c_int(0)
        buf = np.zeros(64, dtype=np.int32)
        _check(self.L.mevo_batch_task_get_goals(self.h, int(env), C.byref(count), C.byref(current), _iptr(buf), 64))
        return {"goals": buf[:count.value].copy(), "current": current.value}

    def render(self, env: int, cam_pos, cam_size, resolution, hide_background=False, hide_grid=False, hide_edges=False,
               hide_boxels=False) -> np.ndarray:
        """``Viewer.render(img_camera)`` + ``Camera.get_image()``: (h, w, 3) uint8, row 0 = top.  ``cam_pos`` / ``cam_size`` in
        engine units (metres), ``resolution`` = (w, h)."""
        w, h = int(resolution[0]), int(resolution[1])
        img = np.zeros((h, w, 3), dtype=np.uint8)
        flags = (1 if hide_background else 0) | (2 if hide_grid else 0) | (4 if hide_edges else 0) | (8 if hide_boxels else 0)
        _check(self.L.mevo_env_render(self.h, int(env), float(cam_pos[0]), float(cam_pos[1]), float(cam_size[0]), float(cam_size[1]),
                                      w, h, flags, _bptr(img)))
        return img

    # -- CUDA graphs (include/mevo_b200.h "CUDA graphs")
    def capture_begin(self) -> None:
        _check(self.L.mevo_batch_capture_begin(self.h))

    def capture_end(self) -> int:
        return _check(self.L.mevo_batch_capture_end(self.h))

    def graph_launch(self, graph: int) -> None:
        _check(self.L.mevo_batch_graph_launch(self.h, int(graph)))

    def graph_destroy(self, graph: int) -> None:
        _check(self.L.mevo_batch_graph_destroy(self.h, int(graph)))

    def object_com_ptr(self, obj: int, host_ptr: int) -> None:
        _check(self.L.mevo_batch_object_com(self.h, int(obj), C.cast(host_ptr, _dp)))

    @property
    def actions_device_ptr(self) -> int:
        return int(self.L.mevo_batch_actions_device_ptr(self.h) or 0)

    @property
    def cuda_stream(self) -> int:
        return int(self.L.mevo_batch_cuda_stream(self.h) or 0)

    @property
    def device_bytes(self) -> int:
        return int(self.L.mevo_batch_device_bytes(self.h))

    # -- test hooks
    def enable_candidate_dump(self, cap_per_env: int = 4096) -> None:
        _check(self.L.mevo_batch_enable_candidate_dump(self.h, int(cap_per_env)))
        self._cand_cap = int(cap_per_env)

    def read_candidates(self, env: int) -> np.ndarray:
        cap = getattr(self, "_cand_cap", 4096)
        out = np.zeros((cap, 4), dtype=np.int32)
        n = _check(self.L.mevo_env_read_candidates(self.h, env, _iptr(out), cap))
        return out[:min(n, cap)].copy()

    def enable_force_dump(self) -> None:
        _check(self.L.mevo_batch_enable_force_dump(self.h))

    def read_collision_forces(self, env: int) -> np.ndarray:
        """(P, 2) ``collision_forces`` of the first substep of the last step call (0 on pruned all-FIXED objects)"""
        out = np.zeros((self.num_points(env), 2))
        _check(self.L.mevo_env_read_collision_forces(self.h, env, _dptr(out)))
        return out

    def error_flags(self, env: int) -> int:
        return _check(self.L.mevo_env_error_flags(self.h, env))


def modular_corner_index(body: np.ndarray) -> Tuple[np.ndarray, int]:
    """``transform_to_modular_obs`` (reference ``evogym/envs/base.py:427-466``) through the library's host code: returns
    ``(index_array (h*w, 4) [TL, TR, BL, BR], voxel_corner_size)``.  Needs no GPU."""
    g = np.ascontiguousarray(body, dtype=np.int32)
    out = np.zeros((g.size, 4), dtype=np.int32)
    loc = _check(load().mevo_modular_corner_index(_iptr(g), g.shape[0], g.shape[1], _iptr(out)))
    return out, loc
