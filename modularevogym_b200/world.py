"""Voxel worlds: the input side of the hot path.

A *world* is an ordered list of voxel objects (robots first, then everything else, each group
alphabetical -- the load order ``EvoSim._init_world_items`` imposes, reference
``evogym/sim.py:52-62``), each described exactly by what the reference hands to
``Sim.read_robot_from_array`` / ``Sim.read_object_from_array`` (``evogym/sim.py:67-97``):

* ``grid``        -- ``(h, w)`` int array of voxel types, **row 0 is the TOP row**,
* ``connections`` -- ``(2, k)`` int array of connected voxel index pairs, indices into ``grid.ravel()``,
* ``x, y``        -- position of the bottom-left voxel in the world grid, in voxels,
* ``is_robot``    -- object contains at least one actuator voxel.

This module is a fresh, small implementation (numpy only) so the package is self-contained on a
machine without the reference checkout; duck-typed ``evogym.EvoWorld`` instances are accepted too
(see :func:`WorldSpec.from_evoworld`).
"""
from __future__ import annotations

import functools
import json
import random as _random
from dataclasses import dataclass, field
from typing import Dict, Iterable, List, Optional, Sequence, Tuple

import numpy as np

VOXEL_TYPES = {"EMPTY": 0, "RIGID": 1, "SOFT": 2, "H_ACT": 3, "V_ACT": 4, "FIXED": 5}
VOXEL_SIZE = 0.1


def full_connectivity(structure: np.ndarray) -> np.ndarray:
    """All 4-adjacent non-empty voxel pairs of ``structure`` as a ``(2, k)`` int array.

    Same contract as reference ``evogym/utils.py:257-292`` (right neighbour, then lower neighbour,
    voxels visited in row-major order)."""
    s = np.asarray(structure)
    h, w = s.shape
    full = s != 0
    idx = np.arange(h * w, dtype=np.int64).reshape(h, w)
    right = np.zeros((h, w), dtype=bool)
    down = np.zeros((h, w), dtype=bool)
    right[:, :-1] = full[:, :-1] & full[:, 1:]
    down[:-1, :] = full[:-1, :] & full[1:, :]
    a = np.concatenate((idx[right], idx[down]))
    if a.size == 0:
        return np.zeros((2, 0), dtype=np.int64)
    b = np.concatenate((idx[right] + 1, idx[down] + w))
    order = np.argsort(np.concatenate((2 * idx[right], 2 * idx[down] + 1)), kind="stable")       # per voxel: right, then lower
    return np.stack((a[order], b[order]))


def is_connected(structure: np.ndarray) -> bool:
    """True when the non-empty voxels form one 4-connected component (``evogym/utils.py:205-238``)."""
    s = np.asarray(structure)
    nz = np.argwhere(s != 0)
    if len(nz) == 0:
        return False
    h, w = s.shape
    seen = np.zeros(s.shape, dtype=bool)
    stack = [tuple(nz[0])]
    seen[stack[0]] = True
    while stack:
        r, c = stack.pop()
        for dr, dc in ((-1, 0), (1, 0), (0, -1), (0, 1)):
            rr, cc = r + dr, c + dc
            if 0 <= rr < h and 0 <= cc < w and s[rr, cc] != 0 and not seen[rr, cc]:
                seen[rr, cc] = True
                stack.append((rr, cc))
    return bool(seen[s != 0].all())


def has_actuator(structure: np.ndarray) -> bool:
    s = np.asarray(structure)
    return bool(((s == VOXEL_TYPES["H_ACT"]) | (s == VOXEL_TYPES["V_ACT"])).any())


def sample_robot(shape: Tuple[int, int], pd: Optional[Sequence[float]] = None,
                 rng=_random) -> Tuple[np.ndarray, np.ndarray]:
    """Random valid (connected, actuated) robot of ``shape`` and its full connectivity.

    Consumes the random stream exactly like reference ``evogym/utils.py:99-152``: one
    ``rng.uniform(0, 1)`` per voxel, inverse-CDF over the voxel-type distribution (default
    ``[0.6, .2, .2, .2, .2]`` normalised), whole-robot rejection until connected and actuated --
    so ``random.seed(i); sample_robot((5, 5))`` reproduces the reference's robot ``i``."""
    if pd is None:
        pd = np.ones(5) / 5
        pd[0] = 0.6
    pd = np.asarray(pd, dtype=np.float64)
    cdf_terms = pd / pd.sum()
    h, w = shape
    while True:
        robot = np.zeros((h, w))
        for r in range(h):
            for c in range(w):
                u = rng.uniform(0, 1)
                acc = 0.0
                pick = None
                for t in range(cdf_terms.size):
                    acc = acc + cdf_terms[t]
                    if u <= acc:
                        pick = t
                        break
                # the reference returns None (-> nan) when rounding leaves u above the last
                # partial sum; that cannot produce a valid robot cell, so treat it as EMPTY.
                robot[r, c] = 0 if pick is None else pick
        if is_connected(robot) and has_actuator(robot):
            return robot, full_connectivity(robot)


def sample_large_robot(shape: Tuple[int, int], rng=_random, p_empty: float = 0.25) -> Tuple[np.ndarray, np.ndarray]:
    """Random connected, actuated robot for shapes where whole-robot rejection (``sample_robot``,
    ``evogym/utils.py:99-152``) practically never terminates (10x10, 20x20: a 43 %-empty random grid
    is almost never one component).  One ``rng.uniform(0, 1)`` per voxel over
    ``[p_empty, rest equally]``, then only the largest 4-connected component is kept."""
    h, w = shape
    edges = [p_empty + (1.0 - p_empty) * k / 4.0 for k in range(5)]
    while True:
        robot = np.zeros((h, w))
        for r in range(h):
            for c in range(w):
                u = rng.uniform(0, 1)
                robot[r, c] = next((t for t in range(5) if u <= edges[t]), 4)
        comp = -np.ones((h, w), dtype=np.int64)
        sizes = []
        for r0 in range(h):
            for c0 in range(w):
                if robot[r0, c0] == 0 or comp[r0, c0] >= 0:
                    continue
                cid, n, stack = len(sizes), 0, [(r0, c0)]
                comp[r0, c0] = cid
                while stack:
                    r, c = stack.pop()
                    n += 1
                    for dr, dc in ((-1, 0), (1, 0), (0, -1), (0, 1)):
                        rr, cc = r + dr, c + dc
                        if 0 <= rr < h and 0 <= cc < w and robot[rr, cc] != 0 and comp[rr, cc] < 0:
                            comp[rr, cc] = cid
                            stack.append((rr, cc))
                sizes.append(n)
        if not sizes:
            continue
        robot[comp != int(np.argmax(sizes))] = 0
        if has_actuator(robot):
            return robot, full_connectivity(robot)


@dataclass
class ObjectSpec:
    """One voxel object, in the form the simulator's ``read_*_from_array`` takes."""
    name: str
    grid: np.ndarray          # (h, w) ints, row 0 = top
    connections: np.ndarray   # (2, k) ints into grid.ravel()
    x: int
    y: int

    @property
    def is_robot(self) -> bool:
        return has_actuator(self.grid)

    @property
    def size(self) -> Tuple[int, int]:
        return int(self.grid.shape[1]), int(self.grid.shape[0])  # (w, h)


@dataclass
class WorldSpec:
    """Ordered voxel world."""
    width: int = 1
    height: int = 1
    objects: Dict[str, ObjectSpec] = field(default_factory=dict)

    # ---- construction -------------------------------------------------------------------
    @classmethod
    def from_json(cls, path: str) -> "WorldSpec":
        """Load a world ``.json`` (format of reference ``evogym/world.py:39-72,367-433``:
        ``grid_width``, ``grid_height``, ``objects{name: {indices, types, neighbors}}`` with voxel
        indices counted row-major from the BOTTOM-left of the world grid)."""
        with open(path, "r") as f:
            state = json.load(f)
        return cls.from_parsed_json(state)

    @classmethod
    def from_parsed_json(cls, state: dict) -> "WorldSpec":
        gw, gh = int(state["grid_width"]), int(state["grid_height"])
        world = cls(width=1, height=1)
        for name, od in state["objects"].items():
            idx = [int(i) for i in od["indices"]]
            types = [int(t) for t in od["types"]]
            if len(idx) != len(types) or len(idx) != len(od["neighbors"]) or not idx:
                raise ValueError(f"corrupted object {name}")
            xs = [i % gw for i in idx]
            ys = [i // gw for i in idx]
            x0, y0 = min(xs), min(ys)
            w, h = max(xs) - x0 + 1, max(ys) - y0 + 1
            grid = np.zeros((h, w), dtype=np.int64)       # row 0 = top
            local = {}
            for i, xx, yy, t in zip(idx, xs, ys, types):
                r, c = (h - 1) - (yy - y0), xx - x0
                grid[r, c] = t
                local[i] = r * w + c
            conn: List[Tuple[int, int]] = []
            for i in idx:
                for n in od["neighbors"][str(i)]:
                    if int(n) not in local:
                        raise ValueError(f"object {name} has voxels with invalid neighbors")
                    conn.append((local[i], local[int(n)]))
            cn = np.asarray(conn, dtype=np.int64).T if conn else np.zeros((2, 0), dtype=np.int64)
            world.add_object(ObjectSpec(name, grid, cn, x0, y0))
        world.width = max(world.width, gw)
        world.height = max(world.height, gh)
        return world

    @classmethod
    def from_evoworld(cls, world) -> "WorldSpec":
        """Convert a duck-typed reference ``EvoWorld`` (``.grid_size``, ``.objects[name].grid`` with
        row 0 = bottom, ``.neighbors`` keyed by ``Pair``) -- mirrors ``evogym/sim.py:64-97``."""
        out = cls(width=int(world.grid_size.x), height=int(world.grid_size.y))
        for name, obj in world.objects.items():
            gw, gh = int(obj.grid_size.x), int(obj.grid_size.y)
            conn: List[Tuple[int, int]] = []
            by_xy = {(int(k.x), int(k.y)): neis for k, neis in obj.neighbors.items()}
            for y in range(gh):
                for x in range(gw):
                    if obj.grid[y][x] == 0:
                        continue
                    for nei in by_xy.get((x, y), ()):
                        conn.append(((gh - 1 - y) * gw + x, (gh - 1 - nei.y) * gw + nei.x))
            cn = np.asarray(conn, dtype=np.int64).T if conn else np.zeros((2, 0), dtype=np.int64)
            out.objects[name] = ObjectSpec(name, np.flipud(np.asarray(obj.grid)).astype(np.int64), cn,
                                           int(obj.pos.x), int(obj.pos.y))
        return out

    def add_object(self, obj: ObjectSpec) -> None:
        if obj.name in self.objects:
            raise ValueError(f"cannot add object with duplicate name: {obj.name}")
        if obj.x < 0 or obj.y < 0:
            raise ValueError(f"object {obj.name} position must be non-negative")
        w, h = obj.size
        for o in self.objects.values():                 # overlap with an existing object: compare the grids where the boxes meet
            ow, oh = o.size
            x0, x1 = max(obj.x, o.x), min(obj.x + w, o.x + ow)
            y0, y1 = max(obj.y, o.y), min(obj.y + h, o.y + oh)
            if x0 >= x1 or y0 >= y1:
                continue
            mine = obj.grid[h - (y1 - obj.y):h - (y0 - obj.y), x0 - obj.x:x1 - obj.x]          # row 0 = top
            theirs = o.grid[oh - (y1 - o.y):oh - (y0 - o.y), x0 - o.x:x1 - o.x]
            if ((mine != 0) & (theirs != 0)).any():
                raise ValueError(f"cannot add object {obj.name}: overlaps an existing object")
        self.width = max(self.width, obj.x + w)
        self.height = max(self.height, obj.y + h)
        self.objects[obj.name] = obj

    def add_from_array(self, name: str, structure: np.ndarray, x: int, y: int,
                       connections: Optional[np.ndarray] = None) -> None:
        """Add an object from a ``(h, w)`` structure with row 0 = TOP (the orientation robots are
        written in; reference ``evogym/world.py:75-94``)."""
        grid = np.asarray(structure).astype(np.int64)
        cn = full_connectivity(grid) if connections is None else np.asarray(connections, dtype=np.int64)
        if cn.size == 0:
            cn = np.zeros((2, 0), dtype=np.int64)
        flat = grid.ravel()
        if cn.size and ((cn < 0).any() or (cn >= flat.size).any() or (flat[np.clip(cn, 0, flat.size - 1)] == 0).any()):
            raise ValueError(f"could not load object {name}: connections contain invalid elements")
        self.add_object(ObjectSpec(name, grid, cn, int(x), int(y)))

    # ---- queries ------------------------------------------------------------------------
    def occupancy(self) -> set:
        occ = set()
        for o in self.objects.values():
            w, h = o.size
            for r in range(h):
                for c in range(w):
                    if o.grid[r, c] != 0:
                        occ.add((o.x + c, o.y + (h - 1 - r)))
        return occ

    def load_order(self) -> List[ObjectSpec]:
        """Robots first, then the rest, each alphabetical (``evogym/sim.py:52-62``); this fixes the
        global point/edge numbering."""
        keyed = sorted(((0 if o.is_robot else 1), name) for name, o in self.objects.items())
        return [self.objects[name] for _, name in keyed]

    def to_dict(self) -> dict:
        """Plain-JSON form (used by the committed test fixtures)."""
        return {"width": int(self.width), "height": int(self.height),
                "objects": [{"name": o.name, "grid": np.asarray(o.grid).astype(int).tolist(),
                             "connections": np.asarray(o.connections).astype(int).reshape(2, -1).tolist(),
                             "x": int(o.x), "y": int(o.y)} for o in self.objects.values()]}

    @classmethod
    def from_dict(cls, d: dict) -> "WorldSpec":
        w = cls(width=int(d["width"]), height=int(d["height"]))
        for o in d["objects"]:
            cn = np.asarray(o["connections"], dtype=np.int64).reshape(2, -1)
            w.objects[o["name"]] = ObjectSpec(o["name"], np.asarray(o["grid"], dtype=np.int64), cn, int(o["x"]), int(o["y"]))
        return w

    def copy(self) -> "WorldSpec":
        w = WorldSpec(self.width, self.height)
        for name, o in self.objects.items():
            w.objects[name] = ObjectSpec(o.name, o.grid.copy(), o.connections.copy(), o.x, o.y)
        return w


def _frozen(w: "WorldSpec") -> "WorldSpec":
    """terrain built once per process and shared by every world a builder hands out: its arrays become read-only"""
    for o in w.objects.values():
        o.grid.setflags(write=False)
        o.connections.setflags(write=False)
    return w


def _shared(terrain: "WorldSpec") -> "WorldSpec":
    w = WorldSpec(width=terrain.width, height=terrain.height)
    w.objects.update(terrain.objects)
    return w


def flat_ground_world(width: int = 100, name: str = "ground") -> WorldSpec:
    """``width`` x 1 strip of FIXED voxels at the bottom of the world: the terrain of Walker-v0
    (reference ``evogym/envs/sim_files/Walker-v0.json``: grid 100 x 1, one object ``ground``)."""
    w = WorldSpec(width=width, height=1)
    grid = np.full((1, width), VOXEL_TYPES["FIXED"], dtype=np.int64)
    w.add_from_array(name, grid, 0, 0)
    return w


@functools.lru_cache(maxsize=None)
def _walker_terrain(width: int = 100) -> WorldSpec:
    w = flat_ground_world(width)
    return _frozen(w)


def walker_world(robot: np.ndarray, connections: Optional[np.ndarray] = None, width: int = 100) -> WorldSpec:
    """Walker-v0: flat fixed ground + ``robot`` placed at voxel (1, 1) (``evogym/envs/walk.py:19-20``)."""
    w = _shared(_walker_terrain(width))
    w.add_from_array("robot", robot, 1, 1, connections)
    return w


def _package() -> np.ndarray:
    """the 3 x 2 SOFT block named ``package`` of Pusher-v0 / Thrower-v0 (``evogym/envs/sim_files/Pusher-v0.json``)"""
    return np.full((2, 3), VOXEL_TYPES["SOFT"], dtype=np.int64)


@functools.lru_cache(maxsize=None)
def _pusher_terrain() -> WorldSpec:
    w = flat_ground_world(100)
    w.add_from_array("package", _package(), 8, 1)
    return _frozen(w)


def pusher_world(robot: np.ndarray, connections: Optional[np.ndarray] = None) -> WorldSpec:
    """Pusher-v0 (``evogym/envs/sim_files/Pusher-v0.json`` + ``evogym/envs/manipulate.py:323-324``): 100 x 1 fixed
    ground, the package at voxel (8, 1), the robot at (1, 1)."""
    w = _shared(_pusher_terrain())
    w.add_from_array("robot", robot, 1, 1, connections)
    return w


@functools.lru_cache(maxsize=None)
def _carrier_terrain() -> WorldSpec:
    w = flat_ground_world(100)
    w.add_from_array("package", _package(), 3, 6)
    return _frozen(w)


def carrier_world(robot: np.ndarray, connections: Optional[np.ndarray] = None) -> WorldSpec:
    """Carrier-v0 (``sim_files/Carrier-v0.json`` + ``evogym/envs/manipulate.py:88-89``): 100 x 1 fixed ground, the SOFT package
    hanging at voxel (3, 6) above the robot at (1, 1)."""
    w = _shared(_carrier_terrain())
    w.add_from_array("robot", robot, 1, 1, connections)
    return w


@functools.lru_cache(maxsize=None)
def _pusher_v1_terrain() -> WorldSpec:
    w = flat_ground_world(100)
    w.add_from_array("package", np.full((2, 3), VOXEL_TYPES["RIGID"], dtype=np.int64), 8, 1)
    return _frozen(w)


def pusher_v1_world(robot: np.ndarray, connections: Optional[np.ndarray] = None) -> WorldSpec:
    """Pusher-v1 (``sim_files/Pusher-v1.json`` + ``manipulate.py:424-425``): 100 x 1 fixed ground, a RIGID package at voxel (8, 1),
    the robot behind it at (13, 1): it has to get to the other side first."""
    w = _shared(_pusher_v1_terrain())
    w.add_from_array("robot", robot, 13, 1, connections)
    return w


@functools.lru_cache(maxsize=None)
def _balancer_terrain(pole: int) -> WorldSpec:
    w = WorldSpec(width=30, height=5)
    grid = np.zeros((5, 30), dtype=np.int64)
    grid[4, :] = VOXEL_TYPES["FIXED"]
    grid[:, 0] = grid[:, 29] = VOXEL_TYPES["FIXED"]
    grid[4 - pole:4, 17] = VOXEL_TYPES["FIXED"]
    w.add_from_array("ground", grid, 0, 0)
    return _frozen(w)


def balancer_world(robot: np.ndarray, connections: Optional[np.ndarray] = None) -> WorldSpec:
    """Balancer-v0 (``sim_files/Balancer-v0.json`` + ``evogym/envs/balance.py:19-20``): a walled 30 x 5 box with a 2-voxel pole
    at column 17, the robot on top of it at voxel (15, 3)."""
    w = _shared(_balancer_terrain(2))
    w.add_from_array("robot", robot, 15, 3, connections)
    return w


def balancer_v1_world(robot: np.ndarray, connections: Optional[np.ndarray] = None) -> WorldSpec:
    """Balancer-v1 (``sim_files/Balancer-v1.json`` + ``balance.py:143-144``): the same box with a 3-voxel pole, the robot on the floor
    at voxel (10, 1): it has to jump onto the pole."""
    w = _shared(_balancer_terrain(3))
    w.add_from_array("robot", robot, 10, 1, connections)
    return w


# (name, width, x, y) of the one-voxel-thick FIXED platforms of GapJumper-v0.json (80 x 6) and PlatformJumper-v0.json (91 x 9)
GAP_JUMPER_PLATFORMS = (("platform_1", 20, 0, 5), ("platform_2", 10, 25, 5), ("platform_3", 10, 40, 5), ("platform_4", 10, 55, 5),
                        ("platform_5", 10, 70, 5))
PLATFORM_JUMPER_PLATFORMS = (("platform_1", 20, 0, 5), ("platform_2", 10, 21, 7), ("platform_3", 10, 33, 8), ("platform_4", 10, 46, 6),
                             ("platform_5", 10, 57, 5), ("platform_6", 10, 70, 6), ("platform_7", 10, 81, 3))


def _platform_terrain(width: int, height: int, platforms) -> WorldSpec:
    w = WorldSpec(width=width, height=height)
    for name, pw, x, y in platforms:
        w.add_from_array(name, np.full((1, pw), VOXEL_TYPES["FIXED"], dtype=np.int64), x, y)
    return _frozen(w)


@functools.lru_cache(maxsize=None)
def _gap_jumper_terrain() -> WorldSpec:
    return _platform_terrain(80, 6, GAP_JUMPER_PLATFORMS)


@functools.lru_cache(maxsize=None)
def _platform_jumper_terrain() -> WorldSpec:
    return _platform_terrain(91, 9, PLATFORM_JUMPER_PLATFORMS)


def gap_jumper_world(robot: np.ndarray, connections: Optional[np.ndarray] = None) -> WorldSpec:
    """GapJumper-v0 (``sim_files/GapJumper-v0.json`` + ``evogym/envs/traverse.py:668-669``): five fixed platforms at one height with
    gaps between them, the robot at voxel (1, 6)."""
    w = _shared(_gap_jumper_terrain())
    w.add_from_array("robot", robot, 1, 6, connections)
    return w


def platform_jumper_world(robot: np.ndarray, connections: Optional[np.ndarray] = None) -> WorldSpec:
    """PlatformJumper-v0 (``sim_files/PlatformJumper-v0.json`` + ``traverse.py:544-545``): seven fixed platforms at different
    heights, the robot at voxel (1, 6)."""
    w = _shared(_platform_jumper_terrain())
    w.add_from_array("robot", robot, 1, 6, connections)
    return w


# rows of the "ground" object of evogym/envs/sim_files/Carrier-v1.json (60 x 4, placed at voxel (0, 1)), top row first
CARRIER_V1_ROWS = (
    "................................................FF..........",
    "FFFFFFFFFFFFFFFFFFFFFFFF........................FF..........",
    ".......................FFFF.....................FF..........",
    "..........................FFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFF",
)


@functools.lru_cache(maxsize=None)
def _carrier_v1_terrain() -> WorldSpec:
    w = WorldSpec(width=60, height=12)
    w.add_from_array("ground", grid_from_rows(CARRIER_V1_ROWS), 0, 1)
    w.add_from_array("package", np.full((2, 3), VOXEL_TYPES["RIGID"], dtype=np.int64), 3, 10)
    return _frozen(w)


def carrier_v1_world(robot: np.ndarray, connections: Optional[np.ndarray] = None) -> WorldSpec:
    """Carrier-v1 (``sim_files/Carrier-v1.json`` + ``evogym/envs/manipulate.py:207-208``): a raised start shelf, two steps down and a
    table post, a RIGID package at voxel (3, 10) above the robot at (1, 4)."""
    w = _shared(_carrier_v1_terrain())
    w.add_from_array("robot", robot, 1, 4, connections)
    return w


# the "ground" object of evogym/envs/sim_files/Lifter-v0.json (11 x 5): a pit between two walled shelves
LIFTER_ROWS = (
    "F.........F",
    "F.........F",
    "FFFF...FFFF",
    "...F...F...",
    "...FFFFF...",
)


@functools.lru_cache(maxsize=None)
def _lifter_terrain() -> WorldSpec:
    w = WorldSpec(width=11, height=5)
    w.add_from_array("ground", grid_from_rows(LIFTER_ROWS), 0, 0)
    w.add_from_array("package", np.full((2, 1), VOXEL_TYPES["RIGID"], dtype=np.int64), 5, 1)
    return _frozen(w)


def lifter_world(robot: np.ndarray, connections: Optional[np.ndarray] = None) -> WorldSpec:
    """Lifter-v0 (``sim_files/Lifter-v0.json`` + ``evogym/envs/manipulate.py:1212-1213``): a RIGID 1 x 2 package at the bottom of a
    pit, the robot above it at voxel (2, 3)."""
    w = _shared(_lifter_terrain())
    w.add_from_array("robot", robot, 2, 3, connections)
    return w


@functools.lru_cache(maxsize=None)
def _flipper_terrain() -> WorldSpec:
    w = WorldSpec(width=70, height=1)
    w.add_from_array("new_object_1", np.full((1, 70), VOXEL_TYPES["FIXED"], dtype=np.int64), 0, 0)
    return _frozen(w)


def flipper_world(robot: np.ndarray, connections: Optional[np.ndarray] = None) -> WorldSpec:
    """Flipper-v0 (``sim_files/Flipper-v0.json`` + ``evogym/envs/flip.py:20-21``): a 70-wide fixed floor ``new_object_1``, the
    robot at voxel (60, 1)."""
    w = _shared(_flipper_terrain())
    w.add_from_array("robot", robot, 60, 1, connections)
    return w


# (name, voxel type, w, h, x, y) of the objects of BeamToppler-v0.json / BeamSlider-v0.json (70 x 7) besides the 70 x 1 FIXED "ground"
BEAM_TOPPLER_OBJECTS = (("beam", "RIGID", 10, 2, 29, 5), ("peg1", "FIXED", 5, 1, 27, 4), ("peg2", "FIXED", 5, 1, 36, 4))
BEAM_SLIDER_OBJECTS = (("beam", "RIGID", 10, 2, 22, 5), ("peg1", "FIXED", 5, 1, 20, 4), ("peg2", "FIXED", 3, 1, 29, 4),
                       ("peg3", "FIXED", 3, 1, 35, 4), ("peg4", "FIXED", 3, 1, 41, 4), ("peg5", "FIXED", 3, 1, 49, 4))


def _beam_terrain(objects) -> WorldSpec:
    w = WorldSpec(width=70, height=7)
    w.add_from_array("ground", np.full((1, 70), VOXEL_TYPES["FIXED"], dtype=np.int64), 0, 0)
    for name, kind, ow, oh, x, y in objects:
        w.add_from_array(name, np.full((oh, ow), VOXEL_TYPES[kind], dtype=np.int64), x, y)
    return _frozen(w)


@functools.lru_cache(maxsize=None)
def _beam_toppler_terrain() -> WorldSpec:
    return _beam_terrain(BEAM_TOPPLER_OBJECTS)


@functools.lru_cache(maxsize=None)
def _beam_slider_terrain() -> WorldSpec:
    return _beam_terrain(BEAM_SLIDER_OBJECTS)


def beam_toppler_world(robot: np.ndarray, connections: Optional[np.ndarray] = None) -> WorldSpec:
    """BeamToppler-v0 (``sim_files/BeamToppler-v0.json`` + ``evogym/envs/manipulate.py:892-893``): a RIGID 10 x 2 beam resting on two
    fixed pegs, the robot at voxel (1, 1)."""
    w = _shared(_beam_toppler_terrain())
    w.add_from_array("robot", robot, 1, 1, connections)
    return w


def beam_slider_world(robot: np.ndarray, connections: Optional[np.ndarray] = None) -> WorldSpec:
    """BeamSlider-v0 (``sim_files/BeamSlider-v0.json`` + ``manipulate.py:1054-1055``): the beam on the first of a row of pegs, the
    robot at voxel (1, 1)."""
    w = _shared(_beam_slider_terrain())
    w.add_from_array("robot", robot, 1, 1, connections)
    return w


@functools.lru_cache(maxsize=None)
def _shape_change_terrain() -> WorldSpec:
    w = WorldSpec(width=20, height=10)
    grid = np.zeros((10, 20), dtype=np.int64)
    grid[9, :] = VOXEL_TYPES["FIXED"]
    grid[:, 0:2] = grid[:, 18:20] = VOXEL_TYPES["FIXED"]
    w.add_from_array("new_object_4", grid, 0, 0)
    return _frozen(w)


def shape_change_world(robot: np.ndarray, connections: Optional[np.ndarray] = None) -> WorldSpec:
    """The ShapeChange environments (``sim_files/ShapeChange.json`` + ``evogym/envs/change_shape.py:292-293``): a walled 20 x 10 box
    ``new_object_4``, the robot at voxel (7, 1)."""
    w = _shared(_shape_change_terrain())
    w.add_from_array("robot", robot, 7, 1, connections)
    return w


@functools.lru_cache(maxsize=None)
def _bidirectional_walker_terrain() -> WorldSpec:
    w = WorldSpec(width=70, height=1)
    w.add_from_array("new_object_1", np.full((1, 70), VOXEL_TYPES["FIXED"], dtype=np.int64), 0, 0)
    return _frozen(w)


def bidirectional_walker_world(robot: np.ndarray, connections: Optional[np.ndarray] = None) -> WorldSpec:
    """BidirectionalWalker-v0 (``sim_files/BidirectionalWalker-v0.json`` + ``evogym/envs/multi_goal.py:96-97``): a flat strip of 70
    fixed voxels ``new_object_1``, the robot in its middle at voxel (33, 1)."""
    w = _shared(_bidirectional_walker_terrain())
    w.add_from_array("robot", robot, 33, 1, connections)
    return w


# BlockSoup.blocks (evogym/envs/traverse.py:792-813): (x, y, edge length) of the loose rigid blocks in the pit
TRAVERSER_BLOCKS = ((20, 1, 3), (24, 1, 3), (27, 1, 2), (31, 1, 3), (34, 1, 3), (38, 1, 2), (41, 1, 3), (45, 1, 3),
                    (21, 4, 3), (24, 4, 2), (26, 4, 1), (29, 4, 2), (31, 4, 3), (34, 4, 2), (36, 4, 2), (40, 4, 3), (43, 4, 2), (46, 4, 3)) + \
                   tuple((x, 7, 2) for x in range(20, 45, 3))


@functools.lru_cache(maxsize=None)
def _traverser_terrain() -> WorldSpec:
    w = WorldSpec(width=60, height=9)
    grid = np.zeros((9, 60), dtype=np.int64)
    fixed = VOXEL_TYPES["FIXED"]
    grid[8 - 8, 0:20] = fixed                    # start ledge (y = 8)
    grid[:, 19] = fixed                          # left wall of the pit
    grid[8 - 0, 19:50] = fixed                   # pit floor (y = 0)
    grid[8 - 6:, 49] = fixed                     # right wall, up to y = 6
    grid[8 - 6, 49:60] = fixed                   # exit ledge (y = 6)
    w.add_from_array("ground", grid, 0, 0)
    for k, (x, y, size) in enumerate(TRAVERSER_BLOCKS):
        w.add_from_array(f"block_{k}", np.full((size, size), VOXEL_TYPES["RIGID"], dtype=np.int64), x, y)
    return _frozen(w)


def traverser_world(robot: np.ndarray, connections: Optional[np.ndarray] = None) -> WorldSpec:
    """Traverser-v0 (``sim_files/Traverser-v0.json`` + ``rigid_{1,2,3}x{1,2,3}.json`` + ``evogym/envs/traverse.py:788-823``): a pit
    between two ledges, filled with 27 loose rigid blocks ``block_0 ... block_26``; the robot on the start ledge at voxel (1, 9)."""
    w = _shared(_traverser_terrain())
    w.add_from_array("robot", robot, 1, 9, connections)
    return w


@functools.lru_cache(maxsize=None)
def _catcher_terrain(offsetx: int, offsety: int) -> WorldSpec:
    w = flat_ground_world(100)
    w.add_from_array("package", _package(), 21 + offsetx, 41 + offsety)
    peg = np.full((1, 5), VOXEL_TYPES["FIXED"], dtype=np.int64)
    w.add_from_array("peg1", peg, 17 + offsetx, 39 + offsety)
    w.add_from_array("peg2", peg, 19 + offsetx, 25 + offsety)
    return _frozen(w)


def catcher_offsets() -> Tuple[int, int]:
    """where CatchSmallRect.random_init hangs the package and its two pegs (``evogym/envs/manipulate.py:652-653``): the reference's
    own two calls of Python's global ``random``"""
    offsetx = _random.randint(-6, 4)
    offsety = _random.randint(0, 5)
    return offsetx, offsety


def catcher_world(robot: np.ndarray, connections: Optional[np.ndarray] = None, offsetx: Optional[int] = None,
                  offsety: Optional[int] = None) -> WorldSpec:
    """Catcher-v0 (``sim_files/Walker-v0.json`` + ``package.json`` + ``peg.json`` + ``evogym/envs/manipulate.py:646-671``): flat
    ground, the robot at voxel (22, 1), high above it a package resting on the end of one fixed peg with a second peg below -- the
    package tips off and tumbles down.  The whole arrangement is shifted by a random offset (drawn here like the reference draws
    it when none is given)."""
    if offsetx is None or offsety is None:
        offsetx, offsety = catcher_offsets()
    w = _shared(_catcher_terrain(int(offsetx), int(offsety)))
    w.add_from_array("robot", robot, 22, 1, connections)
    return w


THROWER_POSTS = {10: 2, 14: 3, 18: 4, 23: 6, 28: 7, 39: 9, 69: 9}     # column -> voxels above the ground row


@functools.lru_cache(maxsize=None)
def _thrower_terrain() -> WorldSpec:
    w = WorldSpec(width=70, height=10)
    grid = np.zeros((10, 70), dtype=np.int64)
    grid[9, :] = VOXEL_TYPES["FIXED"]
    for col, hgt in THROWER_POSTS.items():
        grid[9 - hgt:9, col] = VOXEL_TYPES["FIXED"]
    w.add_from_array("ground", grid, 0, 0)
    w.add_from_array("package", _package(), 3, 7)
    return _frozen(w)


def thrower_world(robot: np.ndarray, connections: Optional[np.ndarray] = None) -> WorldSpec:
    """Thrower-v0 (``evogym/envs/sim_files/Thrower-v0.json`` + ``evogym/envs/manipulate.py:524-525``): 70 x 10 fixed
    terrain (ground row + posts of rising height), the package at voxel (3, 7), the robot at (1, 1)."""
    w = _shared(_thrower_terrain())
    w.add_from_array("robot", robot, 1, 1, connections)
    return w


@functools.lru_cache(maxsize=None)
def _climber_terrain(width: int) -> WorldSpec:
    w = WorldSpec(width=width, height=90)
    grid = np.zeros((90, width), dtype=np.int64)
    grid[:, 0] = grid[:, width - 1] = VOXEL_TYPES["FIXED"]
    grid[0, :] = grid[89, :] = VOXEL_TYPES["FIXED"]
    w.add_from_array("pipe", grid, 0, 0)
    return _frozen(w)


def climber_world(robot: np.ndarray, connections: Optional[np.ndarray] = None) -> WorldSpec:
    """Climber-v0 (``evogym/envs/sim_files/Climber-v0.json`` + ``evogym/envs/climb.py:45-46``): a 7 x 90 fixed ``pipe`` (closed
    at both ends, walls one voxel thick), the robot at voxel (1, 1).  The reference's pipe is 5 voxels wide inside; for a wider
    robot (BASELINE configs[3] names Climber with 10x10 / 20x20 morphologies, which cannot exist in that pipe) the pipe is
    widened to the robot's width + 2 -- a synthetic Climber-style world, same wall contact on both sides."""
    width = max(7, int(np.asarray(robot).shape[1]) + 2)
    w = _shared(_climber_terrain(width))
    w.add_from_array("robot", robot, 1, 1, connections)
    return w


# column heights (voxels of FIXED ground from the bottom) of evogym/envs/sim_files/ObstacleTraverser-v1.json, 60 x 10
OBSTACLE_TRAVERSER_V1_HEIGHTS = (3, 3, 4, 3, 4, 3, 3, 3, 2, 3, 2, 2, 3, 3, 2, 2, 1, 3, 4, 3, 2, 3, 3, 4, 3, 2, 3, 2, 3, 3, 2, 3, 4, 4, 2, 3, 4,
                                 3, 4, 4, 4, 2, 3, 4, 5, 4, 3, 4, 3, 4, 3, 2, 2, 3, 4, 3, 2, 3, 10, 10)


@functools.lru_cache(maxsize=None)
def _obstacle_traverser_terrain() -> WorldSpec:
    w = WorldSpec(width=60, height=10)
    grid = np.zeros((10, 60), dtype=np.int64)
    for c, hgt in enumerate(OBSTACLE_TRAVERSER_V1_HEIGHTS):
        grid[10 - hgt:, c] = VOXEL_TYPES["FIXED"]
    w.add_from_array("ground", grid, 0, 0)
    return _frozen(w)


def obstacle_traverser_world(robot: np.ndarray, connections: Optional[np.ndarray] = None) -> WorldSpec:
    """ObstacleTraverser-v1 (``evogym/envs/sim_files/ObstacleTraverser-v1.json`` + ``evogym/envs/traverse.py:355-356``): 60 x 10
    bumpy fixed ``ground``, the robot at voxel (2, 4)."""
    w = _shared(_obstacle_traverser_terrain())
    w.add_from_array("robot", robot, 2, 4, connections)
    return w


# column heights of evogym/envs/sim_files/ObstacleTraverser-v0.json, 80 x 3
OBSTACLE_TRAVERSER_V0_HEIGHTS = (1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 2, 1, 1, 1, 1, 1, 2, 1, 1, 1, 1, 2, 2, 1, 1, 1, 1,
                                 2, 1, 1, 2, 2, 1, 2, 1, 2, 1, 1, 2, 2, 1, 1, 2, 3, 2, 2, 1, 2, 3, 1, 2, 2, 3, 2, 1, 2, 2, 1, 1, 1, 2, 1, 1, 1, 1,
                                 1, 1, 1)
# one-voxel-thick stair treads of UpStepper-v0.json (70 x 9) / DownStepper-v0.json (75 x 10): tread y spans the columns
# [b[y], b[y + 1]] (ascending) resp. [b[h - y], b[h - 1 - y]] read from the right (descending); neighbouring treads share a column
UP_STEPPER_BREAKS = (0, 20, 29, 36, 42, 47, 51, 55, 61, 69)
DOWN_STEPPER_BREAKS = (0, 20, 28, 35, 41, 47, 52, 57, 61, 65, 74)


@functools.lru_cache(maxsize=None)
def _obstacle_traverser_v0_terrain() -> WorldSpec:
    w = WorldSpec(width=80, height=3)
    grid = np.zeros((3, 80), dtype=np.int64)
    for c, hgt in enumerate(OBSTACLE_TRAVERSER_V0_HEIGHTS):
        grid[3 - hgt:, c] = VOXEL_TYPES["FIXED"]
    w.add_from_array("ground", grid, 0, 0)
    return _frozen(w)


def obstacle_traverser_v0_world(robot: np.ndarray, connections: Optional[np.ndarray] = None) -> WorldSpec:
    """ObstacleTraverser-v0 (``sim_files/ObstacleTraverser-v0.json`` + ``evogym/envs/traverse.py:256-257``): 80 x 3 bumpy fixed
    ``ground``, the robot at voxel (2, 1)."""
    w = _shared(_obstacle_traverser_v0_terrain())
    w.add_from_array("robot", robot, 2, 1, connections)
    return w


@functools.lru_cache(maxsize=None)
def _up_stepper_terrain() -> WorldSpec:
    b = UP_STEPPER_BREAKS
    w = WorldSpec(width=70, height=9)
    grid = np.zeros((9, 70), dtype=np.int64)
    for y in range(9):
        grid[8 - y, b[y]:b[y + 1] + 1] = VOXEL_TYPES["FIXED"]
    w.add_from_array("ground", grid, 0, 0)
    return _frozen(w)


def up_stepper_world(robot: np.ndarray, connections: Optional[np.ndarray] = None) -> WorldSpec:
    """UpStepper-v0 (``sim_files/UpStepper-v0.json`` + ``traverse.py:56-57``): nine ascending treads, the robot at voxel (1, 1)."""
    w = _shared(_up_stepper_terrain())
    w.add_from_array("robot", robot, 1, 1, connections)
    return w


@functools.lru_cache(maxsize=None)
def _down_stepper_terrain() -> WorldSpec:
    b = DOWN_STEPPER_BREAKS
    w = WorldSpec(width=75, height=10)
    grid = np.zeros((10, 75), dtype=np.int64)
    for k in range(10):                               # tread k from the left sits at height 9 - k
        grid[k, b[k]:b[k + 1] + 1] = VOXEL_TYPES["FIXED"]
    w.add_from_array("ground", grid, 0, 0)
    return _frozen(w)


def down_stepper_world(robot: np.ndarray, connections: Optional[np.ndarray] = None) -> WorldSpec:
    """DownStepper-v0 (``sim_files/DownStepper-v0.json`` + ``traverse.py:157-158``): ten descending treads, the robot at voxel (1, 11)."""
    w = _shared(_down_stepper_terrain())
    w.add_from_array("robot", robot, 1, 11, connections)
    return w


# column heights of evogym/envs/sim_files/Hurdler-v0.json, 70 x 6
HURDLER_HEIGHTS = (1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 2, 1, 1, 1, 1, 1, 1, 1, 3, 1, 1, 1, 1, 1, 1, 1, 2, 1, 1, 1, 1, 1, 1, 1, 3, 3, 1, 1, 1, 1, 1, 1, 1, 2,
                   1, 1, 1, 1, 1, 1, 3, 1, 1, 1, 1, 1, 3, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 6)
# rows of evogym/envs/sim_files/Climber-v1.json (7 x 70, top row first) as runs of (left wall, right wall, rows): F = FIXED,
# S = SOFT; "==" = a full FIXED row (the cap and the floor of the pipe)
CLIMBER_V1_WALLS = (("==", 1), ("FF", 4), ("SF", 5), ("FS", 4), ("FF", 2), ("SS", 3), ("FF", 4), ("SS", 2), ("FF", 3), ("SF", 3), ("FF", 1),
                    ("FS", 2), ("SS", 1), ("FF", 2), ("SF", 2), ("FF", 1), ("FS", 1), ("SS", 1), ("SF", 1), ("FF", 2), ("FS", 2), ("SF", 2),
                    ("FF", 3), ("FS", 3), ("FF", 2), ("SF", 2), ("FF", 2), ("FS", 1), ("FF", 7), ("==", 1))


@functools.lru_cache(maxsize=None)
def _hurdler_terrain() -> WorldSpec:
    w = WorldSpec(width=70, height=6)
    grid = np.zeros((6, 70), dtype=np.int64)
    for c, hgt in enumerate(HURDLER_HEIGHTS):
        grid[6 - hgt:, c] = VOXEL_TYPES["FIXED"]
    w.add_from_array("ground", grid, 0, 0)
    return _frozen(w)


def hurdler_world(robot: np.ndarray, connections: Optional[np.ndarray] = None) -> WorldSpec:
    """Hurdler-v0 (``sim_files/Hurdler-v0.json`` + ``evogym/envs/traverse.py:451-452``): fixed ``ground`` with barriers of 2 - 3 voxels
    and a 6-voxel end wall, the robot at voxel (2, 4)."""
    w = _shared(_hurdler_terrain())
    w.add_from_array("robot", robot, 2, 4, connections)
    return w


@functools.lru_cache(maxsize=None)
def _climber_v1_terrain() -> WorldSpec:
    w = WorldSpec(width=7, height=70)
    grid = np.zeros((70, 7), dtype=np.int64)
    t = {"F": VOXEL_TYPES["FIXED"], "S": VOXEL_TYPES["SOFT"]}
    r = 0
    for walls, n in CLIMBER_V1_WALLS:
        if walls == "==":
            grid[r:r + n, :] = VOXEL_TYPES["FIXED"]
        else:
            grid[r:r + n, 0], grid[r:r + n, 6] = t[walls[0]], t[walls[1]]
        r += n
    w.add_from_array("pipe", grid, 0, 0)
    return _frozen(w)


def climber_v1_world(robot: np.ndarray, connections: Optional[np.ndarray] = None) -> WorldSpec:
    """Climber-v1 (``sim_files/Climber-v1.json`` + ``evogym/envs/climb.py:139-140``): a 5-wide pipe whose walls alternate FIXED and
    SOFT stretches -- one object, so the whole pipe is simulated --, the robot at voxel (1, 1)."""
    w = _shared(_climber_v1_terrain())
    w.add_from_array("robot", robot, 1, 1, connections)
    return w


# terrain of evogym/envs/sim_files/CaveCrawler-v0.json (70 x 14, one object "terrain") and Climber-v2.json (27 x 42, "pipe"), top row
# first: F = FIXED, S = SOFT, . = empty
CAVE_CRAWLER_ROWS = (
    "....................F.................................................",
    "....................F.................................................",
    "....................F.................................................",
    "....................F.................................................",
    "....................F................................................F",
    "....................F................................................F",
    "....................FF............FFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFF",
    ".....................SSSSSSSSSSSSSS....SSSSSSSSSSSSSSS....FFF........F",
    ".......................SSSSSSSSSS.........SSSSSSSSSS.......FF........F",
    ".........................SSSSSS............................F.........F",
    ".....................................................................F",
    ".....................................................................F",
    "..............................................SSSSS..................F",
    "FFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFF",
)
CLIMBER_V2_ROWS = (
    ".......FFFFFFFFFFFFFFFFFFFF",
    "......FF..................F",
    ".....FF...................F",
    "....FF....................F",
    "...FF.....................F",
    "...F......................F",
    "...F.......FFFFFFFFFFFFFFFF",
    "..FF......FF...............",
    "..F......FF................",
    "..F.....FF.................",
    "..F.....F..................",
    ".FF.....F..................",
    ".F......F..................",
    ".F......F..................",
    ".F.....FF..................",
    ".F.....F...................",
    ".F.....F...................",
    ".F.....F...................",
    "FF.....F...................",
    "F......F...................",
    "F......F...................",
    "F......F...................",
    "F.....FF...................",
    "F.....F....................",
    "F.....F....................",
    "F.....F....................",
    "F.....F....................",
    "F.....F....................",
    "F.....F....................",
    "F.....F....................",
    "F.....F....................",
    "F.....F....................",
    "F.....F....................",
    "F.....F....................",
    "F.....F....................",
    "F.....F....................",
    "F.....F....................",
    "F.....F....................",
    "F.....F....................",
    "F.....F....................",
    "F.....F....................",
    "FFFFFFF....................",
)


def grid_from_rows(rows: Sequence[str]) -> np.ndarray:
    code = {".": 0, "R": VOXEL_TYPES["RIGID"], "S": VOXEL_TYPES["SOFT"], "H": VOXEL_TYPES["H_ACT"], "V": VOXEL_TYPES["V_ACT"], "F": VOXEL_TYPES["FIXED"]}
    return np.array([[code[ch] for ch in row] for row in rows], dtype=np.int64)


@functools.lru_cache(maxsize=None)
def _cave_crawler_terrain() -> WorldSpec:
    w = WorldSpec(width=70, height=14)
    w.add_from_array("terrain", grid_from_rows(CAVE_CRAWLER_ROWS), 0, 0)
    return _frozen(w)


def cave_crawler_world(robot: np.ndarray, connections: Optional[np.ndarray] = None) -> WorldSpec:
    """CaveCrawler-v0 (``sim_files/CaveCrawler-v0.json`` + ``evogym/envs/walk.py:240-241``): a floor, a fixed ceiling with SOFT
    stalactites, one object ``terrain`` (simulated: it has SOFT voxels); the robot at voxel (1, 2)."""
    w = _shared(_cave_crawler_terrain())
    w.add_from_array("robot", robot, 1, 2, connections)
    return w


@functools.lru_cache(maxsize=None)
def _climber_v2_terrain() -> WorldSpec:
    w = WorldSpec(width=27, height=42)
    w.add_from_array("pipe", grid_from_rows(CLIMBER_V2_ROWS), 0, 0)
    return _frozen(w)


def climber_v2_world(robot: np.ndarray, connections: Optional[np.ndarray] = None) -> WorldSpec:
    """Climber-v2 (``sim_files/Climber-v2.json`` + ``evogym/envs/climb.py:233-234``): a fixed ``pipe`` that bends to the right at the
    top; the robot at voxel (1, 1)."""
    w = _shared(_climber_v2_terrain())
    w.add_from_array("robot", robot, 1, 1, connections)
    return w


@functools.lru_cache(maxsize=None)
def _jumper_terrain() -> WorldSpec:
    w = WorldSpec(width=60, height=8)
    grid = np.zeros((8, 60), dtype=np.int64)
    grid[7, :] = VOXEL_TYPES["FIXED"]
    grid[:, 0] = VOXEL_TYPES["FIXED"]
    grid[:, 59] = VOXEL_TYPES["FIXED"]
    w.add_from_array("ground", grid, 0, 0)
    return _frozen(w)


def jumper_world(robot: np.ndarray, connections: Optional[np.ndarray] = None) -> WorldSpec:
    """Jumper-v0 (``sim_files/Jumper-v0.json`` + ``evogym/envs/jump.py:19-20``): a 60-wide fixed floor between two 8-voxel walls,
    the robot at voxel (32, 1)."""
    w = _shared(_jumper_terrain())
    w.add_from_array("robot", robot, 32, 1, connections)
    return w


BRIDGE_WALKER_PILLARS = (0, 5, 10, 18, 29, 46, 58, 59)     # columns of FIXED voxels from the floor row up to the deck


@functools.lru_cache(maxsize=None)
def _bridge_walker_terrain() -> WorldSpec:
    w = WorldSpec(width=60, height=5)
    grid = np.zeros((5, 60), dtype=np.int64)          # row 0 = top
    grid[4, :] = VOXEL_TYPES["FIXED"]
    grid[0:2, :] = VOXEL_TYPES["SOFT"]
    for c in BRIDGE_WALKER_PILLARS:
        grid[:, c] = VOXEL_TYPES["FIXED"]
    w.add_from_array("new_object_2", grid, 0, 0)
    return _frozen(w)


def bridge_walker_world(robot: np.ndarray, connections: Optional[np.ndarray] = None) -> WorldSpec:
    """BridgeWalker-v0 (``evogym/envs/sim_files/BridgeWalker-v0.json`` + ``evogym/envs/walk.py:152-153``): one 60 x 5 object
    ``new_object_2`` -- a FIXED floor row, FIXED pillars, and a two-voxel-thick SOFT deck hung between the pillars -- with the
    robot at voxel (2, 5).  FIXED and SOFT voxels share one object, so the whole terrain is simulated (SURVEY 7 "dead work")."""
    w = _shared(_bridge_walker_terrain())
    w.add_from_array("robot", robot, 2, 5, connections)
    return w



# env id -> (terrain without the robot, built once per process; robot position).  Climber-v0's pipe depends on the robot's width.
_TERRAINS = {
    "Walker-v0": (lambda robot: _walker_terrain(100), (1, 1)),
    "Pusher-v0": (lambda robot: _pusher_terrain(), (1, 1)),
    "Thrower-v0": (lambda robot: _thrower_terrain(), (1, 1)),
    "Carrier-v0": (lambda robot: _carrier_terrain(), (1, 1)),
    "Carrier-v1": (lambda robot: _carrier_v1_terrain(), (1, 4)),
    "Lifter-v0": (lambda robot: _lifter_terrain(), (2, 3)),
    "BeamToppler-v0": (lambda robot: _beam_toppler_terrain(), (1, 1)),
    "BeamSlider-v0": (lambda robot: _beam_slider_terrain(), (1, 1)),
    "Balancer-v0": (lambda robot: _balancer_terrain(2), (15, 3)),
    "Balancer-v1": (lambda robot: _balancer_terrain(3), (10, 1)),
    "Pusher-v1": (lambda robot: _pusher_v1_terrain(), (13, 1)),
    "Climber-v0": (lambda robot: _climber_terrain(max(7, int(np.asarray(robot).shape[1]) + 2)), (1, 1)),
    "Climber-v1": (lambda robot: _climber_v1_terrain(), (1, 1)),
    "Climber-v2": (lambda robot: _climber_v2_terrain(), (1, 1)),
    "ObstacleTraverser-v0": (lambda robot: _obstacle_traverser_v0_terrain(), (2, 1)),
    "ObstacleTraverser-v1": (lambda robot: _obstacle_traverser_terrain(), (2, 4)),
    "UpStepper-v0": (lambda robot: _up_stepper_terrain(), (1, 1)),
    "DownStepper-v0": (lambda robot: _down_stepper_terrain(), (1, 11)),
    "Hurdler-v0": (lambda robot: _hurdler_terrain(), (2, 4)),
    "BridgeWalker-v0": (lambda robot: _bridge_walker_terrain(), (2, 5)),
    "CaveCrawler-v0": (lambda robot: _cave_crawler_terrain(), (1, 2)),
    "Jumper-v0": (lambda robot: _jumper_terrain(), (32, 1)),
    "Flipper-v0": (lambda robot: _flipper_terrain(), (60, 1)),
    "WingspanMazimizer-v0": (lambda robot: _shape_change_terrain(), (7, 1)),
    "HeightMaximizer-v0": (lambda robot: _shape_change_terrain(), (7, 1)),
    "AreaMaximizer-v0": (lambda robot: _shape_change_terrain(), (7, 1)),
    "AreaMinimizer-v0": (lambda robot: _shape_change_terrain(), (7, 1)),
    "BidirectionalWalker-v0": (lambda robot: _bidirectional_walker_terrain(), (33, 1)),
    "Traverser-v0": (lambda robot: _traverser_terrain(), (1, 9)),
    "Catcher-v0": (lambda robot: _catcher_terrain(*catcher_offsets()), (22, 1)),      # a new draw per environment
    "GapJumper-v0": (lambda robot: _gap_jumper_terrain(), (1, 6)),
    "PlatformJumper-v0": (lambda robot: _platform_jumper_terrain(), (1, 6)),
}


def terrain_for(env_id: str, robot: np.ndarray) -> Tuple[WorldSpec, int, int]:
    """(shared read-only terrain world, robot x, robot y) of a built-in environment: what its ``*_world`` builder puts around ``robot``"""
    make, (rx, ry) = _TERRAINS[env_id]
    return make(robot), rx, ry
