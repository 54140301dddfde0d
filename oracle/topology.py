"""TEST INFRASTRUCTURE ONLY: plain-Python restatement of the reference's world builder.

Follows, step by step, ``ObjectCreator::read_object_from_array``
(``SimulatorCPP/ObjectCreator.cpp:193-641``), ``SimObject::compute_surface``
(``SimulatorCPP/SimObject.cpp:235-307``), the surface-edge orientation pass
(``ObjectCreator.cpp:603-632`` + ``Environment::swap_edge`` ``Environment.cpp:111-116``),
``Environment::set_surface_edge_color`` (``Environment.cpp:24-32``) and ``Robot::init``
(``Robot.cpp:9-22``) -- including the quirk that index 0 doubles as "unset" for point and edge
ids (``Boxel.cpp:6-14``, ``ObjectCreator.cpp:334-423``).

What is NOT reproduced, on purpose: the bounding-box tree (``SimObject::compute_bb_tree``) and the
``unordered_set`` iteration order of the surface lists.  The candidate SET of the dual-tree descent
equals "all leaf pairs whose AABBs overlap" (internal boxes are exact unions), so the oracle
enumerates leaves in ascending (object, boxel) order instead; only the order in which contact
forces are summed differs from the reference (ulp-level, covered by the stated tolerance).

Pure-Python loops: worlds are a few hundred voxels, this runs once per world.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List

import numpy as np

CELL_EMPTY, CELL_RIGID, CELL_SOFT, CELL_ACT_H, CELL_ACT_V, CELL_FIXED = range(6)
TOP, BOT, LEFT, RIGHT, TOP_LEFT, TOP_RIGHT, BOT_LEFT, BOT_RIGHT = range(8)

CELL = 0.1                       # ObjectCreator.h:36
MASS = 1.0                       # ObjectCreator.h:38
K_RIGID_MAIN = 300000000.0 / 3.5  # ObjectCreator.h:40-47
K_RIGID_DIAG = K_RIGID_MAIN / 2.0
K_ACT_MAIN = 300000000.0 / 5
K_ACT_DIAG = K_ACT_MAIN / 2.0
K_SOFT_MAIN = 300000000.0 / 6
K_SOFT_DIAG = K_SOFT_MAIN / 4.0

# boxel record columns (same as ref_driver.cpp::mref_object_boxels)
B_TYPE, B_GRID, B_TL, B_TR, B_BR, B_BL, B_ET, B_ER, B_EB, B_EL = range(10)
B_NEI = 10  # .. 17


@dataclass
class ObjTopo:
    name: str
    is_robot: bool
    min_point: int
    max_point: int
    grid_w: int
    grid_h: int
    boxels: np.ndarray            # (n, 18) int32
    surface_boxels: np.ndarray    # ascending boxel indices
    actuator_grid_indices: np.ndarray = field(default_factory=lambda: np.zeros(0, dtype=np.int32))


@dataclass
class Topology:
    pos0: np.ndarray              # (P, 2)
    mass: np.ndarray              # (P,)
    fixed: np.ndarray             # (P,) uint8
    ea: np.ndarray                # (E,) int32
    eb: np.ndarray
    ek: np.ndarray                # (E,) spring constants
    el0: np.ndarray               # (E,) init_length_eq
    on_surface: np.ndarray        # (E,) uint8
    objects: List[ObjTopo]

    @property
    def P(self) -> int:
        return int(self.pos0.shape[0])

    @property
    def E(self) -> int:
        return int(self.ea.shape[0])

    def object_index(self, name: str) -> int:
        for i, o in enumerate(self.objects):
            if o.name == name:
                return i
        raise KeyError(name)


class _Boxel:
    __slots__ = ("cell_type", "grid_index", "nei", "tl", "tr", "bl", "br", "et", "eb", "el", "er")

    def __init__(self, cell_type: int, grid_index: int):
        self.cell_type = cell_type
        self.grid_index = grid_index
        self.nei = [False] * 8
        self.tl = self.tr = self.bl = self.br = 0     # 0 == NULL == unset (Boxel.cpp:6-14)
        self.et = self.eb = self.el = self.er = 0


class WorldBuilder:
    """Accumulates objects in load order, like one reference ``Sim`` (Sim.cpp:26-99)."""

    def __init__(self, world_w: int, world_h: int):
        self.world_w, self.world_h = int(world_w), int(world_h)
        self.world_grid = [0] * (self.world_w * self.world_h)
        self.pos: List[List[float]] = []
        self.mass: List[float] = []
        self.fixed: List[bool] = []
        self.edges: List[List] = []   # [a, b, l0, k, on_surface]
        self.objects: List[ObjTopo] = []
        self.names: Dict[str, int] = {}

    def add_object(self, grid, connections, name: str, x, y, is_robot: bool) -> bool:
        grid = np.asarray(grid)
        gh, gw = grid.shape
        local = [int(v) for v in grid.ravel()]      # row 0 = top, like the flattened Eigen grid (Sim.cpp:57-66)
        conn = np.asarray(connections).reshape(2, -1)
        P0, E0 = len(self.pos), len(self.edges)
        sx, sy = x * CELL, y * CELL                 # ObjectCreator.cpp:197
        max_y = sy + gh * CELL                      # :201

        def gi(xx, yy):
            return yy * gw + xx

        # overlap / bounds check against the world grid (:204-214, is_valid_in_world_grid :173-191)
        for yy in range(gh):
            for xx in range(gw):
                if local[gi(xx, yy)] == 0:
                    continue
                wx, wy = int(xx + x), int(self.world_h - gh + yy - y)
                if wx < 0 or wx >= self.world_w or wy < 0 or wy >= self.world_h:
                    return False
                if self.world_grid[wy * self.world_w + wx] != 0:
                    return False

        cells = [_Boxel(local[i], i) for i in range(gw * gh)]

        # neighbour flags from the connection list (:244-271)
        for i in range(conn.shape[1]):
            b1, b2 = int(conn[0, i]), int(conn[1, i])
            if b1 > b2:
                b1, b2 = b2, b1
            if not (0 <= b1 < gw * gh and 0 <= b2 < gw * gh):
                continue   # the reference would index out of bounds (UB); the Python layer never sends this
            if abs(b1 - b2) == 1 and gw != 1:
                cells[b1].nei[RIGHT] = True
                cells[b2].nei[LEFT] = True
            if abs(b1 - b2) == gw:
                cells[b1].nei[BOT] = True
                cells[b2].nei[TOP] = True

        new_pos: List[List[float]] = []
        new_edges: List[List] = []

        def make_point(px, py):
            new_pos.append([px, py])
            return len(new_pos) + P0 - 1            # :76-85

        def make_edge(a, b, l0, k):
            new_edges.append([a, b, l0, k, 0])
            return len(new_edges) + E0 - 1          # :87-92

        # points and shared main edges, row-major scan (:275-430)
        for yy in range(gh):
            for xx in range(gw):
                cur = cells[gi(xx, yy)]
                if cur.cell_type == CELL_EMPTY:
                    continue
                top = cells[gi(xx, yy - 1)] if cur.nei[TOP] else None
                bot = cells[gi(xx, yy + 1)] if cur.nei[BOT] else None
                left = cells[gi(xx - 1, yy)] if cur.nei[LEFT] else None
                right = cells[gi(xx + 1, yy)] if cur.nei[RIGHT] else None
                tl = tr = bl = br = None
                if (top is not None and top.nei[LEFT]) or (left is not None and left.nei[TOP]):
                    tl = cells[gi(xx - 1, yy - 1)]
                if (top is not None and top.nei[RIGHT]) or (right is not None and right.nei[TOP]):
                    tr = cells[gi(xx + 1, yy - 1)]
                if (bot is not None and bot.nei[LEFT]) or (left is not None and left.nei[BOT]):
                    bl = cells[gi(xx - 1, yy + 1)]
                if (bot is not None and bot.nei[RIGHT]) or (right is not None and right.nei[BOT]):
                    br = cells[gi(xx + 1, yy + 1)]
                if tl is not None: cur.nei[TOP_LEFT] = True
                if tr is not None: cur.nei[TOP_RIGHT] = True
                if bl is not None: cur.nei[BOT_LEFT] = True
                if br is not None: cur.nei[BOT_RIGHT] = True

                # "!= 0" is the reference's "!= NULL": point/edge id 0 reads as unset (:334-393)
                if top is not None and top.bl != 0: cur.tl = top.bl
                if left is not None and left.tr != 0: cur.tl = left.tr
                if tl is not None and tl.br != 0: cur.tl = tl.br
                if cur.tl == 0:
                    cur.tl = make_point(sx + xx * CELL, max_y - yy * CELL)

                if top is not None and top.br != 0: cur.tr = top.br
                if right is not None and right.tl != 0: cur.tr = right.tl
                if tr is not None and tr.bl != 0: cur.tr = tr.bl
                if cur.tr == 0:
                    cur.tr = make_point(sx + (xx + 1) * CELL, max_y - yy * CELL)

                if bot is not None and bot.tl != 0: cur.bl = bot.tl
                if left is not None and left.br != 0: cur.bl = left.br
                if bl is not None and bl.tr != 0: cur.bl = bl.tr
                if cur.bl == 0:
                    cur.bl = make_point(sx + xx * CELL, max_y - (yy + 1) * CELL)

                if bot is not None and bot.tr != 0: cur.br = bot.tr
                if right is not None and right.bl != 0: cur.br = right.bl
                if br is not None and br.tl != 0: cur.br = br.tl
                if cur.br == 0:
                    cur.br = make_point(sx + (xx + 1) * CELL, max_y - (yy + 1) * CELL)

                if top is not None and top.eb != 0: cur.et = top.eb
                if cur.et == 0:
                    cur.et = make_edge(cur.tl, cur.tr, abs(CELL), K_RIGID_MAIN)
                # :405-406 reads `top->edge_top_index` under a `bot != NULL && bot->edge_top_index != NULL`
                # guard; `bot` lies later in the scan so its edge is still unset and the branch is dead.
                if bot is not None and bot.et != 0:
                    cur.eb = top.et if top is not None else 0
                if cur.eb == 0:
                    cur.eb = make_edge(cur.bl, cur.br, abs(CELL), K_RIGID_MAIN)
                if left is not None and left.er != 0: cur.el = left.er
                if cur.el == 0:
                    cur.el = make_edge(cur.tl, cur.bl, abs(CELL), K_RIGID_MAIN)
                if right is not None and right.el != 0: cur.er = right.el
                if cur.er == 0:
                    cur.er = make_edge(cur.tr, cur.br, abs(CELL), K_RIGID_MAIN)

        # per-material constants + diagonals (:432-494)
        diag = math.sqrt(CELL * CELL + CELL * CELL)
        for yy in range(gh):
            for xx in range(gw):
                cur = cells[gi(xx, yy)]
                t = cur.cell_type
                if t == CELL_EMPTY:
                    continue
                if t == CELL_RIGID or t == CELL_FIXED:
                    kd = K_RIGID_DIAG
                elif t == CELL_SOFT:
                    for e in (cur.et, cur.eb, cur.el, cur.er):
                        new_edges[e - E0][3] = K_SOFT_MAIN
                    kd = K_SOFT_DIAG
                elif t == CELL_ACT_H or t == CELL_ACT_V:
                    for e in (cur.et, cur.eb, cur.el, cur.er):
                        new_edges[e - E0][3] = K_ACT_MAIN
                    kd = K_ACT_DIAG
                else:
                    continue
                make_edge(cur.tr, cur.bl, diag, kd)
                make_edge(cur.tl, cur.br, diag, kd)

        new_fixed = [False] * len(new_pos)
        for c in cells:
            if c.cell_type == CELL_FIXED:
                for p in (c.tl, c.tr, c.bl, c.br):
                    new_fixed[p - P0] = True

        # one connected component only (:19-61, :512-532); flood fill over the 4 cardinal flags
        comp = [-1] * (gw * gh)
        cid = 0
        for yy in range(gh):
            for xx in range(gw):
                if comp[gi(xx, yy)] != -1 or cells[gi(xx, yy)].cell_type == CELL_EMPTY:
                    continue
                stack = [(xx, yy)]
                while stack:
                    cx, cy = stack.pop()
                    if comp[gi(cx, cy)] != -1 or cells[gi(cx, cy)].cell_type == CELL_EMPTY:
                        continue
                    comp[gi(cx, cy)] = cid
                    n = cells[gi(cx, cy)].nei
                    if n[TOP]: stack.append((cx, cy - 1))
                    if n[BOT]: stack.append((cx, cy + 1))
                    if n[LEFT]: stack.append((cx - 1, cy))
                    if n[RIGHT]: stack.append((cx + 1, cy))
                cid += 1
        if cid > 1:
            return False
        if name in self.names:
            return False
        if cid == 0:
            # no voxels: the reference registers the name but creates no object; nothing to simulate
            return False
        self.names[name] = len(self.objects)

        min_point = len(self.pos)
        self.pos.extend(new_pos)
        self.mass.extend([MASS] * len(new_pos))
        self.fixed.extend(new_fixed)
        self.edges.extend(new_edges)
        max_point = len(self.pos) - 1

        for yy in range(gh):
            for xx in range(gw):
                if local[gi(xx, yy)] != 0:
                    self.world_grid[int(self.world_h - gh + yy - y) * self.world_w + int(xx + x)] = local[gi(xx, yy)]

        obj_cells = [c for c in cells if c.cell_type != CELL_EMPTY]
        rec = np.zeros((len(obj_cells), 18), dtype=np.int32)
        for j, c in enumerate(obj_cells):
            rec[j, :10] = (c.cell_type, c.grid_index, c.tl, c.tr, c.br, c.bl, c.et, c.er, c.eb, c.el)
            rec[j, 10:] = [1 if f else 0 for f in c.nei]

        # surface (SimObject.cpp:235-307), orientation (ObjectCreator.cpp:603-632), isOnSurface flag
        surface_boxels = []
        for j, c in enumerate(obj_cells):
            dirs = []
            if not c.nei[TOP]: dirs.append((c.et, (0.0, 1.0)))
            if not c.nei[BOT]: dirs.append((c.eb, (0.0, -1.0)))
            if not c.nei[LEFT]: dirs.append((c.el, (-1.0, 0.0)))
            if not c.nei[RIGHT]: dirs.append((c.er, (1.0, 0.0)))
            if dirs:
                surface_boxels.append(j)
            for e, ref in dirs:
                ed = self.edges[e]
                ed[4] = 1
                pa, pb = self.pos[ed[0]], self.pos[ed[1]]
                dx, dy = pb[0] - pa[0], pb[1] - pa[1]
                nx, ny = -dy, dx
                if ref[0] * nx + ref[1] * ny < 0:
                    ed[0], ed[1] = ed[1], ed[0]

        acts = [c.grid_index for c in obj_cells if c.cell_type in (CELL_ACT_H, CELL_ACT_V)] if is_robot else []
        self.objects.append(ObjTopo(name, bool(is_robot), min_point, max_point, gw, gh, rec,
                                    np.asarray(surface_boxels, dtype=np.int32),
                                    np.asarray(acts, dtype=np.int32)))
        return True

    def finish(self) -> Topology:
        E = len(self.edges)
        return Topology(
            pos0=np.asarray(self.pos, dtype=np.float64).reshape(-1, 2),
            mass=np.asarray(self.mass, dtype=np.float64),
            fixed=np.asarray(self.fixed, dtype=np.uint8),
            ea=np.asarray([e[0] for e in self.edges], dtype=np.int32).reshape(E),
            eb=np.asarray([e[1] for e in self.edges], dtype=np.int32).reshape(E),
            ek=np.asarray([e[3] for e in self.edges], dtype=np.float64).reshape(E),
            el0=np.asarray([e[2] for e in self.edges], dtype=np.float64).reshape(E),
            on_surface=np.asarray([e[4] for e in self.edges], dtype=np.uint8).reshape(E),
            objects=self.objects,
        )


def build_topology(world) -> Topology:
    """``world`` is a ``modularevogym_b200.world.WorldSpec`` (or anything with ``width``, ``height``,
    ``load_order()``)."""
    b = WorldBuilder(world.width, world.height)
    for o in world.load_order():
        if not b.add_object(o.grid, o.connections, o.name, o.x, o.y, o.is_robot):
            raise ValueError(f"object {o.name} rejected")
    return b.finish()
