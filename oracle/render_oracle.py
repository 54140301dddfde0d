"""TEST INFRASTRUCTURE ONLY: CPU restatement (numpy, float64) of what ``Interface::render`` draws, for pixel comparison with the
CUDA rasteriser (``modularevogym_b200/csrc/render.cu``).  Only ``tests/`` may import this module.

The reference draws with fixed-function OpenGL (``SimulatorCPP/Interface.cpp``); the GL path itself is unavailable on this box
(no graphics stack), so PARITY IS UNPINNED for this item: nothing here was checked against an image the reference produced.
What this module pins is that the CUDA kernel implements the stated rules -- the rules are restated here independently, from
the reference's drawing code, in the order it issues its GL calls:

* ``Interface::render`` (``Interface.cpp:100-131``): clear to (244, 245, 247) / 255 (white with ``hide_background``; nothing with
  ``dont_clear``), then ``render_grid``, ``render_boxels``, ``render_edges``; ``glReadPixels`` as RGB8.
* ``Camera::world_to_camera`` (``Camera.cpp:66-70``): ``ndc = (world - pos) * (2 / size)``, y negated for image cameras; the
  viewport maps ndc to window coordinates ``(ndc + 1) * res / 2``; pixel (c, r) has its centre at (c + 0.5, r + 0.5).
* ``render_grid`` (``Interface.cpp:175-220``): GL_LINES at x = i * 0.1 for ``int((pos.x - size.x) * 10) - 1 <= i <=
  int((pos.x + size.x) * 10) + 1`` (full size on each side, as written there), the same in y, colour 0.8 grey.
* ``render_boxels`` (``Interface.cpp:411-517``): one GL_QUADS quad (TL, TR, BR, BL) per boxel, objects and boxels in load order;
  FIXED / RIGID 0.15 grey, SOFT 0.75 grey, actuators through ``setColorInterval`` (``:378-409``) with the mean length of the two
  driven edges over the rest length, into the 100-row ramps of ``Colors.h`` (horizontal: oranges, vertical: blues).
* ``render_edges`` (``Interface.cpp:137-173``): GL_LINES per surface edge, colour (0, 63, 92) / 255.

Coverage (GL leaves line rasterisation and shared-edge ties to the implementation; this is the rule both sides implement): a
pixel belongs to a quad when its centre lies in the closed triangle (TL, TR, BR) or (TL, BR, BL); to a one-pixel line when its
centre is within half a pixel of the segment.  Later primitives paint over earlier ones.
"""
from __future__ import annotations

import numpy as np

from .topology import build_topology

# ColorBrewer 9-class sequential anchors; Colors.h samples each ramp at 100 points, row i = ramp(i / 99)
ORANGES = np.array([[255, 245, 235], [254, 230, 206], [253, 208, 162], [253, 174, 107], [253, 141, 60], [241, 105, 19], [217, 72, 1],
                    [166, 54, 3], [127, 39, 4]], dtype=np.float64)
BLUES = np.array([[247, 251, 255], [222, 235, 247], [198, 219, 239], [158, 202, 225], [107, 174, 214], [66, 146, 198], [33, 113, 181],
                  [8, 81, 156], [8, 48, 107]], dtype=np.float64)


def to_u8(rgb) -> np.ndarray:
    """GL float colour -> unsigned byte: round(clamp(c, 0, 1) * 255)"""
    return np.floor(np.clip(np.asarray(rgb, dtype=np.float64), 0.0, 1.0) * 255.0 + 0.5).astype(np.uint8)


def ramp_row(anchors: np.ndarray, row: int) -> np.ndarray:
    t = min(99, max(0, int(row))) / 99.0 * 8.0
    k = min(int(t), 7)
    f = t - k
    return to_u8((anchors[k] * (1 - f) + anchors[k + 1] * f) / 255.0)


def actuator_colour(actuation: float, vertical: bool) -> np.ndarray:
    """setColorInterval (Interface.cpp:378-409): > 1 -> rows 0..50 by int((2 - a) * 50); else rows 50..99 by int((1 - a) * 50) + 50"""
    ramp = BLUES if vertical else ORANGES
    if actuation > 1:
        return ramp_row(ramp, min(50, max(0, int((2.0 - actuation) * 50.0))))
    return ramp_row(ramp, min(49, max(0, int((1.0 - actuation) * 50.0))) + 50)


def _tri_mask(ax, ay, bx, by, cx, cy, px, py):
    e0 = (bx - ax) * (py - ay) - (by - ay) * (px - ax)
    e1 = (cx - bx) * (py - by) - (cy - by) * (px - bx)
    e2 = (ax - cx) * (py - cy) - (ay - cy) * (px - cx)
    return ((e0 >= 0) & (e1 >= 0) & (e2 >= 0)) | ((e0 <= 0) & (e1 <= 0) & (e2 <= 0))


def render(world, pos: np.ndarray, cam_pos, cam_size, resolution, hide_background=False, hide_grid=False, hide_edges=False,
           hide_boxels=False) -> np.ndarray:
    """(h, w, 3) uint8 image, row 0 = top.  ``pos``: (P, 2) current positions in the reference's numbering; camera in engine units."""
    t = build_topology(world)
    w, h = int(resolution[0]), int(resolution[1])
    cx, cy, cw, ch = float(cam_pos[0]), float(cam_pos[1]), float(cam_size[0]), float(cam_size[1])
    img = np.empty((h, w, 3), dtype=np.uint8)
    img[:] = to_u8([1, 1, 1]) if hide_background else to_u8([244 / 255, 245 / 255, 247 / 255])
    px, py = np.meshgrid(np.arange(w) + 0.5, np.arange(h) + 0.5)

    def X(x):
        return ((x - cx) * (2.0 / cw) + 1.0) * w * 0.5

    def Y(y):
        return (-(y - cy) * (2.0 / ch) + 1.0) * h * 0.5

    if not hide_grid:
        grey = to_u8([0.8, 0.8, 0.8])
        i0, i1 = int((cx - cw) * 10.0) - 1, int((cx + cw) * 10.0) + 1
        j0, j1 = int((cy - ch) * 10.0) - 1, int((cy + ch) * 10.0) + 1
        y_lo, y_hi = sorted((Y(j0 * 0.1), Y(j1 * 0.1)))
        x_lo, x_hi = sorted((X(i0 * 0.1), X(i1 * 0.1)))
        for i in range(i0, i1 + 1):
            xs = X(i * 0.1)
            if -1 <= xs <= w + 1:
                img[(np.abs(px - xs) <= 0.5) & (py >= y_lo - 0.5) & (py <= y_hi + 0.5)] = grey
        for j in range(j0, j1 + 1):
            ys = Y(j * 0.1)
            if -1 <= ys <= h + 1:
                img[(np.abs(py - ys) <= 0.5) & (px >= x_lo - 0.5) & (px <= x_hi + 0.5)] = grey

    def dist(a, b):
        return float(np.hypot(pos[a, 0] - pos[b, 0], pos[a, 1] - pos[b, 1]))

    if not hide_boxels:
        for ob in t.objects:
            for row in ob.boxels:
                cell, tl, tr, br, bl = int(row[0]), int(row[2]), int(row[3]), int(row[4]), int(row[5])
                e_top, e_right, e_bot, e_left = (int(v) for v in row[6:10])
                if cell == 2:
                    rgb = to_u8([0.75, 0.75, 0.75])
                elif cell in (3, 4):
                    vert = cell == 4
                    e1, e2 = (e_left, e_right) if vert else (e_top, e_bot)
                    avg = (dist(t.ea[e1], t.eb[e1]) + dist(t.ea[e2], t.eb[e2])) / 2.0
                    rgb = actuator_colour(avg / t.el0[e1], vert)
                else:
                    rgb = to_u8([0.15, 0.15, 0.15])
                qx = [X(pos[k, 0]) for k in (tl, tr, br, bl)]
                qy = [Y(pos[k, 1]) for k in (tl, tr, br, bl)]
                c0, c1 = max(0, int(np.floor(min(qx))) - 1), min(w, int(np.ceil(max(qx))) + 1)
                r0, r1 = max(0, int(np.floor(min(qy))) - 1), min(h, int(np.ceil(max(qy))) + 1)
                if c0 >= c1 or r0 >= r1:
                    continue
                sx, sy = px[r0:r1, c0:c1], py[r0:r1, c0:c1]
                m = _tri_mask(qx[0], qy[0], qx[1], qy[1], qx[2], qy[2], sx, sy) | _tri_mask(qx[0], qy[0], qx[2], qy[2], qx[3], qy[3], sx, sy)
                img[r0:r1, c0:c1][m] = rgb
    if not hide_edges:
        rgb = to_u8([0.0, 63 / 255, 92 / 255])
        for e in np.nonzero(t.on_surface)[0]:
            x0, y0, x1, y1 = X(pos[t.ea[e], 0]), Y(pos[t.ea[e], 1]), X(pos[t.eb[e], 0]), Y(pos[t.eb[e], 1])
            c0, c1 = max(0, int(np.floor(min(x0, x1))) - 1), min(w, int(np.ceil(max(x0, x1))) + 2)
            r0, r1 = max(0, int(np.floor(min(y0, y1))) - 1), min(h, int(np.ceil(max(y0, y1))) + 2)
            if c0 >= c1 or r0 >= r1:
                continue
            sx, sy = px[r0:r1, c0:c1], py[r0:r1, c0:c1]
            dx, dy = x1 - x0, y1 - y0
            len2 = dx * dx + dy * dy
            tt = np.clip(((sx - x0) * dx + (sy - y0) * dy) / len2, 0.0, 1.0) if len2 > 0 else np.zeros_like(sx)
            ex, ey = sx - (x0 + tt * dx), sy - (y0 + tt * dy)
            img[r0:r1, c0:c1][ex * ex + ey * ey <= 0.25] = rgb
    return img
