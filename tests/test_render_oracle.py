"""The CUDA rasteriser (csrc/render.cu) against the CPU restatement of Interface::render's drawing rules
(oracle/render_oracle.py), pixel for pixel.

PARITY UNPINNED for this item: the reference draws through OpenGL, which is unavailable on this box, so neither side was
compared with an image the reference produced.  What is pinned: (CPU) the restated colour ramps equal the reference's own
100-row tables (Colors.h) as 8-bit colours; (GPU) the kernel's image equals the oracle's on fixture worlds in rest and
deformed states, at three resolutions / cameras and with every flag, except for TIE pixels: pixel centres so close to a
primitive's boundary that float32 (kernel) and float64 (oracle) edge functions disagree in sign.  Those are bounded (<= 0.3 %
of the image) and must sit on a boundary (one of the 8 neighbours of the mismatching pixel has the other side's colour)."""
import os
import re

import numpy as np
import pytest

from helpers import Golden
from oracle import render_oracle as RO

COLORS_H = "/root/reference/evogym/simulator/SimulatorCPP/Colors.h"


@pytest.mark.skipif(not os.path.exists(COLORS_H), reason="reference checkout not present")
def test_colour_ramps_equal_colors_h():
    text = open(COLORS_H).read()
    tables = {}
    for name in ("orange", "blue"):
        body = text[text.index(f"double {name}[100][3]"):]
        body = body[body.index("{") + 1:body.index("};")]
        rows = re.findall(r"\{\s*([0-9.eE+-]+)\s*,\s*([0-9.eE+-]+)\s*,\s*([0-9.eE+-]+)\s*\}", body)
        assert len(rows) == 100
        tables[name] = np.array(rows, dtype=np.float64)
    for name, anchors in (("orange", RO.ORANGES), ("blue", RO.BLUES)):
        mine = np.stack([RO.ramp_row(anchors, i) for i in range(100)])
        assert np.array_equal(mine, RO.to_u8(tables[name])), name


def _compare(img: np.ndarray, ref: np.ndarray):
    diff = (img != ref).any(axis=2)
    n = int(diff.sum())
    # every mismatching pixel must be a boundary pixel: the other image's colour occurs among its 8 neighbours in this image
    bad = 0
    h, w = diff.shape
    for r, c in zip(*np.nonzero(diff)):
        r0, r1, c0, c1 = max(0, r - 1), min(h, r + 2), max(0, c - 1), min(w, c + 2)
        if not (img[r0:r1, c0:c1].reshape(-1, 3) == ref[r, c]).all(axis=1).any():
            bad += 1
    return n, bad


@pytest.mark.gpu
@pytest.mark.parametrize("name,sub", [("walker_s0", None), ("walker_dense", 45), ("pusher_s3", 59), ("thrower_s4", 30), ("bridge_s6", 59),
                                      ("obstacle_10x10", 29), ("squash_s14", 20)])
def test_cuda_image_equals_rule_oracle(name, sub):
    from modularevogym_b200 import capi
    g = Golden(name)
    B = capi.Batch(0)
    B.add_world(g.world)
    B.finalize()
    pos = g["pos0"] if sub is None else g["sub_pos"][sub]
    if sub is not None:
        z = np.zeros_like(pos)
        B.set_state(0, pos, z, pos, z)
    pos = B.get_state(0)[0]                      # pruned all-FIXED terrain keeps its construction positions
    rb = g.obj_info[[i for i in range(g.n_obj) if g.obj_info[i, 0]][0]]
    centre = pos[rb[1]:rb[2] + 1].mean(axis=0)
    worst = 0.0
    for cam_size, res, flags in (((1.6, 0.9), (320, 180), {}), ((0.8, 0.8), (257, 263), {"hide_grid": True}),
                                 ((3.1, 1.3), (200, 96), {"hide_background": True, "hide_edges": True}),
                                 ((1.0, 1.0), (128, 128), {"hide_boxels": True})):
        img = B.render(0, centre, cam_size, res, **flags)
        ref = RO.render(g.world, pos, centre, cam_size, res, **flags)
        n, bad = _compare(img, ref)
        frac = n / (res[0] * res[1])
        worst = max(worst, frac)
        assert bad == 0, (name, cam_size, res, flags, "mismatch away from a primitive boundary", bad)
        assert frac <= 0.003, (name, cam_size, res, flags, frac)
    print(f"\n[render] {name}: worst tie-pixel share {worst:.5f}")
    B.close()
