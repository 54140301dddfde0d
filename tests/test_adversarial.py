"""Adversarial cases for the two deliberate shortcuts of the contact path.

(b) `segments_intersect` decides `0 <= n / det <= 1` without dividing.  tools/check_segment_test.c proves by construction
    and checks on boundary operands (n = det +- k ulp, +-0, denormals, both signs) that this equals the reference's
    rounded-quotient test (PhysicsEngine.cpp:1094-1121) for every |det| < 1 -- the only exception, a quotient that underflows
    to -0.0, needs |det| >= 1, i.e. voxel edges over three metres long.  Not a deviation in any reachable state.

(a) The box prefilter: a corner outside a boxel's bounding box cannot be inside either of its triangles, so it never
    reaches the resolver.  `Boxel::point_in_boxel` (Boxel.cpp:41-65) however also answers "inside" for ANY point on the line
    through TL and BR, because there `c = cross(BR - p, TL - p)` is exactly 0.0 and `c * a >= 0 && c * b >= 0` holds
    vacuously.  The GPU test below builds exactly that state -- a probe voxel whose corner lies on the extended diagonal
    of a block voxel, outside its box, with one of the corner's edges crossing the block -- and records what each engine
    does: the reference applies a contact to the outside corner, this engine does not; one ulp off the diagonal (c != 0)
    both agree again.  This is the single known difference in contact DETECTION; it needs an exactly zero FP64 cross
    product of three points that are not in lattice position.
"""
import os
import subprocess

import numpy as np
import pytest

from helpers import rel_err
from modularevogym_b200 import capi
from modularevogym_b200.world import WorldSpec

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_segment_crossing_predicate_equals_the_rounded_quotient_form(tmp_path):
    exe = str(tmp_path / "cst")
    subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", os.path.join(ROOT, "tools", "check_segment_test.c"), "-lm", "-o", exe])
    out = subprocess.run([exe, "20000000"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout
    assert "mismatches 0 " in out.stdout and out.stdout.rstrip().endswith("|det| < 1: 0"), out.stdout


def _two_voxel_world():
    w = WorldSpec(width=12, height=12)
    w.add_from_array("block", np.array([[2]]), 2, 2)        # one SOFT voxel each: simulated, so their points can be placed freely
    w.add_from_array("probe", np.array([[2]]), 6, 2)
    return w


def _placed_state(B, env, eps_ulps: int):
    """block = axis-aligned square of side 0.125 at (1, 1) (dyadic coordinates: every difference below is exact);
    probe: its BL corner on the block's extended TL-BR diagonal x + y = 2.125, outside the block's box; its TL corner
    inside the block, so its left edge crosses the block's bottom edge."""
    pos = B.get_state(env)[0].copy()
    info = {B.object_info(env, o)["min_point"]: o for o in range(B.num_objects(env))}
    names = {}
    for o in range(B.num_objects(env)):
        names[o] = B.object_boxels(env, o)[0]               # cell_type, grid_index, TL, TR, BR, BL, edges...
    blk, prb = (0, 1) if B.object_index(env, "block") == 0 else (1, 0)
    tl, tr, br, bl = (int(v) for v in names[blk][2:6])
    pos[tl], pos[tr], pos[br], pos[bl] = (1.0, 1.125), (1.125, 1.125), (1.125, 1.0), (1.0, 1.0)
    ptl, ptr, pbr, pbl = (int(v) for v in names[prb][2:6])
    x = 1.1875
    for _ in range(abs(eps_ulps)):
        x = np.nextafter(x, 2.0 if eps_ulps > 0 else 0.0)
    pos[pbl] = (x, 0.9375)                                  # on the diagonal when eps_ulps == 0
    pos[ptl] = (1.05, 1.06)                                 # inside the block
    pos[ptr] = (1.19, 1.07)
    pos[pbr] = (1.30, 0.95)
    return pos, pbl, ptl


@pytest.mark.gpu
def test_corner_exactly_on_extended_diagonal():
    from oracle import ref as oref
    if not oref.available():
        pytest.skip("oracle/_ref/libmevo_ref.so not built")
    w = _two_voxel_world()
    report = {}
    for eps in (0, 1, -1):
        B = capi.Batch(0)
        B.add_world(w)
        B.enable_force_dump()
        B.finalize()
        pos, p_out, p_in = _placed_state(B, 0, eps)
        zero = np.zeros_like(pos)
        B.set_state(0, pos, zero, pos, zero)
        B.step_substeps(1, 1, apply_actions=False)
        cf = B.read_collision_forces(0)
        r = oref.RefSim(w)
        r.set_state(pos, zero, pos, zero)
        r.substep()
        cf_ref = r.collision_forces()
        report[eps] = (cf.copy(), cf_ref.copy())
        # the corner that really is inside the block gets its contact in both engines
        assert np.abs(cf_ref[p_in]).sum() > 0 and rel_err(cf[p_in], cf_ref[p_in]) <= 1e-11
        if eps != 0:
            # one ulp off the line the cross product is non-zero, the reference's triangle test rejects the outside corner, and the
            # engines agree on every force (the outside corner still receives its share of the contacts on the probe's edges)
            assert rel_err(cf, cf_ref) <= 1e-11
            pg, vg = B.get_state(0)[:2]
            pr, vr = r.get_state()[:2]
            assert rel_err(pg, pr) <= 1e-14 and rel_err(vg, vr) <= 1e-11
        B.close()
    # exactly on the line: the reference resolves the outside corner too (point_in_boxel answers "inside" on the whole diagonal line)
    # -- a jump of its collision force against both neighbouring states; this engine's box prefilter drops the corner and stays
    # continuous
    ours0, ref0 = report[0][0][p_out], report[0][1][p_out]
    ours1, ref1 = report[1][0][p_out], report[1][1][p_out]
    assert np.abs(ref0 - ref1).max() > 1e3, "the reference no longer produces the spurious contact: update DESIGN section 5"
    assert np.abs(ours0 - ours1).max() <= 1e-6 * np.abs(ours1).max()
    assert rel_err(report[-1][0], report[1][0]) <= 1e-9
    print("\n[adversarial] collision force on the outside corner -- on the diagonal: engine", ours0, "reference", ref0,
          "| one ulp off: engine", ours1, "reference", ref1)
