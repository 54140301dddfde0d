"""CPU: the C-ABI library loads without a GPU, exports every symbol of include/mevo_b200.h, fails
loudly when asked for a device, and its host-side world builder (csrc/topology.cpp) reproduces the
reference's topology arrays exactly (fixtures from the reference engine)."""
import ctypes

import numpy as np
import pytest

from helpers import Golden, golden_names
from modularevogym_b200 import capi

NAMES = golden_names()


def test_library_exports_every_declared_symbol():
    L = capi.load()
    declared = capi.header_symbols()
    assert len(declared) >= 50
    raw = ctypes.CDLL(capi.LIB_PATH)
    for sym in declared:
        assert hasattr(raw, sym), f"{sym} declared in include/mevo_b200.h but not exported"
    assert sorted(capi.SIGNATURES) == declared
    assert L.mevo_abi_version() == 1


def test_no_cpu_fallback(cuda_available):
    if cuda_available:
        pytest.skip("a CUDA device is visible")
    L = capi.load()
    assert not L.mevo_batch_create(0)
    msg = L.mevo_last_error().decode()
    assert "no CPU fallback" in msg
    with pytest.raises(capi.MevoError):
        capi.Batch(0)


@pytest.mark.parametrize("name", NAMES)
def test_host_builder_matches_reference(name):
    g = Golden(name)
    h = capi.HostWorld(g.world)
    assert h.num_points == g.P and h.num_edges == g.E and h.num_objects == g.n_obj
    t = h.topology()
    for mine, ref in (("pos0", "pos0"), ("mass", "mass"), ("fixed", "fixed"), ("a", "ea"), ("b", "eb"),
                      ("k", "ek"), ("l0", "el0"), ("on_surface", "on_surface")):
        assert np.array_equal(t[mine], g[ref]), ref
    for o in range(g.n_obj):
        info = h.object_info(o)
        want = dict(zip(("is_robot", "min_point", "max_point", "n_boxels", "grid_w", "grid_h", "n_surface"),
                        (int(v) for v in g.obj_info[o])))
        for k, v in want.items():
            assert info[k] == v, (o, k)
        assert np.array_equal(h.object_boxels(o), g[f"boxels_{o}"])
        assert np.array_equal(h.object_surface_boxels(o), g[f"surface_{o}"])
    assert np.array_equal(h.actuator_indices("robot"), g["act_idx"])


def test_host_builder_refusals(capfd):
    """Refusals follow the reference: message on stdout, world unchanged, no exception
    (ObjectCreator.cpp:173-191,512-541)."""
    g = Golden("walker_s0")
    h = capi.HostWorld(g.world)
    P = h.num_points
    robot = g.world.objects["robot"]
    # overlapping an existing object
    assert not h.add_object(robot.grid, robot.connections, "robot2", robot.x, robot.y, True)
    # duplicate name
    assert not h.add_object(robot.grid, robot.connections, "robot", robot.x + 10, robot.y, True)
    # two components
    split = np.array([[3, 0, 3]])
    assert not h.add_object(split, np.zeros((2, 0)), "split", 20, 1, True)
    assert h.num_points == P
    assert "Error" in capfd.readouterr().out
