"""Generates tests/golden/worlds_topology.npz: for every world / object file the reference ships
(``/root/reference/evogym/envs/sim_files/*.json``, 33 files) the topology the REFERENCE builds from it.

TEST INFRASTRUCTURE; runs only where ``/root/reference`` exists.  Each file goes through the reference's own, unmodified
Python loader and load order -- ``evogym.world.EvoWorld.from_json`` (``evogym/world.py:25-72``) and ``EvoSim.__init__`` /
``_init_world_items`` (``evogym/sim.py:29-98``: robots first, then objects, each alphabetical, ``np.flipud``, re-indexed
connections) -- into the compiled reference engine (``oracle/_ref``), and the resulting arrays are stored next to this repo's
own parse of the same file (``WorldSpec.from_json``) so that the loaders can be checked on a box without the reference.

    python tests/golden/make_world_golden.py
"""
from __future__ import annotations

import glob
import importlib.util
import json
import os
import sys
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
SIM_FILES = "/root/reference/evogym/envs/sim_files"


def main():
    warnings.filterwarnings("ignore")
    spec = importlib.util.spec_from_file_location("make_env_golden", os.path.join(HERE, "make_env_golden.py"))
    meg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(meg)
    meg.install_stubs()                      # evogym.simulator_cpp -> the compiled reference (oracle/ref.py), gym / np.float shims
    from evogym import EvoSim, EvoWorld
    from modularevogym_b200.world import WorldSpec

    out = {}
    names = []
    for path in sorted(glob.glob(os.path.join(SIM_FILES, "*.json"))):
        name = os.path.splitext(os.path.basename(path))[0]
        world = EvoWorld.from_json(path)                     # the reference's parser
        sim = EvoSim(world)                                  # the reference's load order, into the reference's engine
        r = sim._r
        ed = r.edges()
        mass, fixed = r.point_consts()
        infos, boxels = [], []
        for o in range(r.num_objects):
            i = r.object_info(o)
            infos.append([i[k] for k in ("is_robot", "min_point", "max_point", "n_boxels", "grid_w", "grid_h", "n_surface")])
            boxels.append(r.object_boxels(o))
        mine = WorldSpec.from_json(path)
        out[name + "/world_json"] = np.array(json.dumps(mine.to_dict()))
        out[name + "/pos0"] = r.get_state()[0]
        out[name + "/fixed"] = fixed.astype(np.uint8)
        out[name + "/ea"] = ed["a"].astype(np.int16)
        out[name + "/eb"] = ed["b"].astype(np.int16)
        out[name + "/ek"] = ed["k"]
        out[name + "/el0"] = ed["l0"]
        out[name + "/on_surface"] = ed["on_surface"].astype(np.uint8)
        out[name + "/obj_info"] = np.asarray(infos, dtype=np.int32)
        out[name + "/obj_names"] = np.array(json.dumps(list(r.object_names)))
        out[name + "/boxels"] = np.concatenate(boxels, axis=0).astype(np.int16)
        names.append(name)
        print(f"{name:28s} objects {r.num_objects}  points {r.num_points:5d}  edges {r.num_edges:5d}  mass==1: {bool((mass == 1.0).all())}")
    out["names"] = np.array(json.dumps(names))
    np.savez_compressed(os.path.join(HERE, "worlds_topology.npz"), **out)
    print(len(names), "worlds ->", os.path.join(HERE, "worlds_topology.npz"), os.path.getsize(os.path.join(HERE, "worlds_topology.npz")), "bytes")


if __name__ == "__main__":
    main()
