"""Generates the committed known-answer fixtures in ``tests/golden/*.npz``.

TEST INFRASTRUCTURE.  Runs only where ``/root/reference`` exists (this container): every number in
the fixtures comes from the UNMODIFIED reference engine compiled by ``oracle/Makefile`` into
``oracle/_ref/libmevo_ref.so`` (``oracle/ref.py``) and driven through its own public entry points
(``Sim::read_*_from_array``, ``Sim::set_action``, ``Environment::step``, ``Sim::step``).  Terrain
comes from the reference's world files ``evogym/envs/sim_files/*.json`` and robots are placed
where the reference's envs place them (walk.py:19-20, manipulate.py:323-324,524-525,
traverse.py:355-356, climb.py:45-46).  The fixtures travel to the GPU box; the reference does not.

    python tests/golden/make_golden.py            # regenerate everything

Fixture layout: see ``record()``.
"""
from __future__ import annotations

import json
import os
import random
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from modularevogym_b200.world import WorldSpec, sample_large_robot, sample_robot  # noqa: E402
from oracle.ref import RefSim  # noqa: E402

SIM_FILES = "/root/reference/evogym/envs/sim_files"
N_SUB = 60          # substeps recorded one by one (2 env-steps)
N_STEPS = 8         # env-steps recorded through Sim::step


def seeded_robot(seed: int, shape=(5, 5)):
    random.seed(seed)
    return sample_robot(shape) if shape[0] * shape[1] <= 36 else sample_large_robot(shape)


def terrain(name: str) -> WorldSpec:
    return WorldSpec.from_json(os.path.join(SIM_FILES, name + ".json"))


def world_with_robot(terrain_name, robot, conn, x, y) -> WorldSpec:
    w = terrain(terrain_name) if terrain_name else WorldSpec(width=30, height=30)
    w.add_from_array("robot", robot, x, y, conn)
    return w


def actions_for(seed: int, n_steps: int, n_act: int) -> np.ndarray:
    """a_t ~ U(-1,1) from default_rng(1234+seed), +1, clipped to [0.6, 1.6]
    (evogym/envs/base.py:42,389)."""
    rng = np.random.default_rng(1234 + seed)
    return np.clip(rng.uniform(-1, 1, (n_steps, n_act)) + 1.0, 0.6, 1.6)


# ------------------------------------------------------------------------------------------------
# state perturbations applied at t = 0 (they steer the run into the rarely taken branches)

def _robot_range(ref: RefSim):
    for o in range(ref.num_objects):
        info = ref.object_info(o)
        if info["is_robot"]:
            return info["min_point"], info["max_point"] + 1
    raise ValueError("no robot")


def perturb_penetrate(ref: RefSim, rng):
    """push the robot 0.03 into the terrain and give it a velocity: contacts with damping + friction (a-6)"""
    pos, vel, _, _ = ref.get_state()
    a, b = _robot_range(ref)
    pos[a:b, 1] -= 0.03
    vel[a:b] = rng.uniform(-0.5, 0.5, (b - a, 2))
    return pos, vel


def perturb_penetrate_xy(ref: RefSim, rng):
    """as above, and 0.03 sideways into the neighbouring object (robot-vs-object contacts)"""
    pos, vel = perturb_penetrate(ref, rng)
    a, b = _robot_range(ref)
    pos[a:b, 0] += 0.03
    return pos, vel


def _scale_robot(ref: RefSim, s: float):
    pos, vel, _, _ = ref.get_state()
    a, b = _robot_range(ref)
    c = pos[a:b].mean(axis=0)
    lowest = pos[a:b, 1].min()
    pos[a:b] = c + (pos[a:b] - c) * s
    pos[a:b, 1] += lowest - pos[a:b, 1].min()      # keep it resting where it was
    return pos, vel


def perturb_stretch(ref: RefSim, rng):
    """strain 1.45 > 1.25: the over-stretch limiter of resolve_edge_constraints (a-10)"""
    return _scale_robot(ref, 1.45)


def perturb_squash(ref: RefSim, rng):
    """strain 0.6 < 0.75: the under-stretch limiter (a-10)"""
    return _scale_robot(ref, 0.6)


def perturb_collapse(ref: RefSim, rng):
    """all voxels on top of each other: trips is_any_robot_self_colliding (a-12)"""
    return _scale_robot(ref, 0.04)


# ------------------------------------------------------------------------------------------------

def record(name: str, world: WorldSpec, seed: int, perturb=None, n_sub=N_SUB, n_steps=N_STEPS) -> dict:
    """One fixture.  Arrays (P = points, E = edges, all in the reference's global numbering):

    world_json                      the WorldSpec, JSON
    pos0,mass,fixed                 points at construction
    ea,eb,ek,el0,on_surface         edges (after the surface-orientation pass)
    obj_info (n_obj,7)              is_robot,min_point,max_point,n_boxels,grid_w,grid_h,n_surface
    boxels_<o> (n,18), surface_<o>  per object (surface list sorted ascending)
    act_idx, actions (n_steps,A)    actuator grid indices (Robot::init order) and per-step action values
    init_pos, init_vel              state the runs start from (pos_last = pos, vel_true = vel)
    sub_pos, sub_vel, sub_vt, sub_cf (n_sub,P,2)   state + collision_forces after each Environment::step
    sub_unstable (n_sub,)           return value of each Environment::step
    cand_<k> (n,4)                  leaf-leaf candidates (obj_a,boxel_a,obj_b,boxel_b) for the positions
                                    BEFORE substep k, reference descent order; k in cand_at
    step_pos, step_vel (n_steps+1,P,2), step_unstable, step_time, step_orient (n_steps,)   the Sim::step run
    """
    rng = np.random.default_rng(99 + seed)
    out = {"world_json": np.array(json.dumps(world.to_dict()))}

    ref = RefSim(world)
    P, E = ref.num_points, ref.num_edges
    ed = ref.edges()
    mass, fixed = ref.point_consts()
    out.update(pos0=ref.get_state()[0], mass=mass, fixed=fixed, ea=ed["a"], eb=ed["b"], ek=ed["k"],
               el0=ed["l0"], on_surface=ed["on_surface"])
    infos = []
    for o in range(ref.num_objects):
        i = ref.object_info(o)
        infos.append([i[k] for k in ("is_robot", "min_point", "max_point", "n_boxels", "grid_w", "grid_h", "n_surface")])
        out[f"boxels_{o}"] = ref.object_boxels(o)
        out[f"surface_{o}"] = np.sort(ref.object_surface_boxels(o))
    out["obj_info"] = np.asarray(infos, dtype=np.int32)
    act_idx = ref.actuator_indices("robot")
    actions = actions_for(seed, n_steps, len(act_idx))
    out.update(act_idx=act_idx, actions=actions)

    if perturb is not None:
        pos, vel = perturb(ref, rng)
    else:
        pos, vel, _, _ = ref.get_state()
    out.update(init_pos=pos.copy(), init_vel=vel.copy())

    # (1) substep by substep through Environment::step
    ref.set_state(pos, vel, pos, vel)
    cand_at = sorted(k for k in set([0, 1, 15, 29, 30, n_sub - 1]) if k < n_sub)
    sub = {k: np.zeros((n_sub, P, 2)) for k in ("pos", "vel", "vt", "cf")}
    sub_unstable = np.zeros(n_sub, dtype=np.uint8)
    for k in range(n_sub):
        if k % 30 == 0:
            ref.set_action_values("robot", actions[k // 30])
        if k in cand_at:
            out[f"cand_{k}"] = ref.candidates()
        sub_unstable[k] = ref.substep()
        p, v, _, vt = ref.get_state()
        sub["pos"][k], sub["vel"][k], sub["vt"][k] = p, v, vt
        sub["cf"][k] = ref.collision_forces()
    out.update(sub_pos=sub["pos"], sub_vel=sub["vel"], sub_vt=sub["vt"], sub_cf=sub["cf"],
               sub_unstable=sub_unstable, cand_at=np.asarray(cand_at, dtype=np.int32))

    # (2) step by step through Sim::step on a fresh Sim (snapshots, sim_time, unstable abort)
    ref2 = RefSim(world)
    if perturb is not None:
        ref2.set_state(pos, vel, pos, vel)
    step_pos = np.zeros((n_steps + 1, P, 2))
    step_vel = np.zeros((n_steps + 1, P, 2))
    step_unstable = np.zeros(n_steps, dtype=np.uint8)
    step_time = np.zeros(n_steps, dtype=np.int32)
    step_orient = np.zeros(n_steps)          # Sim::object_orientation_at_time(get_time(), "robot") after each step
    step_pos[0], step_vel[0] = pos, vel
    for s in range(n_steps):
        ref2.set_action_values("robot", actions[s])
        step_unstable[s] = ref2.step()
        step_time[s] = ref2.get_time()
        p, v, _, _ = ref2.get_state()
        step_pos[s + 1], step_vel[s + 1] = p, v
        step_orient[s] = ref2.object_orientation_at_time(ref2.get_time(), "robot")
    out.update(step_pos=step_pos, step_vel=step_vel, step_unstable=step_unstable, step_time=step_time, step_orient=step_orient)
    n_contacts = int((np.abs(sub["cf"]).sum(axis=(1, 2)) > 0).sum())
    print(f"{name:24s} P={P:4d} E={E:5d} objs={ref.num_objects} A={len(act_idx):3d} "
          f"substeps-with-contact={n_contacts:3d}/{n_sub} unstable(sub)={int(sub_unstable.sum())} "
          f"unstable(step)={int(step_unstable.sum())}")
    return out


def dense_robot():
    """fully dense 5x5 (the P = 238 / E = 611 configuration of SURVEY 8) with every voxel type"""
    g = np.array([[3, 3, 2, 4, 4],
                  [3, 1, 2, 1, 4],
                  [2, 2, 2, 2, 2],
                  [4, 1, 2, 1, 3],
                  [4, 4, 2, 3, 3]])
    return g, None


def cases():
    out = []
    for sd in (0, 1, 2):
        r, c = seeded_robot(sd)
        out.append((f"walker_s{sd}", world_with_robot("Walker-v0", r, c, 1, 1), sd, None))
    r, c = dense_robot()
    out.append(("walker_dense", world_with_robot("Walker-v0", r, c, 1, 1), 10, None))
    r, c = seeded_robot(3)
    out.append(("pusher_s3", world_with_robot("Pusher-v0", r, c, 1, 1), 3, None))
    r, c = seeded_robot(4)
    out.append(("thrower_s4", world_with_robot("Thrower-v0", r, c, 1, 1), 4, None))
    r, c = seeded_robot(5)
    out.append(("climber_s5", world_with_robot("Climber-v0", r, c, 1, 1), 5, None))
    r, c = seeded_robot(6)
    out.append(("bridge_s6", world_with_robot("BridgeWalker-v0", r, c, 2, 5), 6, None))
    r, c = seeded_robot(7, (10, 10))
    out.append(("obstacle_10x10", world_with_robot("ObstacleTraverser-v1", r, c, 2, 4), 7, None))
    r, c = seeded_robot(8, (20, 20))
    out.append(("obstacle_20x20", world_with_robot("ObstacleTraverser-v1", r, c, 2, 4), 8, None))
    r, c = seeded_robot(9)
    out.append(("freefall_s9", world_with_robot(None, r, c, 3, 0), 9, None))       # analytic ground y = 0 only
    r, c = seeded_robot(11)
    out.append(("penetrate_s11", world_with_robot("Walker-v0", r, c, 1, 1), 11, perturb_penetrate))
    r, c = seeded_robot(12)
    out.append(("pusher_penetrate_s12", world_with_robot("Pusher-v0", r, c, 3, 1), 12, perturb_penetrate_xy))
    r, c = seeded_robot(13)
    out.append(("stretch_s13", world_with_robot("Walker-v0", r, c, 1, 1), 13, perturb_stretch))
    r, c = seeded_robot(14)
    out.append(("squash_s14", world_with_robot("Walker-v0", r, c, 1, 1), 14, perturb_squash))
    r, c = seeded_robot(15)
    out.append(("collapse_s15", world_with_robot("Walker-v0", r, c, 1, 1), 15, perturb_collapse))
    # Traverser-v0: 28 simulated objects (the robot and 27 loose rigid blocks that settle onto each other during the rollout)
    from modularevogym_b200.world import traverser_world
    r, c = seeded_robot(16)
    out.append(("blocksoup_s16", traverser_world(r, c), 16, None))
    return out


def main():
    only = set(sys.argv[1:])
    for name, world, seed, perturb in cases():
        if only and name not in only:
            continue
        big = world.objects["robot"].grid.size > 64 or len(world.objects) > 8
        fx = record(name, world, seed, perturb, n_sub=30 if big else N_SUB, n_steps=4 if big else N_STEPS)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **fx)


if __name__ == "__main__":
    main()
