"""Batched vector environment: the front the PPO loop talks to, on the device.

Host-side mirror of what the reference assembles per rollout worker out of processes and pipes
(``examples/transformer_ppo/envs.py:29-77`` ``make_env`` / ``make_vec_envs``: ``gym.make(env_id, body=..., mode='modular')``
-> ``TimeLimit`` -> ``SubprocVecEnv`` (``examples/modular_envs/wrappers/subproc_vec_env.py``: reset on done, return the reset
observation) -> ``VecPyTorch`` (``envs.py:126-159``: numpy -> float32 torch tensors on the device, reward ``(n, 1)``)),
for all 32 environment ids the reference registers (``TASKS``), in ModularEvoGym's modular mode and in EvoGym's flat mode.

One env-step of all environments = three launches on one CUDA stream and no host round trip:
policy output -> actuator targets (``policy_actions_kernel``), 30 substeps (``wstep_kernel`` / ``step_kernel``),
reward / done / observation + reset of finished environments (``task_outputs_kernel``).  PyTorch is used for the
output tensors only; all compute is in ``libmevo_b200.so`` through the C ABI (``include/mevo_b200.h``, "task layer").
"""
from __future__ import annotations

import random
from typing import Dict, Optional, Sequence, Tuple

import numpy as np

from . import capi
from .world import (WorldSpec, balancer_v1_world, balancer_world, beam_slider_world, beam_toppler_world, bidirectional_walker_world, bridge_walker_world, catcher_world, traverser_world, carrier_v1_world, carrier_world, cave_crawler_world,
                    climber_v1_world, climber_v2_world, climber_world, down_stepper_world, flipper_world, full_connectivity, gap_jumper_world, hurdler_world, platform_jumper_world,
                    jumper_world, lifter_world, obstacle_traverser_v0_world, obstacle_traverser_world, pusher_v1_world, pusher_world, shape_change_world,
                    terrain_for, thrower_world, up_stepper_world, walker_world)

# env id -> (task id of the C ABI, world builder, package object, gym registration's max_episode_steps
# (evogym/envs/__init__.py:16-20,125-129,149-153))
TASKS = {
    "Walker-v0": (0, walker_world, None, 500),
    "Pusher-v0": (1, pusher_world, "package", 500),
    "Thrower-v0": (2, thrower_world, "package", 300),
    "Carrier-v0": (14, carrier_world, "package", 500),                        # manipulate.py:83-199, __init__.py:113-117
    "Carrier-v1": (20, carrier_v1_world, "package", 1000),                    # manipulate.py:202-315, __init__.py:119-123
    "Lifter-v0": (21, lifter_world, "package", 300),                          # manipulate.py:1207-1340, __init__.py:107-111
    "BeamToppler-v0": (23, beam_toppler_world, "beam", 1000),                 # manipulate.py:887-1046, __init__.py:137-141
    "BeamSlider-v0": (24, beam_slider_world, "beam", 1000),                   # manipulate.py:1049-1204, __init__.py:143-147
    "Pusher-v1": (15, pusher_v1_world, "package", 600),                       # manipulate.py:419-516, __init__.py:131-135
    "Climber-v0": (3, climber_world, None, 400),                              # evogym/envs/climb.py:38-112, __init__.py:187-191
    "ObstacleTraverser-v1": (4, obstacle_traverser_world, None, 1000),        # evogym/envs/traverse.py:350-426, __init__.py:76-80
    "BridgeWalker-v0": (5, bridge_walker_world, None, 500),                   # evogym/envs/walk.py:122-232, __init__.py:22-26
    "ObstacleTraverser-v0": (6, obstacle_traverser_v0_world, None, 1000),     # traverse.py:251-347, __init__.py:70-74
    "UpStepper-v0": (7, up_stepper_world, None, 600),                         # traverse.py:51-149, __init__.py:58-62
    "DownStepper-v0": (8, down_stepper_world, None, 500),                     # traverse.py:152-248, __init__.py:64-68
    "Hurdler-v0": (10, hurdler_world, None, 1000),                            # traverse.py:446-536, __init__.py:82-86
    "Climber-v1": (11, climber_v1_world, None, 600),                          # climb.py:131-231, __init__.py:193-197
    "CaveCrawler-v0": (12, cave_crawler_world, None, 1000),                   # walk.py:235-345, __init__.py:28-32
    "Climber-v2": (13, climber_v2_world, None, 1000),                         # climb.py:226-320, __init__.py:199-203
    "Balancer-v0": (16, balancer_world, None, 600),                           # balance.py:14-135, __init__.py:46-50
    "Balancer-v1": (17, balancer_v1_world, None, 600),                        # balance.py:138-259, __init__.py:52-56
    "GapJumper-v0": (18, gap_jumper_world, None, 1000),                       # traverse.py:663-782, __init__.py:88-92
    "PlatformJumper-v0": (19, platform_jumper_world, None, 1000),             # traverse.py:539-660, __init__.py:94-98
    "Flipper-v0": (22, flipper_world, None, 600),                             # flip.py:15-140, __init__.py:40-44
    "WingspanMazimizer-v0": (25, shape_change_world, None, 600),              # change_shape.py:287-385, __init__.py:174-178 (sic)
    "HeightMaximizer-v0": (26, shape_change_world, None, 500),                # change_shape.py:388-487, __init__.py:180-184
    "AreaMaximizer-v0": (27, shape_change_world, None, 600),                  # change_shape.py:85-183, __init__.py:162-166
    "AreaMinimizer-v0": (28, shape_change_world, None, 600),                  # change_shape.py:186-284, __init__.py:168-172
    "BidirectionalWalker-v0": (29, bidirectional_walker_world, None, 1000),   # multi_goal.py:91-205, __init__.py:206-210 (flat mode only)
    "Traverser-v0": (30, traverser_world, None, 600),                         # traverse.py:784-911, __init__.py:100-104
    # Catcher-v0: the reference re-draws the package's launch position at EVERY reset (it rebuilds its world and its simulator,
    # manipulate.py:815).  Here it is drawn per environment whenever the batch is built -- construction, update() and reset(), which
    # rebuilds the batch like the reference does -- and an episode restarted on the device takes over the launch position of a
    # uniformly drawn environment of the batch: one of the batch's own draws (mevo_batch_task_launch_sources)
    "Catcher-v0": (31, catcher_world, "package", 400),                        # manipulate.py:630-886, __init__.py:155-159
    "Jumper-v0": (9, jumper_world, None, 500),                                # evogym/envs/jump.py:14-118, __init__.py:34-38
}


def random_goals(lower_bound: int = 20, upper_bound: int = 50, goal_dist: int = 100, start: int = 35, limit: int = 61) -> list:
    """The x goals of one BidirectionalWalker-v0 episode, drawn as ``BiWalk.set_random_goals`` does (multi_goal.py:115-129): uniform
    positions until the path through them is ``goal_dist`` voxels long (``limit``: the device table's capacity, never reached in
    practice -- the mean hop is ten voxels)."""
    goals, curr, dist = [], start, 0
    while dist < goal_dist and len(goals) < limit:
        nxt = random.randrange(lower_bound, upper_bound)
        dist += abs(curr - nxt)
        curr = nxt
        goals.append(nxt)
    return goals


class BatchedEvoGymEnv:
    """``n`` environments of one task, one robot body each, resident on one GPU.

    ``mode='modular'`` (ModularEvoGym, ``evogym/envs/base.py:746-753``): observations are a dict ``modular`` (n, 9 * voxels),
    ``other`` (n, 2 | 6), ``act_mask`` / ``obs_mask`` (n, voxels), ``stage`` (n, 1); actions are one value per voxel.
    ``mode=None`` (EvoGym): the flat observation (n, width) and one action value per actuator; robots with fewer points than the
    widest one have their row zero-padded at the end (sight blocks and goal values follow each robot's OWN relative positions).
    ``worlds`` overrides the built-in terrain (a ``WorldSpec`` per environment that already contains the robot)."""

    def __init__(self, env_id: str, bodies: Sequence[np.ndarray], connections: Optional[Sequence[Optional[np.ndarray]]] = None,
                 device: int = 0, mode: Optional[str] = "modular", max_episode_steps: Optional[int] = None,
                 worlds: Optional[Sequence[WorldSpec]] = None, out_dtype: str = "float32", use_graph: bool = True):
        self._ctor = dict(env_id=env_id, device=device, mode=mode, max_episode_steps=max_episode_steps, out_dtype=out_dtype,
                          use_graph=use_graph)
        import torch

        if env_id not in TASKS:
            raise ValueError(f"Environment not support: {env_id}")          # the reference's message (base.py:744)
        task, builder, package, limit = TASKS[env_id]
        if task == 7 and mode != "modular":
            task = 32           # UpStepper-v0 ends the episode of a tilted robot in flat mode only (traverse.py:124-129): MEVO_TASK_UP_STEPPER_FLAT
        if task == 29 and mode == "modular":
            raise ValueError(f"{env_id} has no modular mode in the reference (evogym/envs/multi_goal.py:93); pass mode=None")
        self.env_id, self.mode, self.device_index = env_id, mode, device
        self.num_envs = len(bodies)
        self.bodies = [np.asarray(b) for b in bodies]
        self._connections, self._own_worlds = connections, worlds is not None
        self.torch = torch
        self.device = torch.device("cuda", device)
        self.dtype = {"float32": torch.float32, "float64": torch.float64}[out_dtype]
        self.B = capi.Batch(device)
        if worlds is not None:
            for i in range(self.num_envs):
                self.B.add_world(worlds[i])
        else:
            conns = [connections[i] if connections is not None and connections[i] is not None else full_connectivity(b)
                     for i, b in enumerate(self.bodies)]
            # environments that share their terrain go in through one call (design loops rebuild the batch every episode); the
            # terrain of Climber-v0 depends on the robot's width, so its environments are grouped by terrain
            placed = [terrain_for(env_id, b) for b in self.bodies]      # once per environment: Catcher-v0's terrain is a random draw
            i = 0
            while i < self.num_envs:
                terrain, rx, ry = placed[i]
                j = i + 1
                while j < self.num_envs and placed[j][0] is terrain:
                    j += 1
                self.B.add_robot_worlds(terrain, self.bodies[i:j], conns[i:j], "robot", rx, ry)
                i = j
        self.B.finalize()
        self.B.set_auto_reset(False)            # the task kernel resets (it needs the flag and the pre-step centres of mass)
        self.B.task_configure(task, "robot", package, limit if max_episode_steps is None else max_episode_steps)
        self.dims = self.B.task_dims()
        n, d = self.num_envs, self.dims
        dev, dt = self.device, self.dtype
        self._flat = torch.zeros((n, d["flat"]), device=dev, dtype=dt)
        self._modular = torch.zeros((n, d["modular"]), device=dev, dtype=dt)
        self._other = torch.zeros((n, d["other"]), device=dev, dtype=dt)
        self._reward = torch.zeros((n, 1), device=dev, dtype=dt)
        self._done = torch.zeros((n,), device=dev, dtype=torch.uint8)
        # done / truncated / unstable as bool rows, written by the task kernel next to the bit field (no kernels here to split it)
        self._flags = torch.zeros((3, n), device=dev, dtype=torch.bool)
        self.B.task_bind_flag_planes(self._flags.data_ptr())
        am = np.ones((n, d["max_voxels"]), dtype=np.int64)
        om = np.ones((n, d["max_voxels"]), dtype=np.int64)
        for i, body in enumerate(self.bodies):          # get_act_mask / get_obs_mask (evogym/envs/base.py:530-560), as task_env_tables reports them
            flat = body.ravel()
            am[i, :flat.size] = ~((flat == 3) | (flat == 4))
            om[i, :int(np.count_nonzero(flat))] = 0
        self.act_mask = torch.from_numpy(am).to(dev)
        self.obs_mask = torch.from_numpy(om).to(dev)
        self._stage = torch.ones((n, 1), device=dev, dtype=dt)
        self._stream = torch.cuda.ExternalStream(self.B.cuda_stream, device=dev)
        self.action_width = d["max_voxels"] if mode == "modular" else d["action_stride"]
        # the three launches of one env-step (+ the fork/join of the size classes) as one CUDA graph: captured after the first
        # eager step, replayed afterwards; the policy output is copied into a persistent buffer the graph reads
        self.use_graph = use_graph
        self._graph = None
        self._graph_epoch = -1
        self._graph_dtype = None
        self._actions = None
        self._eager_steps = 0

    # -- helpers
    def _obs(self):
        if self.mode == "modular":
            return {"modular": self._modular, "other": self._other, "act_mask": self.act_mask, "obs_mask": self.obs_mask,
                    "stage": self._stage}
        return self._flat

    def _ptrs(self) -> Tuple[int, int, int, int, int, int]:
        if self.mode == "modular":
            return 0, 0, self._modular.data_ptr(), self._modular.shape[1], self._other.data_ptr(), self._other.shape[1]
        return self._flat.data_ptr(), self._flat.shape[1], 0, 0, 0, 0

    # -- VecEnv surface
    def reset(self, rebuild: bool = True):
        """``venv.reset()``: every environment to snapshot 0; returns the observation (tensors owned by the env, overwritten
        by the next ``step``).  ``rebuild`` matters for Catcher-v0 only (see ``TASKS``)."""
        if self.env_id == "Catcher-v0" and not self._own_worlds and rebuild:
            self.close()                        # CatchSmallRect.reset: a new draw, a new world, a new simulator
            self.__init__(bodies=self.bodies, connections=self._connections, **self._ctor)
        if self.env_id == "BidirectionalWalker-v0":
            # BiWalk.reset draws new goals from Python's global `random` (multi_goal.py:115-129, 183-184): the same calls, here
            self.B.task_set_goals([random_goals() for _ in range(self.num_envs)])
        f, fs, m, ms, o, os_ = self._ptrs()
        self.B.task_reset_device(f, fs, m, ms, o, os_, self.dtype == self.torch.float32)
        self._wait()
        return self._obs()

    def step(self, actions):
        """``actions``: (n, voxels) in modular mode, (n, actuators) otherwise -- raw policy outputs (the environment adds 1 and
        clips to [0.6, 1.6] like ``BenchmarkBase.step``); a CUDA tensor (float32 / float64) or anything ``torch.as_tensor`` takes.
        Returns ``obs, reward (n, 1), done (n,) bool, infos`` with ``infos = {'truncated': ..., 'unstable': ...}`` tensors (all
        owned by the env and overwritten by the next ``step``, like the observation)."""
        torch = self.torch
        a = torch.as_tensor(actions)
        if a.dtype not in (torch.float32, torch.float64):
            a = a.to(torch.float32)
        a = a.to(self.device).contiguous()
        if a.shape != (self.num_envs, self.action_width):
            raise ValueError(f"actions must have shape {(self.num_envs, self.action_width)}, got {tuple(a.shape)}")
        cur = torch.cuda.current_stream(self.device)
        if self._graph is not None and self.B.finalize_epoch != self._graph_epoch:
            self._graph, self._eager_steps = None, 0            # the batch was re-packed (e.g. translate_object on terrain): re-capture
        if self.use_graph and self._graph is not None and a.dtype == self._graph_dtype:
            self._actions.copy_(a)
            self._stream.wait_stream(cur)
            self.B.graph_launch(self._graph)
        else:
            if self.use_graph and self._eager_steps >= 1 and self._graph is None:
                self._capture(a)
                self._actions.copy_(a)
                self._stream.wait_stream(cur)
                self.B.graph_launch(self._graph)
            else:
                self._stream.wait_stream(cur)       # the engine's stream sees the policy's output
                self._enqueue_step(a.data_ptr(), a.shape[1], a.dtype == torch.float32)
                self._eager_steps += 1
        self._wait()
        infos = {"truncated": self._flags[1], "unstable": self._flags[2]}
        return self._obs(), self._reward, self._flags[0], infos

    def _enqueue_step(self, act_ptr: int, act_stride: int, act_f32: bool) -> None:
        self.B.task_apply_policy_device(act_ptr, act_stride, self.mode == "modular", act_f32)
        self.B.step_async(1, apply_actions=True)
        f, fs, m, ms, o, os_ = self._ptrs()
        self.B.task_outputs_device(self._reward.data_ptr(), self._done.data_ptr(), f, fs, m, ms, o, os_,
                                   self.dtype == self.torch.float32, reset_done=True)

    def _capture(self, like) -> None:
        torch = self.torch
        self._actions = torch.zeros((self.num_envs, self.action_width), device=self.device, dtype=like.dtype)
        self._graph_dtype = like.dtype
        torch.cuda.synchronize(self.device)
        self.B.capture_begin()
        try:
            self._enqueue_step(self._actions.data_ptr(), self.action_width, like.dtype == torch.float32)
        finally:
            self._graph = self.B.capture_end()
            self._graph_epoch = self.B.finalize_epoch

    def _wait(self):
        self.torch.cuda.current_stream(self.device).wait_stream(self._stream)

    def close(self):
        self.torch.cuda.synchronize(self.device)
        self.B.close()

    def update(self, bodies: Sequence[np.ndarray], connections: Optional[Sequence[Optional[np.ndarray]]] = None):
        """Design loops (``env.update(body, connections)``, ``evogym/envs/walk.py:105-120``, called once per episode by the NCA /
        CuCo examples): new robot bodies for every environment.  Like the reference, which builds a new world and a new ``EvoSim``,
        the batch is rebuilt (worlds are packed on all host cores) and the reset observation of the new robots is returned."""
        self.close()
        self.__init__(bodies=bodies, connections=connections, **self._ctor)
        return self.reset(rebuild=False)        # (Catcher-v0: update() draws once, manipulate.py:843-851)


def make_vec_envs(env_name: str, robot_structures: Sequence, device: int = 0, mode: Optional[str] = "modular",
                  max_episode_steps: Optional[int] = None) -> BatchedEvoGymEnv:
    """Signature after ``examples/transformer_ppo/envs.py:52-77``: ``robot_structures[i]`` is either ``(body, connections)`` or
    an object with ``.body`` / ``.connections`` (the reference's ``Structure``)."""
    bodies, conns = [], []
    for s in robot_structures:
        body, cn = (s.body, s.connections) if hasattr(s, "body") else s
        bodies.append(np.asarray(body))
        conns.append(None if cn is None else np.asarray(cn))
    return BatchedEvoGymEnv(env_name, bodies, conns, device=device, mode=mode, max_episode_steps=max_episode_steps)
