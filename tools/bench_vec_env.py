#!/usr/bin/env python
"""Throughput of the batched vec-env front (modularevogym_b200/vec_env.py): policy output (float32, on the device) ->
actuator targets -> 30 substeps -> reward / done / modular observation (float32) + reset of finished environments, three
launches per env-step on one stream, no host round trip.  CUDA events on the engine's stream.

    python tools/bench_vec_env.py [--warmup=W] [--steps=K] [env_id:n_envs ...]
"""
import json
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    from modularevogym_b200.vec_env import BatchedEvoGymEnv
    from modularevogym_b200.world import sample_robot
    opts = dict(a[2:].split("=") for a in sys.argv[1:] if a.startswith("--"))
    specs = [a for a in sys.argv[1:] if not a.startswith("--")] or ["Walker-v0:4096", "Pusher-v0:4096", "Thrower-v0:4096", "Climber-v0:4096", "ObstacleTraverser-v1:4096"]
    for spec in specs:
        env_id, n = spec.split(":")
        n = int(n)
        bodies, conns = [], []
        for i in range(n):
            random.seed(i)
            b, c = sample_robot((5, 5))
            bodies.append(b); conns.append(c)
        env = BatchedEvoGymEnv(env_id, bodies, conns, device=0, mode="modular")
        env.reset()
        K, W = int(opts.get("steps", 100)), int(opts.get("warmup", 50))
        g = torch.Generator(device="cuda").manual_seed(0)
        acts = torch.rand((K + W, n, env.action_width), device="cuda", generator=g) * 2 - 1
        stream = env._stream
        n_done = 0
        print("built", env_id, n, file=sys.stderr, flush=True)
        for s in range(W):                      # the warm-up runs the loop's own consumer of `done` too: its first use loads torch's
            _, _, done, _ = env.step(acts[s])   # reduction kernels, which does not belong in the timed region
            n_done = n_done + done.sum()
        n_done = 0
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        import time
        e0.record(stream)
        t_host = time.perf_counter()
        for s in range(K):
            _, _, done, _ = env.step(acts[W + s])
            n_done = n_done + done.sum()
        t_host = (time.perf_counter() - t_host) / K * 1e3      # host time to enqueue one env-step (no sync inside)
        e1.record(stream)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / K
        # per-launch breakdown of one env-step (events between the three launches, on the engine's stream)
        B, parts = env.B, [0.0, 0.0, 0.0]
        f, fs, m, msd, o, os_ = env._ptrs()
        for s in range(20):
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
            a = acts[W + s]
            ev[0].record(stream)
            B.task_apply_policy_device(a.data_ptr(), a.shape[1], True, True)
            ev[1].record(stream)
            B.step_async(1, apply_actions=True)
            ev[2].record(stream)
            B.task_outputs_device(env._reward.data_ptr(), env._done.data_ptr(), f, fs, m, msd, o, os_, True, reset_done=True)
            ev[3].record(stream)
            torch.cuda.synchronize()
            for k in range(3):
                parts[k] += ev[k].elapsed_time(ev[k + 1]) / 20
        print(json.dumps({"front": "BatchedEvoGymEnv modular f32", "env_id": env_id, "envs": n, "ms_per_step": ms,
                          "env_steps_per_s": n / (ms * 1e-3), "launches_per_step": 3, "episodes_finished": int(n_done),
                          "host_enqueue_ms_per_step": t_host,
                          "ms_policy_actions_kernel": parts[0], "ms_step_kernel": parts[1], "ms_task_outputs_kernel": parts[2]}), flush=True)
        env.close()
        print("closed", env_id, file=sys.stderr, flush=True)


if __name__ == "__main__":
    main()
