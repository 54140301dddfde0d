#!/usr/bin/env python
"""Design-loop rebuild: `BatchedEvoGymEnv.update(bodies)` = what `env.update(body, connections)` does in the
reference once per episode of NCA / CuCo design training (evogym/envs/walk.py:105-120: new EvoWorld + new EvoSim per
environment), for a whole batch: worlds built on the host, packed on all cores, uploaded, reset observation returned.

    python tools/bench_update.py [n_envs]
"""
import json
import os
import random
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    from modularevogym_b200.vec_env import BatchedEvoGymEnv
    from modularevogym_b200.world import sample_robot
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
    random.seed(0)
    gen = lambda: [sample_robot((5, 5)) for _ in range(n)]
    a, b = gen(), gen()
    t0 = time.perf_counter()
    env = BatchedEvoGymEnv("Walker-v0", [r for r, _ in a], [c for _, c in a])
    env.reset()
    torch.cuda.synchronize()
    t_build = time.perf_counter() - t0
    acts = torch.zeros((n, env.action_width), device=env.device)
    for _ in range(3):
        env.step(acts)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    env.update([r for r, _ in b], [c for _, c in b])
    torch.cuda.synchronize()
    t_update = time.perf_counter() - t0
    t0 = time.perf_counter()
    for _ in range(20):
        env.step(acts)
    torch.cuda.synchronize()
    t_step = (time.perf_counter() - t0) / 20
    print(json.dumps({"envs": n, "first_build_s": round(t_build, 3), "update_s": round(t_update, 3), "update_ms_per_env": round(1e3 * t_update / n, 4),
                      "env_step_ms_after_update": round(1e3 * t_step, 3), "update_costs_env_steps": round(t_update / t_step, 1)}))
    env.close()


if __name__ == "__main__":
    main()
