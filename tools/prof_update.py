#!/usr/bin/env python
"""cProfile of `BatchedEvoGymEnv.update` (design-loop rebuild) for 4096 Walker-v0 environments: where the host time goes.

    python tools/prof_update.py
"""
import cProfile, pstats, random, sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from modularevogym_b200.vec_env import BatchedEvoGymEnv
from modularevogym_b200.world import sample_robot
n = 4096
random.seed(0)
a = [sample_robot((5, 5)) for _ in range(n)]
b = [sample_robot((5, 5)) for _ in range(n)]
env = BatchedEvoGymEnv("Walker-v0", [r for r, _ in a], [c for _, c in a])
env.reset()
torch.cuda.synchronize()
pr = cProfile.Profile(); pr.enable()
t0 = time.perf_counter()
env.update([r for r, _ in b], [c for _, c in b])
torch.cuda.synchronize()
print("update s", time.perf_counter() - t0)
pr.disable()
pstats.Stats(pr).sort_stats("cumtime").print_stats(22)
