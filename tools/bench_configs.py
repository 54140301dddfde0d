#!/usr/bin/env python
"""Device-resident throughput of BASELINE.json's other configurations (modularevogym_b200/workloads.py): a different seeded
robot per environment, random clipped actions, CUDA events on the engine's stream.  Development tool; bench.py carries the
judged numbers.

    python tools/bench_configs.py [workload:n_envs[:steps[:settle]] ...]      e.g. obstacle20:1024 pusher:16384 hetero:8192
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    from modularevogym_b200 import capi, workloads
    specs = sys.argv[1:] or ["walker5x5:4096", "pusher:16384", "thrower:16384", "climber5:4096", "bridge:2048", "obstacle10:2048",
                             "obstacle20:1024"]
    dev = torch.device("cuda", 0)
    for spec in specs:
        f = spec.split(":")
        name, n = f[0], int(f[1])
        K = int(f[2]) if len(f) > 2 else 20
        settle = int(f[3]) if len(f) > 3 else 0
        B = capi.Batch(0)
        P = E = 0
        for i in range(n):
            env = B.add_world(workloads.world_for(name, i))
            P += B.num_points(env); E += B.num_edges(env)
        B.finalize()
        B.set_auto_reset(True)
        stride = B.action_stride
        W = 3
        rng = np.random.default_rng(0)
        acts = torch.from_numpy(workloads.clipped_actions(rng, (K + W, n, stride))).to(dev)
        stream = torch.cuda.ExternalStream(B.cuda_stream, device=dev)
        torch.cuda.synchronize()
        for s in range(settle):
            B.bind_actions_device(acts[s % (K + W)].data_ptr(), stride)
            B.step_async(1, True)
        for s in range(W):
            B.bind_actions_device(acts[s].data_ptr(), stride)
            B.step_async(1, True)
        B.sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for s in range(K):
            B.bind_actions_device(acts[W + s].data_ptr(), stride)
            B.step_async(1, True)
        e1.record(stream)
        B.sync()
        ms = e0.elapsed_time(e1) / K
        print(json.dumps({"workload": name, "envs": n, "points": P / n, "edges": E / n, "ms_per_step": round(ms, 4),
                          "env_steps_per_s": round(n / (ms * 1e-3)), "unstable_events": int(B.unstable_events), "launches": B.last_launches,
                          "settle": settle}), flush=True)
        B.bind_actions_device(0, 0)
        B.close()


if __name__ == "__main__":
    main()
