#!/usr/bin/env python
"""Throughput of the other BASELINE.json configurations (parity-test cases, not bench.py lines): N copies of a
golden fixture's world, random clipped actions, device-resident stepping, CUDA events on the engine's stream.

    python tools/bench_configs.py [name:n_envs ...]
"""
import os, sys, json, time
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import torch
    from modularevogym_b200 import capi
    from helpers import Golden
    specs = sys.argv[1:] or ["walker_s0:4096", "pusher_s3:4096", "thrower_s4:4096", "climber_s5:4096", "bridge_s6:2048",
                             "obstacle_10x10:2048", "obstacle_20x20:1024"]
    dev = torch.device("cuda", 0)
    for spec in specs:
        name, n = spec.split(":")
        n = int(n)
        g = Golden(name)
        B = capi.Batch(0)
        for _ in range(n):
            B.add_world(g.world)
        B.finalize()
        B.set_auto_reset(True)
        stride = B.action_stride
        K, W = 20, 3
        rng = np.random.default_rng(0)
        acts = torch.from_numpy(np.clip(rng.uniform(-1, 1, (K + W, n, stride)) + 1.0, 0.6, 1.6)).to(dev)
        stream = torch.cuda.ExternalStream(B.cuda_stream, device=dev)
        torch.cuda.synchronize()
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
        print(json.dumps({"world": name, "envs": n, "points": B.num_points(0), "edges": B.num_edges(0), "ms_per_step": ms,
                          "env_steps_per_s": n / (ms * 1e-3), "unstable_events": int(B.unstable_events)}), flush=True)
        B.bind_actions_device(0, 0)
        B.close()


if __name__ == "__main__":
    main()
