#!/usr/bin/env python
"""Where the time of one BatchedEvoGymEnv.step goes beyond its three kernels: the same 4096 Walker-v0 environments stepped
(a) through env.step with a consumer of `done` (what tools/bench_vec_env.py times), (b) through env.step alone, (c) by copying the
actions and replaying the captured CUDA graph directly.  Every variant starts from a reset + 50 steps and sees the same actions; CUDA
events on the engine's stream, 200 steps each."""
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
    env_id, n = (sys.argv[1].split(":") + ["4096"])[:2] if len(sys.argv) > 1 else ("Walker-v0", "4096")
    n = int(n)
    bodies, conns = [], []
    for i in range(n):
        random.seed(i)
        b, c = sample_robot((5, 5))
        bodies.append(b); conns.append(c)
    env = BatchedEvoGymEnv(env_id, bodies, conns, device=0, mode="modular")
    env.reset()
    K, W = 200, 300
    g = torch.Generator(device="cuda").manual_seed(0)
    acts = torch.rand((64, n, env.action_width), device="cuda", generator=g) * 2 - 1
    for s in range(W):
        env.step(acts[s % 64])
    torch.cuda.synchronize()
    stream = env._stream
    copy_then_replay = lambda s: (env._actions.copy_(acts[s % 64]), env._stream.wait_stream(torch.cuda.current_stream()), env.B.graph_launch(env._graph), env._wait())
    out = {"env_id": env_id, "envs": n}

    def timed(name, body):
        env.reset()                             # every variant starts from the same state and sees the same actions
        for s in range(50):
            env.step(acts[s % 64])
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        t = time.perf_counter()
        for s in range(K):
            body(s)
        host = (time.perf_counter() - t) / K * 1e3
        e1.record(stream)
        torch.cuda.synchronize()
        out[name] = {"ms_per_step": e0.elapsed_time(e1) / K, "host_ms_per_step": host}

    acc = [torch.zeros((), dtype=torch.int64, device="cuda")]

    def with_consumer(s):
        _, _, done, _ = env.step(acts[s % 64])
        acc[0] = acc[0] + done.sum()
    timed("env.step + done.sum()", with_consumer)
    timed("env.step", lambda s: env.step(acts[s % 64]))
    timed("copy + graph replay", copy_then_replay)
    timed("env.step + done.sum() again", with_consumer)
    print(json.dumps(out))
    env.close()


if __name__ == "__main__":
    main()
