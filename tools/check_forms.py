#!/usr/bin/env python
"""Bit-identity of the scheduling forms at benchmark scale: the same rollout stepped (a) one launch per size class and (b) in the
default form (work items for the warp classes, chunked launches on two streams for the group classes where the estimate picks
them) must leave the same states, flags and step counters.

    python tools/check_forms.py [workload:envs[:steps] ...]        default: walker5x5:4096:60 pusher:4096:40 obstacle20:1024:20 hetero:2048:20
"""
import hashlib
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def run(spec):
    from modularevogym_b200 import capi, workloads
    f = spec.split(":")
    name, n, steps = f[0], int(f[1]), int(f[2]) if len(f) > 2 else 20
    B = capi.Batch(0)
    for i in range(n):
        B.add_world(workloads.world_for(name, i))
    B.finalize()
    B.set_auto_reset(True)
    rng = np.random.default_rng(1)
    h = hashlib.sha256()
    launches = 0
    for s in range(steps):
        B.set_actions(workloads.clipped_actions(rng, (n, B.action_stride)))
        _, flags = B.step(1, True)
        launches = B.last_launches
        h.update(flags.tobytes())
    for i in range(n):
        for a in B.get_state(i):
            h.update(a.tobytes())
        h.update(int(B.get_time(i)).to_bytes(4, "little"))
    ev = B.unstable_events
    B.close()
    return {"workload": name, "envs": n, "steps": steps, "launches_per_step": launches, "unstable_events": ev, "sha256": h.hexdigest()}


def main():
    if len(sys.argv) > 2 and sys.argv[1] == "--one":
        print(json.dumps(run(sys.argv[2])))
        return 0
    specs = sys.argv[1:] or ["walker5x5:4096:60", "pusher:4096:40", "obstacle20:1024:20", "hetero:2048:20"]
    ok = True
    for spec in specs:
        out = []
        for env in ({"MEVO_WCHUNKS": "1", "MEVO_CHUNKS": "1"}, {}):
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--one", spec], capture_output=True, text=True, env={**os.environ, **env})
            line = [l for l in r.stdout.splitlines() if l.startswith("{")]
            if not line:
                print(r.stdout[-2000:], r.stderr[-2000:])
                return 1
            out.append(json.loads(line[-1]))
        same = out[0]["sha256"] == out[1]["sha256"] and out[0]["unstable_events"] == out[1]["unstable_events"]
        ok &= same
        print(json.dumps({"spec": spec, "single_launch": out[0], "default_form": out[1], "identical": same}), flush=True)
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
