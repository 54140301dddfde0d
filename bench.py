#!/usr/bin/env python
"""Benchmark of the hot path: batched env-steps/s of Sim::step (30 RK4 substeps) on Walker-v0 5x5.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--envs 4096] [--impl ours|reference]

Workload (BASELINE.json configs[1]): per GPU 4096 Walker-v0 environments (100x1 FIXED
ground + one random 5x5 robot, robot i = sample_robot((5,5)) under random.seed(i), i = global env
index), actions a ~ U(-1,1) + 1 clipped to [0.6, 1.6] per actuator per step, synthetic.  One "step"
= one env-step (30 substeps = 120 force evaluations) of every environment.  Weak scaling: every
rank owns --envs environments, no collective on the step path; one NCCL all-gather of the per-env
reward at the end of the rollout.

Prints ONE JSON line (rank 0).  `value` = device-resident throughput (actions already in HBM),
`e2e` = through the C-ABI with pinned HOST action buffers in and host rewards/flags out every step.
`--impl reference` times the reference's own CPU engine (oracle/_ref/libmevo_ref.so, compiled from
the unmodified sources) on all host cores, on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import random
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from modularevogym_b200.world import sample_robot, walker_world  # noqa: E402

METRIC = "batched env-steps/sec (Walker-v0 5x5, 4096 envs)"
UNIT = "env-steps/s"
SUBSTEPS = 30
WORKLOAD = ("Walker-v0 (100x1 FIXED ground) + random 5x5 robot per env, robot i = sample_robot((5,5)) under "
            "random.seed(i); actions U(-1,1)+1 clipped to [0.6,1.6] per actuator per step")


HETERO_METRIC = "batched env-steps/sec (heterogeneous modular robots, sizes 5x5..20x20, Walker-v0 ground)"
HETERO_WORKLOAD = ("BASELINE.json configs[4] per GPU: Walker-v0 ground + robot i of side s_i ~ U{5..20} (default_rng(i)), "
                   "sample_robot((s,s)) for s <= 6 else sample_large_robot((s,s)) under random.seed(i); same action law")
ACTIVE_WORKLOAD = "walker5x5"


def workload_robot(i: int):
    random.seed(i)
    if ACTIVE_WORKLOAD == "hetero":
        from modularevogym_b200.world import sample_large_robot
        s = int(np.random.default_rng(i).integers(5, 21))
        return sample_robot((s, s)) if s <= 6 else sample_large_robot((s, s))
    return sample_robot((5, 5))


def clipped_actions(rng, shape):
    return np.clip(rng.uniform(-1.0, 1.0, shape) + 1.0, 0.6, 1.6)     # evogym/envs/base.py:42,389


def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


# ------------------------------------------------------------------------------------------------
# CPU arm: the reference engine (or, if it was not built, the oracle port) on the host cores

def _cpu_worker(first_env: int, n_envs: int, kind: str) -> None:
    """Child process: builds n_envs Sims, then serves 'go <n_steps>' requests on stdin: advances every
    Sim by n_steps env-steps and answers with one JSON line (steps done, seconds)."""
    sims = []
    for i in range(first_env, first_env + n_envs):
        r, c = workload_robot(i)
        w = walker_world(r, c)
        if kind == "reference":
            from oracle.ref import RefSim
            s = RefSim(w)
        else:
            from oracle.oracle import OracleSim
            s = OracleSim(w)
        n_act = len(s.actuator_indices("robot"))
        sims.append((s, np.random.default_rng(1234 + i), n_act))
    print("ready", flush=True)
    for line in sys.stdin:
        tok = line.split()
        if not tok or tok[0] != "go":
            break
        n_steps = int(tok[1])
        acts = [clipped_actions(rng, (n_steps, n_act)) for _, rng, n_act in sims]
        t0 = time.perf_counter()
        done = 0
        for (s, _, _), a in zip(sims, acts):
            for k in range(n_steps):
                s.set_action_values("robot", a[k])
                if s.step():
                    s.revert(0)                   # unstable: reset to t = 0
                done += 1
        dt = time.perf_counter() - t0
        print(json.dumps({"env_steps": done, "seconds": dt}), flush=True)


def cpu_kind() -> str:
    from oracle import ref
    return "reference" if ref.available() else "port"


class CpuPool:
    """One process per host core, each owning `envs_per_core` independent Sims of the workload
    (robots 0 .. cores*envs_per_core-1).  sample(n) advances every Sim by n env-steps on all cores at
    once; the aggregate rate is the sum of the per-core rates measured while all cores are busy."""

    def __init__(self, envs_per_core: int = 1, cores: int = 0):
        self.cores = cores or host_cores()
        self.envs_per_core = envs_per_core
        self.kind = cpu_kind()
        if self.kind == "port":
            from oracle import oracle
            oracle.build()
        self.procs = []
        for c in range(self.cores):
            cmd = [sys.executable, os.path.abspath(__file__), "--_cpu_worker", str(c * envs_per_core), str(envs_per_core), self.kind]
            self.procs.append(subprocess.Popen(cmd, stdin=subprocess.PIPE, stdout=subprocess.PIPE, text=True, cwd=ROOT))
        for p in self.procs:
            line = p.stdout.readline()
            if line.strip() != "ready":
                raise RuntimeError(f"cpu worker failed to start: {line!r}")

    def sample(self, steps_per_env: int) -> dict:
        t0 = time.perf_counter()
        for p in self.procs:
            p.stdin.write(f"go {steps_per_env}\n")
            p.stdin.flush()
        rate, total = 0.0, 0
        for p in self.procs:
            res = json.loads(p.stdout.readline())
            rate += res["env_steps"] / res["seconds"]
            total += res["env_steps"]
        return {"value": rate, "wall_s": time.perf_counter() - t0, "env_steps": total}

    def describe(self, steps_per_env: int) -> str:
        src = "unmodified reference sources, g++ -O2" if self.kind == "reference" else "oracle/mevo_oracle.c, gcc -O2"
        return (f"{self.cores} processes x {self.envs_per_core} Walker-v0 5x5 env(s) (robots 0..{self.cores * self.envs_per_core - 1} "
                f"of the workload) x {steps_per_env} env-steps per sample; {src}")

    def close(self) -> None:
        for p in self.procs:
            try:
                p.stdin.write("quit\n"); p.stdin.flush(); p.stdin.close()
                p.wait(timeout=10)
            except Exception:
                p.kill()


# ------------------------------------------------------------------------------------------------
# clocks during the timed region (B200_PROFILING.md)

class ClockSampler:
    """nvidia-smi clocks and throttle reasons, sampled every 20 ms from before the warm-up; stop() keeps the samples
    whose timestamps fall inside the timed region [mark_start(), mark_end()] (all samples under load if the region
    was shorter than the sampling period)."""
    Q = ("timestamp,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.path = tempfile.mktemp(prefix="mevo_clocks_", suffix=".csv")
        self.proc = None
        self.t0 = self.t1 = None
        try:
            self.f = open(self.path, "w")
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "20"], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def mark_start(self) -> None:
        self.t0 = time.time()

    def mark_end(self) -> None:
        self.t1 = time.time()

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.f.close()
        import datetime
        rows = []
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for line in open(self.path):
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                ts = datetime.datetime.strptime(c[0], "%Y/%m/%d %H:%M:%S.%f").timestamp()
                rows.append((ts, float(c[1]), float(c[2]), [n for n, v in zip(names, c[5:9]) if v.lower().startswith("active")]))
            except ValueError:
                continue
        try:
            os.unlink(self.path)
        except OSError:
            pass
        inside = [r for r in rows if self.t0 is not None and self.t1 is not None and self.t0 <= r[0] <= self.t1]
        window = "timed region"
        if not inside:
            inside, window = [r for r in rows if self.t0 is None or r[0] >= self.t0 - 1.0], "warm-up + timed region (timed region shorter than the sampling period)"
        reasons = sorted({n for r in inside for n in r[3]})
        return {"sm_mhz": float(np.median([r[1] for r in inside])) if inside else None,
                "sm_max_mhz": max(r[2] for r in inside) if inside else None,
                "samples": len(inside), "window": window, "reasons": reasons}


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


# ------------------------------------------------------------------------------------------------

def main_reference(args) -> None:
    """--impl reference: the reference's own CPU engine on all host cores.  A 'step' here is a bounded
    sample: every core advances its own Sim by `spe` env-steps (same robots and action law as the
    GPU arm's first environments)."""
    rank, _, world = dist_env()
    if rank != 0:
        return
    pool = CpuPool(1)
    spe = 20
    for _ in range(args.warmup):
        pool.sample(spe)
    vals, secs = [], 0.0
    for _ in range(args.steps):
        r = pool.sample(spe)
        vals.append(r["value"]); secs += r["wall_s"]
    pool.close()
    value = float(np.mean(vals)) if vals else 0.0
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * secs / max(args.steps, 1), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "substeps_per_step": SUBSTEPS,
                   "parallelism": f"{pool.cores} host processes, one Sim each"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": pool.cores, "kind": pool.kind,
                         "sample": pool.describe(spe) + f"; mean of {args.steps} samples"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# dram__bytes_read.sum + dram__bytes_write.sum of one step launch of 4096 envs, from the ncu --set full capture
# summarised in profiles/r01h_wstep_kernel_raw.csv (25.50 MB read + 0.39 MB written): the state stays on chip
NCU_DRAM_BYTES_PER_LAUNCH = 25_889_792


def algorithmic_bytes_per_step(P: int, E: int) -> int:
    """every array the reference touches once in and once out per substep --
    105 B per point (pos, vel, pos_last read+written, mass, fixed) + 32 B per edge (a, b, goal, L0, k),
    in the REFERENCE's numbering (ground included), x 30 substeps."""
    return SUBSTEPS * (105 * P + 32 * E)


def main_ours(args) -> None:
    import torch
    import torch.distributed as dist
    from modularevogym_b200 import capi

    rank, local, world = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the engine has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    n, K, W = args.envs, args.steps, args.warmup
    hetero = args.workload == "hetero"

    # ---- build this rank's shard of the workload (host, untimed)
    t_build = time.perf_counter()
    B = capi.Batch(local)
    P_sum = E_sum = 0
    for i in range(n):
        r, c = workload_robot(rank * n + i)
        env = B.add_world(walker_world(r, c))
        P_sum += B.num_points(env)
        E_sum += B.num_edges(env)
    B.finalize()
    B.set_auto_reset(True)                      # unstable envs are reverted to t = 0 and counted
    t_build = time.perf_counter() - t_build
    stride = B.action_stride
    rng = np.random.default_rng(1234 + rank)
    stream = torch.cuda.ExternalStream(B.cuda_stream, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)       # > 126 MB L2

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def flush_l2():
        with torch.cuda.stream(stream):
            flush.zero_()

    # ---- (1) device-resident arm: actions for all steps already in HBM, one kernel launch per step
    acts_host = torch.from_numpy(clipped_actions(rng, (K + W, n, stride))).pin_memory()
    acts_dev = acts_host.to(dev)
    com_dev = torch.zeros((n, 2), dtype=torch.float64, device=dev)
    torch.cuda.synchronize()

    def resident_step(s):
        B.bind_actions_device(acts_dev[s].data_ptr(), stride)
        B.step_async(1, True)

    clocks = ClockSampler(local) if rank == 0 else None
    for s in range(W):
        resident_step(s)
    barrier()
    ev0 = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
    ev1 = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
    launches = 0
    if clocks:
        clocks.mark_start()
    t_wall = time.perf_counter()
    for s in range(K):
        flush_l2()                                           # between timed iterations, outside the events
        ev0[s].record(stream)
        resident_step(W + s)
        ev1[s].record(stream)
        launches += B.last_launches
    # rollout end: per-env reward (robot COM) gathered over NCCL -- the only collective
    B.object_com_device(0, com_dev.data_ptr())
    launches += 1
    B.sync()
    if world > 1:
        from modularevogym_b200.sharding import gather_env_rows
        gathered = gather_env_rows(com_dev, world * n)       # (world * n, 2) on every rank
        assert gathered.shape[0] == world * n
    barrier()
    t_wall = time.perf_counter() - t_wall
    if clocks:
        clocks.mark_end()
    clock_info = clocks.stop() if clocks else None
    dev_ms = sum(a.elapsed_time(b) for a, b in zip(ev0, ev1))
    events_before = B.unstable_events
    B.bind_actions_device(0, 0)

    # ---- (2) end-to-end arm: pinned host actions in, host rewards + flags out, every step, through the C ABI
    com_host = torch.zeros((n, 2), dtype=torch.float64).pin_memory()
    flags_host = torch.zeros((n,), dtype=torch.uint8).pin_memory()
    e2e_s = 0.0
    for s in range(W + K):
        flush_l2()
        B.sync()
        t0 = time.perf_counter()
        B.set_actions_ptr(acts_host[s].data_ptr(), stride)   # H2D (pinned source, on the engine's stream)
        B.step_async(1, True)                                # kernel
        B.fetch_step_results_ptr(flags_host.data_ptr(), 0, com_host.data_ptr())   # D2H of flags + per-env reward signal, one sync
        dt = time.perf_counter() - t0
        if s >= W:
            e2e_s += dt
    h2d = n * stride * 8
    d2h = n * 2 * 8 + n

    # ---- max over ranks
    t = torch.tensor([dev_ms, e2e_s * 1e3, t_wall * 1e3], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms, wall_ms = (float(x) for x in t.tolist())
    total_env_steps = world * n * K
    value = total_env_steps / (dev_ms * 1e-3)
    e2e_value = total_env_steps / (e2e_ms * 1e-3)

    if rank == 0:
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(peaks_path):
            peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        else:
            peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
        alg_bytes = algorithmic_bytes_per_step(P_sum, E_sum)              # per launch: this rank's n envs, one env-step
        achieved = alg_bytes / (dev_ms * 1e-3 / K) / 1e9
        cpu = None
        if not args.no_cpu_baseline and not hetero:          # the CPU arm times the headline workload only
            pool = CpuPool(1)
            pool.sample(5)
            spe = 40
            r1 = pool.sample(spe)
            r2 = pool.sample(spe)
            pool.close()
            cpu = {"value": 0.5 * (r1["value"] + r2["value"]), "unit": UNIT, "cores": pool.cores, "kind": pool.kind,
                   "sample": pool.describe(spe) + "; mean of 2 samples"}
        line = {
            "metric": HETERO_METRIC if hetero else METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": dev_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": HETERO_WORKLOAD if hetero else WORKLOAD, "envs_per_gpu": n, "substeps_per_step": SUBSTEPS,
                       "points_per_env": P_sum / n, "edges_per_env": E_sum / n, "parallelism": f"env-sharded x{world}, no collective on the step path",
                       "l2": "256 MB buffer written between timed iterations (outside the events)",
                       "timing": "CUDA events on the engine's stream around each step, summed; max over ranks",
                       "wall_ms_incl_flush": wall_ms, "build_s": t_build},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": e2e_ms / K},
            "gpu_launches": launches,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": NCU_DRAM_BYTES_PER_LAUNCH if n == 4096 and not hetero else None,
                         "kernel": "mevo::step_kernel<G> per size class + mevo::wstep_kernel" if hetero else "mevo::wstep_kernel<40,112,24,192,256,112,16,1>",
                         "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": alg_bytes},
            "cpu_baseline": cpu,
            "clocks": clock_info,
            "unstable_env_steps": int(events_before),
        }
        print(json.dumps(line), flush=True)
    B.close()
    if world > 1:
        dist.destroy_process_group()


def main() -> None:
    if len(sys.argv) > 1 and sys.argv[1] == "--_cpu_worker":
        _cpu_worker(int(sys.argv[2]), int(sys.argv[3]), sys.argv[4])
        return
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--envs", type=int, default=4096, help="environments per GPU")
    ap.add_argument("--impl", choices=("ours", "reference"), default="ours")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", choices=("walker5x5", "hetero"), default="walker5x5",
                    help="walker5x5 = the headline (BASELINE.json configs[1]); hetero = configs[4]'s robot mix (our arm only)")
    args = ap.parse_args()
    global ACTIVE_WORKLOAD
    ACTIVE_WORKLOAD = args.workload
    if args.impl == "reference":
        main_reference(args)
    else:
        main_ours(args)


if __name__ == "__main__":
    main()
