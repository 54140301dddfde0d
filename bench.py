#!/usr/bin/env python
"""Benchmark of the hot path: batched env-steps/s of Sim::step (30 RK4 substeps per env-step).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--envs n] [--impl ours|reference]
                    [--workload walker5x5|pusher|thrower|obstacle10|obstacle20|climber20|bridge|hetero]

Headline workload (BASELINE.json configs[1]): per GPU 4096 Walker-v0 environments (100x1 FIXED ground + one
random 5x5 robot, robot i = sample_robot((5,5)) under random.seed(i), i = global env index), actions a ~ U(-1,1) + 1 clipped to
[0.6, 1.6] per actuator per step, synthetic.  One "step" = one env-step (30 substeps = 120 force evaluations) of every
environment.  Weak scaling: every rank owns --envs environments, no collective on the step path; one NCCL all-gather of
per-env rows at the end of the rollout (timed warm, `gather_ms`).  The other workloads (modularevogym_b200/workloads.py) are
BASELINE.json's configs[2..4]; a short run of each, with its own CPU baseline, rides in the headline line's `configs` block.

What is timed (rank 0 prints ONE JSON line):
  * the rollout first runs `settle` (default 500) untimed env-steps with auto-reset -- SURVEY 8d C2 is a 500-step rollout and
    the step kernel's time depends on the state (contact-rich poses cost more than the rest pose) -- then W warm-up steps,
    then K timed steps: `value` = steady-state device-resident throughput (actions already in HBM, CUDA events on the
    engine's stream around each step, 256 MB L2 flush between steps, max over ranks);
  * `value_after_reset` = the same measurement on the first K steps after a reset (what round 1 reported as `value`);
  * `e2e` = through the C ABI with pinned HOST action buffers in and host flags + reward signal out every step;
  * `cpu_baseline` = the reference's own CPU engine (oracle/_ref, compiled from the unmodified sources) on the box's host
    cores, Sims sampled across the whole workload, same settle phase; single-thread and all-core, -O2 and -O3;
  * `--impl reference` times that CPU engine alone (all cores, -O2), each step a bounded sample of the workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from modularevogym_b200 import workloads  # noqa: E402

UNIT = "env-steps/s"
SUBSTEPS = 30
HEADLINE = "walker5x5"
METRIC = "batched env-steps/sec (Walker-v0 5x5, 4096 envs)"
DEFAULT_ENVS = {"walker5x5": 4096, "pusher": 16384, "thrower": 16384, "obstacle10": 2048, "obstacle20": 1024, "climber20": 1024,
                "bridge": 2048, "hetero": 8192}
BASELINE_CONFIG = {"walker5x5": "configs[1]", "pusher": "configs[2]", "thrower": "configs[2]", "obstacle10": "configs[3]",
                   "obstacle20": "configs[3]", "climber20": "configs[3]", "bridge": "mixed FIXED+SOFT terrain (SURVEY 7)", "hetero": "configs[4]"}
SETTLE_DEFAULT = 500          # 500-step rollouts; throughput is quoted on steps 500.. of the rollout
POOL = 64                     # distinct per-step action tensors, cycled (a rollout of any length stays O(POOL) in memory)


def metric_name(workload: str, n: int) -> str:
    return METRIC if workload == HEADLINE and n == 4096 else f"batched env-steps/sec ({workload}, {n} envs per GPU)"


def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


# ------------------------------------------------------------------------------------------------
# CPU arm: the reference engine (or, if it was not built, the oracle port) on the host cores

def cpu_plan(workload: str) -> dict:
    """Bounded CPU sample per workload: Sims per core, settle steps, env-steps per timed sample (about 10-30 s of CPU work per
    build).  The small worlds get the GPU arm's full settle phase; the large ones (35 env-steps/s per core on a 20x20 robot) a
    shorter one, stated in `sample`."""
    if workload in ("walker5x5", "pusher", "thrower", "bridge"):
        return {"sims_per_core": 4, "settle": SETTLE_DEFAULT, "spe": 25}
    if workload in ("obstacle10",):
        return {"sims_per_core": 4, "settle": 120, "spe": 10}
    if workload == "hetero":
        return {"sims_per_core": 4, "settle": 60, "spe": 6}
    return {"sims_per_core": 2, "settle": 60, "spe": 6}


def _cpu_worker(workload: str, n_total: int, core: int, cores: int, sims_per_core: int, settle: int, kind: str, opt: str) -> None:
    """Child process: builds its Sims (environment indices spread evenly over the whole workload), runs the settle phase, then
    serves 'go <n_steps>' requests on stdin: advances every Sim by n_steps env-steps and answers with one JSON line."""
    sims = []
    slots = cores * sims_per_core
    for j in range(sims_per_core):
        i = ((core * sims_per_core + j) * n_total) // slots          # spread over [0, n_total)
        w = workloads.world_for(workload, i)
        if kind == "reference":
            from oracle import ref
            if not sims and opt == "O3":
                ref.use_build("O3")
            s = ref.RefSim(w)
        else:
            from oracle.oracle import OracleSim
            s = OracleSim(w)
        n_act = len(s.actuator_indices("robot"))
        sims.append((s, np.random.default_rng(1234 + i), n_act))

    def advance(n_steps: int):
        acts = [workloads.clipped_actions(rng, (n_steps, n_act)) for _, rng, n_act in sims]
        t0 = time.perf_counter()
        done = resets = 0
        for (s, _, _), a in zip(sims, acts):
            for k in range(n_steps):
                s.set_action_values("robot", a[k])
                if s.step():
                    s.revert(0)                   # unstable: reset to t = 0 and counted
                    resets += 1
                done += 1
        return done, time.perf_counter() - t0, resets

    advance(settle)
    print("ready", flush=True)
    for line in sys.stdin:
        tok = line.split()
        if not tok or tok[0] != "go":
            break
        done, dt, resets = advance(int(tok[1]))
        print(json.dumps({"env_steps": done, "seconds": dt, "resets": resets}), flush=True)


def cpu_kind() -> str:
    from oracle import ref
    return "reference" if ref.available() else "port"


class CpuPool:
    """One process per host core, each owning `sims_per_core` independent Sims sampled across the whole workload.
    sample(n) advances every Sim by n env-steps on all cores at once (aggregate = sum of the per-core rates measured while
    all cores are busy); sample_single(n) runs one process alone (single-thread figure)."""

    def __init__(self, workload: str, n_total: int, opt: str = "O2", cores: int = 0):
        self.workload, self.n_total, self.opt = workload, n_total, opt
        self.cores = cores or host_cores()
        self.plan = cpu_plan(workload)
        self.kind = cpu_kind()
        if self.kind == "port":
            from oracle import oracle
            oracle.build()
            self.opt = "O2"
        else:
            from oracle import ref
            if opt == "O3" and not ref.available("O3"):
                raise FileNotFoundError("oracle/_ref/libmevo_ref_o3.so not built")
        self.procs = []
        for c in range(self.cores):
            cmd = [sys.executable, os.path.abspath(__file__), "--_cpu_worker", workload, str(n_total), str(c), str(self.cores),
                   str(self.plan["sims_per_core"]), str(self.plan["settle"]), self.kind, self.opt]
            self.procs.append(subprocess.Popen(cmd, stdin=subprocess.PIPE, stdout=subprocess.PIPE, text=True, cwd=ROOT))
        for p in self.procs:
            line = p.stdout.readline()
            if line.strip() != "ready":
                raise RuntimeError(f"cpu worker failed to start: {line!r}")

    def _run(self, procs, steps_per_env: int) -> dict:
        t0 = time.perf_counter()
        for p in procs:
            p.stdin.write(f"go {steps_per_env}\n")
            p.stdin.flush()
        rate, total = 0.0, 0
        for p in procs:
            res = json.loads(p.stdout.readline())
            rate += res["env_steps"] / res["seconds"]
            total += res["env_steps"]
        return {"value": rate, "wall_s": time.perf_counter() - t0, "env_steps": total}

    def sample(self, steps_per_env: int = 0) -> dict:
        return self._run(self.procs, steps_per_env or self.plan["spe"])

    def sample_single(self, steps_per_env: int = 0) -> dict:
        return self._run(self.procs[:1], steps_per_env or self.plan["spe"])

    def describe(self, steps_per_env: int = 0) -> str:
        spe = steps_per_env or self.plan["spe"]
        if self.kind == "reference":
            src = "unmodified reference sources, g++ " + ("-O3 -march=x86-64-v3" if self.opt == "O3" else "-O2 (the reference's Release default)")
        else:
            src = "oracle/mevo_oracle.c, gcc -O2"
        return (f"{self.cores} processes x {self.plan['sims_per_core']} Sims of workload {self.workload} (environment indices spread evenly "
                f"over 0..{self.n_total - 1}), {self.plan['settle']} untimed settle env-steps each, then {spe} env-steps per Sim per sample; {src}")

    def close(self) -> None:
        for p in self.procs:
            try:
                p.stdin.write("quit\n"); p.stdin.flush(); p.stdin.close()
                p.wait(timeout=10)
            except Exception:
                p.kill()


def cpu_baseline(workload: str, n_total: int, full: bool) -> dict:
    """The reference engine on this box's host cores.  full: all-core and single-thread, -O2 and -O3 (headline); else all-core -O2."""
    pool = CpuPool(workload, n_total, "O2")
    r = [pool.sample(), pool.sample()]
    out = {"value": 0.5 * (r[0]["value"] + r[1]["value"]), "unit": UNIT, "cores": pool.cores, "kind": pool.kind,
           "sample": pool.describe() + "; mean of 2 samples"}
    if full:
        out["single_thread_O2"] = pool.sample_single()["value"]
    pool.close()
    if full and pool.kind == "reference":
        try:
            p3 = CpuPool(workload, n_total, "O3")
            out["all_cores_O3"] = 0.5 * (p3.sample()["value"] + p3.sample()["value"])
            out["single_thread_O3"] = p3.sample_single()["value"]
            out["O3_flags"] = "-O3 -march=x86-64-v3 (the .so is built off-box, so not -march=native)"
            p3.close()
        except FileNotFoundError:
            out["all_cores_O3"] = None
    return out


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
# the reference arm

def main_reference(args) -> None:
    """--impl reference: the reference's own CPU engine on all host cores (-O2).  A 'step' here is a bounded sample: every Sim
    (several per core, sampled across the workload, settled like the GPU arm's) advances by `spe` env-steps."""
    rank, _, world = dist_env()
    if rank != 0:
        return
    wl = args.workload
    n = args.envs or DEFAULT_ENVS[wl]
    pool = CpuPool(wl, n * max(world, 1), "O2")
    spe = max(2, pool.plan["spe"] // 3)
    for _ in range(args.warmup):
        pool.sample(spe)
    vals, secs = [], 0.0
    for _ in range(args.steps):
        r = pool.sample(spe)
        vals.append(r["value"]); secs += r["wall_s"]
    pool.close()
    value = float(np.mean(vals)) if vals else 0.0
    line = {
        "impl": "reference", "metric": metric_name(wl, n), "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * secs / max(args.steps, 1), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workloads.describe(wl), "workload_name": wl, "baseline_config": BASELINE_CONFIG[wl], "substeps_per_step": SUBSTEPS,
                   "settle_steps": pool.plan["settle"],
                   "parallelism": f"{pool.cores} host processes x {pool.plan['sims_per_core']} Sims, no GPU"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": pool.cores, "kind": pool.kind,
                         "sample": pool.describe(spe) + f"; mean of {args.steps} samples"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# our arm

def algorithmic_bytes_per_step(P: int, E: int) -> int:
    """every array the reference touches once in and once out per substep -- 105 B per point (pos, vel, pos_last
    read+written, mass, fixed) + 32 B per edge (a, b, goal, L0, k) -- x 30 substeps."""
    return SUBSTEPS * (105 * P + 32 * E)


def build_shard(capi, wl: str, n: int, first: int, local: int, count_dynamic: bool):
    """This rank's environments first .. first + n - 1 of workload `wl`.  Returns (batch, sizes): point / edge totals in the
    reference's numbering (terrain included) and of the simulated (non-pruned) objects only."""
    B = capi.Batch(local)
    P = E = Pd = Ed = 0
    for i in range(n):
        env = B.add_world(workloads.world_for(wl, first + i))
        p, e = B.num_points(env), B.num_edges(env)
        P += p; E += e
        if count_dynamic:
            dyn = np.ones(p, dtype=bool)
            for o in range(B.num_objects(env)):
                info = B.object_info(env, o)
                if info["is_static"]:
                    dyn[info["min_point"]:info["max_point"] + 1] = False
            Pd += int(dyn.sum())
            Ed += int(dyn[B.topology(env)["a"]].sum())
    B.finalize()
    B.set_auto_reset(True)                      # unstable envs are reverted to t = 0 and counted
    return B, {"P": P, "E": E, "Pd": Pd, "Ed": Ed}


class Rollout:
    """Device-resident stepping of one batch with a cyclic pool of action tensors already in HBM."""

    def __init__(self, torch, B, dev, seed: int):
        self.torch, self.B, self.dev = torch, B, dev
        self.n, self.stride = B.num_envs, B.action_stride
        rng = np.random.default_rng(seed)
        self.acts_host = torch.from_numpy(workloads.clipped_actions(rng, (POOL, self.n, self.stride))).pin_memory()
        self.acts_dev = self.acts_host.to(dev)
        self.stream = torch.cuda.ExternalStream(B.cuda_stream, device=dev)
        self.t = 0
        self.launches = 0

    def step(self) -> None:
        self.B.bind_actions_device(self.acts_dev[self.t % POOL].data_ptr(), self.stride)
        self.B.step_async(1, True)
        self.t += 1
        self.launches += self.B.last_launches

    def run(self, k: int) -> None:
        for _ in range(k):
            self.step()

    def timed(self, k: int, flush) -> float:
        """k steps, CUDA events on the engine's stream around each step, L2 flushed between steps outside the events; ms summed"""
        torch = self.torch
        ev0 = [torch.cuda.Event(enable_timing=True) for _ in range(k)]
        ev1 = [torch.cuda.Event(enable_timing=True) for _ in range(k)]
        for s in range(k):
            if flush is not None:
                with torch.cuda.stream(self.stream):
                    flush.zero_()
            ev0[s].record(self.stream)
            self.step()
            ev1[s].record(self.stream)
        self.B.sync()
        return sum(a.elapsed_time(b) for a, b in zip(ev0, ev1))

    def e2e(self, k: int, w: int, flush, flags_host, com_host) -> float:
        """k steps through the C ABI with HOST buffers: pinned actions in (H2D on the engine's stream), flags + robot COM out
        (D2H, one synchronisation); seconds of host wall clock over the k steps (w untimed steps first)"""
        torch, B = self.torch, self.B
        B.bind_actions_device(0, 0)
        total = 0.0
        for s in range(w + k):
            if flush is not None:
                with torch.cuda.stream(self.stream):
                    flush.zero_()
            B.sync()
            t0 = time.perf_counter()
            B.set_actions_ptr(self.acts_host[self.t % POOL].data_ptr(), self.stride)
            B.step_async(1, True)
            B.fetch_step_results_ptr(flags_host.data_ptr(), 0, com_host.data_ptr())
            dt = time.perf_counter() - t0
            self.t += 1
            self.launches += B.last_launches + 1
            if s >= w:
                total += dt
        return total


def ncu_summary(kernel_key: str):
    """Measured per-kernel figures of the committed ncu capture (profiles/ncu_summary.json, written by tools/ncu_summary.py from
    an `ncu --set full` report of this kernel; each entry names its report, workload and commit)."""
    path = os.path.join(ROOT, "profiles", "ncu_summary.json")
    try:
        with open(path) as f:
            return json.load(f)["kernels"].get(kernel_key)
    except (OSError, ValueError, KeyError):
        return None


KERNEL_OF = {"walker5x5": "wstep_kernel<40,112,24,192,256,112,16,1,true>", "pusher": "wstep_kernel<52,144,28,320,288,64,12,1,true>",
             "thrower": "wstep_kernel<52,144,28,320,288,64,12,1,true>", "obstacle10": "gstep_kernel<64,8,2,false>",
             "obstacle20": "gstep_kernel<256,2,2,false>", "climber20": "gstep_kernel<256,2,2,false>", "bridge": "gstep_kernel<256,2,2,false>",
             "hetero": "gstep_kernel<256,2,2,false>"}


def short_config_run(torch, capi, dev, wl: str, n: int, rank: int, world: int, local: int, flush, with_cpu: bool, settle: int, K: int):
    """One entry of the headline line's `configs` block: a short steady-state run of another BASELINE.json configuration."""
    import torch.distributed as dist
    t0 = time.perf_counter()
    B, _ = build_shard(capi, wl, n, rank * n, local, count_dynamic=False)
    build_s = time.perf_counter() - t0
    ro = Rollout(torch, B, dev, 4321 + rank)
    ro.run(settle + 3)
    B.sync()
    ms = ro.timed(K, flush)
    events = B.unstable_events
    B.bind_actions_device(0, 0)
    B.close()
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    out = {"baseline_config": BASELINE_CONFIG[wl], "envs_per_gpu": n, "n_gpus": world, "value": world * n * K / (ms * 1e-3), "unit": UNIT,
           "ms_per_step": ms / K, "steps": K, "settle_steps": settle, "unstable_env_steps": int(events), "build_s": round(build_s, 2),
           "kernel": KERNEL_OF.get(wl)}
    if with_cpu and rank == 0:
        cpu = cpu_baseline(wl, n * world, full=False)
        out["cpu_baseline"] = cpu
        out["x_cpu_all_cores"] = out["value"] / cpu["value"] if cpu["value"] else None
    return out


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
    wl = args.workload
    n, K, W = args.envs or DEFAULT_ENVS[wl], args.steps, max(args.warmup, 0)
    settle = args.settle

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(vals):
        t = torch.tensor(vals, dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return [float(x) for x in t.tolist()]

    # ---- build this rank's shard of the workload (host, untimed)
    t_build = time.perf_counter()
    B, sz = build_shard(capi, wl, n, rank * n, local, count_dynamic=True)
    t_build = time.perf_counter() - t_build
    ro = Rollout(torch, B, dev, 1234 + rank)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)       # > 126 MB L2
    clocks = ClockSampler(local) if rank == 0 else None

    # ---- (0) first K steps after a reset (what round 1 reported): W warm-up steps, K timed
    ro.run(W)
    barrier()
    fresh_ms = ro.timed(K, flush)

    # ---- (1) steady state: the rollout continues untimed up to step `settle`, W warm-up steps, then K timed steps
    ro.run(max(0, settle - ro.t))
    ro.run(W)
    barrier()
    launches0 = ro.launches
    if clocks:
        clocks.mark_start()
    t_wall = time.perf_counter()
    dev_ms = ro.timed(K, flush)
    barrier()
    t_wall = time.perf_counter() - t_wall
    if clocks:
        clocks.mark_end()
    clock_info = clocks.stop() if clocks else None
    launches = ro.launches - launches0
    steady_from = ro.t - K
    events = B.unstable_events

    # ---- (2) end to end through the C ABI: pinned host actions in, host flags + reward signal (robot COM) out, every step
    com_host = torch.zeros((n, 2), dtype=torch.float64).pin_memory()
    flags_host = torch.zeros((n,), dtype=torch.uint8).pin_memory()
    e2e_s = ro.e2e(K, W, flush, flags_host, com_host)
    h2d, d2h = n * ro.stride * 8, n * 2 * 8 + n

    # ---- (3) rollout end: the only collective, timed warm: per-env reward rows and a modular-observation-sized block
    gather_ms = None
    com_dev = torch.zeros((n, 2), dtype=torch.float64, device=dev)
    B.object_com_device(0, com_dev.data_ptr())
    B.sync()
    if world > 1:
        from modularevogym_b200.sharding import gather_env_rows
        obs_dev = torch.zeros((n, 9 * 25), dtype=torch.float32, device=dev)      # ModularEvoGym observation of a 5x5 robot: 9 floats per voxel
        gather_env_rows(com_dev, world * n)                                        # throw-away: NCCL's lazy communicator set-up
        torch.cuda.synchronize()
        gms = []
        for tensor in (com_dev, obs_dev):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            gather_env_rows(tensor, world * n)                                     # throw-away at THIS size (output allocation, NCCL buffers)
            torch.cuda.synchronize()
            dist.barrier()
            e0.record()
            for _ in range(5):
                g = gather_env_rows(tensor, world * n)
            e1.record()
            torch.cuda.synchronize()
            assert g.shape[0] == world * n
            gms.append(e0.elapsed_time(e1) / 5)
        gms = max_over_ranks(gms)
        gather_ms = {"reward_rows_f64x2": gms[0], "modular_obs_f32x225": gms[1],
                     "note": "one all_gather_into_tensor over NCCL per rollout, warm, max over ranks; not inside `value`"}

    dev_ms, fresh_ms, e2e_ms, wall_ms = max_over_ranks([dev_ms, fresh_ms, e2e_s * 1e3, t_wall * 1e3])
    total_env_steps = world * n * K
    value = total_env_steps / (dev_ms * 1e-3)
    B.close()
    del ro

    # ---- (4) the other BASELINE.json configurations, short runs (rank 0 adds their CPU baselines)
    configs = None
    if not args.no_configs and wl == HEADLINE:
        configs = {}
        names = ["pusher", "thrower", "obstacle10", "obstacle20", "climber20", "hetero"] if world == 1 else ["hetero"]
        for name in names:
            configs[name] = short_config_run(torch, capi, dev, name, DEFAULT_ENVS[name], rank, world, local, flush,
                                             with_cpu=not args.no_cpu_baseline, settle=100, K=10)

    if rank == 0:
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(peaks_path):
            peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        else:
            peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
        step_s = dev_ms * 1e-3 / K
        alg_bytes = algorithmic_bytes_per_step(sz["P"], sz["E"])              # per launch: this rank's n envs, one env-step
        robot_bytes = algorithmic_bytes_per_step(sz["Pd"], sz["Ed"])          # simulated objects only (all-FIXED terrain pruned)
        achieved = alg_bytes / step_s / 1e9
        kernel = KERNEL_OF.get(wl)
        ncu = ncu_summary(kernel) if kernel else None
        roofline = {
            "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
            "traffic": ncu.get("dram_bytes_per_launch") if ncu and ncu.get("envs") == n else None,
            "kernel": "mevo::" + kernel + (" (largest class; the mix launches one kernel per size class)" if wl == "hetero" else ""),
            "peak_source": peak_src, "algorithmic_bytes_per_launch": alg_bytes,
            "launch_unit": ("one env-step (30 substeps) of this rank's environments; the engine may issue it as chunks of substeps -- warp "
                            "classes: (CTA-group, chunk) items of ONE kernel launch, group classes: several launches on two streams (DESIGN 4, 10); "
                            "`gpu_launches` counts the kernel launches"),
            "note": ("`achieved` is SURVEY 8d's streaming model (105 B/point + 32 B/edge per substep in the reference's numbering, FIXED "
                     "terrain included) / measured step time: an equivalent bandwidth, comparable across designs.  The engine keeps the "
                     "state in shared memory for all 30 substeps and never integrates all-FIXED terrain, so DRAM is nearly idle and the "
                     "kernel is bound by dependent FP64 latency chains and barriers, not by HBM: see robot_only_frac, dram_frac, "
                     "fp64_pipe_frac, issue_frac, stall shares"),
            "bound_observed": "latency / issue (FP64 sqrt-div chains, shared-memory round trips, barriers); DRAM < 1 % of peak",
            "robot_only_bytes_per_launch": robot_bytes, "robot_only_achieved": robot_bytes / step_s / 1e9,
            "robot_only_frac": robot_bytes / step_s / 1e9 / peak,
        }
        if ncu:
            roofline.update({
                "dram_frac": (ncu["dram_bytes_per_launch"] / (ncu["duration_ms"] * 1e-3) / 1e9 / peak) if ncu.get("duration_ms") else None,
                "fp64_pipe_frac": ncu.get("fp64_pipe_frac"), "issue_frac": ncu.get("issue_frac"),
                "stall_shares": ncu.get("stall_shares"), "local_load_store_inst": ncu.get("local_ld_st"),
                "shared_ld_conflict_share": ncu.get("shared_ld_conflict_share"), "occupancy_warps_per_sm": ncu.get("warps_per_sm"),
                "ncu_profile": ncu.get("profile"), "ncu_profile_commit": ncu.get("commit"), "ncu_profile_workload": ncu.get("workload"),
            })
        cpu = None
        if not args.no_cpu_baseline:
            cpu = cpu_baseline(wl, n * world, full=True)
        line = {
            "metric": metric_name(wl, n), "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": dev_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "value_after_reset": total_env_steps / (fresh_ms * 1e-3),
            "config": {"workload": workloads.describe(wl), "workload_name": wl, "baseline_config": BASELINE_CONFIG[wl], "envs_per_gpu": n,
                       "substeps_per_step": SUBSTEPS, "points_per_env": sz["P"] / n, "edges_per_env": sz["E"] / n,
                       "simulated_points_per_env": sz["Pd"] / n, "simulated_edges_per_env": sz["Ed"] / n,
                       "settle_steps": settle, "timed_rollout_steps": [steady_from, steady_from + K],
                       "regime": (f"`value`, `e2e`: steady state, rollout steps {steady_from}..{steady_from + K} after a reset (auto-reset of unstable "
                                  f"environments on); `value_after_reset`: steps {W}..{W + K} after the reset"),
                       "parallelism": f"env-sharded x{world}, no collective on the step path",
                       "l2": "256 MB buffer written between timed iterations (outside the events)",
                       "timing": "CUDA events on the engine's stream around each step, summed; max over ranks",
                       "e2e_returns": "per step: unstable flag (1 B) + robot centre of mass (2 x f64) per env -- the reward signal; observations stay on the device",
                       "wall_ms_incl_flush": wall_ms, "build_s": t_build},
            "e2e": {"value": total_env_steps / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": e2e_ms / K},
            "gpu_launches": launches,
            "roofline": roofline,
            "cpu_baseline": cpu,
            "clocks": clock_info,
            "unstable_env_steps": int(events),
            "gather_ms": gather_ms,
            "configs": configs,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main() -> None:
    if len(sys.argv) > 1 and sys.argv[1] == "--_cpu_worker":
        a = sys.argv[2:]
        _cpu_worker(a[0], int(a[1]), int(a[2]), int(a[3]), int(a[4]), int(a[5]), a[6], a[7])
        return
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--envs", type=int, default=0, help="environments per GPU (default: the workload's BASELINE.json size)")
    ap.add_argument("--impl", choices=("ours", "reference"), default="ours")
    ap.add_argument("--workload", choices=tuple(DEFAULT_ENVS), default=HEADLINE,
                    help="walker5x5 = the headline (BASELINE.json configs[1]); the others = configs[2..4] (modularevogym_b200/workloads.py)")
    ap.add_argument("--settle", type=int, default=SETTLE_DEFAULT, help="untimed env-steps after the reset before the steady-state measurement")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the short runs of the other BASELINE.json configurations")
    args = ap.parse_args()
    if args.impl == "reference":
        main_reference(args)
    else:
        main_ours(args)


if __name__ == "__main__":
    main()
