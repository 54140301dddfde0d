"""GPU: an env-step issued in chunks of consecutive substeps -- group classes: several launches on several streams per class
(``Batch::step_async``: MEVO_CHUNKS x MEVO_LANES); warp classes: one CTA per (CTA-group, chunk) item drawn from
a counter (MEVO_WCHUNKS) -- leaves exactly what the single launch leaves: states bit for
bit, unstable flags, step counters, the reset of an environment whose step aborted in an early chunk (Sim.cpp:105-115)."""
import numpy as np
import pytest

from modularevogym_b200 import capi, workloads

pytestmark = pytest.mark.gpu

N_STEPS = 4


def _run(monkeypatch, chunks: int, lanes: int, auto_reset: bool):
    monkeypatch.setenv("MEVO_CHUNKS", str(chunks))
    monkeypatch.setenv("MEVO_WCHUNKS", str(chunks + 1 if chunks > 1 else 1))
    monkeypatch.setenv("MEVO_LANES", str(lanes))
    monkeypatch.setenv("MEVO_CHUNK_MIN_WAVES", "0")          # chunk although these few environments fit the device at once
    B = capi.Batch(0)
    specs = [("obstacle20", i) for i in range(6)] + [("obstacle10", i) for i in range(8)] + [("bridge", i) for i in range(5)] + \
            [("hetero", i) for i in range(24)] + [("walker5x5", i) for i in range(40)] + [("pusher", i) for i in range(30)]
    for name, i in specs:
        B.add_world(workloads.world_for(name, i))
    B.finalize()
    B.set_auto_reset(auto_reset)
    n = B.num_envs
    rng = np.random.default_rng(5)
    for i in range(0, n, 3):                                 # every third environment folds onto itself: at once (positions
        pos, vel, last, vt = B.get_state(i)                  # squeezed), or some substeps into the step (inward velocities)
        rob = B.object_info(i, 0)
        sl = slice(rob["min_point"], rob["max_point"] + 1)
        c = pos[sl].mean(axis=0)
        if i % 2 == 0:
            pos[sl] = c + 0.3 * (pos[sl] - c)
            B.set_state(i, pos=pos, pos_last=pos)
        else:
            vel[sl] = -(pos[sl] - c) * (250.0 + 150.0 * (i % 5))
            B.set_state(i, vel=vel)
    acts = workloads.clipped_actions(np.random.default_rng(6), (N_STEPS, n, B.action_stride))
    out = []
    for s in range(N_STEPS):
        B.set_actions(acts[s])
        _, flags = B.step(1, True)
        out.append((flags.copy(), [B.get_time(i) for i in range(n)], [B.get_state(i) for i in range(n)], B.last_launches))
    ev = B.unstable_events
    B.close()
    return out, ev


@pytest.mark.parametrize("auto_reset", [False, True])
def test_chunked_step_equals_single_launch(monkeypatch, auto_reset):
    base, ev0 = _run(monkeypatch, 1, 1, auto_reset)
    assert ev0 > 0 and any(f.any() for f, _, _, _ in base), "the test wants steps that abort"
    assert any((~f.astype(bool)).any() for f, _, _, _ in base)
    for chunks, lanes in ((3, 2), (5, 4), (2, 1)):
        got, ev = _run(monkeypatch, chunks, lanes, auto_reset)
        assert ev == ev0
        assert got[0][3] > base[0][3], "more launches than the single-launch run"
        for s in range(N_STEPS):
            assert np.array_equal(got[s][0], base[s][0]), (chunks, lanes, s, "unstable flags")
            assert got[s][1] == base[s][1], (chunks, lanes, s, "sim_time")
            for i, (a, b) in enumerate(zip(got[s][2], base[s][2])):
                for x, y in zip(a, b):
                    assert np.array_equal(x, y, equal_nan=True), (chunks, lanes, s, i)


def test_chunked_step_in_a_cuda_graph(monkeypatch):
    """the vec-env front captures the chunked, laned launches (fork / join over the lane streams) in its CUDA graph: the replay
    equals the eager single-launch stepping bit for bit"""
    import random
    import torch
    from modularevogym_b200.vec_env import BatchedEvoGymEnv
    from modularevogym_b200.world import sample_large_robot
    bodies, conns = [], []
    for s in range(6):
        random.seed(170 + s)
        b, c = sample_large_robot((10, 10)) if s < 4 else sample_large_robot((16, 16))
        bodies.append(b); conns.append(c)
    envs = []
    for chunks, graph in ((3, True), (1, False)):
        monkeypatch.setenv("MEVO_CHUNKS", str(chunks))
        monkeypatch.setenv("MEVO_CHUNK_MIN_WAVES", "0")
        envs.append(BatchedEvoGymEnv("Walker-v0", bodies, conns, mode="modular", out_dtype="float64", use_graph=graph))
    obs = [e.reset() for e in envs]
    assert torch.equal(obs[0]["modular"], obs[1]["modular"])
    gen = torch.Generator(device="cuda").manual_seed(9)
    for step in range(6):
        a = torch.rand((6, envs[0].action_width), device="cuda", generator=gen) * 2 - 1
        out = [e.step(a) for e in envs]
        torch.cuda.synchronize()
        assert torch.equal(out[0][0]["modular"], out[1][0]["modular"]), step
        assert torch.equal(out[0][0]["other"], out[1][0]["other"]) and torch.equal(out[0][1], out[1][1]) and torch.equal(out[0][2], out[1][2])
    assert envs[0]._graph is not None and envs[0].B.last_launches > envs[1].B.last_launches
    for e in envs:
        e.close()


def test_world_without_simulated_points_steps():
    """a world of all-FIXED pieces (nothing to integrate: every copy of the stage-in has size zero) steps and counts time"""
    from modularevogym_b200.world import WorldSpec
    w = WorldSpec(width=12, height=6)
    floor = np.full((1, 12), 5, dtype=np.int64)
    w.add_from_array("floor", floor, 0, 0, np.stack((np.arange(11), np.arange(1, 12))).astype(np.int64))
    B = capi.Batch(0)
    B.add_world(w)
    B.add_world(workloads.world_for("walker5x5", 0))
    B.finalize()
    for s in range(3):
        n, flags = B.step(1, True)
        assert n == 0 and not flags.any()
    assert B.get_time(0) == 3 and B.get_time(1) == 3
    B.close()


def test_two_batches_step_concurrently(monkeypatch):
    """two batches of one GPU enqueue their steps back to back on their own streams: the item-drawing CTAs of one wait only for
    CTAs that are already running, so sharing the SMs with another batch's kernels cannot stall them; results equal the batches
    stepped one after the other"""
    monkeypatch.setenv("MEVO_CHUNK_MIN_WAVES", "0")
    monkeypatch.setenv("MEVO_WCHUNKS", "5")
    monkeypatch.setenv("MEVO_CHUNKS", "3")

    def make(offset):
        B = capi.Batch(0)
        for i in range(600):
            B.add_world(workloads.world_for("walker5x5", offset + i))
        for i in range(40):
            B.add_world(workloads.world_for("obstacle10", offset + i))
        B.finalize()
        return B

    acts = [workloads.clipped_actions(np.random.default_rng(20 + k), (6, 640, 128)) for k in range(2)]
    ref = []
    for k in range(2):                                  # one after the other
        B = make(1000 * k)
        for s in range(6):
            B.set_actions(acts[k][s][:, :B.action_stride])
            B.step(1, True)
        ref.append([B.get_state(i)[0] for i in range(0, 640, 7)])
        B.close()
    Bs = [make(0), make(1000)]
    for s in range(6):                                  # interleaved, no synchronisation between the two
        for k in range(2):
            Bs[k].set_actions(acts[k][s][:, :Bs[k].action_stride])
            Bs[k].step_async(1, True)
    for k in range(2):
        Bs[k].sync()
        got = [Bs[k].get_state(i)[0] for i in range(0, 640, 7)]
        assert all(np.array_equal(a, b, equal_nan=True) for a, b in zip(got, ref[k]))
        Bs[k].close()


def test_build_step_close_cycles_leak_nothing(monkeypatch):
    """design loops rebuild batches all the time: build -> step (item form, chunked lanes) -> close, repeated, leaves the free device
    memory where it was (streams, events, item counters, scratch blocks, snapshot slots are all released)"""
    import torch
    monkeypatch.setenv("MEVO_CHUNK_MIN_WAVES", "0")
    monkeypatch.setenv("MEVO_WCHUNKS", "3")
    monkeypatch.setenv("MEVO_CHUNKS", "2")

    def cycle(k):
        B = capi.Batch(0)
        for i in range(64):
            B.add_world(workloads.world_for("walker5x5", 50 * k + i))
        for i in range(12):
            B.add_world(workloads.world_for("obstacle10", 50 * k + i))
        B.finalize()
        B.step(1, True)
        B.save_snapshot(1)
        B.close()

    cycle(0)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(0)[0]
    for k in range(1, 25):
        cycle(k)
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info(0)[0]
    assert free0 - free1 < 8 << 20, (free0, free1)
