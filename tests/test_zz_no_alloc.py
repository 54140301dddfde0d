"""GPU: the allocation-free stepping entry point ``mevo_batch_step_static_scratch`` (include/mevo_b200.h) against the
existing ``mevo_batch_step_async``, through the C ABI.

The entry point has no counterpart in the reference (it was added for callers that cannot
call cudaMalloc); its neighbour-list scratch is a file-scope ``static __device__`` array of MEVO_STATIC_SCRATCH_BYTES that every
call shares.  Collected last on purpose (file name).

Two identical heterogeneous batches (both kernels, every group size) are stepped side by side, one through each entry point,
with the same device-resident actions, TWO calls in a row (the second call finds the first call's contents in the shared
array); states and unstable flags must agree within the teacher-forced tolerances of tests/test_gpu_parity.py (they are in
fact bit-identical: same kernels, same launch geometry, deterministic).
"""
import numpy as np
import pytest

from helpers import rel_err
from modularevogym_b200 import capi, workloads

pytestmark = pytest.mark.gpu

POS_TOL, VEL_TOL, VT_TOL = 1e-14, 1e-11, 1e-9       # tests/test_gpu_parity.py: TF_POS_TOL, TF_VEL_TOL, TF_VT_TOL
STATIC_SCRATCH_BYTES = 32 << 20                      # MEVO_STATIC_SCRATCH_BYTES

CASES = [("walker5x5", 0), ("walker5x5", 1), ("pusher", 2), ("thrower", 3), ("bridge", 4), ("obstacle7", 5), ("obstacle7", 6),
         ("obstacle10", 7), ("obstacle10", 8), ("obstacle10", 9), ("obstacle15", 10), ("obstacle20", 11), ("climber12", 12)]


def build_batch():
    B = capi.Batch(0)
    for name, i in CASES:
        B.add_world(workloads.world_for(name, i))
    B.finalize()
    return B


def test_static_scratch_step_matches_step_async():
    import torch
    dev = torch.device("cuda", 0)
    A, B = build_batch(), build_batch()
    n, stride = A.num_envs, A.action_stride
    assert B.action_stride == stride
    rng = np.random.default_rng(7)
    flags_b = torch.full((n,), 255, dtype=torch.uint8, device=dev)          # caller-owned output of the new entry point
    worst = [0.0, 0.0, 0.0]
    exact = True
    for call in range(2):                                                   # two calls in a row share the static array
        acts = torch.from_numpy(workloads.clipped_actions(rng, (n, stride))).to(dev)
        torch.cuda.synchronize()
        A.bind_actions_device(acts.data_ptr(), stride)
        A.step_async(2, True)
        A.sync()
        flags_a = np.zeros(n, dtype=np.uint8)
        A.fetch_step_results_ptr(flags_a.ctypes.data, 0, 0)
        B.step_static_scratch(2, acts.data_ptr(), stride, flags_b.data_ptr())
        B.sync()
        assert B.last_launches == A.last_launches >= 2                      # warp kernel + group kernel classes all ran
        assert np.array_equal(flags_b.cpu().numpy(), flags_a), call
        for i in range(n):
            sa, sb = A.get_state(i), B.get_state(i)
            e = (rel_err(sb[0], sa[0]), rel_err(sb[1], sa[1]), rel_err(sb[3], sa[3]))
            assert e[0] <= POS_TOL and e[1] <= VEL_TOL and e[2] <= VT_TOL, (call, CASES[i], e)
            assert np.array_equal(sb[2], sb[0])                             # pos_last = pos after a completed step
            assert A.get_time(i) == B.get_time(i)
            worst = [max(w, v) for w, v in zip(worst, e)]
            exact &= all(np.array_equal(x, y) for x, y in zip(sa, sb))
    print(f"\n[no-alloc] 2 calls x 2 env-steps x {n} envs: worst rel err pos/vel/vel_true {worst}, bit-identical: {exact}")
    # apply_actions = 0 form (NULL actions) and the batch's own flag buffer (NULL output)
    A.bind_actions_device(0, 0)
    A.step_async(1, False)
    B.step_static_scratch(1, 0, 0, 0)
    A.sync(); B.sync()
    for i in range(n):
        assert rel_err(B.get_state(i)[0], A.get_state(i)[0]) <= POS_TOL
    A.close(); B.close()


def test_static_scratch_step_rejects_what_it_cannot_hold():
    # (a) a batch that was never finalized: finalize allocates, the entry point must not
    U = capi.Batch(0)
    U.add_world(workloads.world_for("obstacle10", 0))
    with pytest.raises(capi.MevoError, match="finalize"):
        U.step_static_scratch(1, 0, 0, 0)
    U.close()
    # (b) more neighbour-list scratch than the static array holds: ~33 KB per 20x20 environment -> 1200 of them exceed 32 MiB
    w = workloads.world_for("obstacle20", 0)
    Big = capi.Batch(0)
    for _ in range(1200):
        Big.add_world(w)
    Big.finalize()
    with pytest.raises(capi.MevoError, match="MEVO_STATIC_SCRATCH_BYTES"):
        Big.step_static_scratch(1, 0, 0, 0)
    Big.step_async(1, False)                                                # the allocating entry point still serves it
    Big.sync()
    Big.close()
