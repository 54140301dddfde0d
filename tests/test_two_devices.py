"""GPU (needs two devices; skipped otherwise): batches on two GPUs inside ONE process.  Kernel attributes (the opt-in dynamic
shared-memory size) are per device: every Batch::finalize configures the kernels it launches on its own device, so a second
batch on another GPU must step -- and give bit-identical results to the first, both kernels."""
import numpy as np
import pytest

from modularevogym_b200 import capi, workloads

pytestmark = pytest.mark.gpu

CASES = [("walker5x5", 0), ("pusher", 1), ("obstacle10", 2), ("obstacle20", 3), ("bridge", 4)]


def test_batches_on_two_devices_in_one_process():
    if capi.load().mevo_device_count() < 2:
        pytest.skip("one CUDA device visible")
    worlds = [workloads.world_for(n, i) for n, i in CASES]
    batches = []
    for dev in (0, 1, 0):                       # a third batch back on device 0 after device 1 was configured
        B = capi.Batch(dev)
        for w in worlds:
            B.add_world(w)
        B.finalize()
        batches.append(B)
    rng = np.random.default_rng(1)
    for step in range(3):
        acts = workloads.clipped_actions(rng, (len(CASES), batches[0].action_stride))
        for B in batches:
            B.set_actions(acts)
            B.step_async(1, True)
        for B in batches:
            B.sync()
        for i in range(len(CASES)):
            s0 = batches[0].get_state(i)
            for B in batches[1:]:
                for a, b in zip(s0, B.get_state(i)):
                    assert np.array_equal(a, b), (CASES[i], step, B.device)
    for B in batches:
        B.close()
