"""GPU: SURVEY 8f-3 for real -- the reference's own gym environments (`gym.make(id, body=..., mode='modular')`, unmodified Python
from oracle/_ref/pyref) running on this repo's `simulator_cpp.Sim`, replaying the committed environment fixtures step by step
and the reference's smoke script (`examples/modularevogym_test.py:1-37`).  The work is in tools/gym_e2e.py; it runs in a
subprocess because it rewires `sys.modules` (evogym.simulator_cpp, gym).  Skipped where the reference's Python layer was
not staged (`make -C oracle pyref` needs /root/reference)."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PYREF = os.path.join(ROOT, "oracle", "_ref", "pyref")

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not os.path.isdir(os.path.join(PYREF, "evogym")), reason="oracle/_ref/pyref not staged")]


def test_reference_gym_envs_run_on_this_engine():
    out_dir = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out_dir, exist_ok=True)
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "gym_e2e.py"), "--pyref", PYREF, "--out", os.path.join(out_dir, "gym_e2e.jsonl")],
                         capture_output=True, text=True, timeout=900, cwd=ROOT)
    print(out.stdout[-6000:])
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    lines = [json.loads(l) for l in out.stdout.splitlines() if l.startswith("{")]
    assert lines[-1]["summary"] == "ok"
    assert lines[-1]["fixtures"] >= 41
    assert sum(1 for l in lines if "smoke" in l) == 32
