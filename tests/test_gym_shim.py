"""gym_shim: the reference's Python layer imports on top of this engine's `simulator_cpp` under numpy 2.x
without gym.  Import-level only here (no GPU in this container, no reference checkout on the GPU box); the numbers the
reference's environments produce are pinned by tests/golden/env_*.npz (test_task_layer.py)."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"


def _run(code: str) -> str:
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""), PYTHONWARNINGS="ignore")
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    return out.stdout


def test_minimal_gym_registry_and_time_limit():
    out = _run("""
import numpy as np, sys, types
from modularevogym_b200 import gym_shim
print(gym_shim.install(simulator=False, force_gym=True))
import gym
from gym import spaces
from gym.envs.registration import register
class Count(gym.Env):
    def __init__(self, k=1):
        self.k, self.t = k, 0
        self.action_space = spaces.Box(low=0.6, high=1.6, shape=(3,), dtype=np.float)
        self.observation_space = spaces.Dict({'a': spaces.Box(-1.0, 1.0, shape=(2,))})
    def reset(self):
        self.t = 0
        return 0
    def step(self, a):
        self.t += self.k
        return self.t, 0.0, False, {}
m = types.ModuleType('countmod'); m.Count = Count; sys.modules['countmod'] = m
register(id='Count-v0', entry_point='countmod:Count', max_episode_steps=3)
env = gym.make('Count-v0', k=2)
env.reset()
res = [env.step(None) for _ in range(3)]
print([r[2] for r in res], res[-1][3], res[-1][0], env.action_space.shape, env.unwrapped.k)
""")
    assert "'gym': True" in out
    assert "[False, False, True] {'TimeLimit.truncated': True} 6 (3,) 2" in out


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "evogym")), reason="reference checkout not present")
def test_reference_python_layer_imports_on_this_engine():
    """evogym, evogym.envs and gym.make's registry come up with evogym.simulator_cpp = this repo's pybind11 module; EvoSim
    subclasses its Sim (evogym/sim.py:18)."""
    out = _run(f"""
import sys
from modularevogym_b200 import gym_shim, simulator_cpp
print(gym_shim.install())
sys.path.insert(0, {REF!r})
import evogym, evogym.envs, gym
from evogym.envs.walk import WalkingFlat
print(issubclass(evogym.EvoSim, simulator_cpp.Sim), sys.modules['evogym.simulator_cpp'] is simulator_cpp)
print('Walker-v0' in gym.envs.registration.registry, 'Thrower-v0' in gym.envs.registration.registry)
w = evogym.EvoWorld.from_json({os.path.join(REF, 'evogym/envs/sim_files/Walker-v0.json')!r})
print(w.grid_size.x, w.grid_size.y)
""")
    lines = out.strip().splitlines()
    assert lines[-3] == "True True" and lines[-2] == "True True" and lines[-1] == "100 1"
