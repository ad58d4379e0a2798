"""How often solver_order is called and how often it runs (experiment build -DGGL_ORDER_COUNT: the arenas' overflow counters carry
calls << 16 | runs).   GGL_LIB_PATH=reinforcement-learning_b200/libgigaenv_cuda_count.so python tools/order_call_rate.py [ppt] [arenas]"""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

ppt = int(sys.argv[1]) if len(sys.argv) > 1 else 2
arenas = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
ggl = importlib.import_module("reinforcement-learning_b200")
env = ggl.EnvSet(arenas, ppt, obs_kind=ggl.abi.OBS_ADVANCED, seed=123)
acts = torch.zeros(env.num_players, dtype=torch.int32, device="cuda")
for i in range(160):
    env.random_actions(acts, seed=1000 + i)
    env.step(acts)
env.sync()
n0 = env.contact_overflow()
steps = 60
for i in range(steps):
    env.random_actions(acts, seed=5000 + i)
    env.step(acts)
env.sync()
n = env.contact_overflow() - n0
ticks = arenas * steps * 8
print("%dv%d x %d arenas, %d steps: solver_order called in %.4f %% of the arena-ticks, ran in %.4f %% (raw %d)" %
      (ppt, ppt, arenas, steps, 100.0 * (n >> 16) / ticks, 100.0 * (n & 0xFFFF) / ticks, n))
