"""Where the host-buffer (e2e) step spends its time: python tools/e2e_breakdown.py"""
import importlib, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
ggl = importlib.import_module("reinforcement-learning_b200")
env = ggl.EnvSet(4096, 2, obs_kind=ggl.abi.OBS_ADVANCED, seed=123)
acts = torch.zeros(env.num_players, dtype=torch.int32, device="cuda")
for i in range(160):
    env.random_actions(acts, seed=i); env.step(acts)
torch.cuda.synchronize()
h_act = torch.zeros(env.num_players, dtype=torch.int32).pin_memory()
h_obs = torch.empty(env.num_players, env.obs_size).pin_memory()
h_mask = torch.empty(env.num_players, env.num_actions, dtype=torch.uint8).pin_memory()
h_rew = torch.empty(env.num_players).pin_memory()
h_term = torch.empty(env.num_arenas, dtype=torch.uint8).pin_memory()
env.step_host(h_act, h_obs, h_mask, h_rew, h_term)
def T(f, n=30):
    f(); t0 = time.perf_counter()
    for _ in range(n): f()
    return (time.perf_counter() - t0) / n * 1e3
k = [0]
def pick(): k[0] += 1; env.host_random_actions(h_mask, h_act, k[0])
print("host pick                 %.3f ms" % T(pick))
print("step_host, no outputs     %.3f ms" % T(lambda: env.step_host(h_act)))
print("step_host, all outputs    %.3f ms" % T(lambda: env.step_host(h_act, h_obs, h_mask, h_rew, h_term)))
def dev(): env.step(acts); torch.cuda.synchronize()
print("device step + sync        %.3f ms" % T(dev))
d_obs = env.obs
def d2h(): h_obs.copy_(d_obs, non_blocking=True); torch.cuda.synchronize()
print("obs D2H (10.9 MB) alone   %.3f ms" % T(d2h))
def full(): pick(); env.step_host(h_act, h_obs, h_mask, h_rew, h_term)
print("pick + step_host          %.3f ms" % T(full))
