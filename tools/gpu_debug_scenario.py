"""Dev tool: run one scripted scenario on the GPU box against the reference, printing state diffs and both contact lists
around a tick window.  python tools/gpu_debug_scenario.py name t0 t1"""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import oracle_ref
import lockstep as ls
import scenarios

ggl = importlib.import_module("reinforcement-learning_b200")
name, t0, t1 = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
L = oracle_ref.lib(); L.ref_set_threads(1); L.ref_seed(5)
ref = oracle_ref.RefEnvSet(1, 1)
env = ggl.EnvSet(1, 1, state_setter=ggl.abi.SETTER_INJECTED)


class Sim:
    def set_state(self, st): env.set_state(0, st)
    def get_state(self): return env.get_state(0)
    def set_controls(self, st):
        cur = env.get_state(0)
        for i in range(cur.num_cars):
            cur.cars[i].controls = st.cars[i].controls
        env.set_state(0, cur)
    def step(self, n): env.arena_step(n)


sim = Sim()
st = ref.get_state(0)
ctl = scenarios.SCEN[name](st)
ref.set_state(0, st); sim.set_state(st)
np.set_printoptions(precision=5, suppress=True, linewidth=220)
for t in range(t1):
    sim.set_state(ref.get_state(0))
    w, bad = ls.run_lockstep(ref, sim, 1, 2, controls_fn=(lambda tt, s: ctl(t, s)) if ctl else None, stop_at=1e9, verbose=False)
    if t >= t0:
        print("tick", t, {k: round(v, 5) for k, v in w.items() if v > 1e-4}, bad)
        print(" ref contacts\n", ls.ref_contacts(ref)[:, :12])
        print(" dev contacts\n", env.get_contacts(0)[:, :12])
