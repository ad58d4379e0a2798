"""Divergence curves of the CUDA tick against the compiled reference (north star: per-tick parity within 1e-3 uu over a short
horizon, divergence curves reported beyond it). Every scenario starts from one injected state on both sides and then runs FREE
(no re-synchronisation) for 480 ticks = 4 s; printed is the largest position / velocity difference of any body at tick
1, 2, 4, ... 480 and whether any integer flag (on-ground, wheel contacts, demoed, pads ...) differs at that tick.
Runs on the GPU box:   python tools/divergence_curves.py > profiles/<tag>_divergence_curves.txt"""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import lockstep as ls     # noqa: E402
import oracle_ref         # noqa: E402
import scenarios          # noqa: E402
import test_env_gpu as T  # noqa: E402  (scenario tables and the one-arena adapter)

MARKS = [1, 2, 4, 8, 16, 32, 64, 120, 240, 480]


def run(ggl, name, init, ctl, scen):
    L = oracle_ref.lib(); L.ref_set_threads(1); L.ref_seed(7)
    ref = oracle_ref.RefEnvSet(1, 1)
    sim = T._DevSim(ggl)
    st = ref.get_state(0)
    if scen is not None:
        ctl = scen(st)
    elif init:
        init(st)
    ref.set_state(0, st); sim.set_state(st)
    out = []
    t = 0
    for m in MARKS:
        worst, flags = {}, 0
        while t < m:
            if ctl:
                cur = ref.get_state(0)
                ctl(t, cur)
                ref.set_controls(0, cur); sim.set_controls(cur)
            ref.arena_step(1, 0); sim.step(1)
            t += 1
        d, bad = ls.diff_states(ref.get_state(0), sim.get_state(), 2)
        pos = max(v for k, v in d.items() if k.endswith(".pos"))
        vel = max(v for k, v in d.items() if k.endswith(".vel"))
        out.append((m, pos, vel, len(bad)))
    ref.close(); sim.env.close()
    print("%-12s " % name + "  ".join("%9.2e" % p for _, p, _, _ in out))
    print("%-12s " % "  vel uu/s" + "  ".join("%9.2e" % v for _, _, v, _ in out))
    print("%-12s " % "  flags" + "  ".join("%9d" % b for _, _, _, b in out))


def main():
    ggl = importlib.import_module("reinforcement-learning_b200")
    print("largest position difference of any body (uu) against the compiled reference, free-running from one injected state")
    print("%-12s " % "tick" + "  ".join("%9d" % m for m in MARKS))
    for name, init, ctl in T._scenario_states():
        run(ggl, name, init, ctl, None)
    for name in sorted(scenarios.SCEN):
        run(ggl, name, None, None, scenarios.SCEN[name])


if __name__ == "__main__":
    main()
