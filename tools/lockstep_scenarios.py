"""Dev tool (test infrastructure): lock-step the host compile of the device simulator against the compiled reference on
scripted scenarios and print the divergence per scenario.  python tools/lockstep_scenarios.py [name] [ticks]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import oracle_ref
import hostsim_binding as hb
import lockstep as ls


from scenarios import SCEN, car_at, yaw_rot  # noqa: E402,F401


def run(name, ticks=240, horizon=1, verbose=True):
    """horizon = ticks between re-synchronisations of the simulator under test to the reference state
    (1 = pure per-tick error, 16 = the north-star horizon, 0 = free run)."""
    L = oracle_ref.lib(); L.ref_set_threads(1); L.ref_seed(5)
    env = oracle_ref.RefEnvSet(1, 1)
    sim = hb.HostSim()
    st = env.get_state(0)
    ctl = SCEN[name](st)
    env.set_state(0, st); sim.set_state(st)
    first_bad = None
    worst = {}
    nbad = 0
    for t in range(ticks):
        if horizon and t % horizon == 0:
            sim.set_state(env.get_state(0))
        w, bad = ls.run_lockstep(env, sim, 1, 2, controls_fn=(lambda tt, s: ctl(t, s)) if ctl else None, stop_at=1e9, verbose=False)
        for k, v in w.items():
            worst[k] = max(worst.get(k, 0), v)
        mx = max(v for k, v in w.items() if k.endswith(".pos"))
        if bad is not None:
            nbad += 1
        if first_bad is None and (mx > 1e-3 or bad is not None):
            first_bad = t
            if verbose:
                print("  first bad at tick", t, {k: round(v, 5) for k, v in w.items() if v > 1e-3})
    pos = max(v for k, v in worst.items() if k.endswith(".pos"))
    vel = max(v for k, v in worst.items() if k.endswith(".vel"))
    print("%-12s ticks=%d horizon=%d worst pos %.5f uu vel %.4f uu/s  flag-mismatch ticks %d  first bad: %s" % (name, ticks, horizon, pos, vel, nbad, first_bad))
    env.close()
    return pos, first_bad, nbad


if __name__ == "__main__":
    names = [sys.argv[1]] if len(sys.argv) > 1 and sys.argv[1] != "all" else list(SCEN)
    ticks = int(sys.argv[2]) if len(sys.argv) > 2 else 240
    horizon = int(sys.argv[3]) if len(sys.argv) > 3 else 1
    for n in names:
        run(n, ticks, horizon)
