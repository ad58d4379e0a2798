"""How much of the reference's next env step is decided by state the snapshot (ggl_arena_state = RocketSim's CarState / BallState /
BoostPadState + the wheels' hidden fields) does not carry (TEST INFRASTRUCTURE, CPU only).

In steady-state play every arena-step is done three times from the same snapshot and the same actions:
  R   the running reference (which keeps whatever else it remembers: broadphase boxes of the last tick, the order of its pair
      and manifold arrays, ...), exactly the arm the parity tests compare against
  F   a FRESH one-arena reference that is handed the snapshot (same car order)
  D   the host compile of the device code, handed the snapshot
and F and D are both compared with R on the physical state with the obs tolerance translated to state units (1e-4 of the obs
scale: 0.23 uu, 0.23 uu/s, 5.5e-4 rad/s, 1e-4 on the orientation). "F differs from R" is a floor no function of the snapshot can
get under.    python tools/reference_self_consistency.py [ppt] [arenas] [steps] [scripted 0/1]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "tools")]
import numpy as np  # noqa: E402

import classify_divergence as cd  # noqa: E402
import env_parity  # noqa: E402
import hostsim_binding as hb  # noqa: E402
import lockstep as ls  # noqa: E402
import oracle_ref  # noqa: E402
import scenarios  # noqa: E402

abi = oracle_ref.abi
TOL = {".pos": 0.23, ".vel": 0.23, ".ang_vel": 5.5e-4, ".rot": 1e-4}


def outside(d, bad):
    return bool(bad) or any(v > t for k, v in d.items() for s, t in TOL.items() if k.endswith(s))


def main():
    ppt = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    arenas = int(sys.argv[2]) if len(sys.argv) > 2 else 64
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 20
    scripted = (int(sys.argv[4]) if len(sys.argv) > 4 else 1) != 0
    ppa = 2 * ppt
    L = oracle_ref.lib(); L.ref_set_threads(1); L.ref_seed(23)
    ref = oracle_ref.RefEnvSet(arenas, ppt, obs_kind=1, state_setter=abi.SETTER_INJECTED)
    rng = np.random.default_rng(23)
    for _ in range(100):
        ref.step(env_parity.random_valid_actions(rng, ref.outputs()[1]))
    if scripted:
        for a, f in scenarios.env_scripts(arenas, ppt, 2).items():
            st = ref.get_state(a); f(st); ref.set_state(a, st)
    ids_all = ref.slot_info()[0]
    table = oracle_ref.action_table()
    n = nf = nd = both = 0
    for s in range(steps):
        acts = env_parity.random_valid_actions(rng, ref.outputs()[1])
        snaps = [ref.get_state(a) for a in range(arenas)]
        termb = ref.outputs()[3].copy()
        ref.step(acts)
        for a in range(arenas):
            if termb[a]:
                continue
            ids = ids_all[a * ppa:(a + 1) * ppa]
            r1, same = cd.one_arena_ref(ppt, ids)
            if not same:
                r1.close(); continue
            r1.arena_step(1, 0)
            sim = hb.HostSim()
            st0 = abi.ArenaState.from_buffer_copy(bytes(snaps[a]))
            r1.set_state(0, st0); sim.set_state(st0); sim.set_slot_order(ids)
            r1.arena_step(7, 0); sim.step(7)
            st = r1.get_state(0)
            for slot in range(ppa):
                row = table[acts[a * ppa + slot]]; ci = ids[slot] - 1
                ls.set_controls(st, ci, throttle=row[0], steer=row[1], pitch=row[2], yaw=row[3], roll=row[4], jump=int(row[5]),
                                boost=int(row[6]), handbrake=int(row[7]))
            r1.set_controls(0, st); sim.set_controls(st)
            r1.arena_step(1, 0); sim.step(1)
            R = ref.get_state(a)
            f_out = outside(*ls.diff_states(R, r1.get_state(0), ppa))
            d_out = outside(*ls.diff_states(R, sim.get_state(), ppa))
            n += 1; nf += f_out; nd += d_out; both += f_out and d_out
            r1.close()
        print("after step %d: %d arena-steps, fresh reference outside the tolerance %d, device code outside %d, both %d" % (s, n, nf, nd, both), flush=True)
    print("\n%dv%d %s: %d arena-steps from the running reference's snapshots" % (ppt, ppt, "scripted" if scripted else "unscripted", n))
    print("  a FRESH REFERENCE given the snapshot leaves the tolerance in %d (%.2f %%)" % (nf, 100.0 * nf / max(1, n)))
    print("  the device code given the snapshot leaves it in              %d (%.2f %%), %d of them together with the fresh reference" % (nd, 100.0 * nd / max(1, n), both))


if __name__ == "__main__":
    main()
