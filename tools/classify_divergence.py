"""What the arena-steps that leave the obs tolerance in steady-state play are made of (TEST INFRASTRUCTURE, runs on the CPU: the host
compile of the device code against the compiled reference, both deterministic).

Steady-state play as in tests/env_parity.run_steady_state (physics re-synchronised before every env step). Every arena-step whose
obs differ by more than the tolerance is replayed tick by tick on a one-arena reference and the host simulator, re-synchronised
before every tick, and the FIRST tick whose state differs by more than 1e-3 uu (or in a flag) is classified by its solver input:
  same rows, same order     both sides solved the same contact rows in the same order: arithmetic (rounding next to a discontinuity)
  same rows, other order    the same set of rows in another order (the reference's manifold order inside an island)
  other rows                the contact sets differ (a contact more or less, another normal): narrow phase / internal edges
  no rows                   no contact row on either side: rays, suspension, car logic
  not reproduced            the replay agrees on every tick (the one-arena reference has another car order than the dumped arena)
    python tools/classify_divergence.py [ppt] [arenas] [steps] [scripted 0/1]"""
import collections
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import numpy as np  # noqa: E402

import env_parity  # noqa: E402
import hostsim_binding as hb  # noqa: E402
import lockstep as ls  # noqa: E402
import oracle_ref  # noqa: E402
import scenarios  # noqa: E402

abi = oracle_ref.abi


def row_key(r):
    # bodies, normal, point on B, distance (the reference's harness reports the point on A in world space, the device relative to A)
    return tuple(np.round(np.r_[r[0:5], r[8:12]], 3).tolist())


def one_arena_ref(ppt, ids):
    keep = []
    for _ in range(300):
        r = oracle_ref.RefEnvSet(1, ppt, obs_kind=1, state_setter=abi.SETTER_INJECTED)
        if list(r.slot_info()[0]) == list(ids):
            for k in keep:
                k.close()
            return r, True
        keep.append(r)
    for k in keep[:-1]:
        k.close()
    return keep[-1], False


def classify(ppt, state, acts, ids, table):
    nc = 2 * ppt
    ref1, same_order = one_arena_ref(ppt, ids)
    ref1.arena_step(1, 0)                       # the solver's time step is only set by the first tick
    sim = hb.HostSim()
    st0 = abi.ArenaState.from_buffer_copy(state)
    ref1.set_state(0, st0); sim.set_state(st0)
    sim.set_slot_order(ref1.slot_info()[0])
    out = ("not reproduced" if same_order else "not reproduced (other car order)", -1, "")
    for t in range(8):
        st = ref1.get_state(0)
        if t == 7:
            for slot in range(nc):
                row = table[acts[slot]]; ci = ids[slot] - 1
                ls.set_controls(st, ci, throttle=row[0], steer=row[1], pitch=row[2], yaw=row[3], roll=row[4], jump=int(row[5]),
                                boost=int(row[6]), handbrake=int(row[7]))
            ref1.set_controls(0, st); st = ref1.get_state(0)
        sim.set_state(st)
        ref1.arena_step(1, 0); sim.step(1)
        df, bad = ls.diff_states(ref1.get_state(0), sim.get_state(), nc)
        worst = max(v for k, v in df.items() if k.endswith(".pos") or k.endswith(".vel") and v > 0.05)
        if bad or worst > 1e-3:
            rc, hc = ls.ref_contacts(ref1, 0), sim.contacts()
            def body_pos(i):
                i = int(i)
                return np.array(st.ball.pos if i == 0 else st.cars[i - 1].pos) / 50.0

            def same(a, b):
                # a = reference row (points on A and B in world space), b = device row (points relative to the bodies): bodies, normal,
                # distance and the point on A in world space
                if a[0] != b[0] or a[1] != b[1]:
                    return False
                pa_dev = body_pos(b[0]) + b[5:8]
                return max(np.max(np.abs(a[2:5] - b[2:5])), abs(a[11] - b[11]), np.max(np.abs(a[5:8] - pa_dev))) < 2e-3
            top = sorted(((v, k) for k, v in df.items() if k.endswith((".pos", ".vel", ".ang_vel"))), reverse=True)[:2]
            why = " [%s]" % ", ".join("%s %.4g" % (k, v) for v, k in top) + (" flags %s" % bad[:2] if bad else "")
            if len(rc) == 0 and len(hc) == 0:
                out = ("no rows", t, why)
            elif len(rc) == len(hc) and all(same(a, b) for a, b in zip(rc, hc)):
                out = ("same rows, same order", t, why)
            elif len(rc) == len(hc) and all(any(same(a, b) for b in hc) for a in rc) and all(any(same(a, b) for a in rc) for b in hc):
                out = ("same rows, other order", t, why)
            else:
                out = ("other rows", t, why + " ref %d dev %d rows" % (len(rc), len(hc)))
                if os.environ.get("GGL_CLASSIFY_ROWS"):
                    np.set_printoptions(precision=4, suppress=True, linewidth=200)
                    print("   ref rows (bodies | normal | point on A (world), distance)\n", np.c_[rc[:, 0:8], rc[:, 11]])
                    print("   dev rows\n", np.c_[hc[:, 0:5], np.array([body_pos(h[0]) + h[5:8] for h in hc]).reshape(-1, 3), hc[:, 11]])
            break
    ref1.close()
    return out


def main():
    ppt = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    arenas = int(sys.argv[2]) if len(sys.argv) > 2 else 64
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    scripted = (int(sys.argv[4]) if len(sys.argv) > 4 else 1) != 0
    ppa = 2 * ppt
    hb.order_mode(int(os.environ.get("GGL_ORDER_MODE", "0")))      # 2 = without the reference's manifold order (solver_order)
    L = oracle_ref.lib(); L.ref_set_threads(1); L.ref_seed(21 + ppt)
    kw = dict(obs_kind=1, state_setter=abi.SETTER_INJECTED)
    ref = oracle_ref.RefEnvSet(arenas, ppt, **kw)
    dev = hb.HostEnvSet(arenas, ppt, seed=1, **kw)
    rng = np.random.default_rng(21 + ppt)
    for _ in range(100):                        # pre-roll the reference
        ref.step(env_parity.random_valid_actions(rng, ref.outputs()[1]))
    if scripted:
        for a, f in scenarios.env_scripts(arenas, ppt, 2).items():
            st = ref.get_state(a); f(st); ref.set_state(a, st)
    dev.set_slot_order(ref.slot_info()[0])
    ids_all = ref.slot_info()[0]
    table = oracle_ref.action_table()
    counts = collections.Counter()
    total = 0
    for s in range(steps):
        acts = env_parity.random_valid_actions(rng, ref.outputs()[1])
        snaps = [ref.get_state(a) for a in range(arenas)]
        term_before = ref.outputs()[3].copy()
        for a in range(arenas):
            dev.set_state(a, snaps[a])
        ref.step(acts); dev.step_np(acts)
        ro, do = ref.outputs()[0], dev.obs.cpu().numpy()
        ok = env_parity.close_rel(do, ro).reshape(arenas, -1).all(1)
        total += arenas
        for a in np.flatnonzero(~ok):
            if term_before[a]:
                continue                        # the step began with a reset: the snapshot is not what was stepped
            kind, tick, why = classify(ppt, bytes(snaps[a]), acts[a * ppa:(a + 1) * ppa], ids_all[a * ppa:(a + 1) * ppa], table)
            counts[kind] += 1
            print("step %3d arena %3d: %s (tick %d)%s" % (s, a, kind, tick, why), flush=True)
    print("\n%dv%d %s, %d arena-steps, %d outside the obs tolerance:" % (ppt, ppt, "scripted" if scripted else "unscripted", total, sum(counts.values())))
    for k, v in counts.most_common():
        print("  %-40s %d" % (k, v))


if __name__ == "__main__":
    main()
