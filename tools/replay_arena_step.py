"""Debug aid (TEST INFRASTRUCTURE): replays one env step of one arena tick by tick on the compiled reference and on the host
compile of the device simulator, from a dump written by tests/env_parity.py (GGL_PARITY_DUMP=/tmp/x.pkl python -m pytest ...),
re-synchronising before every tick, and prints where the two differ (state fields, contact rows, wheel rays).
    python tools/replay_arena_step.py /tmp/x.pkl [detail] [free]"""
import ctypes as C
import os
import pickle
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402

import hostsim_binding as hb  # noqa: E402
import lockstep as ls  # noqa: E402
import oracle_ref  # noqa: E402


def main():
    d = pickle.load(open(sys.argv[1], "rb"))
    detail = "detail" in sys.argv[2:]
    free = "free" in sys.argv[2:]          # no re-synchronisation between the ticks: what the env step itself does
    np.set_printoptions(precision=4, suppress=True, linewidth=220)
    nc = d["ppa"]; ppt = nc // 2
    st0 = oracle_ref.abi.ArenaState.from_buffer_copy(d["state"])
    L = oracle_ref.lib(); L.ref_set_threads(1)
    L.ref_arena_wheel_info.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
    HL = hb.lib(); HL.hostsim_wheel_info.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    # the reference's car order (iteration order of an unordered_set of pointers) cannot be set: make arenas until one has
    # the order of the dump
    ref1, keep = None, []
    for attempt in range(400):
        r = oracle_ref.RefEnvSet(1, ppt, obs_kind=1, state_setter=oracle_ref.abi.SETTER_INJECTED)
        if list(r.slot_info()[0]) == list(d["ids"]):
            ref1 = r
            break
        keep.append(r)
    if ref1 is None:
        print("no one-arena reference with the dumped car order after 400 attempts; using", keep[-1].slot_info()[0])
        ref1 = keep[-1]
    ref1.arena_step(1, 0)          # the solver's time step is only set by the first tick
    sim = hb.HostSim()
    ref1.set_state(0, st0); sim.set_state(st0)
    order = ref1.slot_info()[0]
    sim.set_slot_order(order)      # the one-arena reference has its own car order: both sides use it
    print("step", d["step"], "arena", d["arena"], "dumped car order", d["ids"], "replay car order", order)
    table = oracle_ref.action_table()
    for t in range(8):
        st = ref1.get_state(0)
        if t == 7:
            for slot in range(nc):
                row = table[d["acts"][slot]]; ci = d["ids"][slot] - 1
                ls.set_controls(st, ci, throttle=row[0], steer=row[1], pitch=row[2], yaw=row[3], roll=row[4], jump=int(row[5]),
                                boost=int(row[6]), handbrake=int(row[7]))
            ref1.set_controls(0, st); st = ref1.get_state(0)
        if free:
            if t == 7:
                sim.set_controls(st)
        else:
            sim.set_state(st)
        ref1.arena_step(1, 0); sim.step(1)
        a = ref1.get_state(0); b = sim.get_state()
        df, bad = ls.diff_states(a, b, nc)
        big = {k: round(v, 6) for k, v in df.items() if v > 3e-4}
        rc = ls.ref_contacts(ref1, 0); hc = sim.contacts()
        print("tick", t, "flags", bad[:6], big, "contacts ref %d host %d" % (len(rc), len(hc)))
        if (big or bad or len(rc) != len(hc)) and detail:
            print("ref contacts (bodyA bodyB normal | posB dist friction restitution special | manifold)\n", np.c_[rc[:, :5], rc[:, 8:15], rc[:, 19]])
            print("host contacts\n", np.c_[hc[:, :5], hc[:, 8:15]])
            for ci in range(nc):
                ra = np.zeros((4, 12), np.float32); ha = np.zeros((4, 12), np.float32)
                L.ref_arena_wheel_info(ref1.h, 0, ci + 1, ra.ctypes.data); HL.hostsim_wheel_info(sim.h, ci + 1, ha.ctypes.data)
                if np.abs(ra - ha).max() > 1e-4:
                    print("car", ci, "wheels (normal | point | suspLen ground | impulse | suspRelVel) ref\n", ra, "\nhost\n", ha)
                print("car", ci, "pos", np.array(st.cars[ci].pos), "vel", np.array(st.cars[ci].vel), "demoed", st.cars[ci].is_demoed, "team", st.cars[ci].team)
            print("ball", np.array(st.ball.pos), np.array(st.ball.vel))


if __name__ == "__main__":
    main()
