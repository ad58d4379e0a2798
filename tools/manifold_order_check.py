"""Does the device predict the reference's manifold array? (TEST INFRASTRUCTURE, CPU: the host compile of the device code against
the compiled reference.)

solver_order (csrc/sim_world.cuh) rebuilds, from the bodies' boxes and the broadphase registry, the array of contact manifolds
the reference's dispatcher holds in a tick - empty manifolds included - because Bullet sorts THAT array by island with an unstable
quickSort before the solver walks it. This tool plays arenas on the reference, re-synchronises a host simulator before every
tick, and compares the predicted array with the dispatcher's (oracle/ref_harness.cpp: ref_arena_manifolds) on every tick in
which something touched something.
    python tools/manifold_order_check.py [ppt] [arenas] [env steps] [scripted 0/1]"""
import collections
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import ctypes as C  # noqa: E402

import numpy as np  # noqa: E402

import env_parity  # noqa: E402
import hostsim_binding as hb  # noqa: E402
import lockstep as ls  # noqa: E402
import oracle_ref  # noqa: E402
import scenarios  # noqa: E402

abi = oracle_ref.abi


def ref_manifolds(ref, arena, maxn=128):
    out = np.zeros((maxn, 3), np.int32)
    L = oracle_ref.lib()
    L.ref_arena_manifolds.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
    n = L.ref_arena_manifolds(ref.h, arena, out.ctypes.data_as(C.c_void_p), maxn)
    return out[:n]


def canon(rows, num_meshes_offset):
    """(dynamic body or static code) pairs with the dynamic body of lower code first, statics as -(100 + index)"""
    out = []
    for a, b in rows:
        a, b = int(a), int(b)
        if (a < 0 and b >= 0) or (a >= 0 and b >= 0 and b < a):
            a, b = b, a
        out.append((a, b))
    return out


def run(ppt=2, arenas=32, steps=30, scripted=True, show=0, show_state=False):
    """Counter of outcomes over every tick with a contact (see the module text)"""
    ppa = 2 * ppt
    L = oracle_ref.lib(); L.ref_set_threads(1); L.ref_seed(33 + ppt)
    ref = oracle_ref.RefEnvSet(arenas, ppt, obs_kind=1, state_setter=abi.SETTER_INJECTED)
    rng = np.random.default_rng(33 + ppt)
    for _ in range(60):
        ref.step(env_parity.random_valid_actions(rng, ref.outputs()[1]))
    if scripted:
        for a, f in scenarios.env_scripts(arenas, ppt, 2).items():
            st = ref.get_state(a); f(st); ref.set_state(a, st)
    ids_all = ref.slot_info()[0]
    table = oracle_ref.action_table()
    hb.order_mode(1)
    sims = []
    for a in range(arenas):
        s = hb.HostSim()
        s.set_state(ref.get_state(a))
        s.set_slot_order(ids_all[a * ppa:(a + 1) * ppa])
        sims.append(s)
    # static index of the reference = position in the world's object array = the device's static index (meshes, then planes)
    stats = collections.Counter()
    shown = 0
    for s_i in range(steps):
        acts = env_parity.random_valid_actions(rng, ref.outputs()[1])
        for a in range(arenas):
            st = ref.get_state(a)
            for slot in range(ppa):
                row = table[acts[a * ppa + slot]]; ci = ids_all[a * ppa + slot] - 1
                ls.set_controls(st, ci, throttle=row[0], steer=row[1], pitch=row[2], yaw=row[3], roll=row[4], jump=int(row[5]),
                                boost=int(row[6]), handbrake=int(row[7]))
            ref.set_controls(a, st)
        for t in range(8):
            for a in range(arenas):
                st = ref.get_state(a)
                sims[a].set_state(st)
                ref.arena_step(1, a)
                sims[a].step(1)
                pred = sims[a].predicted_manifolds()
                rm = ref_manifolds(ref, a)
                nonempty = int((rm[:, 2] > 0).sum())
                if pred is None:
                    stats["ticks without a device contact"] += 1
                    if nonempty:
                        stats["  ... but the reference has points"] += 1
                        if shown < show:
                            shown += 1
                            print("step %d tick %d arena %d: no device contact, reference manifolds %s" % (s_i, t, a, rm.tolist()), flush=True)
                    continue
                pm, reordered, unknown = pred
                stats["ticks with a contact"] += 1
                # manifolds of a demolished car exist in the dispatcher but never reach the sort (no contact response:
                # btCollisionDispatcher::needsResponse), and the prediction is of the sort's input
                after = ref.get_state(a)
                gone = {i + 1 for i in range(ppa) if st.cars[i].is_demoed or after.cars[i].is_demoed}
                want = [w for w in canon(rm[:, :2], 0) if w[0] not in gone and w[1] not in gone]
                got = [(int(x), int(y)) for x, y in pm[:, :2]]
                if unknown:
                    stats["a row's manifold missing from the prediction"] += 1
                if want == got:
                    stats["array predicted exactly"] += 1
                elif sorted(want) == sorted(got):
                    stats["same manifolds, other order"] += 1
                    if shown < show:
                        shown += 1
                        print("step %d tick %d arena %d (order)\n  ref  %s\n  pred %s" % (s_i, t, a, want, got), flush=True)
                else:
                    stats["other manifolds"] += 1
                    extra = sorted(set(got) - set(want)); missing = sorted(set(want) - set(got))
                    stats["  predicted but absent: %d, present but not predicted: %d" % (min(len(extra), 3), min(len(missing), 3))] += 1
                    if shown < show:
                        shown += 1
                        print("step %d tick %d arena %d\n  ref  %s\n  pred %s" % (s_i, t, a, want, got), flush=True)
                        if show_state:
                            for x, y in extra + missing:
                                for b in (x, y):
                                    if b == 0:
                                        after = ref.get_state(a).ball
                                        print("    ball pos %s vel %s angvel %s -> after: pos %s vel %s" % (list(st.ball.pos), list(st.ball.vel), list(st.ball.ang_vel), list(after.pos), list(after.vel)))
                                    if b > 0:
                                        c = st.cars[b - 1]
                                        pb = np.zeros((8, 6), np.float32)
                                        hb.lib().hostsim_predicted_boxes(C.c_void_p(sims[a].h), pb.ctypes.data_as(C.c_void_p))
                                        slot = list(ids_all[a * ppa:(a + 1) * ppa]).index(b)
                                        after = ref.get_state(a).cars[b - 1]
                                        print("    box predicted %s\n        reference %s" % (pb[b].tolist(), list(after.ray_aabb_min) + list(after.ray_aabb_max)))
                                        print("    car %d pos %s vel %s angvel %s demoed %d fwd %s up %s" % (b, list(c.pos), list(c.vel), list(c.ang_vel), c.is_demoed, list(c.rot[0:3]), list(c.rot[6:9])))
                if reordered:
                    stats["ticks whose rows were reordered"] += 1
    hb.order_mode(0)
    for s in sims:
        s.close() if hasattr(s, "close") else None
    ref.close()
    return stats


def main():
    ppt = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    arenas = int(sys.argv[2]) if len(sys.argv) > 2 else 32
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 30
    scripted = (int(sys.argv[4]) if len(sys.argv) > 4 else 1) != 0
    stats = run(ppt, arenas, steps, scripted, show=int(os.environ.get("GGL_SHOW", "12")), show_state=bool(os.environ.get("GGL_SHOW_STATE")))
    print("\n%dv%d, %d arenas x %d env steps (%s)" % (ppt, ppt, arenas, steps, "scripted" if scripted else "unscripted"))
    for k, v in sorted(stats.items()):
        print("  %-60s %d" % (k, v))


if __name__ == "__main__":
    main()
