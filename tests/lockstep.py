"""TEST INFRASTRUCTURE: lock-step comparison helpers (reference oracle vs a simulator under test)."""
import ctypes as C

import numpy as np

import oracle_ref

CAR_FLOAT_FIELDS = ["pos", "vel", "ang_vel", "rot", "boost", "jump_time", "flip_time", "air_time", "air_time_since_jump",
                    "time_spent_boosting", "supersonic_time", "handbrake_val", "auto_flip_timer", "flip_rel_torque",
                    "world_contact_normal", "car_contact_cooldown", "demo_respawn_timer", "hit_extra_vel", "hit_rel_pos"]
CAR_INT_FIELDS = ["is_on_ground", "has_jumped", "has_double_jumped", "has_flipped", "is_flipping", "is_jumping",
                  "is_supersonic", "is_auto_flipping", "world_contact_has", "is_demoed", "hit_valid", "hit_tick",
                  "hit_tick_extra", "car_contact_other_id"]


def _val(x):
    try:
        return np.array(list(x), dtype=np.float64)
    except TypeError:
        return np.array([x], dtype=np.float64)


def diff_states(a, b, ncars):
    """Returns (dict field -> max abs difference for float fields, list of mismatching integer fields)."""
    d = {}
    bad = []
    for f in ("pos", "vel", "ang_vel"):
        d["ball." + f] = float(np.max(np.abs(_val(getattr(a.ball, f)) - _val(getattr(b.ball, f)))))
    for f in ("hs_y_target_dir", "hs_cur_target_speed", "hs_time_since_hit"):     # BallState::HeatseekerInfo
        d["ball." + f] = abs(getattr(a.ball, f) - getattr(b.ball, f))
    for i in range(ncars):
        ca, cb = a.cars[i], b.cars[i]
        for f in CAR_FLOAT_FIELDS:
            d["car%d.%s" % (i, f)] = float(np.max(np.abs(_val(getattr(ca, f)) - _val(getattr(cb, f)))))
        for f in CAR_INT_FIELDS:
            if getattr(ca, f) != getattr(cb, f):
                bad.append(("car%d.%s" % (i, f), getattr(ca, f), getattr(cb, f)))
        for w in range(4):
            for f in ("steer_angle", "engine_force", "brake", "lat_friction", "long_friction", "extra_pushback"):
                d["car%d.w%d.%s" % (i, w, f)] = abs(getattr(ca.wheels[w], f) - getattr(cb.wheels[w], f))
            if ca.wheels_with_contact[w] != cb.wheels_with_contact[w]:
                bad.append(("car%d.wheel%d" % (i, w), ca.wheels_with_contact[w], cb.wheels_with_contact[w]))
    for p in range(34):
        if a.pads[p].is_active != b.pads[p].is_active:
            bad.append(("pad%d.active" % p, a.pads[p].is_active, b.pads[p].is_active))
        d["pads.cooldown"] = max(d.get("pads.cooldown", 0.0), abs(a.pads[p].cooldown - b.pads[p].cooldown))
    return d, bad


def set_controls(st, car, throttle=0, steer=0, pitch=0, yaw=0, roll=0, jump=0, boost=0, handbrake=0):
    k = st.cars[car].controls
    k.throttle, k.steer, k.pitch, k.yaw, k.roll = throttle, steer, pitch, yaw, roll
    k.jump, k.boost, k.handbrake = jump, boost, handbrake


def ref_contacts(env, arena=0, maxn=64):
    L = oracle_ref.lib()
    L.ref_arena_contacts.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
    out = np.zeros((maxn, 20), np.float32)
    n = L.ref_arena_contacts(env.h, arena, out.ctypes.data, maxn)
    return out[:n]


def run_lockstep(env, sim, ticks, ncars, controls_fn=None, stop_at=1e-2, verbose=True):
    """Steps the reference arena 0 and `sim` tick by tick; returns (worst float diffs, first mismatching tick or None)."""
    worst = {}
    for t in range(ticks):
        if controls_fn is not None:
            st = env.get_state(0)
            controls_fn(t, st)
            # only the controls change: push them into both sides without touching the rest of the state
            env.set_controls(0, st)
            sim.set_controls(st)
        env.arena_step(1, 0)
        sim.step(1)
        a = env.get_state(0)
        b = sim.get_state()
        d, bad = diff_states(a, b, ncars)
        for k, v in d.items():
            worst[k] = max(worst.get(k, 0), v)
        if bad or max(d.values()) > stop_at:
            if verbose:
                print("tick", t, "bad", bad[:6], {k: round(v, 5) for k, v in d.items() if v > stop_at / 10})
            return worst, t
    return worst, None
