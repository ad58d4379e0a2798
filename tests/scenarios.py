"""TEST INFRASTRUCTURE: scripted physics scenarios (initial state + per-tick controls) on a 1v1 arena that exercise the
arena meshes (walls, corners, ramps), car-car bumps and demolitions. Used by tests/test_env_gpu.py (CUDA kernels through the
C ABI), tests/test_hostsim_cpu.py (host compile of the same headers) and tools/lockstep_scenarios.py."""
import numpy as np

import lockstep as ls


def yaw_rot(yaw):
    c, s = np.cos(yaw), np.sin(yaw)
    return [c, s, 0, -s, c, 0, 0, 0, 1]


def car_at(st, i, pos, yaw, vel=(0, 0, 0)):
    c = st.cars[i]
    c.pos[:] = pos; c.rot[:] = yaw_rot(yaw); c.vel[:] = vel


SCEN = {}


def scen(f):
    SCEN[f.__name__] = f
    return f


@scen
def wall_hit(st):
    car_at(st, 0, [2000.0, 4600.0, 17.0], np.pi / 2, [0, 1500.0, 0])
    return lambda t, s: ls.set_controls(s, 0, throttle=1, boost=1)


@scen
def corner_hit(st):
    car_at(st, 0, [3000.0, 4000.0, 17.0], np.pi / 4, [900.0, 900.0, 0])
    return lambda t, s: ls.set_controls(s, 0, throttle=1, steer=0.2)


@scen
def ramp_drive(st):
    car_at(st, 0, [3500.0, 0.0, 17.0], 0.0, [1200.0, 0, 0])
    return lambda t, s: ls.set_controls(s, 0, throttle=1)


@scen
def ball_wall(st):
    st.ball.pos[:] = [1500.0, 4700.0, 400.0]; st.ball.vel[:] = [400.0, 2500.0, 300.0]; st.ball.ang_vel[:] = [1, 2, 3]
    return None


@scen
def ball_corner(st):
    st.ball.pos[:] = [3200.0, 4200.0, 200.0]; st.ball.vel[:] = [1500.0, 1500.0, -100.0]
    return None


@scen
def ball_ramp(st):
    st.ball.pos[:] = [3600.0, 100.0, 300.0]; st.ball.vel[:] = [1500.0, 200.0, -400.0]
    return None


@scen
def car_car(st):
    car_at(st, 0, [1000.0, -1000.0, 17.0], np.pi / 2, [0, 1200.0, 0])
    car_at(st, 1, [1030.0, 0.0, 17.0], -np.pi / 2, [0, -800.0, 0])
    def ctl(t, s):
        ls.set_controls(s, 0, throttle=1); ls.set_controls(s, 1, throttle=1)
    return ctl


@scen
def demo(st):
    car_at(st, 0, [1000.0, -2500.0, 17.0], np.pi / 2, [0, 2250.0, 0])
    car_at(st, 1, [1000.0, 0.0, 17.0], 0.0, [0, 0, 0])
    def ctl(t, s):
        ls.set_controls(s, 0, throttle=1, boost=1)
    return ctl


# ------------------------------------------------------------------------------------------------------------------
# Scripted ARENA states for the env-level parity runs (tests/env_parity.py: run_steady_state): each function rewrites the
# reference's mid-play state of one arena so that, within a few env steps of random actions, the rollout goes through a goal,
# a shot, a save, a bump, a demolition or a boost-pad pickup. `ppt` = players per team; cars [0, ppt) are blue (they attack
# +y), cars [ppt, 2 ppt) orange. Cars the script does not use are parked in far corners, on the ground, out of the way.
def _park_rest(st, used, ppt):
    spots = [(-3000.0, -4000.0), (3000.0, -4000.0), (-3000.0, 4200.0), (3000.0, 4200.0), (-3500.0, 0.0), (3500.0, 1500.0)]
    k = 0
    for i in range(2 * ppt):
        if i in used:
            continue
        car_at(st, i, [spots[k][0], spots[k][1], 17.0], 0.3 * k, [0, 0, 0])
        _ground(st.cars[i]); k += 1


def _ground(c):
    c.ang_vel[:] = [0, 0, 0]
    c.is_on_ground = 1; c.has_jumped = 0; c.has_double_jumped = 0; c.has_flipped = 0; c.is_flipping = 0; c.is_jumping = 0
    c.is_demoed = 0; c.demo_respawn_timer = 0.0; c.air_time = 0.0; c.air_time_since_jump = 0.0; c.jump_time = 0.0; c.flip_time = 0.0
    c.is_auto_flipping = 0; c.car_contact_cooldown = 0.0; c.car_contact_other_id = 0
    for w in range(4):
        c.wheels_with_contact[w] = 1


def _ball(st, pos, vel=(0, 0, 0)):
    st.ball.pos[:] = pos; st.ball.vel[:] = vel; st.ball.ang_vel[:] = [0, 0, 0]


def env_script_goal(ppt):
    def f(st):
        _ball(st, [200.0, 4950.0, 120.0], [0.0, 1800.0, 0.0])
        _park_rest(st, set(), ppt)
    return f


def env_script_shot(ppt):
    def f(st):
        _ball(st, [0.0, 4300.0, 93.15])
        car_at(st, 0, [0.0, 4000.0, 17.0], np.pi / 2, [0, 1900.0, 0]); _ground(st.cars[0])
        _park_rest(st, {0}, ppt)
    return f


def env_script_save(ppt):
    def f(st):
        # orange car shoots at the blue goal (-y), a blue car drives out of the goal mouth into the ball
        _ball(st, [0.0, -3600.0, 93.15])
        car_at(st, ppt, [0.0, -3250.0, 17.0], -np.pi / 2, [0, -2000.0, 0]); _ground(st.cars[ppt])
        car_at(st, 0, [0.0, -4900.0, 17.0], np.pi / 2, [0, 1400.0, 0]); _ground(st.cars[0])
        _park_rest(st, {0, ppt}, ppt)
    return f


def env_script_bump(ppt):
    def f(st):
        _ball(st, [-2000.0, 2000.0, 93.15])
        car_at(st, 0, [1000.0, -500.0, 17.0], np.pi / 2, [0, 1300.0, 0]); _ground(st.cars[0])
        car_at(st, ppt, [1030.0, 0.0, 17.0], -np.pi / 2, [0, -800.0, 0]); _ground(st.cars[ppt])
        _park_rest(st, {0, ppt}, ppt)
    return f


def env_script_demo(ppt):
    def f(st):
        _ball(st, [-2000.0, -2000.0, 93.15])
        car_at(st, 0, [1000.0, -700.0, 17.0], np.pi / 2, [0, 2290.0, 0]); _ground(st.cars[0])
        st.cars[0].is_supersonic = 1; st.cars[0].supersonic_time = 0.5; st.cars[0].boost = 100.0
        car_at(st, ppt, [1000.0, 0.0, 17.0], 0.0, [0, 0, 0]); _ground(st.cars[ppt])
        _park_rest(st, {0, ppt}, ppt)
    return f


def env_script_pad(ppt):
    def f(st):
        _ball(st, [0.0, 0.0, 93.15])
        car_at(st, 0, [3584.0 - 450.0, 0.0, 17.0], 0.0, [1200.0, 0, 0]); _ground(st.cars[0]); st.cars[0].boost = 5.0
        car_at(st, ppt, [-1788.0 + 300.0, 2300.0, 17.0], np.pi, [-900.0, 0, 0]); _ground(st.cars[ppt]); st.cars[ppt].boost = 0.0
        for p in range(34):
            st.pads[p].is_active = 1; st.pads[p].cooldown = 0.0; st.pads[p].prev_locked_car_id = 0
        _park_rest(st, {0, ppt}, ppt)
    return f


def env_script_pileup(ppt):
    """every car of the arena converges on one spot: car-car manifolds of several pairs + floor contacts at once (the contact
    window overflow case of ADVICE r1)"""
    def f(st):
        _ball(st, [0.0, 2500.0, 93.15])
        n = 2 * ppt
        for i in range(n):
            ang = 2 * np.pi * i / n
            r = 330.0
            car_at(st, i, [r * np.cos(ang), r * np.sin(ang) - 1500.0, 17.0], ang + np.pi, [-1500.0 * np.cos(ang), -1500.0 * np.sin(ang), 0]); _ground(st.cars[i])
    return f


ENV_SCRIPTS = [env_script_goal, env_script_shot, env_script_save, env_script_bump, env_script_demo, env_script_pad, env_script_pileup]


def env_scripts(num_arenas, ppt, every=2):
    """{arena: script}: every `every`-th arena gets one of the scripts (round robin), the others keep the pre-rolled state"""
    out = {}
    k = 0
    for a in range(0, num_arenas, every):
        out[a] = ENV_SCRIPTS[k % len(ENV_SCRIPTS)](ppt); k += 1
    return out
