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
