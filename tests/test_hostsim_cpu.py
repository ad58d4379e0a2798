"""CPU suite: the device simulator headers, compiled for the host (tests/hostsim — test infrastructure), lock-stepped
against the compiled reference when oracle/_ref is present. Catches logic regressions without a GPU; the GPU suite
repeats the comparison with the real CUDA kernels."""
import pytest

oracle_ref = pytest.importorskip("oracle_ref")
if not oracle_ref.available():
    pytest.skip("oracle/_ref not built (needs /root/reference)", allow_module_level=True)

import env_parity              # noqa: E402
import hostsim_binding as hb   # noqa: E402
import lockstep as ls          # noqa: E402


def test_ball_bounce_and_resting_cars_lockstep():
    L = oracle_ref.lib(); L.ref_set_threads(1); L.ref_seed(3)
    env = oracle_ref.RefEnvSet(1, 1)
    sim = hb.HostSim()
    st = env.get_state(0)
    st.ball.pos[:] = [200.0, -300.0, 600.0]; st.ball.vel[:] = [300.0, 250.0, -100.0]; st.ball.ang_vel[:] = [1.0, -2.0, 0.5]
    env.set_state(0, st); sim.set_state(st)
    worst, bad = ls.run_lockstep(env, sim, 240, 2, stop_at=1e9, verbose=False)
    assert max(v for k, v in worst.items() if k.endswith(".pos")) < 5e-3, worst
    env.close()


def test_driving_and_ball_hit_lockstep():
    L = oracle_ref.lib(); L.ref_set_threads(1); L.ref_seed(4)
    env = oracle_ref.RefEnvSet(1, 1)
    sim = hb.HostSim()
    st = env.get_state(0)
    c = st.cars[0]
    c.pos[:] = [0.0, -600.0, 17.0]; c.rot[:] = [0, 1, 0, -1, 0, 0, 0, 0, 1]; c.vel[:] = [0, 1000.0, 0]
    env.set_state(0, st); sim.set_state(st)

    def ctl(t, s):
        ls.set_controls(s, 0, throttle=1, boost=1, steer=0.05)
        ls.set_controls(s, 1, throttle=1, steer=-0.5, jump=1 if t in range(20, 30) else 0)
    worst, bad = ls.run_lockstep(env, sim, 200, 2, controls_fn=ctl, stop_at=1e9, verbose=False)
    assert max(v for k, v in worst.items() if k.endswith(".pos")) < 2e-2, worst
    assert worst["ball.vel"] < 0.05, worst
    env.close()


@pytest.mark.parametrize("ppt,obs_kind", [(1, 0), (2, 1)])
def test_env_layer_matches_reference(ppt, obs_kind):
    r = env_parity.run_env_steps(None, num_arenas=3, ppt=ppt, steps=20, seed=ppt, obs_kind=obs_kind, env_cls=hb.HostEnvSet)
    assert r["mask_mismatch"] == 0 and r["term_mismatch"] == 0 and r["event_mismatch"] == 0 and r["obs_bad"] == 0, r
    assert r["rew"] < 1e-4, r


@pytest.mark.parametrize("name", ["corner_hit", "ball_ramp", "car_car", "demo"])
def test_contact_scenarios_per_tick(name):
    """Host compile of the device headers on the mesh / car-car / demolition scenarios, re-synchronised every tick."""
    import scenarios
    L = oracle_ref.lib(); L.ref_set_threads(1); L.ref_seed(5)
    env = oracle_ref.RefEnvSet(1, 1)
    sim = hb.HostSim()
    st = env.get_state(0)
    ctl = scenarios.SCEN[name](st)
    env.set_state(0, st); sim.set_state(st)
    worst, nbad = 0.0, 0
    for t in range(160):
        sim.set_state(env.get_state(0))
        w, bad = ls.run_lockstep(env, sim, 1, 2, controls_fn=(lambda tt, s: ctl(t, s)) if ctl else None, stop_at=1e9, verbose=False)
        worst = max(worst, max(v for k, v in w.items() if k.endswith(".pos")))
        nbad += bad is not None
    assert nbad == 0 and worst <= 1e-3, (name, worst, nbad)
    env.close()
