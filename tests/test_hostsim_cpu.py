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


def test_steady_state_scripted_events_match_reference():
    """Host compile, ADVICE r1: parity where the game actually happens. The reference is pre-rolled 100 steps from the kickoff,
    every arena is then put into a scripted situation (goal, shot, save attempt, bump, demolition, pad pickup, pile-up) or left
    in mid-play, and 40 env steps are compared with the physics re-synchronised before every step. Integer outputs (masks,
    terminals, all nine event flags) must be exact with goals, shots, bumps and demolitions actually occurring; obs / rewards
    hold 1e-4 except on the arena-steps that depend on state the reference keeps outside its snapshots or on a knife-edge
    decision, which are counted and bounded: 3 of 560 measured with the reference's manifold order
    reproduced (solver_order), 2 % allowed."""
    import scenarios
    r = env_parity.run_steady_state(None, num_arenas=14, ppt=2, preroll=100, steps=40, seed=2, env_cls=hb.HostEnvSet,
                                    scripts=scenarios.env_scripts(14, 2, 1))
    assert r["mask_mismatch"] == 0 and r["term_mismatch"] == 0 and r["event_mismatch"] == 0, r
    assert r["contact_overflow"] == 0, r
    ev = r["events"]
    assert ev[0] > 0 and ev[3] > 0 and ev[5] > 0 and ev[6] > 0, r          # goal, shot, bump, bumped
    assert r["touches"] > 0 and r["goals_seen"] > 0 and r["terminals_seen"] > 0, r
    assert r["bad_arena_steps"] <= 0.02 * r["arena_steps"], r


def test_steady_state_3v3_pileups_do_not_overflow_contact_rows():
    import scenarios
    r = env_parity.run_steady_state(None, num_arenas=14, ppt=3, preroll=60, steps=30, seed=3, env_cls=hb.HostEnvSet,
                                    scripts=scenarios.env_scripts(14, 3, 1), verbose=bool(__import__("os").environ.get("GGL_TEST_VERBOSE")))
    assert r["mask_mismatch"] == 0 and r["term_mismatch"] == 0 and r["event_mismatch"] == 0, r
    assert r["contact_overflow"] == 0, r
    assert r["events"][7] > 0 and r["events"][8] > 0, r                      # demo, demoed
    assert r["bad_arena_steps"] <= 0.03 * r["arena_steps"], r                 # 2 of 420 measured


def test_suspension_ray_vs_car_hitbox_matches_reference_cast():
    """The convex cast a suspension ray runs against another car's hitbox (btSubsimplexConvexCast through
    btCollisionWorld::rayTestSingle, reference shapes) against the device's restatement: hit / miss, fraction and normal."""
    import ctypes as C
    import numpy as np
    L = oracle_ref.lib(); L.ref_set_threads(1)
    ref = oracle_ref.RefEnvSet(1, 1)
    H = hb.lib()
    fp = C.c_void_p
    L.ref_convex_ray_cast.argtypes = [C.c_void_p, C.c_int, fp, fp, fp, fp, fp]
    H.hostsim_convex_ray_cast.argtypes = [C.c_int, fp, fp, fp, fp, fp]
    rng = np.random.default_rng(0)
    hits = 0
    for it in range(3000):
        q = rng.standard_normal(4); q /= np.linalg.norm(q); w, x, y, z = q
        basis = np.array([1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w), 2 * (x * y + z * w), 1 - 2 * (x * x + z * z),
                          2 * (y * z - x * w), 2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)], np.float32)
        pos = (rng.standard_normal(3) * 20).astype(np.float32)
        d = rng.standard_normal(3); d /= np.linalg.norm(d)
        target = pos + (rng.standard_normal(3) * 0.7).astype(np.float32)
        frm = (target - d * rng.uniform(0.2, 1.5)).astype(np.float32); to = (frm + d * 1.3).astype(np.float32)
        o1 = np.zeros(4, np.float32); o2 = np.zeros(4, np.float32)
        h1 = L.ref_convex_ray_cast(ref.h, 0, frm.ctypes.data, to.ctypes.data, pos.ctypes.data, basis.ctypes.data, o1.ctypes.data)
        h2 = H.hostsim_convex_ray_cast(0, frm.ctypes.data, to.ctypes.data, pos.ctypes.data, basis.ctypes.data, o2.ctypes.data)
        assert h1 == h2, (it, h1, h2)
        if h1:
            hits += 1
            assert o1[0] == o2[0] and np.abs(o1[1:] - o2[1:]).max() <= 5e-7, (it, o1, o2)
    assert hits > 500
    ref.close()


def test_mesh_triangles_are_stored_in_reference_bvh_leaf_order():
    """The device meets the triangles of a mesh in index order; the reference in the leaf order of the mesh's quantized BVH
    (row order inside a multi-triangle manifold, which contact normal a car keeps). The host build of the world stores every
    mesh's triangles in that order (sim_convert.cuh: bvh_leaf_order): checked against the compiled reference's own trees."""
    import ctypes as C
    import struct
    import numpy as np
    L = oracle_ref.lib()
    L.ref_mesh_leaf_order.argtypes = [C.c_int, C.c_void_p, C.c_int]
    sim = hb.HostSim()
    H = hb.lib()
    H.hostsim_tri.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    H.hostsim_num_tris.argtypes = [C.c_void_p]
    meshes = oracle_ref.arena_mesh.soccar_meshes()
    base, multi = 0, 0
    for m, blob in enumerate(meshes):
        nt, nv = struct.unpack_from("<ii", blob, 0)
        idx = np.frombuffer(blob, np.int32, nt * 3, 8).reshape(nt, 3)
        vert = np.frombuffer(blob, np.float32, nv * 3, 8 + nt * 12).reshape(nv, 3)
        order = np.zeros(nt, np.int32)
        assert L.ref_mesh_leaf_order(m, order.ctypes.data, nt) == nt
        multi += int((order != np.arange(nt)).any())
        for k in range(nt):
            out = np.zeros(9, np.float32)
            assert H.hostsim_tri(sim.h, base + k, out.ctypes.data) == m
            assert np.array_equal(out.reshape(3, 3), vert[idx[order[k]]]), (m, k, order[k])
        base += nt
    assert base == H.hostsim_num_tris(sim.h)
    assert multi > 0      # at least one mesh whose BVH order is not the file order, so the check means something


def test_free_box_of_the_world_touches_no_triangle():
    """Cars, ball and suspension rays whose AABB lies inside the world's empty-cell box skip the triangle grid (WorldDev::freeMin /
    freeMax): no triangle's AABB, grown by the 0.1 the grid registers it with, may reach into that box, and the box is not trivial."""
    import ctypes as C
    import numpy as np
    sim = hb.HostSim()
    H = hb.lib()
    H.hostsim_free_box.argtypes = [C.c_void_p, C.c_void_p]
    H.hostsim_tri.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    H.hostsim_num_tris.argtypes = [C.c_void_p]
    box = np.zeros(6, np.float32)
    H.hostsim_free_box(sim.h, box.ctypes.data)
    lo, hi = box[:3], box[3:]
    assert (hi - lo > np.array([100.0, 100.0, 30.0])).all(), box          # > 5000 x 5000 x 1500 uu of open field
    out = np.zeros(9, np.float32)
    for i in range(H.hostsim_num_tris(sim.h)):
        H.hostsim_tri(sim.h, i, out.ctypes.data)
        v = out.reshape(3, 3)
        mn, mx = v.min(0) - 0.1, v.max(0) + 0.1
        assert not ((mx > lo).all() and (mn < hi).all()), (i, v, box)


def test_kickoff_proximity_reward_matches_reference():
    """KickoffProximityReward2v2Enhanced (row a8): goer / cheater branches off the kickoff, 2v2, against the compiled reference."""
    A = oracle_ref.abi
    r = env_parity.run_env_steps(None, num_arenas=6, ppt=2, steps=12, seed=7, obs_kind=1, env_cls=hb.HostEnvSet,
                                 rewards=[(A.REW_KICKOFF_PROXIMITY_2V2, 1.0, [1.2, 0.6, 0.3, 0.2])])
    assert r["rew"] <= 1e-5 and r["rew_nonzero"] > 20 and r["rew_abs_max"] > 0.1, r
    assert r["mask_mismatch"] == 0 and r["obs_bad"] == 0, r


def heatseeker_per_tick(make_sim, ticks=900):
    """One blue car drives into the kickoff-side ball; the ball then seeks the orange goal, hits the back wall, turns round ...
    Every tick starts from the reference's state (re-synchronised); returns the worst per-tick differences and what happened."""
    L = oracle_ref.lib(); L.ref_set_threads(1); L.ref_seed(11)
    ref = oracle_ref.RefEnvSet(1, 1, game_mode=oracle_ref.abi.MODE_HEATSEEKER)
    sim = make_sim()
    st = ref.get_state(0)
    st.ball.pos[:] = [0.0, -1500.0, 92.75]; st.ball.vel[:] = [0.0, 0.0, 0.0]
    c = st.cars[0]; c.pos[:] = [100.0, -2600.0, 17.0]; c.rot[:] = [0, 1, 0, -1, 0, 0, 0, 0, 1]; c.vel[:] = [0, 1200.0, 0]
    ref.set_state(0, st)
    worst = {}
    flips, speedups, nbad, last_dir, last_speed = 0, 0, 0, 0.0, st.ball.hs_cur_target_speed
    for t in range(ticks):
        cur = ref.get_state(0)
        sim.set_state(cur)
        ls.set_controls(cur, 0, throttle=1 if t < 150 else -1, boost=1 if t < 150 else 0, steer=0.02 if t < 150 else 1.0)      # hit the ball, then circle in reverse
        ls.set_controls(cur, 1, throttle=1, boost=1, steer=-0.3 if t > 300 else 0)
        ref.set_controls(0, cur); sim.set_controls(cur)
        ref.arena_step(1, 0); sim.step(1)
        a, b = ref.get_state(0), sim.get_state()
        d, bad = ls.diff_states(a, b, 2)
        nbad += len(bad) > 0
        for k, v in d.items():
            worst[k] = max(worst.get(k, 0.0), v)
        if a.ball.hs_y_target_dir != last_dir:
            flips += 1; last_dir = a.ball.hs_y_target_dir
        if a.ball.hs_cur_target_speed != last_speed:
            speedups += 1; last_speed = a.ball.hs_cur_target_speed
    boost_end = ref.get_state(0).cars[1].boost
    ref.close()
    return worst, nbad, flips, speedups, boost_end


def test_heatseeker_ball_per_tick():
    def make():
        s = hb.HostSim(); s.set_game_mode(oracle_ref.abi.MODE_HEATSEEKER); return s
    worst, nbad, flips, speedups, boost_end = heatseeker_per_tick(make)
    assert nbad == 0
    assert flips >= 2 and speedups >= 1, (flips, speedups)          # the car's hit, then at least one back-wall turn
    assert max(v for k, v in worst.items() if k.endswith(".pos")) <= 2e-3, worst
    assert worst["ball.hs_y_target_dir"] == 0 and worst["ball.hs_cur_target_speed"] == 0 and worst["ball.hs_time_since_hit"] <= 1e-5, worst
    assert boost_end >= 100.0 / 3 - 1e-3                             # unlimited boost: 900 ticks of boosting used nothing (pads may add)


def test_island_sort_is_bullets_quicksort():
    """order_sort_keys (csrc/sim_world.cuh) against btAlignedObjectArray::quickSort itself (oracle/ref_harness.cpp:
    ref_quicksort_keys): same permutation for every key array - the sort is unstable, so the permutation of equal keys is
    the whole point (btSimulationIslandManager.cpp:370-384)."""
    import ctypes as C
    import numpy as np
    R, H = oracle_ref.lib(), hb.lib()
    rng = np.random.default_rng(5)
    cases = [np.zeros(n, np.int32) for n in range(2, 20)]
    cases += [rng.integers(0, k, n).astype(np.int32) for n in range(2, 41) for k in (2, 3, 5, 7)]
    for keys in cases:
        a, b = np.zeros(keys.size, np.int32), np.zeros(keys.size, np.int32)
        R.ref_quicksort_keys(keys.ctypes.data_as(C.c_void_p), a.ctypes.data_as(C.c_void_p), int(keys.size))
        H.hostsim_sort_keys(keys.ctypes.data_as(C.c_void_p), b.ctypes.data_as(C.c_void_p), int(keys.size))
        assert np.array_equal(a, b), (keys.tolist(), a.tolist(), b.tolist())
        assert np.all(np.diff(keys[a]) >= 0)


def test_manifold_array_of_the_reference_is_predicted():
    """solver_order's prediction of the dispatcher's manifold array (pairs of the custom broadphase in handle order, manifolds
    made with the pair or by the compound child test) against the reference's array on every tick with a contact, in scripted
    play with the contact scenarios of tests/scenarios.py. The one documented gap - the order in which dynamic bodies arrived
    in a broadphase cell - only shows in pile-ups of three or more cars."""
    import os
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tools"))
    import manifold_order_check
    stats = manifold_order_check.run(ppt=2, arenas=16, steps=6, scripted=True)
    ticks = stats["ticks with a contact"]
    assert ticks >= 60
    # (the 4-car pile-up of the scenarios is a tenth of this sample's contact ticks; over unscripted play the array is exact)
    assert stats["array predicted exactly"] >= 0.85 * ticks, dict(stats)
    assert stats["array predicted exactly"] + stats["same manifolds, other order"] >= 0.98 * ticks, dict(stats)
    assert stats["a row's manifold missing from the prediction"] == 0
    assert stats["ticks whose rows were reordered"] > 0
