"""CPU parity of the random plugins: the host compile of the device code (tests/hostsim) against the compiled
reference, draw for draw - FuzzedKickoffState, RandomState, CombinedState, the kickoff shuffle, DefaultObsPadded."""
import pytest

oracle_ref = pytest.importorskip("oracle_ref")
if not oracle_ref.available():
    pytest.skip("oracle/_ref not built (needs /root/reference)", allow_module_level=True)

import hostsim_binding      # noqa: E402
import setter_parity        # noqa: E402
from oracle_ref import abi  # noqa: E402

STATE_ATOL = 1e-4      # uu / unit vectors: both sides run the same float operations on the host, differences are rounding of uu<->bt
OBS_ATOL = 1e-6


@pytest.mark.parametrize("ppt", [1, 2, 3])
def test_kickoff_shuffle_draws_like_std_shuffle(ppt):
    r = setter_parity.run_resets(hostsim_binding.HostEnvSet, ppt, trials=60, seed=ppt, state_setter=abi.SETTER_KICKOFF, obs_kind=abi.OBS_DEFAULT)
    assert r["state"] <= STATE_ATOL and r["obs"] <= OBS_ATOL, r


@pytest.mark.parametrize("ppt", [1, 3])
def test_fuzzed_kickoff_state(ppt):
    r = setter_parity.run_resets(hostsim_binding.HostEnvSet, ppt, trials=40, seed=10 + ppt, state_setter=abi.SETTER_FUZZED_KICKOFF)
    assert r["state"] <= STATE_ATOL and r["obs"] <= OBS_ATOL, r


@pytest.mark.parametrize("flags", [0, 1, 2, 3, 4, 7])
def test_random_state(flags):
    r = setter_parity.run_resets(hostsim_binding.HostEnvSet, 2, trials=40, seed=20 + flags, state_setter=abi.SETTER_RANDOM, state_setter_flags=flags)
    assert r["state"] <= STATE_ATOL and r["obs"] <= OBS_ATOL, r
    assert r["ball_z_values"] > 20


def test_combined_state_picks_children_by_weight():
    kids = [(abi.SETTER_KICKOFF, 1.0), (abi.SETTER_RANDOM, 2.0, 3), (abi.SETTER_FUZZED_KICKOFF, 0.5), (abi.SETTER_RANDOM, 1.5, 4)]
    r = setter_parity.run_resets(hostsim_binding.HostEnvSet, 2, trials=80, seed=31, state_setter=abi.SETTER_COMBINED, setter_children=kids)
    assert r["state"] <= STATE_ATOL and r["obs"] <= OBS_ATOL, r
    assert r["ball_z_values"] > 10          # both kickoff children (one value) and random children were drawn


@pytest.mark.parametrize("ppt,maxp", [(1, 1), (1, 3), (2, 2), (2, 3), (3, 3), (2, 4)])
def test_default_obs_padded_resets(ppt, maxp):
    r = setter_parity.run_resets(hostsim_binding.HostEnvSet, ppt, trials=40, seed=40 + ppt + maxp, state_setter=abi.SETTER_RANDOM, state_setter_flags=3,
                                 obs_kind=abi.OBS_DEFAULT_PADDED, obs_max_players=maxp)
    assert r["state"] <= STATE_ATOL and r["obs"] <= OBS_ATOL, r


@pytest.mark.parametrize("ppt,maxp", [(2, 3), (3, 3)])
def test_default_obs_padded_steps(ppt, maxp):
    r = setter_parity.run_padded_steps(hostsim_binding.HostEnvSet, ppt, maxp, steps=25, seed=50 + ppt)
    assert r["obs"] <= 2e-4, r
    if maxp > ppt:          # with every slot occupied the layout is one; the values still tell players apart
        assert r["distinct_layouts"] > 5


def test_custom_obs_coefficients():
    """DefaultObs(posCoef, velCoef, angVelCoef) and DefaultObsPadded(maxPlayers, ...) with non-default coefficients (DefaultObs.h:9-17)."""
    import env_parity
    coefs = [1 / 5000.0, 1 / 6000.0, 1 / 2500.0, 1 / 3000.0, 0.25]
    r = env_parity.run_env_steps(None, num_arenas=3, ppt=2, steps=12, seed=3, obs_kind=abi.OBS_DEFAULT, env_cls=hostsim_binding.HostEnvSet, obs_coefs=coefs)
    assert r["obs_bad"] == 0 and r["mask_mismatch"] == 0, r
    r = setter_parity.run_resets(hostsim_binding.HostEnvSet, 2, trials=20, seed=77, state_setter=abi.SETTER_RANDOM, state_setter_flags=3,
                                 obs_kind=abi.OBS_DEFAULT_PADDED, obs_max_players=3, obs_coefs=coefs)
    assert r["state"] <= STATE_ATOL and r["obs"] <= OBS_ATOL, r


# ---- GameMode::HEATSEEKER: kickoff, seeking ball, hit and wall-bounce rules, unlimited boost ----
def test_heatseeker_kickoff_matches_reference():
    """Arena::ResetToRandomKickoff in heatseeker mode: four slots, partial shuffle, ball start side from one more draw."""
    for ppt in (1, 2, 3):
        r = setter_parity.run_resets(hostsim_binding.HostEnvSet, ppt, trials=40, seed=60 + ppt, state_setter=abi.SETTER_KICKOFF, game_mode=abi.MODE_HEATSEEKER)
        assert r["state"] <= STATE_ATOL and r["obs"] <= OBS_ATOL, r
