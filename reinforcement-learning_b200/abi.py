"""ctypes mirror of include/gigaenv.h (struct layouts, enums). Shared by the product binding
(`__init__.py`) and by the test-only oracle binding (tests/oracle_ref.py)."""
import ctypes as C

GGL_MAX_CARS = 6
GGL_NUM_PADS = 34
GGL_NUM_ACTIONS = 90
GGL_MAX_REWARDS = 16
GGL_MAX_TERMINALS = 4
GGL_MAX_SETTER_CHILDREN = 4
GGL_MAX_OBS_PLAYERS = 6

OBS_DEFAULT, OBS_ADVANCED, OBS_DEFAULT_PADDED = 0, 1, 2

(REW_PLAYER_GOAL, REW_ASSIST, REW_SHOT, REW_SHOT_PASS, REW_SAVE, REW_BUMP, REW_BUMPED_PENALTY, REW_DEMO,
 REW_DEMOED_PENALTY, REW_GOAL, REW_VELOCITY, REW_VELOCITY_BALL_TO_GOAL, REW_VELOCITY_PLAYER_TO_BALL,
 REW_FACE_BALL, REW_TOUCH_BALL, REW_SPEED, REW_WAVEDASH, REW_PICKUP_BOOST, REW_SAVE_BOOST, REW_AIR,
 REW_TOUCH_ACCEL, REW_STRONG_TOUCH, REW_KICKOFF_PROXIMITY_2V2) = range(23)

TERM_GOAL_SCORE, TERM_NO_TOUCH = 0, 1
MODE_SOCCAR, MODE_HEATSEEKER = 0, 1
SETTER_KICKOFF, SETTER_INJECTED, SETTER_FUZZED_KICKOFF, SETTER_RANDOM, SETTER_COMBINED = 0, 1, 2, 3, 4
RANDOM_BALL_SPEED, RANDOM_CAR_SPEED, RANDOM_CARS_ON_GROUND = 1, 2, 4


class RewardDesc(C.Structure):
    _fields_ = [("kind", C.c_int32), ("weight", C.c_float), ("params", C.c_float * 4), ("zero_sum", C.c_int32)]


class TerminalDesc(C.Structure):
    _fields_ = [("kind", C.c_int32), ("params", C.c_float * 2)]


class MeshBlob(C.Structure):
    _fields_ = [("data", C.c_void_p), ("size", C.c_size_t)]


class SetterChild(C.Structure):
    _fields_ = [("kind", C.c_int32), ("flags", C.c_int32), ("weight", C.c_float)]


class EnvSetDesc(C.Structure):
    _fields_ = [
        ("device", C.c_int32), ("num_arenas", C.c_int32), ("players_per_team", C.c_int32),
        ("tick_skip", C.c_int32), ("action_delay", C.c_int32), ("obs_kind", C.c_int32),
        ("num_rewards", C.c_int32), ("rewards", RewardDesc * GGL_MAX_REWARDS),
        ("num_terminals", C.c_int32), ("terminals", TerminalDesc * GGL_MAX_TERMINALS),
        ("state_setter", C.c_int32), ("num_meshes", C.c_int32), ("meshes", C.POINTER(MeshBlob)),
        ("seed", C.c_uint64), ("slot_order_reversed", C.c_int32),
        ("obs_max_players", C.c_int32), ("state_setter_flags", C.c_int32), ("num_setter_children", C.c_int32),
        ("setter_children", SetterChild * GGL_MAX_SETTER_CHILDREN), ("game_mode", C.c_int32), ("obs_coefs", C.c_float * 5),
    ]


class Controls(C.Structure):
    _fields_ = [("throttle", C.c_float), ("steer", C.c_float), ("pitch", C.c_float), ("yaw", C.c_float),
                ("roll", C.c_float), ("jump", C.c_uint8), ("boost", C.c_uint8), ("handbrake", C.c_uint8),
                ("_pad", C.c_uint8)]


class WheelState(C.Structure):
    _fields_ = [("steer_angle", C.c_float), ("engine_force", C.c_float), ("brake", C.c_float),
                ("lat_friction", C.c_float), ("long_friction", C.c_float), ("extra_pushback", C.c_float)]


class CarState(C.Structure):
    _fields_ = [
        ("pos", C.c_float * 3), ("rot", C.c_float * 9), ("vel", C.c_float * 3), ("ang_vel", C.c_float * 3),
        ("flip_rel_torque", C.c_float * 3),
        ("jump_time", C.c_float), ("flip_time", C.c_float), ("air_time", C.c_float), ("air_time_since_jump", C.c_float),
        ("boost", C.c_float), ("time_spent_boosting", C.c_float), ("supersonic_time", C.c_float), ("handbrake_val", C.c_float),
        ("auto_flip_timer", C.c_float), ("auto_flip_torque_scale", C.c_float),
        ("world_contact_normal", C.c_float * 3), ("car_contact_cooldown", C.c_float), ("demo_respawn_timer", C.c_float),
        ("hit_rel_pos", C.c_float * 3), ("hit_ball_pos", C.c_float * 3), ("hit_extra_vel", C.c_float * 3),
        ("hit_tick", C.c_uint64), ("hit_tick_extra", C.c_uint64),
        ("car_contact_other_id", C.c_uint32), ("id", C.c_uint32),
        ("team", C.c_uint8),
        ("is_on_ground", C.c_uint8), ("has_jumped", C.c_uint8), ("has_double_jumped", C.c_uint8),
        ("has_flipped", C.c_uint8), ("is_flipping", C.c_uint8), ("is_jumping", C.c_uint8),
        ("is_supersonic", C.c_uint8), ("is_auto_flipping", C.c_uint8), ("world_contact_has", C.c_uint8),
        ("is_demoed", C.c_uint8), ("hit_valid", C.c_uint8),
        ("wheels_with_contact", C.c_uint8 * 4),
        ("last_controls", Controls), ("controls", Controls),
        ("wheels", WheelState * 4),
        ("update_counter", C.c_uint64), ("ray_aabb_min", C.c_float * 3), ("ray_aabb_max", C.c_float * 3),
    ]


class BallState(C.Structure):
    _fields_ = [("pos", C.c_float * 3), ("vel", C.c_float * 3), ("ang_vel", C.c_float * 3), ("update_counter", C.c_uint64),
                ("hs_y_target_dir", C.c_float), ("hs_cur_target_speed", C.c_float), ("hs_time_since_hit", C.c_float), ("_pad", C.c_float)]


class PadState(C.Structure):
    _fields_ = [("is_active", C.c_uint8), ("_pad", C.c_uint8 * 3), ("cooldown", C.c_float), ("prev_locked_car_id", C.c_uint32)]


class ArenaState(C.Structure):
    _fields_ = [("tick_count", C.c_uint64), ("num_cars", C.c_int32), ("_pad", C.c_int32), ("ball", BallState),
                ("cars", CarState * GGL_MAX_CARS), ("pads", PadState * GGL_NUM_PADS)]


class ExpBufTensors(C.Structure):
    _fields_ = [("states", C.c_void_p), ("actions", C.c_void_p), ("log_probs", C.c_void_p), ("target_values", C.c_void_p),
                ("action_masks", C.c_void_p), ("advantages", C.c_void_p)]


class TrajDesc(C.Structure):
    _fields_ = [("device", C.c_int32), ("num_players", C.c_int32), ("players_per_arena", C.c_int32), ("obs_size", C.c_int32),
                ("num_actions", C.c_int32), ("max_episode_length", C.c_int32), ("capacity_timesteps", C.c_int64),
                ("capacity_truncated", C.c_int64), ("capacity_player_rows", C.c_int64)]


class TrajViews(C.Structure):
    _fields_ = [("states", C.c_void_p), ("action_masks", C.c_void_p), ("actions", C.c_void_p), ("log_probs", C.c_void_p),
                ("rewards", C.c_void_p), ("terminals", C.c_void_p), ("next_states", C.c_void_p)]


class EnvSetViews(C.Structure):
    _fields_ = [("num_arenas", C.c_int32), ("num_players", C.c_int32), ("obs_size", C.c_int32),
                ("num_actions", C.c_int32), ("players_per_arena", C.c_int32),
                ("d_obs", C.c_void_p), ("d_action_masks", C.c_void_p), ("d_rewards", C.c_void_p),
                ("d_terminals", C.c_void_p), ("d_slot_car_id", C.c_void_p), ("d_slot_team", C.c_void_p),
                ("d_events", C.c_void_p), ("d_last_rewards", C.c_void_p)]


def make_desc(num_arenas, players_per_team, obs_kind=OBS_ADVANCED, rewards=None, terminals=None,
              tick_skip=8, action_delay=7, state_setter=SETTER_KICKOFF, meshes=None, seed=0, device=0,
              slot_order_reversed=1, obs_max_players=0, state_setter_flags=0, setter_children=None, obs_coefs=None, game_mode=0):
    """rewards: list of (kind, weight[, params[, zero_sum]]); terminals: list of (kind[, params]);
    setter_children (SETTER_COMBINED): list of (kind, weight[, flags]).
    Returns (desc, keepalive) — keepalive holds the mesh buffers the desc points into."""
    d = EnvSetDesc()
    d.device = device
    d.num_arenas = num_arenas
    d.players_per_team = players_per_team
    d.tick_skip = tick_skip
    d.action_delay = action_delay
    d.obs_kind = obs_kind
    rewards = rewards if rewards is not None else [(REW_VELOCITY_PLAYER_TO_BALL, 1.0), (REW_TOUCH_BALL, 1.0), (REW_GOAL, 10.0, [-1.0])]
    terminals = terminals if terminals is not None else [(TERM_GOAL_SCORE,), (TERM_NO_TOUCH, [10.0])]
    d.num_rewards = len(rewards)
    for i, r in enumerate(rewards):
        d.rewards[i].kind = r[0]
        d.rewards[i].weight = r[1]
        ps = r[2] if len(r) > 2 and r[2] is not None else []
        for j, p in enumerate(ps):
            d.rewards[i].params[j] = p
        d.rewards[i].zero_sum = int(r[3]) if len(r) > 3 else 0
    d.num_terminals = len(terminals)
    for i, t in enumerate(terminals):
        d.terminals[i].kind = t[0]
        ps = t[1] if len(t) > 1 else []
        for j, p in enumerate(ps):
            d.terminals[i].params[j] = p
    d.state_setter = state_setter
    keep = []
    if meshes is None:
        from . import arena_mesh
        meshes = arena_mesh.soccar_meshes()
    arr = (MeshBlob * len(meshes))()
    for i, b in enumerate(meshes):
        buf = C.create_string_buffer(b, len(b))
        keep.append(buf)
        arr[i].data = C.cast(buf, C.c_void_p)
        arr[i].size = len(b)
    keep.append(arr)
    d.num_meshes = len(meshes)
    d.meshes = C.cast(arr, C.POINTER(MeshBlob))
    d.seed = seed
    d.slot_order_reversed = slot_order_reversed
    d.obs_max_players = obs_max_players
    d.state_setter_flags = state_setter_flags
    d.game_mode = game_mode
    for i, k in enumerate(obs_coefs or []):      # DefaultObs(posCoef xyz, velCoef, angVelCoef); None = the defaults
        d.obs_coefs[i] = k
    setter_children = setter_children or []
    d.num_setter_children = len(setter_children)
    for i, ch in enumerate(setter_children):
        d.setter_children[i].kind = ch[0]
        d.setter_children[i].weight = ch[1]
        d.setter_children[i].flags = ch[2] if len(ch) > 2 else 0
    return d, keep
