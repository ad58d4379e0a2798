// TEST INFRASTRUCTURE — NOT PART OF THE PRODUCT.
// Host (g++) compile of the device simulator headers under reinforcement-learning_b200/csrc so that the tick logic
// can be lock-stepped against the compiled reference (oracle/_ref) in a container WITHOUT a GPU.  The product
// library (libgigaenv_cuda.so) never links or loads this; it only launches CUDA kernels built from the same headers.
//   g++ -O2 -ffp-contract=off -shared -fPIC -I../../include hostsim.cpp -o _build/libhostsim.so
#include <cstdio>
#define GGL_ORDER_DEBUG 1      // keeps the manifold array solver_order predicted (hostsim_predicted_manifolds)
#include "../../reinforcement-learning_b200/csrc/sim_convert.cuh"
#include "../../reinforcement-learning_b200/csrc/env_reset.cuh"

using namespace ggl;

struct HostSim {
    WorldHost world;
    ArenaDev arena;
    bool orderSet = false;
};

extern "C" {

void* hostsim_create(const ggl_mesh_blob* meshes, int n) {
    HostSim* h = new HostSim();
    for (int i = 0; i < n; i++)
        if (!world_add_cmf(h->world, meshes[i].data, meshes[i].size)) { delete h; return nullptr; }
    world_build_grid(h->world);
    memset(&h->arena, 0, sizeof(h->arena));
    h->arena.rng = 1;
    return h;
}
void hostsim_destroy(void* p) { delete (HostSim*)p; }
// vertices of stored triangle i (9 floats) and its mesh index: the storage order is the reference's BVH leaf order per mesh
int hostsim_tri(void* p, int i, float* out9) {
    const Tri& t = ((HostSim*)p)->world.tris[i];
    out9[0] = t.v0.x; out9[1] = t.v0.y; out9[2] = t.v0.z; out9[3] = t.v1.x; out9[4] = t.v1.y; out9[5] = t.v1.z;
    out9[6] = t.v2.x; out9[7] = t.v2.y; out9[8] = t.v2.z;
    return (int)(t.flags >> TRI_MESH_SHIFT);
}
// the empty-cell box of the world (Bullet units): out6 = min xyz, max xyz
void hostsim_free_box(void* p, float* out6) {
    const WorldHost& W = ((HostSim*)p)->world;
    out6[0] = W.freeMin.x; out6[1] = W.freeMin.y; out6[2] = W.freeMin.z; out6[3] = W.freeMax.x; out6[4] = W.freeMax.y; out6[5] = W.freeMax.z;
}
int hostsim_num_tris(void* p) { return (int)((HostSim*)p)->world.tris.size(); }

void hostsim_set_state(void* p, const ggl_arena_state* st) {
    HostSim* h = (HostSim*)p;
    arena_phys_from_host(*st, h->arena);
    if (!h->orderSet)
        for (int s = 0; s < h->arena.numCars; s++) h->arena.slotCar[s] = (uint8_t)(h->arena.numCars - 1 - s);
}
// players[] slot order (= the reference's car-set iteration order, which also orders the cars' pre-ticks): ids[slot] = carId
void hostsim_set_slot_order(void* p, const int32_t* ids) {
    HostSim* h = (HostSim*)p;
    for (int s = 0; s < h->arena.numCars; s++) h->arena.slotCar[s] = (uint8_t)(ids[s] - 1);
    h->orderSet = true;
}
void hostsim_get_state(void* p, ggl_arena_state* st) { arena_phys_to_host(((HostSim*)p)->arena, *st); }
void hostsim_set_game_mode(void* p, int mode) { ((HostSim*)p)->arena.gameMode = (uint8_t)mode; }
void hostsim_seed(void* p, uint32_t seed) { ((HostSim*)p)->arena.rng = seed; }

void hostsim_step(void* p, int ticks) {
    HostSim* h = (HostSim*)p;
    WorldDev W = world_view(h->world);
    ExecSerial ex;
    for (int i = 0; i < ticks; i++) { order_debug().n = -1; arena_tick(h->arena, W, ex); }
}

// contacts of the last tick: [n][20] like ref_arena_contacts (bodyA, bodyB, normal, posA, posB, dist, friction, restitution, special)
int hostsim_contacts(void* p, float* out, int max) {
    HostSim* h = (HostSim*)p;
    const ArenaDev& A = h->arena;
    int n = 0;
    for (int k = 0; k < A.numContacts && n < max; k++) {
        const Contact& c = A.contacts[A.contactOrder[k]];
        float* o = out + 20 * n++;
        memset(o, 0, 80);
        o[0] = c.bodyA; o[1] = c.bodyB;
        o[2] = c.normal.x; o[3] = c.normal.y; o[4] = c.normal.z;
        o[11] = c.dist; o[12] = c.friction; o[13] = c.restitution; o[14] = c.special;
        o[5] = c.relA.x; o[6] = c.relA.y; o[7] = c.relA.z; o[8] = c.relB.x; o[9] = c.relB.y; o[10] = c.relB.z;
    }
    return n;
}


// The manifold array solver_order (sim_world.cuh) predicted for the last tick, coded like ref_arena_manifolds: out[i] = {body0,
// body1, island key}; body = 0 ball, car id, or -(100 + static index) (planes 1000 further down). mode: 0 = as shipped (only
// when an island mixes manifolds), 1 = predict on every tick with a contact, 2 = never reorder. Returns -1 if the last tick
// did not predict; flags[0] = rows were reordered, flags[1] = a row's manifold was not in the array.
void hostsim_order_mode(int mode) { order_debug().always = mode == 1; order_debug().off = mode == 2; }
int hostsim_predicted_manifolds(void* p, int* out, int max, int* flags) {
    HostSim* h = (HostSim*)p;
    const OrderDebug& D = order_debug();
    flags[0] = D.reordered; flags[1] = D.unknownRow;
    if (D.n < 0) return -1;
    const int planeBase = h->world.numStatics - 4;
    auto code = [&](int b) { return b == 0 ? 0 : (b > 0 ? (int)h->arena.cars[b - 1].id : (-(1 + b) >= planeBase ? -(1100 + (-(1 + b))) : -(100 + (-(1 + b))))); };
    for (int k = 0; k < D.n && k < max; k++) { out[3 * k] = code(D.mf[k].dyn); out[3 * k + 1] = code(D.mf[k].other); out[3 * k + 2] = D.key[k]; }
    return D.n;
}

// order_sort_keys (sim_world.cuh) on its own: idx[] = the permutation it applies to n manifolds with these island ids
void hostsim_sort_keys(const int* keys, int* idx, int n) {
    uint8_t k[kMaxOrderManifolds], ix[kMaxOrderManifolds], lo[kMaxOrderManifolds], hi[kMaxOrderManifolds];
    for (int i = 0; i < n; i++) { k[i] = (uint8_t)keys[i]; ix[i] = (uint8_t)i; }
    if (n > 1) order_sort_keys(k, ix, n, lo, hi);
    for (int i = 0; i < n; i++) idx[i] = ix[i];
}

// the registry's static boxes ([n][6], like ref_arena_static_boxes) and the boxes the last prediction used for the ball and the
// cars in slot order ([1 + cars][6])
int hostsim_static_boxes(void* p, float* out, int max) {
    HostSim* h = (HostSim*)p;
    int n = h->world.numStatics < max ? h->world.numStatics : max;
    memcpy(out, h->world.bpStaticBox.data(), sizeof(float) * 6 * n);
    return n;
}
void hostsim_predicted_boxes(void* p, float* out) {
    HostSim* h = (HostSim*)p;
    const OrderDebug& D = order_debug();
    for (int b = 0; b <= h->arena.numCars; b++) {
        float* o = out + 6 * b;
        o[0] = D.mn[b].x; o[1] = D.mn[b].y; o[2] = D.mn[b].z; o[3] = D.mx[b].x; o[4] = D.mx[b].y; o[5] = D.mx[b].z;
    }
}

// same call as ref_convex_ray_cast (oracle/ref_harness.cpp): the device's ray-vs-hitbox / ray-vs-ball cast
int hostsim_convex_ray_cast(int kind, const float* from, const float* to, const float* pos, const float* basis, float* out) {
    RayHit h; h.t = 1.f; h.ground = GROUND_NONE; h.normal = v3(0, 0, 0);
    V3 f = v3(from[0], from[1], from[2]), t = v3(to[0], to[1], to[2]);
    if (kind == 0) {
        CarDev c; memset(&c, 0, sizeof(c));
        c.pos = v3(pos[0], pos[1], pos[2]);
        c.basis.r0 = v3(basis[0], basis[1], basis[2]); c.basis.r1 = v3(basis[3], basis[4], basis[5]); c.basis.r2 = v3(basis[6], basis[7], basis[8]);
        ray_vs_car_box(f, t, c, GROUND_CAR0, h);
    } else {
        ray_vs_sphere(f, t, v3(pos[0], pos[1], pos[2]), rl::BALL_RADIUS * kUU2BT, GROUND_BALL, h);
    }
    if (h.ground == GROUND_NONE) return 0;
    out[0] = h.t; out[1] = h.normal.x; out[2] = h.normal.y; out[3] = h.normal.z;
    return 1;
}

// same layout as ref_arena_wheel_info (oracle/ref_harness.cpp)
void hostsim_wheel_info(void* p, int carId, float* out) {
    HostSim* h = (HostSim*)p;
    const CarDev& c = h->arena.cars[carId - 1];
    for (int i = 0; i < 4; i++) {
        const Wheel& w = c.w[i];
        float* o = out + 12 * i;
        o[0] = w.contactNormal.x; o[1] = w.contactNormal.y; o[2] = w.contactNormal.z;
        o[3] = w.contactPoint.x; o[4] = w.contactPoint.y; o[5] = w.contactPoint.z;
        o[6] = w.suspLen; o[7] = (float)w.ground;
        o[8] = w.impulse.x; o[9] = w.impulse.y; o[10] = w.impulse.z; o[11] = w.suspRelVel;
    }
}

void hostsim_set_controls(void* p, const ggl_arena_state* st) {
    HostSim* h = (HostSim*)p;
    for (int i = 0; i < h->arena.numCars; i++) h->arena.cars[i].controls = controls_from_host(st->cars[i].controls);
}

size_t hostsim_sizeof_arena() { return sizeof(ArenaDev); }

}  // extern "C"

// ------------------------------------------------------------------------------------------------------------
// Host mirror of env_kernels.cuh's orchestration (same phase functions, one lane) for lock-stepping the env layer.
struct HostEnv {
    EnvCfg cfg;
    WorldHost world;
    std::vector<ArenaDev> arenas, resets;
    std::vector<float> obs, rewards, lastRewards, table;
    std::vector<uint8_t> masks, terminals, events;
};

static void host_write_rows(HostEnv* h, int a) {
    ArenaDev& A = h->arenas[a];
    for (int s = 0; s < A.numCars; s++) {
        int row = a * h->cfg.numCars + s;
        uint32_t pm = 0, po = 0;
        if (h->cfg.obsKind == GGL_OBS_DEFAULT_PADDED) padded_obs_draw(A.rng, h->cfg.obsMaxPlayers, pm, po);
        build_obs_row(A, h->cfg, s, h->obs.data() + (size_t)row * h->cfg.obsSize, pm, po);
        action_mask_row(A.cars[A.slotCar[s]], h->cfg.actionBits, h->masks.data() + (size_t)row * kNumActions);
    }
}

static void host_reset_arena(HostEnv* h, int a, bool force, int setter) {
    ArenaDev& A = h->arenas[a];
    if (!force && A.terminal == 0) return;
    EnvCfg cfg = h->cfg;
    cfg.stateSetter = setter;
    arena_state_set(A, cfg, &h->resets[a]);
    env_after_state_set(A);
    h->terminals[a] = 0;
    host_write_rows(h, a);
    for (int s = 0; s < A.numCars; s++)
        for (int e = 0; e < 9; e++) h->events[(size_t)(a * h->cfg.numCars + s) * 9 + e] = 0;
}

extern "C" {

void* hostenv_create(const ggl_envset_desc* d) {
    HostEnv* h = new HostEnv();
    EnvCfg& c = h->cfg;
    memset(&c, 0, sizeof(c));
    c.numArenas = d->num_arenas; c.playersPerTeam = d->players_per_team; c.numCars = 2 * d->players_per_team;
    c.tickSkip = d->tick_skip; c.actionDelay = d->action_delay;
    if (const char* why = env_cfg_plugins(c, d)) { fprintf(stderr, "hostenv_create: %s\n", why); delete h; return nullptr; }
    c.numRewards = d->num_rewards; c.numTerminals = d->num_terminals;
    for (int i = 0; i < d->num_rewards; i++) {
        c.rewards[i].kind = d->rewards[i].kind; c.rewards[i].weight = d->rewards[i].weight;
        for (int k = 0; k < 4; k++) c.rewards[i].params[k] = d->rewards[i].params[k];
    }
    for (int i = 0; i < d->num_terminals; i++) {
        c.terminals[i].kind = d->terminals[i].kind; c.terminals[i].params[0] = d->terminals[i].params[0];
    }
    for (int i = 0; i < d->num_meshes; i++) world_add_cmf(h->world, d->meshes[i].data, d->meshes[i].size);
    world_build_grid(h->world);
    const int np = c.numArenas * c.numCars;
    h->arenas.resize(c.numArenas); h->resets.resize(c.numArenas);
    memset(h->arenas.data(), 0, h->arenas.size() * sizeof(ArenaDev));
    memset(h->resets.data(), 0, h->resets.size() * sizeof(ArenaDev));
    h->obs.assign((size_t)np * c.obsSize, 0.f); h->rewards.assign(np, 0.f); h->lastRewards.assign((size_t)c.numArenas * (c.numRewards + 1), 0.f);
    h->masks.assign((size_t)np * kNumActions, 0); h->terminals.assign(c.numArenas, 0); h->events.assign((size_t)np * 9, 0);
    h->table.resize(kNumActions * 8);
    for (int i = 0; i < kNumActions; i++) action_from_index(i, h->table.data() + 8 * i);
    action_bits_from_table(h->table.data(), c.actionBits);
    for (int a = 0; a < c.numArenas; a++) {
        ArenaDev& A = h->arenas[a];
        A.numCars = c.numCars; A.gameMode = (uint8_t)c.gameMode; A.rng = (uint32_t)(d->seed * 7919u + a * 104729u + 1u) % 2147483646u + 1u;
        for (int i = 0; i < c.numCars; i++) {
            CarDev& car = A.cars[i];
            car.id = i + 1; car.team = i < c.playersPerTeam ? 0 : 1; car.basis = m3_identity();
            car_reset_state_defaults(car); car_update_inertia(car);
            for (int w = 0; w < 4; w++) car.w[w].ground = GROUND_NONE;
        }
        for (int s = 0; s < c.numCars; s++) A.slotCar[s] = (uint8_t)(d->slot_order_reversed ? c.numCars - 1 - s : s);
        A.lastTouchCarId = -1;
        host_reset_arena(h, a, true, c.stateSetter == GGL_SETTER_INJECTED ? GGL_SETTER_KICKOFF : c.stateSetter);
    }
    return h;
}
void hostenv_destroy(void* p) { delete (HostEnv*)p; }

void hostenv_set_state(void* p, int a, const ggl_arena_state* st) { arena_phys_from_host(*st, ((HostEnv*)p)->arenas[a]); }
void hostenv_get_state(void* p, int a, ggl_arena_state* st) { arena_phys_to_host(((HostEnv*)p)->arenas[a], *st); }
void hostenv_set_reset_state(void* p, int a, const ggl_arena_state* st) {
    HostEnv* h = (HostEnv*)p;
    memset(&h->resets[a], 0, sizeof(ArenaDev));
    arena_phys_from_host(*st, h->resets[a]);
    h->resets[a].hasResetState = 1;
}
uint64_t hostenv_contact_overflow(void* p) {
    HostEnv* h = (HostEnv*)p;
    uint64_t n = 0;
    for (auto& A : h->arenas) n += A.contactOverflow;
    return n;
}
void hostenv_reset_arena(void* p, int a) {
    HostEnv* h = (HostEnv*)p;
    for (int i = 0; i < h->cfg.numArenas; i++)
        if (a < 0 || i == a) host_reset_arena(h, i, true, h->cfg.stateSetter);
}
void hostenv_set_rng(void* p, int a, uint32_t state) {
    HostEnv* h = (HostEnv*)p;
    for (int i = 0; i < h->cfg.numArenas; i++) if (a < 0 || i == a) h->arenas[i].rng = state;
}
uint32_t hostenv_get_rng(void* p, int a) { return ((HostEnv*)p)->arenas[a].rng; }
void hostenv_set_slot_order(void* p, const int32_t* ids) {
    HostEnv* h = (HostEnv*)p;
    for (int a = 0; a < h->cfg.numArenas; a++)
        for (int s = 0; s < h->cfg.numCars; s++) h->arenas[a].slotCar[s] = (uint8_t)(ids[a * h->cfg.numCars + s] - 1);
}

// Reset + StepFirstHalf + StepSecondHalf (the order env_reset_kernel / env_step_kernel run)
void hostenv_step(void* p, const int32_t* actions) {
    HostEnv* h = (HostEnv*)p;
    const EnvCfg& cfg = h->cfg;
    WorldDev W = world_view(h->world);
    ExecSerial ex;
    for (int a = 0; a < cfg.numArenas; a++) {
        host_reset_arena(h, a, false, cfg.stateSetter);
        ArenaDev& A = h->arenas[a];
        env_begin_step(A);
        for (int t = 0; t < cfg.actionDelay; t++) arena_tick(A, W, ex);
        env_apply_actions(A, actions + a * cfg.numCars, h->table.data());
        for (int t = cfg.actionDelay; t < cfg.tickSkip; t++) arena_tick(A, W, ex);
        env_after_ticks(A, cfg);
        h->terminals[a] = A.terminal;
        int lowSlot = 0;
        for (int s = 1; s < A.numCars; s++)
            if (A.cars[A.slotCar[s]].id < A.cars[A.slotCar[lowSlot]].id) lowSlot = s;
        for (int s = 0; s < A.numCars; s++) {
            float raw[kMaxRewards];
            int row = a * cfg.numCars + s;
            h->rewards[row] = env_player_reward(A, cfg, s, raw);
            if (s == lowSlot) for (int r = 0; r < cfg.numRewards; r++) h->lastRewards[(size_t)a * cfg.numRewards + r] = raw[r];
            for (int e = 0; e < 9; e++) h->events[(size_t)row * 9 + e] = A.players[s].events[e];
        }
        host_write_rows(h, a);
        env_end_step(A);
    }
}

void hostenv_copy_out(void* p, float* obs, uint8_t* masks, float* rewards, uint8_t* terminals, uint8_t* events, float* lastRewards) {
    HostEnv* h = (HostEnv*)p;
    if (obs) memcpy(obs, h->obs.data(), h->obs.size() * 4);
    if (masks) memcpy(masks, h->masks.data(), h->masks.size());
    if (rewards) memcpy(rewards, h->rewards.data(), h->rewards.size() * 4);
    if (terminals) memcpy(terminals, h->terminals.data(), h->terminals.size());
    if (events) memcpy(events, h->events.data(), h->events.size());
    if (lastRewards) memcpy(lastRewards, h->lastRewards.data(), (size_t)h->cfg.numArenas * h->cfg.numRewards * 4);
}

}  // extern "C"
