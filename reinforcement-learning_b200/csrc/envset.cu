// C ABI of the device-resident EnvSet (include/gigaenv.h). Host side: descriptor translation, device memory,
// kernel launches on the caller's stream. No CPU fallback: every entry point that computes launches a kernel.
#include <cstdlib>
#include "common.cuh"
#include "sim_convert.cuh"
#include "env_kernels.cuh"

#include <new>
#include <vector>

using namespace ggl;

struct LaunchGeom { int warps, threads, grid; size_t smem; };

struct ggl_envset {
    EnvCfg cfg;
    EnvDevPtrs P;
    int device;
    int numPlayers;
    LaunchGeom geom;
    WorldHost worldHost;
    Tri* d_tris; int* d_cellStart; int* d_cellTris; uint32_t* d_bpCellStatics; float* d_bpStaticBox;
    ArenaDev* d_reset;
    float* d_table;
    int32_t* d_actionsScratch;
    int32_t* d_arenaOfPos; int32_t* d_balanceOrder; uint32_t* d_stepCost;      // block load balancing (balance_kernel)
    bool balance;
    // pinned staging for the host-buffer wrapper
    int32_t* h_actions; float* h_obs; uint8_t* h_masks; float* h_rewards; uint8_t* h_terminals;
    cudaStream_t ownStream, copyStream;
    cudaEvent_t evReset, evCopied;
    uint64_t seed;
};

namespace {

int grid_for(const EnvCfg& cfg) { return (cfg.numArenas + kArenasPerBlock - 1) / kArenasPerBlock; }   // reset kernel
int slot_bytes(const EnvCfg& cfg) { return (int)arena_slot_bytes(cfg.numCars); }

// Launch geometry of the step kernels: as many warps per block (<= kMaxWarps, <= what fits in 227 KB of shared memory) as it
// takes to cover the arenas with one block per SM, so that 4096 arenas of 2v2 run as 147 blocks of 7 warps in a single wave.
LaunchGeom launch_geom(const EnvCfg& cfg, int numSMs) {
    // unit = 4 arenas (one warp at 8 lanes per arena, half a warp at 4, two warps at 16); at most 7 units per block
    const int unitThreads = 4 * kGroup;
    const int totalUnits = (cfg.numArenas + 3) / 4;
    const size_t perUnit = (size_t)4 * slot_bytes(cfg);
    int fit = (int)((227 * 1024 - 1024) / perUnit);
    if (fit > 7) fit = 7;
    if (fit < 1) fit = 1;
    int units = (totalUnits + numSMs - 1) / numSMs;
    if (units > fit) units = fit;
    if (const char* cap = getenv("GGL_DEBUG_MAX_BLOCK_WARPS")) {     // test hook: same arenas under another block geometry
        const int c = atoi(cap);
        if (c >= 1 && c < units) units = c;
    }
    if (units < 1) units = 1;
    LaunchGeom g;
    g.warps = units;
    g.threads = units * unitThreads;
    g.grid = (totalUnits + units - 1) / units;
    g.smem = perUnit * units;
    return g;
}


template <class T>
int dmalloc(T** p, size_t n) {
    cudaError_t e = cudaMalloc((void**)p, n * sizeof(T));
    if (e != cudaSuccess) { ggl::set_error("cudaMalloc(%zu bytes) -> %s", n * sizeof(T), cudaGetErrorString(e)); return GGL_ERR_NOMEM; }
    return GGL_OK;
}

void free_env(ggl_envset* e) {
    if (!e) return;
    cudaFree(e->P.arenas); cudaFree(e->d_reset); cudaFree(e->P.obs); cudaFree(e->P.masks); cudaFree(e->P.rewards);
    cudaFree(e->P.terminals); cudaFree(e->P.slotCarId); cudaFree(e->P.slotTeam); cudaFree(e->P.events); cudaFree(e->P.lastRewards);
    cudaFree(e->d_tris); cudaFree(e->d_cellStart); cudaFree(e->d_cellTris); cudaFree(e->d_bpCellStatics); cudaFree(e->d_bpStaticBox); cudaFree(e->d_table); cudaFree(e->d_actionsScratch);
    cudaFree(e->d_arenaOfPos); cudaFree(e->d_balanceOrder); cudaFree(e->d_stepCost);
    if (e->h_actions) cudaFreeHost(e->h_actions);
    if (e->h_obs) cudaFreeHost(e->h_obs);
    if (e->h_masks) cudaFreeHost(e->h_masks);
    if (e->h_rewards) cudaFreeHost(e->h_rewards);
    if (e->h_terminals) cudaFreeHost(e->h_terminals);
    if (e->ownStream) cudaStreamDestroy(e->ownStream);
    if (e->copyStream) cudaStreamDestroy(e->copyStream);
    if (e->evReset) cudaEventDestroy(e->evReset);
    if (e->evCopied) cudaEventDestroy(e->evCopied);
    delete e;
}

}  // namespace

extern "C" int ggl_action_table(float* out_table, int* out_num_actions) {
    if (!out_table) { ggl::set_error("ggl_action_table: null"); return GGL_ERR_INVALID; }
    for (int i = 0; i < kNumActions; i++) action_from_index(i, out_table + 8 * i);
    if (out_num_actions) *out_num_actions = kNumActions;
    return GGL_OK;
}

extern "C" int ggl_envset_create(const ggl_envset_desc* d, ggl_envset** out) {
    if (!d || !out) { ggl::set_error("ggl_envset_create: null argument"); return GGL_ERR_INVALID; }
    if (d->num_arenas <= 0 || d->players_per_team < 1 || d->players_per_team > 3 || d->tick_skip <= 0 || d->action_delay < 0 ||
        d->action_delay > d->tick_skip || d->num_rewards < 0 || d->num_rewards > GGL_MAX_REWARDS || d->num_terminals < 0 ||
        d->num_terminals > GGL_MAX_TERMINALS) {
        ggl::set_error("ggl_envset_create: descriptor out of range"); return GGL_ERR_INVALID;
    }
    {
        EnvCfg probe;
        memset(&probe, 0, sizeof(probe));
        if (const char* why = env_cfg_plugins(probe, d)) { ggl::set_error("ggl_envset_create: %s", why); return GGL_ERR_UNSUPPORTED; }
    }
    for (int i = 0; i < d->num_rewards; i++)
        if (d->rewards[i].kind < 0 || d->rewards[i].kind >= GGL_REW_KIND_COUNT) { ggl::set_error("unsupported reward kind %d", d->rewards[i].kind); return GGL_ERR_UNSUPPORTED; }
    for (int i = 0; i < d->num_terminals; i++)
        if (d->terminals[i].kind != GGL_TERM_GOAL_SCORE && d->terminals[i].kind != GGL_TERM_NO_TOUCH) { ggl::set_error("unsupported terminal kind"); return GGL_ERR_UNSUPPORTED; }
    GGL_CUDA(cudaSetDevice(d->device));
    ggl_envset* e = new (std::nothrow) ggl_envset();
    if (!e) return GGL_ERR_NOMEM;
    memset(&e->cfg, 0, sizeof(e->cfg)); memset(&e->P, 0, sizeof(e->P));
    e->d_tris = nullptr; e->d_cellStart = e->d_cellTris = nullptr; e->d_bpCellStatics = nullptr; e->d_bpStaticBox = nullptr; e->d_reset = nullptr; e->d_table = nullptr; e->d_actionsScratch = nullptr;
    e->d_arenaOfPos = e->d_balanceOrder = nullptr; e->d_stepCost = nullptr; e->balance = false;
    e->h_actions = nullptr; e->h_obs = nullptr; e->h_masks = nullptr; e->h_rewards = nullptr; e->h_terminals = nullptr; e->ownStream = nullptr; e->copyStream = nullptr; e->evReset = nullptr; e->evCopied = nullptr;
    e->device = d->device; e->seed = d->seed;
    EnvCfg& c = e->cfg;
    c.numArenas = d->num_arenas; c.playersPerTeam = d->players_per_team; c.numCars = 2 * d->players_per_team;
    c.tickSkip = d->tick_skip; c.actionDelay = d->action_delay; env_cfg_plugins(c, d);
    c.numRewards = d->num_rewards; c.numTerminals = d->num_terminals;
    for (int i = 0; i < d->num_rewards; i++) {
        c.rewards[i].kind = d->rewards[i].kind; c.rewards[i].weight = d->rewards[i].weight;
        for (int k = 0; k < 4; k++) c.rewards[i].params[k] = d->rewards[i].params[k];
    }
    for (int i = 0; i < d->num_terminals; i++) {
        c.terminals[i].kind = d->terminals[i].kind;
        c.terminals[i].params[0] = d->terminals[i].params[0]; c.terminals[i].params[1] = d->terminals[i].params[1];
    }
    e->numPlayers = c.numArenas * c.numCars;
    // arena meshes -> triangle grid
    for (int i = 0; i < d->num_meshes; i++)
        if (!d->meshes || !world_add_cmf(e->worldHost, d->meshes[i].data, d->meshes[i].size)) {
            ggl::set_error("ggl_envset_create: malformed .cmf mesh %d", i); free_env(e); return GGL_ERR_INVALID;
        }
    world_build_grid(e->worldHost);
    int rc = GGL_OK;
    const size_t nt = e->worldHost.tris.size() ? e->worldHost.tris.size() : 1, ncs = e->worldHost.cellStart.size(),
                 nct = e->worldHost.cellTris.size() ? e->worldHost.cellTris.size() : 1;
    if ((rc = dmalloc(&e->d_bpCellStatics, e->worldHost.bpCellStatics.size())) || (rc = dmalloc(&e->d_bpStaticBox, e->worldHost.bpStaticBox.size())) || (rc = dmalloc(&e->d_tris, nt)) || (rc = dmalloc(&e->d_cellStart, ncs)) || (rc = dmalloc(&e->d_cellTris, nct)) ||
        (rc = dmalloc(&e->P.arenas, (size_t)c.numArenas)) || (rc = dmalloc(&e->P.obs, (size_t)e->numPlayers * c.obsSize)) ||
        (rc = dmalloc(&e->P.masks, (size_t)e->numPlayers * kNumActions)) || (rc = dmalloc(&e->P.rewards, (size_t)e->numPlayers)) ||
        (rc = dmalloc(&e->P.terminals, (size_t)c.numArenas)) || (rc = dmalloc(&e->P.slotCarId, (size_t)e->numPlayers)) ||
        (rc = dmalloc(&e->P.slotTeam, (size_t)e->numPlayers)) || (rc = dmalloc(&e->P.events, (size_t)e->numPlayers * 9)) ||
        (rc = dmalloc(&e->P.lastRewards, (size_t)c.numArenas * (c.numRewards ? c.numRewards : 1))) ||
        (rc = dmalloc(&e->d_table, (size_t)kNumActions * 8)) || (rc = dmalloc(&e->d_actionsScratch, (size_t)e->numPlayers))) {
        free_env(e); return rc;
    }
    if (e->worldHost.tris.size()) cudaMemcpy(e->d_tris, e->worldHost.tris.data(), e->worldHost.tris.size() * sizeof(Tri), cudaMemcpyHostToDevice);
    cudaMemcpy(e->d_cellStart, e->worldHost.cellStart.data(), ncs * sizeof(int), cudaMemcpyHostToDevice);
    if (e->worldHost.cellTris.size()) cudaMemcpy(e->d_cellTris, e->worldHost.cellTris.data(), e->worldHost.cellTris.size() * sizeof(int), cudaMemcpyHostToDevice);
    WorldDev w = world_view(e->worldHost);
    w.tris = e->d_tris; w.cellStart = e->d_cellStart; w.cellTris = e->d_cellTris;
    cudaMemcpy(e->d_bpCellStatics, e->worldHost.bpCellStatics.data(), e->worldHost.bpCellStatics.size() * sizeof(uint32_t), cudaMemcpyHostToDevice);
    cudaMemcpy(e->d_bpStaticBox, e->worldHost.bpStaticBox.data(), e->worldHost.bpStaticBox.size() * sizeof(float), cudaMemcpyHostToDevice);
    w.bpCellStatics = e->d_bpCellStatics; w.bpStaticBox = e->d_bpStaticBox;
    e->P.world = w;
    float table[kNumActions * 8];
    ggl_action_table(table, nullptr);
    action_bits_from_table(table, e->cfg.actionBits);
    cudaMemcpy(e->d_table, table, sizeof(table), cudaMemcpyHostToDevice);
    e->P.actionTable = e->d_table;
    // initial arena blobs: cars exist (AddCar order: blue first), everything else comes from the first reset
    {
        std::vector<ArenaDev> init((size_t)c.numArenas);
        memset(init.data(), 0, init.size() * sizeof(ArenaDev));
        for (int a = 0; a < c.numArenas; a++) {
            ArenaDev& A = init[a];
            A.numCars = c.numCars;
            A.gameMode = (uint8_t)c.gameMode;
            uint64_t s = (d->seed * 6364136223846793005ull + 1442695040888963407ull) ^ ((uint64_t)(a + 1) * 0x9E3779B97F4A7C15ull);
            A.rng = (uint32_t)(s % 2147483646ull) + 1u;
            for (int i = 0; i < c.numCars; i++) {
                CarDev& car = A.cars[i];
                car.id = (uint32_t)(i + 1);
                car.team = i < c.playersPerTeam ? 0 : 1;
                car.basis = m3_identity();
                car_reset_state_defaults(car);
                car_update_inertia(car);
                for (int w2 = 0; w2 < 4; w2++) car.w[w2].ground = GROUND_NONE;
            }
            for (int s2 = 0; s2 < c.numCars; s2++) A.slotCar[s2] = (uint8_t)(d->slot_order_reversed ? c.numCars - 1 - s2 : s2);
            A.lastTouchCarId = -1;
        }
        cudaMemcpy(e->P.arenas, init.data(), init.size() * sizeof(ArenaDev), cudaMemcpyHostToDevice);
    }
    cudaMemset(e->P.terminals, 0, (size_t)c.numArenas);
    cudaMemset(e->P.rewards, 0, (size_t)e->numPlayers * sizeof(float));
    cudaMemset(e->P.lastRewards, 0, (size_t)c.numArenas * (c.numRewards ? c.numRewards : 1) * sizeof(float));
    GGL_CUDA(cudaStreamCreateWithFlags(&e->ownStream, cudaStreamNonBlocking));
    GGL_CUDA(cudaStreamCreateWithFlags(&e->copyStream, cudaStreamNonBlocking));
    GGL_CUDA(cudaEventCreateWithFlags(&e->evReset, cudaEventDisableTiming));
    GGL_CUDA(cudaEventCreateWithFlags(&e->evCopied, cudaEventDisableTiming));
    {
        int sms = 148;
        GGL_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, d->device));
        e->geom = launch_geom(e->cfg, sms);
        // Load balancing between blocks only matters once a launch has several multi-warp blocks; GGL_BALANCE=0 turns it off
        // (arenas then sit in index order, as in round 1).
#if defined(GGL_COST_SWEEP)
        if (const char* cw = getenv("GGL_COST_W")) {
            int w[3] = {4, 6, 1};
            sscanf(cw, "%d,%d,%d", &w[0], &w[1], &w[2]);
            cudaMemcpyToSymbol(ggl::g_costW, w, sizeof(w));
        }
#endif
        const char* bal = getenv("GGL_BALANCE");
        e->balance = e->geom.grid > 1 && !(bal && atoi(bal) == 0);
        if (e->balance) {
            int rc2;
            if ((rc2 = dmalloc(&e->d_arenaOfPos, (size_t)e->cfg.numArenas)) || (rc2 = dmalloc(&e->d_balanceOrder, (size_t)e->cfg.numArenas)) ||
                (rc2 = dmalloc(&e->d_stepCost, (size_t)e->cfg.numArenas))) { free_env(e); return rc2; }
            std::vector<int32_t> ident(e->cfg.numArenas);
            for (int i = 0; i < e->cfg.numArenas; i++) ident[i] = i;
            cudaMemcpy(e->d_arenaOfPos, ident.data(), ident.size() * sizeof(int32_t), cudaMemcpyHostToDevice);
            cudaMemset(e->d_stepCost, 0, (size_t)e->cfg.numArenas * sizeof(uint32_t));
            e->P.arenaOfPos = e->d_arenaOfPos; e->P.stepCost = e->d_stepCost;
        }
        // the attribute is per function, not per envset: always allow the largest block any envset may ask for
        GGL_CUDA(cudaFuncSetAttribute(env_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 1024));
        GGL_CUDA(cudaFuncSetAttribute(arena_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 1024));
    }
    // EnvSet ctor: reset every arena once (EnvSet.cpp:106-110)
    EnvCfg kc = e->cfg;
    if (kc.stateSetter == GGL_SETTER_INJECTED) kc.stateSetter = GGL_SETTER_KICKOFF;     // no states supplied yet
    env_reset_kernel<<<grid_for(e->cfg), kBlock, 0, 0>>>(kc, e->P, 1);
    ggl::count_launch();
    cudaError_t err = cudaDeviceSynchronize();
    if (err != cudaSuccess) { ggl::set_error("env_reset_kernel -> %s", cudaGetErrorString(err)); free_env(e); return GGL_ERR_CUDA; }
    *out = e;
    return GGL_OK;
}

extern "C" int ggl_envset_destroy(ggl_envset* e) {
    if (!e) return GGL_OK;
    cudaSetDevice(e->device);
    cudaDeviceSynchronize();
    free_env(e);
    return GGL_OK;
}

#define ENV_CHECK(e) if (!(e)) { ggl::set_error("null envset"); return GGL_ERR_INVALID; }

// new arena -> block assignment from the work estimates of the env step that was just launched (same stream, no host sync)
static int rebalance(ggl_envset* e, cudaStream_t s) {
    if (!e->balance) return GGL_OK;
    balance_kernel<<<1, 1024, 0, s>>>(e->d_stepCost, e->d_arenaOfPos, e->cfg.numArenas, e->geom.threads / kGroup, e->geom.grid, e->d_balanceOrder);
    GGL_LAUNCH_CHECK(); ggl::count_launch();
    return GGL_OK;
}

extern "C" int ggl_envset_reset(ggl_envset* e, void* stream) {
    ENV_CHECK(e);
    env_reset_kernel<<<grid_for(e->cfg), kBlock, 0, (cudaStream_t)stream>>>(e->cfg, e->P, 0);
    GGL_LAUNCH_CHECK(); ggl::count_launch();
    return GGL_OK;
}
extern "C" int ggl_envset_reset_arena(ggl_envset* e, int arena, void* stream) {
    ENV_CHECK(e);
    if (arena < -1 || arena >= e->cfg.numArenas) { ggl::set_error("reset_arena: arena %d out of range", arena); return GGL_ERR_INVALID; }
    env_reset_kernel<<<grid_for(e->cfg), kBlock, 0, (cudaStream_t)stream>>>(e->cfg, e->P, arena < 0 ? 1 : 2 + arena);
    GGL_LAUNCH_CHECK(); ggl::count_launch();
    return GGL_OK;
}
extern "C" int ggl_envset_step_first_half(ggl_envset* e, void* stream) {
    ENV_CHECK(e);
    env_step_kernel<<<e->geom.grid, e->geom.threads, e->geom.smem, (cudaStream_t)stream>>>(e->cfg, e->P, nullptr, 1, 0, slot_bytes(e->cfg));
    GGL_LAUNCH_CHECK(); ggl::count_launch();
    return GGL_OK;
}
extern "C" int ggl_envset_step_second_half(ggl_envset* e, const int32_t* d_actions, void* stream) {
    ENV_CHECK(e);
    if (!d_actions) { ggl::set_error("null actions"); return GGL_ERR_INVALID; }
    env_step_kernel<<<e->geom.grid, e->geom.threads, e->geom.smem, (cudaStream_t)stream>>>(e->cfg, e->P, d_actions, 0, 1, slot_bytes(e->cfg));
    GGL_LAUNCH_CHECK(); ggl::count_launch();
    return rebalance(e, (cudaStream_t)stream);
}
extern "C" int ggl_envset_step(ggl_envset* e, const int32_t* d_actions, void* stream) {
    ENV_CHECK(e);
    if (!d_actions) { ggl::set_error("null actions"); return GGL_ERR_INVALID; }
    env_step_kernel<<<e->geom.grid, e->geom.threads, e->geom.smem, (cudaStream_t)stream>>>(e->cfg, e->P, d_actions, 1, 1, slot_bytes(e->cfg));
    GGL_LAUNCH_CHECK(); ggl::count_launch();
    return rebalance(e, (cudaStream_t)stream);
}
extern "C" int ggl_arena_step(ggl_envset* e, int ticks, void* stream) {
    ENV_CHECK(e);
    arena_step_kernel<<<e->geom.grid, e->geom.threads, e->geom.smem, (cudaStream_t)stream>>>(e->cfg, e->P, ticks, slot_bytes(e->cfg));
    GGL_LAUNCH_CHECK(); ggl::count_launch();
    return GGL_OK;
}
extern "C" int ggl_envset_random_actions(ggl_envset* e, int32_t* d_actions, uint64_t seed, void* stream) {
    ENV_CHECK(e);
    random_actions_kernel<<<(e->numPlayers + 255) / 256, 256, 0, (cudaStream_t)stream>>>(e->P.masks, d_actions, e->numPlayers, seed);
    GGL_LAUNCH_CHECK(); ggl::count_launch();
    return GGL_OK;
}
extern "C" int ggl_envset_sync(ggl_envset* e) {
    ENV_CHECK(e);
    GGL_CUDA(cudaSetDevice(e->device));
    GGL_CUDA(cudaDeviceSynchronize());
    return GGL_OK;
}
extern "C" int ggl_envset_views_get(ggl_envset* e, ggl_envset_views* v) {
    ENV_CHECK(e);
    if (!v) return GGL_ERR_INVALID;
    v->num_arenas = e->cfg.numArenas; v->num_players = e->numPlayers; v->obs_size = e->cfg.obsSize; v->num_actions = kNumActions;
    v->players_per_arena = e->cfg.numCars;
    v->d_obs = e->P.obs; v->d_action_masks = e->P.masks; v->d_rewards = e->P.rewards; v->d_terminals = e->P.terminals;
    v->d_slot_car_id = e->P.slotCarId; v->d_slot_team = e->P.slotTeam; v->d_events = e->P.events; v->d_last_rewards = e->P.lastRewards;
    return GGL_OK;
}

// ---- state injection / extraction (host <-> device, one arena at a time; test and StateSetter path) ----
extern "C" int ggl_envset_set_state(ggl_envset* e, int arena, const ggl_arena_state* st) {
    ENV_CHECK(e);
    if (!st || arena < 0 || arena >= e->cfg.numArenas || st->num_cars != e->cfg.numCars) { ggl::set_error("set_state: bad argument"); return GGL_ERR_INVALID; }
    GGL_CUDA(cudaSetDevice(e->device));
    GGL_CUDA(cudaDeviceSynchronize());
    ArenaDev* A = new ArenaDev();
    cudaError_t err = cudaMemcpy(A, e->P.arenas + arena, sizeof(ArenaDev), cudaMemcpyDeviceToHost);
    if (err == cudaSuccess) {
        arena_phys_from_host(*st, *A);
        err = cudaMemcpy(e->P.arenas + arena, A, sizeof(ArenaDev), cudaMemcpyHostToDevice);
    }
    delete A;
    if (err != cudaSuccess) { ggl::set_error("set_state -> %s", cudaGetErrorString(err)); return GGL_ERR_CUDA; }
    return GGL_OK;
}

extern "C" int ggl_envset_get_state(ggl_envset* e, int arena, ggl_arena_state* st) {
    ENV_CHECK(e);
    if (!st || arena < 0 || arena >= e->cfg.numArenas) { ggl::set_error("get_state: bad argument"); return GGL_ERR_INVALID; }
    GGL_CUDA(cudaSetDevice(e->device));
    GGL_CUDA(cudaDeviceSynchronize());
    ArenaDev* A = new ArenaDev();
    cudaError_t err = cudaMemcpy(A, e->P.arenas + arena, sizeof(ArenaDev), cudaMemcpyDeviceToHost);
    if (err == cudaSuccess) arena_phys_to_host(*A, *st);
    delete A;
    if (err != cudaSuccess) { ggl::set_error("get_state -> %s", cudaGetErrorString(err)); return GGL_ERR_CUDA; }
    return GGL_OK;
}

// ordered contact rows of the arena's last tick (test / debug): out[max][20] floats per row = bodyA, bodyB, normal xyz,
// relA xyz, relB xyz, distance, friction, restitution, special; returns the row count or a negative status
extern "C" int ggl_envset_get_contacts(ggl_envset* e, int arena, float* out, int max) {
    ENV_CHECK(e);
    if (!out || arena < 0 || arena >= e->cfg.numArenas) { ggl::set_error("get_contacts: bad argument"); return GGL_ERR_INVALID; }
    GGL_CUDA(cudaSetDevice(e->device));
    GGL_CUDA(cudaDeviceSynchronize());
    ArenaDev* A = new ArenaDev();
    cudaError_t err = cudaMemcpy(A, e->P.arenas + arena, sizeof(ArenaDev), cudaMemcpyDeviceToHost);
    int n = 0;
    if (err == cudaSuccess)
        for (int k = 0; k < A->numContacts && n < max; k++) {
            const Contact& c = A->contacts[A->contactOrder[k]];
            float* o = out + 20 * n++;
            memset(o, 0, 80);
            o[0] = c.bodyA; o[1] = c.bodyB; o[2] = c.normal.x; o[3] = c.normal.y; o[4] = c.normal.z;
            o[5] = c.relA.x; o[6] = c.relA.y; o[7] = c.relA.z; o[8] = c.relB.x; o[9] = c.relB.y; o[10] = c.relB.z;
            o[11] = c.dist; o[12] = c.friction; o[13] = c.restitution; o[14] = c.special;
        }
    delete A;
    if (err != cudaSuccess) { ggl::set_error("get_contacts -> %s", cudaGetErrorString(err)); return GGL_ERR_CUDA; }
    return n;
}

extern "C" int ggl_envset_set_reset_state(ggl_envset* e, int arena, const ggl_arena_state* st) {
    ENV_CHECK(e);
    if (!st || arena < 0 || arena >= e->cfg.numArenas || st->num_cars != e->cfg.numCars) { ggl::set_error("set_reset_state: bad argument"); return GGL_ERR_INVALID; }
    GGL_CUDA(cudaSetDevice(e->device));
    GGL_CUDA(cudaDeviceSynchronize());
    if (!e->d_reset) {
        int rc = dmalloc(&e->d_reset, (size_t)e->cfg.numArenas);
        if (rc) return rc;
        GGL_CUDA(cudaMemset(e->d_reset, 0, (size_t)e->cfg.numArenas * sizeof(ArenaDev)));
        e->P.resetArenas = e->d_reset;
    }
    ArenaDev* A = new ArenaDev();
    memset(A, 0, sizeof(ArenaDev));
    arena_phys_from_host(*st, *A);
    A->hasResetState = 1;
    cudaError_t err = cudaMemcpy(e->d_reset + arena, A, sizeof(ArenaDev), cudaMemcpyHostToDevice);
    delete A;
    if (err != cudaSuccess) { ggl::set_error("set_reset_state -> %s", cudaGetErrorString(err)); return GGL_ERR_CUDA; }
    return GGL_OK;
}

extern "C" int ggl_envset_set_rng(ggl_envset* e, int arena, uint32_t state) {
    ENV_CHECK(e);
    if (arena < -1 || arena >= e->cfg.numArenas || state < 1u || state > 2147483646u) { ggl::set_error("set_rng: bad argument"); return GGL_ERR_INVALID; }
    GGL_CUDA(cudaSetDevice(e->device));
    GGL_CUDA(cudaDeviceSynchronize());
    for (int a = (arena < 0 ? 0 : arena); a < (arena < 0 ? e->cfg.numArenas : arena + 1); a++)
        GGL_CUDA(cudaMemcpy((char*)(e->P.arenas + a) + offsetof(ArenaDev, rng), &state, sizeof(uint32_t), cudaMemcpyHostToDevice));
    return GGL_OK;
}

extern "C" int ggl_envset_get_rng(ggl_envset* e, int arena, uint32_t* out_state) {
    ENV_CHECK(e);
    if (!out_state || arena < 0 || arena >= e->cfg.numArenas) { ggl::set_error("get_rng: bad argument"); return GGL_ERR_INVALID; }
    GGL_CUDA(cudaSetDevice(e->device));
    GGL_CUDA(cudaDeviceSynchronize());
    GGL_CUDA(cudaMemcpy(out_state, (const char*)(e->P.arenas + arena) + offsetof(ArenaDev, rng), sizeof(uint32_t), cudaMemcpyDeviceToHost));
    return GGL_OK;
}

extern "C" int ggl_envset_set_slot_order(ggl_envset* e, const int32_t* h_car_ids) {
    ENV_CHECK(e);
    if (!h_car_ids) return GGL_ERR_INVALID;
    GGL_CUDA(cudaSetDevice(e->device));
    GGL_CUDA(cudaDeviceSynchronize());
    const int P = e->cfg.numCars;
    std::vector<ArenaDev> all((size_t)e->cfg.numArenas);
    GGL_CUDA(cudaMemcpy(all.data(), e->P.arenas, all.size() * sizeof(ArenaDev), cudaMemcpyDeviceToHost));
    for (int a = 0; a < e->cfg.numArenas; a++) {
        bool seen[kMaxCars] = {};
        for (int s = 0; s < P; s++) {
            int id = h_car_ids[a * P + s];
            if (id < 1 || id > P || seen[id - 1]) { ggl::set_error("set_slot_order: arena %d is not a permutation of 1..%d", a, P); return GGL_ERR_INVALID; }
            seen[id - 1] = true;
            all[a].slotCar[s] = (uint8_t)(id - 1);
        }
    }
    GGL_CUDA(cudaMemcpy(e->P.arenas, all.data(), all.size() * sizeof(ArenaDev), cudaMemcpyHostToDevice));
    // obs / mask rows follow the slot order: rebuild them without touching the arena state
    return GGL_OK;
}

// ---- host-buffer wrapper: the reference-facing call (EnvState vectors live on the host, EnvSet.h:35-60) ----
extern "C" int ggl_envset_step_host(ggl_envset* e, const int32_t* h_actions, float* h_obs, uint8_t* h_masks, float* h_rewards,
                                    uint8_t* h_terminals) {
    ENV_CHECK(e);
    if (!h_actions) { ggl::set_error("null actions"); return GGL_ERR_INVALID; }
    GGL_CUDA(cudaSetDevice(e->device));
    cudaStream_t s = e->ownStream;
    const size_t np = (size_t)e->numPlayers;
    GGL_CUDA(cudaMemcpyAsync(e->d_actionsScratch, h_actions, np * sizeof(int32_t), cudaMemcpyHostToDevice, s));
    int rc = ggl_envset_reset(e, s);
    if (rc) return rc;
    rc = ggl_envset_step(e, e->d_actionsScratch, s);
    if (rc) return rc;
    if (h_obs) GGL_CUDA(cudaMemcpyAsync(h_obs, e->P.obs, np * e->cfg.obsSize * sizeof(float), cudaMemcpyDeviceToHost, s));
    if (h_masks) GGL_CUDA(cudaMemcpyAsync(h_masks, e->P.masks, np * kNumActions, cudaMemcpyDeviceToHost, s));
    if (h_rewards) GGL_CUDA(cudaMemcpyAsync(h_rewards, e->P.rewards, np * sizeof(float), cudaMemcpyDeviceToHost, s));
    if (h_terminals) GGL_CUDA(cudaMemcpyAsync(h_terminals, e->P.terminals, (size_t)e->cfg.numArenas, cudaMemcpyDeviceToHost, s));
    GGL_CUDA(cudaStreamSynchronize(s));
    return GGL_OK;
}

// Pipelined host-buffer step (the Learner loop): the obs/mask copy and the caller's policy overlap the first half.
extern "C" int ggl_envset_begin_step_host(ggl_envset* e, float* h_obs, uint8_t* h_masks) {
    ENV_CHECK(e);
    GGL_CUDA(cudaSetDevice(e->device));
    const size_t np = (size_t)e->numPlayers;
    int rc = ggl_envset_reset(e, e->ownStream);
    if (rc) return rc;
    GGL_CUDA(cudaEventRecord(e->evReset, e->ownStream));
    rc = ggl_envset_step_first_half(e, e->ownStream);                 // runs while the copies below are in flight
    if (rc) return rc;
    GGL_CUDA(cudaStreamWaitEvent(e->copyStream, e->evReset, 0));
    if (h_obs) GGL_CUDA(cudaMemcpyAsync(h_obs, e->P.obs, np * e->cfg.obsSize * sizeof(float), cudaMemcpyDeviceToHost, e->copyStream));
    if (h_masks) GGL_CUDA(cudaMemcpyAsync(h_masks, e->P.masks, np * kNumActions, cudaMemcpyDeviceToHost, e->copyStream));
    GGL_CUDA(cudaEventRecord(e->evCopied, e->copyStream));
    GGL_CUDA(cudaEventSynchronize(e->evCopied));
    return GGL_OK;
}
extern "C" int ggl_envset_finish_step_host(ggl_envset* e, const int32_t* h_actions, float* h_rewards, uint8_t* h_terminals, float* h_final_obs) {
    ENV_CHECK(e);
    if (!h_actions) { ggl::set_error("null actions"); return GGL_ERR_INVALID; }
    GGL_CUDA(cudaSetDevice(e->device));
    cudaStream_t s = e->ownStream;
    const size_t np = (size_t)e->numPlayers;
    GGL_CUDA(cudaMemcpyAsync(e->d_actionsScratch, h_actions, np * sizeof(int32_t), cudaMemcpyHostToDevice, s));
    int rc = ggl_envset_step_second_half(e, e->d_actionsScratch, s);
    if (rc) return rc;
    if (h_rewards) GGL_CUDA(cudaMemcpyAsync(h_rewards, e->P.rewards, np * sizeof(float), cudaMemcpyDeviceToHost, s));
    if (h_terminals) GGL_CUDA(cudaMemcpyAsync(h_terminals, e->P.terminals, (size_t)e->cfg.numArenas, cudaMemcpyDeviceToHost, s));
    if (h_final_obs) GGL_CUDA(cudaMemcpyAsync(h_final_obs, e->P.obs, np * e->cfg.obsSize * sizeof(float), cudaMemcpyDeviceToHost, s));
    GGL_CUDA(cudaStreamSynchronize(s));
    return GGL_OK;
}

// StepSecondHalf with host action indices + host read-back of the EnvState buffers (the C++ host mirror's path). Work is
// issued on the legacy default stream like the stream-less calls of the mirror (StepFirstHalf(nullptr)), so it orders after them.
extern "C" int ggl_envset_step_second_half_host(ggl_envset* e, const int32_t* h_actions) {
    ENV_CHECK(e);
    if (!h_actions) { ggl::set_error("null actions"); return GGL_ERR_INVALID; }
    GGL_CUDA(cudaSetDevice(e->device));
    GGL_CUDA(cudaMemcpyAsync(e->d_actionsScratch, h_actions, (size_t)e->numPlayers * sizeof(int32_t), cudaMemcpyHostToDevice, 0));
    int rc = ggl_envset_step_second_half(e, e->d_actionsScratch, nullptr);
    if (rc) return rc;
    GGL_CUDA(cudaStreamSynchronize(0));
    return GGL_OK;
}
extern "C" int ggl_envset_read_host(ggl_envset* e, float* h_obs, uint8_t* h_masks, float* h_rewards, uint8_t* h_terminals) {
    ENV_CHECK(e);
    GGL_CUDA(cudaSetDevice(e->device));
    GGL_CUDA(cudaDeviceSynchronize());
    const size_t np = (size_t)e->numPlayers;
    if (h_obs) GGL_CUDA(cudaMemcpy(h_obs, e->P.obs, np * e->cfg.obsSize * sizeof(float), cudaMemcpyDeviceToHost));
    if (h_masks) GGL_CUDA(cudaMemcpy(h_masks, e->P.masks, np * kNumActions, cudaMemcpyDeviceToHost));
    if (h_rewards) GGL_CUDA(cudaMemcpy(h_rewards, e->P.rewards, np * sizeof(float), cudaMemcpyDeviceToHost));
    if (h_terminals) GGL_CUDA(cudaMemcpy(h_terminals, e->P.terminals, (size_t)e->cfg.numArenas, cudaMemcpyDeviceToHost));
    return GGL_OK;
}

// profiling build only (-DGGL_PHASE_TIMING): per-phase SM cycles summed over arena leaders; zeros in the product library
extern "C" int ggl_debug_phase_cycles(uint64_t* out16, int reset) {
    if (!out16) return GGL_ERR_INVALID;
    memset(out16, 0, 16 * sizeof(uint64_t));
#if defined(GGL_PHASE_TIMING)
    GGL_CUDA(cudaDeviceSynchronize());
    GGL_CUDA(cudaMemcpyFromSymbol(out16, g_phase_cycles, 16 * sizeof(uint64_t)));
    if (reset == 2) {      // sub-phase table instead (collide_body: car-ball, car-mesh, car-planes, car-car, ball-mesh)
        GGL_CUDA(cudaMemcpyFromSymbol(out16, g_sub_cycles, 16 * sizeof(uint64_t)));
        uint64_t z[16] = {}; GGL_CUDA(cudaMemcpyToSymbol(g_sub_cycles, z, sizeof(z)));
        return 1;
    }
    if (reset) { uint64_t z[16] = {}; GGL_CUDA(cudaMemcpyToSymbol(g_phase_cycles, z, sizeof(z))); }
    return 1;
#else
    (void)reset;
    return 0;
#endif
}

// profiling build only: per-arena phase cycles of the first `n` (<= 8192) arenas, out[n][16]
extern "C" int ggl_debug_arena_phase_cycles(uint64_t* out, int n, int reset) {
#if defined(GGL_PHASE_TIMING)
    if (!out || n < 1 || n > 8192) return GGL_ERR_INVALID;
    GGL_CUDA(cudaDeviceSynchronize());
    GGL_CUDA(cudaMemcpyFromSymbol(out, g_arena_phase_cycles, (size_t)n * 16 * sizeof(uint64_t)));
    if (reset) {
        void* p = nullptr;
        GGL_CUDA(cudaGetSymbolAddress(&p, g_arena_phase_cycles));
        GGL_CUDA(cudaMemset(p, 0, sizeof(uint64_t) * 8192 * 16));
    }
    return 1;
#else
    (void)out; (void)n; (void)reset;
    return 0;
#endif
}

// sizes of the C structs (the ctypes mirror in abi.py is checked against these)
extern "C" size_t ggl_abi_sizeof(int which) {
    switch (which) {
    case 0: return sizeof(ggl_envset_desc);
    case 1: return sizeof(ggl_arena_state);
    case 2: return sizeof(ggl_car_state);
    case 3: return sizeof(ggl_ball_state);
    case 4: return sizeof(ggl_pad_state);
    case 5: return sizeof(ggl_envset_views);
    case 6: return sizeof(ggl_reward_desc);
    case 7: return sizeof(ggl_terminal_desc);
    case 8: return sizeof(ggl_controls);
    case 9: return sizeof(ggl_wheel_state);
    case 100: return sizeof(ArenaDev);
    default: return 0;
    }
}

// Workload census of the last raw tick (debug; call after ggl_arena_step(e, 1, ...), which writes the per-tick scratch back):
// out[8][40] histograms over arenas: 0 solver rows, 1 rows of the largest island, 2 (car, triangle) candidates, 3 islands
// with rows, 4 wheels in contact, 5 cars with world / car-car rows, 6 mesh hits, 7 largest per-car row window
namespace {
__global__ void tick_stats_kernel(const ArenaDev* arenas, int numArenas, unsigned* out) {
    const int a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= numArenas) return;
    const ArenaDev& A = arenas[a];
    auto bump = [&](int h, int v) { atomicAdd(out + h * 40 + (v < 0 ? 0 : v > 39 ? 39 : v), 1u); };
    bump(0, A.numContacts);
    int rowsOf[8] = {}, nIsl = 0, maxRows = 0;
    for (int k = 0; k < A.numContacts; k++) rowsOf[A.island[A.contacts[A.contactOrder[k]].bodyA]]++;
    for (int b = 0; b < 8; b++) { if (rowsOf[b]) nIsl++; if (rowsOf[b] > maxRows) maxRows = rowsOf[b]; }
    bump(1, maxRows); bump(3, nIsl);
    int cand = 0, wheels = 0, carsRest = 0, hits = 0, maxWin = 0;
    for (int c = 0; c < A.numCars; c++) {
        cand += A.meshCandN[c]; hits += (int)A.meshHitN[c];
        for (int w = 0; w < 4; w++) wheels += A.cars[c].wheelsWithContact[w] != 0;      // (the wheels' per-tick scratch holds solver rows by now)
        carsRest += A.cntRest[c] != 0;
        if (A.cntRest[c] > maxWin) maxWin = A.cntRest[c];
    }
    bump(2, cand); bump(4, wheels); bump(5, carsRest); bump(6, hits); bump(7, maxWin);
}
}  // namespace
// Contact points dropped since create because an arena's windows and overflow pool were full (ArenaDev::contactOverflow,
// sticky), summed over the arenas. 0 = every contact the narrow phase found reached the solver.
namespace {
__global__ void overflow_sum_kernel(const ArenaDev* arenas, int numArenas, unsigned long long* out) {
    const int a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a < numArenas && arenas[a].contactOverflow) atomicAdd(out, (unsigned long long)arenas[a].contactOverflow);
}
}  // namespace
extern "C" int ggl_envset_contact_overflow(ggl_envset* e, uint64_t* out) {
    ENV_CHECK(e);
    if (!out) { ggl::set_error("contact_overflow: null output"); return GGL_ERR_INVALID; }
    unsigned long long* d = nullptr;
    GGL_CUDA(cudaSetDevice(e->device));
    GGL_CUDA(cudaDeviceSynchronize());
    GGL_CUDA(cudaMalloc(&d, sizeof(*d)));
    GGL_CUDA(cudaMemset(d, 0, sizeof(*d)));
    overflow_sum_kernel<<<(e->cfg.numArenas + 255) / 256, 256>>>(e->P.arenas, e->cfg.numArenas, d);
    ggl::count_launch();
    unsigned long long h = 0;
    cudaError_t err = cudaMemcpy(&h, d, sizeof(h), cudaMemcpyDeviceToHost);
    cudaFree(d);
    if (err != cudaSuccess) { ggl::set_error("contact_overflow -> %s", cudaGetErrorString(err)); return GGL_ERR_CUDA; }
    *out = (uint64_t)h;
    return GGL_OK;
}

extern "C" int ggl_debug_tick_stats(ggl_envset* e, uint32_t* out320) {
    ENV_CHECK(e);
    if (!out320) return GGL_ERR_INVALID;
    GGL_CUDA(cudaSetDevice(e->device));
    unsigned* d = nullptr;
    GGL_CUDA(cudaMalloc(&d, 320 * sizeof(unsigned)));
    GGL_CUDA(cudaMemset(d, 0, 320 * sizeof(unsigned)));
    tick_stats_kernel<<<(e->cfg.numArenas + 127) / 128, 128>>>(e->P.arenas, e->cfg.numArenas, d);
    cudaError_t err = cudaMemcpy(out320, d, 320 * sizeof(unsigned), cudaMemcpyDeviceToHost);
    cudaFree(d);
    if (err != cudaSuccess) { ggl::set_error("tick_stats -> %s", cudaGetErrorString(err)); return GGL_ERR_CUDA; }
    return GGL_OK;
}
