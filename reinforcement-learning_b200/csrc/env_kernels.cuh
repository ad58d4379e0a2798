// CUDA kernels of the rollout hot path (sm_100a).
//
// Mapping: one lane group of kGroup = 8 threads owns one arena, a warp owns 4 arenas, and a block is 1..kMaxWarps warps chosen
// at launch so that the grid is ONE wave with one block per SM (4096 arenas -> 147 blocks of 7 warps on 148 SMs). Each arena's
// blob ([header][cars x numCars], sim_types.cuh) is staged HBM -> shared memory by its own lane group, all tick_skip ticks + the
// env layer run out of shared memory, and the state part of the blob goes back to HBM once: one HBM round trip per env step.
// Inside a tick, lanes 0..P-1 run the per-car phases and the per-player outputs, the wheel / (car, triangle) / pad tasks are
// spread over all 8 lanes, contact islands are solved one per lane, and lane 0 runs the arena-serial bookkeeping. Groups never
// straddle a warp, so phase boundaries are __syncwarp()s on the group's lanes; on top of that the warps of a block meet with
// __syncthreads() at four points of every tick (ExecGroup::phase_sync). The kernel is ~750 KB of SASS and in steady-state
// play ncu showed 58 % of all stall samples waiting for instruction fetch when every warp wandered through it on its own:
// walking the code together lets the warps of an SM share the instruction cache (1.98 -> 1.28 ms per 4096-arena step).
#pragma once
#include <cuda_runtime.h>
#include "env_reset.cuh"

namespace ggl {

#ifndef GGL_GROUP
#define GGL_GROUP 8
#endif
constexpr int kGroup = GGL_GROUP;                // lanes per arena (8 or 16)
constexpr int kArenasPerWarp = 32 / kGroup;
constexpr int kMaxWarps = (7 * kGroup + 7) / 8;  // warps per block of the step kernels: 7 units of 4 arenas (see launch_geom)
constexpr int kBlock = 32;                       // the reset kernel: one warp per block
constexpr int kArenasPerBlock = kBlock / kGroup;

// GGL_PSYNC_GROUPS > 1 (experiment): the warps of a block meet in that many independent sets (warp w in set w % G, named
// barriers 1..G) instead of all together: fewer arenas wait for the slowest one, more code regions live at once.
#ifndef GGL_PSYNC_GROUPS
#define GGL_PSYNC_GROUPS 1
#endif
struct ExecGroup {
    int lane_, mask_;
#if GGL_PSYNC_GROUPS > 1
    int barId_, barThreads_;
#endif
#if defined(GGL_PHASE_TIMING)
    int arena_;
#endif
    __device__ __forceinline__ int lane() const { return lane_; }
    __device__ __forceinline__ int lanes() const { return kGroup; }
    __device__ __forceinline__ void sync() const { __syncwarp(mask_); }
    // The warps of a block meet at a few phase boundaries of the tick, so that the warps of an SM walk through the same code
    // together and share the instruction cache. Lane groups of arenas past the end have left the kernel before the first one.
#if GGL_PSYNC_GROUPS > 1
    __device__ __forceinline__ void phase_sync() const { asm volatile("bar.sync %0, %1;" ::"r"(barId_), "r"(barThreads_) : "memory"); }
#else
    __device__ __forceinline__ void phase_sync() const { __syncthreads(); }
#endif
};

struct EnvDevPtrs {
    ArenaDev* arenas;
    const ArenaDev* resetArenas;     // injected reset states (may be null)
    float* obs; uint8_t* masks; float* rewards; uint8_t* terminals;
    int32_t* slotCarId; uint8_t* slotTeam; uint8_t* events; float* lastRewards;
    const float* actionTable;        // [90*8]
    WorldDev world;
    // load balancing between the blocks of the step kernel: position (block, lane group) -> arena, rebuilt after every env step
    // from the arenas' work estimates of that step (balance_kernel); null = identity
    const int32_t* arenaOfPos;
    uint32_t* stepCost;              // [numArenas] ArenaDev::stepCost of the last completed env step
};

__device__ __forceinline__ int arena_of(const EnvDevPtrs& P, int pos) { return P.arenaOfPos ? __ldg(P.arenaOfPos + pos) : pos; }

// liveWarps: warps of this block that hold at least one arena (the others have left the kernel and never reach a barrier)
__device__ __forceinline__ ExecGroup make_exec(int liveWarps = 0) {
    ExecGroup ex;
    ex.lane_ = threadIdx.x % kGroup;
    int base = (threadIdx.x % 32) / kGroup * kGroup;
    ex.mask_ = (int)(((1u << kGroup) - 1u) << base);
#if GGL_PSYNC_GROUPS > 1
    {
        const int w = threadIdx.x / 32, g = w % GGL_PSYNC_GROUPS;
        ex.barId_ = 1 + g;
        ex.barThreads_ = 32 * ((liveWarps - g + GGL_PSYNC_GROUPS - 1) / GGL_PSYNC_GROUPS);
    }
#endif
#if defined(GGL_PHASE_TIMING)
    ex.arena_ = blockIdx.x * (blockDim.x / kGroup) + threadIdx.x / kGroup;
#endif
    return ex;
}

// obs row + mask row (+ slot identity) of one player
__device__ __forceinline__ void write_player_rows(const ArenaDev& A, const EnvCfg& cfg, const EnvDevPtrs& P, int arena, int slot) {
    const int row = arena * cfg.numCars + slot;
    uint32_t permMates = 0, permOpps = 0;
    if (cfg.obsKind == GGL_OBS_DEFAULT_PADDED) padded_obs_draws_upto(A.rng, cfg.obsMaxPlayers, slot, permMates, permOpps);
    build_obs_row(A, cfg, slot, P.obs + (size_t)row * cfg.obsSize, permMates, permOpps);
    action_mask_row(A.cars[A.slotCar[slot]], cfg.actionBits, P.masks + (size_t)row * kNumActions);
    P.slotCarId[row] = (int)A.cars[A.slotCar[slot]].id;
    P.slotTeam[row] = A.cars[A.slotCar[slot]].team;
}

// after every row of the arena is out: the stream moves past the draws of all of them (DefaultObsPadded only)
__device__ __forceinline__ void rows_advance_rng(ArenaDev& A, const EnvCfg& cfg) {
    if (cfg.obsKind != GGL_OBS_DEFAULT_PADDED) return;
    uint32_t pm, po;
    A.rng = padded_obs_draws_upto(A.rng, cfg.obsMaxPlayers, A.numCars - 1, pm, po);
}

// ---- blob staging: the whole warp moves each of its arenas with 16-byte accesses (512 contiguous bytes per instruction)
extern __shared__ __align__(16) unsigned char g_arena_smem[];

// Every lane group moves its own arena (8 lanes x 16 B = 128 contiguous bytes per instruction). `full` = 0 skips the per-tick
// scratch in the middle of the blob (contact rows, car-mesh staging: ~3.7 KB of 7.7 KB for 2v2), which is dead between env
// steps; the raw-tick debug path (ggl_arena_step + ggl_envset_get_contacts) moves everything.
__device__ __forceinline__ void stage_in(const EnvDevPtrs& P, int arena, const ExecGroup& ex, int slotBytes, int full) {
    const int n16 = slotBytes >> 4;
    const int s0 = full ? n16 : (int)(arena_persist_bytes() >> 4), s1 = (int)(arena_hdr_bytes() >> 4);
    const uint4* src = reinterpret_cast<const uint4*>(P.arenas + arena);
    uint4* dst = reinterpret_cast<uint4*>(g_arena_smem + (size_t)(threadIdx.x / kGroup) * slotBytes);
    for (int i = ex.lane(); i < s0; i += kGroup) dst[i] = __ldg(src + i);
    for (int i = s1 + ex.lane(); i < n16; i += kGroup) dst[i] = __ldg(src + i);
    ex.sync();
}
__device__ __forceinline__ void stage_out(const EnvDevPtrs& P, int arena, const ExecGroup& ex, int slotBytes, int full) {
    ex.sync();
    const int n16 = slotBytes >> 4;
    const int s0 = full ? n16 : (int)(arena_persist_bytes() >> 4), s1 = (int)(arena_hdr_bytes() >> 4);
    uint4* dst = reinterpret_cast<uint4*>(P.arenas + arena);
    const uint4* src = reinterpret_cast<const uint4*>(g_arena_smem + (size_t)(threadIdx.x / kGroup) * slotBytes);
    for (int i = ex.lane(); i < s0; i += kGroup) dst[i] = src[i];
    for (int i = s1 + ex.lane(); i < n16; i += kGroup) dst[i] = src[i];
}
__device__ __forceinline__ ArenaDev& my_arena(int slotBytes) {
    return *reinterpret_cast<ArenaDev*>(g_arena_smem + (size_t)(threadIdx.x / kGroup) * slotBytes);
}

// EnvSet::ResetArena for arenas whose terminal flag is set; `force` = 1 resets every arena (EnvSet ctor), `force` = 2 + i
// resets arena i whatever its terminal flag (EnvSet::ResetArena(i), EnvSet.cpp:275-331, called directly)
__global__ void __launch_bounds__(kBlock) env_reset_kernel(EnvCfg cfg, EnvDevPtrs P, int force) {
    const int arena = blockIdx.x * kArenasPerBlock + threadIdx.x / kGroup;
    if (arena >= cfg.numArenas) return;
    const ExecGroup ex = make_exec();
    ArenaDev& A = P.arenas[arena];
    const bool doReset = force == 1 || (force >= 2 && arena == force - 2) || (force == 0 && A.terminal != 0);
    if (!doReset) return;
    if (ex.lane() == 0) {
        arena_state_set(A, cfg, P.resetArenas ? &P.resetArenas[arena] : nullptr);
        env_after_state_set(A);
        P.terminals[arena] = 0;
    }
    ex.sync();
    if (ex.lane() < A.numCars) {
        write_player_rows(A, cfg, P, arena, ex.lane());
        for (int e = 0; e < 9; e++) P.events[(size_t)(arena * cfg.numCars + ex.lane()) * 9 + e] = 0;
    }
    if (cfg.obsKind == GGL_OBS_DEFAULT_PADDED) {
        ex.sync();
        if (ex.lane() == 0) rows_advance_rng(A, cfg);
    }
}

// StepFirstHalf and/or StepSecondHalf of every arena in one launch. With doReset, arenas whose terminal flag is set are
// reset first (EnvSet::Reset, the Learner loop order of Learner.cpp:669-793) inside the same launch.
__global__ void __launch_bounds__(32 * kMaxWarps, 1) env_step_kernel(EnvCfg cfg, EnvDevPtrs P, const int32_t* __restrict__ actions, int doFirst,
                                                          int doSecond, int slotBytes) {
    GGL_PH_BEGIN();
    // The world descriptor is handed by reference to out-of-line device functions; taking the address of a kernel parameter
    // makes the compiler keep a copy on every thread's stack (ncu: local loads on every W.field access, e.g. the empty-cell box
    // test of every ray). One copy per block in shared memory instead.
    __shared__ WorldDev s_world;
    if (threadIdx.x == 0) s_world = P.world;
    __syncthreads();
    const int pos = blockIdx.x * (blockDim.x / kGroup) + threadIdx.x / kGroup;
    if (pos >= cfg.numArenas) return;            // whole groups leave: nobody waits for them at a group or block barrier
    const int arena = arena_of(P, pos);
    const int groupsHere = min((int)(blockDim.x / kGroup), cfg.numArenas - (int)(blockIdx.x * (blockDim.x / kGroup)));
    const ExecGroup ex = make_exec((groupsHere + kArenasPerWarp - 1) / kArenasPerWarp);
    stage_in(P, arena, ex, slotBytes, 0);
    {
        ArenaDev& A = my_arena(slotBytes);
        const WorldDev& W = s_world;
        GGL_PH(13);
        // ONE tick loop (a single inlined copy of arena_tick keeps the kernel's code half the size, which matters at ~7 warps
        // per SM that all sit in different places of it): the parsed actions take effect at tick `actionDelay`.
        const int tBegin = doFirst ? 0 : cfg.actionDelay, tEnd = doSecond ? cfg.tickSkip : cfg.actionDelay;
        if (doFirst) {
            if (ex.lane() == 0) env_begin_step(A);
            ex.sync();
        }
        bool applied = !doSecond;
#pragma unroll 1
        for (int t = tBegin; t < tEnd; t++) {
            if (!applied && t == cfg.actionDelay) {
                if (ex.lane() == 0) env_apply_actions(A, actions + arena * cfg.numCars, P.actionTable);
                ex.sync();
                applied = true;
            }
            arena_tick(A, W, ex);
        }
        if (doSecond) {
            if (!applied) {      // actionDelay == tickSkip: nothing left to tick after the actions are parsed
                if (ex.lane() == 0) env_apply_actions(A, actions + arena * cfg.numCars, P.actionTable);
                ex.sync();
            }
            GGL_PH_RESET();
            if (ex.lane() == 0) {
                env_after_ticks(A, cfg);
                P.terminals[arena] = A.terminal;
            }
            ex.sync();
            const int slot = ex.lane();
            if (slot < A.numCars) {
                const int row = arena * cfg.numCars + slot;
                // saved raw rewards: the player with the lowest car id (EnvSet.cpp:235-243, shuffleRewardSampling = false)
                int lowSlot = 0;
                for (int s = 1; s < A.numCars; s++)
                    if (A.cars[A.slotCar[s]].id < A.cars[A.slotCar[lowSlot]].id) lowSlot = s;
                float raw[kMaxRewards];
                P.rewards[row] = env_player_reward(A, cfg, slot, raw);
                if (slot == lowSlot)
                    for (int r = 0; r < cfg.numRewards; r++) P.lastRewards[(size_t)arena * cfg.numRewards + r] = raw[r];
                write_player_rows(A, cfg, P, arena, slot);
                for (int e = 0; e < 9; e++) P.events[(size_t)row * 9 + e] = A.players[slot].events[e];
            }
            ex.sync();
            if (ex.lane() == 0) {
                env_end_step(A);
                rows_advance_rng(A, cfg);
                if (P.stepCost) P.stepCost[arena] = A.stepCost;
            }
            GGL_PH(12);
        }
    }
    stage_out(P, arena, ex, slotBytes, 0);
}

// Which arena sits where in the next launch. The step kernel's time is that of its slowest block, and a block pays, phase by
// phase, for the slowest of its 28 arenas (block barriers); with arenas in index order the blocks of a 4096-arena 2v2 rollout
// spread from 0.78 M to 2.26 M cycles per step around a mean of 1.49 M (profiles/r2g_phase_profile.txt), and heavy situations
// (a pile-up, a car on the wall, the ball rolling along the floor) last many steps. So the arenas are dealt to the blocks by
// the work estimate of the step that just ended: sorted by cost (counting sort on a log scale), then
// dealt round by round across the blocks in snake order, every round landing in another warp of the block, so that each block
// and each warp gets the same mix. One block; runs on the step's stream right after the step kernel.
constexpr int kBalanceBuckets = 64;
__global__ void __launch_bounds__(1024) balance_kernel(const uint32_t* __restrict__ cost, int32_t* __restrict__ arenaOfPos, int numArenas,
                                                        int groupsPerBlock, int numBlocks, int32_t* __restrict__ order) {
    __shared__ int s_cnt[kBalanceBuckets], s_start[kBalanceBuckets];
    const int tid = threadIdx.x;
    if (tid < kBalanceBuckets) s_cnt[tid] = 0;
    __syncthreads();
    auto bucket = [](uint32_t c) { return c == 0 ? kBalanceBuckets - 1 : max(0, kBalanceBuckets - 2 - (int)(4.f * __log2f((float)c))); };   // heavy first
    for (int a = tid; a < numArenas; a += blockDim.x) atomicAdd(&s_cnt[bucket(cost[a])], 1);
    __syncthreads();
    if (tid == 0) { int run = 0; for (int b = 0; b < kBalanceBuckets; b++) { s_start[b] = run; run += s_cnt[b]; } }
    __syncthreads();
    // arenas of one bucket in whatever order the atomics hand out (which block an arena runs in does not change its results:
    // tests/test_env_gpu.py::test_block_geometry_does_not_change_results)
    for (int a = tid; a < numArenas; a += blockDim.x) order[atomicAdd(&s_start[bucket(cost[a])], 1)] = a;
    __syncthreads();
    // deal: rounds over the lane groups of a block; round r sits in lane group (r % W) * G + r / W (W warps per block, G groups
    // per warp), visits the blocks in snake order, and skips the groups of the last block that lie past numArenas
    const int G = 4, W = groupsPerBlock / G;          // units of 4 arenas (a warp at 8 lanes per arena)
    const int rem = numArenas - (numBlocks - 1) * groupsPerBlock;         // valid groups of the last block
    for (int i = tid; i < numArenas; i += blockDim.x) {
        int r = 0, cum = 0, group = 0, nb = numBlocks;
        for (; r < groupsPerBlock; r++) {
            group = W > 0 ? (r % W) * G + r / W : r;
            nb = numBlocks - (group >= rem ? 1 : 0);
            if (i < cum + nb) break;
            cum += nb;
        }
        const int j = i - cum, k = (r & 1) ? nb - 1 - j : j;
        arenaOfPos[k * groupsPerBlock + group] = order[i];
    }
}

// raw Arena::Step(ticks)
__global__ void __launch_bounds__(32 * kMaxWarps, 1) arena_step_kernel(EnvCfg cfg, EnvDevPtrs P, int ticks, int slotBytes) {
    __shared__ WorldDev s_world;
    if (threadIdx.x == 0) s_world = P.world;
    __syncthreads();
    const int arena = blockIdx.x * (blockDim.x / kGroup) + threadIdx.x / kGroup;
    if (arena >= cfg.numArenas) return;
    const int groupsHere = min((int)(blockDim.x / kGroup), cfg.numArenas - (int)(blockIdx.x * (blockDim.x / kGroup)));
    const ExecGroup ex = make_exec((groupsHere + kArenasPerWarp - 1) / kArenasPerWarp);
    stage_in(P, arena, ex, slotBytes, 1);
    ArenaDev& A = my_arena(slotBytes);
    for (int t = 0; t < ticks; t++) arena_tick(A, s_world, ex);
    stage_out(P, arena, ex, slotBytes, 1);
}

// uniformly random mask-valid action per player (bench / smoke helper): k-th set bit of the mask row, k from a hash
__global__ void random_actions_kernel(const uint8_t* __restrict__ masks, int32_t* __restrict__ actions, int numPlayers, uint64_t seed) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= numPlayers) return;
    uint64_t z = seed + 0x9E3779B97F4A7C15ull * (uint64_t)(i + 1);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; z ^= z >> 31;
    const uint8_t* m = masks + (size_t)i * kNumActions;
    int cnt = 0;
    for (int a = 0; a < kNumActions; a++) cnt += m[a] != 0;
    int sel = 0;
    if (cnt > 0) {
        int k = (int)(z % (uint64_t)cnt);
        for (int a = 0; a < kNumActions; a++)
            if (m[a]) { if (k-- == 0) { sel = a; break; } }
    }
    actions[i] = sel;
}

}  // namespace ggl
