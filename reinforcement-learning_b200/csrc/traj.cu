// Trajectory assembly on the device — the order of the flat experience stream (sm_100a).
//
// Replaces the host-side bookkeeping inside GGL::Learner::Start
// (GigaLearnCPP/src/public/GigaLearnCPP/Learner.cpp:504-547 `Trajectory`, :701-712 obs/mask append, :823-863 result append,
// terminal decision and `combinedTraj.Append`). The reference keeps one std::vector-backed buffer per player, copies every obs
// row into it on the host, and appends a finished buffer to the combined vectors (a second host copy of every row).
// Here the per-player buffers and the combined stream are device tensors:
//   per player   PAGES of TRAJ_ROWS steps out of one pool (pStates f32[pool rows][obs], pMasks u8[pool rows][A], pActions i32 /
//                pLogProbs f32 / pRewards f32 / pTerminals i8 [pool rows]); pageOf[N][pages per episode] says which page holds
//                steps 16 k .. 16 k + 15 of a player's running episode. A page is taken off the free stack when a player's
//                length reaches a multiple of TRAJ_ROWS and goes back when the finished buffer has been moved, so the pool
//                holds what is in flight (players x mean episode length), not players x maxEpisodeLength
//   combined     the same six tensors with capacity_timesteps rows + nextStates f32[capacity_truncated][obs]
// and a step costs four small launches: the row append (HBM-bound, one row in and out per player), the scalar append, an
// ordered scan over the players that finished this step (segment offsets in player order, exactly the order of the
// reference's loop at :844-862), and the segment copy (HBM-bound, every row moves once more, like the reference's second copy).
#include <cstdlib>
#include "common.cuh"

namespace {

constexpr int TRAJ_ROWS = 16;           // rows of a finished segment one warp moves per work item

struct TrajDev {
    int N, ppa, obs, A, cap, maxEpLen;
    int pagesPerPlayer, numPages;                       // page table width; pages in the pool
    int* pageOf;                                        // [N][pagesPerPlayer] page of the player's k-th run of TRAJ_ROWS steps
    int* freePages; int* freeTop;                       // stack of free pages; allocation and release never share a launch
    long long capL, capK;
    float* pStates; uint8_t* pMasks; int32_t* pActions; float* pLogProbs; float* pRewards; int8_t* pTerminals; int* pLen;
    int* segLen;                                        // per player, this step: length of the finished buffer (0 = still running)
    int* finList; long long* finOff; long long* finNext;  // per finished player in player order: id, first row in the stream, nextStates row or -1
    float* cStates; uint8_t* cMasks; int32_t* cActions; float* cLogProbs; float* cRewards; int8_t* cTerminals; float* cNext;
    long long* counts;                                  // [0] combined length, [1] nextStates rows, [2] overflow flag, [3] finished this step
};

// warp copy of n 4-byte (or 2-byte) words, eight loads in flight per lane before the first store (source and destination may
// alias as far as the compiler knows, so a plain loop would pay one HBM round trip per word)
template <class W>
__device__ __forceinline__ void warp_copy(W* __restrict__ d, const W* __restrict__ s, int n, int lane) {
    for (int base = 0; base < n; base += 256) {
        W v[8];
#pragma unroll
        for (int j = 0; j < 8; j++) { const int i = base + j * 32 + lane; v[j] = i < n ? __ldcs(s + i) : W(0); }
#pragma unroll
        for (int j = 0; j < 8; j++) { const int i = base + j * 32 + lane; if (i < n) __stcs(d + i, v[j]); }
    }
}
__device__ __forceinline__ void warp_copy_bytes(uint8_t* d, const uint8_t* s, int n, int lane) {
    if ((n & 1) == 0 && ((reinterpret_cast<uintptr_t>(d) | reinterpret_cast<uintptr_t>(s)) & 1) == 0)
        warp_copy(reinterpret_cast<unsigned short*>(d), reinterpret_cast<const unsigned short*>(s), n >> 1, lane);
    else
        warp_copy(d, s, n, lane);
}

// :701-712 — one warp per player appends the obs row and the mask row at that player's current length
__global__ void __launch_bounds__(256) traj_push_obs_kernel(TrajDev T, const float* __restrict__ obs, const uint8_t* __restrict__ masks) {
    const int lane = threadIdx.x & 31;
    const int p = (int)(((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    if (p >= T.N) return;
    const int len = T.pLen[p];
    int page;
    if (len % TRAJ_ROWS == 0) {                                     // first step of a page: take one off the free stack
        page = 0;
        if (lane == 0) {
            const int top = atomicSub(T.freeTop, 1) - 1;
            if (top >= 0) page = T.freePages[top];
            else { page = -1; T.counts[2] = 2; }                    // pool exhausted: sticky, the host call fails (ggl_traj_counts)
            T.pageOf[(long long)p * T.pagesPerPlayer + len / TRAJ_ROWS] = page;
        }
        page = __shfl_sync(0xffffffffu, page, 0);
    } else {
        page = T.pageOf[(long long)p * T.pagesPerPlayer + len / TRAJ_ROWS];
    }
    if (page < 0) return;
    const long long row = (long long)page * TRAJ_ROWS + len % TRAJ_ROWS;
    warp_copy(T.pStates + row * T.obs, obs + (long long)p * T.obs, T.obs, lane);
    warp_copy_bytes(T.pMasks + row * T.A, masks + (long long)p * T.A, T.A, lane);
}

// :823-850 — per player: scalars, terminal type, forced truncation; returns the length of the finished buffer (0 = still running)
__device__ __forceinline__ int traj_push_scalars(const TrajDev& T, int p, const int32_t* __restrict__ actions, const float* __restrict__ logProbs,
                                                 const float* __restrict__ rewards, const uint8_t* __restrict__ terminals, int8_t& termOut) {
    const int len = T.pLen[p];
    const int page = T.pageOf[(long long)p * T.pagesPerPlayer + len / TRAJ_ROWS];
    const long long row = (long long)(page < 0 ? 0 : page) * TRAJ_ROWS + len % TRAJ_ROWS;
    if (page >= 0) {
        T.pActions[row] = actions[p];
        T.pRewards[row] = rewards[p];
        T.pLogProbs[row] = logProbs[p];
    }
    const int newLen = len + 1;                                     // traj.Length() after the push_backs of :826-828
    int8_t term = (int8_t)terminals[p / T.ppa];                     // :831-841
    if (!term && newLen >= T.maxEpLen) term = 2;                    // :848-850 forced TRUNCATED
    if (page >= 0) T.pTerminals[row] = term;
    T.segLen[p] = term == 2 ? -newLen : (term ? newLen : 0);        // TRUNCATED rides in the sign
    T.pLen[p] = term ? 0 : newLen;                                  // traj.Clear() at :861
    termOut = term;
    return term ? newLen : 0;
}

__global__ void __launch_bounds__(256) traj_push_scalars_kernel(TrajDev T, const int32_t* __restrict__ actions, const float* __restrict__ logProbs,
                                                                 const float* __restrict__ rewards, const uint8_t* __restrict__ terminals) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    int8_t term;
    if (p < T.N) traj_push_scalars(T, p, actions, logProbs, rewards, terminals, term);
}

// :844-862 — where each finished buffer lands: exclusive sums over the players in index order. One block; N is at most ~1e5.
__global__ void __launch_bounds__(1024) traj_scan_kernel(TrajDev T) {
    __shared__ int wLen[32], wFin[32], wTr[32];
    __shared__ long long carryLen;
    __shared__ int carryFin, carryTr;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) { carryLen = 0; carryFin = 0; carryTr = 0; }
    __syncthreads();
    const long long baseLen = T.counts[0], baseK = T.counts[1];
    for (int c0 = 0; c0 < T.N; c0 += 1024) {
        const int p = c0 + tid;
        const int sl = p < T.N ? T.segLen[p] : 0;
        const int len = abs(sl), fin = sl != 0, tr = sl < 0;
        int sLen = len, sFin = fin, sTr = tr;                       // inclusive warp scans
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int a = __shfl_up_sync(0xffffffffu, sLen, o), b = __shfl_up_sync(0xffffffffu, sFin, o), c = __shfl_up_sync(0xffffffffu, sTr, o);
            if (lane >= o) { sLen += a; sFin += b; sTr += c; }
        }
        if (lane == 31) { wLen[warp] = sLen; wFin[warp] = sFin; wTr[warp] = sTr; }
        __syncthreads();
        if (warp == 0) {
            int a = wLen[lane], b = wFin[lane], c = wTr[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int x = __shfl_up_sync(0xffffffffu, a, o), y = __shfl_up_sync(0xffffffffu, b, o), z = __shfl_up_sync(0xffffffffu, c, o);
                if (lane >= o) { a += x; b += y; c += z; }
            }
            wLen[lane] = a; wFin[lane] = b; wTr[lane] = c;          // inclusive over warps
        }
        __syncthreads();
        const long long exLen = carryLen + (warp ? wLen[warp - 1] : 0) + sLen - len;
        const int exFin = carryFin + (warp ? wFin[warp - 1] : 0) + sFin - fin;
        const int exTr = carryTr + (warp ? wTr[warp - 1] : 0) + sTr - tr;
        if (fin) {
            T.finList[exFin] = p;
            T.finOff[exFin] = baseLen + exLen;
            T.finNext[exFin] = tr ? baseK + exTr : -1;
        }
        __syncthreads();
        if (tid == 1023) { carryLen += wLen[31]; carryFin += wFin[31]; carryTr += wTr[31]; }
        __syncthreads();
    }
    if (tid == 0) {
        if (baseLen + carryLen > T.capL || baseK + carryTr > T.capK) {
            if (T.counts[2] == 0) T.counts[2] = 1;                  // out of capacity: nothing is copied, the host call fails
            T.counts[3] = 0;
        } else {
            T.counts[0] = baseLen + carryLen;
            T.counts[1] = baseK + carryTr;
            T.counts[3] = carryFin;
        }
    }
}

// combinedTraj.Append(traj) (:523-531, :860) + nextStates (:855-858): a warp moves TRAJ_ROWS rows of one finished buffer
__global__ void __launch_bounds__(256) traj_move_kernel(TrajDev T, const float* __restrict__ obsFinal) {
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    const int chunks = T.pagesPerPlayer;
    const long long items = T.counts[3] * chunks;
    for (long long item = warp; item < items; item += nwarps) {
        const int k = (int)(item / chunks), c = (int)(item - (long long)k * chunks);
        const int p = T.finList[k];
        const int len = abs(T.segLen[p]);
        const int r0 = c * TRAJ_ROWS;
        if (r0 >= len) continue;
        const int rows = min(TRAJ_ROWS, len - r0);
        const int page = T.pageOf[(long long)p * T.pagesPerPlayer + c];
        if (page < 0) continue;                                     // pool was exhausted while this buffer grew (flagged)
        const long long src = (long long)page * TRAJ_ROWS, dst = T.finOff[k] + r0;
        warp_copy(T.cStates + dst * T.obs, T.pStates + src * T.obs, rows * T.obs, lane);
        warp_copy_bytes(T.cMasks + dst * T.A, T.pMasks + src * T.A, rows * T.A, lane);
        if (lane < rows) {
            T.cActions[dst + lane] = T.pActions[src + lane];
            T.cLogProbs[dst + lane] = T.pLogProbs[src + lane];
            T.cRewards[dst + lane] = T.pRewards[src + lane];
            T.cTerminals[dst + lane] = T.pTerminals[src + lane];
        }
        if (c == 0 && T.finNext[k] >= 0) {
            const float* s = obsFinal + (long long)p * T.obs;
            float* d = T.cNext + T.finNext[k] * T.obs;
            warp_copy(d, s, T.obs, lane);
        }
        __syncwarp();
        if (lane == 0) T.freePages[atomicAdd(T.freeTop, 1)] = page;     // the page is free again (no allocation runs in this launch)
    }
}

__global__ void traj_init_pages_kernel(TrajDev T) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < T.numPages) T.freePages[i] = T.numPages - 1 - i;        // page 0 on top
    if (i == 0) *T.freeTop = T.numPages;
}

}  // namespace

struct ggl_traj {
    TrajDev T;
    int device, sms;
    long long* h_counts;            // pinned
    void* allocs[32];
    int nalloc;
};

namespace {

template <class X>
int talloc(ggl_traj* t, X** p, size_t n) {
    cudaError_t e = cudaMalloc((void**)p, n * sizeof(X));
    if (e != cudaSuccess) { ggl::set_error("ggl_traj_create: cudaMalloc(%zu bytes) -> %s", n * sizeof(X), cudaGetErrorString(e)); return GGL_ERR_NOMEM; }
    t->allocs[t->nalloc++] = *p;
    return GGL_OK;
}

void tfree(ggl_traj* t) {
    if (!t) return;
    for (int i = 0; i < t->nalloc; i++) cudaFree(t->allocs[i]);
    if (t->h_counts) cudaFreeHost(t->h_counts);
    delete t;
}

}  // namespace

#define TRAJ_TRY(x) do { int _rc = (x); if (_rc != GGL_OK) { tfree(t); return _rc; } } while (0)

extern "C" int ggl_traj_create(const ggl_traj_desc* d, ggl_traj** out) {
    if (!d || !out || d->num_players <= 0 || d->players_per_arena <= 0 || d->obs_size <= 0 || d->num_actions <= 0 || d->capacity_timesteps <= 0 ||
        d->capacity_truncated <= 0 || d->num_players % d->players_per_arena != 0) {
        ggl::set_error("ggl_traj_create: bad descriptor"); return GGL_ERR_INVALID;
    }
    GGL_CUDA(cudaSetDevice(d->device));
    ggl_traj* t = new ggl_traj();
    t->nalloc = 0; t->h_counts = nullptr; t->device = d->device; t->sms = 148;
    cudaDeviceGetAttribute(&t->sms, cudaDevAttrMultiProcessorCount, d->device);
    TrajDev& T = t->T;
    T.N = d->num_players; T.ppa = d->players_per_arena; T.obs = d->obs_size; T.A = d->num_actions;
    T.maxEpLen = d->max_episode_length;
    T.cap = d->max_episode_length > 1 ? d->max_episode_length : 1;   // a buffer is flushed as soon as it holds maxEpisodeLength steps
    T.capL = d->capacity_timesteps; T.capK = d->capacity_truncated;
    // Per-player storage: pages of TRAJ_ROWS steps from one pool. capacity_player_rows = 0 sizes the pool for the worst case
    // (every player at maxEpisodeLength: C2's shape, 16 384 players x 1 800 steps x obs 167, is 22 GB; 98 304 players x 1 800 x
    // obs 225 would be 177 GB); a positive value sizes it for the steps that are in flight at any one time (players x the
    // episode length one expects, with margin) - running out of pages is reported by ggl_traj_counts, not silently dropped.
    T.pagesPerPlayer = (T.cap + TRAJ_ROWS - 1) / TRAJ_ROWS;
    const long long worstPages = (long long)T.N * T.pagesPerPlayer;
    long long pages = worstPages;
    if (d->capacity_player_rows > 0) {
        pages = (d->capacity_player_rows + TRAJ_ROWS - 1) / TRAJ_ROWS;
        if (pages < T.N) pages = T.N;                                // every player holds at least its current page
        if (pages > worstPages) pages = worstPages;
    }
    if (pages > 0x7FFFFFF0ll / TRAJ_ROWS) { ggl::set_error("ggl_traj_create: per-player pool of %lld pages is too large", pages); tfree(t); return GGL_ERR_INVALID; }
    T.numPages = (int)pages;
    const size_t rows = (size_t)pages * TRAJ_ROWS;
    {
        const size_t perRow = (size_t)T.obs * 4 + (size_t)T.A + 4 + 4 + 4 + 1;
        const size_t need = rows * perRow + (size_t)T.capL * perRow + (size_t)T.capK * T.obs * 4 + (size_t)worstPages * 4;
        size_t freeB = 0, totalB = 0;
        if (cudaMemGetInfo(&freeB, &totalB) == cudaSuccess && need > freeB) {
            ggl::set_error("ggl_traj_create: a per-player pool of %zu steps (%zu players, up to %d steps per episode) x %zu B per step + the "
                           "combined stream need %.1f GB, %.1f GB are free on device %d: set capacity_player_rows to the steps in flight "
                           "(players x expected episode length) or lower max_episode_length",
                           rows, (size_t)T.N, T.cap, perRow, need / 1e9, freeB / 1e9, d->device);
            tfree(t);
            return GGL_ERR_NOMEM;
        }
    }
    TRAJ_TRY(talloc(t, &T.pageOf, (size_t)worstPages)); TRAJ_TRY(talloc(t, &T.freePages, (size_t)pages)); TRAJ_TRY(talloc(t, &T.freeTop, (size_t)1));
    TRAJ_TRY(talloc(t, &T.pStates, rows * T.obs)); TRAJ_TRY(talloc(t, &T.pMasks, rows * T.A)); TRAJ_TRY(talloc(t, &T.pActions, rows));
    TRAJ_TRY(talloc(t, &T.pLogProbs, rows)); TRAJ_TRY(talloc(t, &T.pRewards, rows)); TRAJ_TRY(talloc(t, &T.pTerminals, rows));
    TRAJ_TRY(talloc(t, &T.pLen, (size_t)T.N)); TRAJ_TRY(talloc(t, &T.segLen, (size_t)T.N));
    TRAJ_TRY(talloc(t, &T.finList, (size_t)T.N)); TRAJ_TRY(talloc(t, &T.finOff, (size_t)T.N)); TRAJ_TRY(talloc(t, &T.finNext, (size_t)T.N));
    TRAJ_TRY(talloc(t, &T.cStates, (size_t)T.capL * T.obs)); TRAJ_TRY(talloc(t, &T.cMasks, (size_t)T.capL * T.A));
    TRAJ_TRY(talloc(t, &T.cActions, (size_t)T.capL)); TRAJ_TRY(talloc(t, &T.cLogProbs, (size_t)T.capL)); TRAJ_TRY(talloc(t, &T.cRewards, (size_t)T.capL));
    TRAJ_TRY(talloc(t, &T.cTerminals, (size_t)T.capL)); TRAJ_TRY(talloc(t, &T.cNext, (size_t)T.capK * T.obs));
    TRAJ_TRY(talloc(t, &T.counts, (size_t)4));
    traj_init_pages_kernel<<<(T.numPages + 255) / 256, 256>>>(T);
    if (cudaMemset(T.pageOf, 0xFF, sizeof(int) * (size_t)worstPages) != cudaSuccess || cudaMemset(T.pLen, 0, sizeof(int) * T.N) != cudaSuccess || cudaMemset(T.counts, 0, sizeof(long long) * 4) != cudaSuccess ||
        cudaMallocHost((void**)&t->h_counts, sizeof(long long) * 4) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) {
        ggl::set_error("ggl_traj_create: initialisation failed: %s", cudaGetErrorString(cudaGetLastError())); tfree(t); return GGL_ERR_CUDA;
    }
    *out = t;
    return GGL_OK;
}

extern "C" int ggl_traj_destroy(ggl_traj* t) {
    if (!t) return GGL_OK;
    cudaSetDevice(t->device);
    cudaDeviceSynchronize();
    tfree(t);
    return GGL_OK;
}

#define TRAJ_CHECK(t) if (!(t)) { ggl::set_error("null trajectory store"); return GGL_ERR_INVALID; }

extern "C" int ggl_traj_views_get(ggl_traj* t, ggl_traj_views* out) {
    TRAJ_CHECK(t);
    if (!out) { ggl::set_error("ggl_traj_views_get: null output"); return GGL_ERR_INVALID; }
    out->states = t->T.cStates; out->action_masks = t->T.cMasks; out->actions = t->T.cActions; out->log_probs = t->T.cLogProbs;
    out->rewards = t->T.cRewards; out->terminals = t->T.cTerminals; out->next_states = t->T.cNext;
    return GGL_OK;
}

extern "C" int ggl_traj_begin_iteration(ggl_traj* t, void* stream) {
    TRAJ_CHECK(t);
    GGL_CUDA(cudaMemsetAsync(t->T.counts, 0, sizeof(long long) * 2, (cudaStream_t)stream));      // length and nextStates rows; the overflow flag is sticky
    return GGL_OK;
}

extern "C" int ggl_traj_push_obs(ggl_traj* t, const float* d_obs, const uint8_t* d_masks, void* stream) {
    TRAJ_CHECK(t);
    if (!d_obs || !d_masks) { ggl::set_error("ggl_traj_push_obs: null input"); return GGL_ERR_INVALID; }
    const int blocks = (t->T.N + 7) / 8;
    traj_push_obs_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(t->T, d_obs, d_masks);
    GGL_LAUNCH_CHECK(); ggl::count_launch();
    return GGL_OK;
}

extern "C" int ggl_traj_push_result(ggl_traj* t, const int32_t* d_actions, const float* d_log_probs, const float* d_rewards,
                                    const uint8_t* d_terminals, const float* d_obs_final, void* stream) {
    TRAJ_CHECK(t);
    if (!d_actions || !d_log_probs || !d_rewards || !d_terminals || !d_obs_final) { ggl::set_error("ggl_traj_push_result: null input"); return GGL_ERR_INVALID; }
    cudaStream_t s = (cudaStream_t)stream;
    traj_push_scalars_kernel<<<(t->T.N + 255) / 256, 256, 0, s>>>(t->T, d_actions, d_log_probs, d_rewards, d_terminals);
    GGL_LAUNCH_CHECK();
    traj_scan_kernel<<<1, 1024, 0, s>>>(t->T);
    GGL_LAUNCH_CHECK();
    traj_move_kernel<<<t->sms * 4, 256, 0, s>>>(t->T, d_obs_final);
    GGL_LAUNCH_CHECK(); ggl::count_launch(3);
    return GGL_OK;
}

extern "C" int ggl_traj_counts(ggl_traj* t, void* stream, int64_t* out_length, int64_t* out_truncated) {
    TRAJ_CHECK(t);
    GGL_CUDA(cudaMemcpyAsync(t->h_counts, t->T.counts, sizeof(long long) * 4, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    GGL_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    if (t->h_counts[2]) {
        if (t->h_counts[2] == 2)
            ggl::set_error("trajectory store: the per-player pool ran out of pages (%d pages of %d steps for %d players): raise "
                           "capacity_player_rows", t->T.numPages, TRAJ_ROWS, t->T.N);
        else
            ggl::set_error("trajectory store out of capacity (%lld timesteps / %lld truncated rows)", t->T.capL, t->T.capK);
        return GGL_ERR_INVALID;
    }
    if (out_length) *out_length = t->h_counts[0];
    if (out_truncated) *out_truncated = t->h_counts[1];
    return GGL_OK;
}

extern "C" int ggl_traj_player_lengths(ggl_traj* t, int32_t* h_out) {
    TRAJ_CHECK(t);
    if (!h_out) { ggl::set_error("ggl_traj_player_lengths: null output"); return GGL_ERR_INVALID; }
    GGL_CUDA(cudaMemcpy(h_out, t->T.pLen, sizeof(int) * t->T.N, cudaMemcpyDeviceToHost));
    return GGL_OK;
}
