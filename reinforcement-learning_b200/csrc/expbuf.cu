// ExperienceBuffer batching on the device — one fused permuted gather per epoch (sm_100a).
//
// Replaces GGL::ExperienceBuffer::_GetSamples / GetAllBatchesShuffled
// (GigaLearnCPP/src/private/GigaLearnCPP/PPO/ExperienceBuffer.cpp:15-98, 100-163), which per epoch shuffles an index vector
// on the host and then, for EVERY minibatch, runs six index_select gathers on CPU tensors followed by an H2D copy
// (PPOLearner.cpp:319-342). The minibatches of an epoch are consecutive slices of ONE permutation of the stream, so here the
// whole permuted stream is gathered once, on the device, and every minibatch is a view (start, size) into the result:
//     out.states[i]  = states[perm[i]]      f32[obs]         out.actionMasks[i] = masks[perm[i]]   u8[numActions]
//     out.actions[i] = clamp(actions[perm[i]])   and the three f32 scalars per sample
// HBM-bound: 2 x (4*obs + numActions + 16) + 8 bytes per sample (read + write + the index), e.g. 1556 B at obs 167.
// One warp moves one sample: 32 lanes x 4 B over the row (rows are only 4-byte aligned: obs is odd), 2-byte lanes over the
// mask row, lanes 0..3 the scalars; all of a sample's loads are issued before its first store, and the next pass's row index is
// prefetched, so a warp pays one HBM round trip per sample (grid-stride loop over 148 x 16 blocks).
// The permutation is an input: std::shuffle over std::default_random_engine is libstdc++'s, the host keeps producing it.
#include <cstdlib>
#include "common.cuh"

namespace {

struct ExpPtrs {
    const float* states; const int32_t* actions; const float* logProbs; const float* targetValues; const uint8_t* masks; const float* advantages;
    float* oStates; int32_t* oActions; float* oLogProbs; float* oTargetValues; uint8_t* oMasks; float* oAdvantages;
};

#ifndef EXP_UNROLL_CFG
#define EXP_UNROLL_CFG 2
#endif
constexpr int EXP_UNROLL = EXP_UNROLL_CFG;      // samples a warp has in flight (strided by the number of warps)
constexpr int EXP_KMAX = 8;        // 32-float chunks of a state row held in registers (obs <= 256; wider rows loop)
#ifndef EXP_MMAX_CFG
#define EXP_MMAX_CFG 2
#endif
constexpr int EXP_MMAX = EXP_MMAX_CFG;   // mask units (halfwords for even rows, bytes for odd ones) per lane held in registers

// Everything one sample needs is LOADED before anything is stored: the source and destination pointers may alias as far as the
// compiler knows, so a store between two loads would put one HBM round trip per tensor on the warp's critical path
// (ncu r1i: 62 % long-scoreboard stalls, the six tensors' latencies back to back).
template <int K>
struct Sample {
    float st[K];
    unsigned short mk[EXP_MMAX];
    uint32_t scalar;               // lane 0: action, lane 1: log prob, lane 2: target value, lane 3: advantage (raw 32 bits)
};

template <int K>
__device__ __forceinline__ void load_sample(Sample<K>& S, const ExpPtrs& p, long long r, int lane, int obs, int numActions, bool regStates,
                                            bool regMasks) {
    if (p.states && regStates) {
        const float* __restrict__ s = p.states + r * obs;
#pragma unroll
        for (int k = 0; k < K; k++) S.st[k] = (k * 32 + lane < obs) ? __ldcs(s + k * 32 + lane) : 0.f;
    }
    if (p.masks && regMasks) {
        const uint8_t* s = p.masks + r * numActions;
        if ((numActions & 1) == 0) {                        // even rows are 2-byte aligned: move halfwords
            const unsigned short* s2 = reinterpret_cast<const unsigned short*>(s);
#pragma unroll
            for (int k = 0; k < EXP_MMAX; k++) S.mk[k] = (k * 32 + lane < (numActions >> 1)) ? __ldcs(s2 + k * 32 + lane) : 0;
        } else {
#pragma unroll
            for (int k = 0; k < EXP_MMAX; k++) S.mk[k] = (k * 32 + lane < numActions) ? s[k * 32 + lane] : 0;
        }
    }
    const void* sp = lane == 0 ? (const void*)p.actions : lane == 1 ? (const void*)p.logProbs : lane == 2 ? (const void*)p.targetValues
                   : lane == 3 ? (const void*)p.advantages : nullptr;
    S.scalar = sp ? __ldcs(reinterpret_cast<const uint32_t*>(sp) + r) : 0u;
}

template <int K>
__device__ __forceinline__ void store_sample(const Sample<K>& S, const ExpPtrs& p, long long r, long long i, int lane, int obs, int numActions,
                                             long long maxAction, bool regStates, bool regMasks) {
    if (p.states) {
        float* __restrict__ d = p.oStates + i * obs;
        if (regStates) {
#pragma unroll
            for (int k = 0; k < K; k++)
                if (k * 32 + lane < obs) __stcs(d + k * 32 + lane, S.st[k]);
        } else {
            const float* __restrict__ s = p.states + r * obs;
            for (int k = lane; k < obs; k += 32) __stcs(d + k, __ldcs(s + k));
        }
    }
    if (p.masks) {
        uint8_t* d = p.oMasks + i * numActions;
        if (regMasks) {
            if ((numActions & 1) == 0) {
                unsigned short* d2 = reinterpret_cast<unsigned short*>(d);
#pragma unroll
                for (int k = 0; k < EXP_MMAX; k++)
                    if (k * 32 + lane < (numActions >> 1)) __stcs(d2 + k * 32 + lane, S.mk[k]);
            } else {
#pragma unroll
                for (int k = 0; k < EXP_MMAX; k++)
                    if (k * 32 + lane < numActions) d[k * 32 + lane] = (uint8_t)S.mk[k];
            }
        } else {
            const uint8_t* s = p.masks + r * numActions;
            for (int k = lane; k < numActions; k += 32) d[k] = s[k];
        }
    }
    if (lane == 0 && p.actions) {
        long long a = (int32_t)S.scalar;
        if (maxAction >= 0) a = a < 0 ? 0 : (a > maxAction ? maxAction : a);     // ExperienceBuffer.cpp:88-90
        p.oActions[i] = (int32_t)a;
    }
    if (lane == 1 && p.logProbs) reinterpret_cast<uint32_t*>(p.oLogProbs)[i] = S.scalar;
    if (lane == 2 && p.targetValues) reinterpret_cast<uint32_t*>(p.oTargetValues)[i] = S.scalar;
    if (lane == 3 && p.advantages) reinterpret_cast<uint32_t*>(p.oAdvantages)[i] = S.scalar;
}

// K = 32-float chunks per state row kept in registers (2, 4, 6 or 8: the register count, and with it the number of resident
// warps, follows the row width)
template <int K>
__global__ void __launch_bounds__(256) expbuf_gather_kernel(ExpPtrs p, const int64_t* __restrict__ idx, long long count, long long rowLimit,
                                                            int obs, int numActions, long long maxAction) {
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    const bool regStates = obs <= 32 * K;
    const bool regMasks = ((numActions & 1) == 0 ? (numActions >> 1) : numActions) <= 32 * EXP_MMAX;
    // the row index of the NEXT pass is fetched while this pass's rows are in flight
    long long rn[EXP_UNROLL];
#pragma unroll
    for (int u = 0; u < EXP_UNROLL; u++) { const long long i = warp + u * nwarps; rn[u] = i < count ? __ldg(idx + i) : 0; }
    for (long long i0 = warp; i0 < count; i0 += EXP_UNROLL * nwarps) {
        long long r[EXP_UNROLL];
        Sample<K> S[EXP_UNROLL];
#pragma unroll
        for (int u = 0; u < EXP_UNROLL; u++) {
            r[u] = rn[u] >= rowLimit ? rowLimit - 1 : rn[u];            // ExperienceBuffer.cpp:42-69
            const long long in = i0 + (u + EXP_UNROLL) * nwarps;
            rn[u] = in < count ? __ldg(idx + in) : 0;
        }
#pragma unroll
        for (int u = 0; u < EXP_UNROLL; u++)
            if (i0 + u * nwarps < count) load_sample(S[u], p, r[u], lane, obs, numActions, regStates, regMasks);
#pragma unroll
        for (int u = 0; u < EXP_UNROLL; u++) {
            const long long i = i0 + u * nwarps;
            if (i < count) store_sample(S[u], p, r[u], i, lane, obs, numActions, maxAction, regStates, regMasks);
        }
    }
}


// ---- staged variant: aligned 16-byte copies in, packed tiles out -----------------------------------------------------------------
// The register kernel above is bound by the load/store unit, not by HBM: rows start on 4-byte boundaries, so every 128-byte warp
// request of it straddles two cache lines (stores-only 0.19 ms + loads-only 0.22 ms = the 0.42 ms it takes, gpurun_out/exp36).
// Here a block owns a TILE of 32 consecutive output samples. The 16-byte-aligned span that contains each source row is copied to
// shared memory with cp.async (16 bytes per lane, no registers held, a whole tile in flight per stage while the previous one is
// written), and the tile, contiguous in every destination and 128-byte aligned (32 rows), leaves as whole-line stores.
#ifndef EXP_TILE_CFG
#define EXP_TILE_CFG 32
#endif
#ifndef EXP_STAGES_CFG
#define EXP_STAGES_CFG 2
#endif
constexpr int EXP_TILE = EXP_TILE_CFG;        // 16 or 32 samples: the tile starts 16-byte aligned in every destination
constexpr int EXP_STAGES = EXP_STAGES_CFG;    // tiles in flight per block = EXP_STAGES - 1 while one is written out

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, int srcBytes) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(dst), "l"(src), "r"(srcBytes) : "memory");
}
__device__ __forceinline__ void cp_async4(uint32_t dst, const void* src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }

struct StagedGeom {
    int slotS, slotM;          // bytes per staged state / mask row (aligned span, multiple of 16)
    int stageBytes;            // one stage: EXP_TILE state slots, EXP_TILE mask slots, 4 x EXP_TILE scalars, 2 x (EXP_TILE + 4) row bases
};

__host__ __device__ inline StagedGeom staged_geom(int obs, int numActions) {
    StagedGeom g;
    g.slotS = (obs * 4 + 12 + 15) & ~15;
    g.slotM = (numActions + 15 + 15) & ~15;
    g.stageBytes = EXP_TILE * (g.slotS + g.slotM) + 4 * EXP_TILE * 4 + 2 * (EXP_TILE + 4) * 4;
    return g;
}

__global__ void __launch_bounds__(256) expbuf_gather_staged_kernel(ExpPtrs p, const int64_t* __restrict__ idx, long long count, long long rowLimit,
                                                                   int obs, int numActions, long long maxAction) {
    extern __shared__ __align__(128) unsigned char smem[];
    const StagedGeom G = staged_geom(obs, numActions);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long tiles = (count + EXP_TILE - 1) / EXP_TILE;
    constexpr int PER_WARP = EXP_TILE / 8;                       // samples a warp stages
    auto stage_ptr = [&](int st) { return smem + (size_t)st * G.stageBytes; };

    // lanes 0..PER_WARP-1 of each warp hold the row index of "their" sample of the tile after next
    auto fetch_row = [&](long long tile) -> long long {
        const long long i = tile * EXP_TILE + warp * PER_WARP + lane;
        if (lane < PER_WARP && tile < tiles && i < count) return __ldg(idx + i);      // clamped where it is used: no wait on the load here
        return -1;
    };
    auto issue = [&](int st, long long rowMine) {
        if (rowMine >= rowLimit) rowMine = rowLimit - 1;              // ExperienceBuffer.cpp:42-69
        unsigned char* base = stage_ptr(st);
        unsigned char* sS = base;
        unsigned char* sM = base + EXP_TILE * G.slotS;
        uint32_t* sScal = reinterpret_cast<uint32_t*>(sM + EXP_TILE * G.slotM);          // [4][EXP_TILE]
        int* sOffS = reinterpret_cast<int*>(sScal + 4 * EXP_TILE);
        int* sOffM = sOffS + EXP_TILE + 4;
#pragma unroll
        for (int q = 0; q < PER_WARP; q++) {
            const long long r = __shfl_sync(0xffffffffu, rowMine, q);
            const int s = warp * PER_WARP + q;
            if (r < 0) continue;                                      // (rowMine was clamped by the caller)
            if (p.states) {
                const unsigned char* g0 = reinterpret_cast<const unsigned char*>(p.states + r * obs);
                const int off = (int)(reinterpret_cast<uintptr_t>(g0) & 15);
                const unsigned char* a0 = g0 - off;
                const int span = off + obs * 4;                       // bytes from a0 to the end of the row
                const uint32_t d0 = smem_u32(sS + s * G.slotS);
                for (int v = lane; v * 16 < span; v += 32) cp_async16(d0 + v * 16, a0 + v * 16, min(16, span - v * 16));
                if (lane == 0) sOffS[s] = s * G.slotS + off;           // byte position of the row inside the state slots
            }
        }
        if (p.masks) {                                                 // 8 lanes per sample: spans of up to 128 bytes
            const int q = lane >> 3, v0 = lane & 7;
            const long long r = __shfl_sync(0xffffffffu, rowMine, q < PER_WARP ? q : 0);
            if (q < PER_WARP && r >= 0) {
                const int s = warp * PER_WARP + q;
                const unsigned char* g0 = p.masks + r * numActions;
                const int off = (int)(reinterpret_cast<uintptr_t>(g0) & 15);
                const unsigned char* a0 = g0 - off;
                const int span = off + numActions;
                const uint32_t d0 = smem_u32(sM + s * G.slotM);
                for (int v = v0; v * 16 < span; v += 8) cp_async16(d0 + v * 16, a0 + v * 16, min(16, span - v * 16));
                if (v0 == 0) sOffM[s] = s * G.slotM + off;
            }
        }
        {                                                               // 4 lanes per sample: the four scalars
            const int q = lane >> 2, which = lane & 3;
            const long long r = __shfl_sync(0xffffffffu, rowMine, q < PER_WARP ? q : 0);
            const void* sp = which == 0 ? (const void*)p.actions : which == 1 ? (const void*)p.logProbs : which == 2 ? (const void*)p.targetValues
                                                                                                        : (const void*)p.advantages;
            if (q < PER_WARP && r >= 0 && sp) cp_async4(smem_u32(sScal + which * EXP_TILE + warp * PER_WARP + q), reinterpret_cast<const uint32_t*>(sp) + r);
        }
        cp_async_commit();
    };

    // floor(n / d) = umulhi(n, ceil(2^32 / d)) for n * d < 2^32, which the shared-memory limit of this kernel guarantees
    const unsigned magicObs = 0xFFFFFFFFu / (unsigned)obs + 1u, magicAct = 0xFFFFFFFFu / (unsigned)numActions + 1u;
    const bool dstMask16 = (reinterpret_cast<uintptr_t>(p.oMasks) & 15) == 0;
    long long tile = blockIdx.x;
    if (tile >= tiles) return;
    const long long stride = gridDim.x;
    // prologue: EXP_STAGES - 1 tiles in flight (tiles past the end issue nothing but still commit a group)
#pragma unroll
    for (int j = 0; j < EXP_STAGES - 1; j++) issue(j, fetch_row(tile + j * stride));
    long long rowPref = fetch_row(tile + (EXP_STAGES - 1) * stride);
    int st = 0;
    for (; tile < tiles; tile += stride, st = (st + 1 == EXP_STAGES ? 0 : st + 1)) {
        const long long rowLater = fetch_row(tile + EXP_STAGES * stride);
        cp_async_wait<EXP_STAGES - 2>();                               // this tile's group has landed (for this thread)
        __syncthreads();                                               // ... for every thread; and everyone is done with the previous tile
        issue(st == 0 ? EXP_STAGES - 1 : st - 1, rowPref);             // refill the stage the previous iteration read
        rowPref = rowLater;
        // ---- write the tile out: every destination range of the tile is contiguous
        const unsigned char* base = stage_ptr(st);
        const unsigned char* sS = base;
        const unsigned char* sM = base + EXP_TILE * G.slotS;
        const uint32_t* sScal = reinterpret_cast<const uint32_t*>(sM + EXP_TILE * G.slotM);
        const int* sOffS = reinterpret_cast<const int*>(sScal + 4 * EXP_TILE);
        const int* sOffM = sOffS + EXP_TILE + 4;
        const long long i0 = tile * EXP_TILE;
        const int nvalid = (int)min((long long)EXP_TILE, count - i0);
        if (p.states) {
            float* __restrict__ d = p.oStates + i0 * obs;
            const int W = nvalid * obs;
            if (nvalid == EXP_TILE) {
                // whole tile: 16 bytes per thread per pass; a float4 crosses a row boundary at most once (obs >= 4 here)
                for (int q = tid; q * 4 < W; q += 256) {
                    const int e = q * 4, r0 = (int)__umulhi((unsigned)e, magicObs), col = e - r0 * obs;
                    const int b0 = sOffS[r0] + col * 4, b1 = sOffS[r0 + 1] - (obs - col) * 4;      // sOffS has slack entries past the tile
                    float4 v;
                    v.x = *reinterpret_cast<const float*>(sS + b0);
                    v.y = *reinterpret_cast<const float*>(sS + (col + 1 < obs ? b0 : b1) + 4);
                    v.z = *reinterpret_cast<const float*>(sS + (col + 2 < obs ? b0 : b1) + 8);
                    v.w = *reinterpret_cast<const float*>(sS + (col + 3 < obs ? b0 : b1) + 12);
                    __stcs(reinterpret_cast<float4*>(d) + q, v);
                }
            } else {
                int r0 = tid / obs, col = tid - r0 * obs;
                for (int w = tid; w < W; w += 256) {
                    __stcs(d + w, *reinterpret_cast<const float*>(sS + sOffS[r0] + col * 4));
                    col += 256;
                    while (col >= obs) { col -= obs; r0++; }
                }
            }
        }
        if (p.masks) {
            uint8_t* d = p.oMasks + i0 * numActions;
            const int B = nvalid * numActions;                       // bytes of the tile; the tile starts 32-byte aligned
            if (nvalid == EXP_TILE && dstMask16 && numActions >= 16) {
                // 32 rows: B is a multiple of 16; 16 output bytes cross a row boundary at most once
                const bool even = (numActions & 1) == 0;             // then every row starts on a 2-byte boundary
                for (int q = tid; q * 16 < B; q += 256) {
                    const int b = q * 16, r0 = (int)__umulhi((unsigned)b, magicAct), col = b - r0 * numActions;
                    const int kx = numActions - col;                 // first byte of the 16 that belongs to the next row
                    const int pos0 = sOffM[r0] + col, pos1 = sOffM[r0 + 1] - kx;
                    uint32_t wds[4] = {0, 0, 0, 0};
                    if (even) {
#pragma unroll
                        for (int k = 0; k < 16; k += 2)
                            wds[k >> 2] |= (uint32_t)*reinterpret_cast<const unsigned short*>(sM + (k < kx ? pos0 : pos1) + k) << (8 * (k & 3));
                    } else {
#pragma unroll
                        for (int k = 0; k < 16; k++) wds[k >> 2] |= (uint32_t)sM[(k < kx ? pos0 : pos1) + k] << (8 * (k & 3));
                    }
                    __stcs(reinterpret_cast<uint4*>(d) + q, make_uint4(wds[0], wds[1], wds[2], wds[3]));
                }
            } else {
                for (int w = tid; w * 4 < B; w += 256) {
                    uint32_t word = 0;
                    int b = w * 4, r0 = b / numActions, col = b - r0 * numActions;
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        if (b + k < B) word |= (uint32_t)sM[sOffM[r0] + col] << (8 * k);
                        if (++col == numActions) { col = 0; r0++; }
                    }
                    if (b + 4 <= B) __stcs(reinterpret_cast<uint32_t*>(d) + w, word);
                    else for (int k = 0; b + k < B; k++) d[b + k] = (uint8_t)(word >> (8 * k));
                }
            }
        }
        if (tid < 4 * EXP_TILE) {
            const int which = tid / EXP_TILE, j = tid % EXP_TILE;
            if (j < nvalid) {
                const uint32_t v = sScal[which * EXP_TILE + j];
                if (which == 0) {
                    if (p.actions) {
                        long long a = (int32_t)v;
                        if (maxAction >= 0) a = a < 0 ? 0 : (a > maxAction ? maxAction : a);     // ExperienceBuffer.cpp:88-90
                        p.oActions[i0 + j] = (int32_t)a;
                    }
                } else {
                    uint32_t* o = reinterpret_cast<uint32_t*>(which == 1 ? (void*)p.oLogProbs : which == 2 ? (void*)p.oTargetValues : (void*)p.oAdvantages);
                    const void* sp = which == 1 ? (const void*)p.logProbs : which == 2 ? (const void*)p.targetValues : (const void*)p.advantages;
                    if (sp) o[i0 + j] = v;
                }
            }
        }
    }
    cp_async_wait<0>();
}

}  // namespace

// GetAllBatchesShuffled's partition of the permuted stream (ExperienceBuffer.cpp:143-160). Host arithmetic, no device work.
extern "C" int64_t ggl_expbuf_plan_batches(int64_t exp_size, int64_t batch_size, int overbatching, int64_t* out_starts, int64_t* out_sizes,
                                           int64_t max_batches) {
    if (batch_size <= 0 || exp_size < 0) { ggl::set_error("ggl_expbuf_plan_batches: bad size"); return GGL_ERR_INVALID; }
    int64_t n = 0;
    for (int64_t start = 0; start < exp_size; start += batch_size) {
        int64_t end = start + batch_size;
        if (end + batch_size > exp_size && overbatching) end = exp_size;
        if (end > exp_size) break;
        if (end - start <= 0) break;
        if (n < max_batches) { if (out_starts) out_starts[n] = start; if (out_sizes) out_sizes[n] = end - start; }
        n++;
        if (end == exp_size) break;
    }
    return n;
}

extern "C" int ggl_expbuf_gather(const ggl_expbuf_tensors* src, int64_t rows_states, int64_t rows_actions, int32_t obs_size, int32_t num_actions,
                                 const int64_t* d_indices, int64_t count, int64_t max_action_index, const ggl_expbuf_tensors* dst,
                                 void* stream) {
    if (!src || !dst || count < 0 || obs_size <= 0 || num_actions <= 0 || (count > 0 && !d_indices)) { ggl::set_error("ggl_expbuf_gather: bad argument"); return GGL_ERR_INVALID; }
    int64_t rowLimit = rows_states;                                              // ExperienceBuffer.cpp:24-33
    if (src->actions && rows_actions < rowLimit) rowLimit = rows_actions;
    if (!src->states || rowLimit <= 0 || count == 0) return GGL_OK;               // :19-21, :31-33: nothing to gather
    if ((src->states && !dst->states) || (src->actions && !dst->actions) || (src->log_probs && !dst->log_probs) ||
        (src->target_values && !dst->target_values) || (src->action_masks && !dst->action_masks) || (src->advantages && !dst->advantages)) {
        ggl::set_error("ggl_expbuf_gather: destination missing for a defined source tensor"); return GGL_ERR_INVALID;
    }
    ExpPtrs p;
    p.states = src->states; p.actions = src->actions; p.logProbs = src->log_probs; p.targetValues = src->target_values;
    p.masks = src->action_masks; p.advantages = src->advantages;
    p.oStates = dst->states; p.oActions = dst->actions; p.oLogProbs = dst->log_probs; p.oTargetValues = dst->target_values;
    p.oMasks = dst->action_masks; p.oAdvantages = dst->advantages;
    int dev = 0, sms = 148;
    GGL_CUDA(cudaGetDevice(&dev));
    GGL_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    // staged kernel: needs 16-byte-aligned source bases (its aligned spans then never leave the tensors) and two stages in shared memory
    const StagedGeom G = staged_geom(obs_size, num_actions);
    const size_t stagedSmem = (size_t)EXP_STAGES * G.stageBytes;
    const bool aligned = ((uintptr_t)src->states & 15) == 0 && ((uintptr_t)src->action_masks & 15) == 0 && ((uintptr_t)dst->states & 15) == 0 &&
                         ((uintptr_t)dst->action_masks & 3) == 0;
    const bool forceRegs = getenv("GGL_EXPBUF_REGISTER_KERNEL") != nullptr;      // test hook: exercise the general kernel
    if (aligned && !forceRegs && stagedSmem <= 100 * 1024 && count >= EXP_TILE && obs_size >= 4) {
        GGL_CUDA(cudaFuncSetAttribute(expbuf_gather_staged_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
        const long long tiles = (count + EXP_TILE - 1) / EXP_TILE;
        int perSM = (int)((227 * 1024) / (stagedSmem + 1024));
        if (perSM > 8) perSM = 8;
        if (perSM < 1) perSM = 1;
        long long grid = (long long)sms * perSM;
        if (grid > tiles) grid = tiles;
        expbuf_gather_staged_kernel<<<(int)grid, 256, stagedSmem, (cudaStream_t)stream>>>(p, d_indices, count, rowLimit, obs_size, num_actions, max_action_index);
        GGL_LAUNCH_CHECK();
        ggl::count_launch();
        return GGL_OK;
    }
    long long blocks = (count + 7) / 8;                                          // 8 warps per block, a sample per warp per pass
    if (blocks > (long long)sms * 16) blocks = (long long)sms * 16;
    auto kern = obs_size <= 64 ? expbuf_gather_kernel<2> : obs_size <= 128 ? expbuf_gather_kernel<4>
              : obs_size <= 192 ? expbuf_gather_kernel<6> : expbuf_gather_kernel<EXP_KMAX>;
    kern<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(p, d_indices, count, rowLimit, obs_size, num_actions, max_action_index);
    GGL_LAUNCH_CHECK();
    ggl::count_launch();
    return GGL_OK;
}
