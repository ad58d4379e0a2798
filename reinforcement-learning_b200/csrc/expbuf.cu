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
    long long blocks = (count + 7) / 8;                                          // 8 warps per block, a sample per warp per pass
    if (blocks > (long long)sms * 16) blocks = (long long)sms * 16;
    auto kern = obs_size <= 64 ? expbuf_gather_kernel<2> : obs_size <= 128 ? expbuf_gather_kernel<4>
              : obs_size <= 192 ? expbuf_gather_kernel<6> : expbuf_gather_kernel<EXP_KMAX>;
    kern<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(p, d_indices, count, rowLimit, obs_size, num_actions, max_action_index);
    GGL_LAUNCH_CHECK();
    ggl::count_launch();
    return GGL_OK;
}
