// Host mirror (g++ only, no CUDA headers) of the learner-side pieces of the hot path, over the C ABI of include/gigaenv.h:
//   GGL::ExperienceBuffer   GigaLearnCPP/src/private/GigaLearnCPP/PPO/ExperienceBuffer.h / .cpp:8-13, 100-163
//   GGL::TrajectoryStore    the `Trajectory` buffers of GigaLearnCPP/src/public/GigaLearnCPP/Learner.cpp:504-547, 701-712, 823-863
//   GGL::GAE::Compute       GigaLearnCPP/src/private/GigaLearnCPP/PPO/GAE.cpp:7-208 (device tensors in, device tensors out)
// Same names, argument meaning and error behaviour (exceptions where the reference calls RG_ERR_CLOSE); tensors are raw DEVICE
// pointers with explicit sizes where the reference has torch::Tensor (libtorch wraps them with from_blob).
#pragma once
#include <algorithm>
#include <cstdint>
#include <numeric>
#include <random>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../../include/gigaenv.h"

namespace GGL {

inline void Check(int rc, const char* what) {
    if (rc < 0) throw std::runtime_error(std::string(what) + ": " + ggl_last_error());      // RG_ERR_CLOSE (Framework.h:16-21)
}

// ExperienceBuffer.h:8-15 — six tensors; `size` rows (`actionsSize` rows of actions, which may be shorter: .cpp:24-33)
struct ExperienceTensors {
    float* states = nullptr; int32_t* actions = nullptr; float* logProbs = nullptr; float* targetValues = nullptr;
    uint8_t* actionMasks = nullptr; float* advantages = nullptr;
    int64_t size = 0, actionsSize = 0;
};

class ExperienceBuffer {
public:
    int seed, device;
    ExperienceTensors data;                 // set by the learner (Learner.cpp:975-980); device pointers, not owned
    int obsSize = 0, numActions = 0;
    int64_t maxActionIndex = -1;            // ExperienceBuffer.cpp:88-90
    std::default_random_engine rng;         // .cpp:8-9: seeded once, stateful across epochs
    std::vector<int64_t> shuffledIndices;

    ExperienceBuffer(int seed, int device) : seed(seed), device(device), rng(seed) { shuffledIndices.reserve(200000); }
    ExperienceBuffer(const ExperienceBuffer&) = delete;
    ~ExperienceBuffer() { Release(); }

    // .cpp:100-163. The minibatches are views into ONE gathered copy of the permuted stream (valid until the next call).
    std::vector<ExperienceTensors> GetAllBatchesShuffled(int64_t batchSize, bool overbatching, void* stream = nullptr) {
        if (!data.states || data.size <= 0) return {};
        int64_t expSize = data.size;
        if (data.actions) expSize = std::min(expSize, data.actionsSize);
        if (expSize <= 0) return {};
        if (shuffledIndices.size() != static_cast<size_t>(expSize)) {           // :121-128: refilled only when the size changes
            shuffledIndices.resize(expSize);
            std::iota(shuffledIndices.begin(), shuffledIndices.end(), 0);
        }
        std::shuffle(shuffledIndices.begin(), shuffledIndices.end(), rng);       // :130
        Reserve(expSize);
        Check(ggl_dev_upload(dPerm, shuffledIndices.data(), sizeof(int64_t) * expSize, stream), "ExperienceBuffer: permutation upload");
        ggl_expbuf_tensors src = {data.states, data.actions, data.logProbs, data.targetValues, data.actionMasks, data.advantages};
        ggl_expbuf_tensors dst = {own.states, data.actions ? own.actions : nullptr, data.logProbs ? own.logProbs : nullptr,
                                  data.targetValues ? own.targetValues : nullptr, data.actionMasks ? own.actionMasks : nullptr,
                                  data.advantages ? own.advantages : nullptr};
        Check(ggl_expbuf_gather(&src, data.size, data.actions ? data.actionsSize : data.size, obsSize, numActions, dPerm, expSize, maxActionIndex,
                                &dst, stream), "ExperienceBuffer: gather");
        const int64_t nb = ggl_expbuf_plan_batches(expSize, batchSize, overbatching ? 1 : 0, nullptr, nullptr, 0);
        Check((int)std::min<int64_t>(nb, 0), "ExperienceBuffer: batch plan");
        std::vector<int64_t> starts(nb), sizes(nb);
        ggl_expbuf_plan_batches(expSize, batchSize, overbatching ? 1 : 0, starts.data(), sizes.data(), nb);
        std::vector<ExperienceTensors> result;
        result.reserve(nb);
        for (int64_t b = 0; b < nb; b++) {
            ExperienceTensors t;
            const int64_t s = starts[b];
            t.states = dst.states + s * obsSize;
            t.actions = dst.actions ? dst.actions + s : nullptr;
            t.logProbs = dst.log_probs ? dst.log_probs + s : nullptr;
            t.targetValues = dst.target_values ? dst.target_values + s : nullptr;
            t.actionMasks = dst.action_masks ? dst.action_masks + s * numActions : nullptr;
            t.advantages = dst.advantages ? dst.advantages + s : nullptr;
            t.size = t.actionsSize = sizes[b];
            result.push_back(t);
        }
        return result;
    }

private:
    ExperienceTensors own;                  // the gathered stream
    int64_t* dPerm = nullptr;
    int64_t capacity = 0;

    void Release() {
        ggl_dev_free(own.states); ggl_dev_free(own.actions); ggl_dev_free(own.logProbs); ggl_dev_free(own.targetValues);
        ggl_dev_free(own.actionMasks); ggl_dev_free(own.advantages); ggl_dev_free(dPerm);
        own = ExperienceTensors(); dPerm = nullptr; capacity = 0;
    }
    void Reserve(int64_t rows) {
        if (rows <= capacity) return;
        Release();
        auto alloc = [&](auto** p, size_t bytes) { Check(ggl_dev_malloc(device, bytes, reinterpret_cast<void**>(p)), "ExperienceBuffer: allocation"); };
        alloc(&own.states, sizeof(float) * rows * obsSize); alloc(&own.actions, sizeof(int32_t) * rows); alloc(&own.logProbs, sizeof(float) * rows);
        alloc(&own.targetValues, sizeof(float) * rows); alloc(&own.actionMasks, (size_t)rows * numActions); alloc(&own.advantages, sizeof(float) * rows);
        alloc(&dPerm, sizeof(int64_t) * rows);
        capacity = rows;
    }
};

// `trajectories[]` + `combinedTraj` of Learner::Start as one device store (RAII over ggl_traj_*)
class TrajectoryStore {
public:
    TrajectoryStore(int device, int numPlayers, int playersPerArena, int obsSize, int numActions, int maxEpisodeLength, int64_t capacityTimesteps,
                    int64_t capacityTruncated, int64_t capacityPlayerRows = 0) {
        // capacityPlayerRows: steps all per-player buffers hold together (0 = numPlayers x maxEpisodeLength, the worst case)
        ggl_traj_desc d = {device, numPlayers, playersPerArena, obsSize, numActions, maxEpisodeLength, capacityTimesteps, capacityTruncated, capacityPlayerRows};
        Check(ggl_traj_create(&d, &h), "TrajectoryStore");
        Check(ggl_traj_views_get(h, &views), "TrajectoryStore views");
    }
    TrajectoryStore(const TrajectoryStore&) = delete;
    ~TrajectoryStore() { ggl_traj_destroy(h); }

    void Clear(void* stream = nullptr) { Check(ggl_traj_begin_iteration(h, stream), "TrajectoryStore::Clear"); }             // Learner.cpp:631
    void PushObs(const float* dObs, const uint8_t* dMasks, void* stream = nullptr) {                                          // :701-712
        Check(ggl_traj_push_obs(h, dObs, dMasks, stream), "TrajectoryStore::PushObs");
    }
    void PushResult(const int32_t* dActions, const float* dLogProbs, const float* dRewards, const uint8_t* dTerminals, const float* dObsFinal,
                    void* stream = nullptr) {                                                                                  // :823-863
        Check(ggl_traj_push_result(h, dActions, dLogProbs, dRewards, dTerminals, dObsFinal, stream), "TrajectoryStore::PushResult");
    }
    int64_t Length(void* stream = nullptr) {                                                                                   // combinedTraj.Length(), :665
        int64_t L = 0, K = 0;
        Check(ggl_traj_counts(h, stream, &L, &K), "TrajectoryStore::Length");
        numTruncated = K;
        return L;
    }
    ggl_traj_views views;                   // device pointers of the combined stream
    int64_t numTruncated = 0;               // rows of views.next_states as of the last Length()
    ggl_traj* handle() { return h; }

private:
    ggl_traj* h = nullptr;
};

namespace GAE {
// GAE.cpp:7-19 with device tensors; `workspace` = ggl_gae_workspace_bytes(L) bytes of device memory. Throws on the truncation-count
// mismatch of GAE.cpp:196-197 when checked by the caller through dTruncCount, like the reference's RG_ERR_CLOSE.
inline void Compute(const float* dRews, const int8_t* dTerminals, const float* dValPreds, const float* dTruncValPreds, int64_t L, int64_t K,
                    float* dOutAdvantages, float* dOutTargetValues, float* dOutReturns, float* dOutRewClipPortion, int32_t* dTruncCount,
                    float gamma, float lambda, float returnStd, float clipRange, void* workspace, void* stream = nullptr) {
    Check(ggl_gae_compute(dRews, dTerminals, dValPreds, dTruncValPreds, L, K, gamma, lambda, returnStd, clipRange, dOutAdvantages, dOutReturns,
                          dOutTargetValues, dOutRewClipPortion, dTruncCount, workspace, stream), "GAE::Compute");
}
}  // namespace GAE

}  // namespace GGL
