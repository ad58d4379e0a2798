// TEST INFRASTRUCTURE — NOT PART OF THE PRODUCT.
// C-ABI wrapper around the reference's own GGL::ExperienceBuffer (GigaLearnCPP/src/private/GigaLearnCPP/PPO/ExperienceBuffer.cpp:
// _GetSamples :15-98, GetAllBatchesShuffled :100-163), compiled unmodified against the pip libtorch into
// oracle/_ref/libref_expbuf.so. The buffer object is kept between calls because the reference re-shuffles the PREVIOUS
// permutation when the size is unchanged (:121-130).
#include <GigaLearnCPP/PPO/ExperienceBuffer.h>
#include <cstring>

extern "C" void* ref_expbuf_create(int seed, int64_t maxActionIndex) {
    auto* b = new GGL::ExperienceBuffer(seed, torch::kCPU);
    b->maxActionIndex = maxActionIndex;
    return b;
}
extern "C" void ref_expbuf_destroy(void* p) { delete (GGL::ExperienceBuffer*)p; }

// installs copies of the six tensors (Learner.cpp:975-980); rowsActions may differ from rows to exercise the min() rule
extern "C" void ref_expbuf_set_data(void* p, int64_t rows, int64_t rowsActions, int64_t obs, int64_t numActions, const float* states,
                                    const int32_t* actions, const float* logProbs, const float* targetValues, const uint8_t* masks,
                                    const float* advantages) {
    auto* b = (GGL::ExperienceBuffer*)p;
    auto f32 = torch::TensorOptions().dtype(torch::kFloat32);
    b->data.states = torch::from_blob((void*)states, {rows, obs}, f32).clone();
    b->data.actions = torch::from_blob((void*)actions, {rowsActions}, torch::TensorOptions().dtype(torch::kInt32)).clone();
    b->data.logProbs = torch::from_blob((void*)logProbs, {rows}, f32).clone();
    b->data.targetValues = torch::from_blob((void*)targetValues, {rows}, f32).clone();
    b->data.actionMasks = torch::from_blob((void*)masks, {rows, numActions}, torch::TensorOptions().dtype(torch::kUInt8)).clone();
    b->data.advantages = torch::from_blob((void*)advantages, {rows}, f32).clone();
}

// One GetAllBatchesShuffled call. Batches are concatenated into the out* arrays (caller sizes them for `rows` samples);
// outBatchSizes[maxBatches] receives the size of every batch. Returns the number of batches, or -1.
extern "C" int64_t ref_expbuf_get_batches(void* p, int64_t batchSize, int overbatching, int64_t maxBatches, int64_t* outBatchSizes,
                                          float* outStates, int32_t* outActions, float* outLogProbs, float* outTargetValues,
                                          uint8_t* outMasks, float* outAdvantages) {
    try {
        auto* b = (GGL::ExperienceBuffer*)p;
        auto batches = b->GetAllBatchesShuffled(batchSize, overbatching != 0);
        if ((int64_t)batches.size() > maxBatches) return -1;
        int64_t at = 0;
        const int64_t obs = b->data.states.size(1), na = b->data.actionMasks.size(1);
        for (size_t i = 0; i < batches.size(); i++) {
            auto& t = batches[i];
            const int64_t n = t.states.size(0);
            outBatchSizes[i] = n;
            std::memcpy(outStates + at * obs, t.states.contiguous().data_ptr<float>(), n * obs * sizeof(float));
            std::memcpy(outActions + at, t.actions.contiguous().data_ptr<int32_t>(), n * sizeof(int32_t));
            std::memcpy(outLogProbs + at, t.logProbs.contiguous().data_ptr<float>(), n * sizeof(float));
            std::memcpy(outTargetValues + at, t.targetValues.contiguous().data_ptr<float>(), n * sizeof(float));
            std::memcpy(outMasks + at * na, t.actionMasks.contiguous().data_ptr<uint8_t>(), n * na);
            std::memcpy(outAdvantages + at, t.advantages.contiguous().data_ptr<float>(), n * sizeof(float));
            at += n;
        }
        return (int64_t)batches.size();
    } catch (std::exception& e) {
        fprintf(stderr, "ref_expbuf_get_batches: %s\n", e.what());
        return -1;
    }
}
