// TEST INFRASTRUCTURE — NOT PART OF THE PRODUCT.
// C-ABI wrapper around the reference's own GGL::GAE::Compute (GigaLearnCPP/src/private/GigaLearnCPP/PPO/GAE.cpp:7-208),
// compiled unmodified against the pip libtorch into oracle/_ref/libref_gae.so.
#include <GigaLearnCPP/PPO/GAE.h>
#include <cstring>

extern "C" int ref_gae_compute(const float* rews, const int8_t* terminals, const float* valPreds,
                               const float* truncValPreds, int64_t L, int64_t K,
                               float gamma, float lambda, float returnStd, float clipRange,
                               float* outAdv, float* outRet, float* outTarget, float* outClipPortion) {
    try {
        auto f32 = torch::TensorOptions().dtype(torch::kFloat32);
        auto i8 = torch::TensorOptions().dtype(torch::kInt8);
        torch::Tensor tR = torch::from_blob((void*)rews, {L}, f32);
        torch::Tensor tT = torch::from_blob((void*)terminals, {L}, i8);
        torch::Tensor tV = torch::from_blob((void*)valPreds, {L}, f32);
        torch::Tensor tTV;
        if (truncValPreds) tTV = torch::from_blob((void*)truncValPreds, {K}, f32);
        torch::Tensor adv, tgt, ret;
        float clip = 0;
        GGL::GAE::Compute(tR, tT, tV, tTV, adv, tgt, ret, clip, gamma, lambda, returnStd, clipRange);
        std::memcpy(outAdv, adv.data_ptr<float>(), L * sizeof(float));
        std::memcpy(outRet, ret.data_ptr<float>(), L * sizeof(float));
        tgt = tgt.contiguous();
        std::memcpy(outTarget, tgt.data_ptr<float>(), L * sizeof(float));
        if (outClipPortion) *outClipPortion = clip;
        return 0;
    } catch (std::exception& e) {
        fprintf(stderr, "ref_gae_compute: %s\n", e.what());
        return -1;
    }
}
