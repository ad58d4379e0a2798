// TEST INFRASTRUCTURE (oracle): C harness over the UNMODIFIED reference GGL::BatchedWelfordStat
// (GigaLearnCPP/src/private/GigaLearnCPP/Util/WelfordStat.h:69-243), compiled from the reference sources where they lie.
// The obs-normalisation statistics of Learner.cpp:680-692: IncrementRow over sampled rows, NormalizeInPlace over all rows.
#include <GigaLearnCPP/Util/WelfordStat.h>

#include <cstdint>
#include <cstring>

extern "C" {
void* ref_welford_create(int width) { return new GGL::BatchedWelfordStat(width); }
void ref_welford_destroy(void* h) { delete static_cast<GGL::BatchedWelfordStat*>(h); }
void ref_welford_increment_rows(void* h, float* obs, int width, const int32_t* idx, int n) {
    auto* s = static_cast<GGL::BatchedWelfordStat*>(h);
    for (int k = 0; k < n; k++) s->IncrementRow(obs + (int64_t)idx[k] * width);       // Learner.cpp:683-686
}
void ref_welford_normalize(void* h, float* obs, int rows, int width, double meanClamp, double minSTD) {
    static_cast<GGL::BatchedWelfordStat*>(h)->NormalizeInPlace(obs, rows, width, meanClamp, minSTD);   // Learner.cpp:688-694
}
int64_t ref_welford_state(void* h, double* means, double* vars) {
    auto* s = static_cast<GGL::BatchedWelfordStat*>(h);
    std::memcpy(means, s->runningMeans.data(), sizeof(double) * s->width);
    std::memcpy(vars, s->runningVariances.data(), sizeof(double) * s->width);
    return s->count;
}
}
