// Drives the g++-only learner-side mirror (host/GigaLearnCPP/Experience.h) the way PPOLearner::Learn and Learner::Start drive the
// reference classes: two epochs of GetAllBatchesShuffled on one ExperienceBuffer (the RNG is stateful across epochs,
// ExperienceBuffer.cpp:8-9,130), and a short collection loop through the TrajectoryStore. Everything is checked on the host.
#include <GigaLearnCPP/Experience.h>

#include <cmath>
#include <cstdio>
#include <cstring>

template <class T>
static T* to_dev(const std::vector<T>& v) {
    void* d = nullptr;
    GGL::Check(ggl_dev_malloc(0, sizeof(T) * v.size(), &d), "alloc");
    GGL::Check(ggl_dev_upload(d, v.data(), sizeof(T) * v.size(), nullptr), "upload");
    return static_cast<T*>(d);
}
template <class T>
static std::vector<T> to_host(const T* d, size_t n) {
    std::vector<T> v(n);
    GGL::Check(ggl_dev_download(v.data(), d, sizeof(T) * n, nullptr), "download");
    return v;
}

static int test_experience() {
    const int64_t rows = 5000, batch = 1024;
    const int obs = 167, na = 90, seed = 3;
    std::vector<float> states(rows * obs), logp(rows), tv(rows), adv(rows);
    std::vector<int32_t> actions(rows);
    std::vector<uint8_t> masks(rows * na);
    for (int64_t i = 0; i < rows; i++) {
        for (int k = 0; k < obs; k++) states[i * obs + k] = (float)(i * 7 + k);
        for (int k = 0; k < na; k++) masks[i * na + k] = (uint8_t)((i + k) % 3 == 0);
        actions[i] = (int32_t)(i % 97) - 2; logp[i] = (float)i; tv[i] = (float)i * 0.5f; adv[i] = -(float)i;
    }
    GGL::ExperienceBuffer buf(seed, 0);
    buf.obsSize = obs; buf.numActions = na; buf.maxActionIndex = na - 1;
    buf.data.states = to_dev(states); buf.data.actions = to_dev(actions); buf.data.logProbs = to_dev(logp);
    buf.data.targetValues = to_dev(tv); buf.data.actionMasks = to_dev(masks); buf.data.advantages = to_dev(adv);
    buf.data.size = buf.data.actionsSize = rows;
    // the reference's own permutation stream: same engine, same seed, iota once, shuffled again every epoch
    std::default_random_engine rng(seed);
    std::vector<int64_t> perm(rows);
    std::iota(perm.begin(), perm.end(), 0);
    int ok = 1;
    for (int epoch = 0; epoch < 2; epoch++) {
        std::shuffle(perm.begin(), perm.end(), rng);
        auto batches = buf.GetAllBatchesShuffled(batch, true);
        ok &= batches.size() == 4;                                    // 5000 / 1024 with overbatching: 1024 x 3 + 1928
        int64_t at = 0;
        for (auto& b : batches) {
            auto s = to_host(b.states, b.size * obs); auto a = to_host(b.actions, b.size); auto l = to_host(b.logProbs, b.size);
            auto m = to_host(b.actionMasks, b.size * na); auto t = to_host(b.targetValues, b.size); auto d = to_host(b.advantages, b.size);
            for (int64_t i = 0; i < b.size; i++) {
                const int64_t r = perm[at + i];
                int32_t ea = actions[r] < 0 ? 0 : (actions[r] > na - 1 ? na - 1 : actions[r]);
                ok &= l[i] == logp[r] && a[i] == ea && t[i] == tv[r] && d[i] == adv[r];
                ok &= std::memcmp(&s[i * obs], &states[r * obs], sizeof(float) * obs) == 0;
                ok &= std::memcmp(&m[i * na], &masks[r * na], na) == 0;
            }
            at += b.size;
        }
        ok &= at == rows;
    }
    GGL::ExperienceBuffer empty(1, 0);
    ok &= empty.GetAllBatchesShuffled(256, true).empty();             // ExperienceBuffer.cpp:105-107
    std::printf("experience ok=%d\n", ok);
    return ok ? 0 : 1;
}

static int test_trajectory() {
    // 2 arenas x 1 player, maxEpisodeLength 3: arena 1 ends (NORMAL) at step 1, player 0 is force-truncated at step 2
    GGL::TrajectoryStore store(0, 2, 1, 1, 1, 3, 64, 16);
    store.Clear();
    const uint8_t terms[4][2] = {{0, 0}, {0, 1}, {0, 0}, {0, 0}};
    std::vector<uint8_t> mask = {1, 1};
    std::vector<float> lp = {0.5f, 0.5f}, rew = {1.f, 2.f};
    int64_t len = 0;
    for (int t = 0; t < 4; t++) {
        std::vector<float> obs = {10.f * t, 10.f * t + 1}, fin = {10.f * t + 0.5f, 10.f * t + 1.5f};
        std::vector<int32_t> act = {t, t};
        std::vector<uint8_t> tm = {terms[t][0], terms[t][1]};
        float* dObs = to_dev(obs); float* dFin = to_dev(fin); uint8_t* dM = to_dev(mask); int32_t* dA = to_dev(act);
        float* dL = to_dev(lp); float* dR = to_dev(rew); uint8_t* dT = to_dev(tm);
        store.PushObs(dObs, dM);
        store.PushResult(dA, dL, dR, dT, dFin);
        len = store.Length();
        for (void* p : {(void*)dObs, (void*)dFin, (void*)dM, (void*)dA, (void*)dL, (void*)dR, (void*)dT}) ggl_dev_free(p);
    }
    auto s = to_host(store.views.states, len); auto tr = to_host(store.views.terminals, len); auto nx = to_host(store.views.next_states, 1);
    const float es[5] = {1, 11, 0, 10, 20}; const int8_t et[5] = {0, 1, 0, 0, 2};
    int ok = len == 5 && store.numTruncated == 1 && nx[0] == 20.5f;
    for (int i = 0; i < 5 && ok; i++) ok &= s[i] == es[i] && tr[i] == et[i];
    bool threw = false;
    try { GGL::TrajectoryStore bad(0, 3, 2, 1, 1, 3, 64, 16); } catch (const std::runtime_error&) { threw = true; }    // players not a multiple of the arena size
    std::printf("trajectory ok=%d bad_desc_throws=%d\n", ok, (int)threw);
    return ok && threw ? 0 : 1;
}

int main() {
    try {
        return test_experience() | test_trajectory();
    } catch (const std::exception& e) {
        std::printf("exception: %s\n", e.what());
        return 2;
    }
}
