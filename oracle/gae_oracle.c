/* TEST INFRASTRUCTURE — NOT PART OF THE PRODUCT. Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline leg may call this.
 *
 * Plain-C restatement of GGL::GAE::Compute
 * (reference: GigaLearnCPP/src/private/GigaLearnCPP/PPO/GAE.cpp:7-208).  Pinned against the reference itself:
 * tests/test_oracle_cpu.py compares it bit-for-bit with oracle/_ref/libref_gae.so (the unmodified GAE.cpp compiled
 * against libtorch) and with the golden vectors in tests/golden/gae_*.npz generated from that library.
 *
 * terminals: 0 NOT_TERMINAL, 1 NORMAL, 2 TRUNCATED (RLGymCPP TerminalCondition.h:6-11).
 * Returns 0, or -1 on the truncation-count mismatch the reference treats as fatal (GAE.cpp:196-197).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

static float clampf(float v, float lo, float hi) { /* RS_CLAMP = RS_MIN(RS_MAX(v, lo), hi) */
    float m = (v > lo) ? v : lo;
    return (m < hi) ? m : hi;
}

int oracle_gae_compute(const float* rews, const int8_t* terminals, const float* val_preds,
                       const float* trunc_val_preds /* NULL = undefined tensor */, int64_t L, int64_t K,
                       float gamma, float lambda, float return_std, float clip_range,
                       float* out_adv, float* out_ret, float* out_target, float* out_clip_portion) {
    if (L == 0) { if (out_clip_portion) *out_clip_portion = 0.f; return 0; }           /* GAE.cpp:14-20 */
    const int has_trunc = trunc_val_preds != NULL;
    const int should_norm = (return_std != 0.f && return_std != 1.f);                    /* :53 */
    const int should_clip = clip_range > 0.f;                                            /* :54 */
    const float inv_std = should_norm ? (1.0f / return_std) : 1.0f;                      /* :55 */
    const float gamma_lambda = gamma * lambda;                                           /* :56 */
    const int64_t last = L - 1;

    float* next_val = (float*)malloc(sizeof(float) * (size_t)L);
    float* nd = (float*)malloc(sizeof(float) * (size_t)L);
    float* norm = should_norm ? (float*)malloc(sizeof(float) * (size_t)L) : NULL;

    /* pass 1+2 (:68-102): rank of each TRUNCATED step among the truncations (the reference finds it by linear
     * search in the sorted list of truncated indices = the running count), next value prediction, continuation mask */
    int64_t trunc_seen = 0;
    for (int64_t i = 0; i < L; i++) {
        const int8_t t = terminals[i];
        const float done = (t == 1) ? 1.f : 0.f, trunc = (t == 2) ? 1.f : 0.f;
        nd[i] = (1.f - done) * (1.f - trunc);
        if (t == 1) {
            next_val[i] = 0.f;
        } else if (t == 2 && has_trunc) {
            const int64_t tidx = trunc_seen;
            next_val[i] = (tidx >= 0 && tidx < K) ? trunc_val_preds[tidx] : 0.f;
        } else if (i < last) {
            next_val[i] = val_preds[i + 1];
        } else {
            next_val[i] = 0.f;
        }
        if (t == 2) trunc_seen++;
    }

    /* pass 3 (:109-164): reward normalisation + clipping, float accumulation in stream order */
    float tot = 0.f, tot_clipped = 0.f;
    const float* rptr = rews;
    if (should_norm) {
        /* the reference unrolls by 8 and adds the 8 magnitudes together BEFORE accumulating (:117-147);
         * float addition is not associative, so the grouping is part of the result */
        int64_t i = 0;
        const int64_t unroll_end = L - (L % 8);
        for (; i < unroll_end; i += 8) {
            float n[8], a = 0.f, c = 0.f;
            for (int e = 0; e < 8; e++) n[e] = rews[i + e] * inv_std;
            a = fabsf(n[0]) + fabsf(n[1]) + fabsf(n[2]) + fabsf(n[3]) + fabsf(n[4]) + fabsf(n[5]) + fabsf(n[6]) + fabsf(n[7]);
            tot += a;
            if (should_clip) for (int e = 0; e < 8; e++) n[e] = clampf(n[e], -clip_range, clip_range);
            c = fabsf(n[0]) + fabsf(n[1]) + fabsf(n[2]) + fabsf(n[3]) + fabsf(n[4]) + fabsf(n[5]) + fabsf(n[6]) + fabsf(n[7]);
            tot_clipped += c;
            for (int e = 0; e < 8; e++) norm[i + e] = n[e];
        }
        for (; i < L; i++) {
            float n = rews[i] * inv_std;
            tot += fabsf(n);
            if (should_clip) n = clampf(n, -clip_range, clip_range);
            tot_clipped += fabsf(n);
            norm[i] = n;
        }
        rptr = norm;
    }

    /* pass 4 (:174-193) */
    float prev_lambda = 0.f, prev_ret = 0.f;
    for (int64_t i = last; i >= 0; i--) {
        const float pred_return = rptr[i] + gamma * next_val[i];
        const float delta = pred_return - val_preds[i];
        const float cur_ret = rews[i] + prev_ret * gamma * nd[i];
        out_ret[i] = cur_ret;
        prev_lambda = delta + gamma_lambda * nd[i] * prev_lambda;
        out_adv[i] = prev_lambda;
        prev_ret = cur_ret;
    }
    for (int64_t i = 0; i < L; i++) out_target[i] = val_preds[i] + out_adv[i];           /* :200 */

    if (out_clip_portion) {                                                              /* :203-207 */
        if (should_norm) {
            const float d = tot > 1e-7f ? tot : 1e-7f;
            *out_clip_portion = (tot - tot_clipped) / d;
        } else {
            *out_clip_portion = 0.f;
        }
    }
    free(next_val); free(nd); if (norm) free(norm);
    if (has_trunc && trunc_seen != K) return -1;
    return 0;
}
