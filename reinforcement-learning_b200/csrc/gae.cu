// GAE / returns over the flat experience stream — single-pass segmented reverse scan for sm_100a.
//
// Replaces GGL::GAE::Compute (GigaLearnCPP/src/private/GigaLearnCPP/PPO/GAE.cpp:7-208), which walks the
// stream backwards on one CPU thread:
//     delta_i = r~_i + gamma * nextV_i - V_i            (r~ = reward / returnStd, clipped to +-clipRange)
//     adv_i   = delta_i + (gamma*lambda*nd_i) * adv_{i+1}
//     ret_i   = r_i     + (gamma*nd_i)        * ret_{i+1}      (raw reward, never bootstrapped — GAE.cpp:180-182)
//     target_i = V_i + adv_i
// nd_i = 1 only when terminals[i] == 0.  nextV_i is 0 after a NORMAL terminal, truncValPreds[rank] after a
// TRUNCATED one (rank = number of TRUNCATED entries before i, GAE.cpp:68-95), else V_{i+1} (0 past the end).
//
// Both recurrences are first-order linear, x_i = a_i * x_{i+1} + b_i, so a span of elements is the affine
// map (A, B) and spans compose associatively.  The kernel is a decoupled-look-back chained scan running from
// the END of the stream: tiles of 4096 elements are handed out through an atomic ticket in scan order (so a
// tile's predecessors are always already running; the next ticket is requested while the current tile is processed), each
// tile reads its 9 B/element once through coalesced streaming loads (held in registers while the integer chain
// resolves, then transposed through padded shared memory), scans with warp shuffles, publishes its aggregate as a
// self-validating 16-byte record, resolves its carry-in with 128-tile look-back windows (one global round trip per
// window), and writes 12 B/element once.  Algorithmic traffic = 21 B/element — the
// HBM roofline of this path.  The truncation rank is obtained in the same pass as
// K - (suffix count of TRUNCATED entries), a second (integer) chain with single-word status records.
//
// Compiled with -fmad=false: per-element arithmetic then rounds exactly like the reference's scalar loop; only
// the cross-thread carry (composition order) differs, and its effect decays by gamma*lambda per step.
#include <mutex>
#include "common.cuh"

namespace {

#ifndef GAE_CFG_THREADS
#define GAE_CFG_THREADS 256
#endif
#ifndef GAE_CFG_ITEMS
#define GAE_CFG_ITEMS 16
#endif
#ifndef GAE_CFG_MINBLOCKS
#define GAE_CFG_MINBLOCKS 4
#endif
constexpr int GAE_THREADS = GAE_CFG_THREADS;
constexpr int GAE_ITEMS = GAE_CFG_ITEMS;                // multiple of 16 (terminals are fetched 16 per load)
constexpr int GAE_TILE = GAE_THREADS * GAE_ITEMS;       // 4096 elements
constexpr int GAE_PAD = GAE_TILE / 32;
constexpr int GAE_WARPS = GAE_THREADS / 32;

struct Lin2 {            // x_adv' = Aa*x_adv + Ba ; x_ret' = Ar*x_ret + Br
    float Aa, Ba, Ar, Br;
};

__device__ __forceinline__ Lin2 lin_identity() { return Lin2{1.f, 0.f, 1.f, 0.f}; }
// `later` is applied after `earlier` (earlier = closer to the end of the stream)
__device__ __forceinline__ Lin2 compose(const Lin2& later, const Lin2& earlier) {
    Lin2 r;
    r.Aa = later.Aa * earlier.Aa;
    r.Ba = later.Aa * earlier.Ba + later.Ba;
    r.Ar = later.Ar * earlier.Ar;
    r.Br = later.Ar * earlier.Br + later.Br;
    return r;
}
__device__ __forceinline__ Lin2 shfl_down_lin(const Lin2& v, int off) {
    Lin2 r;
    r.Aa = __shfl_down_sync(0xffffffffu, v.Aa, off);
    r.Ba = __shfl_down_sync(0xffffffffu, v.Ba, off);
    r.Ar = __shfl_down_sync(0xffffffffu, v.Ar, off);
    r.Br = __shfl_down_sync(0xffffffffu, v.Br, off);
    return r;
}

struct GaeHeader {
    unsigned int tile_ticket;
    unsigned int done_ctas;
    unsigned int pre_done;            // blocks of the count pre-pass that have finished
    unsigned int trunc_total;         // number of TRUNCATED entries in the stream
    double sum_abs;
    double sum_abs_clipped;
};

struct GaeParams {
    const float* rews;
    const int8_t* terms;
    const float* vals;
    const float* trunc_vals;
    long long L;
    long long K;
    float gamma, gamma_lambda, inv_std, clip;
    float* out_adv;
    float* out_ret;
    float* out_tgt;
    float* out_clip_portion;
    int* out_trunc_count;
    // workspace
    GaeHeader* hdr;
    const unsigned int* cnt_in;       // [tiles]  TRUNCATED entries in the tiles before this one in scan order (pre-pass)
    ulonglong2* val_rec;              // [tiles]  self-validating (advantage, return) records, see enc_agg / enc_inc
    int num_tiles;
};

__device__ __forceinline__ int pidx(int i) { return i + (i >> 5); }

__device__ __forceinline__ unsigned long long ld_volatile_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_volatile_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ ulonglong2 ld_volatile_rec(const ulonglong2* p) {
    ulonglong2 v;
    asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_volatile_rec(ulonglong2* p, ulonglong2 v) {
    asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(v.x), "l"(v.y) : "memory");
}


// Look-back records of the (advantage, return) chain. One record = two 64-bit words, word0 = (Aa | Ba<<32) for the
// advantage recurrence and word1 = (Ar | Br<<32) for the return recurrence. The multiplier A of a span is a product of
// gamma*lambda*nd >= 0, so its sign bit is free: an AGGREGATE word carries A with the sign bit set (never 0), an
// INCLUSIVE word carries 0xFFFFFFFF in the A half (B = the finished prefix value), and an all-zero word means "not
// published yet". Each 64-bit half is written atomically, so a record validates itself: one global round trip per
// look-back window instead of flag-then-data.
constexpr unsigned int REC_INC = 0xFFFFFFFFu;
#ifndef GAE_LB_K0
#define GAE_LB_K0 1
#endif
#ifndef GAE_LB_K
#define GAE_LB_K 1
#endif
#ifndef GAE_SLEEP_NS
#define GAE_SLEEP_NS 100
#endif
constexpr int LB_K0 = GAE_LB_K0;              // records per lane in the first (nearest) look-back window
constexpr int LB_K = GAE_LB_K;                // records per lane in the deeper windows (128 tiles per poll)
constexpr int LB_BIG = 1 << 20;

__device__ __forceinline__ ulonglong2 enc_agg(const Lin2& m) {
    ulonglong2 w;
    w.x = (unsigned long long)(__float_as_uint(m.Aa) | 0x80000000u) | ((unsigned long long)__float_as_uint(m.Ba) << 32);
    w.y = (unsigned long long)(__float_as_uint(m.Ar) | 0x80000000u) | ((unsigned long long)__float_as_uint(m.Br) << 32);
    return w;
}
__device__ __forceinline__ ulonglong2 enc_inc(float adv, float ret) {
    ulonglong2 w;
    w.x = (unsigned long long)REC_INC | ((unsigned long long)__float_as_uint(adv) << 32);
    w.y = (unsigned long long)REC_INC | ((unsigned long long)__float_as_uint(ret) << 32);
    return w;
}
// 0 = not ready (or torn between two states), 1 = aggregate, 2 = inclusive
__device__ __forceinline__ int rec_state(const ulonglong2& w) {
    if (w.x == 0ull || w.y == 0ull) return 0;
    const bool i0 = (unsigned int)w.x == REC_INC, i1 = (unsigned int)w.y == REC_INC;
    if (i0 != i1) return 0;
    return i0 ? 2 : 1;
}
__device__ __forceinline__ Lin2 rec_decode(const ulonglong2& w, int state) {
    Lin2 m;
    m.Aa = (state == 2) ? 0.f : __uint_as_float((unsigned int)w.x & 0x7FFFFFFFu);
    m.Ba = __uint_as_float((unsigned int)(w.x >> 32));
    m.Ar = (state == 2) ? 0.f : __uint_as_float((unsigned int)w.y & 0x7FFFFFFFu);
    m.Br = __uint_as_float((unsigned int)(w.y >> 32));
    return m;
}

// One look-back window of K records per lane starting at tile `win` and going down; composes the records up to the first
// INCLUSIVE one into `acc` (valid on lane 0) and returns true when it found one. Spins (with a short back-off, so that
// ~600 polling warps do not eat the L2 bandwidth the streaming loads need) while a record it needs is unpublished.
template <int K>
__device__ __forceinline__ bool lb_vals_window(const ulonglong2* rec, int win, int lane, Lin2& acc) {
    for (;;) {
        ulonglong2 w[K];
        int st[K];
        int my_inc = LB_BIG, my_none = LB_BIG;
#pragma unroll
        for (int j = 0; j < K; j++) {
            const int q = win - (lane * K + j);
            if (q >= 0) w[j] = ld_volatile_rec(rec + q);
            else { w[j].x = (unsigned long long)REC_INC; w[j].y = (unsigned long long)REC_INC; }   // before the stream end: x = 0
        }
#pragma unroll
        for (int j = K - 1; j >= 0; j--) {
            st[j] = rec_state(w[j]);
            if (st[j] == 2) my_inc = lane * K + j;
            if (st[j] == 0) my_none = lane * K + j;
        }
        const int first_inc = __reduce_min_sync(0xffffffffu, my_inc);
        const int first_none = __reduce_min_sync(0xffffffffu, my_none);
        if (first_none < first_inc) { __nanosleep(GAE_SLEEP_NS); continue; }
        Lin2 c = lin_identity();
#pragma unroll
        for (int j = 0; j < K; j++)
            if (lane * K + j <= first_inc) c = compose(c, rec_decode(w[j], st[j]));
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const Lin2 o = shfl_down_lin(c, off);
            if (lane + off < 32) c = compose(c, o);
        }
        acc = compose(acc, c);
        return first_inc < LB_BIG;
    }
}

// warp 0: carry-in (advantage, return) of tile t from the tiles before it in scan order (valid on lane 0). The nearest
// 32*LB_K0 tiles are polled first (they are the ones still being computed); deeper windows are 32*LB_K wide.
__device__ __forceinline__ void lookback_vals(const ulonglong2* rec, int t, int lane, float& adv_in, float& ret_in) {
    Lin2 acc = lin_identity();
    int win = t - 1;
    if (!lb_vals_window<LB_K0>(rec, win, lane, acc)) {
        win -= 32 * LB_K0;
        while (!lb_vals_window<LB_K>(rec, win, lane, acc)) win -= 32 * LB_K;
    }
    adv_in = acc.Ba;
    ret_in = acc.Br;
}

// ---- truncation ranks: a pre-pass over the terminals only (1 B/element, +4.8 % traffic) ----
// truncValPreds[rank] needs, for every TRUNCATED element, the number of TRUNCATED elements after it in the stream
// (GAE.cpp:68-95). Resolving that count with a second look-back chain inside the scan kernel made 7 of 8 warps of every tile
// wait at a barrier for it (ncu: 42 % of all stall samples), because the tiles in flight all reach that point together and
// none of them is inclusive yet. Here every warp counts whole tiles, and the last block to finish turns the per-tile counts
// into exclusive suffix counts in scan order; the scan kernel then reads one word per tile.
constexpr int CNT_THREADS = 256;
__device__ __forceinline__ int count_twos(unsigned int w) {
    const unsigned int x = w ^ 0x02020202u;                                                   // bytes equal to 2 become 0
    const unsigned int z = ~(((x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | x | 0x7F7F7F7Fu);           // 0x80 in every zero byte of x
    return __popc(z);
}
__global__ void __launch_bounds__(CNT_THREADS) gae_trunc_count_kernel(const int8_t* __restrict__ terms, long long L, int num_tiles,
                                                                       unsigned int* tile_cnt, unsigned int* cnt_in, GaeHeader* hdr) {
    const int lane = threadIdx.x & 31;
    const int wid = (blockIdx.x * CNT_THREADS + threadIdx.x) >> 5, nw = (gridDim.x * CNT_THREADS) >> 5;
    for (int t = wid; t < num_tiles; t += nw) {
        const long long base = (long long)(num_tiles - 1 - t) * GAE_TILE;
        int c = 0;
#pragma unroll
        for (int k = 0; k < GAE_TILE / (32 * 16); k++) {
            const long long gi = base + ((long long)k * 32 + lane) * 16;
            if (gi + 15 < L) {
                const uint4 w = __ldcs(reinterpret_cast<const uint4*>(terms + gi));
                c += count_twos(w.x) + count_twos(w.y) + count_twos(w.z) + count_twos(w.w);
            } else {
                for (int e = 0; e < 16; e++)
                    if (gi + e < L) c += terms[gi + e] == 2;
            }
        }
        c = __reduce_add_sync(0xffffffffu, c);
        if (lane == 0) __stcg(&tile_cnt[t], (unsigned int)c);
    }
    // last block: exclusive prefix over the tiles in scan order
    __shared__ unsigned int s_last;
    __shared__ unsigned int s_part[CNT_THREADS];
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = (atomicAdd(&hdr->pre_done, 1u) == gridDim.x - 1) ? 1u : 0u;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    const int per = (num_tiles + CNT_THREADS - 1) / CNT_THREADS;
    const int b0 = threadIdx.x * per, b1 = min(num_tiles, b0 + per);
    // the thread's run of per-tile counts: up to 32 of them are fetched as independent loads and kept in registers
    unsigned int vals[32];
    unsigned int sum = 0;
    if (per <= 32) {
#pragma unroll
        for (int k = 0; k < 32; k++) vals[k] = (b0 + k < b1) ? __ldcg(&tile_cnt[b0 + k]) : 0u;
#pragma unroll
        for (int k = 0; k < 32; k++) sum += vals[k];
    } else {
#pragma unroll 8
        for (int t = b0; t < b1; t++) sum += __ldcg(&tile_cnt[t]);
    }
    s_part[threadIdx.x] = sum;
    __syncthreads();
    {   // exclusive scan of the CNT_THREADS partial sums: warp shuffles + one pass over the warp totals
        __shared__ unsigned int s_wsum[CNT_THREADS / 32];
        unsigned int inc = sum;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const unsigned int o = __shfl_up_sync(0xffffffffu, inc, off);
            if (lane >= off) inc += o;
        }
        if (lane == 31) s_wsum[threadIdx.x >> 5] = inc;
        __syncthreads();
        unsigned int wbase = 0, total = 0;
#pragma unroll
        for (int w = 0; w < CNT_THREADS / 32; w++) {
            if (w < (int)(threadIdx.x >> 5)) wbase += s_wsum[w];
            total += s_wsum[w];
        }
        s_part[threadIdx.x] = wbase + inc - sum;
        if (threadIdx.x == 0) hdr->trunc_total = total;
    }
    __syncthreads();
    unsigned int run = s_part[threadIdx.x];
    if (per <= 32) {
#pragma unroll
        for (int k = 0; k < 32; k++)
            if (b0 + k < b1) { cnt_in[b0 + k] = run; run += vals[k]; }
    } else {
        for (int t = b0; t < b1; t++) { cnt_in[t] = run; run += __ldcg(&tile_cnt[t]); }
    }
}

template <bool HAS_TRUNC, bool NORM, bool CLIP>
__global__ void __launch_bounds__(GAE_THREADS, GAE_CFG_MINBLOCKS) gae_scan_kernel(GaeParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* s_r = reinterpret_cast<float*>(smem_raw);                 // rewards -> advantages
    float* s_v = s_r + (GAE_TILE + GAE_PAD + 4);                     // values (+1 lookahead) -> returns
    float* s_o = s_v + (GAE_TILE + GAE_PAD + 4);                     // b_adv -> target values

    __shared__ int s_tile;
    __shared__ Lin2 s_warp[GAE_WARPS];
    __shared__ int s_wcnt[GAE_WARPS];
    __shared__ float s_x_in[2];
    __shared__ double s_red[2][GAE_WARPS];

    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int warp = tid >> 5;
    const float g = p.gamma, gl = p.gamma_lambda;

    if (tid == 0) s_tile = (int)atomicAdd(&p.hdr->tile_ticket, 1u);
    __syncthreads();
    for (;;) {
        const int t = s_tile;
        if (t >= p.num_tiles) break;
        const long long base = (long long)(p.num_tiles - 1 - t) * GAE_TILE;
        // the NEXT ticket is requested now and consumed at the bottom of the loop: its round trip hides behind this tile
        unsigned int next_ticket = 0;
        if (tid == 0) next_ticket = atomicAdd(&p.hdr->tile_ticket, 1u);

        // terminals first (1 B/element): the integer chain below only needs them, and resolves while the 8 B/element of
        // rewards and values are still in flight. Every thread owns GAE_ITEMS consecutive elements; the HIGHEST thread id comes first in scan order
        unsigned int tword[GAE_ITEMS / 4];
#pragma unroll
        for (int k = 0; k < GAE_ITEMS / 16; k++) {
            const long long gi = base + (long long)tid * GAE_ITEMS + 16 * k;
            uint4 tw;
            if (gi + 15 < p.L) {
                tw = __ldcs(reinterpret_cast<const uint4*>(p.terms + gi));
            } else {
                unsigned int w[4] = {0, 0, 0, 0};
                for (int e = 0; e < 16; e++)
                    if (gi + e < p.L) w[e >> 2] |= (unsigned int)(unsigned char)p.terms[gi + e] << (8 * (e & 3));
                tw = make_uint4(w[0], w[1], w[2], w[3]);
            }
            tword[4 * k] = tw.x; tword[4 * k + 1] = tw.y; tword[4 * k + 2] = tw.z; tword[4 * k + 3] = tw.w;
        }
        // ---- coalesced streaming loads, issued up front into registers (they stay in flight during the count chain) ----
        float4 r4[GAE_ITEMS / 4], v4[GAE_ITEMS / 4];
#pragma unroll
        for (int c = 0; c < GAE_ITEMS / 4; c++) {
            const int li = (c * GAE_THREADS + tid) * 4;
            const long long gi = base + li;
            if (gi + 3 < p.L) {
                r4[c] = __ldcs(reinterpret_cast<const float4*>(p.rews + gi));
                v4[c] = __ldcs(reinterpret_cast<const float4*>(p.vals + gi));
            } else {
                float rr[4], vv[4];
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    const bool ok = gi + e < p.L;
                    rr[e] = ok ? p.rews[gi + e] : 0.f;
                    vv[e] = ok ? p.vals[gi + e] : 0.f;
                }
                r4[c] = make_float4(rr[0], rr[1], rr[2], rr[3]);
                v4[c] = make_float4(vv[0], vv[1], vv[2], vv[3]);
            }
        }
        float v_look = 0.f;
        if (tid == 0) {
            const long long gi = base + GAE_TILE;
            v_look = (gi < p.L) ? p.vals[gi] : 0.f;
        }

        // ---- truncation rank: tile offset from the pre-pass + suffix count inside the tile ----
        int cnt_before = 0;   // TRUNCATED entries at higher stream indices than this thread's span
        if (HAS_TRUNC) {
            int c_loc = 0;
#pragma unroll
            for (int k = 0; k < GAE_ITEMS / 4; k++) c_loc += count_twos(tword[k]);
            int c = c_loc;
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                const int o = __shfl_down_sync(0xffffffffu, c, off);
                if (lane + off < 32) c += o;
            }
            if (lane == 0) s_wcnt[warp] = c;
            __syncthreads();
            int wcarry = 0;
#pragma unroll
            for (int w = 0; w < GAE_WARPS; w++)
                if (w > warp) wcarry += s_wcnt[w];
            cnt_before = (int)__ldg(&p.cnt_in[t]) + wcarry + (c - c_loc);
        }

        // ---- transpose through padded shared memory ----
#pragma unroll
        for (int c = 0; c < GAE_ITEMS / 4; c++) {
            const int pi = pidx((c * GAE_THREADS + tid) * 4);
            s_r[pi] = r4[c].x; s_r[pi + 1] = r4[c].y; s_r[pi + 2] = r4[c].z; s_r[pi + 3] = r4[c].w;
            s_v[pi] = v4[c].x; s_v[pi + 1] = v4[c].y; s_v[pi + 2] = v4[c].z; s_v[pi + 3] = v4[c].w;
        }
        if (tid == 0) s_v[pidx(GAE_TILE)] = v_look;
        __syncthreads();

        // ---- per-thread pass 1: build (a, b) of every element and the thread's affine map ----
        Lin2 loc = lin_identity();
        float sum_abs = 0.f, sum_clip = 0.f;
        {
            int running = cnt_before;
            const int e0 = tid * GAE_ITEMS;
            float vnext = s_v[pidx(e0 + GAE_ITEMS)];
#pragma unroll
            for (int j = GAE_ITEMS - 1; j >= 0; j--) {
                const int pi = pidx(e0 + j);
                const unsigned int term = (tword[j >> 2] >> (8 * (j & 3))) & 0xffu;
                const float r = s_r[pi];
                const float v = s_v[pi];
                float rn = r;
                if (NORM) {
                    rn = r * p.inv_std;
                    sum_abs += fabsf(rn);
                    if (CLIP) rn = fminf(fmaxf(rn, -p.clip), p.clip);
                    sum_clip += fabsf(rn);
                }
                float next_v = vnext;
                if (term == 1u) next_v = 0.f;
                if (HAS_TRUNC && term == 2u) {
                    running++;
                    const long long tidx = p.K - running;
                    next_v = (tidx >= 0 && tidx < p.K) ? __ldg(p.trunc_vals + tidx) : 0.f;
                }
                const float nd = (term == 0u) ? 1.f : 0.f;
                const float b_adv = (rn + g * next_v) - v;
                s_o[pi] = b_adv;
                Lin2 e;
                e.Aa = gl * nd; e.Ba = b_adv; e.Ar = g * nd; e.Br = r;
                loc = compose(e, loc);
                vnext = v;
            }
        }

        // ---- block scan (reverse thread order) with warp shuffles ----
        Lin2 incl = loc;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const Lin2 o = shfl_down_lin(incl, off);
            if (lane + off < 32) incl = compose(incl, o);
        }
        Lin2 excl = shfl_down_lin(incl, 1);
        if (lane == 31) excl = lin_identity();
        if (lane == 0) s_warp[warp] = incl;
        __syncthreads();
        Lin2 wcarry = lin_identity();
        Lin2 tile_agg = lin_identity();
#pragma unroll
        for (int w = GAE_WARPS - 1; w >= 0; w--) {
            if (w == warp) wcarry = tile_agg;
            tile_agg = compose(s_warp[w], tile_agg);
        }
        const Lin2 thread_in = compose(excl, wcarry);

        // ---- chain 2 (float): carry-in of (advantage, return) from the tiles after this one ----
        // A tile that contains a terminal has multiplier 0 in both recurrences: what it hands on does not depend on its own
        // carry-in, so it publishes its INCLUSIVE record at once and later tiles stop their look-back right there (with
        // episodes shorter than a tile -- the usual PPO stream -- every look-back is one window deep).
        if (warp == 0) {
            float adv_in = 0.f, ret_in = 0.f;
            const bool closed = (tile_agg.Aa == 0.f && tile_agg.Ar == 0.f);
            if (t > 0) {
                if (lane == 0) st_volatile_rec(&p.val_rec[t], closed ? enc_inc(tile_agg.Ba, tile_agg.Br) : enc_agg(tile_agg));
#ifndef GAE_DBG_NOCHAIN2
                lookback_vals(p.val_rec, t, lane, adv_in, ret_in);
#endif
            }
            if (lane == 0) {
                if (!closed || t == 0) {
                    const float adv_out = tile_agg.Aa * adv_in + tile_agg.Ba;
                    const float ret_out = tile_agg.Ar * ret_in + tile_agg.Br;
                    st_volatile_rec(&p.val_rec[t], enc_inc(adv_out, ret_out));
                }
                s_x_in[0] = adv_in;
                s_x_in[1] = ret_in;
            }
        }
        __syncthreads();

        // ---- per-thread pass 2: replay the recurrence with the exact carry-in, stage outputs ----
        {
            float x_adv = thread_in.Aa * s_x_in[0] + thread_in.Ba;
            float x_ret = thread_in.Ar * s_x_in[1] + thread_in.Br;
            const int e0 = tid * GAE_ITEMS;
#pragma unroll
            for (int j = GAE_ITEMS - 1; j >= 0; j--) {
                const int pi = pidx(e0 + j);
                const unsigned int term = (tword[j >> 2] >> (8 * (j & 3))) & 0xffu;
                const float nd = (term == 0u) ? 1.f : 0.f;
                const float r = s_r[pi];
                const float v = s_v[pi];
                const float b_adv = s_o[pi];
                x_ret = r + (x_ret * g) * nd;
                x_adv = b_adv + (gl * nd) * x_adv;
                s_r[pi] = x_adv;
                s_v[pi] = x_ret;
                s_o[pi] = v + x_adv;
            }
        }
        __syncthreads();

        // ---- coalesced streaming stores ----
#pragma unroll
        for (int c = 0; c < GAE_ITEMS / 4; c++) {
            const int li = (c * GAE_THREADS + tid) * 4;
            const long long gi = base + li;
            const int pi = pidx(li);
            if (gi + 3 < p.L) {
                __stcs(reinterpret_cast<float4*>(p.out_adv + gi), make_float4(s_r[pi], s_r[pi + 1], s_r[pi + 2], s_r[pi + 3]));
                __stcs(reinterpret_cast<float4*>(p.out_ret + gi), make_float4(s_v[pi], s_v[pi + 1], s_v[pi + 2], s_v[pi + 3]));
                __stcs(reinterpret_cast<float4*>(p.out_tgt + gi), make_float4(s_o[pi], s_o[pi + 1], s_o[pi + 2], s_o[pi + 3]));
            } else {
#pragma unroll
                for (int e = 0; e < 4; e++)
                    if (gi + e < p.L) {
                        p.out_adv[gi + e] = s_r[pi + e];
                        p.out_ret[gi + e] = s_v[pi + e];
                        p.out_tgt[gi + e] = s_o[pi + e];
                    }
            }
        }

        if (NORM) {
            double a = (double)sum_abs, b = (double)sum_clip;
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) {
                a += __shfl_xor_sync(0xffffffffu, a, off);
                b += __shfl_xor_sync(0xffffffffu, b, off);
            }
            if (lane == 0) { s_red[0][warp] = a; s_red[1][warp] = b; }
        }
        if (tid == 0) s_tile = (int)next_ticket;
        __syncthreads();   // shared tiles are reused by the next ticket; s_tile / s_red are visible
        if (NORM && tid == 0) {
            double ta = 0, tb = 0;
            for (int w = 0; w < GAE_WARPS; w++) { ta += s_red[0][w]; tb += s_red[1][w]; }
            atomicAdd(&p.hdr->sum_abs, ta);
            atomicAdd(&p.hdr->sum_abs_clipped, tb);
        }
    }

    // ---- the last CTA to leave finalises the scalars (GAE.cpp:196-207) ----
    if (tid == 0) {
        __threadfence();
        const unsigned int prev = atomicAdd(&p.hdr->done_ctas, 1u);
        if (prev == gridDim.x - 1) {
            __threadfence();
            if (p.out_clip_portion) {
                float portion = 0.f;
                if (NORM) {
                    const float tot = (float)*((volatile double*)&p.hdr->sum_abs);
                    const float clipped = (float)*((volatile double*)&p.hdr->sum_abs_clipped);
                    portion = (tot - clipped) / fmaxf(tot, 1e-7f);
                }
                *p.out_clip_portion = portion;
            }
            if (p.out_trunc_count) {
                int n = -1;
                if (HAS_TRUNC) n = (int)*((volatile unsigned int*)&p.hdr->trunc_total);
                *p.out_trunc_count = n;
            }
        }
    }
}

constexpr size_t gae_smem_bytes() { return 3 * (size_t)(GAE_TILE + GAE_PAD + 4) * sizeof(float); }

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct WsLayout {
    size_t off_rec, off_tcnt, off_cin, zero_bytes, total;
};
WsLayout ws_layout(long long L) {
    const size_t tiles = (size_t)((L + GAE_TILE - 1) / GAE_TILE);
    WsLayout w;
    size_t o = align_up(sizeof(GaeHeader), 16);
    w.off_rec = o;  o += tiles * sizeof(ulonglong2);
    w.zero_bytes = o;                                   // header + look-back records start as zeros
    w.off_tcnt = o; o += align_up(tiles * sizeof(unsigned int), 16);
    w.off_cin = o;  o += align_up(tiles * sizeof(unsigned int), 16);
    w.total = align_up(o, 256);
    return w;
}

// The opt-in shared memory size is a per-device function attribute: one process may drive several GPUs (the reference's
// Learner is one process), so it is set on every launch (cheap) rather than remembered per process.
template <bool HT, bool NM, bool CL>
int launch(const GaeParams& p, int grid, cudaStream_t s) {
    GGL_CUDA(cudaFuncSetAttribute(gae_scan_kernel<HT, NM, CL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gae_smem_bytes()));
    gae_scan_kernel<HT, NM, CL><<<grid, GAE_THREADS, gae_smem_bytes(), s>>>(p);
    GGL_LAUNCH_CHECK();
    ggl::count_launch();
    return GGL_OK;
}

// scratch of the host-buffer wrapper and occupancy of the scan kernel, cached PER DEVICE ordinal behind a mutex
struct HostScratch {
    void* d = nullptr;
    size_t bytes = 0;
    cudaStream_t stream = nullptr;
    int occ = 0;
};
constexpr int kMaxDevices = 64;
HostScratch g_scratch_dev[kMaxDevices];
std::mutex g_scratch_mutex, g_occ_mutex;

}  // namespace

extern "C" size_t ggl_gae_workspace_bytes(int64_t L) { return ws_layout(L < 1 ? 1 : L).total; }

extern "C" int ggl_gae_compute(const float* d_rews, const int8_t* d_terminals, const float* d_val_preds,
                               const float* d_trunc_val_preds, int64_t L, int64_t K,
                               float gamma, float lambda, float return_std, float clip_range,
                               float* d_out_advantages, float* d_out_returns, float* d_out_target_values,
                               float* d_out_clip_portion, int32_t* d_out_trunc_count,
                               void* d_workspace, void* stream) {
    cudaStream_t s = (cudaStream_t)stream;
    if (L < 0 || K < 0) { ggl::set_error("ggl_gae_compute: negative size"); return GGL_ERR_INVALID; }
    if (!(gamma >= 0.f) || !(lambda >= 0.f)) {   // the look-back records keep their state in the sign bit of gamma*lambda products
        ggl::set_error("ggl_gae_compute: gamma and lambda must be >= 0"); return GGL_ERR_INVALID;
    }
    if (L == 0) {   // GAE.cpp:14-20
        if (d_out_clip_portion) GGL_CUDA(cudaMemsetAsync(d_out_clip_portion, 0, sizeof(float), s));
        return GGL_OK;
    }
    if (!d_rews || !d_terminals || !d_val_preds || !d_out_advantages || !d_out_returns || !d_out_target_values || !d_workspace) {
        ggl::set_error("ggl_gae_compute: null pointer"); return GGL_ERR_INVALID;
    }
    if (((uintptr_t)d_rews | (uintptr_t)d_val_preds | (uintptr_t)d_terminals | (uintptr_t)d_out_advantages |
         (uintptr_t)d_out_returns | (uintptr_t)d_out_target_values | (uintptr_t)d_workspace) & 15) {
        ggl::set_error("ggl_gae_compute: buffers must be 16-byte aligned"); return GGL_ERR_INVALID;
    }
    const WsLayout w = ws_layout(L);
    GGL_CUDA(cudaMemsetAsync(d_workspace, 0, w.zero_bytes, s));   // header + value records
    ggl::count_launch();

    GaeParams p;
    p.rews = d_rews; p.terms = d_terminals; p.vals = d_val_preds; p.trunc_vals = d_trunc_val_preds;
    p.L = L; p.K = K;
    p.gamma = gamma; p.gamma_lambda = gamma * lambda;
    const bool norm = (return_std != 0.f && return_std != 1.f);     // GAE.cpp:53
    const bool clip = clip_range > 0.f;                             // GAE.cpp:54
    p.inv_std = norm ? 1.0f / return_std : 1.0f;
    p.clip = clip_range;
    p.out_adv = d_out_advantages; p.out_ret = d_out_returns; p.out_tgt = d_out_target_values;
    p.out_clip_portion = d_out_clip_portion; p.out_trunc_count = d_out_trunc_count;
    unsigned char* wsb = (unsigned char*)d_workspace;
    p.hdr = (GaeHeader*)wsb;
    p.cnt_in = (const unsigned int*)(wsb + w.off_cin);
    p.val_rec = (ulonglong2*)(wsb + w.off_rec);
    p.num_tiles = (int)((L + GAE_TILE - 1) / GAE_TILE);

    int dev = 0, sms = 148;
    GGL_CUDA(cudaGetDevice(&dev));
    GGL_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    int per_sm = GAE_CFG_MINBLOCKS;                                    // persistent: as many CTAs as stay resident
    if (dev >= 0 && dev < kMaxDevices) {
        std::lock_guard<std::mutex> lock(g_occ_mutex);
        int& occ = g_scratch_dev[dev].occ;
        if (!occ) {
            GGL_CUDA(cudaFuncSetAttribute(gae_scan_kernel<true, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gae_smem_bytes()));
            GGL_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, gae_scan_kernel<true, true, true>, GAE_THREADS, gae_smem_bytes()));
            if (occ < 1) occ = 1;
        }
        per_sm = occ;
    }
    const int grid = p.num_tiles < sms * per_sm ? p.num_tiles : sms * per_sm;

    const bool ht = d_trunc_val_preds != nullptr;
    if (ht) {
        const int warps_per_block = CNT_THREADS / 32;
        int cgrid = (p.num_tiles + warps_per_block - 1) / warps_per_block;
        if (cgrid > sms * 8) cgrid = sms * 8;
        gae_trunc_count_kernel<<<cgrid, CNT_THREADS, 0, s>>>(d_terminals, L, p.num_tiles, (unsigned int*)(wsb + w.off_tcnt),
                                                            (unsigned int*)(wsb + w.off_cin), p.hdr);
        GGL_LAUNCH_CHECK();
        ggl::count_launch();
        if (norm) return clip ? launch<true, true, true>(p, grid, s) : launch<true, true, false>(p, grid, s);
        return launch<true, false, false>(p, grid, s);
    }
    if (norm) return clip ? launch<false, true, true>(p, grid, s) : launch<false, true, false>(p, grid, s);
    return launch<false, false, false>(p, grid, s);
}

// ---- allocation-free variant (no counterpart in the reference) ----
// The scan's scratch in a file-scope device array: every GPU that loads the library holds GGL_GAE_STATIC_SCRATCH_BYTES until
// exit. In HOST code the name `g_gae_static_scratch` is only the host-side stand-in of the symbol, not a device address: the
// address a kernel may use comes from cudaGetSymbolAddress, resolved once per device with that device current.
static __device__ __align__(256) unsigned char g_gae_static_scratch[GGL_GAE_STATIC_SCRATCH_BYTES];
namespace {
void* g_static_scratch_ptr[kMaxDevices];          // device address of the array on device i, null until resolved
std::mutex g_static_scratch_mutex;
}  // namespace

extern "C" int ggl_gae_compute_static(const float* d_rews, const int8_t* d_terminals, const float* d_val_preds,
                                      const float* d_trunc_val_preds, int64_t L, int64_t K,
                                      float gamma, float lambda, float return_std, float clip_range,
                                      float* d_out_advantages, float* d_out_returns, float* d_out_target_values,
                                      float* d_out_clip_portion, int32_t* d_out_trunc_count, void* stream) {
    if (L > (int64_t)GGL_GAE_STATIC_MAX_L || (L > 0 && ws_layout(L).total > (size_t)GGL_GAE_STATIC_SCRATCH_BYTES)) {
        ggl::set_error("ggl_gae_compute_static: L = %lld exceeds GGL_GAE_STATIC_MAX_L = %lld (static scratch of %d bytes); use "
                       "ggl_gae_compute with a caller workspace", (long long)L, (long long)GGL_GAE_STATIC_MAX_L, (int)GGL_GAE_STATIC_SCRATCH_BYTES);
        return GGL_ERR_INVALID;
    }
    int dev = 0;
    GGL_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) { ggl::set_error("ggl_gae_compute_static: device ordinal %d out of range", dev); return GGL_ERR_INVALID; }
    void* scratch = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_static_scratch_mutex);
        if (!g_static_scratch_ptr[dev]) {
            void* ptr = nullptr;
            GGL_CUDA(cudaGetSymbolAddress(&ptr, g_gae_static_scratch));
            if (!ptr) { ggl::set_error("ggl_gae_compute_static: cudaGetSymbolAddress returned a null address"); return GGL_ERR_CUDA; }
            g_static_scratch_ptr[dev] = ptr;
        }
        scratch = g_static_scratch_ptr[dev];
    }
    // ggl_gae_compute clears the part of the scratch that must start at zero (header + look-back records) on `stream` before
    // its kernels and allocates nothing itself (memset, attribute queries and launches only)
    return ggl_gae_compute(d_rews, d_terminals, d_val_preds, d_trunc_val_preds, L, K, gamma, lambda, return_std, clip_range,
                           d_out_advantages, d_out_returns, d_out_target_values, d_out_clip_portion, d_out_trunc_count, scratch, stream);
}

extern "C" int ggl_gae_compute_host(const float* h_rews, const int8_t* h_terminals, const float* h_val_preds,
                                    const float* h_trunc_val_preds, int64_t L, int64_t K,
                                    float gamma, float lambda, float return_std, float clip_range,
                                    float* h_out_advantages, float* h_out_returns, float* h_out_target_values,
                                    float* h_out_clip_portion) {
    if (L < 0 || K < 0) { ggl::set_error("ggl_gae_compute_host: negative size"); return GGL_ERR_INVALID; }
    if (L == 0) { if (h_out_clip_portion) *h_out_clip_portion = 0.f; return GGL_OK; }
    int dev = 0;
    GGL_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) { ggl::set_error("ggl_gae_compute_host: device ordinal %d not supported", dev); return GGL_ERR_INVALID; }
    std::lock_guard<std::mutex> lock(g_scratch_mutex);      // one host-buffer call per process at a time: the scratch is shared
    HostScratch& g_scratch = g_scratch_dev[dev];
    if (!g_scratch.stream) GGL_CUDA(cudaStreamCreateWithFlags(&g_scratch.stream, cudaStreamNonBlocking));
    cudaStream_t s = g_scratch.stream;
    const size_t nL = align_up((size_t)L * 4, 256), nT = align_up((size_t)L, 256), nK = align_up((size_t)(K ? K : 1) * 4, 256);
    const size_t ws = ggl_gae_workspace_bytes(L);
    const size_t need = 5 * nL + nT + nK + 256 + ws;
    if (need > g_scratch.bytes) {
        if (g_scratch.d) cudaFree(g_scratch.d);
        g_scratch.d = nullptr; g_scratch.bytes = 0;
        GGL_CUDA(cudaMalloc(&g_scratch.d, need));
        g_scratch.bytes = need;
    }
    unsigned char* b = (unsigned char*)g_scratch.d;
    float* d_r = (float*)b;               b += nL;
    float* d_v = (float*)b;               b += nL;
    float* d_a = (float*)b;               b += nL;
    float* d_ret = (float*)b;             b += nL;
    float* d_tg = (float*)b;              b += nL;
    int8_t* d_t = (int8_t*)b;             b += nT;
    float* d_tv = (float*)b;              b += nK;
    float* d_scal = (float*)b;            b += 256;     // [0] clip portion, [1] trunc count (int)
    void* d_ws = b;
    GGL_CUDA(cudaMemcpyAsync(d_r, h_rews, (size_t)L * 4, cudaMemcpyHostToDevice, s));
    GGL_CUDA(cudaMemcpyAsync(d_t, h_terminals, (size_t)L, cudaMemcpyHostToDevice, s));
    GGL_CUDA(cudaMemcpyAsync(d_v, h_val_preds, (size_t)L * 4, cudaMemcpyHostToDevice, s));
    const bool ht = h_trunc_val_preds != nullptr;
    if (ht && K > 0) GGL_CUDA(cudaMemcpyAsync(d_tv, h_trunc_val_preds, (size_t)K * 4, cudaMemcpyHostToDevice, s));
    int rc = ggl_gae_compute(d_r, d_t, d_v, ht ? d_tv : nullptr, L, K, gamma, lambda, return_std, clip_range,
                             d_a, d_ret, d_tg, d_scal, (int32_t*)(d_scal + 1), d_ws, s);
    if (rc != GGL_OK) return rc;
    GGL_CUDA(cudaMemcpyAsync(h_out_advantages, d_a, (size_t)L * 4, cudaMemcpyDeviceToHost, s));
    GGL_CUDA(cudaMemcpyAsync(h_out_returns, d_ret, (size_t)L * 4, cudaMemcpyDeviceToHost, s));
    GGL_CUDA(cudaMemcpyAsync(h_out_target_values, d_tg, (size_t)L * 4, cudaMemcpyDeviceToHost, s));
    float scal[2] = {0.f, 0.f};
    GGL_CUDA(cudaMemcpyAsync(scal, d_scal, 8, cudaMemcpyDeviceToHost, s));
    GGL_CUDA(cudaStreamSynchronize(s));
    if (h_out_clip_portion) *h_out_clip_portion = scal[0];
    if (ht) {   // GAE.cpp:196-197: truncation count mismatch is fatal in the reference
        int n; memcpy(&n, &scal[1], 4);
        if (n != (int)K) { ggl::set_error("GAE: truncation count mismatch (%d/%lld)", n, (long long)K); return GGL_ERR_INVALID; }
    }
    return GGL_OK;
}
