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
// tile's predecessors are always already running), each tile reads its 9 B/element once through coalesced
// streaming loads into padded shared memory, scans with warp shuffles, publishes its aggregate, resolves its
// carry-in from predecessor tiles, and writes 12 B/element once.  Algorithmic traffic = 21 B/element — the
// HBM roofline of this path.  The truncation rank is obtained in the same pass as
// K - (suffix count of TRUNCATED entries), a second (integer) chain with single-word status records.
//
// Compiled with -fmad=false: per-element arithmetic then rounds exactly like the reference's scalar loop; only
// the cross-thread carry (composition order) differs, and its effect decays by gamma*lambda per step.
#include "common.cuh"

namespace {

constexpr int GAE_THREADS = 256;
constexpr int GAE_ITEMS = 16;
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
    unsigned long long* cnt_status;   // [tiles]  status<<62 | count
    unsigned int* val_flag;           // [tiles]  0 none, 1 aggregate, 2 inclusive
    float4* val_agg;                  // [tiles]
    float4* val_inc;                  // [tiles]
    int num_tiles;
};

__device__ __forceinline__ int pidx(int i) { return i + (i >> 5); }

__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_u32(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_u32(unsigned int* p, unsigned int v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

constexpr unsigned long long ST_AGG = 1ull << 62;
constexpr unsigned long long ST_INC = 2ull << 62;

template <bool HAS_TRUNC, bool NORM, bool CLIP>
__global__ void __launch_bounds__(GAE_THREADS, 4) gae_scan_kernel(GaeParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* s_r = reinterpret_cast<float*>(smem_raw);                 // rewards -> advantages
    float* s_v = s_r + (GAE_TILE + GAE_PAD + 4);                     // values (+1 lookahead) -> returns
    float* s_o = s_v + (GAE_TILE + GAE_PAD + 4);                     // b_adv -> target values
    uint4* s_t = reinterpret_cast<uint4*>(s_o + (GAE_TILE + GAE_PAD + 4));   // terminals, 16 per thread

    __shared__ int s_tile;
    __shared__ Lin2 s_warp[GAE_WARPS];
    __shared__ int s_wcnt[GAE_WARPS];
    __shared__ int s_cnt_in;
    __shared__ float s_x_in[2];
    __shared__ double s_red[2][GAE_WARPS];

    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int warp = tid >> 5;
    const float g = p.gamma, gl = p.gamma_lambda;

    for (;;) {
        if (tid == 0) s_tile = (int)atomicAdd(&p.hdr->tile_ticket, 1u);
        __syncthreads();
        const int t = s_tile;
        if (t >= p.num_tiles) break;
        const long long base = (long long)(p.num_tiles - 1 - t) * GAE_TILE;

        // ---- coalesced streaming loads into padded shared memory ----
#pragma unroll
        for (int c = 0; c < GAE_ITEMS / 4; c++) {
            const int li = (c * GAE_THREADS + tid) * 4;
            const long long gi = base + li;
            float4 r4, v4;
            if (gi + 3 < p.L) {
                r4 = __ldcs(reinterpret_cast<const float4*>(p.rews + gi));
                v4 = __ldcs(reinterpret_cast<const float4*>(p.vals + gi));
            } else {
                float rr[4], vv[4];
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    const bool ok = gi + e < p.L;
                    rr[e] = ok ? p.rews[gi + e] : 0.f;
                    vv[e] = ok ? p.vals[gi + e] : 0.f;
                }
                r4 = make_float4(rr[0], rr[1], rr[2], rr[3]);
                v4 = make_float4(vv[0], vv[1], vv[2], vv[3]);
            }
            const int pi = pidx(li);
            s_r[pi] = r4.x; s_r[pi + 1] = r4.y; s_r[pi + 2] = r4.z; s_r[pi + 3] = r4.w;
            s_v[pi] = v4.x; s_v[pi + 1] = v4.y; s_v[pi + 2] = v4.z; s_v[pi + 3] = v4.w;
        }
        {
            const long long gi = base + (long long)tid * 16;
            uint4 tw;
            if (gi + 15 < p.L) {
                tw = __ldcs(reinterpret_cast<const uint4*>(p.terms + gi));
            } else {
                unsigned int w[4] = {0, 0, 0, 0};
                for (int e = 0; e < 16; e++)
                    if (gi + e < p.L) w[e >> 2] |= (unsigned int)(unsigned char)p.terms[gi + e] << (8 * (e & 3));
                tw = make_uint4(w[0], w[1], w[2], w[3]);
            }
            s_t[tid] = tw;
        }
        if (tid == 0) {
            const long long gi = base + GAE_TILE;
            s_v[pidx(GAE_TILE)] = (gi < p.L) ? p.vals[gi] : 0.f;
        }
        __syncthreads();

        // every thread owns GAE_ITEMS consecutive elements; the HIGHEST thread id comes first in scan order
        const uint4 tw = s_t[tid];
        const unsigned int tword[4] = {tw.x, tw.y, tw.z, tw.w};

        // ---- chain 1 (integer): suffix count of TRUNCATED entries -> truncValPreds rank ----
        int cnt_before = 0;   // TRUNCATED entries at higher stream indices than this thread's span
        if (HAS_TRUNC) {
            int c_loc = 0;
#pragma unroll
            for (int j = 0; j < 16; j++) c_loc += (((tword[j >> 2] >> (8 * (j & 3))) & 0xffu) == 2u);
            int c = c_loc;
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                const int o = __shfl_down_sync(0xffffffffu, c, off);
                if (lane + off < 32) c += o;
            }
            if (lane == 0) s_wcnt[warp] = c;
            __syncthreads();
            int wcarry = 0, total = 0;
#pragma unroll
            for (int w = 0; w < GAE_WARPS; w++) {
                total += s_wcnt[w];
                if (w > warp) wcarry += s_wcnt[w];
            }
            if (warp == 0) {
                int acc = 0;
                if (t > 0) {
                    if (lane == 0) st_relaxed_u64(&p.cnt_status[t], ST_AGG | (unsigned long long)(unsigned int)total);
                    int win = t - 1;
                    for (;;) {
                        const int q = win - lane;
                        unsigned long long s = (q >= 0) ? ld_relaxed_u64(&p.cnt_status[q]) : ST_INC;
                        const unsigned int st = (unsigned int)(s >> 62);
                        const unsigned int inc_mask = __ballot_sync(0xffffffffu, st == 2u);
                        const unsigned int none_mask = __ballot_sync(0xffffffffu, st == 0u);
                        const int first_inc = inc_mask ? (__ffs(inc_mask) - 1) : 32;
                        const unsigned int relevant = (first_inc >= 31) ? 0xffffffffu : ((2u << first_inc) - 1u);
                        if (none_mask & relevant) continue;
                        int contrib = (lane <= first_inc) ? (int)(unsigned int)(s & 0xffffffffull) : 0;
#pragma unroll
                        for (int off = 16; off > 0; off >>= 1) contrib += __shfl_xor_sync(0xffffffffu, contrib, off);
                        acc += contrib;
                        if (first_inc < 32) break;
                        win -= 32;
                    }
                }
                if (lane == 0) {
                    st_relaxed_u64(&p.cnt_status[t], ST_INC | (unsigned long long)(unsigned int)(acc + total));
                    s_cnt_in = acc;
                }
            }
            __syncthreads();
            cnt_before = s_cnt_in + wcarry + (c - c_loc);
        }

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
        if (warp == 0) {
            Lin2 acc = lin_identity();
            bool resolved = (t == 0);
            if (t > 0) {
                if (lane == 0) {
                    __stcg(&p.val_agg[t], make_float4(tile_agg.Aa, tile_agg.Ba, tile_agg.Ar, tile_agg.Br));
                    st_release_u32(&p.val_flag[t], 1u);
                }
                int win = t - 1;
                for (;;) {
                    const int q = win - lane;
                    const unsigned int f = (q >= 0) ? ld_acquire_u32(&p.val_flag[q]) : 2u;
                    const unsigned int inc_mask = __ballot_sync(0xffffffffu, f == 2u);
                    const unsigned int none_mask = __ballot_sync(0xffffffffu, f == 0u);
                    const int first_inc = inc_mask ? (__ffs(inc_mask) - 1) : 32;
                    const unsigned int relevant = (first_inc >= 31) ? 0xffffffffu : ((2u << first_inc) - 1u);
                    if (none_mask & relevant) continue;
                    Lin2 c = lin_identity();
                    if (lane <= first_inc) {
                        if (q < 0) {
                            c = Lin2{0.f, 0.f, 0.f, 0.f};
                        } else {
                            const float4 d = __ldcg((f == 2u) ? &p.val_inc[q] : &p.val_agg[q]);
                            c = Lin2{d.x, d.y, d.z, d.w};
                        }
                    }
#pragma unroll
                    for (int off = 1; off < 32; off <<= 1) {
                        const Lin2 o = shfl_down_lin(c, off);
                        if (lane + off < 32) c = compose(c, o);
                    }
                    acc = compose(acc, c);   // valid on lane 0
                    if (first_inc < 32) { resolved = true; break; }
                    win -= 32;
                }
            }
            if (lane == 0) {
                const float adv_in = (t == 0) ? 0.f : acc.Ba;
                const float ret_in = (t == 0) ? 0.f : acc.Br;
                const float adv_out = tile_agg.Aa * adv_in + tile_agg.Ba;
                const float ret_out = tile_agg.Ar * ret_in + tile_agg.Br;
                __stcg(&p.val_inc[t], make_float4(0.f, adv_out, 0.f, ret_out));
                st_release_u32(&p.val_flag[t], 2u);
                s_x_in[0] = adv_in;
                s_x_in[1] = ret_in;
            }
            (void)resolved;
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
            __syncthreads();
            if (tid == 0) {
                double ta = 0, tb = 0;
                for (int w = 0; w < GAE_WARPS; w++) { ta += s_red[0][w]; tb += s_red[1][w]; }
                atomicAdd(&p.hdr->sum_abs, ta);
                atomicAdd(&p.hdr->sum_abs_clipped, tb);
            }
        }
        __syncthreads();   // shared tiles are reused by the next ticket
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
                if (HAS_TRUNC && p.num_tiles > 0)
                    n = (int)(unsigned int)(ld_relaxed_u64(&p.cnt_status[p.num_tiles - 1]) & 0xffffffffull);
                *p.out_trunc_count = n;
            }
        }
    }
}

constexpr size_t gae_smem_bytes() { return 3 * (size_t)(GAE_TILE + GAE_PAD + 4) * sizeof(float) + GAE_TILE; }

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct WsLayout {
    size_t off_cnt, off_flag, off_agg, off_inc, total;
};
WsLayout ws_layout(long long L) {
    const size_t tiles = (size_t)((L + GAE_TILE - 1) / GAE_TILE);
    WsLayout w;
    size_t o = align_up(sizeof(GaeHeader), 16);
    w.off_cnt = o;  o += align_up(tiles * sizeof(unsigned long long), 16);
    w.off_flag = o; o += align_up(tiles * sizeof(unsigned int), 16);
    w.off_agg = o;  o += tiles * sizeof(float4);
    w.off_inc = o;  o += tiles * sizeof(float4);
    w.total = align_up(o, 256);
    return w;
}

template <bool HT, bool NM, bool CL>
int launch(const GaeParams& p, int grid, cudaStream_t s) {
    static bool attr_done = false;
    if (!attr_done) {
        GGL_CUDA(cudaFuncSetAttribute(gae_scan_kernel<HT, NM, CL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gae_smem_bytes()));
        attr_done = true;
    }
    gae_scan_kernel<HT, NM, CL><<<grid, GAE_THREADS, gae_smem_bytes(), s>>>(p);
    GGL_LAUNCH_CHECK();
    ggl::count_launch();
    return GGL_OK;
}

// cached per-device scratch for the host-buffer wrapper
struct HostScratch {
    void* d = nullptr;
    size_t bytes = 0;
    cudaStream_t stream = nullptr;
} g_scratch;

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
    GGL_CUDA(cudaMemsetAsync(d_workspace, 0, w.off_agg, s));   // header + status words + flags
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
    p.cnt_status = (unsigned long long*)(wsb + w.off_cnt);
    p.val_flag = (unsigned int*)(wsb + w.off_flag);
    p.val_agg = (float4*)(wsb + w.off_agg);
    p.val_inc = (float4*)(wsb + w.off_inc);
    p.num_tiles = (int)((L + GAE_TILE - 1) / GAE_TILE);

    int dev = 0, sms = 148;
    GGL_CUDA(cudaGetDevice(&dev));
    GGL_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int grid = p.num_tiles < sms * 4 ? p.num_tiles : sms * 4;   // persistent: 4 resident CTAs per SM

    const bool ht = d_trunc_val_preds != nullptr;
    if (ht) {
        if (norm) return clip ? launch<true, true, true>(p, grid, s) : launch<true, true, false>(p, grid, s);
        return launch<true, false, false>(p, grid, s);
    }
    if (norm) return clip ? launch<false, true, true>(p, grid, s) : launch<false, true, false>(p, grid, s);
    return launch<false, false, false>(p, grid, s);
}

extern "C" int ggl_gae_compute_host(const float* h_rews, const int8_t* h_terminals, const float* h_val_preds,
                                    const float* h_trunc_val_preds, int64_t L, int64_t K,
                                    float gamma, float lambda, float return_std, float clip_range,
                                    float* h_out_advantages, float* h_out_returns, float* h_out_target_values,
                                    float* h_out_clip_portion) {
    if (L < 0 || K < 0) { ggl::set_error("ggl_gae_compute_host: negative size"); return GGL_ERR_INVALID; }
    if (L == 0) { if (h_out_clip_portion) *h_out_clip_portion = 0.f; return GGL_OK; }
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
