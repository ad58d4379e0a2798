// Running obs statistics and in-place obs normalisation on the device (sm_100a; built with -fmad=false).
//
// Replaces GGL::BatchedWelfordStat (GigaLearnCPP/src/private/GigaLearnCPP/Util/WelfordStat.h:69-243) as the Learner uses it
// every collection step (Learner.cpp:680-694): IncrementRow over up to maxObsSamples randomly chosen obs rows, then
// NormalizeInPlace over all rows — on the host that is a pass over numPlayers x obsSize floats between the env step and the
// policy; here the obs never leave the GPU.
//   * increment: Welford in double, one thread per column walking the sampled rows IN ORDER (columns are independent and the
//     update is order-dependent, so this is the reference's arithmetic exactly: same IEEE double operations in the same order).
//     The row choice is an input (Math::RandInt lives on the host).
//   * normalise: coefficients as in :131-160 (clamped mean, std floored at minSTD, inv = 1.0f / (float)std,
//     negMeanDivSTD = -(float)mean * inv), then row[j] = row[j] * inv[j] + neg[j] as a separate multiply and add (the x86-64
//     build of the reference has no FMA contraction). HBM-bound: 8 B per obs element (read + write).
#include "common.cuh"

namespace {

struct StatDev {
    int width;
    double* means; double* vars; long long* count;
};

// WelfordStat.h:89-97 for every sampled row, in order
__global__ void obsstat_increment_kernel(StatDev S, const float* __restrict__ obs, const int32_t* __restrict__ idx, int n) {
    const long long count0 = *S.count;
    __syncthreads();                                       // everyone has read the count before thread 0 moves it
    if (threadIdx.x == 0) *S.count = count0 + n;           // count++ per row (:96)
    // row offsets once into shared memory (one round trip instead of one per sample)
    extern __shared__ long long rowOff[];
    for (int k = threadIdx.x; k < n; k += blockDim.x) rowOff[k] = (long long)__ldg(idx + k) * S.width;
    __syncthreads();
    for (int j = threadIdx.x; j < S.width; j += blockDim.x) {
        double mean = S.means[j], var = S.vars[j];
        // the samples of 16 rows are fetched together and one chunk ahead (independent loads), then folded in one after the
        // other: the update chain is serial by definition, the HBM round trips need not be
        float x[16], xn[16];
#pragma unroll
        for (int u = 0; u < 16; u++) x[u] = (u < n) ? __ldg(obs + rowOff[u] + j) : 0.f;
        for (int k0 = 0; k0 < n; k0 += 16) {
#pragma unroll
            for (int u = 0; u < 16; u++) xn[u] = (k0 + 16 + u < n) ? __ldg(obs + rowOff[k0 + 16 + u] + j) : 0.f;
#pragma unroll
            for (int u = 0; u < 16; u++) {
                if (k0 + u < n) {
                    const long long count = count0 + k0 + u;
                    const double delta = (double)x[u] - mean;
                    const double deltaN = delta / (double)(count + 1);
                    mean += deltaN;
                    var += delta * deltaN * (double)count;
                }
            }
#pragma unroll
            for (int u = 0; u < 16; u++) x[u] = xn[u];
        }
        S.means[j] = mean; S.vars[j] = var;
    }
}

// WelfordStat.h:107-127 (GetSTD) and :135-160 (clamps and float coefficients) for column j
__device__ __forceinline__ void obsstat_coeff(const StatDev& S, int j, long long count, double meanClamp, double minSTD, float& inv, float& neg) {
    const double var = S.vars[j] / (double)(count - 1);
    const double sd = (var > 0) ? sqrt(var) : 1.0;
    const double m = S.means[j];
    const double cm = m < -meanClamp ? -meanClamp : (m > meanClamp ? meanClamp : m);       // RS_CLAMP
    const double cs = sd > minSTD ? sd : minSTD;                                           // RS_MAX
    inv = 1.0f / (float)cs;
    neg = -(float)cm * inv;
}

// :163-216 — every element once; 16 bytes per thread
__global__ void __launch_bounds__(256) obsstat_apply_kernel(StatDev S, float* __restrict__ obs, long long total, double meanClamp, double minSTD) {
    extern __shared__ float coef[];                      // inv[width], neg[width]: every block derives them itself (a few hundred sqrt)
    const long long count = *S.count;
    if (count < 2) return;                               // :133 — not enough data: the obs stay as they are, bit for bit
    for (int j = threadIdx.x; j < S.width; j += blockDim.x) obsstat_coeff(S, j, count, meanClamp, minSTD, coef[j], coef[S.width + j]);
    __syncthreads();
    const float* inv = coef; const float* neg = coef + S.width;
    const long long nvec = total >> 2;
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < nvec; q += (long long)gridDim.x * blockDim.x) {
        const long long e = q * 4;
        int col = (int)(e % S.width);
        float4 v = __ldcs(reinterpret_cast<const float4*>(obs) + q);
        float* c = reinterpret_cast<float*>(&v);
#pragma unroll
        for (int k = 0; k < 4; k++) {
            c[k] = c[k] * inv[col] + neg[col];
            if (++col == S.width) col = 0;
        }
        __stcs(reinterpret_cast<float4*>(obs) + q, v);
    }
    if (blockIdx.x == 0 && threadIdx.x < (int)(total & 3)) {          // tail of a stream whose length is not a multiple of 4
        const long long e = (nvec << 2) + threadIdx.x;
        const int col = (int)(e % S.width);
        obs[e] = obs[e] * inv[col] + neg[col];
    }
}

}  // namespace

struct ggl_obsstat {
    StatDev S;
    int device, sms;
};

extern "C" int ggl_obsstat_create(int device, int width, ggl_obsstat** out) {
    if (!out || width <= 0) { ggl::set_error("ggl_obsstat_create: bad argument"); return GGL_ERR_INVALID; }
    GGL_CUDA(cudaSetDevice(device));
    ggl_obsstat* s = new ggl_obsstat();
    s->device = device; s->sms = 148;
    cudaDeviceGetAttribute(&s->sms, cudaDevAttrMultiProcessorCount, device);
    s->S.width = width;
    void* blob = nullptr;
    const size_t bytes = sizeof(double) * 2 * width + sizeof(long long) + sizeof(float) * 2 * width + 64;
    cudaError_t e = cudaMalloc(&blob, bytes);
    if (e != cudaSuccess) { ggl::set_error("ggl_obsstat_create: cudaMalloc -> %s", cudaGetErrorString(e)); delete s; return GGL_ERR_NOMEM; }
    cudaMemset(blob, 0, bytes);
    s->S.means = static_cast<double*>(blob);
    s->S.vars = s->S.means + width;
    s->S.count = reinterpret_cast<long long*>(s->S.vars + width);
    GGL_CUDA(cudaDeviceSynchronize());
    *out = s;
    return GGL_OK;
}

extern "C" int ggl_obsstat_destroy(ggl_obsstat* s) {
    if (!s) return GGL_OK;
    cudaSetDevice(s->device);
    cudaDeviceSynchronize();
    cudaFree(s->S.means);
    delete s;
    return GGL_OK;
}

#define STAT_CHECK(s) if (!(s)) { ggl::set_error("null obs statistics"); return GGL_ERR_INVALID; }

extern "C" int ggl_obsstat_increment_rows(ggl_obsstat* s, const float* d_obs, const int32_t* d_row_indices, int32_t n, void* stream) {
    STAT_CHECK(s);
    if (n < 0 || (n > 0 && (!d_obs || !d_row_indices))) { ggl::set_error("ggl_obsstat_increment_rows: bad argument"); return GGL_ERR_INVALID; }
    if (n == 0) return GGL_OK;
    const int threads = s->S.width >= 1024 ? 1024 : ((s->S.width + 31) / 32) * 32;
    if ((size_t)n * sizeof(long long) > 200 * 1024) { ggl::set_error("ggl_obsstat_increment_rows: at most 25600 rows per call"); return GGL_ERR_INVALID; }
    if ((size_t)n * sizeof(long long) > 48 * 1024) GGL_CUDA(cudaFuncSetAttribute(obsstat_increment_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    obsstat_increment_kernel<<<1, threads, (size_t)n * sizeof(long long), (cudaStream_t)stream>>>(s->S, d_obs, d_row_indices, n);
    GGL_LAUNCH_CHECK(); ggl::count_launch();
    return GGL_OK;
}

extern "C" int ggl_obsstat_normalize(ggl_obsstat* s, float* d_obs, int64_t num_rows, double mean_clamp, double min_std, void* stream) {
    STAT_CHECK(s);
    if (num_rows <= 0) return GGL_OK;                                                     // :133
    if (!d_obs || ((uintptr_t)d_obs & 15)) { ggl::set_error("ggl_obsstat_normalize: obs must be a 16-byte aligned device pointer"); return GGL_ERR_INVALID; }
    const long long total = (long long)num_rows * s->S.width;
    long long blocks = ((total >> 2) + 255) / 256;
    if (blocks > (long long)s->sms * 8) blocks = (long long)s->sms * 8;
    if (blocks < 1) blocks = 1;
    obsstat_apply_kernel<<<(int)blocks, 256, sizeof(float) * 2 * s->S.width, (cudaStream_t)stream>>>(s->S, d_obs, total, mean_clamp, min_std);
    GGL_LAUNCH_CHECK(); ggl::count_launch();
    return GGL_OK;
}

extern "C" int ggl_obsstat_get(ggl_obsstat* s, double* h_means, double* h_vars, int64_t* h_count) {
    STAT_CHECK(s);
    long long c = 0;
    if (h_means) GGL_CUDA(cudaMemcpy(h_means, s->S.means, sizeof(double) * s->S.width, cudaMemcpyDeviceToHost));
    if (h_vars) GGL_CUDA(cudaMemcpy(h_vars, s->S.vars, sizeof(double) * s->S.width, cudaMemcpyDeviceToHost));
    GGL_CUDA(cudaMemcpy(&c, s->S.count, sizeof(long long), cudaMemcpyDeviceToHost));
    if (h_count) *h_count = c;
    return GGL_OK;
}

extern "C" int ggl_obsstat_set(ggl_obsstat* s, const double* h_means, const double* h_vars, int64_t count) {
    STAT_CHECK(s);
    if (!h_means || !h_vars || count < 0) { ggl::set_error("ggl_obsstat_set: bad argument"); return GGL_ERR_INVALID; }
    const long long c = count;
    GGL_CUDA(cudaMemcpy(s->S.means, h_means, sizeof(double) * s->S.width, cudaMemcpyHostToDevice));
    GGL_CUDA(cudaMemcpy(s->S.vars, h_vars, sizeof(double) * s->S.width, cudaMemcpyHostToDevice));
    GGL_CUDA(cudaMemcpy(s->S.count, &c, sizeof(long long), cudaMemcpyHostToDevice));
    return GGL_OK;
}
