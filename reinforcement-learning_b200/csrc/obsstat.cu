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
    float* inv; float* neg;             // coefficients of the last normalise call
};

// WelfordStat.h:89-97 for every sampled row, in order
__global__ void obsstat_increment_kernel(StatDev S, const float* __restrict__ obs, const int32_t* __restrict__ idx, int n) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const long long count0 = *S.count;
    if (j < S.width) {
        double mean = S.means[j], var = S.vars[j];
        for (int k = 0; k < n; k++) {
            const long long count = count0 + k;
            const double sample = (double)obs[(long long)idx[k] * S.width + j];
            const double delta = sample - mean;
            const double deltaN = delta / (double)(count + 1);
            mean += deltaN;
            var += delta * deltaN * (double)count;
        }
        S.means[j] = mean; S.vars[j] = var;
    }
}
__global__ void obsstat_bump_count_kernel(StatDev S, int n) { *S.count += n; }

// WelfordStat.h:107-127 (GetSTD) and :135-160 (clamps and float coefficients)
__global__ void obsstat_coeff_kernel(StatDev S, double meanClamp, double minSTD) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const long long count = *S.count;
    if (j >= S.width || count < 2) return;
    const double var = S.vars[j] / (double)(count - 1);
    const double sd = (var > 0) ? sqrt(var) : 1.0;
    const double m = S.means[j];
    const double cm = m < -meanClamp ? -meanClamp : (m > meanClamp ? meanClamp : m);       // RS_CLAMP
    const double cs = sd > minSTD ? sd : minSTD;                                           // RS_MAX
    const float inv = 1.0f / (float)cs;
    S.inv[j] = inv;
    S.neg[j] = -(float)cm * inv;
}

// :163-216 — every element once; 16 bytes per thread
__global__ void __launch_bounds__(256) obsstat_apply_kernel(StatDev S, float* __restrict__ obs, long long total) {
    extern __shared__ float coef[];                      // inv[width], neg[width]
    if (*S.count < 2) return;                            // :133 — not enough data: the obs stay as they are, bit for bit
    for (int j = threadIdx.x; j < S.width; j += blockDim.x) { coef[j] = S.inv[j]; coef[S.width + j] = S.neg[j]; }
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
    s->S.inv = reinterpret_cast<float*>(s->S.count + 1);
    s->S.neg = s->S.inv + width;
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
    obsstat_increment_kernel<<<(s->S.width + 63) / 64, 64, 0, (cudaStream_t)stream>>>(s->S, d_obs, d_row_indices, n);
    GGL_LAUNCH_CHECK();
    obsstat_bump_count_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(s->S, n);
    GGL_LAUNCH_CHECK(); ggl::count_launch(2);
    return GGL_OK;
}

extern "C" int ggl_obsstat_normalize(ggl_obsstat* s, float* d_obs, int64_t num_rows, double mean_clamp, double min_std, void* stream) {
    STAT_CHECK(s);
    if (num_rows <= 0) return GGL_OK;                                                     // :133
    if (!d_obs || ((uintptr_t)d_obs & 15)) { ggl::set_error("ggl_obsstat_normalize: obs must be a 16-byte aligned device pointer"); return GGL_ERR_INVALID; }
    const long long total = (long long)num_rows * s->S.width;
    obsstat_coeff_kernel<<<(s->S.width + 63) / 64, 64, 0, (cudaStream_t)stream>>>(s->S, mean_clamp, min_std);
    GGL_LAUNCH_CHECK();
    long long blocks = ((total >> 2) + 255) / 256;
    if (blocks > (long long)s->sms * 8) blocks = (long long)s->sms * 8;
    if (blocks < 1) blocks = 1;
    obsstat_apply_kernel<<<(int)blocks, 256, sizeof(float) * 2 * s->S.width, (cudaStream_t)stream>>>(s->S, d_obs, total);
    GGL_LAUNCH_CHECK(); ggl::count_launch(2);
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
