// Profiling hooks of the tick (compiled out of the product library).
#pragma once
#if defined(__CUDACC__)
#include <cuda_runtime.h>
#endif
namespace ggl {
// Phase timing (profiling build only, -DGGL_PHASE_TIMING -> libgigaenv_cuda_prof.so): the leader lane of every arena adds
// the SM cycles it spent in each tick phase (including the time its warp ran other arenas' divergent paths) to a global
// table read back with ggl_debug_phase_cycles. Compiled out of the product library.
#if defined(GGL_PHASE_TIMING) && defined(__CUDACC__)
__device__ unsigned long long g_phase_cycles[16];
__device__ unsigned long long g_sub_cycles[16];                   // lane-summed sub-phase cycles inside collide_body
__device__ unsigned long long g_arena_phase_cycles[8192][16];     // per-arena breakdown of the first 8192 arenas
#endif
#if defined(GGL_PHASE_TIMING) && defined(__CUDA_ARCH__)
#define GGL_SUB_BEGIN() long long _sub_t = clock64()
#define GGL_SUB(k) do { long long _n = clock64(); atomicAdd(&g_sub_cycles[k], (unsigned long long)(_n - _sub_t)); _sub_t = _n; } while (0)
#define GGL_PH_BEGIN() long long _ph_t = clock64()
#define GGL_PH_RESET() _ph_t = clock64()
#define GGL_PH(k) do { long long _n = clock64(); if (ex.lane() == 0) { atomicAdd(&g_phase_cycles[k], (unsigned long long)(_n - _ph_t)); \
        if (ex.arena_ < 8192) g_arena_phase_cycles[ex.arena_][k] += (unsigned long long)(_n - _ph_t); } _ph_t = _n; } while (0)
#else
#define GGL_SUB_BEGIN() do {} while (0)
#define GGL_SUB(k) do {} while (0)
#define GGL_PH_BEGIN() do {} while (0)
#define GGL_PH_RESET() do {} while (0)
#define GGL_PH(k) do {} while (0)
#endif

}  // namespace ggl
