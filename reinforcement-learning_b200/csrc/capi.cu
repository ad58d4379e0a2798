// Process-wide pieces of the C ABI: last-error text, launch counter, version.
#include "common.cuh"
#include <stdarg.h>
#include <string.h>

namespace ggl {
static thread_local char g_err[512] = "";
std::atomic<uint64_t> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
}  // namespace ggl

extern "C" const char* ggl_last_error(void) { return ggl::g_err; }
extern "C" uint64_t ggl_launch_count(void) { return ggl::g_launches.load(); }
extern "C" const char* ggl_version(void) { return "gigaenv-b200 0.1 (sm_100a)"; }
