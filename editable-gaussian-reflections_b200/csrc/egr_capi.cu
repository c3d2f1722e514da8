// Library-level entry points of the C ABI (include/egr_b200.h): error message, version and the compile-time
// constants the reference exports as static methods (csrc/raytracer.cpp:145-148, csrc/flags.h:5-17).
#include <stdarg.h>

#include "egr_common.cuh"

namespace egr {
static thread_local char g_err[512] = "";
void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
}  // namespace egr

extern "C" int egr_abi_version(void) { return EGR_ABI_VERSION; }
extern "C" const char *egr_last_error(void) { return egr::g_err; }
extern "C" double egr_max_alpha(void) { return (double)egr::A_MAX; }
extern "C" int egr_roughness_downweight_grad(void) { return 1; }
extern "C" double egr_roughness_downweight_grad_power(void) { return (double)egr::ROUGH_DOWNWEIGHT_POW; }
extern "C" int egr_buffer_size(void) { return egr::KBUF; }
extern "C" int egr_max_iterations(void) { return egr::MAX_ITERATIONS; }
