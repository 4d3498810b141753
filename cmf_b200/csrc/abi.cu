// Version + error-string plumbing of the C ABI.
#include <stdarg.h>
#include <string.h>

#include <atomic>

#include "common.cuh"

namespace cmf {
static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
static std::atomic<unsigned long long> g_launches{0};
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
}  // namespace cmf

extern "C" int cmf_version(void) { return CMF_B200_ABI_VERSION; }
extern "C" const char* cmf_last_error_string(void) { return cmf::g_err; }
extern "C" unsigned long long cmf_launch_count(void) { return cmf::g_launches.load(); }
