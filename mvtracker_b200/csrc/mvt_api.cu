// Error reporting and version of libmvt.so.
#include "mvt_common.cuh"
#include <stdarg.h>
#include <atomic>

namespace mvt {
static thread_local char g_err[512] = "";
void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
static std::atomic<unsigned long long> g_launches{0};
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
static unsigned long long *g_stats_ptr = nullptr;
unsigned long long *debug_stats_ptr() { return g_stats_ptr; }
void set_debug_stats_ptr(unsigned long long *p) { g_stats_ptr = p; }
}  // namespace mvt

MVT_EXPORT int mvt_version(void) { return MVT_VERSION; }
MVT_EXPORT const char *mvt_last_error(void) { return mvt::g_err; }
MVT_EXPORT unsigned long long mvt_launch_count(void) { return mvt::g_launches.load(); }
MVT_EXPORT int mvt_debug_set_stats(void *dev_counters) { mvt::set_debug_stats_ptr((unsigned long long *)dev_counters); return MVT_OK; }
