// Error string + launch counter shared by all translation units of libgsplat_b200.so.
#include "common.cuh"

#include <atomic>
#include <cstdarg>
#include <cstdio>

namespace {
thread_local char g_error[512] = "";
std::atomic<long long> g_launches{0};
// kernel-variant switches for A/B measurements (gsb_set_option); the defaults are the tuned path
std::atomic<int> g_options[GSB_OPT_COUNT] = {{2}, {2}, {0}, {2}, {0}, {0}, {2}, {0}, {37}};
}  // namespace

int gsb_option(int which) { return g_options[which].load(std::memory_order_relaxed); }

void gsb_set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_error, sizeof(g_error), fmt, ap);
    va_end(ap);
}

void gsb_count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

extern "C" const char *gsb_last_error(void) { return g_error; }
extern "C" int gsb_abi_version(void) { return 3; }
extern "C" int gsb_set_option(int which, int value) {
    if (which < 0 || which >= GSB_OPT_COUNT) return GSB_ERR_INVALID;
    g_options[which].store(value, std::memory_order_relaxed);
    return GSB_OK;
}
extern "C" int gsb_get_option(int which) {
    return (which < 0 || which >= GSB_OPT_COUNT) ? -1 : g_options[which].load(std::memory_order_relaxed);
}
extern "C" int64_t gsb_launch_count(void) { return (int64_t)g_launches.load(std::memory_order_relaxed); }
extern "C" void gsb_reset_launch_count(void) { g_launches.store(0, std::memory_order_relaxed); }
