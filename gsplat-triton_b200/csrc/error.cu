// Error string + launch counter shared by all translation units of libgsplat_b200.so.
#include "common.cuh"

#include <atomic>
#include <cstdarg>
#include <cstdio>

namespace {
thread_local char g_error[512] = "";
std::atomic<long long> g_launches{0};
}  // namespace

void gsb_set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_error, sizeof(g_error), fmt, ap);
    va_end(ap);
}

void gsb_count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

extern "C" const char *gsb_last_error(void) { return g_error; }
extern "C" int gsb_abi_version(void) { return 2; }
extern "C" int64_t gsb_launch_count(void) { return (int64_t)g_launches.load(std::memory_order_relaxed); }
extern "C" void gsb_reset_launch_count(void) { g_launches.store(0, std::memory_order_relaxed); }
