// Library-wide plumbing of the C ABI: thread-local error string, version, launch counter.
#include "common.cuh"
#include "../../include/mdx.h"
#include <atomic>
#include <stdarg.h>

std::atomic<long long> g_mdx_launches{0};

static thread_local char g_err[512] = "";

void mdx_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" const char* mdx_last_error(void) { return g_err; }
extern "C" int mdx_version(void) { return 100; }
extern "C" int64_t mdx_launch_count(void) { return g_mdx_launches.load(std::memory_order_relaxed); }
