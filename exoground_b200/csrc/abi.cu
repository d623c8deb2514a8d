// abi.cu — process-wide plumbing of the C ABI: version, thread-local error text, launch counter.
#include "common.cuh"
#include <atomic>
#include <string.h>

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

void exg_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void exg_count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

extern "C" int exg_version(void) { return 100; }
extern "C" const char* exg_last_error(void) { return g_err; }
extern "C" long long exg_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }
