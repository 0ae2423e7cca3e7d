// Library-level state: last error (thread-local), launch counter, version.
#include <stdarg.h>

#include <atomic>

#include "common.cuh"

namespace iit {
static thread_local char g_err[512] = "";
static std::atomic<unsigned long long> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add((unsigned long long)n, std::memory_order_relaxed); }
}  // namespace iit

extern "C" const char* iit_last_error(void) { return iit::g_err; }
extern "C" int iit_version(void) { return 100; }
extern "C" unsigned long long iit_launch_count(void) { return iit::g_launches.load(); }
