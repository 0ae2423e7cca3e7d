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
static std::atomic<unsigned long long> g_variants[KV_COUNT];
void count_variant(int id) { if (id >= 0 && id < KV_COUNT) g_variants[id].fetch_add(1ull, std::memory_order_relaxed); }
void count_launch(int n) { g_launches.fetch_add((unsigned long long)n, std::memory_order_relaxed); }
}  // namespace iit

extern "C" const char* iit_last_error(void) { return iit::g_err; }
extern "C" int iit_version(void) { return 100; }
extern "C" unsigned long long iit_launch_count(void) { return iit::g_launches.load(); }

static const char* const kVariantNames[iit::KV_COUNT] = {
    "conv3x3_tc_kernel<64,7,1>", "conv3x3_tc_kernel<128,4,1>", "conv3x3_tc_kernel<128,4,2>", "conv3x3_tc_kernel<256,3,1>",
    "conv3x3_tc2_kernel<256,6>", "conv3x3_tc2_kernel<128,8>", "conv_win_kernel<128,9,4>", "conv_win_kernel<32,16,4>",
    "wgrad3x3_tc_kernel", "wgrad_win_kernel", "stem_wgrad_win_kernel"};
extern "C" int iit_kernel_variant_slots(void) { return iit::KV_COUNT; }
extern "C" const char* iit_kernel_variant_name(int id) { return (id >= 0 && id < iit::KV_COUNT) ? kVariantNames[id] : ""; }
extern "C" unsigned long long iit_kernel_variant_count(int id) {
  return (id >= 0 && id < iit::KV_COUNT) ? iit::g_variants[id].load() : 0ull;
}
