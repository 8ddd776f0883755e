// Library-wide C-ABI pieces: version, error text, launch counter.
#include "common.cuh"
#include <atomic>

namespace mappo {
static thread_local char g_err[512] = "";
static std::atomic<uint64_t> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }
}  // namespace mappo

extern "C" int mappo_abi_version(void) { return MAPPO_B200_ABI_VERSION; }
extern "C" int mappo_sm_arch(void) { return 100; }
extern "C" const char* mappo_last_error(void) { return mappo::g_err; }
extern "C" uint64_t mappo_launch_count(void) { return mappo::g_launches.load(); }
