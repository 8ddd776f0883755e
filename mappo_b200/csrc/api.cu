// Library-wide C-ABI pieces: version, error text, launch counter.
#include "common.cuh"
#include <atomic>
#include <string.h>

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

// Host-side helper (no CUDA): the runner's one-hot `actions_env = np.eye(n)[actions]`
// (onpolicy/runner/shared/mpe_runner.py:110-121; MultiDiscrete: the heads' one-hots concatenated) written as uint8
// into a caller buffer (pinned, so that envs.step sends 1 byte per logit).  ids: int64 [rows][n_heads] on the HOST.
extern "C" int mappo_host_onehot_u8(const int64_t* ids_host, long long rows, int n_heads, const int* head_dim_host,
                                    uint8_t* out_host) {
  MAPPO_REQUIRE(ids_host && out_host && head_dim_host && rows > 0 && n_heads >= 1 && n_heads <= MAPPO_MAX_HEADS,
                "host_onehot_u8: bad arguments");
  int width = 0, off[MAPPO_MAX_HEADS];
  for (int h = 0; h < n_heads; ++h) { off[h] = width; width += head_dim_host[h]; }
  memset(out_host, 0, (size_t)rows * width);
  for (long long r = 0; r < rows; ++r) {
    uint8_t* row = out_host + r * width;
    for (int h = 0; h < n_heads; ++h) {
      const int64_t a = ids_host[r * n_heads + h];
      MAPPO_REQUIRE(a >= 0 && a < head_dim_host[h], "host_onehot_u8: action id %lld out of range at row %lld", (long long)a, r);
      row[off[h] + a] = 1;
    }
  }
  return 0;
}

extern "C" int mappo_host_onehot_u8_f32(const float* ids_host, long long rows, int n_heads, const int* head_dim_host,
                                    uint8_t* out_host) {
  MAPPO_REQUIRE(ids_host && out_host && head_dim_host && rows > 0 && n_heads >= 1 && n_heads <= MAPPO_MAX_HEADS,
                "host_onehot_u8_f32: bad arguments");
  int width = 0, off[MAPPO_MAX_HEADS];
  for (int h = 0; h < n_heads; ++h) { off[h] = width; width += head_dim_host[h]; }
  memset(out_host, 0, (size_t)rows * width);
  for (long long r = 0; r < rows; ++r) {
    uint8_t* row = out_host + r * width;
    for (int h = 0; h < n_heads; ++h) {
      const long long a = (long long)ids_host[r * n_heads + h];
      MAPPO_REQUIRE(a >= 0 && a < head_dim_host[h], "host_onehot_u8_f32: action id %lld out of range at row %lld", (long long)a, r);
      row[off[h] + a] = 1;
    }
  }
  return 0;
}
