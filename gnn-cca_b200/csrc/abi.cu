// Version / status plumbing of the C ABI (include/gnncca.h).
#include <string.h>
#include <atomic>
#include "common.cuh"

namespace gnncca {
static thread_local char g_last_error[256] = "";
void set_cuda_error(cudaError_t e) {
  strncpy(g_last_error, cudaGetErrorString(e), sizeof(g_last_error) - 1);
  g_last_error[sizeof(g_last_error) - 1] = 0;
}
void set_error_text(const char* text) {   // same slot as set_cuda_error: what gnncca_last_cuda_error() returns
  strncpy(g_last_error, text, sizeof(g_last_error) - 1);
  g_last_error[sizeof(g_last_error) - 1] = 0;
}
thread_local bool tl_launch_as_dependent = false;
static std::atomic<long long> g_launches{0};
void count_launches(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
}  // namespace gnncca

extern "C" int64_t gnncca_launch_count(void) { return (int64_t)gnncca::g_launches.load(std::memory_order_relaxed); }

extern "C" int gnncca_abi_version(void) { return GNNCCA_ABI_VERSION; }

extern "C" const char* gnncca_status_string(int status) {
  switch (status) {
    case GNNCCA_OK: return "ok";
    case GNNCCA_ERR_BAD_ARG: return "bad argument";
    case GNNCCA_ERR_UNSUPPORTED: return "unsupported configuration (no kernel built for it; there is no fallback)";
    case GNNCCA_ERR_LAUNCH: return "CUDA launch failure";
    case GNNCCA_ERR_WORKSPACE: return "workspace too small";
    default: return "unknown status";
  }
}

extern "C" const char* gnncca_last_cuda_error(void) { return gnncca::g_last_error; }
