// b200ps — error text, launch counter, device queries, driver entry points.
#include "common.h"

#include <mutex>

namespace b200ps {

static thread_local char tl_error[512] = "";
std::atomic<uint64_t> g_launches{0};

int set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(tl_error, sizeof(tl_error), fmt, ap);
  va_end(ap);
  return code;
}

int sm_count() {
  static int cached[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

bool pdl_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = std::getenv("B200PS_PDL");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v == 1;
}

tensormap_encode_fn get_tensormap_encode() {
  static tensormap_encode_fn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
            cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<tensormap_encode_fn>(p);
  });
  return fn;
}

}  // namespace b200ps

extern "C" {
const char* b200ps_last_error(void) { return b200ps::tl_error; }
const char* b200ps_version(void) { return "b200ps 0.1 sm_100a"; }
uint64_t b200ps_launch_count(void) { return b200ps::g_launches.load(); }
}
