#include "workspace.h"

#include <mutex>
#include <vector>

namespace b200ps {

static Workspace g_ws[64];
static std::mutex g_ws_mu;
// Blocks that were outgrown.  Kernel nodes of already captured CUDA graphs keep their addresses, so they stay
// allocated until the process exits (a few MiB at most: the workspace only ever grows).
static std::vector<void*> g_ws_retired;

int get_workspace(size_t bytes, Workspace* out) {
  int dev = 0;
  B200PS_CUDA_CHECK(cudaGetDevice(&dev));
  B200PS_REQUIRE(dev >= 0 && dev < 64, "workspace: device index %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_ws_mu);
  Workspace& w = g_ws[dev];
  if (w.counters == nullptr) {
    B200PS_CUDA_CHECK(cudaMalloc(reinterpret_cast<void**>(&w.counters), 64 * sizeof(unsigned int)));
    B200PS_CUDA_CHECK(cudaMemset(w.counters, 0, 64 * sizeof(unsigned int)));
  }
  if (bytes > w.capacity) {
    // growing allocates, which is illegal inside a stream capture: cudaMalloc fails there and the
    // error text below tells the caller what to do
    size_t want = (bytes + (size_t(1) << 20) - 1) & ~((size_t(1) << 20) - 1);
    void* p = nullptr;
    cudaError_t e = cudaMalloc(&p, want);
    if (e != cudaSuccess) {
      return set_error(B200PS_ERR_CUDA,
                       "workspace: cudaMalloc(%zu) failed: %s (size the workspace with an eager call "
                       "before CUDA-graph capture)", want, cudaGetErrorString(e));
    }
    if (w.data) g_ws_retired.push_back(w.data);   // never freed: captured graphs may still point into it
    w.data = p;
    w.capacity = want;
  }
  *out = w;
  return 0;
}

unsigned long long* bad_id_counter() {
  Workspace w;
  if (get_workspace(0, &w) != 0) return nullptr;
  return reinterpret_cast<unsigned long long*>(w.counters + 62);   // 8-byte aligned tail of the counter block
}

}  // namespace b200ps
