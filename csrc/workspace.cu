#include "workspace.h"

#include <mutex>

namespace b200ps {

static Workspace g_ws[64];
static std::mutex g_ws_mu;

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
    if (w.data) cudaFree(w.data);
    w.data = p;
    w.capacity = want;
  }
  *out = w;
  return 0;
}

}  // namespace b200ps
