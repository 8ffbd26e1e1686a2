#include "workspace.h"

#include <mutex>
#include <vector>

namespace b200ps {

static Workspace g_ws[64];
struct WsUse { cudaStream_t stream = nullptr; cudaEvent_t ev = nullptr; bool armed = false; };
static WsUse g_use[64];

static bool capturing(cudaStream_t s) {
  cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(s, &st) != cudaSuccess) { cudaGetLastError(); return false; }
  return st != cudaStreamCaptureStatusNone;
}
static std::mutex g_ws_mu;
// Blocks that were outgrown.  Kernel nodes of already captured CUDA graphs keep their addresses, so they stay
// allocated until the process exits (a few MiB at most: the workspace only ever grows).
static std::vector<void*> g_ws_retired;

int get_workspace(size_t bytes, Workspace* out, cudaStream_t stream, bool check_stream) {
  int dev = 0;
  B200PS_CUDA_CHECK(cudaGetDevice(&dev));
  B200PS_REQUIRE(dev >= 0 && dev < 64, "workspace: device index %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_ws_mu);
  Workspace& w = g_ws[dev];
  if (check_stream && !capturing(stream)) {
    WsUse& u = g_use[dev];
    if (u.armed && u.stream != stream && cudaEventQuery(u.ev) == cudaErrorNotReady)
      return set_error(B200PS_ERR_INVALID,
                       "workspace: a kernel on another stream is still using the library workspace of device %d "
                       "(launches that use it must not overlap: one training stream per process)", dev);
    cudaGetLastError();
  }
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

void workspace_mark_used(cudaStream_t stream) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64 || capturing(stream)) return;
  std::lock_guard<std::mutex> lock(g_ws_mu);
  WsUse& u = g_use[dev];
  if (u.ev == nullptr && cudaEventCreateWithFlags(&u.ev, cudaEventDisableTiming) != cudaSuccess) { cudaGetLastError(); return; }
  if (cudaEventRecord(u.ev, stream) == cudaSuccess) { u.stream = stream; u.armed = true; }
  else cudaGetLastError();
}

unsigned long long* bad_id_counter() {
  Workspace w;
  if (get_workspace(0, &w) != 0) return nullptr;
  return reinterpret_cast<unsigned long long*>(w.counters + 62);   // 8-byte aligned tail of the counter block
}

}  // namespace b200ps
