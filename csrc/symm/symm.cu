// b200ps N1 — symmetric memory: identical buffers on every rank, mapped into every rank's
// address space through CUDA IPC, so kernels can load/store/atomically-add peer HBM over
// NVLink directly.  This replaces the per-task gRPC server of the reference
// (example.py:38-40, lr2.py:331): "connecting to the ps" becomes "mapping its shard".
// Handle exchange (64 bytes per buffer) is done by the Python control plane over
// torch.distributed; nothing here talks to the network.
#include "../common.h"

#include <cstring>

using namespace b200ps;

extern "C" int b200ps_symm_alloc(uint64_t bytes, void** ptr_out, void* handle_out /* 64 bytes */) {
  B200PS_REQUIRE(bytes > 0 && ptr_out, "symm_alloc: bad args");
  void* p = nullptr;
  B200PS_CUDA_CHECK(cudaMalloc(&p, bytes));
  B200PS_CUDA_CHECK(cudaMemset(p, 0, bytes));
  if (handle_out) {
    cudaIpcMemHandle_t h;
    cudaError_t e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) {
      cudaFree(p);
      return set_error(B200PS_ERR_CUDA, "cudaIpcGetMemHandle failed: %s", cudaGetErrorString(e));
    }
    static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
    std::memcpy(handle_out, &h, sizeof(h));
  }
  *ptr_out = p;
  return 0;
}

extern "C" int b200ps_symm_open(const void* handle /* 64 bytes */, void** ptr_out) {
  B200PS_REQUIRE(handle && ptr_out, "symm_open: bad args");
  cudaIpcMemHandle_t h;
  std::memcpy(&h, handle, sizeof(h));
  void* p = nullptr;
  B200PS_CUDA_CHECK(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  *ptr_out = p;
  return 0;
}

extern "C" int b200ps_symm_close(void* ptr) {
  if (ptr) B200PS_CUDA_CHECK(cudaIpcCloseMemHandle(ptr));
  return 0;
}

extern "C" int b200ps_symm_free(void* ptr) {
  if (ptr) B200PS_CUDA_CHECK(cudaFree(ptr));
  return 0;
}

// Plain device allocation helpers so C embedders need no CUDA headers of their own.
extern "C" int b200ps_device_count(int32_t* n) {
  int c = 0;
  cudaError_t e = cudaGetDeviceCount(&c);
  if (e != cudaSuccess) { *n = 0; cudaGetLastError(); return 0; }
  *n = c;
  return 0;
}

extern "C" int b200ps_can_access_peer(int32_t dev, int32_t peer, int32_t* ok) {
  int c = 0;
  B200PS_CUDA_CHECK(cudaDeviceCanAccessPeer(&c, dev, peer));
  *ok = c;
  return 0;
}
