// b200ps — host-side plumbing shared by every translation unit of libb200ps.so.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <atomic>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <utility>

#include "../include/b200ps.h"

namespace b200ps {

// Thread-local error text + global launch counter (definitions in common.cu).
int set_error(int code, const char* fmt, ...);
extern std::atomic<uint64_t> g_launches;
inline void count_launch(uint64_t n = 1) { g_launches.fetch_add(n, std::memory_order_relaxed); }

#define B200PS_CUDA_CHECK(expr)                                                              \
  do {                                                                                       \
    cudaError_t _e = (expr);                                                                 \
    if (_e != cudaSuccess)                                                                   \
      return ::b200ps::set_error(B200PS_ERR_CUDA, "%s failed: %s (%s:%d)", #expr,            \
                                 cudaGetErrorString(_e), __FILE__, __LINE__);                \
  } while (0)

#define B200PS_REQUIRE(cond, ...)                                                            \
  do {                                                                                       \
    if (!(cond)) return ::b200ps::set_error(B200PS_ERR_INVALID, __VA_ARGS__);                \
  } while (0)

// Checks the launch itself (not completion) and bumps the launch counter.
#define B200PS_LAUNCH_CHECK()                                                                \
  do {                                                                                       \
    ::b200ps::count_launch();                                                                \
    B200PS_CUDA_CHECK(cudaGetLastError());                                                   \
  } while (0)

// Number of SMs of the current device (cached per device).
int sm_count();

// cuTensorMapEncodeTiled resolved through the runtime (no link-time libcuda dependency,
// so the library also loads on the GPU-less build box).
typedef CUresult (*tensormap_encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                        const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                        const cuuint32_t*, CUtensorMapInterleave,
                                        CUtensorMapSwizzle, CUtensorMapL2promotion,
                                        CUtensorMapFloatOOBfill);
tensormap_encode_fn get_tensormap_encode();

inline cudaStream_t as_stream(b200ps_stream_t s) { return reinterpret_cast<cudaStream_t>(s); }

// Is programmatic dependent launch enabled?  (B200PS_PDL=0 in the environment turns it off.)
bool pdl_enabled();

// Launch `kernel` with the programmatic-stream-serialization attribute (when enabled), so its
// prologue can overlap the tail of the previous kernel in the stream.  The kernel must execute
// griddepcontrol.wait (pdl_wait()) before touching anything its predecessor produced.
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                              cudaStream_t stream, Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(std::forward<Args>(args))...);
}

}  // namespace b200ps
