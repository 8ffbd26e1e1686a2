// b200ps — host-side plumbing shared by every translation unit of libb200ps.so.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <atomic>
#include <cstdarg>
#include <cstdint>
#include <cstdio>

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

}  // namespace b200ps
