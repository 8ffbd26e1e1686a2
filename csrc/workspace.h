// b200ps — per-device scratch workspace owned by the library.
//
// A few kernels need global memory beyond their inputs and outputs (cross-CTA partial sums,
// arrival counters).  The workspace is allocated with cudaMalloc on first use and grown on
// demand, one per device; it is NOT keyed by stream, so launches that use it must not overlap
// on the GPU (they do not in this framework: one training stream per process).  Growth is an
// allocation and therefore must happen in an eager call before any CUDA-graph capture; a block that
// was outgrown is kept allocated (never freed), so graphs captured against it stay valid.
#pragma once
#include "common.h"

namespace b200ps {

struct Workspace {
  void* data = nullptr;         // scratch bytes (contents undefined between launches)
  size_t capacity = 0;
  unsigned int* counters = nullptr;  // 64 zero-initialised uint32 arrival counters (self-resetting)
};

// Returns the current device's workspace with at least `bytes` of scratch.  With a stream (the one the kernel is
// about to be launched on) the single-stream rule is CHECKED on eager launches: if the previous eager user ran on a
// different stream and has not finished, the call fails instead of letting two kernels share the scratch.  (Launches
// under stream capture are not tracked: a captured graph replays wherever its owner puts it.)
int get_workspace(size_t bytes, Workspace* out, cudaStream_t stream = nullptr, bool check_stream = false);
// Call right after the launch that uses the workspace (records "in use until here" for the check above).
void workspace_mark_used(cudaStream_t stream);

// Device counter of feature ids that the sparse kernels refused (id < 0 or id >= features); counters[62..63].
unsigned long long* bad_id_counter();

}  // namespace b200ps
