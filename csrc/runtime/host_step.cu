// b200ps runtime — the host side of one end-to-end training step, as ONE native call.
//
// A user-level step is: copy this step's inputs from pinned host memory to the device, replay
// the captured step graph, copy the loss back, mark completion.  Issued from Python these are 5-6
// framework calls (~100 us of interpreter and dispatcher time, 4x the GPU time of the MNIST step);
// issued from here they are a handful of driver calls (~10 us).  This replaces the per-step
// `sess.run(..., feed_dict=...)` round trip of the reference (example.py:168-170).
#include "../common.h"

using namespace b200ps;

extern "C" int b200ps_host_step(void* dst0, const void* src0, uint64_t bytes0, void* dst1, const void* src1,
                                uint64_t bytes1, void* graph_exec, void* result_host, const void* result_dev,
                                uint64_t result_bytes, void* done_event, b200ps_stream_t stream) {
  cudaStream_t s = as_stream(stream);
  if (bytes0) B200PS_CUDA_CHECK(cudaMemcpyAsync(dst0, src0, bytes0, cudaMemcpyHostToDevice, s));
  if (bytes1) B200PS_CUDA_CHECK(cudaMemcpyAsync(dst1, src1, bytes1, cudaMemcpyHostToDevice, s));
  B200PS_REQUIRE(graph_exec != nullptr, "host_step: null graph");
  B200PS_CUDA_CHECK(cudaGraphLaunch(reinterpret_cast<cudaGraphExec_t>(graph_exec), s));
  if (result_bytes)
    B200PS_CUDA_CHECK(cudaMemcpyAsync(result_host, result_dev, result_bytes, cudaMemcpyDeviceToHost, s));
  if (done_event) B200PS_CUDA_CHECK(cudaEventRecord(reinterpret_cast<cudaEvent_t>(done_event), s));
  return 0;
}

extern "C" int b200ps_event_create(void** ev) {
  cudaEvent_t e;
  B200PS_CUDA_CHECK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  *ev = e;
  return 0;
}
extern "C" int b200ps_event_synchronize(void* ev) {
  B200PS_CUDA_CHECK(cudaEventSynchronize(reinterpret_cast<cudaEvent_t>(ev)));
  return 0;
}
extern "C" int b200ps_event_destroy(void* ev) {
  if (ev) B200PS_CUDA_CHECK(cudaEventDestroy(reinterpret_cast<cudaEvent_t>(ev)));
  return 0;
}
