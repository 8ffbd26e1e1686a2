// b200ps runtime — HostFeeder: a pipelined host->device input feeder around captured step graphs.
//
// The reference feeds every step through `sess.run(..., feed_dict={x: batch_x, y_: batch_y})`
// (example.py:165-170): host batch in, loss out, nothing overlapped.  Here a ring of S *slots*
// decouples the three engines that a step touches:
//
//   copy stream     H2D(slot s)            -- waits until the graph that last read slot s is done
//   compute stream  graph[s]               -- waits until slot s is filled; graphs run in ticket order
//   host            wait(ticket)           -- blocks on the slot's completion event only
//
// so the PCIe copy of step t+1 runs under the kernels of step t and the host never blocks unless it
// asks for a result.  A slot's graph may contain one step or a chunk of K steps over the slot's
// staging buffer; the step kernels write their losses straight into the slot's pinned result
// block (mapped host memory), so there is no D2H copy to queue.  The feeder neither allocates
// device memory nor owns the graphs: the trainer binds (graph, staging pointers) per slot.
#include <chrono>
#include <cstring>
#include <vector>

#include "../common.h"

using namespace b200ps;

namespace {

struct Slot {
  cudaGraphExec_t graph = nullptr;
  void* dst[2] = {nullptr, nullptr};
  cudaEvent_t ready = nullptr;  // staging filled (copy stream)
  cudaEvent_t done = nullptr;   // graph finished (compute stream): results valid, staging reusable
  int64_t ticket = -1;          // last ticket that used this slot
};

struct Feeder {
  int device = 0;
  uint64_t bytes[2] = {0, 0};
  uint64_t result_bytes = 0;
  cudaStream_t copy = nullptr;
  char* results = nullptr;  // pinned + mapped, slots * result_bytes
  std::vector<Slot> slots;
  int64_t next_ticket = 0;
  // host-side cost accounting of submit(): wait-event | copies | record+wait | graph launch | record
  double phase_ns[5] = {0, 0, 0, 0, 0};
};

}  // namespace

extern "C" int b200ps_feeder_destroy(void* feeder);

extern "C" int b200ps_feeder_create(void** out, int32_t slots, uint64_t bytes0, uint64_t bytes1,
                                    uint64_t result_bytes) {
  B200PS_REQUIRE(out && slots >= 1 && slots <= 64, "feeder_create: slots must be in [1, 64]");
  B200PS_REQUIRE(result_bytes >= 4 && result_bytes % 4 == 0, "feeder_create: result_bytes must be a multiple of 4");
  Feeder* f = new Feeder();
  f->bytes[0] = bytes0;
  f->bytes[1] = bytes1;
  f->result_bytes = result_bytes;
  f->slots.resize(slots);
  cudaError_t e = cudaGetDevice(&f->device);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&f->copy, cudaStreamNonBlocking);
  if (e == cudaSuccess)
    e = cudaHostAlloc(reinterpret_cast<void**>(&f->results), slots * result_bytes, cudaHostAllocPortable | cudaHostAllocMapped);
  for (int i = 0; i < slots && e == cudaSuccess; ++i) {
    e = cudaEventCreateWithFlags(&f->slots[i].ready, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&f->slots[i].done, cudaEventDisableTiming);
  }
  if (e != cudaSuccess) {
    b200ps_feeder_destroy(f);      // releases whatever was created so far
    return set_error(B200PS_ERR_CUDA, "feeder_create: %s", cudaGetErrorString(e));
  }
  memset(f->results, 0, slots * result_bytes);
  *out = f;
  return 0;
}

// Address (host == device under UVA) of slot `slot`'s result block; the step kernels of that slot's
// graph write their loss words there.
extern "C" int b200ps_feeder_result_ptr(void* feeder, int32_t slot, void** ptr) {
  Feeder* f = static_cast<Feeder*>(feeder);
  B200PS_REQUIRE(f && slot >= 0 && slot < (int)f->slots.size() && ptr, "feeder_result_ptr: bad slot");
  *ptr = f->results + slot * f->result_bytes;
  return 0;
}

extern "C" int b200ps_feeder_bind(void* feeder, int32_t slot, void* graph_exec, void* dst0, void* dst1) {
  Feeder* f = static_cast<Feeder*>(feeder);
  B200PS_REQUIRE(f && slot >= 0 && slot < (int)f->slots.size(), "feeder_bind: bad slot");
  B200PS_REQUIRE(graph_exec != nullptr, "feeder_bind: null graph");
  B200PS_REQUIRE((dst0 || !f->bytes[0]) && (dst1 || !f->bytes[1]), "feeder_bind: null staging pointer");
  Slot& s = f->slots[slot];
  s.graph = reinterpret_cast<cudaGraphExec_t>(graph_exec);
  s.dst[0] = dst0;
  s.dst[1] = dst1;
  return 0;
}

// Queue one unit of work: copy src0/src1 (pinned host) into the next slot's staging buffers on the
// copy stream, then replay that slot's graph on `stream`.  Never blocks the host: reuse of a slot is
// ordered on the device (the copy waits for the graph that last read the slot).  A result nobody
// waited for within `slots` submissions is overwritten.  `*ticket` identifies the unit.
extern "C" int b200ps_feeder_submit(void* feeder, const void* src0, const void* src1, b200ps_stream_t stream,
                                    int64_t* ticket) {
  Feeder* f = static_cast<Feeder*>(feeder);
  B200PS_REQUIRE(f && ticket, "feeder_submit: null argument");
  const int64_t t = f->next_ticket;
  Slot& s = f->slots[t % (int64_t)f->slots.size()];
  B200PS_REQUIRE(s.graph != nullptr, "feeder_submit: slot %d has no graph bound", (int)(t % (int64_t)f->slots.size()));
  cudaStream_t cs = as_stream(stream);
  using clk = std::chrono::steady_clock;
  auto t0 = clk::now();
  auto lap = [&](int i) {
    auto t1 = clk::now();
    f->phase_ns[i] += std::chrono::duration<double, std::nano>(t1 - t0).count();
    t0 = t1;
  };
  if (s.ticket >= 0) B200PS_CUDA_CHECK(cudaStreamWaitEvent(f->copy, s.done, 0));
  lap(0);
  if (f->bytes[0]) B200PS_CUDA_CHECK(cudaMemcpyAsync(s.dst[0], src0, f->bytes[0], cudaMemcpyHostToDevice, f->copy));
  if (f->bytes[1]) B200PS_CUDA_CHECK(cudaMemcpyAsync(s.dst[1], src1, f->bytes[1], cudaMemcpyHostToDevice, f->copy));
  lap(1);
  B200PS_CUDA_CHECK(cudaEventRecord(s.ready, f->copy));
  B200PS_CUDA_CHECK(cudaStreamWaitEvent(cs, s.ready, 0));
  lap(2);
  B200PS_CUDA_CHECK(cudaGraphLaunch(s.graph, cs));
  lap(3);
  B200PS_CUDA_CHECK(cudaEventRecord(s.done, cs));
  lap(4);
  s.ticket = t;
  f->next_ticket = t + 1;
  *ticket = t;
  return 0;
}

// Block until unit `ticket` has finished and copy its result block to `out` (result_bytes).  Only
// the most recent `slots` tickets are retrievable: older results have been overwritten.
extern "C" int b200ps_feeder_wait(void* feeder, int64_t ticket, void* out) {
  Feeder* f = static_cast<Feeder*>(feeder);
  B200PS_REQUIRE(f && ticket >= 0 && ticket < f->next_ticket, "feeder_wait: unknown ticket");
  const int idx = (int)(ticket % (int64_t)f->slots.size());
  Slot& s = f->slots[idx];
  B200PS_REQUIRE(s.ticket == ticket, "feeder_wait: ticket %lld was overwritten (ring of %d slots)", (long long)ticket,
                 (int)f->slots.size());
  B200PS_CUDA_CHECK(cudaEventSynchronize(s.done));
  if (out) memcpy(out, f->results + idx * f->result_bytes, f->result_bytes);
  return 0;
}

// Host-side cost of submit() so far: nanoseconds spent in [wait-event, H2D copies, ready record+wait,
// graph launch, done record] and the number of units submitted.
extern "C" int b200ps_feeder_stats(void* feeder, double* phase_ns /* 5 */, int64_t* submitted) {
  Feeder* f = static_cast<Feeder*>(feeder);
  B200PS_REQUIRE(f && phase_ns && submitted, "feeder_stats: null argument");
  for (int i = 0; i < 5; ++i) phase_ns[i] = f->phase_ns[i];
  *submitted = f->next_ticket;
  return 0;
}

extern "C" int b200ps_feeder_destroy(void* feeder) {
  Feeder* f = static_cast<Feeder*>(feeder);
  if (!f) return 0;
  if (f->copy) cudaStreamSynchronize(f->copy);
  for (Slot& s : f->slots) {
    if (s.done) cudaEventSynchronize(s.done);
    if (s.ready) cudaEventDestroy(s.ready);
    if (s.done) cudaEventDestroy(s.done);
  }
  if (f->copy) cudaStreamDestroy(f->copy);
  if (f->results) cudaFreeHost(f->results);
  delete f;
  return 0;
}
