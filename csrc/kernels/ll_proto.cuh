// b200ps — pieces of the push protocols shared by ps_comm.cu and the fused MLP step (mlp_fused.cu):
// the signal-pad layout and the 8-byte {value, epoch} "LL" packets of the low-latency push.
#pragma once
#include "ptx.cuh"

namespace b200ps {

// layout of one rank's signal pad (uint32 words)
constexpr int kPadWords = 256;
constexpr int kMaxWorld = 32;
constexpr int kFlagGradsReady = 0;    // [0,32)   written by peer r at slot r
constexpr int kFlagParamsDone = 32;   // [32,64)
constexpr int kFlagBarrier = 64;      // [64,96)
constexpr int kSlotEpoch = 96;        // push epoch (local)
constexpr int kSlotBarrierEpoch = 97; // barrier epoch (local)
constexpr int kSlotCtaDone = 98;      // CTA completion counter (local)


// An aligned 8-byte store is single-copy atomic, so a receiver that sees the expected epoch in the
// upper word also sees the value: data and flag travel in one NVLink write, no fence, no second hop.
__device__ __forceinline__ void st_packet(uint2* p, float v, uint32_t epoch) {
#ifdef B200PS_CHECKED
  {   // a slot is written exactly once per epoch, and epochs only move forward
    uint32_t ov, oe;
    asm volatile("ld.relaxed.sys.global.v2.u32 {%0, %1}, [%2];" : "=r"(ov), "=r"(oe) : "l"(p) : "memory");
    B200PS_DEV_CHECK((int32_t)(oe - epoch) < 0 /* wrap-safe: epochs are 32-bit counters */, "packet slot %p rewritten: holds epoch %u, writing epoch %u\n", (void*)p, oe, epoch);
  }
#endif
  asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(__float_as_uint(v)), "r"(epoch) : "memory");
}
// One look at a packet (no spinning): lets a thread put many polls in flight before it checks any of them.
__device__ __forceinline__ uint2 peek_packet(const uint2* p) {
  uint2 r;
  asm volatile("ld.relaxed.sys.global.v2.u32 {%0, %1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ float poll_packet(const uint2* p, uint32_t epoch) {
  uint32_t v, f;
  asm volatile("ld.relaxed.sys.global.v2.u32 {%0, %1}, [%2];" : "=r"(v), "=r"(f) : "l"(p) : "memory");
  if (f != epoch) {
    const uint64_t t0 = globaltimer_ns();
    uint32_t spins = 0;
    do {
      asm volatile("ld.relaxed.sys.global.v2.u32 {%0, %1}, [%2];" : "=r"(v), "=r"(f) : "l"(p) : "memory");
      B200PS_DEV_CHECK((int32_t)(f - epoch) <= 0, "packet %p is from the future: epoch %u while waiting for %u\n", (void*)p, f, epoch);
      if ((++spins & 0xff) == 0 && globaltimer_ns() - t0 > B200PS_WAIT_TIMEOUT_NS) {
        printf("b200ps: LL packet wait timed out (block %d, want epoch %u, have %u)\n", blockIdx.x, epoch, f);
        __trap();
      }
    } while (f != epoch);
  }
  return __uint_as_float(v);
}


}  // namespace b200ps
