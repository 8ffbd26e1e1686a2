// b200ps K1-NVLS — the synchronous push (reduce-scatter + mean + SGD + broadcast) through NVSwitch multicast.
//
// With the symmetric buffers bound to an NVLS multicast object (b200ps/parallel/nvls.py: torch's
// symmetric-memory allocator provides the peer pointers and the multicast pointer), the owner of a slice of
// the flat parameter space needs ONE instruction per 16 bytes for each direction:
//
//   multimem.ld_reduce.add.v4.f32  [mc_grads + i]   the switch sums the replicas' gradients on the way in
//   w = master[i] - lr/world * sum ; master[i] = w   (fp32 master on the owner)
//   multimem.st.v4.f32             [mc_work + i], w  the switch fans the fresh values out to every replica
//
// instead of world-1 peer loads and world-1 peer stores per element (ps_comm.cu: push_sync_sgd_kernel).  The
// two orderings around it ("all gradients written" before, "all parameters landed" after) are the caller's:
// symmetric-memory barriers in b200ps/parallel/nvls.py.  Replaces the gRPC Send/Recv + ApplyGradientDescent
// edges E1/E2 of SURVEY §2.5.3 (example.py:108-123) for large dense groups.
#include "../common.h"
#include "ptx.cuh"

namespace b200ps {

__device__ __forceinline__ float4 multimem_ld_reduce_add_v4(const float* mc) {
  float4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(mc)
               : "memory");
  return r;
}
__device__ __forceinline__ void multimem_st_v4(float* mc, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w)
               : "memory");
}

// Elements [lo, hi) of the flat space are owned by this rank; lo and hi are multiples of 4.
__global__ void __launch_bounds__(256) push_sync_nvls_kernel(const float* __restrict__ mc_grads, float* __restrict__ mc_work,
                                                            float* __restrict__ master, long long lo, long long hi,
                                                            float step, unsigned long long* global_step) {
  pdl_wait();
  pdl_trigger();
  const long long nvec = (hi - lo) >> 2;
  for (long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += (long long)gridDim.x * blockDim.x) {
    const long long i = lo + 4 * v;
    const float4 g = multimem_ld_reduce_add_v4(mc_grads + i);
    float4 w = *reinterpret_cast<const float4*>(master + i);
    w.x = fmaf(-step, g.x, w.x); w.y = fmaf(-step, g.y, w.y);
    w.z = fmaf(-step, g.z, w.z); w.w = fmaf(-step, g.w, w.w);
    *reinterpret_cast<float4*>(master + i) = w;
    multimem_st_v4(mc_work + i, w);
  }
  if (global_step && blockIdx.x == 0 && threadIdx.x == 0) *global_step += 1ull;
}

}  // namespace b200ps

using namespace b200ps;

// mc_grads / mc_work: MULTICAST addresses of the (fp32) gradient and working-copy buffers; master: this rank's
// fp32 master, indexable by the global element index; [lo, hi): the range this rank owns (multiples of 4).
// step = lr / replicas.  Every rank launches it between two cross-rank barriers.
extern "C" int b200ps_push_sync_nvls(const float* mc_grads, float* mc_work, float* master, int64_t lo, int64_t hi,
                                     float step, uint64_t* global_step, int32_t max_ctas, b200ps_stream_t stream) {
  B200PS_REQUIRE(mc_grads && mc_work && master && lo >= 0 && hi >= lo && lo % 4 == 0 && hi % 4 == 0,
                 "push_sync_nvls: bad arguments (owner range must be 4-aligned)");
  if (hi == lo) return 0;
  int ctas = (int)std::min<long long>(((hi - lo) / 4 + 255) / 256, (long long)sm_count() * 4);
  if (max_ctas > 0) ctas = std::min(ctas, max_ctas);
  B200PS_CUDA_CHECK(launch_pdl(push_sync_nvls_kernel, dim3(ctas), dim3(256), 0, as_stream(stream), mc_grads, mc_work,
                               master, (long long)lo, (long long)hi, step,
                               reinterpret_cast<unsigned long long*>(global_step)));
  B200PS_LAUNCH_CHECK();
  return 0;
}
