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
#include "ll_proto.cuh"

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

// ---------------------------------------------------------------------------------------------------------------
// The ParameterStore's own NVLS path: the same update as push_sync_sgd_kernel (ps_comm.cu) — both flag exchanges of
// the synchronous barrier included — with the replica loop replaced by the switch.  G / O: gradient and working-copy
// element types (fp32 groups: float/float with master == working copy; bf16 groups: bf16/bf16 + fp32 master).
//   phase 0  "my gradients are written" to every peer's pad, wait for everybody's
//   phase 1  owner: multimem.ld_reduce (the switch sums the replicas; bf16 inputs accumulate in fp32 inside the
//            switch) -> mean -> SGD on the fp32 master -> multimem.st of the fresh values into EVERY replica
//            (broadcast = 0: plain store into the owner's own copy only, for the fused-pull forward)
//   phase 2  last CTA: one system-scope fence, "my slice has landed" to every peer, wait for everybody's
// The reduction order inside the switch is not the rank order of the P2P kernel, so results are not bit-identical
// to that path (replicas still receive identical values: there is one owner and one broadcast per element).
template <typename T> struct Mm;
template <> struct Mm<float> {
  static constexpr int kPer = 4;                    // elements per 16-byte multimem access
  static __device__ __forceinline__ void ld_reduce(const float* mc, float (&v)[4]) {
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]) : "l"(mc) : "memory");
  }
  static __device__ __forceinline__ void st(float* mc, const float (&v)[4]) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]) : "memory");
  }
  static __device__ __forceinline__ void st_local(float* p, const float (&v)[4]) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  }
};
template <> struct Mm<__nv_bfloat16> {
  static constexpr int kPer = 8;
  static __device__ __forceinline__ void ld_reduce(const __nv_bfloat16* mc, float (&v)[8]) {
    uint32_t r[4];
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "l"(mc) : "memory");
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float2 f = __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&r[k]));
      v[2 * k] = f.x; v[2 * k + 1] = f.y;
    }
  }
  static __device__ __forceinline__ void pack(const float (&v)[8], uint32_t (&r)[4]) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      __nv_bfloat162 b = __floats2bfloat162_rn(v[2 * k], v[2 * k + 1]);
      r[k] = *reinterpret_cast<uint32_t*>(&b);
    }
  }
  static __device__ __forceinline__ void st(__nv_bfloat16* mc, const float (&v)[8]) {
    uint32_t r[4];
    pack(v, r);
    asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1, %2, %3, %4};" ::"l"(mc), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]) : "memory");
  }
  static __device__ __forceinline__ void st_local(__nv_bfloat16* p, const float (&v)[8]) {
    uint32_t r[4];
    pack(v, r);
    *reinterpret_cast<uint4*>(p) = make_uint4(r[0], r[1], r[2], r[3]);
  }
};

struct NvlsArgs {
  uint32_t* const* pads;      // [world] signal pads (peer-mapped)
  const void* mc_grads;       // multicast address of the gradient buffers (null: pull only)
  void* mc_work;              // multicast address of the working copies
  void* work_local;           // this rank's own working copy (unicast)
  float* master;              // this rank's fp32 master, indexable by the global element index
  long long lo, hi;           // owned range (multiples of 8)
  float step;
  unsigned long long* global_step;
  int* cursor; int cursor_step, cursor_wrap;
  int world, rank, broadcast;
  int do_reduce;              // 1: push (reduce + SGD + broadcast); 0: pull (broadcast the master slice)
};

template <typename G, typename O>
__global__ void __launch_bounds__(256) store_nvls_kernel(const NvlsArgs a) {
  constexpr int P = Mm<O>::kPer;
  const int tid = threadIdx.x, world = a.world, rank = a.rank;
  uint32_t* pad = a.pads[rank];
  pdl_wait();
  pdl_trigger();
  const uint32_t epoch = pad[kSlotEpoch] + 1;
  // -- phase 0: everybody's gradients are written (pull: nothing to wait for — masters are always current)
  if (a.do_reduce) {
    if (blockIdx.x == 0 && tid < world) st_relaxed_sys(a.pads[tid] + kFlagGradsReady + rank, epoch);
    if (tid < world) wait_flag_geq(pad + kFlagGradsReady + tid, epoch);
    __syncthreads();
  }
  // -- phase 1 (kU independent 16-byte accesses per thread per trip: all switch reads are issued before any use)
  const long long nvec = (a.hi - a.lo) / P;
  const long long stride = (long long)gridDim.x * blockDim.x;
  constexpr int kU = 4;
  for (long long v0 = (long long)blockIdx.x * blockDim.x + tid; v0 < nvec; v0 += kU * stride) {
    float g[kU][P], w[kU][P];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const long long v = v0 + u * stride;
      if (v < nvec && a.do_reduce) Mm<G>::ld_reduce(reinterpret_cast<const G*>(a.mc_grads) + a.lo + v * P, g[u]);
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const long long v = v0 + u * stride;
      if (v >= nvec) continue;
      const long long i = a.lo + v * P;
#pragma unroll
      for (int k = 0; k < P; k += 4) {
        const float4 m = __ldcs(reinterpret_cast<const float4*>(a.master + i + k));
        w[u][k] = m.x; w[u][k + 1] = m.y; w[u][k + 2] = m.z; w[u][k + 3] = m.w;
      }
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const long long v = v0 + u * stride;
      if (v >= nvec) continue;
      const long long i = a.lo + v * P;
      if (a.do_reduce) {
#pragma unroll
        for (int k = 0; k < P; ++k) w[u][k] = fmaf(-a.step, g[u][k], w[u][k]);
#pragma unroll
        for (int k = 0; k < P; k += 4)
          __stcs(reinterpret_cast<float4*>(a.master + i + k), make_float4(w[u][k], w[u][k + 1], w[u][k + 2], w[u][k + 3]));
      }
      if (a.broadcast) Mm<O>::st(reinterpret_cast<O*>(a.mc_work) + i, w[u]);
      else Mm<O>::st_local(reinterpret_cast<O*>(a.work_local) + i, w[u]);
    }
  }
  // -- phase 2: when every CTA of this rank is done, tell the peers and wait for theirs
  __syncthreads();
  __shared__ bool s_last;
  if (tid == 0) {
    __threadfence();
    s_last = gridDim.x == 1 || (atomicAdd(pad + kSlotCtaDone, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (s_last) {
    if (tid == 0) __threadfence_system();
    __syncthreads();
    if (tid < world) st_relaxed_sys(a.pads[tid] + kFlagParamsDone + rank, epoch);
    if (tid < world) wait_flag_geq(pad + kFlagParamsDone + tid, epoch);
    __syncthreads();
    if (tid == 0) {
      pad[kSlotCtaDone] = 0;
      pad[kSlotEpoch] = epoch;
      if (a.global_step) *a.global_step += 1ull;
      if (a.cursor) *a.cursor = (*a.cursor + a.cursor_step) % a.cursor_wrap;
    }
  }
}

}  // namespace b200ps

using namespace b200ps;

// The ParameterStore's NVLS push / pull.  dtype: B200PS_DT_F32 (master may alias work_local) or B200PS_DT_BF16.
// do_reduce = 1: synchronous push (mean of ALL replicas); 0: pull (owners broadcast their master slices).
extern "C" int b200ps_store_nvls(uint32_t* const* pads, const void* mc_grads, void* mc_work, void* work_local,
                                 float* master, int64_t lo, int64_t hi, int32_t dtype, int32_t world, int32_t rank,
                                 float lr, int32_t do_reduce, int32_t broadcast, uint64_t* global_step,
                                 int32_t* cursor, int32_t cursor_step, int32_t cursor_wrap, b200ps_stream_t stream) {
  B200PS_REQUIRE(pads && mc_work && work_local && master && world >= 2 && world <= kMaxWorld && rank >= 0 && rank < world,
                 "store_nvls: bad arguments");
  B200PS_REQUIRE(!do_reduce || mc_grads, "store_nvls: a push needs the multicast address of the gradients");
  B200PS_REQUIRE(lo >= 0 && hi >= lo && lo % 8 == 0 && hi % 8 == 0, "store_nvls: the owned range must be 8-aligned");
  B200PS_REQUIRE(dtype == B200PS_DT_F32 || dtype == B200PS_DT_BF16, "store_nvls: dtype must be fp32 or bf16");
  NvlsArgs a;
  a.pads = pads; a.mc_grads = mc_grads; a.mc_work = mc_work; a.work_local = work_local; a.master = master;
  a.lo = lo; a.hi = hi; a.step = lr / (float)world;
  a.global_step = reinterpret_cast<unsigned long long*>(global_step);
  a.cursor = cursor; a.cursor_step = cursor_step; a.cursor_wrap = cursor_wrap > 0 ? cursor_wrap : 1;
  a.world = world; a.rank = rank; a.broadcast = broadcast; a.do_reduce = do_reduce;
  const int per = dtype == B200PS_DT_F32 ? 4 : 8;
  const long long nvec = (hi - lo) / per;
  // every rank launches at least one CTA: the flag exchanges need all of them even when a slice is empty
  // as many CTAs as the P2P kernel by default (one per SM, 256 threads); B200PS_NVLS_CTAS caps it for pushes that
  // run on a side stream under GEMMs (read per launch: an experiment knob)
  const char* cap_env = getenv("B200PS_NVLS_CTAS");
  const int cap = cap_env && atoi(cap_env) > 0 ? atoi(cap_env) : sm_count();
  const int ctas = (int)std::max<long long>(1, std::min<long long>((nvec + 1023) / 1024, (long long)cap));
  cudaStream_t s = as_stream(stream);
  if (dtype == B200PS_DT_F32)
    B200PS_CUDA_CHECK(launch_pdl(store_nvls_kernel<float, float>, dim3(ctas), dim3(256), 0, s, a));
  else
    B200PS_CUDA_CHECK(launch_pdl(store_nvls_kernel<__nv_bfloat16, __nv_bfloat16>, dim3(ctas), dim3(256), 0, s, a));
  B200PS_LAUNCH_CHECK();
  return 0;
}

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
