// b200ps K1/K3/K12 + pull — the parameter-server "wire", as kernels over NVLink peer memory.
//
// In the reference every sess.run partitions the graph into a worker part and a ps part
// joined by gRPC Send/Recv: variables are pulled (example.py:95,97), gradients pushed and
// ApplyGradientDescent runs on the ps (example.py:108,118).  Here each GPU is worker k AND
// parameter-server shard k.  All ranks hold peer-mapped pointers to each other's buffers
// ("symmetric memory") and the kernels below issue the traffic themselves:
//
//   push_sync  (K1): reduce-scatter of the replicas' gradients (the shard owner loads its
//                    slice from every replica in a fixed order), mean over the aggregated
//                    replicas, SGD apply on the owner's fp32 master, and broadcast of the
//                    fresh values into every replica's working copy (the next "pull" is
//                    already done).  Two flag exchanges bracket it: "my gradients are
//                    ready" and "my slice has landed everywhere" — together they are the
//                    token barrier of SyncReplicasOptimizer (example.py:110-123,147-152).
//   push_async (K3): Hogwild.  -lr*g is added straight into the owner's master with
//                    red.global.add.v4.f32 at system scope; no flag, no barrier;
//                    global_step += 1 on the chief's counter (example.py:118, README.md:3).
//   pull            : copy every peer-owned slice into the local working copy (stale reads
//                    are allowed in async mode, exactly like un-locked variable reads).
//   barrier   (K12): stand-alone flag barrier (Supervisor ready barrier, done tokens).
//
// Several "virtual ranks" can be driven by one launch (blockIdx.y = local rank): that is
// how the multi-rank protocol is exercised on a single GPU, where ranks that wait on each
// other must not be separate launches.
#include "../common.h"
#include "ptx.cuh"
#include "ll_proto.cuh"

namespace b200ps {

struct CommArgs {
  uint32_t* const* pads;      // [world] signal pads
  float* const* masters;      // [world] fp32 master, indexable by GLOBAL element index on its owner
  void* const* outs;          // [world] working copies (full size), dtype O
  const void* const* grads;   // [world] gradient buffers (full size), dtype G
  const long long* bounds;    // [world+1] owner ranges (multiples of 8 elements)
  unsigned long long* const* global_steps;  // [world] per-rank int64 counters
  int world, rank_base;
  int replicas_to_aggregate;
  float lr;
  int* cursor;                // resident-dataset batch pointer advanced at the end of the step
  int broadcast;              // 1: write fresh values into every replica; 0: owner's copy only (pull is fused elsewhere)
  int cursor_step, cursor_wrap;
};

template <typename T> struct Vec4;
template <> struct Vec4<float> {
  // L2-only (never a stale L1 line; peer data is not cached in the local L2 at all); ordered
  // after the flag acquire, and — unlike a volatile sys-scope load — free to be batched
  static __device__ __forceinline__ float4 load(const float* p) {
    return __ldcg(reinterpret_cast<const float4*>(p));
  }
  static __device__ __forceinline__ void store(float* p, float4 v) {
    *reinterpret_cast<float4*>(p) = v;
  }
};
template <> struct Vec4<__nv_bfloat16> {
  static __device__ __forceinline__ float4 load(const __nv_bfloat16* p) {
    uint2 u = __ldcg(reinterpret_cast<const uint2*>(p));
    const float2 a = __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&u.x));
    const float2 b = __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&u.y));
    return make_float4(a.x, a.y, b.x, b.y);
  }
  static __device__ __forceinline__ void store(__nv_bfloat16* p, float4 v) {
    uint2 u;
    *reinterpret_cast<__nv_bfloat162*>(&u.x) = __floats2bfloat162_rn(v.x, v.y);
    *reinterpret_cast<__nv_bfloat162*>(&u.y) = __floats2bfloat162_rn(v.z, v.w);
    *reinterpret_cast<uint2*>(p) = u;
  }
};

// Is replica r part of the aggregated set of this step?  With replicas_to_aggregate == world
// every replica is.  With fewer ("backup workers") TF takes the first R gradients to arrive
// and drops the rest; arrival order is not reproducible, so the set rotates with the step.
__device__ __forceinline__ bool replica_counts(int r, uint32_t epoch, int world, int R) {
  if (R >= world) return true;
  const int d = (r - (int)(epoch % (uint32_t)world) + world) % world;
  return d < R;
}

template <typename G, typename O>
__global__ void __launch_bounds__(1024) push_sync_sgd_kernel(const CommArgs a) {
  const int rank = a.rank_base + blockIdx.y;
  const int world = a.world;
  const int tid = threadIdx.x;
  uint32_t* pad = a.pads[rank];
  pdl_wait();      // the gradients come from the kernels just before us
  pdl_trigger();
  const uint32_t epoch = pad[kSlotEpoch] + 1;

  // A single shard has nobody to wait for: no flags, no system-scope fences, no arrival count.
  const bool solo = (world == 1);

  // -- phase 0: my gradients (written by earlier kernels of this stream) are ready
  // (relaxed store: the gradients were written by EARLIER kernels, i.e. are already visible)
  if (!solo) {
    if (blockIdx.x == 0 && tid < world) st_relaxed_sys(a.pads[tid] + kFlagGradsReady + rank, epoch);
    if (tid < world) wait_flag_geq(pad + kFlagGradsReady + tid, epoch);
    __syncthreads();
  }

  // -- phase 1: reduce my slice over the replicas, apply, broadcast
  const long long lo = a.bounds[rank], hi = a.bounds[rank + 1];
  const float step = a.lr / (float)min(a.replicas_to_aggregate, world);
  float* master = a.masters[rank];
  const long long nvec = (hi - lo) >> 2;
  const long long stride = (long long)gridDim.x * blockDim.x;
  constexpr int kU = 4;   // independent float4s per thread per trip: all loads issue before any use
  for (long long v0 = (long long)blockIdx.x * blockDim.x + tid; v0 < nvec; v0 += kU * stride) {
    float4 s[kU], w[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      s[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      const long long v = v0 + u * stride;
      if (v < nvec) w[u] = *reinterpret_cast<const float4*>(master + lo + (v << 2));
    }
    for (int r = 0; r < world; ++r) {  // fixed order => bitwise reproducible
      if (!replica_counts(r, epoch, world, a.replicas_to_aggregate)) continue;
      const G* gsrc = reinterpret_cast<const G*>(a.grads[r]) + lo;
      float4 g[kU];
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const long long v = v0 + u * stride;
        if (v < nvec) g[u] = Vec4<G>::load(gsrc + (v << 2));
      }
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const long long v = v0 + u * stride;
        if (v < nvec) { s[u].x += g[u].x; s[u].y += g[u].y; s[u].z += g[u].z; s[u].w += g[u].w; }
      }
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const long long v = v0 + u * stride;
      if (v >= nvec) continue;
      const long long i = lo + (v << 2);
      float4 x = w[u];
      x.x -= step * s[u].x; x.y -= step * s[u].y; x.z -= step * s[u].z; x.w -= step * s[u].w;
      *reinterpret_cast<float4*>(master + i) = x;
      for (int r = 0; r < world; ++r) {
        if (!a.broadcast && r != rank) continue;
        O* dst = reinterpret_cast<O*>(a.outs[r]) + i;
        if (reinterpret_cast<void*>(dst) != reinterpret_cast<void*>(master + i)) Vec4<O>::store(dst, x);
      }
    }
  }

  if (solo) {
    // the step counters are only read by later kernels / the host: any one thread may bump them
    if (blockIdx.x == 0 && tid == 0) {
      pad[kSlotEpoch] = epoch;
      if (a.global_steps) *a.global_steps[rank] += 1ull;
      if (a.cursor) *a.cursor = (*a.cursor + a.cursor_step) % a.cursor_wrap;
    }
    return;
  }

  // -- phase 2: when every CTA of this rank is done, tell the peers and wait for theirs.
  // One system-scope fence per CTA (thread 0, after the block barrier: fences are cumulative
  // over everything that happened-before them) instead of one per thread.
  __syncthreads();
  __shared__ bool s_last;
  if (tid == 0) {
    if (gridDim.x == 1) {
      s_last = true;   // a small shard is handled by ONE wide CTA: no arrival counting at all
    } else {
      __threadfence();   // gpu scope here; the last CTA issues the one system-scope fence
      s_last = (atomicAdd(pad + kSlotCtaDone, 1u) == gridDim.x - 1);
    }
  }
  __syncthreads();
  if (s_last) {
    if (tid == 0) __threadfence_system();   // the one system-scope fence ordering every CTA's stores
    __syncthreads();
    if (tid < world) st_relaxed_sys(a.pads[tid] + kFlagParamsDone + rank, epoch);
    if (tid < world) wait_flag_geq(pad + kFlagParamsDone + tid, epoch);
    __syncthreads();
    if (tid == 0) {
      pad[kSlotCtaDone] = 0;
      pad[kSlotEpoch] = epoch;
      if (a.global_steps) *a.global_steps[rank] += 1ull;  // one step per aggregated update
      if (a.cursor && blockIdx.y == 0) *a.cursor = (*a.cursor + a.cursor_step) % a.cursor_wrap;
    }
  }
}

// ------------------------------------------------------------------------------------
// push_sync, low-latency form for small fp32 models (the MNIST MLP: 80 k parameters).
//
// The flag-based kernel above pays two flag round trips, a remote read round trip and a
// system-scope fence per step — ~10 us on NVLink regardless of size.  Here every 4-byte value
// travels in an 8-byte store together with the step's epoch ("LL" packets): the receiver polls
// the packet itself, so each hop costs ONE one-way NVLink latency and no fence is needed
// (an aligned 8-byte store is single-copy atomic).
//
//   phase 1  every rank stores its gradient slice for owner o into o's landing buffer
//   phase 2  the owner polls the packets of all replicas (fixed order), applies SGD to its fp32
//            master, and stores the fresh value as a packet into every replica's landing buffer
//   phase 3  every rank polls the packets of the other owners and unpacks them into its plain
//            fp32 working copy (what the TMA-fed GEMMs read)
//
// Epoch tags make clearing unnecessary; a slot is only rewritten after its previous packet was
// consumed (the sender cannot finish step e before the receiver's phase 2/3 of step e sent it
// the data it is waiting for).  Semantics are those of push_sync_sgd with all replicas
// aggregated: mean gradient, one global step, nobody continues before the update has landed.
// ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) push_sync_ll_kernel(const CommArgs a, uint2* const* ll_grads,
                                                          uint2* const* ll_params, long long max_shard) {
  const int rank = a.rank_base + blockIdx.y;
  const int world = a.world;
  const int tid = threadIdx.x;
  uint32_t* pad = a.pads[rank];
  pdl_wait();
  pdl_trigger();
  const uint32_t epoch = pad[kSlotEpoch] + 1;
  const float* g = reinterpret_cast<const float*>(a.grads[rank]);
  float* work = a.masters[rank];                       // fp32 group: master == working copy
  const long long gtid = (long long)blockIdx.x * blockDim.x + tid;
  const long long gstride = (long long)gridDim.x * blockDim.x;
  const long long total = a.bounds[world];
  const long long lo = a.bounds[rank], hi = a.bounds[rank + 1];

  // phase 1: my gradient slices -> their owners (walk the whole flat space, skip my own range)
  for (long long i = gtid; i < total; i += gstride) {
    if (i >= lo && i < hi) continue;
    int o = 0;
    while (i >= a.bounds[o + 1]) ++o;
    st_packet(ll_grads[o] + (long long)rank * max_shard + (i - a.bounds[o]), g[i], epoch);
  }
  // phase 2: reduce my range over the replicas in rank order, apply, publish
  const float step = a.lr / (float)world;
  for (long long i = lo + gtid; i < hi; i += gstride) {
    float sum = 0.f;
    for (int r = 0; r < world; ++r)
      sum += (r == rank) ? g[i] : poll_packet(ll_grads[rank] + (long long)r * max_shard + (i - lo), epoch);
    const float w = work[i] - step * sum;
    work[i] = w;
    for (int p = 0; p < world; ++p)
      if (p != rank) st_packet(ll_params[p] + i, w, epoch);
  }
  // phase 3: the other owners' fresh values -> my plain fp32 working copy
  for (long long i = gtid; i < total; i += gstride) {
    if (i >= lo && i < hi) continue;
    work[i] = poll_packet(ll_params[rank] + i, epoch);
  }
  // bookkeeping by the last CTA of this rank (gpu-scope only: nothing here crosses the link)
  __syncthreads();
  if (tid == 0) {
    __threadfence();
    if (atomicAdd(pad + kSlotCtaDone, 1u) == gridDim.x - 1) {
      pad[kSlotCtaDone] = 0;
      pad[kSlotEpoch] = epoch;
      if (a.global_steps) *a.global_steps[rank] += 1ull;
      if (a.cursor && blockIdx.y == 0) *a.cursor = (*a.cursor + a.cursor_step) % a.cursor_wrap;
    }
  }
}

__global__ void __launch_bounds__(256) push_async_kernel(const CommArgs a, long long n) {
  const int rank = a.rank_base + blockIdx.y;
  pdl_wait();      // the gradients come from the kernels just before us
  pdl_trigger();
  const float* g = reinterpret_cast<const float*>(a.grads[rank]);
  const float nlr = -a.lr;
  for (int r = 0; r < a.world; ++r) {
    const long long lo = a.bounds[r], hi = a.bounds[r + 1];
    float* master = a.masters[r];
    const long long nvec = (hi - lo) >> 2;
    for (long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x; v < nvec;
         v += (long long)gridDim.x * blockDim.x) {
      const long long i = lo + (v << 2);
      const float4 gv = *reinterpret_cast<const float4*>(g + i);
      red_add_v4f32_sys(master + i, make_float4(nlr * gv.x, nlr * gv.y, nlr * gv.z, nlr * gv.w));
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    if (a.global_steps) atomicAdd_system(a.global_steps[0], 1ull);  // the chief's shard holds global_step
    // (safe to advance here: every kernel that reads the cursor in this step has already completed)
    if (a.cursor) *a.cursor = (*a.cursor + a.cursor_step) % a.cursor_wrap;
  }
}

template <typename O>
__global__ void __launch_bounds__(256) pull_kernel(const CommArgs a, int include_self) {
  const int rank = a.rank_base + blockIdx.y;
  pdl_wait();
  pdl_trigger();
  O* dst = reinterpret_cast<O*>(a.outs[rank]);
  for (int rr = 0; rr < a.world; ++rr) {
    const int r = (rank + rr) % a.world;  // start with the local slice, then walk the peers
    if (r == rank && !include_self) continue;
    const long long lo = a.bounds[r], hi = a.bounds[r + 1];
    const float* src = a.masters[r];
    const long long nvec = (hi - lo) >> 2;
    const long long stride = (long long)gridDim.x * blockDim.x;
    if (reinterpret_cast<const void*>(dst + lo) == reinterpret_cast<const void*>(src + lo)) continue;
    constexpr int kU = 8;   // independent 16-byte loads in flight per thread (peer latency is ~2 us)
    for (long long v0 = (long long)blockIdx.x * blockDim.x + threadIdx.x; v0 < nvec; v0 += kU * stride) {
      float4 x[kU];
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const long long v = v0 + u * stride;
        if (v < nvec) x[u] = __ldcg(reinterpret_cast<const float4*>(src + lo + (v << 2)));
      }
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const long long v = v0 + u * stride;
        if (v < nvec) Vec4<O>::store(dst + lo + (v << 2), x[u]);
      }
    }
  }
}

__global__ void __launch_bounds__(32) barrier_kernel(uint32_t* const* pads, int world, int rank_base) {
  const int rank = rank_base + blockIdx.y;
  uint32_t* pad = pads[rank];
  const uint32_t epoch = pad[kSlotBarrierEpoch] + 1;
  const int tid = threadIdx.x;
  __threadfence_system();
  if (tid < world) st_release_sys(pads[tid] + kFlagBarrier + rank, epoch);
  if (tid < world) wait_flag_geq(pad + kFlagBarrier + tid, epoch);
  __syncwarp();
  if (tid == 0) pad[kSlotBarrierEpoch] = epoch;
}

}  // namespace b200ps

using namespace b200ps;

static int comm_grid(long long max_elems, int local_ranks, int want) {
  // ~4 float4 per thread: few CTAs for small payloads (fences/atomics are per CTA)
  int ctas = (int)std::min<long long>(std::max<long long>((max_elems / 16 + 255) / 256, 1), want);
  if (local_ranks > 1) ctas = std::max(1, std::min(ctas, sm_count() / local_ranks));  // co-resident
  return ctas;
}

extern "C" int b200ps_push_sync_sgd(uint32_t* const* pads, float* const* masters, void* const* outs,
                                    const void* const* grads, const int64_t* bounds,
                                    uint64_t* const* global_steps, int32_t world, int32_t rank_base,
                                    int32_t local_ranks, int64_t max_shard_elems, int32_t grad_dtype,
                                    int32_t out_dtype, int32_t replicas_to_aggregate, float lr,
                                    int32_t max_ctas, int32_t* cursor, int32_t cursor_step,
                                    int32_t cursor_wrap, int32_t broadcast, b200ps_stream_t stream) {
  B200PS_REQUIRE(world >= 1 && world <= kMaxWorld, "push_sync: world=%d out of range", world);
  B200PS_REQUIRE(local_ranks >= 1 && rank_base >= 0 && rank_base + local_ranks <= world, "push_sync: bad ranks");
  B200PS_REQUIRE(grad_dtype == out_dtype, "push_sync: gradient and working-copy dtypes must match");
  B200PS_REQUIRE(replicas_to_aggregate >= 1, "push_sync: replicas_to_aggregate < 1");
  CommArgs a;
  a.pads = pads; a.masters = masters; a.outs = outs; a.grads = grads;
  a.bounds = reinterpret_cast<const long long*>(bounds);
  a.global_steps = reinterpret_cast<unsigned long long* const*>(global_steps);
  a.world = world; a.rank_base = rank_base; a.replicas_to_aggregate = replicas_to_aggregate; a.lr = lr;
  a.cursor = cursor; a.cursor_step = cursor_step; a.cursor_wrap = cursor_wrap > 0 ? cursor_wrap : 1;
  a.broadcast = broadcast ? 1 : 0;
  const int want = max_ctas > 0 ? max_ctas : sm_count();
  // latency-bound regime (small shards, several ranks): one 1024-thread CTA per rank does the whole
  // shard in one trip of 4 float4 per thread, and the per-CTA fence/atomic/last-CTA machinery vanishes
  const bool one_cta = world > 1 && max_shard_elems <= 16384;
  dim3 grid(one_cta ? 1 : comm_grid(max_shard_elems, local_ranks, want), local_ranks);
  dim3 block(one_cta ? 1024 : 256);
  cudaStream_t s = as_stream(stream);
  if (grad_dtype == B200PS_DT_F32)
    B200PS_CUDA_CHECK(launch_pdl(push_sync_sgd_kernel<float, float>, grid, block, 0, s, a));
  else
    B200PS_CUDA_CHECK(launch_pdl(push_sync_sgd_kernel<__nv_bfloat16, __nv_bfloat16>, grid, block, 0, s, a));
  B200PS_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200ps_push_sync_ll(uint32_t* const* pads, float* const* masters, const void* const* grads,
                                   const int64_t* bounds, uint64_t* const* global_steps, void* const* ll_grads,
                                   void* const* ll_params, int32_t world, int32_t rank_base, int32_t local_ranks,
                                   int64_t max_shard_elems, int64_t total_elems, float lr, int32_t* cursor,
                                   int32_t cursor_step, int32_t cursor_wrap, b200ps_stream_t stream) {
  B200PS_REQUIRE(world >= 2 && world <= kMaxWorld, "push_sync_ll: world=%d out of range (needs >= 2)", world);
  B200PS_REQUIRE(local_ranks >= 1 && rank_base >= 0 && rank_base + local_ranks <= world, "push_sync_ll: bad ranks");
  CommArgs a{};
  a.pads = pads; a.masters = masters; a.grads = grads;
  a.bounds = reinterpret_cast<const long long*>(bounds);
  a.global_steps = reinterpret_cast<unsigned long long* const*>(global_steps);
  a.world = world; a.rank_base = rank_base; a.replicas_to_aggregate = world; a.lr = lr;
  a.cursor = cursor; a.cursor_step = cursor_step; a.cursor_wrap = cursor_wrap > 0 ? cursor_wrap : 1;
  int ctas = (int)std::min<long long>(std::max<long long>((total_elems + 255) / 256, 1), sm_count());
  if (local_ranks > 1) ctas = std::max(1, std::min(ctas, sm_count() / local_ranks));   // co-resident virtual ranks
  B200PS_CUDA_CHECK(launch_pdl(push_sync_ll_kernel, dim3(ctas, local_ranks), dim3(256), 0, as_stream(stream), a,
                               reinterpret_cast<uint2* const*>(ll_grads), reinterpret_cast<uint2* const*>(ll_params),
                               (long long)max_shard_elems));
  B200PS_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200ps_push_async(float* const* masters, const void* const* grads, const int64_t* bounds,
                                 uint64_t* const* global_steps, int32_t world, int32_t rank_base,
                                 int32_t local_ranks, int64_t total_elems, float lr, int32_t max_ctas,
                                 int32_t* cursor, int32_t cursor_step, int32_t cursor_wrap,
                                 b200ps_stream_t stream) {
  B200PS_REQUIRE(world >= 1 && world <= kMaxWorld, "push_async: world=%d out of range", world);
  CommArgs a{};
  a.masters = masters; a.grads = grads; a.bounds = reinterpret_cast<const long long*>(bounds);
  a.global_steps = reinterpret_cast<unsigned long long* const*>(global_steps);
  a.world = world; a.rank_base = rank_base; a.lr = lr;
  a.cursor = cursor; a.cursor_step = cursor_step; a.cursor_wrap = cursor_wrap > 0 ? cursor_wrap : 1;
  const int want = max_ctas > 0 ? max_ctas : sm_count();
  dim3 grid(comm_grid(total_elems / std::max(world, 1), 1, want), local_ranks);
  B200PS_CUDA_CHECK(launch_pdl(push_async_kernel, grid, dim3(256), 0, as_stream(stream), a, (long long)total_elems));
  B200PS_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200ps_pull(float* const* masters, void* const* outs, const int64_t* bounds, int32_t world,
                           int32_t rank_base, int32_t local_ranks, int64_t max_shard_elems,
                           int32_t out_dtype, int32_t include_self, int32_t max_ctas,
                           b200ps_stream_t stream) {
  B200PS_REQUIRE(world >= 1 && world <= kMaxWorld, "pull: world=%d out of range", world);
  CommArgs a{};
  a.masters = masters; a.outs = outs; a.bounds = reinterpret_cast<const long long*>(bounds);
  a.world = world; a.rank_base = rank_base;
  const int want = max_ctas > 0 ? max_ctas : sm_count();
  dim3 grid(comm_grid(max_shard_elems, 1, want), local_ranks);
  if (out_dtype == B200PS_DT_F32)
    B200PS_CUDA_CHECK(launch_pdl(pull_kernel<float>, grid, dim3(256), 0, as_stream(stream), a, (int)include_self));
  else
    B200PS_CUDA_CHECK(launch_pdl(pull_kernel<__nv_bfloat16>, grid, dim3(256), 0, as_stream(stream), a, (int)include_self));
  B200PS_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200ps_barrier(uint32_t* const* pads, int32_t world, int32_t rank_base, int32_t local_ranks,
                              b200ps_stream_t stream) {
  B200PS_REQUIRE(world >= 1 && world <= kMaxWorld, "barrier: world=%d out of range", world);
  barrier_kernel<<<dim3(1, local_ranks), 32, 0, as_stream(stream)>>>(pads, world, rank_base);
  B200PS_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200ps_signal_pad_words(void) { return kPadWords; }
