// Micro-benchmarks behind the cluster formulation of the one-launch MLP step (K14):
//   - does a cluster of 16 (non-portable) / 8 CTAs with ~200 KB of shared memory per CTA launch?
//   - barrier.cluster arrive+wait latency
//   - DSMEM read / write throughput per CTA (ld/st.shared::cluster.v4.f32, all CTAs active)
//   - one-way signalling latency through a remote mbarrier arrive
//   - mma.sync.m16n8k8 tf32 issue rate of one warp / of 8 warps
//   - cp.async.bulk of 100 rows x 784 B from L2 into shared memory
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o cluster_probe cluster_probe.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstdlib>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s failed: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t cluster_rank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ uint32_t cluster_size() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t mapa(uint32_t addr, uint32_t rank) {
  uint32_t r; asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank)); return r;
}
__device__ __forceinline__ float4 ld_dsmem4(uint32_t a) {
  float4 v; asm volatile("ld.shared::cluster.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a)); return v;
}
__device__ __forceinline__ void st_dsmem4(uint32_t a, float4 v) {
  asm volatile("st.shared::cluster.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(a), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

struct Out { long long bar, rd, wr, sig, mma1, mma8, bulk; float sink; };

// mode bits: kernel measures everything in sequence; KB = bytes each CTA reads/writes remotely
__global__ void __launch_bounds__(256) probe(Out* out, const float* g, int kb, int iters) {
  extern __shared__ __align__(16) float sm[];
  __shared__ __align__(8) uint64_t mbar[2];
  const uint32_t rank = cluster_rank(), cs = cluster_size();
  const int tid = threadIdx.x;
  for (int i = tid; i < kb * 256; i += 256) sm[i] = (float)(i + rank);
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&mbar[0])));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&mbar[1])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  cluster_sync();
  // ---- cluster barrier latency
  long long t0 = clock64();
  for (int i = 0; i < iters; ++i) cluster_sync();
  long long t1 = clock64();
  const long long bar = (t1 - t0) / iters;
  // ---- DSMEM reads: thread reads float4 i of peer (rank + 1 + i / chunk) ...
  float acc = 0.f;
  const int n4 = kb * 64;                     // float4 per CTA
  cluster_sync();
  t0 = clock64();
  {
    const int per_peer = n4 / (cs - 1);
    for (int i = tid; i < per_peer * (cs - 1); i += 256 * 4) {
      float4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int j = i + u * 256;
        const int peer = (rank + 1 + j / per_peer) % cs;
        v[u] = j < per_peer * (cs - 1) ? ld_dsmem4(mapa(smem_u32(sm + 4 * (j % per_peer)), peer)) : make_float4(0, 0, 0, 0);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) acc += v[u].x + v[u].y + v[u].z + v[u].w;
    }
  }
  __syncthreads();
  t1 = clock64();
  const long long rd = t1 - t0;
  cluster_sync();
  // ---- DSMEM writes
  t0 = clock64();
  {
    const int per_peer = n4 / (cs - 1);
    for (int j = tid; j < per_peer * (cs - 1); j += 256) {
      const int peer = (rank + 1 + j / per_peer) % cs;
      st_dsmem4(mapa(smem_u32(sm + 4 * (j % per_peer) + 4 * per_peer * 0), peer), make_float4(acc, 1.f, 2.f, 3.f));
    }
  }
  __syncthreads();
  cluster_sync();
  t1 = clock64();
  const long long wr = t1 - t0;
  // ---- one-way signal: rank r arrives on mbar[0] of rank r+1 ; each waits for its own (ring) `iters` times
  long long sig = 0;
  if (tid == 0) {
    const uint32_t remote = mapa(smem_u32(&mbar[0]), (rank + 1) % cs);
    t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      // ping-pong around the ring: rank 0 starts, everyone forwards after receiving
      if (!(rank == 0)) {
        uint32_t ok = 0;
        while (!ok) asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0,1,0,p;}" : "=r"(ok) : "r"(smem_u32(&mbar[0])), "r"(i & 1) : "memory");
      }
      asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
      if (rank == 0) {
        uint32_t ok = 0;
        while (!ok) asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0,1,0,p;}" : "=r"(ok) : "r"(smem_u32(&mbar[0])), "r"(i & 1) : "memory");
      }
    }
    t1 = clock64();
    sig = (t1 - t0) / iters / cs;             // per hop
  }
  __syncthreads();
  cluster_sync();
  // ---- mma.sync tf32 m16n8k8: 13 independent accumulators (the n-tiles of one m-tile), 64 rounds
  float c[13][4];
#pragma unroll
  for (int n = 0; n < 13; ++n) c[n][0] = c[n][1] = c[n][2] = c[n][3] = 0.f;
  uint32_t a0 = __float_as_uint(sm[tid]), a1 = __float_as_uint(sm[tid + 1]), a2 = a0 ^ 1, a3 = a1 ^ 1, b0 = a0 ^ 2, b1 = a1 ^ 2;
  auto mma_round = [&]() {
#pragma unroll
    for (int n = 0; n < 13; ++n)
      asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                   : "+f"(c[n][0]), "+f"(c[n][1]), "+f"(c[n][2]), "+f"(c[n][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  };
  __syncthreads();
  long long mma1 = 0, mma8 = 0;
  if (tid < 32) {
    t0 = clock64();
    for (int i = 0; i < 64; ++i) mma_round();
    t1 = clock64();
    mma1 = t1 - t0;
  }
  __syncthreads();
  t0 = clock64();
  for (int i = 0; i < 64; ++i) mma_round();
  __syncthreads();
  t1 = clock64();
  mma8 = t1 - t0;
#pragma unroll
  for (int n = 0; n < 13; ++n) acc += c[n][0] + c[n][1] + c[n][2] + c[n][3];
  // ---- bulk copy: 100 rows x 784 B from global (L2-resident after the first round)
  long long bulk = 0;
  for (int round = 0; round < 3; ++round) {
    __syncthreads();
    t0 = clock64();
    if (tid == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&mbar[1])), "r"(100 * 784) : "memory");
    __syncthreads();
    if (tid < 100)
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(smem_u32(sm + tid * 196)), "l"(g + (size_t)tid * 784 + rank * 196), "r"(784), "r"(smem_u32(&mbar[1])) : "memory");
    {
      uint32_t ok = 0;
      while (!ok) asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0,1,0,p;}" : "=r"(ok) : "r"(smem_u32(&mbar[1])), "r"(round & 1) : "memory");
    }
    t1 = clock64();
    bulk = t1 - t0;
  }
  acc += sm[tid];
  cluster_sync();
  if (tid == 0) {
    Out* o = out + blockIdx.x;
    o->bar = bar; o->rd = rd; o->wr = wr; o->sig = sig; o->mma1 = mma1; o->mma8 = mma8; o->bulk = bulk; o->sink = acc;
  }
}

int main(int argc, char** argv) {
  const int kb = argc > 1 ? atoi(argv[1]) : 40;
  Out* out; float* g;
  CK(cudaMalloc(&out, sizeof(Out) * 64));
  CK(cudaMalloc(&g, 4 * 784 * 128));
  CK(cudaMemset(g, 0, 4 * 784 * 128));
  for (int cs : {8, 16}) {
    for (int smem_kb : {100, 200}) {
      const size_t smem = (size_t)smem_kb * 1024;
      CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      if (cs > 8) CK(cudaFuncSetAttribute(probe, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
      cudaLaunchConfig_t cfg{};
      cfg.gridDim = dim3(cs); cfg.blockDim = dim3(256); cfg.dynamicSmemBytes = smem;
      cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = cs; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
      cfg.attrs = at; cfg.numAttrs = 1;
      int nclusters = -1;
      cudaError_t e = cudaOccupancyMaxActiveClusters(&nclusters, probe, &cfg);
      printf("cluster=%d smem=%dKB: maxActiveClusters=%d (%s)\n", cs, smem_kb, nclusters, cudaGetErrorString(e));
      if (e != cudaSuccess || nclusters < 1) { cudaGetLastError(); continue; }
      for (int rep = 0; rep < 2; ++rep) {
        e = cudaLaunchKernelEx(&cfg, probe, out, (const float*)g, kb, 50);
        if (e != cudaSuccess) { printf("  launch failed: %s\n", cudaGetErrorString(e)); cudaGetLastError(); break; }
        CK(cudaDeviceSynchronize());
      }
      Out h[16];
      CK(cudaMemcpy(h, out, sizeof(Out) * cs, cudaMemcpyDeviceToHost));
      long long rd = 0, wr = 0, bulk = 0;
      for (int i = 0; i < cs; ++i) { rd = rd > h[i].rd ? rd : h[i].rd; wr = wr > h[i].wr ? wr : h[i].wr; bulk = bulk > h[i].bulk ? bulk : h[i].bulk; }
      printf("  cluster.sync %lld cyc | DSMEM read %d KB/CTA: %lld cyc (%.1f B/cyc) | write(+sync): %lld cyc (%.1f B/cyc) | remote mbarrier hop %lld cyc\n",
             h[0].bar, kb, rd, kb * 1024.0 / rd, wr, kb * 1024.0 / wr, h[0].sig);
      printf("  mma.sync tf32 m16n8k8 x832: one warp %lld cyc (%.2f cyc/mma), 8 warps %lld cyc (%.2f cyc/mma/warp) | bulk 100x784B: %lld cyc\n",
             h[0].mma1, h[0].mma1 / 832.0, h[0].mma8, h[0].mma8 / 832.0, bulk);
    }
  }
  return 0;
}
