// Micro-benchmark behind the design of the fused MLP step (csrc/kernels/mlp_fused.cu): how long does one
// CTA per SM (256 threads) need to pull N KB that another kernel just wrote, by access pattern; and what
// does a device-wide barrier cost.  Prints SM-clock cycles (median over CTAs of the per-CTA time, and max).
#include <cuda_runtime.h>
#include <algorithm>
#include <cstdio>
#include <vector>

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void cp16(void* s, const void* g) { asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(s)), "l"(g) : "memory"); }
__device__ __forceinline__ void cp4(void* s, const void* g) { asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(s)), "l"(g) : "memory"); }
__device__ __forceinline__ void cp_commit_wait() { asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ unsigned ld_acq(const unsigned* p) { unsigned v; asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }

__global__ void fill(float* p, size_t n) { for (size_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = (float)(i & 1023); }

// mode 0: cp.async 16B contiguous (every CTA the same `bytes`)   1: same, each CTA its own region
// mode 2: ld.global.cg.v4 into registers, 8 per batch              3: cp.async 4B
// mode 4: ld.global.cg.f32, 32 per lane in one batch (slab pattern: 8 warps x rows of 400 B, stride 40 KB)
// mode 5: bulk copy (cp.async.bulk) of `bytes` by one thread
__global__ void __launch_bounds__(256) fetch(const float* src, int bytes, int mode, long long* out, float* sink) {
  extern __shared__ __align__(128) float sm[];
  __shared__ __align__(8) unsigned long long bar;
  const int tid = threadIdx.x;
  const int nf4 = bytes / 16;
  if (mode == 5 && tid == 0) { asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar))); asm volatile("fence.mbarrier_init.release.cluster;"); }
  __syncthreads();
  const long long t0 = clock64();
  float acc = 0.f;
  if (mode == 0 || mode == 1) {
    const float* s = src + (mode == 1 ? (size_t)blockIdx.x * (bytes / 4) : 0);
    for (int i = tid; i < nf4; i += 256) cp16(sm + 4 * i, s + 4 * i);
    cp_commit_wait();
  } else if (mode == 2) {
    const float4* s = reinterpret_cast<const float4*>(src);
    for (int i0 = tid; i0 < nf4; i0 += 256 * 8) {
      float4 v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = (i0 + u * 256 < nf4) ? __ldcg(s + i0 + u * 256) : make_float4(0, 0, 0, 0);
#pragma unroll
      for (int u = 0; u < 8; ++u) if (i0 + u * 256 < nf4) reinterpret_cast<float4*>(sm)[i0 + u * 256] = v[u];
    }
  } else if (mode == 3) {
    for (int i = tid; i < bytes / 4; i += 256) cp4(sm + i, src + i);
    cp_commit_wait();
  } else if (mode == 4) {
    const int lane = tid & 31, warp = tid >> 5;
    const float* s = src + (size_t)(blockIdx.x * 8 + warp) * 100 + lane;
    float v[32];
    const int nb = bytes / (8 * 400);   // "slabs"
#pragma unroll
    for (int u = 0; u < 8; ++u)
#pragma unroll
      for (int k = 0; k < 4; ++k) v[u * 4 + k] = (u < nb) ? __ldcg(s + (size_t)u * 10000 + min(32 * k, 68)) : 0.f;
#pragma unroll
    for (int u = 0; u < 32; ++u) acc += v[u];
  } else if (mode == 5) {
    if (tid == 0) {
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(bytes) : "memory");
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(sm)), "l"(src), "r"(bytes), "r"(smem_u32(&bar)) : "memory");
    }
    unsigned done = 0;
    while (!done) asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(smem_u32(&bar)) : "memory");
  }
  __syncthreads();
  const long long t1 = clock64();
  if (tid == 0) out[blockIdx.x] = t1 - t0;
  if (acc == 123.456f) sink[0] = acc + sm[tid];
}

// write-then-read inside ONE kernel: every CTA writes its share of `bytes`, device-wide barrier, then every CTA
// fetches all of it (mode 0: cp.async16, 1: bulk copy, 2: ld.cg.v4 batches, 3: plain ld.v4 batches).
__global__ void __launch_bounds__(256) wr_fetch(float* buf, int bytes, int mode, unsigned* ctr, long long* out, float* sink) {
  extern __shared__ __align__(128) float sm[];
  __shared__ __align__(8) unsigned long long bar;
  const int tid = threadIdx.x, G = gridDim.x;
  const int nf4 = bytes / 16;
  if (tid == 0) { asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar))); asm volatile("fence.mbarrier_init.release.cluster;"); }
  for (int i = blockIdx.x * 256 + tid; i < nf4; i += G * 256) reinterpret_cast<float4*>(buf)[i] = make_float4(i, 1, 2, 3);
  if (mode == 1) asm volatile("fence.proxy.async;" ::: "memory");
  __syncthreads();
  if (tid == 0) {
    asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(ctr) : "memory");
    while (ld_acq(ctr) < (unsigned)G) {}
  }
  __syncthreads();
  const long long t0 = clock64();
  float acc = 0.f;
  if (mode == 0) {
    for (int i = tid; i < nf4; i += 256) cp16(sm + 4 * i, buf + 4 * i);
    cp_commit_wait();
  } else if (mode == 1) {
    if (tid == 0) {
      asm volatile("fence.proxy.async;" ::: "memory");
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(bytes) : "memory");
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(sm)), "l"(buf), "r"(bytes), "r"(smem_u32(&bar)) : "memory");
    }
    unsigned done = 0;
    while (!done) asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(smem_u32(&bar)) : "memory");
  } else {
    const float4* s = reinterpret_cast<const float4*>(buf);
    for (int i0 = tid; i0 < nf4; i0 += 256 * 8) {
      float4 v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int i = min(i0 + u * 256, nf4 - 1);
        if (mode == 2) v[u] = __ldcg(s + i); else v[u] = s[i];
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) if (i0 + u * 256 < nf4) reinterpret_cast<float4*>(sm)[i0 + u * 256] = v[u];
    }
  }
  __syncthreads();
  const long long t1 = clock64();
  if (tid == 0) out[blockIdx.x] = t1 - t0;
  if (acc == 123.456f) sink[0] = acc + sm[tid];
}

// device-wide barrier round trip: all CTAs arrive (fence + atomic), thread 0 spins
__global__ void __launch_bounds__(256) gbar(unsigned* ctr, int rounds, int variant, long long* out) {
  const int G = gridDim.x;
  __syncthreads();
  const long long t0 = clock64();
  for (int r = 1; r <= rounds; ++r) {
    __syncthreads();
    if (threadIdx.x == 0) {
      if (variant == 0) { __threadfence(); atomicAdd(ctr, 1u); }
      else asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(ctr) : "memory");
      while (ld_acq(ctr) < (unsigned)(r * G)) {}
      if (variant == 0) __threadfence();
    }
    __syncthreads();
  }
  const long long t1 = clock64();
  if (threadIdx.x == 0) out[blockIdx.x] = (t1 - t0) / rounds;
}

static void report(const char* name, long long* d_out, int G) {
  std::vector<long long> h(G);
  cudaMemcpy(h.data(), d_out, G * sizeof(long long), cudaMemcpyDeviceToHost);
  std::sort(h.begin(), h.end());
  printf("%-46s median %6lld  max %6lld cycles\n", name, h[G / 2], h[G - 1]);
}

int main() {
  const int G = 98;
  float *buf, *sink; long long* out; unsigned* ctr;
  const size_t n = 64u << 20;
  cudaMalloc(&buf, n * 4); cudaMalloc(&sink, 4096); cudaMalloc(&out, 256 * 8); cudaMalloc(&ctr, 4);
  cudaFuncSetAttribute(fetch, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  const char* names[] = {"cp.async16 shared region", "cp.async16 per-CTA region", "ld.cg.v4 x8 batch -> st.shared", "cp.async4", "ld.cg.f32 x32 per lane (slab pattern)", "cp.async.bulk by one thread"};
  for (int kb : {4, 24, 40}) {
    for (int mode = 0; mode < 6; ++mode) {
      for (int rep = 0; rep < 3; ++rep) {
        fill<<<296, 256>>>(buf, 8u << 20);       // freshly written by another kernel -> resident in L2
        fetch<<<G, 256, 100 * 1024>>>(buf, kb * 1024, mode, out, sink);
      }
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
      char nm[128]; snprintf(nm, sizeof nm, "%2d KB  %s", kb, names[mode]);
      report(nm, out, G);
    }
  }
  cudaFuncSetAttribute(wr_fetch, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  const char* wnames[] = {"cp.async16", "cp.async.bulk", "ld.cg.v4 x8", "ld.v4 x8"};
  for (int mode = 0; mode < 4; ++mode) {
    for (int rep = 0; rep < 3; ++rep) {
      cudaMemset(ctr, 0, 4);
      wr_fetch<<<G, 256, 100 * 1024>>>(buf, 40 * 1024, mode, ctr, out, sink);
    }
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
    char nm[128]; snprintf(nm, sizeof nm, "40 KB written in-kernel, then %s", wnames[mode]);
    report(nm, out, G);
  }
  for (int variant = 0; variant < 2; ++variant)
    for (int g : {13, 98, 148}) {
      cudaMemset(ctr, 0, 4);
      gbar<<<g, 256>>>(ctr, 50, variant, out);
      cudaDeviceSynchronize();
      char nm[128]; snprintf(nm, sizeof nm, "grid barrier, %3d CTAs, %s", g, variant ? "red.release + poll" : "fence + atomicAdd + poll + fence");
      report(nm, out, g);
    }
  return 0;
}
