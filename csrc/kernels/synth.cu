// b200ps K9 — on-device synthetic input pipelines (Philox4x32-10 counter RNG), replacing the
// MNIST download (example.py:60-62,165), the HDFS libsvm readers (lr2.py:87-155) and the
// queue-runner demos (input_pipeline*.py).  Every generator is a pure function of
// (seed, example index), so any worker can produce any shard: example i of worker k of n is
// global example (offset + i), with offset chosen by the caller — the same
// files[task_index::num_workers] split rule as lr2.py:302-304, applied to a counter.
#include "../common.h"
#include "ptx.cuh"

namespace b200ps {

struct Philox {
  uint32_t c[4];
};
__device__ __forceinline__ Philox philox4x32(uint64_t ctr_lo, uint64_t ctr_hi, uint64_t key) {
  uint32_t c0 = (uint32_t)ctr_lo, c1 = (uint32_t)(ctr_lo >> 32), c2 = (uint32_t)ctr_hi,
           c3 = (uint32_t)(ctr_hi >> 32);
  uint32_t k0 = (uint32_t)key, k1 = (uint32_t)(key >> 32);
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    const uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  return Philox{{c0, c1, c2, c3}};
}
__device__ __forceinline__ float u01(uint32_t x) { return (x >> 8) * (1.0f / 16777216.0f); }
__device__ __forceinline__ uint32_t mix32(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
  return x;
}

// ---- MNIST-like: 784 pixels in [0,1]; class = example-dependent; each class lights a fixed
// pseudo-random third of the pixels, plus uniform noise, so the task is learnable.
__global__ void __launch_bounds__(256) synth_mnist_kernel(float* __restrict__ x, float* __restrict__ onehot,
                                                          int* __restrict__ labels, int B, int D, int C,
                                                          uint64_t seed, uint64_t offset) {
  const int quads = (D + 3) / 4;
  const size_t total = (size_t)B * quads;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (size_t)gridDim.x * blockDim.x) {
    const int b = (int)(i / quads), q = (int)(i - (size_t)b * quads);
    const uint64_t ex = offset + b;
    const int label = (int)(philox4x32(ex, 0x6c61626cull, seed).c[0] % (uint32_t)C);
    const Philox r = philox4x32(ex, 1 + q, seed);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int d = q * 4 + e;
      if (d < D) {
        const bool lit = (mix32((uint32_t)d * 131u + (uint32_t)label * 7919u + (uint32_t)seed) % 3u) == 0u;
        const float n = u01(r.c[e]);
        x[(size_t)b * D + d] = lit ? 0.55f + 0.45f * n : 0.25f * n * n;
      }
    }
    if (q == 0) {
      if (labels) labels[b] = label;
      if (onehot)
        for (int c = 0; c < C; ++c) onehot[(size_t)b * C + c] = (c == label) ? 1.f : 0.f;
    }
  }
}

// ---- dense bf16 features for the wide stress model: x = pattern(label) + noise, zero-mean.
__global__ void __launch_bounds__(256) synth_dense_bf16_kernel(__nv_bfloat16* __restrict__ x,
                                                               int* __restrict__ labels, int B, int D,
                                                               int C, uint64_t seed, uint64_t offset) {
  const int quads = D / 4;
  const size_t total = (size_t)B * quads;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (size_t)gridDim.x * blockDim.x) {
    const int b = (int)(i / quads), q = (int)(i - (size_t)b * quads);
    const uint64_t ex = offset + b;
    const int label = (int)(philox4x32(ex, 0x6c61626cull, seed).c[0] % (uint32_t)C);
    const Philox r = philox4x32(ex, 1 + q, seed);
    __nv_bfloat16 o[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int d = q * 4 + e;
      const float sgn = (mix32((uint32_t)d * 2654435761u ^ ((uint32_t)label * 40503u)) & 1u) ? 0.5f : -0.5f;
      o[e] = __float2bfloat16(sgn + (u01(r.c[e]) - 0.5f));
    }
    *reinterpret_cast<uint2*>(x + (size_t)b * D + q * 4) = *reinterpret_cast<uint2*>(o);
    if (q == 0 && labels) labels[b] = label;
  }
}

// ---- sparse hashed-feature batches in CSR (lr2.py:69-84 builds the same thing as python lists).
// nnz_i in [nnz_mean - spread, nnz_mean + spread]; ids uniform (or power-law skewed) in [0,F);
// label ~ Bernoulli(sigmoid(teacher . x)) with a fixed hash-defined teacher vector.
__device__ __forceinline__ float teacher_w(int64_t fid, uint64_t seed) {
  const uint32_t h = mix32((uint32_t)fid ^ mix32((uint32_t)(fid >> 32) + (uint32_t)seed * 97u));
  return (u01(h) - 0.5f) * 2.f;
}

__global__ void __launch_bounds__(1024) synth_sparse_rowptr_kernel(int* __restrict__ row_ptr, int B,
                                                                   int nnz_mean, int spread,
                                                                   uint64_t seed, uint64_t offset) {
  // single CTA exclusive scan of the per-row lengths
  __shared__ int s_part[1024];
  const int tid = threadIdx.x;
  const int per = (B + 1023) / 1024;
  int local = 0;
  for (int k = 0; k < per; ++k) {
    const int b = tid * per + k;
    if (b < B) {
      const uint32_t r = philox4x32(offset + b, 0x6e6e7aull, seed).c[1];
      local += nnz_mean - spread + (int)(r % (uint32_t)(2 * spread + 1));
    }
  }
  s_part[tid] = local;
  __syncthreads();
  for (int o = 1; o < 1024; o <<= 1) {
    const int v = tid >= o ? s_part[tid - o] : 0;
    __syncthreads();
    s_part[tid] += v;
    __syncthreads();
  }
  int run = s_part[tid] - local;
  for (int k = 0; k < per; ++k) {
    const int b = tid * per + k;
    if (b < B) {
      row_ptr[b] = run;
      const uint32_t r = philox4x32(offset + b, 0x6e6e7aull, seed).c[1];
      run += nnz_mean - spread + (int)(r % (uint32_t)(2 * spread + 1));
    }
  }
  if (tid == 1023) row_ptr[B] = s_part[1023];
}

__global__ void __launch_bounds__(128) synth_sparse_fill_kernel(const int* __restrict__ row_ptr,
                                                                int64_t* __restrict__ fid,
                                                                float* __restrict__ fval,
                                                                float* __restrict__ labels, int B,
                                                                int64_t F, float skew, uint64_t seed,
                                                                uint64_t offset) {
  // one warp per example
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= B) return;
  const uint64_t ex = offset + warp;
  const int beg = row_ptr[warp], end = row_ptr[warp + 1];
  float z = 0.f;
  for (int j = beg + lane; j < end; j += 32) {
    const Philox r = philox4x32(ex, 16 + (j - beg), seed);
    double u = ((double)r.c[0] * 4294967296.0 + (double)r.c[1]) * (1.0 / 18446744073709551616.0);
    if (skew != 1.f) u = pow(u, (double)skew);
    int64_t id = (int64_t)(u * (double)F);
    if (id >= F) id = F - 1;
    const float v = 0.5f + u01(r.c[2]);
    fid[j] = id;
    fval[j] = v;
    z += teacher_w(id, seed) * v;
  }
  z = warp_sum(z);
  if (lane == 0) {
    const float n = (float)(end - beg);
    const float logit = 3.f * z * rsqrtf(fmaxf(n, 1.f));
    const float pr = 1.f / (1.f + __expf(-logit));
    const float u = u01(philox4x32(ex, 0x6c62ull, seed).c[3]);
    labels[warp] = u < pr ? 1.f : 0.f;
  }
}

// N(0,1) initialiser (tf.random_normal, example.py:84-85; lr2.py:384): Box-Muller on Philox,
// deterministic per (seed, element index).
__global__ void __launch_bounds__(256) init_normal_kernel(float* __restrict__ out, size_t n, float mean,
                                                          float stddev, uint64_t seed, uint64_t stream_id,
                                                          uint64_t quad_offset) {
  const size_t quads = (n + 3) / 4;
  for (size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x; q < quads;
       q += (size_t)gridDim.x * blockDim.x) {
    const Philox r = philox4x32(q + quad_offset, stream_id, seed);
    float g[4];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const float u1 = fmaxf(u01(r.c[2 * h]), 5.9604645e-8f), u2 = u01(r.c[2 * h + 1]);
      const float rad = sqrtf(-2.f * __logf(u1));
      float sn, cs;
      __sincosf(6.28318530718f * u2, &sn, &cs);
      g[2 * h] = rad * cs;
      g[2 * h + 1] = rad * sn;
    }
#pragma unroll
    for (int e = 0; e < 4; ++e)
      if (q * 4 + e < n) out[q * 4 + e] = mean + stddev * g[e];
  }
}

__global__ void __launch_bounds__(256) cast_f32_bf16_kernel(const float* __restrict__ in,
                                                           __nv_bfloat16* __restrict__ out, size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (size_t)gridDim.x * blockDim.x)
    out[i] = __float2bfloat16(in[i]);
}

}  // namespace b200ps

using namespace b200ps;

static int grid_for(size_t work, int threads) {
  return (int)std::min<size_t>((work + threads - 1) / threads, (size_t)sm_count() * 16);
}

extern "C" int b200ps_synth_mnist(float* x, float* onehot, int32_t* labels, int32_t B, int32_t D,
                                  int32_t C, uint64_t seed, uint64_t offset, b200ps_stream_t stream) {
  B200PS_REQUIRE(B > 0 && D > 0 && C > 0, "synth_mnist: bad shape");
  const size_t work = (size_t)B * ((D + 3) / 4);
  synth_mnist_kernel<<<grid_for(work, 256), 256, 0, as_stream(stream)>>>(x, onehot, labels, B, D, C, seed, offset);
  B200PS_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200ps_synth_dense_bf16(void* x, int32_t* labels, int32_t B, int32_t D, int32_t C,
                                       uint64_t seed, uint64_t offset, b200ps_stream_t stream) {
  B200PS_REQUIRE(B > 0 && D > 0 && D % 4 == 0 && C > 0, "synth_dense: D must be a multiple of 4");
  const size_t work = (size_t)B * (D / 4);
  synth_dense_bf16_kernel<<<grid_for(work, 256), 256, 0, as_stream(stream)>>>((__nv_bfloat16*)x, labels, B, D, C, seed, offset);
  B200PS_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200ps_synth_sparse(int32_t* row_ptr, int64_t* fid, float* fval, float* labels, int32_t B,
                                   int32_t nnz_mean, int32_t spread, int64_t F, float skew, uint64_t seed,
                                   uint64_t offset, b200ps_stream_t stream) {
  B200PS_REQUIRE(B > 0 && nnz_mean > 0 && spread >= 0 && spread < nnz_mean && F > 0, "synth_sparse: bad args");
  cudaStream_t s = as_stream(stream);
  synth_sparse_rowptr_kernel<<<1, 1024, 0, s>>>(row_ptr, B, nnz_mean, spread, seed, offset);
  B200PS_LAUNCH_CHECK();
  synth_sparse_fill_kernel<<<(B * 32 + 127) / 128, 128, 0, s>>>(row_ptr, fid, fval, labels, B, F, skew, seed, offset);
  B200PS_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200ps_init_normal(float* out, uint64_t n, float mean, float stddev, uint64_t seed,
                                  uint64_t stream_id, uint64_t elem_offset, b200ps_stream_t stream) {
  if (n == 0) return 0;
  B200PS_REQUIRE(elem_offset % 4 == 0, "init_normal: elem_offset must be a multiple of 4");
  init_normal_kernel<<<grid_for((n + 3) / 4, 256), 256, 0, as_stream(stream)>>>(out, n, mean, stddev, seed, stream_id, elem_offset / 4);
  B200PS_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200ps_cast_f32_bf16(const float* in, void* out, uint64_t n, b200ps_stream_t stream) {
  if (n == 0) return 0;
  cast_f32_bf16_kernel<<<grid_for(n, 256), 256, 0, as_stream(stream)>>>(in, (__nv_bfloat16*)out, n);
  B200PS_LAUNCH_CHECK();
  return 0;
}
