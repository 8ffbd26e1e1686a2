// b200ps K6/K8/K11 — the non-GEMM pieces of the MLP step, each fused as far as the data
// dependencies allow:
//   mlp_tail      : bias1 + sigmoid, the narrow head GEMM (H x C, C <= 32), stable log-softmax
//                   cross-entropy, accuracy count, dZ3, dW2, db2, dA2 * sigmoid', db1
//                   (example.py:96-103,127-128 and the autodiff of those lines)
//   softmax_ce    : wide-head variant (one CTA per row) for the 8192-class stress model
//   colsum        : bias gradients of the wide layers
//   sgd / adam    : owner-side optimiser apply (example.py:108, model_export.py:38)
#include "../common.h"
#include "ptx.cuh"

namespace b200ps {

__device__ __forceinline__ float sigmoidf_acc(float x) { return 1.f / (1.f + __expf(-x)); }

// ------------------------------------------------------------------------------------
// mlp_tail: rows are processed in tiles of kTailRows; one CTA owns a contiguous row range.
// With a single CTA (B <= kTailRows * iterations) every reduction has a fixed order, so
// the sync-replica path is bitwise reproducible; with several CTAs partial sums are merged
// with fp32 atomics (the caller zeroes the outputs first).
// ------------------------------------------------------------------------------------
constexpr int kTailThreads = 512;
constexpr int kTailRows = 128;
constexpr int kMaxClasses = 32;

struct TailParams {
  const float* __restrict__ z2;     // [B,H] pre-activation WITHOUT bias (split-K GEMM output)
  const float* __restrict__ b1;     // [H]
  const float* __restrict__ w2;     // [H,C]
  const float* __restrict__ b2;     // [C]
  const int* __restrict__ labels;   // [B] class ids
  float* a2_out;       // [B,H] or null
  float* dz2;          // [B,H] (train)
  float* dw2;          // [H,C] (train)
  float* db2;          // [C]   (train)
  float* db1;          // [H]   (train)
  float* loss_sum;     // += sum_b CE_b * loss_scale
  int* correct;        // += #(argmax == label)
  float* probs_out;    // [B,C] or null
  int B, H, C;
  int train;
  int rows_per_cta;
  float grad_scale;    // 1/B_global for the mean loss
  float loss_scale;    // 1/B for a mean, 1 for a sum
  int atomic_out;      // gridDim.x > 1
};

__global__ void __launch_bounds__(kTailThreads, 1) mlp_tail_kernel(const TailParams p) {
  extern __shared__ float sm[];
  const int H = p.H, C = p.C;
  const int HP = H + 1;  // pad: conflict-free column walks
  float* s_w2 = sm;                       // [H*C]
  float* s_b2 = s_w2 + H * C;             // [C]
  float* s_b1 = s_b2 + kMaxClasses;       // [H]
  float* s_a2 = s_b1 + H;                 // [kTailRows * HP]   activations, later dz2
  float* s_dz3 = s_a2 + kTailRows * HP;   // [kTailRows * kMaxClasses] logits -> probs -> dz3
  float* s_red = s_dz3 + kTailRows * kMaxClasses;  // [32] block reduction scratch
  __shared__ int s_correct;

  const int tid = threadIdx.x;
  for (int i = tid; i < H * C; i += kTailThreads) s_w2[i] = p.w2[i];
  for (int i = tid; i < C; i += kTailThreads) s_b2[i] = p.b2[i];
  for (int i = tid; i < H; i += kTailThreads) s_b1[i] = p.b1[i];
  if (tid == 0) s_correct = 0;

  const int row_begin = blockIdx.x * p.rows_per_cta;
  const int row_end = min(p.B, row_begin + p.rows_per_cta);

  // per-thread accumulators for the batch reductions (fixed ownership => fixed order)
  //   dW2: element e = tid + k*kTailThreads for k < kDw2PerThread
  constexpr int kDw2PerThread = 8;  // supports H*C <= 4096
  float acc_dw2[kDw2PerThread];
#pragma unroll
  for (int k = 0; k < kDw2PerThread; ++k) acc_dw2[k] = 0.f;
  float acc_db1 = 0.f;  // thread h < H owns db1[h]
  float acc_db2 = 0.f;  // thread c < C owns db2[c]
  float loss_local = 0.f;
  __syncthreads();

  for (int r0 = row_begin; r0 < row_end; r0 += kTailRows) {
    const int nrows = min(kTailRows, row_end - r0);
    // (1) a2 = sigmoid(z2 + b1).  The tile is one contiguous run of nrows*H floats: issue all of
    // a thread's loads before the first use so the global latency is paid once, not per element.
    {
      const float* __restrict__ zsrc = p.z2 + (size_t)r0 * H;
      const int n = nrows * H;
      constexpr int kU = 8;
      for (int base = 0; base < n; base += kU * kTailThreads) {
        float v[kU];
#pragma unroll
        for (int u = 0; u < kU; ++u) {
          const int i = base + u * kTailThreads + tid;
          v[u] = i < n ? __ldg(zsrc + i) : 0.f;
        }
#pragma unroll
        for (int u = 0; u < kU; ++u) {
          const int i = base + u * kTailThreads + tid;
          if (i < n) {
            const int r = i / H, h = i - r * H;
            const float a = sigmoidf_acc(v[u] + s_b1[h]);
            s_a2[r * HP + h] = a;
            if (p.a2_out) p.a2_out[(size_t)(r0 + r) * H + h] = a;
          }
        }
      }
    }
    __syncthreads();
    // (2) logits z3 = a2 . W2 + b2 : one thread per (row, class)
    for (int i = tid; i < nrows * C; i += kTailThreads) {
      const int r = i / C, c = i - r * C;
      float z = s_b2[c];
      const float* a = s_a2 + r * HP;
#pragma unroll 4
      for (int h = 0; h < H; ++h) z = fmaf(a[h], s_w2[h * C + c], z);
      s_dz3[r * kMaxClasses + c] = z;
    }
    __syncthreads();
    // (3) stable log-softmax, loss, accuracy, dz3 = (p - y) * grad_scale : one thread per row
    if (tid < nrows) {
      float* z = s_dz3 + tid * kMaxClasses;
      const int label = p.labels[r0 + tid];
      float m = z[0];
      int arg = 0;
      for (int c = 1; c < C; ++c)
        if (z[c] > m) { m = z[c]; arg = c; }
      float se = 0.f;
      for (int c = 0; c < C; ++c) se += __expf(z[c] - m);
      const float lse = m + __logf(se);
      loss_local += (lse - z[label]);
      if (arg == label) atomicAdd(&s_correct, 1);
      const float inv = 1.f / se;
      for (int c = 0; c < C; ++c) {
        const float pr = __expf(z[c] - m) * inv;
        if (p.probs_out) p.probs_out[(size_t)(r0 + tid) * C + c] = pr;
        z[c] = (pr - (c == label ? 1.f : 0.f)) * p.grad_scale;
      }
    }
    __syncthreads();
    if (p.train) {
      // (4) dW2 += a2^T dz3 ; db2 += sum_r dz3
#pragma unroll
      for (int k = 0; k < kDw2PerThread; ++k) {
        const int e = tid + k * kTailThreads;
        if (e < H * C) {
          const int h = e / C, c = e - h * C;
          float s = 0.f;
          for (int r = 0; r < nrows; ++r) s = fmaf(s_a2[r * HP + h], s_dz3[r * kMaxClasses + c], s);
          acc_dw2[k] += s;
        }
      }
      if (tid < C) {
        float s = 0.f;
        for (int r = 0; r < nrows; ++r) s += s_dz3[r * kMaxClasses + tid];
        acc_db2 += s;
      }
      __syncthreads();
      // (5) dz2 = (dz3 . W2^T) * a2 (1 - a2)   (overwrites s_a2 in place)
      for (int i = tid; i < nrows * H; i += kTailThreads) {
        const int r = i / H, h = i - r * H;
        const float* dz = s_dz3 + r * kMaxClasses;
        const float* w = s_w2 + h * C;
        float d = 0.f;
        for (int c = 0; c < C; ++c) d = fmaf(dz[c], w[c], d);
        const float a = s_a2[r * HP + h];
        d *= a * (1.f - a);
        s_a2[r * HP + h] = d;
        p.dz2[(size_t)(r0 + r) * H + h] = d;
      }
      __syncthreads();
      // (6) db1 += sum_r dz2
      if (tid < H) {
        float s = 0.f;
        for (int r = 0; r < nrows; ++r) s += s_a2[r * HP + tid];
        acc_db1 += s;
      }
    }
    __syncthreads();
  }

  // block-level loss reduction in a fixed order
  float l = warp_sum(loss_local);
  if ((tid & 31) == 0) s_red[tid >> 5] = l;
  __syncthreads();
  if (tid == 0) {
    float tot = 0.f;
    for (int w = 0; w < kTailThreads / 32; ++w) tot += s_red[w];
    tot *= p.loss_scale;
    if (p.loss_sum) { if (p.atomic_out) atomicAdd(p.loss_sum, tot); else *p.loss_sum += tot; }
    if (p.correct) atomicAdd(p.correct, s_correct);
  }
  if (p.train) {
#pragma unroll
    for (int k = 0; k < kDw2PerThread; ++k) {
      const int e = tid + k * kTailThreads;
      if (e < H * C) { if (p.atomic_out) atomicAdd(p.dw2 + e, acc_dw2[k]); else p.dw2[e] = acc_dw2[k]; }
    }
    if (tid < C) { if (p.atomic_out) atomicAdd(p.db2 + tid, acc_db2); else p.db2[tid] = acc_db2; }
    if (tid < H) { if (p.atomic_out) atomicAdd(p.db1 + tid, acc_db1); else p.db1[tid] = acc_db1; }
  }
}

// ------------------------------------------------------------------------------------
// softmax_ce: logits [B,C] (bf16 or fp32) -> dlogits (same dtype, may alias), loss, accuracy.
// One CTA per row; the row is held in registers between the two passes.
// ------------------------------------------------------------------------------------
constexpr int kCeThreads = 256;
constexpr int kCeMaxPerThread = 64;  // C <= 16384

template <typename T>
__device__ __forceinline__ float to_f(T v);
template <> __device__ __forceinline__ float to_f<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <typename T>
__device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f<__nv_bfloat16>(float v) { return __float2bfloat16(v); }

template <typename T>
__global__ void __launch_bounds__(kCeThreads) softmax_ce_kernel(const T* __restrict__ logits,
                                                               const int* __restrict__ labels,
                                                               T* dlogits, float* loss_sum,
                                                               int* correct, int B, int C,
                                                               long long ld, float grad_scale,
                                                               float loss_scale) {
  __shared__ float s_val[kCeThreads / 32];
  __shared__ int s_idx[kCeThreads / 32];
  __shared__ float s_bcast[2];
  const int row = blockIdx.x;
  const int tid = threadIdx.x;
  const T* z = logits + (size_t)row * ld;
  float v[kCeMaxPerThread];
  float m = -INFINITY;
  int arg = 0;
  const int per = (C + kCeThreads - 1) / kCeThreads;
#pragma unroll
  for (int k = 0; k < kCeMaxPerThread; ++k) {
    if (k < per) {
      const int c = tid + k * kCeThreads;
      v[k] = c < C ? to_f<T>(z[c]) : -INFINITY;
      if (v[k] > m) { m = v[k]; arg = c; }
    }
  }
  // block argmax (lowest index wins ties, like tf.argmax)
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float om = __shfl_xor_sync(0xffffffffu, m, o);
    const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
    if (om > m || (om == m && oa < arg)) { m = om; arg = oa; }
  }
  if ((tid & 31) == 0) { s_val[tid >> 5] = m; s_idx[tid >> 5] = arg; }
  __syncthreads();
  if (tid == 0) {
    float bm = s_val[0]; int ba = s_idx[0];
    for (int w = 1; w < kCeThreads / 32; ++w)
      if (s_val[w] > bm || (s_val[w] == bm && s_idx[w] < ba)) { bm = s_val[w]; ba = s_idx[w]; }
    s_bcast[0] = bm;
    s_idx[0] = ba;
  }
  __syncthreads();
  m = s_bcast[0];
  const int amax = s_idx[0];
  __syncthreads();
  float se = 0.f;
#pragma unroll
  for (int k = 0; k < kCeMaxPerThread; ++k)
    if (k < per) { v[k] = __expf(v[k] - m); se += v[k]; }
  se = warp_sum(se);
  if ((tid & 31) == 0) s_val[tid >> 5] = se;
  __syncthreads();
  if (tid == 0) {
    float t = 0.f;
    for (int w = 0; w < kCeThreads / 32; ++w) t += s_val[w];
    s_bcast[1] = t;
  }
  __syncthreads();
  se = s_bcast[1];
  const int label = labels[row];
  const float inv = 1.f / se;
  if (dlogits) {
    T* d = dlogits + (size_t)row * ld;
#pragma unroll
    for (int k = 0; k < kCeMaxPerThread; ++k) {
      if (k < per) {
        const int c = tid + k * kCeThreads;
        if (c < C) d[c] = from_f<T>((v[k] * inv - (c == label ? 1.f : 0.f)) * grad_scale);
      }
    }
  }
  if (tid == 0) {
    const float zl = to_f<T>(z[label]);
    if (loss_sum) atomicAdd(loss_sum, (m + __logf(se) - zl) * loss_scale);
    if (correct && amax == label) atomicAdd(correct, 1);
  }
}

// ------------------------------------------------------------------------------------
// colsum: out[n] (+)= scale * sum_m x[m,n]   — bias gradients.
// ------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) colsum_kernel(const T* __restrict__ x, float* out, int M, int N,
                                                     long long ld, int rows_per_cta, float scale) {
  __shared__ float s[8][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int col = blockIdx.x * 32 + tx;
  const int r0 = blockIdx.y * rows_per_cta;
  const int r1 = min(M, r0 + rows_per_cta);
  float acc = 0.f;
  if (col < N)
    for (int r = r0 + ty; r < r1; r += 8) acc += to_f<T>(x[(size_t)r * ld + col]);
  s[ty][tx] = acc;
  __syncthreads();
  if (ty == 0 && col < N) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += s[k][tx];
    atomicAdd(out + col, t * scale);
  }
}

// ------------------------------------------------------------------------------------
// optimisers
// ------------------------------------------------------------------------------------
template <typename G>
__global__ void __launch_bounds__(256) sgd_apply_kernel(float* __restrict__ p, const G* __restrict__ g,
                                                        __nv_bfloat16* lowp, size_t n, float lr,
                                                        float gscale) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float w = p[i] - lr * gscale * to_f<G>(g[i]);
    p[i] = w;
    if (lowp) lowp[i] = __float2bfloat16(w);
  }
}

// Adam exactly as TF's ApplyAdam: lr_t = lr * sqrt(1-b2^t)/(1-b1^t);
// m = b1 m + (1-b1) g ; v = b2 v + (1-b2) g^2 ; p -= lr_t * m / (sqrt(v) + eps).
// beta powers live on the device (beta_pow[0] = b1^t, beta_pow[1] = b2^t) and are
// advanced by the launch's last block, which mirrors TF's beta1_power/beta2_power slots.
__global__ void __launch_bounds__(256) adam_apply_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                         float* __restrict__ m, float* __restrict__ v,
                                                         size_t n, float lr, float b1, float b2,
                                                         float eps, const float* beta_pow) {
  const float b1p = beta_pow[0], b2p = beta_pow[1];
  const float lr_t = lr * sqrtf(1.f - b2p) / (1.f - b1p);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float gi = g[i];
    const float mi = b1 * m[i] + (1.f - b1) * gi;
    const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    p[i] -= lr_t * mi / (sqrtf(vi) + eps);
  }
}
__global__ void adam_advance_kernel(float* beta_pow, float b1, float b2) {
  beta_pow[0] *= b1;
  beta_pow[1] *= b2;
}

// tiny linear regression (model_export.py:27-38): y_pred = x*w + b ; loss = sum((y-yp)^2)/n
// grads for w (sum over features of x^T r) and b.  x:[B,D], w:[D,1]
__global__ void __launch_bounds__(256) linreg_fwd_bwd_kernel(const float* __restrict__ x,
                                                            const float* __restrict__ y,
                                                            const float* __restrict__ w,
                                                            const float* __restrict__ b, float* ypred,
                                                            float* loss, float* gw, float* gb, int B,
                                                            int D, float inv_n) {
  // single CTA: B and D are tiny in the export fixture
  __shared__ float s_red[256];
  const int tid = threadIdx.x;
  float l = 0.f, gbl = 0.f;
  for (int r = tid; r < B; r += blockDim.x) {
    float yp = b[0];
    for (int d = 0; d < D; ++d) yp = fmaf(x[(size_t)r * D + d], w[d], yp);
    if (ypred) ypred[r] = yp;
    const float e = y[r] - yp;
    l += e * e * inv_n;
    gbl += -2.f * e * inv_n;
  }
  s_red[tid] = l;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) { if (tid < o) s_red[tid] += s_red[tid + o]; __syncthreads(); }
  if (tid == 0 && loss) *loss = s_red[0];
  __syncthreads();
  s_red[tid] = gbl;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) { if (tid < o) s_red[tid] += s_red[tid + o]; __syncthreads(); }
  if (tid == 0 && gb) *gb = s_red[0];
  __syncthreads();
  if (gw) {
    for (int d = 0; d < D; ++d) {
      float g = 0.f;
      for (int r = tid; r < B; r += blockDim.x) {
        float yp = b[0];
        for (int k = 0; k < D; ++k) yp = fmaf(x[(size_t)r * D + k], w[k], yp);
        g += -2.f * (y[r] - yp) * inv_n * x[(size_t)r * D + d];
      }
      s_red[tid] = g;
      __syncthreads();
      for (int o = 128; o > 0; o >>= 1) { if (tid < o) s_red[tid] += s_red[tid + o]; __syncthreads(); }
      if (tid == 0) gw[d] = s_red[0];
      __syncthreads();
    }
  }
}

}  // namespace b200ps

using namespace b200ps;

extern "C" int b200ps_mlp_tail(const float* z2, const float* b1, const float* w2, const float* b2,
                               const int32_t* labels, float* a2_out, float* dz2, float* dw2,
                               float* db2, float* db1, float* loss_sum, int32_t* correct,
                               float* probs_out, int32_t B, int32_t H, int32_t C, int32_t train,
                               float grad_scale, float loss_scale, b200ps_stream_t stream) {
  B200PS_REQUIRE(B > 0 && H > 0 && C > 0, "mlp_tail: bad shape");
  B200PS_REQUIRE(C <= kMaxClasses, "mlp_tail: C=%d > %d (use gemm + softmax_ce)", C, kMaxClasses);
  B200PS_REQUIRE(H <= 256 && H * C <= 4096, "mlp_tail: H=%d too wide for the fused head", H);
  B200PS_REQUIRE(!train || (dz2 && dw2 && db2 && db1), "mlp_tail: training needs gradient outputs");
  TailParams p;
  p.z2 = z2; p.b1 = b1; p.w2 = w2; p.b2 = b2; p.labels = labels; p.a2_out = a2_out; p.dz2 = dz2;
  p.dw2 = dw2; p.db2 = db2; p.db1 = db1; p.loss_sum = loss_sum; p.correct = correct;
  p.probs_out = probs_out; p.B = B; p.H = H; p.C = C; p.train = train;
  p.grad_scale = grad_scale; p.loss_scale = loss_scale;
  // training on small batches: one CTA => fixed reduction order (bitwise reproducible)
  int ctas = 1;
  if (B > 2 * kTailRows) ctas = std::min((B + kTailRows - 1) / kTailRows, 4 * sm_count());
  p.rows_per_cta = ((B + ctas - 1) / ctas + kTailRows - 1) / kTailRows * kTailRows;
  ctas = (B + p.rows_per_cta - 1) / p.rows_per_cta;
  p.atomic_out = ctas > 1 ? 1 : 0;
  cudaStream_t s = as_stream(stream);
  if (p.atomic_out && train) {
    B200PS_CUDA_CHECK(cudaMemsetAsync(dw2, 0, sizeof(float) * H * C, s));
    B200PS_CUDA_CHECK(cudaMemsetAsync(db2, 0, sizeof(float) * C, s));
    B200PS_CUDA_CHECK(cudaMemsetAsync(db1, 0, sizeof(float) * H, s));
  }
  const size_t smem = sizeof(float) * ((size_t)H * C + kMaxClasses + H + (size_t)kTailRows * (H + 1) +
                                       (size_t)kTailRows * kMaxClasses + 32);
  static size_t smem_set = 0;
  if (smem > smem_set) {
    B200PS_CUDA_CHECK(cudaFuncSetAttribute(mlp_tail_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)smem));
    smem_set = smem;
  }
  mlp_tail_kernel<<<ctas, kTailThreads, smem, s>>>(p);
  B200PS_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200ps_softmax_ce(const void* logits, const int32_t* labels, void* dlogits,
                                 float* loss_sum, int32_t* correct, int32_t B, int32_t C, int64_t ld,
                                 int32_t dtype, float grad_scale, float loss_scale,
                                 b200ps_stream_t stream) {
  B200PS_REQUIRE(B > 0 && C > 0 && C <= kCeThreads * kCeMaxPerThread, "softmax_ce: bad shape");
  cudaStream_t s = as_stream(stream);
  if (dtype == B200PS_DT_F32)
    softmax_ce_kernel<float><<<B, kCeThreads, 0, s>>>((const float*)logits, labels, (float*)dlogits,
                                                     loss_sum, correct, B, C, ld, grad_scale, loss_scale);
  else
    softmax_ce_kernel<__nv_bfloat16><<<B, kCeThreads, 0, s>>>((const __nv_bfloat16*)logits, labels,
                                                             (__nv_bfloat16*)dlogits, loss_sum, correct,
                                                             B, C, ld, grad_scale, loss_scale);
  B200PS_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200ps_colsum(const void* x, float* out, int32_t M, int32_t N, int64_t ld, int32_t dtype,
                             float scale, b200ps_stream_t stream) {
  B200PS_REQUIRE(M > 0 && N > 0, "colsum: bad shape");
  cudaStream_t s = as_stream(stream);
  B200PS_CUDA_CHECK(cudaMemsetAsync(out, 0, sizeof(float) * N, s));
  const int rows_per_cta = 256;
  dim3 grid((N + 31) / 32, (M + rows_per_cta - 1) / rows_per_cta);
  if (dtype == B200PS_DT_F32)
    colsum_kernel<float><<<grid, 256, 0, s>>>((const float*)x, out, M, N, ld, rows_per_cta, scale);
  else
    colsum_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>((const __nv_bfloat16*)x, out, M, N, ld, rows_per_cta, scale);
  B200PS_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200ps_sgd_apply(float* params, const void* grads, int32_t grad_dtype, void* lowp,
                                uint64_t n, float lr, float grad_scale, b200ps_stream_t stream) {
  if (n == 0) return 0;
  cudaStream_t s = as_stream(stream);
  const int blocks = (int)std::min<uint64_t>((n + 255) / 256, (uint64_t)sm_count() * 8);
  if (grad_dtype == B200PS_DT_F32)
    sgd_apply_kernel<float><<<blocks, 256, 0, s>>>(params, (const float*)grads, (__nv_bfloat16*)lowp, n, lr, grad_scale);
  else
    sgd_apply_kernel<__nv_bfloat16><<<blocks, 256, 0, s>>>(params, (const __nv_bfloat16*)grads, (__nv_bfloat16*)lowp, n, lr, grad_scale);
  B200PS_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200ps_adam_apply(float* params, const float* grads, float* m, float* v, uint64_t n,
                                 float lr, float beta1, float beta2, float eps, const float* beta_pow,
                                 b200ps_stream_t stream) {
  if (n == 0) return 0;
  cudaStream_t s = as_stream(stream);
  const int blocks = (int)std::min<uint64_t>((n + 255) / 256, (uint64_t)sm_count() * 8);
  adam_apply_kernel<<<blocks, 256, 0, s>>>(params, grads, m, v, n, lr, beta1, beta2, eps, beta_pow);
  B200PS_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200ps_adam_advance(float* beta_pow, float beta1, float beta2, b200ps_stream_t stream) {
  adam_advance_kernel<<<1, 1, 0, as_stream(stream)>>>(beta_pow, beta1, beta2);
  B200PS_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200ps_linreg_fwd_bwd(const float* x, const float* y, const float* w, const float* b,
                                     float* ypred, float* loss, float* gw, float* gb, int32_t B,
                                     int32_t D, float inv_n, b200ps_stream_t stream) {
  B200PS_REQUIRE(B > 0 && D > 0, "linreg: bad shape");
  linreg_fwd_bwd_kernel<<<1, 256, 0, as_stream(stream)>>>(x, y, w, b, ypred, loss, gw, gb, B, D, inv_n);
  B200PS_LAUNCH_CHECK();
  return 0;
}
