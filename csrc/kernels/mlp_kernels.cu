// b200ps K6/K8/K11 — the non-GEMM pieces of the MLP step, each fused as far as the data
// dependencies allow:
//   mlp_tail      : bias1 + sigmoid, the narrow head GEMM (H x C, C <= 32), stable log-softmax
//                   cross-entropy, accuracy count, dZ3, dW2, db2, dA2 * sigmoid', db1
//                   (example.py:96-103,127-128 and the autodiff of those lines)
//   softmax_ce    : wide-head variant (one CTA per row) for the 8192-class stress model
//   colsum        : bias gradients of the wide layers
//   sgd / adam    : owner-side optimiser apply (example.py:108, model_export.py:38)
#include "../common.h"
#include "../workspace.h"
#include "ptx.cuh"

namespace b200ps {

__device__ __forceinline__ float sigmoidf_acc(float x) { return 1.f / (1.f + __expf(-x)); }

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
  int flags;           // B200PS_TAIL_*
  const int* label_offset;   // device scalar added to the label row index, or null
  float* z2_clear;     // == z2 when B200PS_TAIL_CLEAR_Z2
  int z2_splits;       // >= 1: number of split-K slabs stacked in z2
  long long z2_split_stride;
};

// ------------------------------------------------------------------------------------
// mlp_tail: one SM cannot issue the ~80k warp instructions of the head in less than ~20 us
// (measured), so rows are dealt to CTAs in tiles of 8
// (one warp per row, lanes over hidden units, class vectors in registers, warp shuffles for
// the row reductions).  Cross-row sums (dW2, db2, db1, loss) are accumulated per CTA in a
// fixed order, written to the library workspace, and merged in CTA order by the last CTA to
// arrive — deterministic, no fp32 atomics.
// ------------------------------------------------------------------------------------
constexpr int kT2Threads = 256;
constexpr int kT2Rows = 8;
constexpr int kT2MaxPer = 16;   // outputs of dW2 per thread: H*C <= 4096

template <int CP>
__global__ void __launch_bounds__(kT2Threads) mlp_tail2_kernel(const TailParams p, float* __restrict__ scratch,
                                                              unsigned int* counter) {
  extern __shared__ __align__(16) float sm[];
  const int H = p.H, C = p.C;
  constexpr int CPP = CP + 4;                // padded row: float4 loads of 8 consecutive rows are conflict-free
  float* s_w2 = sm;                          // [H][CPP], zero padded classes
  float* s_b2 = s_w2 + H * CPP;              // [CP]
  float* s_b1 = s_b2 + CP;                   // [H] (+pad)
  float* s_a2 = s_b1 + ((H + 3) & ~3);       // [8][H]
  float* s_dz2 = s_a2 + kT2Rows * H;         // [8][H]
  float* s_dz3 = s_dz2 + kT2Rows * H;        // [8][CP]
  __shared__ float s_loss[kT2Rows];
  __shared__ int s_corr[kT2Rows];
  __shared__ bool s_last;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < H * CP; i += kT2Threads) {
    const int h = i / CP, c = i - h * CP;
    s_w2[h * CPP + c] = c < C ? __ldg(p.w2 + h * C + c) : 0.f;
  }
  for (int i = tid; i < CP; i += kT2Threads) s_b2[i] = i < C ? __ldg(p.b2 + i) : 0.f;
  for (int i = tid; i < H; i += kT2Threads) s_b1[i] = __ldg(p.b1 + i);

  const int row_begin = blockIdx.x * p.rows_per_cta;
  const int row_end = min(p.B, row_begin + p.rows_per_cta);
  const int HK = (H + 31) >> 5;   // hidden units per lane (<= 4)
  const int P = H * C + C + H + 2;   // dW2 | db2 | db1 | loss | correct

  // gradient outputs owned by this thread (fixed ownership => fixed summation order)
  int eh[kT2MaxPer], ec[kT2MaxPer];
  float acc[kT2MaxPer];
#pragma unroll
  for (int j = 0; j < kT2MaxPer; ++j) {
    const int e = tid + j * kT2Threads;
    eh[j] = e < H * C ? e / C : -1;
    ec[j] = e < H * C ? e - (e / C) * C : 0;
    acc[j] = 0.f;
  }
  float acc_b2 = 0.f, acc_b1a = 0.f, acc_b1b = 0.f;   // thread c < C ; threads h = tid, tid + 256 < H
  float loss_w = 0.f;
  int corr_w = 0;
  // W2/b1/b2 were last written by the previous step's push, which every kernel between it and
  // us has already waited on; only z2 comes from our immediate predecessor.
  pdl_wait();
  pdl_trigger();
  const int label0 = p.label_offset ? __ldg(p.label_offset) : 0;
  __syncthreads();

  for (int r0 = row_begin; r0 < row_end; r0 += kT2Rows) {
    const int r = r0 + warp;
    const bool row_ok = r < row_end;
    // (a) a = sigmoid(z2 + b1) for this lane's hidden units (label fetched in the same batch of loads)
    const int label = row_ok ? __ldg(p.labels + label0 + r) : 0;
    float zk[4], ak[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int h = lane + 32 * k;
      zk[k] = (row_ok && k < HK && h < H) ? __ldcg(p.z2 + (size_t)r * H + h) : 0.f;
    }
    for (int sp = 1; sp < p.z2_splits; ++sp) {   // split-K slabs, summed in slab order
      float part[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int h = lane + 32 * k;
        part[k] = (row_ok && k < HK && h < H) ? __ldcg(p.z2 + sp * p.z2_split_stride + (size_t)r * H + h) : 0.f;
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) zk[k] += part[k];
    }
    if (p.z2_clear) {   // leave z2 zeroed for the next step's split-K accumulation
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int h = lane + 32 * k;
        if (row_ok && k < HK && h < H) p.z2_clear[(size_t)r * H + h] = 0.f;
      }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int h = lane + 32 * k;
      const bool ok = row_ok && k < HK && h < H;
      ak[k] = ok ? sigmoidf_acc(zk[k] + s_b1[h]) : 0.f;
      if (k < HK && h < H) {
        s_a2[warp * H + h] = ak[k];
        if (ok && p.a2_out) p.a2_out[(size_t)r * H + h] = ak[k];
      }
    }
    // (b) logits: per-lane partial over its hidden units, then a recursive-halving reduce that
    // leaves class c's total in lanes c and c+16 (CP=16) / lane c (CP=32): 16+ shuffles, not 80
    float zc[CP];
#pragma unroll
    for (int c = 0; c < CP; ++c) zc[c] = 0.f;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int h = lane + 32 * k;
      if (k < HK && h < H) {
        const float4* w = reinterpret_cast<const float4*>(s_w2 + h * CPP);
#pragma unroll
        for (int c4 = 0; c4 < CP / 4; ++c4) {
          const float4 wv = w[c4];
          zc[4 * c4] = fmaf(ak[k], wv.x, zc[4 * c4]);
          zc[4 * c4 + 1] = fmaf(ak[k], wv.y, zc[4 * c4 + 1]);
          zc[4 * c4 + 2] = fmaf(ak[k], wv.z, zc[4 * c4 + 2]);
          zc[4 * c4 + 3] = fmaf(ak[k], wv.w, zc[4 * c4 + 3]);
        }
      }
    }
    if (CP == 16) {   // fold the two half-warps first so that 16 lanes x 16 classes remain
#pragma unroll
      for (int c = 0; c < CP; ++c) zc[c] += __shfl_xor_sync(0xffffffffu, zc[c], 16);
    }
    // halving over lane bits: after the step for bit b, a lane keeps only the classes whose
    // bit b equals its own lane bit b
#pragma unroll
    for (int bit = (CP == 32 ? 16 : 8), n = CP / 2; bit >= 1; bit >>= 1, n >>= 1) {
      const bool up = (lane & bit) != 0;
#pragma unroll
      for (int j = 0; j < n; ++j) {
        const float send = up ? zc[j] : zc[j + n];
        const float keep = up ? zc[j + n] : zc[j];
        zc[j] = keep + __shfl_xor_sync(0xffffffffu, send, bit);
      }
    }
    const int myc = lane & (CP - 1);            // the class this lane now owns
    float z = zc[0] + s_b2[myc];
    // (c) stable log-softmax across the lanes of the (half-)warp that own real classes
    const bool cls_ok = myc < C;
    float m = cls_ok ? z : -INFINITY;
    int arg = myc;
#pragma unroll
    for (int o = CP / 2; o > 0; o >>= 1) {
      const float om = __shfl_xor_sync(0xffffffffu, m, o);
      const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
      if (om > m || (om == m && oa < arg)) { m = om; arg = oa; }
    }
    const float e = cls_ok ? __expf(z - m) : 0.f;
    float se = e;
#pragma unroll
    for (int o = CP / 2; o > 0; o >>= 1) se += __shfl_xor_sync(0xffffffffu, se, o);
    const float zl = __shfl_sync(0xffffffffu, z, label & (CP - 1));
    const float pr = e / se;
    if (row_ok && lane == 0) {
      loss_w += m + __logf(se) - zl;
      corr_w += (arg == label) ? 1 : 0;
    }
    const float mine = (row_ok && cls_ok) ? (pr - (myc == label ? 1.f : 0.f)) * p.grad_scale : 0.f;
    if (row_ok && p.probs_out && lane < CP && cls_ok) p.probs_out[(size_t)r * C + myc] = pr;
    if (lane < CP) s_dz3[warp * CP + lane] = mine;
    __syncwarp();
    {   // every lane needs the whole dz3 row for (d): broadcast reads back from shared memory
      const float4* dz = reinterpret_cast<const float4*>(s_dz3 + warp * CP);
#pragma unroll
      for (int c4 = 0; c4 < CP / 4; ++c4) {
        const float4 d = dz[c4];
        zc[4 * c4] = d.x; zc[4 * c4 + 1] = d.y; zc[4 * c4 + 2] = d.z; zc[4 * c4 + 3] = d.w;
      }
    }
    if (p.train) {
      // (d) dz2 = (dz3 . W2^T) * a (1 - a)
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int h = lane + 32 * k;
        if (k < HK && h < H) {
          const float4* w = reinterpret_cast<const float4*>(s_w2 + h * CPP);
          float d = 0.f;
#pragma unroll
          for (int c4 = 0; c4 < CP / 4; ++c4) {
            const float4 wv = w[c4];
            d = fmaf(zc[4 * c4], wv.x, d);
            d = fmaf(zc[4 * c4 + 1], wv.y, d);
            d = fmaf(zc[4 * c4 + 2], wv.z, d);
            d = fmaf(zc[4 * c4 + 3], wv.w, d);
          }
          d *= ak[k] * (1.f - ak[k]);
          s_dz2[warp * H + h] = d;
          if (row_ok) p.dz2[(size_t)r * H + h] = d;
        }
      }
      __syncthreads();
      // (e) this tile's contribution to dW2 / db2 / db1 (padded rows hold zeros)
#pragma unroll
      for (int j = 0; j < kT2MaxPer; ++j) {
        if (eh[j] >= 0) {
          float t = 0.f;
#pragma unroll
          for (int w8 = 0; w8 < kT2Rows; ++w8) t = fmaf(s_a2[w8 * H + eh[j]], s_dz3[w8 * CP + ec[j]], t);
          acc[j] += t;
        }
      }
      if (tid < C) {
        float t = 0.f;
#pragma unroll
        for (int w8 = 0; w8 < kT2Rows; ++w8) t += s_dz3[w8 * CP + tid];
        acc_b2 += t;
      }
      if (tid < H) {
        float t = 0.f;
#pragma unroll
        for (int w8 = 0; w8 < kT2Rows; ++w8) t += s_dz2[w8 * H + tid];
        acc_b1a += t;
      }
      if (tid + kT2Threads < H) {
        float t = 0.f;
#pragma unroll
        for (int w8 = 0; w8 < kT2Rows; ++w8) t += s_dz2[w8 * H + tid + kT2Threads];
        acc_b1b += t;
      }
      __syncthreads();
    }
  }

  if (lane == 0) { s_loss[warp] = loss_w; s_corr[warp] = corr_w; }
  __syncthreads();
  float loss_cta = 0.f;
  int corr_cta = 0;
  if (tid == 0)
    for (int w8 = 0; w8 < kT2Rows; ++w8) { loss_cta += s_loss[w8]; corr_cta += s_corr[w8]; }
  const bool overwrite = (p.flags & B200PS_TAIL_OVERWRITE_STATS) != 0;

  if (gridDim.x == 1) {
    if (tid == 0) {
      if (p.loss_sum) *p.loss_sum = (overwrite ? 0.f : *p.loss_sum) + loss_cta * p.loss_scale;
      if (p.correct) *p.correct = (overwrite ? 0 : *p.correct) + corr_cta;
    }
    if (p.train) {
#pragma unroll
      for (int j = 0; j < kT2MaxPer; ++j)
        if (eh[j] >= 0) p.dw2[tid + j * kT2Threads] = acc[j];
      if (tid < C) p.db2[tid] = acc_b2;
      if (tid < H) p.db1[tid] = acc_b1a;
      if (tid + kT2Threads < H) p.db1[tid + kT2Threads] = acc_b1b;
    }
    return;
  }

  // partials -> workspace, last CTA merges them in CTA order
  float* mine = scratch + (size_t)blockIdx.x * P;
  if (p.train) {
#pragma unroll
    for (int j = 0; j < kT2MaxPer; ++j)
      if (eh[j] >= 0) mine[tid + j * kT2Threads] = acc[j];
    if (tid < C) mine[H * C + tid] = acc_b2;
    if (tid < H) mine[H * C + C + tid] = acc_b1a;
    if (tid + kT2Threads < H) mine[H * C + C + tid + kT2Threads] = acc_b1b;
  }
  if (tid == 0) { mine[P - 2] = loss_cta; mine[P - 1] = (float)corr_cta; }   // counts < 2^24: exact
  __syncthreads();
  if (tid == 0) {
    __threadfence();
    s_last = (atomicAdd(counter, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  const int first = p.train ? 0 : P - 2;
  // 4 outputs x 8 partials = 32 independent loads in flight per thread; sums stay in CTA order
  for (int o0 = first + tid; o0 < P; o0 += 4 * kT2Threads) {
    float t[4] = {0.f, 0.f, 0.f, 0.f};
    for (unsigned int b0 = 0; b0 < gridDim.x; b0 += 8) {
      float v[4][8];
#pragma unroll
      for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int o = o0 + q * kT2Threads;
          v[q][u] = (o < P && b0 + u < gridDim.x) ? __ldcg(scratch + (size_t)(b0 + u) * P + o) : 0.f;
        }
#pragma unroll
      for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int u = 0; u < 8; ++u) t[q] += v[q][u];
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int o = o0 + q * kT2Threads;
      if (o >= P) continue;
      if (o < H * C) p.dw2[o] = t[q];
      else if (o < H * C + C) p.db2[o - H * C] = t[q];
      else if (o < P - 2) p.db1[o - H * C - C] = t[q];
      else if (o == P - 2) { if (p.loss_sum) *p.loss_sum = (overwrite ? 0.f : *p.loss_sum) + t[q] * p.loss_scale; }
      else if (p.correct) *p.correct = (overwrite ? 0 : *p.correct) + (int)(t[q] + 0.5f);
    }
  }
  if (tid == 0) *counter = 0;   // self-resetting for the next launch
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

// bf16 rows with C a multiple of 8 and 16-byte aligned rows: every thread owns 8-element (16 B)
// vectors, so a warp reads 512 contiguous bytes per instruction instead of 64.
constexpr int kCeVecMax = 8;   // vectors per thread: C <= 256 * 8 * 8 = 16384
__global__ void __launch_bounds__(kCeThreads) softmax_ce_bf16_vec_kernel(const __nv_bfloat16* __restrict__ logits,
                                                                        const int* __restrict__ labels,
                                                                        __nv_bfloat16* dlogits, float* loss_sum,
                                                                        int* correct, int B, int C, long long ld,
                                                                        float grad_scale, float loss_scale) {
  __shared__ float s_val[kCeThreads / 32];
  __shared__ int s_idx[kCeThreads / 32];
  __shared__ float s_bcast[2];
  const int row = blockIdx.x, tid = threadIdx.x;
  const uint4* z = reinterpret_cast<const uint4*>(logits + (size_t)row * ld);
  const int nvec = C >> 3;
  float v[kCeVecMax][8];
  float m = -INFINITY;
  int arg = 0;
#pragma unroll
  for (int k = 0; k < kCeVecMax; ++k) {
    const int i = tid + k * kCeThreads;
    if (i < nvec) {
      const uint4 u = z[i];
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 f = __bfloat1622float2(h[e]);
        v[k][2 * e] = f.x; v[k][2 * e + 1] = f.y;
      }
#pragma unroll
      for (int e = 0; e < 8; ++e)
        if (v[k][e] > m) { m = v[k][e]; arg = i * 8 + e; }
    }
  }
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
    s_bcast[0] = bm; s_idx[0] = ba;
  }
  __syncthreads();
  m = s_bcast[0];
  const int amax = s_idx[0];
  __syncthreads();
  float se = 0.f;
#pragma unroll
  for (int k = 0; k < kCeVecMax; ++k)
    if (tid + k * kCeThreads < nvec)
#pragma unroll
      for (int e = 0; e < 8; ++e) { v[k][e] = __expf(v[k][e] - m); se += v[k][e]; }
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
    uint4* d = reinterpret_cast<uint4*>(dlogits + (size_t)row * ld);
#pragma unroll
    for (int k = 0; k < kCeVecMax; ++k) {
      const int i = tid + k * kCeThreads;
      if (i < nvec) {
        uint4 u;
        __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int c = i * 8 + 2 * e;
          h[e] = __floats2bfloat162_rn((v[k][2 * e] * inv - (c == label ? 1.f : 0.f)) * grad_scale,
                                       (v[k][2 * e + 1] * inv - (c + 1 == label ? 1.f : 0.f)) * grad_scale);
        }
        d[i] = u;
      }
    }
  }
  if (tid == 0) {
    const float zl = __bfloat162float(logits[(size_t)row * ld + label]);
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
                               float grad_scale, float loss_scale, int32_t flags,
                               const int32_t* label_offset, int32_t z2_splits, int64_t z2_split_stride,
                               b200ps_stream_t stream) {
  B200PS_REQUIRE(B > 0 && H > 0 && C > 0, "mlp_tail: bad shape");
  B200PS_REQUIRE(C <= kMaxClasses, "mlp_tail: C=%d > %d (use gemm + softmax_ce)", C, kMaxClasses);
  B200PS_REQUIRE(H <= 128 && H * C <= 4096, "mlp_tail: H=%d too wide for the fused head (<= 128)", H);
  B200PS_REQUIRE(!train || (dz2 && dw2 && db2 && db1), "mlp_tail: training needs gradient outputs");
  TailParams p;
  p.z2 = z2; p.b1 = b1; p.w2 = w2; p.b2 = b2; p.labels = labels; p.a2_out = a2_out; p.dz2 = dz2;
  p.dw2 = dw2; p.db2 = db2; p.db1 = db1; p.loss_sum = loss_sum; p.correct = correct;
  p.probs_out = probs_out; p.B = B; p.H = H; p.C = C; p.train = train;
  p.grad_scale = grad_scale; p.loss_scale = loss_scale;
  p.flags = flags; p.label_offset = label_offset;
  p.z2_splits = z2_splits < 1 ? 1 : z2_splits; p.z2_split_stride = z2_split_stride;
  p.z2_clear = (flags & B200PS_TAIL_CLEAR_Z2) ? const_cast<float*>(z2) : nullptr;
  // rows are dealt to CTAs in tiles of 8; at most one CTA per SM so the merge stays short
  int ctas = std::min((B + kT2Rows - 1) / kT2Rows, sm_count());
  p.rows_per_cta = ((B + ctas - 1) / ctas + kT2Rows - 1) / kT2Rows * kT2Rows;
  ctas = (B + p.rows_per_cta - 1) / p.rows_per_cta;
  cudaStream_t s = as_stream(stream);
  const int CP = C <= 16 ? 16 : 32;
  const int P = H * C + C + H + 2;
  Workspace ws;
  int rc = get_workspace(sizeof(float) * (size_t)ctas * P, &ws, s, true);
  if (rc) return rc;
  const size_t smem = sizeof(float) * ((size_t)H * (CP + 4) + CP + ((H + 3) & ~3) + 2 * (size_t)kT2Rows * H +
                                       (size_t)kT2Rows * CP);
  if (CP == 16)
    B200PS_CUDA_CHECK(launch_pdl(mlp_tail2_kernel<16>, dim3(ctas), dim3(kT2Threads), smem, s, p,
                                 reinterpret_cast<float*>(ws.data), ws.counters));
  else
    B200PS_CUDA_CHECK(launch_pdl(mlp_tail2_kernel<32>, dim3(ctas), dim3(kT2Threads), smem, s, p,
                                 reinterpret_cast<float*>(ws.data), ws.counters));
  B200PS_LAUNCH_CHECK();
  workspace_mark_used(s);
  return 0;
}

// ------------------------------------------------------------------------------------
// b200ps_mlp_tail_static: the same head computation for hosts that must not allocate.
//
// b200ps_mlp_tail above takes its cross-CTA scratch from the library workspace, which is
// cudaMalloc'ed (and grown) on demand.  Some embedders cannot tolerate that: a serving or
// simulation process that has handed the whole device to its own memory plan (a later
// cudaMalloc from a plug-in fails, or silently eats another tenant's budget), and control
// loops with a latency bound, where cudaMalloc's implicit device-wide synchronisation is not
// acceptable at step time.  For them the partial sums live in a fixed array inside this
// library's device image instead.
//
//   * capacity: kTailStaticFloats floats (1 MiB) of partials + one arrival counter, resident
//     on every GPU that loads libb200ps.so until the process exits;
//   * a launch needs ctas * (H*C + C + H + 2) floats (ctas = row tiles, at most one per SM);
//     larger problems are refused with B200PS_ERR_INVALID — use b200ps_mlp_tail for those;
//   * the array is shared by every call and keeps the previous call's contents: the counter
//     is cleared on the stream before each launch, the partials are always fully rewritten
//     before they are read.  Two calls whose kernels overlap on the GPU (different streams,
//     no ordering between them) race on it — serialise them;
//   * the first call on each device looks the array's device address up with
//     cudaGetSymbolAddress (one-time set-up, cached per device).  Make that first call during
//     start-up, before any stream capture is begun;
//   * no cudaMalloc / cudaFree / allocating library routine is reached from this entry point.
// ------------------------------------------------------------------------------------
constexpr size_t kTailStaticFloats = 256 * 1024;
constexpr size_t kTailStaticCounterWords = 64;
// In host code this name is only a handle for cudaGetSymbolAddress, never a device pointer.
static __device__ __align__(256) float g_tail_static_scratch[kTailStaticFloats + kTailStaticCounterWords];
static std::atomic<float*> g_tail_static_addr[64];   // per-device result of the symbol look-up (zero-initialised)

extern "C" int b200ps_mlp_tail_static(const float* z2, const float* b1, const float* w2, const float* b2,
                                      const int32_t* labels, float* a2_out, float* dz2, float* dw2,
                                      float* db2, float* db1, float* loss_sum, int32_t* correct,
                                      float* probs_out, int32_t B, int32_t H, int32_t C, int32_t train,
                                      float grad_scale, float loss_scale, int32_t flags,
                                      const int32_t* label_offset, int32_t z2_splits,
                                      int64_t z2_split_stride, b200ps_stream_t stream) {
  B200PS_REQUIRE(B > 0 && H > 0 && C > 0, "mlp_tail_static: bad shape");
  B200PS_REQUIRE(C <= kMaxClasses, "mlp_tail_static: C=%d > %d", C, kMaxClasses);
  B200PS_REQUIRE(H <= 128 && H * C <= 4096, "mlp_tail_static: H=%d too wide for the fused head (<= 128)", H);
  B200PS_REQUIRE(!train || (dz2 && dw2 && db2 && db1), "mlp_tail_static: training needs gradient outputs");
  TailParams p;
  p.z2 = z2; p.b1 = b1; p.w2 = w2; p.b2 = b2; p.labels = labels; p.a2_out = a2_out; p.dz2 = dz2;
  p.dw2 = dw2; p.db2 = db2; p.db1 = db1; p.loss_sum = loss_sum; p.correct = correct;
  p.probs_out = probs_out; p.B = B; p.H = H; p.C = C; p.train = train;
  p.grad_scale = grad_scale; p.loss_scale = loss_scale;
  p.flags = flags; p.label_offset = label_offset;
  p.z2_splits = z2_splits < 1 ? 1 : z2_splits; p.z2_split_stride = z2_split_stride;
  p.z2_clear = (flags & B200PS_TAIL_CLEAR_Z2) ? const_cast<float*>(z2) : nullptr;
  int ctas = std::min((B + kT2Rows - 1) / kT2Rows, sm_count());
  p.rows_per_cta = ((B + ctas - 1) / ctas + kT2Rows - 1) / kT2Rows * kT2Rows;
  ctas = (B + p.rows_per_cta - 1) / p.rows_per_cta;
  const int CP = C <= 16 ? 16 : 32;
  const size_t P = (size_t)H * C + C + H + 2;
  const size_t need = (size_t)ctas * P;
  B200PS_REQUIRE(need <= kTailStaticFloats,
                 "mlp_tail_static: B=%d H=%d C=%d needs %zu bytes of static scratch, the built-in maximum is "
                 "%zu bytes (use b200ps_mlp_tail, which allocates a workspace)",
                 B, H, C, need * sizeof(float), kTailStaticFloats * sizeof(float));

  // one-time, per-device set-up: resolve the array's device address with that device current
  int dev = 0;
  B200PS_CUDA_CHECK(cudaGetDevice(&dev));
  B200PS_REQUIRE(dev >= 0 && dev < 64, "mlp_tail_static: device index %d out of range", dev);
  float* scratch = g_tail_static_addr[dev].load(std::memory_order_acquire);
  if (scratch == nullptr) {
    float* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_tail_static_scratch);
    if (e != cudaSuccess || ptr == nullptr)
      return set_error(B200PS_ERR_CUDA, "mlp_tail_static: cudaGetSymbolAddress failed on device %d: %s", dev,
                       cudaGetErrorString(e));
    scratch = ptr;
    g_tail_static_addr[dev].store(ptr, std::memory_order_release);
  }
  unsigned int* counter = reinterpret_cast<unsigned int*>(scratch + kTailStaticFloats);
  cudaStream_t s = as_stream(stream);
  // the arrival counter must start at zero; a previous call that died mid-way may have left it dirty
  B200PS_CUDA_CHECK(cudaMemsetAsync(counter, 0, sizeof(unsigned int), s));
  const size_t smem = sizeof(float) * ((size_t)H * (CP + 4) + CP + ((H + 3) & ~3) + 2 * (size_t)kT2Rows * H +
                                       (size_t)kT2Rows * CP);
  if (CP == 16)
    mlp_tail2_kernel<16><<<ctas, kT2Threads, smem, s>>>(p, scratch, counter);
  else
    mlp_tail2_kernel<32><<<ctas, kT2Threads, smem, s>>>(p, scratch, counter);
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
  else if (C % 8 == 0 && C <= kCeThreads * 8 * kCeVecMax && (ld * 2) % 16 == 0 &&
           (reinterpret_cast<uintptr_t>(logits) & 15) == 0 && (!dlogits || (reinterpret_cast<uintptr_t>(dlogits) & 15) == 0))
    softmax_ce_bf16_vec_kernel<<<B, kCeThreads, 0, s>>>((const __nv_bfloat16*)logits, labels, (__nv_bfloat16*)dlogits,
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

// ------------------------------------------------------------------------------------
// copy_words: n 32-bit words device -> (possibly pinned, device-mapped host) memory as a graph-capturable
// kernel node: how a step graph delivers its loss into a HostFeeder result block without a memcpy node.
// ------------------------------------------------------------------------------------
namespace b200ps {
__global__ void copy_words_kernel(uint32_t* __restrict__ dst, const uint32_t* __restrict__ src, long long n) {
  pdl_wait();
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    dst[i] = __ldcg(src + i);
}
}  // namespace b200ps

extern "C" int b200ps_copy_words(void* dst, const void* src, int64_t n_words, b200ps_stream_t stream) {
  B200PS_REQUIRE(dst && src && n_words > 0, "copy_words: bad arguments");
  const int ctas = (int)std::min<long long>((n_words + 255) / 256, sm_count());
  B200PS_CUDA_CHECK(launch_pdl(copy_words_kernel, dim3(ctas), dim3(256), 0, as_stream(stream),
                               static_cast<uint32_t*>(dst), static_cast<const uint32_t*>(src), (long long)n_words));
  B200PS_LAUNCH_CHECK();
  return 0;
}
