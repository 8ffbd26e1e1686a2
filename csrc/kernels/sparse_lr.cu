// b200ps K4/K5/K10 — sparse logistic regression on a row-sharded [F,1] weight.
//
// The reference keeps W on the parameter server and, every step, ships the batch's unique
// ids up, the gathered rows down and the IndexedSlices gradient up again over gRPC
// (lr2.py:390,395-396).  Here W is row-sharded over the GPUs' HBM and each warp walks one
// example's CSR row, loading W[fid] straight from the owning GPU through its peer-mapped
// pointer and (when training) scattering -lr*dz*fval back with fire-and-forget
// red.global.add.f32 — no id message, no gathered copy, no barrier in async mode.
//
//   z_i = sum_j W[fid_ij]*fval_ij + b ; p_i = sigmoid(z_i)
//   loss_i = max(z,0) - z*y + log1p(exp(-|z|))      (tf.nn.sigmoid_cross_entropy_with_logits)
//   dz_i = (p_i - y_i) * grad_scale
#include "../common.h"
#include "../workspace.h"
#include "ptx.cuh"

namespace b200ps {

constexpr int kSparseThreads = 256;  // 8 warps = 8 examples in flight per CTA

struct SparseParams {
  const int* row_ptr;      // [B+1]
  const int64_t* fid;      // [nnz]
  const float* fval;       // [nnz]
  const float* labels;     // [B]
  float* const* w_shards;  // [world] peer-mapped shard bases
  float* bias;             // owner's bias (peer-mapped), 1 element
  float* pred;             // [B] or null
  float* dz_out;           // [B] or null
  float* loss_sum;         // += sum_i loss_i * loss_scale
  unsigned long long* global_step;  // += 1 (by one thread) or null
  int B;
  int64_t rows_per_shard;
  int64_t features;        // valid ids are [0, features); anything else is skipped and counted
  unsigned long long* bad_ids;
  int shard_shift;         // >= 0 when rows_per_shard is a power of two
  int do_forward, do_update;
  const float* dz_in;      // [B] (update-only launches)
  float lr, grad_scale, loss_scale;
};

__device__ __forceinline__ float* w_addr(const SparseParams& p, int64_t id) {
  int owner;
  int64_t local;
  if (p.shard_shift >= 0) {
    owner = (int)(id >> p.shard_shift);
    local = id & (p.rows_per_shard - 1);
  } else {
    owner = (int)(id / p.rows_per_shard);
    local = id - (int64_t)owner * p.rows_per_shard;
  }
  return p.w_shards[owner] + local;
}

// An id outside [0, features) must never reach w_addr(): its owner index would run off the pointer table
// (embedding_lookup_sparse raises InvalidArgument there, lr2.py:390).  Such non-zeros contribute nothing and
// are counted; the host raises when it next looks at the counter (b200ps_sparse_bad_ids).
__device__ __forceinline__ bool id_ok(const SparseParams& p, int64_t id) {
  return (unsigned long long)id < (unsigned long long)p.features;
}

__global__ void __launch_bounds__(kSparseThreads) sparse_lr_kernel(const SparseParams p) {
  __shared__ float s_loss[kSparseThreads / 32];
  __shared__ float s_dz[kSparseThreads / 32];
  const int lane = threadIdx.x & 31;
  const int warp_in_cta = threadIdx.x >> 5;
  const int warps_total = gridDim.x * (kSparseThreads / 32);
  float loss_acc = 0.f, dz_acc = 0.f;
  float b = 0.f;
  int bad = 0;
  if (p.do_forward) b = __ldcg(p.bias);  // one scalar from the owner (may be remote)

  for (int ex = blockIdx.x * (kSparseThreads / 32) + warp_in_cta; ex < p.B; ex += warps_total) {
    const int beg = p.row_ptr[ex], end = p.row_ptr[ex + 1];
    float dz = p.do_forward ? 0.f : p.dz_in[ex];
    // Rows are walked in batches of 128 non-zeros (4 per lane): ids and values are fetched first so the four row
    // gathers of a lane leave back to back — the gather is pure latency, 7 of 8 rows live on another GPU.  A row
    // that fits one batch (the common case, ~100 non-zeros) keeps its ids / values / validity in registers for
    // the scatter, so the fused step reads the CSR arrays once.
    const bool one_batch = end - beg <= 128;
    int64_t id[4];
    float fv[4];
    bool ok[4];
    float z = 0.f;
    for (int j0 = beg; j0 < end; j0 += 128) {
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int j = j0 + lane + 32 * u;
        id[u] = j < end ? p.fid[j] : -1;
        fv[u] = j < end ? p.fval[j] : 0.f;
        ok[u] = id_ok(p, id[u]);
        if (j < end && !ok[u]) ++bad;
      }
      if (p.do_forward) {
        float w[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) w[u] = ok[u] ? __ldcg(w_addr(p, id[u])) : 0.f;
#pragma unroll
        for (int u = 0; u < 4; ++u) z = fmaf(w[u], fv[u], z);
      } else if (p.do_update) {        // update-only launch: scatter this batch right away
        const float s = -p.lr * dz;
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (ok[u]) red_add_f32_sys(w_addr(p, id[u]), s * fv[u]);
      }
    }
    if (p.do_forward) {
      z = warp_sum(z) + b;
      const float y = p.labels[ex];
      const float pr = 1.f / (1.f + __expf(-z));
      dz = (pr - y) * p.grad_scale;
      if (lane == 0) {
        loss_acc += fmaxf(z, 0.f) - z * y + log1pf(__expf(-fabsf(z)));
        if (p.pred) p.pred[ex] = pr;
        if (p.dz_out) p.dz_out[ex] = dz;
      }
      if (p.do_update) {
        const float s = -p.lr * dz;
        if (one_batch) {
#pragma unroll
          for (int u = 0; u < 4; ++u)
            if (ok[u]) red_add_f32_sys(w_addr(p, id[u]), s * fv[u]);
        } else {
          for (int j = beg + lane; j < end; j += 32) {
            const int64_t f = p.fid[j];
            if (id_ok(p, f)) red_add_f32_sys(w_addr(p, f), s * p.fval[j]);
          }
        }
      }
    }
    if (p.do_update && lane == 0) dz_acc += dz;
  }
  if (bad && p.bad_ids) atomicAdd(p.bad_ids, (unsigned long long)bad);
  if (lane == 0) { s_loss[warp_in_cta] = loss_acc; s_dz[warp_in_cta] = dz_acc; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float l = 0.f, d = 0.f;
    for (int w = 0; w < kSparseThreads / 32; ++w) { l += s_loss[w]; d += s_dz[w]; }
    if (p.do_forward && p.loss_sum) atomicAdd(p.loss_sum, l * p.loss_scale);
    if (p.do_update) {
      red_add_f32_sys(p.bias, -p.lr * d);
      if (blockIdx.x == 0 && p.global_step) atomicAdd_system(p.global_step, 1ull);
    }
  }
}

// Streaming AUC accumulators (tf.contrib.metrics.streaming_auc, lr2.py:400): the bucket of a
// prediction is the number of thresholds strictly below it, so TP/FP/TN/FN at every one of
// the T thresholds follow from two (T+1)-bin histograms by a suffix sum on the host.
__global__ void __launch_bounds__(256) auc_hist_kernel(const float* __restrict__ pred,
                                                       const float* __restrict__ labels, int B,
                                                       const float* __restrict__ thresholds, int T,
                                                       unsigned long long* hist_pos,
                                                       unsigned long long* hist_neg) {
  extern __shared__ unsigned int s_hist[];  // [2*(T+1)] then thresholds [T]
  unsigned int* s_pos = s_hist;
  unsigned int* s_neg = s_hist + (T + 1);
  float* s_thr = reinterpret_cast<float*>(s_hist + 2 * (T + 1));
  for (int i = threadIdx.x; i < 2 * (T + 1); i += blockDim.x) s_hist[i] = 0;
  for (int i = threadIdx.x; i < T; i += blockDim.x) s_thr[i] = thresholds[i];
  __syncthreads();
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < B; i += gridDim.x * blockDim.x) {
    const float pr = pred[i];
    int lo = 0, hi = T;  // first index with thr >= pr  == number of thresholds < pr
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (s_thr[mid] < pr) lo = mid + 1; else hi = mid;
    }
    if (labels[i] > 0.5f) atomicAdd(&s_pos[lo], 1u); else atomicAdd(&s_neg[lo], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < T + 1; i += blockDim.x) {
    if (s_pos[i]) atomicAdd(hist_pos + i, (unsigned long long)s_pos[i]);
    if (s_neg[i]) atomicAdd(hist_neg + i, (unsigned long long)s_neg[i]);
  }
}

}  // namespace b200ps

using namespace b200ps;

extern "C" int b200ps_sparse_lr(const int32_t* row_ptr, const int64_t* fid, const float* fval,
                                const float* labels, float* const* w_shards, float* bias, float* pred,
                                float* dz_out, const float* dz_in, float* loss_sum,
                                uint64_t* global_step, int32_t B, int64_t rows_per_shard, int64_t features,
                                int32_t do_forward, int32_t do_update, float lr, float grad_scale,
                                float loss_scale, b200ps_stream_t stream) {
  B200PS_REQUIRE(B > 0 && rows_per_shard > 0 && features > 0, "sparse_lr: bad shape");
  B200PS_REQUIRE(do_forward || (do_update && dz_in), "sparse_lr: update-only needs dz_in");
  SparseParams p;
  p.row_ptr = row_ptr; p.fid = fid; p.fval = fval; p.labels = labels; p.w_shards = w_shards;
  p.bias = bias; p.pred = pred; p.dz_out = dz_out; p.dz_in = dz_in; p.loss_sum = loss_sum;
  p.global_step = reinterpret_cast<unsigned long long*>(global_step);
  p.B = B; p.rows_per_shard = rows_per_shard; p.features = features;
  p.bad_ids = bad_id_counter();
  p.shard_shift = -1;
  if ((rows_per_shard & (rows_per_shard - 1)) == 0) {
    int s = 0;
    while (((int64_t)1 << s) < rows_per_shard) ++s;
    p.shard_shift = s;
  }
  p.do_forward = do_forward; p.do_update = do_update; p.lr = lr; p.grad_scale = grad_scale;
  p.loss_scale = loss_scale;
  const int warps_per_cta = kSparseThreads / 32;
  int ctas = std::min((B + warps_per_cta - 1) / warps_per_cta, sm_count() * 8);
  sparse_lr_kernel<<<ctas, kSparseThreads, 0, as_stream(stream)>>>(p);
  B200PS_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200ps_sparse_bad_ids(uint64_t* count, int32_t reset) {
  B200PS_REQUIRE(count, "sparse_bad_ids: null output");
  unsigned long long* ctr = bad_id_counter();
  B200PS_REQUIRE(ctr, "sparse_bad_ids: no workspace");
  unsigned long long v = 0;
  B200PS_CUDA_CHECK(cudaMemcpy(&v, ctr, sizeof(v), cudaMemcpyDeviceToHost));
  if (reset && v) B200PS_CUDA_CHECK(cudaMemset(ctr, 0, sizeof(v)));
  *count = v;
  return 0;
}

extern "C" int b200ps_auc_hist(const float* pred, const float* labels, int32_t B, const float* thresholds,
                               int32_t T, uint64_t* hist_pos, uint64_t* hist_neg, b200ps_stream_t stream) {
  B200PS_REQUIRE(B > 0 && T > 1 && T <= 4096, "auc_hist: bad args");
  const int ctas = std::min((B + 255) / 256, sm_count() * 4);
  const size_t smem = sizeof(unsigned) * 2 * (T + 1) + sizeof(float) * T;
  auc_hist_kernel<<<ctas, 256, smem, as_stream(stream)>>>(pred, labels, B, thresholds, T,
                                                         (unsigned long long*)hist_pos,
                                                         (unsigned long long*)hist_neg);
  B200PS_LAUNCH_CHECK();
  return 0;
}
