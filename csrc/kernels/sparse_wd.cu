// b200ps K4w/K5w — sparse wide & deep logistic model on row-sharded tables.
//
// The reference sketches this extension of the LR trainer in a comment (lr2_debug.py:423-428):
//
//   py_x  = embedding_lookup_sparse(W[F,1],  ids, vals, 'sum')                     (wide)
//   py_x += embedding_lookup_sparse(E[F,D],  ids, vals, 'sum') @ w2[D,1]           (deep, D = 128)
//
// In TF that is two gathers on the parameter server, two gRPC round trips for the rows and two
// IndexedSlices pushes per step.  Here W, E and w2 are row-sharded over the GPUs' HBM and one warp
// owns one example: the 32 lanes read a 512-byte embedding row as one coalesced float4 access per
// 128 columns straight from the owning GPU (peer-mapped pointer), accumulate h = sum_j val_j*E[fid_j]
// in registers, dot it with w2, and — when training — scatter
//
//   E[fid_j] -= lr*dz*val_j*w2      (red.global.add.v4.f32, fire and forget, Hogwild)
//   W[fid_j] -= lr*dz*val_j
//   w2       -= lr*sum_i dz_i*h_i   (per-warp register partials -> smem -> one red per column per CTA)
//   b        -= lr*sum_i dz_i
//
// back into the owners, all in the same launch.  No id message, no gathered copy, no barrier (async).
//
// Deferred mode (the default across GPUs).  The embedding update of a step is rank one: every row moves
// along the same vector, dE[fid,:] = c[fid] * w2 with c[fid] = -lr * sum dz*val — and c is exactly the
// wide update dW.  So a worker does not have to ship 512-byte atomic rows over NVLink: it adds the SCALAR
// c into a pending-coefficient table P laid out like W (red.global.add.f32, 4 bytes per non-zero, one lane
// per non-zero), and the gradient of w2 into a pending vector P2.  The owner of a shard folds P in before
// its next step (sparse_wd_apply_kernel): a float4 scan of its slice of P, and for every touched row
// c = atomicExch(P[row], 0), W[row] += c, E[row,:] += c * w2 — plain stores, because only the owner writes W
// and E any more; then w2 += P2 in a second pass so that every update of a step used the old w2.
// Remote traffic per non-zero drops from 516 to 4 bytes, and the update pass of this kernel from a
// serialised row-by-row scatter to one coalesced pass over the non-zeros.
#include "../common.h"
#include "../workspace.h"
#include "ptx.cuh"

namespace b200ps {

constexpr int kWdThreads = 128;            // 4 warps = 4 examples per CTA: a batch of 500 spreads over 125 SMs
constexpr int kWdWarps = kWdThreads / 32;

struct ShardedTable {
  float* const* shards;    // [world] peer-mapped shard bases
  int64_t rows_per_shard;
  int shift;               // >= 0 when rows_per_shard is a power of two
  __device__ __forceinline__ float* row(int64_t id, int dim) const {
    int owner;
    int64_t local;
    if (shift >= 0) {
      owner = (int)(id >> shift);
      local = id & (rows_per_shard - 1);
    } else {
      owner = (int)(id / rows_per_shard);
      local = id - (int64_t)owner * rows_per_shard;
    }
    return shards[owner] + local * dim;
  }
};

struct WdParams {
  const int* row_ptr;      // [B+1]
  const int64_t* fid;      // [nnz]
  const float* fval;       // [nnz]
  const float* labels;     // [B]
  ShardedTable wide;       // [F,1]   (shards == nullptr: no wide part)
  ShardedTable emb;        // [F,D]
  ShardedTable w2;         // [D,1]
  ShardedTable pend;       // [F,1] pending coefficients (shards == nullptr: direct mode)
  ShardedTable pend2;      // [D,1] pending w2 update
  float* bias;             // owner's bias (peer-mapped), 1 element
  float* pred;             // [B] or null
  float* dz_out;           // [B] or null
  const float* dz_in;      // [B] (update-only launches)
  float* loss_sum;
  unsigned long long* global_step;
  int B;
  int do_forward, do_update;
  float lr, grad_scale, loss_scale;
  int64_t features;        // valid ids are [0, features); others are skipped (weight 0) and counted
  unsigned long long* bad_ids;
};

__device__ __forceinline__ float4 ldcg4(const float* p) { return __ldcg(reinterpret_cast<const float4*>(p)); }

template <int NCH>   // D = 128 * NCH columns; lane l owns columns [(c*32+l)*4, +4) of chunk c
__global__ void __launch_bounds__(kWdThreads) sparse_wd_kernel(const WdParams p) {
  constexpr int D = 128 * NCH;
  __shared__ float s_g2[kWdWarps][D];
  __shared__ float s_loss[kWdWarps];
  __shared__ float s_dz[kWdWarps];
  const int lane = threadIdx.x & 31;
  const int warp_in_cta = threadIdx.x >> 5;
  const int warps_total = gridDim.x * kWdWarps;
  const bool has_wide = p.wide.shards != nullptr;

  float4 w2r[NCH], g2[NCH];
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    w2r[c] = ldcg4(p.w2.row((c * 32 + lane) * 4, 1));   // a 4-aligned group never straddles shards
    g2[c] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  const float b = p.do_forward ? __ldcg(p.bias) : 0.f;
  float loss_acc = 0.f, dz_acc = 0.f;
  int bad = 0;

  for (int ex = blockIdx.x * kWdWarps + warp_in_cta; ex < p.B; ex += warps_total) {
    const int beg = p.row_ptr[ex], end = p.row_ptr[ex + 1];
    float4 h[NCH];
#pragma unroll
    for (int c = 0; c < NCH; ++c) h[c] = make_float4(0.f, 0.f, 0.f, 0.f);
    float zw = 0.f;
    // ---- gather: lanes fetch 32 (id, value) pairs, then the warp walks them FL rows at a time
    for (int base = beg; base < end; base += 32) {
      const int j = base + lane;
      bool live = j < end;
      int64_t id = live ? p.fid[j] : 0;                // dead lanes name row 0 with weight 0
      if (live && (unsigned long long)id >= (unsigned long long)p.features) { live = false; id = 0; ++bad; }
      const float v = live ? p.fval[j] : 0.f;
      if (has_wide && live && p.do_forward) zw = fmaf(__ldcg(p.wide.row(id, 1)), v, zw);
      const unsigned long long rowp = reinterpret_cast<unsigned long long>(p.emb.row(id, D));
      const int n = min(32, end - base);
      constexpr int FL = NCH == 1 ? 16 : (NCH == 2 ? 8 : 4);         // embedding rows in flight per warp (the gather is pure latency)
      for (int k = 0; k < n; k += FL) {
        float4 e[FL][NCH];
        float vk[FL];
#pragma unroll
        for (int u = 0; u < FL; ++u) {
          const float* r = reinterpret_cast<const float*>(__shfl_sync(0xffffffffu, rowp, (k + u) & 31));
          vk[u] = __shfl_sync(0xffffffffu, v, (k + u) & 31);
#pragma unroll
          for (int c = 0; c < NCH; ++c) e[u][c] = ldcg4(r + (c * 32 + lane) * 4);
        }
#pragma unroll
        for (int u = 0; u < FL; ++u) {
#pragma unroll
          for (int c = 0; c < NCH; ++c) {
            h[c].x = fmaf(vk[u], e[u][c].x, h[c].x);
            h[c].y = fmaf(vk[u], e[u][c].y, h[c].y);
            h[c].z = fmaf(vk[u], e[u][c].z, h[c].z);
            h[c].w = fmaf(vk[u], e[u][c].w, h[c].w);
          }
        }
      }
    }
    float dz;
    if (p.do_forward) {
      float part = zw;
#pragma unroll
      for (int c = 0; c < NCH; ++c)
        part += h[c].x * w2r[c].x + h[c].y * w2r[c].y + h[c].z * w2r[c].z + h[c].w * w2r[c].w;
      const float z = warp_sum(part) + b;
      const float y = p.labels[ex];
      const float pr = 1.f / (1.f + __expf(-z));
      dz = (pr - y) * p.grad_scale;
      if (lane == 0) {
        loss_acc += fmaxf(z, 0.f) - z * y + log1pf(__expf(-fabsf(z)));
        if (p.pred) p.pred[ex] = pr;
        if (p.dz_out) p.dz_out[ex] = dz;
      }
    } else {
      dz = p.dz_in[ex];
    }
    if (p.do_update) {
      const float s = -p.lr * dz;
#pragma unroll
      for (int c = 0; c < NCH; ++c) {
        g2[c].x = fmaf(dz, h[c].x, g2[c].x);
        g2[c].y = fmaf(dz, h[c].y, g2[c].y);
        g2[c].z = fmaf(dz, h[c].z, g2[c].z);
        g2[c].w = fmaf(dz, h[c].w, g2[c].w);
      }
      if (p.pend.shards) {                            // deferred: one scalar per non-zero, lane-parallel
        for (int j = beg + lane; j < end; j += 32) {
          const int64_t id = p.fid[j];
          if ((unsigned long long)id < (unsigned long long)p.features) red_add_f32_sys(p.pend.row(id, 1), s * p.fval[j]);
        }
      } else {
        for (int base = beg; base < end; base += 32) {
          const int j = base + lane;
          bool live = j < end;
          int64_t id = live ? p.fid[j] : 0;
          if (live && (unsigned long long)id >= (unsigned long long)p.features) { live = false; id = 0; }
          const float v = live ? p.fval[j] : 0.f;      // a refused id adds 0 to row 0
          if (has_wide && live) red_add_f32_sys(p.wide.row(id, 1), s * v);
          const unsigned long long rowp = reinterpret_cast<unsigned long long>(p.emb.row(id, D));
          const int n = min(32, end - base);
          for (int k = 0; k < n; ++k) {
            float* r = reinterpret_cast<float*>(__shfl_sync(0xffffffffu, rowp, k));
            const float sv = s * __shfl_sync(0xffffffffu, v, k);
  #pragma unroll
            for (int c = 0; c < NCH; ++c)
              red_add_v4f32_sys(r + (c * 32 + lane) * 4,
                                make_float4(sv * w2r[c].x, sv * w2r[c].y, sv * w2r[c].z, sv * w2r[c].w));
          }
        }
      }
      if (lane == 0) dz_acc += dz;
    }
  }

  if (bad && p.bad_ids) atomicAdd(p.bad_ids, (unsigned long long)bad);
  if (p.do_update) {
#pragma unroll
    for (int c = 0; c < NCH; ++c)
      *reinterpret_cast<float4*>(&s_g2[warp_in_cta][(c * 32 + lane) * 4]) = g2[c];
  }
  if (lane == 0) { s_loss[warp_in_cta] = loss_acc; s_dz[warp_in_cta] = dz_acc; }
  __syncthreads();
  if (p.do_update) {
    for (int col = threadIdx.x; col < D; col += kWdThreads) {
      float g = 0.f;
#pragma unroll
      for (int w = 0; w < kWdWarps; ++w) g += s_g2[w][col];
      if (g != 0.f) red_add_f32_sys(p.pend2.shards ? p.pend2.row(col, 1) : p.w2.row(col, 1), -p.lr * g);
    }
  }
  if (threadIdx.x == 0) {
    float l = 0.f, d = 0.f;
    for (int w = 0; w < kWdWarps; ++w) { l += s_loss[w]; d += s_dz[w]; }
    if (p.do_forward && p.loss_sum) atomicAdd(p.loss_sum, l * p.loss_scale);
    if (p.do_update) {
      red_add_f32_sys(p.bias, -p.lr * d);
      if (blockIdx.x == 0 && p.global_step) atomicAdd_system(p.global_step, 1ull);
    }
  }
}

// Owner-side fold of the pending coefficients (deferred mode).  phase 1: rows of this rank's shard;
// phase 2: this rank's slice of w2.  One lane looks at one row of P; the warp then visits the touched rows
// of its 32 together (ballot), every lane updating 4 columns of each 128-column chunk.
// With `pend2` (owner-side step, sparse_wd_owner.cu) the fold also accumulates the gradient of w2 that the workers
// could not form (they never saw the rows): dw2 = sum_f c_f * E_old[f], added into the pending w2 vector.
template <int NCH>
__global__ void __launch_bounds__(256) sparse_wd_apply_kernel(float* __restrict__ pend, float* __restrict__ wide,
                                                              float* __restrict__ emb, long long rows,
                                                              const ShardedTable w2, const ShardedTable pend2) {
  constexpr int D = 128 * NCH;
  const int lane = threadIdx.x & 31;
  float4 w2r[NCH], g2[NCH];
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    w2r[c] = ldcg4(w2.row((c * 32 + lane) * 4, 1));
    g2[c] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
  // 128 rows of P per warp iteration (one float4 per lane): the scan is a plain streaming read, ~1 % of the
  // rows are touched in a step
  for (long long r0 = warp * 128; r0 < rows; r0 += nwarps * 128) {
    const long long r = r0 + lane * 4;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r + 3 < rows) v = __ldcg(reinterpret_cast<const float4*>(pend + r));
    else {
      if (r < rows) v.x = __ldcg(pend + r);
      if (r + 1 < rows) v.y = __ldcg(pend + r + 1);
      if (r + 2 < rows) v.z = __ldcg(pend + r + 2);
    }
    float c[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (c[k] != 0.f) {
        c[k] = atomicExch_system(pend + r + k, 0.f);   // claim what has landed so far (peers add with sys-scope reds)
        if (wide) wide[r + k] += c[k];
      }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      unsigned m = __ballot_sync(0xffffffffu, c[k] != 0.f);
      while (m) {
        const int src = __ffs(m) - 1;
        m &= m - 1;
        const float cv = __shfl_sync(0xffffffffu, c[k], src);
        float* row = emb + (r0 + src * 4 + k) * D;
#pragma unroll
        for (int ch = 0; ch < NCH; ++ch) {
          float4* q = reinterpret_cast<float4*>(row + (ch * 32 + lane) * 4);
          float4 e = *q;
          g2[ch].x = fmaf(cv, e.x, g2[ch].x); g2[ch].y = fmaf(cv, e.y, g2[ch].y);
          g2[ch].z = fmaf(cv, e.z, g2[ch].z); g2[ch].w = fmaf(cv, e.w, g2[ch].w);
          e.x = fmaf(cv, w2r[ch].x, e.x); e.y = fmaf(cv, w2r[ch].y, e.y);
          e.z = fmaf(cv, w2r[ch].z, e.z); e.w = fmaf(cv, w2r[ch].w, e.w);
          *q = e;
        }
      }
    }
  }
  if (pend2.shards) {                  // one red per column per CTA into the owners of the pending w2 vector
    __shared__ float s_g2[8][D];
    const int wic = threadIdx.x >> 5;
#pragma unroll
    for (int c = 0; c < NCH; ++c) *reinterpret_cast<float4*>(&s_g2[wic][(c * 32 + lane) * 4]) = g2[c];
    __syncthreads();
    for (int col = threadIdx.x; col < D; col += 256) {
      float g = 0.f;
#pragma unroll
      for (int w = 0; w < 8; ++w) g += s_g2[w][col];
      if (g != 0.f) red_add_f32_sys(pend2.row(col, 1), g);
    }
  }
}

__global__ void sparse_wd_apply_w2_kernel(float* __restrict__ pend2, float* __restrict__ w2, long long n) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    if (__ldcg(pend2 + i) != 0.f) w2[i] += atomicExch_system(pend2 + i, 0.f);
  }
}

static ShardedTable make_table(float* const* shards, int64_t rps) {
  ShardedTable t;
  t.shards = shards;
  t.rows_per_shard = rps;
  t.shift = -1;
  if (rps > 0 && (rps & (rps - 1)) == 0) {
    int s = 0;
    while (((int64_t)1 << s) < rps) ++s;
    t.shift = s;
  }
  return t;
}

}  // namespace b200ps

using namespace b200ps;

extern "C" int b200ps_sparse_wd(const int32_t* row_ptr, const int64_t* fid, const float* fval, const float* labels,
                                float* const* wide_shards, int64_t wide_rows_per_shard, float* const* emb_shards,
                                int64_t emb_rows_per_shard, int32_t dim, float* const* w2_shards,
                                int64_t w2_rows_per_shard, float* bias, float* pred, float* dz_out,
                                const float* dz_in, float* loss_sum, uint64_t* global_step, int32_t B,
                                int32_t do_forward, int32_t do_update, float lr, float grad_scale,
                                float loss_scale, float* const* pend_shards, int64_t pend_rows_per_shard,
                                float* const* pend2_shards, int64_t pend2_rows_per_shard, int64_t features,
                                b200ps_stream_t stream) {
  B200PS_REQUIRE(B > 0 && emb_shards && w2_shards && bias && features > 0, "sparse_wd: bad arguments");
  B200PS_REQUIRE((pend_shards == nullptr) == (pend2_shards == nullptr), "sparse_wd: pend and pend2 go together");
  B200PS_REQUIRE(!pend_shards || (pend_rows_per_shard == emb_rows_per_shard && pend2_rows_per_shard == w2_rows_per_shard),
                 "sparse_wd: pending tables must be sharded like the embedding / w2");
  B200PS_REQUIRE(dim == 128 || dim == 256 || dim == 512, "sparse_wd: dim must be 128, 256 or 512 (got %d)", dim);
  B200PS_REQUIRE(emb_rows_per_shard > 0 && w2_rows_per_shard > 0 && w2_rows_per_shard % 4 == 0,
                 "sparse_wd: w2 rows_per_shard must be a positive multiple of 4");
  B200PS_REQUIRE(!wide_shards || wide_rows_per_shard > 0, "sparse_wd: bad wide rows_per_shard");
  B200PS_REQUIRE(do_forward || (do_update && dz_in), "sparse_wd: update-only needs dz_in");
  WdParams p;
  p.row_ptr = row_ptr; p.fid = fid; p.fval = fval; p.labels = labels;
  p.wide = make_table(wide_shards, wide_rows_per_shard);
  p.emb = make_table(emb_shards, emb_rows_per_shard);
  p.w2 = make_table(w2_shards, w2_rows_per_shard);
  p.pend = make_table(pend_shards, pend_rows_per_shard);
  p.pend2 = make_table(pend2_shards, pend2_rows_per_shard);
  p.bias = bias; p.pred = pred; p.dz_out = dz_out; p.dz_in = dz_in; p.loss_sum = loss_sum;
  p.global_step = reinterpret_cast<unsigned long long*>(global_step);
  p.B = B; p.do_forward = do_forward; p.do_update = do_update;
  p.lr = lr; p.grad_scale = grad_scale; p.loss_scale = loss_scale;
  p.features = features; p.bad_ids = bad_id_counter();
  const int ctas = std::min((B + kWdWarps - 1) / kWdWarps, sm_count() * 4);
  cudaStream_t s = as_stream(stream);
  if (dim == 128) sparse_wd_kernel<1><<<ctas, kWdThreads, 0, s>>>(p);
  else if (dim == 256) sparse_wd_kernel<2><<<ctas, kWdThreads, 0, s>>>(p);
  else sparse_wd_kernel<4><<<ctas, kWdThreads, 0, s>>>(p);
  B200PS_LAUNCH_CHECK();
  return 0;
}

// Deferred mode, owner side.  phase 1: fold this rank's slice of the pending coefficients into its rows of W
// (optional) and E; phase 2: fold its slice of the pending w2 update.  All pointers are LOCAL shard bases.
extern "C" int b200ps_sparse_wd_apply(float* pend_local, float* wide_local, float* emb_local, int64_t rows_local,
                                      int32_t dim, float* const* w2_shards, int64_t w2_rows_per_shard,
                                      float* pend2_local, float* w2_local, int64_t w2_rows_local, int32_t phase,
                                      float* const* pend2_shards, b200ps_stream_t stream) {
  B200PS_REQUIRE(dim == 128 || dim == 256 || dim == 512, "sparse_wd_apply: dim must be 128, 256 or 512");
  cudaStream_t s = as_stream(stream);
  if (phase == 1) {
    if (rows_local <= 0) return 0;
    B200PS_REQUIRE(pend_local && emb_local && w2_shards && w2_rows_per_shard > 0, "sparse_wd_apply: bad arguments");
    const ShardedTable w2 = make_table(w2_shards, w2_rows_per_shard);
    const ShardedTable p2 = make_table(pend2_shards, w2_rows_per_shard);   // null shards: no w2-gradient accumulation
    const int ctas = (int)std::min<long long>((rows_local + 1023) / 1024, (long long)sm_count() * 8);
    if (dim == 128) sparse_wd_apply_kernel<1><<<ctas, 256, 0, s>>>(pend_local, wide_local, emb_local, rows_local, w2, p2);
    else if (dim == 256) sparse_wd_apply_kernel<2><<<ctas, 256, 0, s>>>(pend_local, wide_local, emb_local, rows_local, w2, p2);
    else sparse_wd_apply_kernel<4><<<ctas, 256, 0, s>>>(pend_local, wide_local, emb_local, rows_local, w2, p2);
  } else {
    if (w2_rows_local <= 0) return 0;
    B200PS_REQUIRE(pend2_local && w2_local, "sparse_wd_apply: bad arguments");
    sparse_wd_apply_w2_kernel<<<(int)((w2_rows_local + 255) / 256), 256, 0, s>>>(pend2_local, w2_local, w2_rows_local);
  }
  B200PS_LAUNCH_CHECK();
  return 0;
}
