// b200ps K4o/K5o — the sparse wide & deep step across GPUs with OWNER-SIDE row work (lr2_debug.py:423-428;
// embedding_lookup_sparse runs its gather on the parameter server, lr2.py:390).
//
// The direct kernel (sparse_wd.cu) moves every 512-byte embedding row to the worker and every row update back as
// remote atomics: ~1 KB of NVLink traffic per non-zero, read-latency bound — 10 M examples/s flat from 2 to 8 GPUs.
// But the model only needs a SCALAR per non-zero in each direction:
//
//   z_i = b + sum_j v_ij * (W[f_ij] + E[f_ij] . w2)            forward: s_f = W[f] + E[f] . w2 is computed by f's OWNER
//   dE[f] = c_f * w2,  dW[f] = c_f,  dw2 = sum_f c_f * E[f]    backward: c_f = -lr * sum dz_i v_ij is a scalar, the
//                                                              rank-one row update and the w2 gradient are OWNER work
//
// so a step becomes collective phases separated by the cross-GPU barrier kernel; the remote traffic is posted writes
// of 4-8 bytes, no row ever crosses a link, and the only remote READ is the (sharded) w2 vector, once per CTA of the
// serve and fold kernels (once per warp it was half of the step at 8 GPUs):
//
//   request   every rank buckets its batch's non-zeros by owner (warp-aggregated slot allocation) and writes
//             {local row, non-zero index} into the owner's inbox; the last CTA publishes the per-owner counts
//   serve     every rank walks its inbox: a warp reads one LOCAL row (coalesced 512 B), dots it with w2, adds the
//             wide weight and writes the scalar into the requester's response array in INBOX order (a warp's 32
//             answers are one contiguous 256-B remote write; the requester kept non-zero -> slot in slot_tab)
//   loss      warp per example: z, sigmoid, logistic loss, dz; c = -lr*dz*v goes as a scalar red.add into the
//             owner's pending table (the deferred-update table of sparse_wd.cu); bias and global_step as before
//   fold      (b200ps_sparse_wd_apply, phase 1 extended) the owner folds the pending scalars into W and E and
//             accumulates dw2 = sum_f c_f * E_old[f] into the pending w2 vector; phase 2 folds that into w2
//
// 24 bytes on the wire per non-zero instead of ~1 KB.
#include "../common.h"
#include "../workspace.h"
#include "ptx.cuh"

namespace b200ps {

constexpr int kOwnThreads = 256;
constexpr int kMaxWorldOwn = 32;

struct OwnerParams {
  const int* row_ptr; const int64_t* fid; const float* fval; const float* labels;
  int B, world, rank, cap;                 // cap: capacity of the batch's non-zero arrays (inbox stride)
  int64_t rows_per_shard, features;
  int shift;                               // >= 0 when rows_per_shard is a power of two
  uint2* const* inbox;                     // [world] peer-mapped: inbox[o] + src*cap + slot = {local row, nz index}
  unsigned int* const* inbox_cnt;          // [world] peer-mapped: inbox_cnt[o][src]
  float2* const* resp;                     // [world] peer-mapped: resp[r][owner*cap + slot] = {E[f].w2, W[f]} (inbox order)
  int* slot_tab;                           // [cap] local: non-zero index -> owner*cap + slot (-1: refused id)
  unsigned int* count;                     // [world + 1] local: slots handed out per owner | CTA arrival counter
  unsigned long long* bad_ids;
};

__device__ __forceinline__ int own_owner(const OwnerParams& p, int64_t id, int64_t& local) {
  int o;
  if (p.shift >= 0) { o = (int)(id >> p.shift); local = id & (p.rows_per_shard - 1); }
  else { o = (int)(id / p.rows_per_shard); local = id - (int64_t)o * p.rows_per_shard; }
  return o;
}

// The (sharded, mostly remote) w2 vector into shared memory: D/4 threads of the CTA fetch one float4 each.
template <int D>
__device__ __forceinline__ void stage_w2(float* s_w2, float* const* w2_shards, int64_t w2_rows_per_shard) {
  for (int q = threadIdx.x; q < D / 4; q += kOwnThreads) {
    const int col = q * 4;                             // a 4-aligned group never straddles w2 shards
    *reinterpret_cast<float4*>(s_w2 + col) =
        __ldcg(reinterpret_cast<const float4*>(w2_shards[col / w2_rows_per_shard] + col % w2_rows_per_shard));
  }
}

// ---- phase 1: bucket the batch's non-zeros by owner and post the requests
__global__ void __launch_bounds__(kOwnThreads) wd_request_kernel(const OwnerParams p) {
  const int nnz = p.row_ptr[p.B];
  const int lane = threadIdx.x & 31;
  int bad = 0;
  for (int j0 = blockIdx.x * kOwnThreads; j0 < nnz; j0 += gridDim.x * kOwnThreads) {
    const int j = j0 + threadIdx.x;
    int o = -1;
    int64_t local = 0;
    if (j < nnz) {
      const int64_t id = p.fid[j];
      if ((unsigned long long)id < (unsigned long long)p.features) o = own_owner(p, id, local);
      else { ++bad; p.slot_tab[j] = -1; }
    }
    // warp-aggregated slot allocation: one atomic per (warp, owner)
    const unsigned peers = __match_any_sync(0xffffffffu, o);
    const int leader = __ffs(peers) - 1;
    unsigned base = 0;
    if (o >= 0 && lane == leader) base = atomicAdd(p.count + o, (unsigned)__popc(peers));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (o >= 0) {
      const unsigned slot = base + (unsigned)__popc(peers & ((1u << lane) - 1u));
      p.inbox[o][(size_t)p.rank * p.cap + slot] = make_uint2((unsigned)local, (unsigned)j);
      p.slot_tab[j] = o * p.cap + (int)slot;       // the answer comes back at the same (owner, slot): coalesced remote writes
    }
  }
  if (bad && p.bad_ids) atomicAdd(p.bad_ids, (unsigned long long)bad);
  // the last CTA out publishes this rank's per-owner request counts (and re-arms the local counters)
  __syncthreads();
  __shared__ bool s_last;
  if (threadIdx.x == 0) {
    __threadfence();
    s_last = atomicAdd(p.count + p.world, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (s_last && threadIdx.x < p.world) {
    const unsigned c = atomicExch(p.count + threadIdx.x, 0u);
    p.inbox_cnt[threadIdx.x][p.rank] = c;
    if (threadIdx.x == 0) p.count[p.world] = 0;
  }
}

// ---- phase 2: serve the inbox — one warp per requested row, D = 128 * NCH
template <int NCH>
__global__ void __launch_bounds__(kOwnThreads) wd_serve_kernel(const OwnerParams p, const float* __restrict__ emb,
                                                              const float* __restrict__ wide, float* const* w2_shards,
                                                              int64_t w2_rows_per_shard) {
  constexpr int D = 128 * NCH;
  const int lane = threadIdx.x & 31;
  // w2 lives on world shards: ONE remote read of it per CTA (staged in shared memory), not one per warp — remote reads
  // are round trips, 2368 warps x 7 peers of them were a visible part of the kernel at 8 GPUs
  __shared__ __align__(16) float s_w2[D];
  stage_w2<D>(s_w2, w2_shards, w2_rows_per_shard);
  // flattened request space: [src 0 requests | src 1 requests | ...]
  __shared__ unsigned s_pref[kMaxWorldOwn + 1];
  if (threadIdx.x == kOwnThreads - 1) {
    unsigned acc = 0;
    for (int r = 0; r < p.world; ++r) { s_pref[r] = acc; acc += __ldcg(p.inbox_cnt[p.rank] + r); }
    s_pref[p.world] = acc;
  }
  __syncthreads();
  float4 w2r[NCH];
#pragma unroll
  for (int c = 0; c < NCH; ++c) w2r[c] = *reinterpret_cast<const float4*>(&s_w2[(c * 32 + lane) * 4]);
  const unsigned total = s_pref[p.world];
  const uint2* inbox = p.inbox[p.rank];
  const unsigned warp = (blockIdx.x * kOwnThreads + threadIdx.x) >> 5, nwarps = (gridDim.x * kOwnThreads) >> 5;
  constexpr int FL = NCH == 1 ? 8 : 4;                // rows in flight per warp
  // A warp takes 32 consecutive requests: lane l fetches request l (one coalesced load for the chunk — a row load
  // per request behind its own dependent request load would serialise 8 L2 latencies), the rows are then visited
  // FL at a time, and lane l finally posts the answer of request l (32 stores in one instruction).
  for (unsigned t0 = warp * 32; t0 < total; t0 += nwarps * 32) {
    const unsigned t = t0 + lane;
    uint2 rq = make_uint2(0u, 0u);
    int src = -1;
    if (t < total) {
      int r = 0;
      while (t >= s_pref[r + 1]) ++r;
      src = r;
      rq = __ldcg(inbox + (size_t)r * p.cap + (t - s_pref[r]));
    }
    const int n = (int)min(32u, total - t0);
    float mine = 0.f;
    for (int k0 = 0; k0 < n; k0 += FL) {
      float4 e[FL][NCH];
#pragma unroll
      for (int u = 0; u < FL; ++u) {
        const unsigned row = __shfl_sync(0xffffffffu, rq.x, (k0 + u) & 31);
        if (k0 + u < n) {
          const float* rp = emb + (size_t)row * D;
#pragma unroll
          for (int c = 0; c < NCH; ++c) e[u][c] = __ldcg(reinterpret_cast<const float4*>(rp + (c * 32 + lane) * 4));
        }
      }
#pragma unroll
      for (int u = 0; u < FL; ++u) {
        float part = 0.f;
        if (k0 + u < n) {
#pragma unroll
          for (int c = 0; c < NCH; ++c)
            part += e[u][c].x * w2r[c].x + e[u][c].y * w2r[c].y + e[u][c].z * w2r[c].z + e[u][c].w * w2r[c].w;
        }
        const float sdot = warp_sum(part);
        if (lane == k0 + u) mine = sdot;
      }
    }
    if (src >= 0)
      p.resp[src][(size_t)p.rank * p.cap + (t - s_pref[src])] = make_float2(mine, wide ? __ldcg(wide + rq.x) : 0.f);
  }
}

// ---- phase 3: loss and the scalar updates
struct OwnerLoss {
  float* const* pend; int64_t pend_rows_per_shard; int pend_shift;     // pending coefficient table (sharded like E)
  float* bias; float* bias_pend; float* pred; float* dz_out; float* loss_sum; unsigned long long* global_step;
  float lr, grad_scale, loss_scale;
  int do_update;
};

__global__ void __launch_bounds__(kOwnThreads) wd_loss_kernel(const OwnerParams p, const OwnerLoss q) {
  __shared__ float s_loss[kOwnThreads / 32], s_dz[kOwnThreads / 32];
  const int lane = threadIdx.x & 31, wic = threadIdx.x >> 5;
  const int warps_total = gridDim.x * (kOwnThreads / 32);
  const float2* resp = p.resp[p.rank];
  const float b = __ldcg(q.bias);
  float loss_acc = 0.f, dz_acc = 0.f;
  for (int ex = blockIdx.x * (kOwnThreads / 32) + wic; ex < p.B; ex += warps_total) {
    const int beg = p.row_ptr[ex], end = p.row_ptr[ex + 1];
    float z = 0.f;
    for (int j = beg + lane; j < end; j += 32) {
      const int st = p.slot_tab[j];
      if (st >= 0) {
        const float2 r = __ldcg(resp + st);
        z = fmaf(p.fval[j], r.x + r.y, z);
      }
    }
    z = warp_sum(z) + b;
    const float y = p.labels[ex];
    const float pr = 1.f / (1.f + __expf(-z));
    const float dz = (pr - y) * q.grad_scale;
    if (lane == 0) {
      loss_acc += fmaxf(z, 0.f) - z * y + log1pf(__expf(-fabsf(z)));
      if (q.pred) q.pred[ex] = pr;
      if (q.dz_out) q.dz_out[ex] = dz;
      dz_acc += dz;
    }
    if (q.do_update) {
      const float s = -q.lr * dz;
      for (int j = beg + lane; j < end; j += 32) {
        const int64_t id = p.fid[j];
        if ((unsigned long long)id < (unsigned long long)p.features) {
          int64_t local;
          int o;
          if (q.pend_shift >= 0) { o = (int)(id >> q.pend_shift); local = id & (q.pend_rows_per_shard - 1); }
          else { o = (int)(id / q.pend_rows_per_shard); local = id - (int64_t)o * q.pend_rows_per_shard; }
          red_add_f32_sys(q.pend[o] + local, s * p.fval[j]);
        }
      }
    }
  }
  if (lane == 0) { s_loss[wic] = loss_acc; s_dz[wic] = dz_acc; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float l = 0.f, d = 0.f;
    for (int w = 0; w < kOwnThreads / 32; ++w) { l += s_loss[w]; d += s_dz[w]; }
    if (q.loss_sum) atomicAdd(q.loss_sum, l * q.loss_scale);
    if (q.do_update) {
      // the bias moves only in the fold: every rank of the round must read the same value
      red_add_f32_sys(q.bias_pend ? q.bias_pend : q.bias, -q.lr * d);
      if (blockIdx.x == 0 && q.global_step) atomicAdd_system(q.global_step, 1ull);
    }
  }
}

// ---- phase 4: the owner folds the round's coefficients into its rows — driven by the inbox (exactly the rows
// that were requested this round), not by a scan of the whole pending table.  The first visitor of a row claims its
// coefficient (local atomicExch; later duplicates read 0 and skip): W[row] += c, dw2 += c * E_old[row],
// E[row] += c * w2.  One red per column per CTA adds dw2 into the pending w2 vector.
template <int NCH>
__global__ void __launch_bounds__(kOwnThreads) wd_fold_kernel(const OwnerParams p, float* __restrict__ pend,
                                                             float* __restrict__ emb, float* __restrict__ wide,
                                                             float* const* w2_shards, float* const* pend2_shards,
                                                             int64_t w2_rows_per_shard) {
  constexpr int D = 128 * NCH;
  __shared__ float s_g2[kOwnThreads / 32][D];
  __shared__ unsigned s_pref[kMaxWorldOwn + 1];
  const int lane = threadIdx.x & 31, wic = threadIdx.x >> 5;
  __shared__ __align__(16) float s_w2[D];
  stage_w2<D>(s_w2, w2_shards, w2_rows_per_shard);
  if (threadIdx.x == kOwnThreads - 1) {
    unsigned acc = 0;
    for (int r = 0; r < p.world; ++r) { s_pref[r] = acc; acc += __ldcg(p.inbox_cnt[p.rank] + r); }
    s_pref[p.world] = acc;
  }
  __syncthreads();
  float4 w2r[NCH], g2[NCH];
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    w2r[c] = *reinterpret_cast<const float4*>(&s_w2[(c * 32 + lane) * 4]);
    g2[c] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  const unsigned total = s_pref[p.world];
  const uint2* inbox = p.inbox[p.rank];
  const unsigned warp = (blockIdx.x * kOwnThreads + threadIdx.x) >> 5, nwarps = (gridDim.x * kOwnThreads) >> 5;
  constexpr int FL = 4;
  for (unsigned t0 = warp * 32; t0 < total; t0 += nwarps * 32) {     // lane l claims the row of request l
    const unsigned t = t0 + lane;
    unsigned row = 0;
    float cmine = 0.f;
    if (t < total) {
      int r = 0;
      while (t >= s_pref[r + 1]) ++r;
      row = __ldcg(inbox + (size_t)r * p.cap + (t - s_pref[r])).x;
      cmine = atomicExch(pend + row, 0.f);                            // duplicates of a row read 0 and skip
      if (wide && cmine != 0.f) wide[row] += cmine;
    }
    unsigned live = __ballot_sync(0xffffffffu, cmine != 0.f);
    while (live) {
      unsigned rows[FL];
      float cv[FL];
      float4 e[FL][NCH];
#pragma unroll
      for (int u = 0; u < FL; ++u) {
        cv[u] = 0.f;
        rows[u] = 0;
        if (live) {
          const int srcl = __ffs(live) - 1;
          live &= live - 1;
          rows[u] = __shfl_sync(0xffffffffu, row, srcl);
          cv[u] = __shfl_sync(0xffffffffu, cmine, srcl);
          const float* er = emb + (size_t)rows[u] * D;
#pragma unroll
          for (int c = 0; c < NCH; ++c) e[u][c] = *reinterpret_cast<const float4*>(er + (c * 32 + lane) * 4);
        }
      }
#pragma unroll
      for (int u = 0; u < FL; ++u) {
        if (cv[u] == 0.f) continue;                                    // warp-uniform
        float* er = emb + (size_t)rows[u] * D;
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
          float4 v = e[u][c];
          g2[c].x = fmaf(cv[u], v.x, g2[c].x); g2[c].y = fmaf(cv[u], v.y, g2[c].y);
          g2[c].z = fmaf(cv[u], v.z, g2[c].z); g2[c].w = fmaf(cv[u], v.w, g2[c].w);
          v.x = fmaf(cv[u], w2r[c].x, v.x); v.y = fmaf(cv[u], w2r[c].y, v.y);
          v.z = fmaf(cv[u], w2r[c].z, v.z); v.w = fmaf(cv[u], w2r[c].w, v.w);
          *reinterpret_cast<float4*>(er + (c * 32 + lane) * 4) = v;
        }
      }
    }
  }
#pragma unroll
  for (int c = 0; c < NCH; ++c) *reinterpret_cast<float4*>(&s_g2[wic][(c * 32 + lane) * 4]) = g2[c];
  __syncthreads();
  for (int col = threadIdx.x; col < D; col += kOwnThreads) {
    float g = 0.f;
#pragma unroll
    for (int w = 0; w < kOwnThreads / 32; ++w) g += s_g2[w][col];
    if (g != 0.f) red_add_f32_sys(pend2_shards[col / w2_rows_per_shard] + col % w2_rows_per_shard, g);
  }
}

static void fill_params(OwnerParams& p, const int32_t* row_ptr, const int64_t* fid, const float* fval, const float* labels,
                        int32_t B, int32_t world, int32_t rank, int32_t cap, int64_t rows_per_shard, int64_t features,
                        void* const* inbox, uint32_t* const* inbox_cnt, void* const* resp, uint32_t* count,
                        int32_t* slot_tab) {
  p.row_ptr = row_ptr; p.fid = fid; p.fval = fval; p.labels = labels;
  p.B = B; p.world = world; p.rank = rank; p.cap = cap; p.rows_per_shard = rows_per_shard; p.features = features;
  p.shift = -1;
  if (rows_per_shard > 0 && (rows_per_shard & (rows_per_shard - 1)) == 0) {
    int s = 0;
    while (((int64_t)1 << s) < rows_per_shard) ++s;
    p.shift = s;
  }
  p.inbox = reinterpret_cast<uint2* const*>(inbox);
  p.inbox_cnt = reinterpret_cast<unsigned int* const*>(inbox_cnt);
  p.resp = reinterpret_cast<float2* const*>(resp);
  p.count = count;
  p.slot_tab = slot_tab;
  p.bad_ids = bad_id_counter();
}

}  // namespace b200ps

using namespace b200ps;

#define OWN_COMMON_CHECKS(name)                                                                                   \
  B200PS_REQUIRE(B > 0 && world >= 1 && world <= kMaxWorldOwn && rank >= 0 && rank < world && cap > 0 &&         \
                     rows_per_shard > 0 && features > 0 && inbox && inbox_cnt && resp && count,                   \
                 name ": bad arguments")

// phase 1 of the owner-side wide & deep step: post this rank's row requests (see the file header).
extern "C" int b200ps_wd_owner_request(const int32_t* row_ptr, const int64_t* fid, int32_t B, int32_t world, int32_t rank,
                                       int32_t cap, int64_t rows_per_shard, int64_t features, void* const* inbox,
                                       uint32_t* const* inbox_cnt, void* const* resp, uint32_t* count,
                                       int32_t* slot_tab, b200ps_stream_t stream) {
  OWN_COMMON_CHECKS("wd_owner_request");
  B200PS_REQUIRE(slot_tab, "wd_owner_request: null slot table");
  OwnerParams p;
  fill_params(p, row_ptr, fid, nullptr, nullptr, B, world, rank, cap, rows_per_shard, features, inbox, inbox_cnt, resp, count,
              slot_tab);
  const int ctas = std::min((cap + kOwnThreads - 1) / kOwnThreads, sm_count() * 2);
  wd_request_kernel<<<ctas, kOwnThreads, 0, as_stream(stream)>>>(p);
  B200PS_LAUNCH_CHECK();
  return 0;
}

// phase 2: serve the requests that landed in this rank's inbox from its LOCAL shards of E [rows, dim] and W [rows].
extern "C" int b200ps_wd_owner_serve(int32_t world, int32_t rank, int32_t cap, void* const* inbox,
                                     uint32_t* const* inbox_cnt, void* const* resp, const float* emb_local,
                                     const float* wide_local, int32_t dim, float* const* w2_shards,
                                     int64_t w2_rows_per_shard, b200ps_stream_t stream) {
  B200PS_REQUIRE(world >= 1 && world <= kMaxWorldOwn && rank >= 0 && rank < world && cap > 0 && inbox && inbox_cnt && resp &&
                     emb_local && w2_shards && w2_rows_per_shard > 0 && w2_rows_per_shard % 4 == 0,
                 "wd_owner_serve: bad arguments");
  B200PS_REQUIRE(dim == 128 || dim == 256 || dim == 512, "wd_owner_serve: dim must be 128, 256 or 512");
  OwnerParams p{};
  p.world = world; p.rank = rank; p.cap = cap;
  p.inbox = reinterpret_cast<uint2* const*>(inbox);
  p.inbox_cnt = reinterpret_cast<unsigned int* const*>(inbox_cnt);
  p.resp = reinterpret_cast<float2* const*>(resp);
  const int ctas = sm_count() * 2;
  cudaStream_t s = as_stream(stream);
  if (dim == 128) wd_serve_kernel<1><<<ctas, kOwnThreads, 0, s>>>(p, emb_local, wide_local, w2_shards, w2_rows_per_shard);
  else if (dim == 256) wd_serve_kernel<2><<<ctas, kOwnThreads, 0, s>>>(p, emb_local, wide_local, w2_shards, w2_rows_per_shard);
  else wd_serve_kernel<4><<<ctas, kOwnThreads, 0, s>>>(p, emb_local, wide_local, w2_shards, w2_rows_per_shard);
  B200PS_LAUNCH_CHECK();
  return 0;
}

// phase 3: loss, dz and (do_update) the scalar coefficients into the owners' pending tables, bias, global_step.
extern "C" int b200ps_wd_owner_loss(const int32_t* row_ptr, const int64_t* fid, const float* fval, const float* labels,
                                    int32_t B, int32_t world, int32_t rank, int32_t cap, int64_t rows_per_shard,
                                    int64_t features, void* const* resp, const int32_t* slot_tab, float* const* pend_shards, float* bias,
                                    float* bias_pend, float* pred, float* dz_out, float* loss_sum, uint64_t* global_step,
                                    int32_t do_update, float lr, float grad_scale, float loss_scale, b200ps_stream_t stream) {
  B200PS_REQUIRE(B > 0 && world >= 1 && world <= kMaxWorldOwn && rank >= 0 && rank < world && rows_per_shard > 0 && features > 0 &&
                     resp && bias && (!do_update || pend_shards),
                 "wd_owner_loss: bad arguments");
  OwnerParams p;
  uint32_t dummy = 0;
  B200PS_REQUIRE(slot_tab, "wd_owner_loss: null slot table");
  fill_params(p, row_ptr, fid, fval, labels, B, world, rank, cap, rows_per_shard, features, resp /* unused */, nullptr, resp, &dummy,
              const_cast<int32_t*>(slot_tab));
  p.count = nullptr; p.inbox = nullptr; p.inbox_cnt = nullptr;
  OwnerLoss q;
  q.pend = pend_shards; q.pend_rows_per_shard = rows_per_shard; q.pend_shift = p.shift;
  q.bias = bias; q.bias_pend = bias_pend; q.pred = pred; q.dz_out = dz_out; q.loss_sum = loss_sum;
  q.global_step = reinterpret_cast<unsigned long long*>(global_step);
  q.lr = lr; q.grad_scale = grad_scale; q.loss_scale = loss_scale; q.do_update = do_update;
  const int wpc = kOwnThreads / 32;
  const int ctas = std::min((B + wpc - 1) / wpc, sm_count() * 4);
  wd_loss_kernel<<<ctas, kOwnThreads, 0, as_stream(stream)>>>(p, q);
  B200PS_LAUNCH_CHECK();
  return 0;
}

// phase 4: fold the round's coefficients into this rank's rows, driven by its inbox (the rows requested this round);
// adds dw2 into the pending w2 vector (pend2_shards, sharded like w2).  After a barrier, b200ps_sparse_wd_apply
// phase 2 folds the pending w2 vector.
extern "C" int b200ps_wd_owner_fold(int32_t world, int32_t rank, int32_t cap, void* const* inbox, uint32_t* const* inbox_cnt,
                                    float* pend_local, float* emb_local, float* wide_local, int32_t dim,
                                    float* const* w2_shards, float* const* pend2_shards, int64_t w2_rows_per_shard,
                                    b200ps_stream_t stream) {
  B200PS_REQUIRE(world >= 1 && world <= kMaxWorldOwn && rank >= 0 && rank < world && cap > 0 && inbox && inbox_cnt &&
                     pend_local && emb_local && w2_shards && pend2_shards && w2_rows_per_shard > 0 && w2_rows_per_shard % 4 == 0,
                 "wd_owner_fold: bad arguments");
  B200PS_REQUIRE(dim == 128 || dim == 256 || dim == 512, "wd_owner_fold: dim must be 128, 256 or 512");
  OwnerParams p{};
  p.world = world; p.rank = rank; p.cap = cap;
  p.inbox = reinterpret_cast<uint2* const*>(inbox);
  p.inbox_cnt = reinterpret_cast<unsigned int* const*>(inbox_cnt);
  const int ctas = sm_count() * 2;
  cudaStream_t s = as_stream(stream);
  if (dim == 128) wd_fold_kernel<1><<<ctas, kOwnThreads, 0, s>>>(p, pend_local, emb_local, wide_local, w2_shards, pend2_shards, w2_rows_per_shard);
  else if (dim == 256) wd_fold_kernel<2><<<ctas, kOwnThreads, 0, s>>>(p, pend_local, emb_local, wide_local, w2_shards, pend2_shards, w2_rows_per_shard);
  else wd_fold_kernel<4><<<ctas, kOwnThreads, 0, s>>>(p, pend_local, emb_local, wide_local, w2_shards, pend2_shards, w2_rows_per_shard);
  B200PS_LAUNCH_CHECK();
  return 0;
}
