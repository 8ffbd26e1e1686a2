// b200ps K4u/K5u — sparse logistic regression with the batch's ids made UNIQUE first.
//
// The reference gathers with embedding_lookup_sparse over the batch's ids (lr2.py:390): TensorFlow makes the ids
// unique, fetches every distinct row once from the parameter server and sends ONE IndexedSlices entry per distinct
// row back (lr2.py:395-396).  The direct kernel (sparse_lr.cu) skips that: every non-zero loads its weight and
// posts its own red.add, which is the faster plan while ids hardly repeat (2^24 uniform ids: 0.1 % repeats).  Click
// logs are not like that — a few hot features sit in most rows — and then the direct plan reads the same remote
// word thousands of times and serialises thousands of remote atomics on one address.  This file is the other plan:
//
// one cooperative launch, up to two CTAs per SM, four phases separated by grid syncs:
//   0. unique   every non-zero inserts its id into an open-addressing hash table in this GPU's memory (64-bit keys
//               tagged with a 24-bit step epoch, so the table is never cleared; a slot is rewritten every few steps,
//               an entry surviving a whole epoch cycle is not a practical event); the winners of an insert take the
//               next slots of their CTA's OWN segment of the unique list (shared-memory counter: no global one).
//               Each non-zero remembers its table position.
//   1. gather   wbuf[u] = W[uniq[u]]  for the U distinct ids (the only remote reads), gbuf[u] = 0; the non-zeros'
//               table positions become indices into the unique list
//   2. examples warp per example: z from wbuf (local), loss, dz; dz*val added into one of 8 private copies of gbuf
//               (local atomics; a hot id's slot is a single address)
//   3. scatter  one red.add of -lr*sum(gbuf copies)[u] per distinct id (the only remote writes)
//
// Remote traffic per step: U reads + U reds instead of nnz + nnz.  With nothing repeated it costs three grid syncs
// and the hash pass more than the direct kernel, so it is a choice (SparseLR(dedupe=True)), not the default.
#include <cooperative_groups.h>

#include "../common.h"
#include "../workspace.h"
#include "ptx.cuh"

namespace cg = cooperative_groups;

namespace b200ps {

constexpr int kDdThreads = 256;
constexpr int kDdGradCopies = 8;                       // privatised gradient lists (a hot id's slot is one address)
constexpr int kDdCtasPerSm = 2;                        // 75 k threads: one round of the hash pass for ~60 k non-zeros
constexpr int kDdEpochShift = 40;                      // ids < 2^40; the upper 24 bits of a key carry the epoch
constexpr unsigned long long kDdIdMask = (1ull << kDdEpochShift) - 1ull;

// state[1] = epoch, state[3] = non-zeros of the last finished step, state[4..11] = CTA 0's globaltimer stamps
struct DedupScratch {
  unsigned long long* keys;   // [T]  (epoch << 40) | id ; anything with another epoch counts as empty
  int* slot_of;               // [T]  table position -> index in uniq
  int* hpos;                  // [cap] non-zero -> table position after the unique pass, -> index in uniq once the
                              //       step kernel has converted it (-1: refused id)
  long long* uniq;            // [grid*seg] distinct ids, CTA c's in [c*seg, c*seg + seg_cnt[c])
  float* wbuf;                // [grid*seg] gathered weights
  float* gbuf;                // [kDdGradCopies][gstride] accumulated dz*val: CTA c adds into copy c % kDdGradCopies —
                              //       the red.adds on a hot id's slot retire one after another (20 us of a 54 us step
                              //       with one copy and an id in half of the non-zeros), the scatter sums the copies
  long long gstride;          // list_cap
  unsigned* state;            // [16]
  unsigned* seg_cnt;          // [grid] distinct ids in each CTA's segment of uniq / wbuf / gbuf
  int seg;                    // segment length: rounds of the hash pass * 256 (a CTA cannot win more than it visits)
  unsigned tmask;             // T - 1
};

struct DedupParams {
  const int* row_ptr; const long long* fid; const float* fval; const float* labels;
  float* const* w_shards; float* bias; float* pred; float* dz_out; const float* dz_in; float* loss_sum;
  unsigned long long* global_step; unsigned long long* bad_ids;
  int B; long long rows_per_shard; long long features; int shard_shift;
  int do_forward, do_update, finish, convert;
  float lr, grad_scale, loss_scale;
  DedupScratch s;
};

__device__ __forceinline__ unsigned long long dd_mix(unsigned long long x) {   // splitmix64 finaliser
  x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ull;
  x ^= x >> 27; x *= 0x94d049bb133111ebull;
  return x ^ (x >> 31);
}

__device__ __forceinline__ float* dd_w_addr(const DedupParams& p, long long id) {
  int owner;
  long long local;
  if (p.shard_shift >= 0) { owner = (int)(id >> p.shard_shift); local = id & (p.rows_per_shard - 1); }
  else { owner = (int)(id / p.rows_per_shard); local = id - (long long)owner * p.rows_per_shard; }
  return p.w_shards[owner] + local;
}

// ---- unique: insert every id; the winners of an insert append it to their CTA's segment of the list ----------
// No global list counter: 148 CTAs bumping one word with returning atomics measured 3.4 us per round here (same-
// address atomics retire one after the other, ~23 ns each), so each CTA fills its own segment [c*seg, c*seg + n_c)
// from a shared-memory counter and leaves n_c in seg_cnt[c].
__device__ __forceinline__ int dd_unique_pass(const long long* __restrict__ fid, int nnz, long long features,
                                              const DedupScratch& s, unsigned long long* bad_ids) {
  __shared__ unsigned s_cnt;
  const unsigned long long epoch = (unsigned long long)s.state[1] << kDdEpochShift;
  const int lane = threadIdx.x & 31;
  const unsigned seg0 = blockIdx.x * (unsigned)s.seg;
  if (threadIdx.x == 0) s_cnt = 0u;
  __syncthreads();
  int bad = 0;
  for (int j0 = blockIdx.x * kDdThreads; j0 < nnz; j0 += gridDim.x * kDdThreads) {   // CTA-uniform trip count
    const int j = j0 + threadIdx.x;
    const long long id = j < nnz ? fid[j] : -1;
    const bool valid = (unsigned long long)id < (unsigned long long)features;
    if (j < nnz && !valid) ++bad;
    int where = -1;
    bool won = false;
    if (valid) {
      const unsigned long long mine = epoch | (unsigned long long)id;
      unsigned h = (unsigned)dd_mix((unsigned long long)id) & s.tmask;
      // the table has at least twice as many positions as non-zeros, so a probe sequence ends; tmask + 1 probes
      // without an end would mean a caller broke that, and the non-zero is then dropped and counted like a bad id
      for (unsigned probe = 0; probe <= s.tmask; ++probe, h = (h + 1) & s.tmask) {
        const unsigned long long cur = __ldcg(s.keys + h);
        if (cur == mine) { where = (int)h; break; }
        if ((cur & ~kDdIdMask) != epoch) {                             // stale or never used: try to take it
          const unsigned long long seen = atomicCAS(s.keys + h, cur, mine);
          if (seen == cur) { where = (int)h; won = true; break; }      // inserted: this id is new in the batch
          if (seen == mine) { where = (int)h; break; }                 // a duplicate of ours got there first
          if ((seen & ~kDdIdMask) != epoch) { --probe; h = (h - 1) & s.tmask; }   // lost to a stale rewrite: look again
        }
      }
      if (where < 0) ++bad;
    }
    const unsigned wm = __ballot_sync(0xffffffffu, won);
    if (wm) {                                          // one shared-memory atomic per warp
      const int leader = __ffs(wm) - 1;
      unsigned base = 0;
      if (lane == leader) base = atomicAdd(&s_cnt, (unsigned)__popc(wm));
      base = __shfl_sync(0xffffffffu, base, leader);
      if (won) {
        const unsigned slot = seg0 + base + (unsigned)__popc(wm & ((1u << lane) - 1u));
        s.uniq[slot] = id;
        s.slot_of[where] = (int)slot;
      }
    }
    if (j < nnz) s.hpos[j] = where;
  }
  if (bad && bad_ids) atomicAdd(bad_ids, (unsigned long long)bad);
  __syncthreads();
  if (threadIdx.x == 0) s.seg_cnt[blockIdx.x] = s_cnt;
  return (int)s_cnt;
}

// ---- the step over the unique list ------------------------------------------------------------------------------
__global__ void __launch_bounds__(kDdThreads) sparse_lr_unique_kernel(const DedupParams p) {
  cg::grid_group grid = cg::this_grid();
  __shared__ float s_loss[kDdThreads / 32], s_dz[kDdThreads / 32];
  const int lane = threadIdx.x & 31, wic = threadIdx.x >> 5;
  const int tid = blockIdx.x * kDdThreads + threadIdx.x, nthreads = gridDim.x * kDdThreads;
  const int nnz = p.row_ptr[p.B];
  auto stamp = [&](int i) { if (tid == 0) p.s.state[4 + i] = (unsigned)globaltimer_ns(); };
  stamp(0);
  int U;                                              // distinct ids in this CTA's segment
  if (p.convert) {                                    // phase 0: this batch's ids -> unique list
    U = dd_unique_pass(p.fid, nnz, p.features, p.s, p.bad_ids);
    stamp(1);
    grid.sync();
  } else {
    U = (int)p.s.seg_cnt[blockIdx.x];                 // the launch that ran the hash pass had this very grid
  }
  stamp(2);
  const int seg0 = blockIdx.x * p.s.seg;

  // phase 1: the distinct weights of this CTA's segment (the gather is latency: most rows are remote) ...
  for (int u = threadIdx.x; u < U; u += kDdThreads) {
    const float w = p.do_forward ? __ldcg(dd_w_addr(p, p.s.uniq[seg0 + u])) : 0.f;
    p.s.wbuf[seg0 + u] = w;
#pragma unroll
    for (int r = 0; r < kDdGradCopies; ++r) p.s.gbuf[r * p.s.gstride + seg0 + u] = 0.f;
  }
  float* const gmine = p.s.gbuf + (blockIdx.x % kDdGradCopies) * p.s.gstride;
  // ... and every non-zero's table position becomes its index in the unique list (once per batch: the launch that
  // ran the unique pass), so that phase 2 is one dependent load shorter
  if (p.convert) {
    for (int j = tid; j < nnz; j += nthreads) {
      const int h = p.s.hpos[j];
      if (h >= 0) p.s.hpos[j] = p.s.slot_of[h];
    }
  }
  const float b = p.do_forward ? __ldcg(p.bias) : 0.f;
  stamp(3);
  grid.sync();
  stamp(4);

  // phase 2: warp per example, everything local; 4 non-zeros per lane in flight, kept in registers for the update
  float loss_acc = 0.f, dz_acc = 0.f;
  for (int ex = blockIdx.x * (kDdThreads / 32) + wic; ex < p.B; ex += gridDim.x * (kDdThreads / 32)) {
    const int beg = p.row_ptr[ex], end = p.row_ptr[ex + 1];
    const bool one_batch = end - beg <= 128;
    float dz = p.do_forward ? 0.f : p.dz_in[ex];
    int sl[4];
    float fv[4];
    float z = 0.f;
    for (int j0 = beg; j0 < end; j0 += 128) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int j = j0 + lane + 32 * k;
        sl[k] = j < end ? __ldcg(p.s.hpos + j) : -1;
        fv[k] = j < end ? p.fval[j] : 0.f;
      }
      if (p.do_forward) {
        float w[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) w[k] = sl[k] >= 0 ? __ldcg(p.s.wbuf + sl[k]) : 0.f;
#pragma unroll
        for (int k = 0; k < 4; ++k) z = fmaf(w[k], fv[k], z);
      } else if (p.do_update) {
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (sl[k] >= 0) atomicAdd(gmine + sl[k], dz * fv[k]);
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
        if (one_batch) {
#pragma unroll
          for (int k = 0; k < 4; ++k)
            if (sl[k] >= 0) atomicAdd(gmine + sl[k], dz * fv[k]);
        } else {
          for (int j = beg + lane; j < end; j += 32) {
            const int u = __ldcg(p.s.hpos + j);
            if (u >= 0) atomicAdd(gmine + u, dz * p.fval[j]);
          }
        }
      }
    }
    if (p.do_update && lane == 0) dz_acc += dz;
  }
  if (lane == 0) { s_loss[wic] = loss_acc; s_dz[wic] = dz_acc; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float l = 0.f, d = 0.f;
    for (int w = 0; w < kDdThreads / 32; ++w) { l += s_loss[w]; d += s_dz[w]; }
    if (p.do_forward && p.loss_sum) atomicAdd(p.loss_sum, l * p.loss_scale);
    if (p.do_update) {
      red_add_f32_sys(p.bias, -p.lr * d);
      if (blockIdx.x == 0 && p.global_step) atomicAdd_system(p.global_step, 1ull);
    }
  }
  stamp(5);
  if (p.do_update) {
    grid.sync();
    stamp(6);
    // phase 3: one red per distinct id, every CTA its segment
    for (int u = threadIdx.x; u < U; u += kDdThreads) {
      float g = 0.f;
#pragma unroll
      for (int r = 0; r < kDdGradCopies; ++r) g += __ldcg(p.s.gbuf + r * p.s.gstride + seg0 + u);
      if (g != 0.f) red_add_f32_sys(dd_w_addr(p, p.s.uniq[seg0 + u]), -p.lr * g);
    }
  }
  // hand the table to the next batch: a new epoch.  Every thread read the epoch in phase 0, at least one grid sync
  // ago, and the next hash pass is a later launch on this stream.
  if (p.finish && tid == 0) {
    p.s.state[3] = (unsigned)nnz;
    p.s.state[11] = (unsigned)globaltimer_ns();
    const unsigned e = (p.s.state[1] + 1u) & 0xffffffu;
    p.s.state[1] = e ? e : 1u;                      // epoch 0 is what a zero-filled table holds
  }
}

}  // namespace b200ps

using namespace b200ps;

extern "C" int64_t b200ps_sparse_dedup_table_size(int64_t cap) {
  int64_t t = 1024;
  while (t < 4 * cap) t <<= 1;      // load <= 0.25: probe chains of a warp's 32 lanes stay at 1-2 positions
  return t;
}

// entries of uniq / wbuf (gbuf: b200ps_sparse_dedup_grad_copies() times as many) for `cap` non-zeros on this device
extern "C" int64_t b200ps_sparse_dedup_list_size(int64_t cap) { return cap + (int64_t)kDdCtasPerSm * sm_count() * kDdThreads; }

extern "C" int32_t b200ps_sparse_dedup_grad_copies(void) { return kDdGradCopies; }

extern "C" int b200ps_sparse_lr_unique(const int32_t* row_ptr, const int64_t* fid, const float* fval, const float* labels,
                                       float* const* w_shards, float* bias, float* pred, float* dz_out,
                                       const float* dz_in, float* loss_sum, uint64_t* global_step, int32_t B,
                                       int32_t cap, int64_t rows_per_shard, int64_t features, int32_t do_unique,
                                       int32_t do_forward, int32_t do_update, int32_t finish, float lr,
                                       float grad_scale, float loss_scale, uint64_t* keys, int32_t* slot_of,
                                       int64_t table_size, int32_t* hpos, int64_t* uniq, float* wbuf, float* gbuf,
                                       int64_t list_cap, uint32_t* seg_cnt, uint32_t* state, b200ps_stream_t stream) {
  B200PS_REQUIRE(B > 0 && cap > 0 && rows_per_shard > 0 && features > 0, "sparse_lr_unique: bad shape");
  B200PS_REQUIRE(features <= (int64_t)1 << kDdEpochShift, "sparse_lr_unique: ids must fit 40 bits");
  B200PS_REQUIRE(do_forward || (do_update && dz_in), "sparse_lr_unique: update-only needs dz_in");
  B200PS_REQUIRE(keys && slot_of && hpos && uniq && wbuf && gbuf && seg_cnt && state, "sparse_lr_unique: null scratch");
  B200PS_REQUIRE(table_size >= 2 * (int64_t)cap && (table_size & (table_size - 1)) == 0 && table_size <= ((int64_t)1 << 31),
                 "sparse_lr_unique: the table needs a power-of-two size of at least twice the non-zero capacity");
  cudaStream_t s = as_stream(stream);
  DedupParams p;
  p.row_ptr = row_ptr; p.fid = reinterpret_cast<const long long*>(fid); p.fval = fval; p.labels = labels;
  p.w_shards = w_shards; p.bias = bias; p.pred = pred; p.dz_out = dz_out; p.dz_in = dz_in; p.loss_sum = loss_sum;
  p.global_step = reinterpret_cast<unsigned long long*>(global_step);
  p.bad_ids = bad_id_counter();
  p.B = B; p.rows_per_shard = rows_per_shard; p.features = features;
  p.shard_shift = -1;
  if ((rows_per_shard & (rows_per_shard - 1)) == 0) {
    int sh = 0;
    while (((int64_t)1 << sh) < rows_per_shard) ++sh;
    p.shard_shift = sh;
  }
  p.do_forward = do_forward; p.do_update = do_update; p.finish = finish; p.convert = do_unique;
  p.lr = lr; p.grad_scale = grad_scale; p.loss_scale = loss_scale;
  p.s.keys = reinterpret_cast<unsigned long long*>(keys); p.s.slot_of = slot_of; p.s.hpos = hpos;
  p.s.uniq = reinterpret_cast<long long*>(uniq); p.s.wbuf = wbuf; p.s.gbuf = gbuf; p.s.state = state;
  p.s.tmask = (unsigned)(table_size - 1);
  p.s.seg_cnt = seg_cnt;
  p.s.gstride = list_cap;

  // the grid syncs need the whole grid resident: cooperative launch, sized below what the device can hold
  int dev = 0;
  B200PS_CUDA_CHECK(cudaGetDevice(&dev));
  static int per_sm[64] = {0};
  if (per_sm[dev & 63] == 0) {
    int n = 0;
    B200PS_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, sparse_lr_unique_kernel, kDdThreads, 0));
    per_sm[dev & 63] = n > 0 ? n : -1;
  }
  B200PS_REQUIRE(per_sm[dev & 63] > 0, "sparse_lr_unique: the kernel does not fit an SM");
  // the list passes (insert, gather, convert, scatter) are chains of dependent L2 round trips (~0.45 us each under
  // this load): as many threads as non-zeros, up to two CTAs per SM; the example pass needs B/8 CTAs
  const int want = std::max((B + kDdThreads / 32 - 1) / (kDdThreads / 32), (cap + kDdThreads - 1) / kDdThreads);
  const int ctas = std::max(1, std::min(want, std::min(kDdCtasPerSm, per_sm[dev & 63]) * sm_count()));
  p.s.seg = ((cap + ctas * kDdThreads - 1) / (ctas * kDdThreads)) * kDdThreads;
  B200PS_REQUIRE(list_cap >= (int64_t)ctas * p.s.seg,
                 "sparse_lr_unique: uniq / wbuf / gbuf hold %lld entries, %d CTAs x %d are needed (b200ps_sparse_dedup_list_size)",
                 (long long)list_cap, ctas, p.s.seg);
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(ctas); cfg.blockDim = dim3(kDdThreads); cfg.dynamicSmemBytes = 0; cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeCooperative; at[0].val.cooperative = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  B200PS_CUDA_CHECK(cudaLaunchKernelEx(&cfg, sparse_lr_unique_kernel, p));
  B200PS_LAUNCH_CHECK();
  return 0;
}
