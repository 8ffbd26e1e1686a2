// b200ps K13 — the whole training step of the narrow sigmoid MLP (example.py:95-118) as ONE launch.
//
// The 784-100-10 step is 48 MFLOP: on a B200 it is bounded by dependent-launch and pipeline
// set-up latency, not by arithmetic.  As a chain of four kernels (tcgen05 GEMM, head, tcgen05 GEMM,
// push) it costs ~25 us, of which < 2 us is math.  This kernel runs the step as three phases of one
// co-resident grid separated by two device-wide barriers:
//
//   A  Z2 partials      CTA (row tile, K slice): X[16 x ks] . W1[ks x H]  -> slab[slice][B][H]
//   --------------------------------------------------------------- grid barrier 1
//   B  head             warp per example: sum slabs + b1, sigmoid, 100x10 head, stable
//                       log-softmax CE, accuracy, dZ3, dZ2 -> dz2[B][H]; per-CTA partials of
//                       dW2 / db2 / db1 / loss / correct -> part[cta][P]
//   --------------------------------------------------------------- grid barrier 2
//   C  dW1 + apply      CTA (8 rows of W1): X^T[8 x B] . dZ2[B x H]; every CTA also merges a
//                       chunk of the head partials in fixed CTA order; then either
//                         apply = 1 (one replica): W -= lr * dW in place, global_step += 1, or
//                         apply = 0: gradients written for the push kernel that follows, or
//                         apply = 2 (synchronous replicas, one rank per GPU): the low-latency push runs
//                         inside this launch — see "in-kernel push" below.
//
// In-kernel push (apply = 2).  Phase C does not store its gradients: every value leaves as an 8-byte
// {value, epoch} packet straight into the landing buffer of the rank that owns it (peer memory over
// NVLink; its own slot for the part it owns itself).  The packets are the synchronisation, so no third grid
// barrier and no flag exchange is needed:
//   D  owner            every CTA takes a stripe of this rank's owned range, polls the packets of all
//                       replicas in rank order, applies the mean-gradient SGD step to the fp32 master and
//                       publishes the fresh value as a packet to every other rank
//   E  replica          polls the other owners' packets into the local working copy
// i.e. forward, backward, reduce-scatter, optimiser and all-gather of one synchronous step are ONE kernel,
// and a step costs two one-way NVLink latencies on top of the single-GPU step (protocol: ll_proto.cuh,
// same landing buffers and epochs as b200ps_push_sync_ll, so the two can be mixed).
//
// In-kernel Hogwild (apply = 3, the reference's default asynchronous mode, example.py:108,118).  There is no
// local copy of the parameters at all: phase A and the head fetch W1 / W2 / b1 / b2 straight from the fp32
// masters of the ranks that own them (bulk copies and cp.async over peer-mapped pointers — the pull fused into
// the first GEMM), and phase C adds -lr * gradient into those masters with fire-and-forget
// red.global.add.v4.f32 over NVLink, unlocked; one atomicAdd_system on the chief's counter keeps global_step
// counting every worker's update.  No barrier between ranks, one launch per worker step.
//
// fp32 FMA on the CUDA cores, exact fp32 accumulation in a fixed order: the step is bit-reproducible
// and matches the reference's fp32 CPU arithmetic more closely than a tf32 tensor-core product.  A
// tensor-core tile (tcgen05, M=128) would spend longer allocating TMEM and fetching its tensor maps
// than these 100-row products take; the tcgen05 GEMM remains the path for every shape where math
// matters (b200ps_gemm) and for this model when batch or width outgrow the limits below.
//
// Every operand a phase needs that does not depend on the previous phase is fetched with cp.async at
// kernel entry (three commit groups), so each phase pays at most one L2 round trip after its barrier.
//
// Limits: B <= 128, H <= 128 (multiple of 4), C <= 16, D multiple of 8.
#include "../common.h"
#include "../workspace.h"
#include "ptx.cuh"
#include "ll_proto.cuh"

namespace b200ps {

constexpr int kFsThreads = 256;
constexpr int kFsWarps = kFsThreads / 32;
constexpr int kFsRowTile = 16;     // phase A: batch rows per CTA (one per thread row group)
constexpr int kFsDRows = 8;        // phase C: rows of W1 per CTA
constexpr int kFsCP = 16;          // padded class count
constexpr int kFsCPP = 20;         // padded W2 row in shared memory
constexpr int kFsMaxPer = 8;       // dW2 outputs per thread: H*C <= 2048

struct FusedParams {
  const float* x;            // [rows, D]; the batch is rows [*cursor, *cursor + B)
  const int* labels;         // [rows]
  const int* cursor;         // device scalar row offset, or null (= 0)
  float* w1; float* b1; float* w2; float* b2;       // parameters (updated in place when apply)
  float* gw1; float* gb1; float* gw2; float* gb2;   // gradients (written when !apply)
  float* slabs;              // [KS][B][H]
  float* dz2;                // [B][H]
  float* part;               // [NT][P]
  unsigned int* bar;         // two barrier counters
  float* loss; int* correct; // step statistics (overwritten)
  unsigned long long* global_step;
  int* cursor_rw; int cursor_step, cursor_wrap;
  int B, D, H, C;
  int RT, KS, ks, xpitch;    // phase A: row tiles, K slices, slice width, smem pitch of the X tile
  int NT, NC, cpitch;        // phase B / C active CTAs, smem pitch of the X^T tile
  int apply;
  float lr, grad_scale, loss_scale;
  long long* timing;         // debug: SM-clock stamps of CTA 0 at the phase boundaries (or null)
  // in-kernel low-latency push (apply == 2)
  uint32_t* const* pads; float* const* masters; const long long* bounds;
  uint2* const* ll_grads; uint2* const* ll_params; unsigned long long* const* global_steps;
  long long max_shard, off_w1, off_b1, off_w2, off_b2;
  int world, rank;
  int skew_ns;               // debug: random per-CTA delay (< skew_ns) at entry and before the push phases
};

__device__ __forceinline__ void fs_skew(const FusedParams& p, uint32_t salt) {
  if (p.skew_ns > 0) {       // stress knob (b200ps_debug_set_skew_ns): shakes the arrival order of CTAs and ranks
    if (threadIdx.x == 0) {
      const uint32_t h = (salt * 2654435761u) ^ ((uint32_t)p.rank * 40503u) ^ (blockIdx.x * 9973u) ^ (uint32_t)clock64();
      __nanosleep(h % (uint32_t)p.skew_ns);
    }
    __syncthreads();
  }
}

__device__ __forceinline__ float fs_sigmoid(float x) { return 1.f / (1.f + __expf(-x)); }

__device__ __forceinline__ void fs_grid_barrier(unsigned int* ctr, unsigned int expected) {
  __syncthreads();
  if (threadIdx.x == 0) {
    // release-arrive (orders this CTA's writes, made visible to thread 0 by the bar.sync above) + acquire-poll
    asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(ctr) : "memory");
    if (ld_acquire_gpu(ctr) < expected) {
      const uint64_t t0 = globaltimer_ns();
      uint32_t spins = 0;
      while (ld_acquire_gpu(ctr) < expected) {
        if ((++spins & 0xff) == 0 && globaltimer_ns() - t0 > B200PS_WAIT_TIMEOUT_NS) {
          printf("b200ps: fused step grid barrier timed out (block %d, want %u, have %u)\n", blockIdx.x, expected,
                 ld_acquire_gpu(ctr));
          __trap();
        }
      }
    }
  }
  __syncthreads();
}

// acc[i][:] += sum_k xs[rg + RG i][k] * ws[k][4 cg .. 4 cg + 3]   for k in [0, K), K a multiple of 4.
// xs rows are K-contiguous with a pitch whose quarter is odd, so the 8 row groups of a warp hit 8
// different 16-byte bank groups; the 4 column groups of a warp read 4 consecutive float4 of ws.
template <int RPT, int RG>
__device__ __forceinline__ void fs_tile_fma(const float* __restrict__ xs, int xpitch, const float* __restrict__ ws,
                                            int wpitch, int K, int rg, int cg, float (&acc)[RPT][4]) {
  const float* xr = xs + rg * xpitch;
  const float* wc = ws + cg * 4;
#pragma unroll 2
  for (int k = 0; k < K; k += 4) {
    float4 xv[RPT], wv[4];
#pragma unroll
    for (int i = 0; i < RPT; ++i) xv[i] = *reinterpret_cast<const float4*>(xr + i * RG * xpitch + k);
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) wv[kk] = *reinterpret_cast<const float4*>(wc + (k + kk) * wpitch);
#pragma unroll
    for (int i = 0; i < RPT; ++i) {
      const float xk[4] = {xv[i].x, xv[i].y, xv[i].z, xv[i].w};
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        acc[i][0] = fmaf(xk[kk], wv[kk].x, acc[i][0]);
        acc[i][1] = fmaf(xk[kk], wv[kk].y, acc[i][1]);
        acc[i][2] = fmaf(xk[kk], wv[kk].z, acc[i][2]);
        acc[i][3] = fmaf(xk[kk], wv[kk].w, acc[i][3]);
      }
    }
  }
}

// gradient element(s) of flat index i -> the owner's landing buffer as {value, epoch} packets
__device__ __forceinline__ int fs_owner(const long long* s_bounds, long long i) {
  int o = 0;
  while (i >= s_bounds[o + 1]) {
    ++o;
    B200PS_DEV_CHECK(o < kMaxWorld, "flat index %lld has no owner\n", i);
  }
  return o;
}
__device__ __forceinline__ void fs_send1(const FusedParams& p, const long long* s_bounds, long long i, float v, uint32_t epoch) {
  const int o = fs_owner(s_bounds, i);
  st_packet(p.ll_grads[o] + (long long)p.rank * p.max_shard + (i - s_bounds[o]), v, epoch);
}
__device__ __forceinline__ void fs_red1(const FusedParams& p, const long long* s_bounds, long long i, float g) {
  red_add_f32_sys(p.masters[fs_owner(s_bounds, i)] + i, -p.lr * g);
}
__device__ __forceinline__ void fs_send4(const FusedParams& p, const long long* s_bounds, long long i, const float (&v)[4], uint32_t epoch) {
  const int o = fs_owner(s_bounds, i);      // owner ranges are multiples of 8: a 4-aligned group has one owner
  uint2* dst = p.ll_grads[o] + (long long)p.rank * p.max_shard + (i - s_bounds[o]);
#pragma unroll
  for (int j = 0; j < 4; ++j) st_packet(dst + j, v[j], epoch);
}

#define FS_STAMP(i) do { if (p.timing && cta == 0 && tid == 0) p.timing[i] = clock64(); } while (0)

__global__ void __launch_bounds__(kFsThreads) mlp_fused_step_kernel(const FusedParams p) {
  extern __shared__ __align__(16) float sm[];
  const int B = p.B, D = p.D, H = p.H, C = p.C;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int cta = blockIdx.x, G = gridDim.x;
  const int H4 = (H + 3) & ~3, B4 = (B + 3) & ~3;
  // persistent regions
  float* s_w2 = sm;                               // [H][CPP]
  float* s_b2 = s_w2 + H * kFsCPP;                // [16]
  float* s_b1 = s_b2 + kFsCP;                     // [H4]
  int* s_lab = reinterpret_cast<int*>(s_b1 + H4); // [8]       labels of this CTA's examples
  float* s_w2raw = s_b1 + H4 + kFsWarps;          // [H*C]     W2 as it lies in memory (re-laid into s_w2)
  float* s_xc = s_w2raw + H * C;                  // [B4][8]   X columns of phase C (prefetched)
  float* s_w1c = s_xc + B4 * kFsDRows;            // [8][H]    W1 rows of phase C (apply only)
  float* s_u = s_w1c + kFsDRows * H;              // union of the per-phase buffers
  __shared__ float s_loss[kFsWarps];
  __shared__ int s_corr[kFsWarps];
  __shared__ __align__(8) uint64_t s_mbar[2];   // bulk-copy completion: [0] W1 slice (phase A), [1] dZ2 (phase C)
  if (tid == 0) {
    mbar_init(&s_mbar[0], 1);
    mbar_init(&s_mbar[1], 1);
    fence_mbar_init();
  }
  __syncthreads();

  FS_STAMP(0);
  pdl_wait();        // parameters and the cursor were written by the previous step
  pdl_trigger();
  FS_STAMP(1);
  const int row0 = p.cursor ? __ldcg(p.cursor) : 0;      // consumed late: the parameter fetches below go out first
  __shared__ long long s_bounds[kMaxWorld + 1];
  uint32_t epoch = 0;
  if (p.apply >= 2) {
    if (tid <= p.world) s_bounds[tid] = p.bounds[tid];
    if (p.apply == 2) epoch = __ldcg(p.pads[p.rank] + kSlotEpoch) + 1;
    __syncthreads();
  }
  fs_skew(p, epoch);
  if (cta == 0 && tid == 0) p.bar[1] = 0;         // nobody can reach barrier 2 before we reach barrier 1

  // ---------------------------------------------------------------- operand fetch: three cp.async groups
  // group 0: phase A tiles; group 1: head constants + labels; group 2: X^T tile and W1 rows of phase C.
  const bool in_a = cta < p.RT * p.KS;
  const int rt = cta % p.RT, sl_a = cta / p.RT;
  const int r0 = rt * kFsRowTile, k0 = sl_a * p.ks;
  const int kw = in_a ? min(p.ks, D - k0) : 0;    // multiple of 4
  float* s_xs = s_u;                              // [16][xpitch]
  float* s_ws = s_u + kFsRowTile * p.xpitch;      // [ks][H]
  if (in_a && tid == 0) {                         // W1[k0 : k0 + kw, :] is one contiguous run: a single bulk copy
    mbar_arrive_expect_tx(&s_mbar[0], (uint32_t)(kw * H) * 4u);
    if (p.apply == 3) {                           // ... from the masters of the owner(s) of that run of the flat space
      long long i = p.off_w1 + (long long)k0 * H;
      const long long end = i + (long long)kw * H;
      float* dst = s_ws;
      while (i < end) {
        const int o = fs_owner(s_bounds, i);
        const long long seg = min(end, s_bounds[o + 1]) - i;
        bulk_load_g2s(dst, p.masters[o] + i, (uint32_t)seg * 4u, &s_mbar[0]);
        dst += seg;
        i += seg;
      }
    } else {
      bulk_load_g2s(s_ws, p.w1 + (size_t)k0 * H, (uint32_t)(kw * H) * 4u, &s_mbar[0]);
    }
  }
  const float* xb = p.x + (size_t)row0 * D;
  if (in_a) {
    const int QK = kw >> 2;
#pragma unroll 1
    for (int q = tid & 15; q < QK; q += 16) {                      // thread -> (row = tid / 16, quads q, q + 16, ...)
      const int row = tid >> 4;
      float* dst = s_xs + row * p.xpitch + 4 * q;
      if (r0 + row < B) cp_async_16(dst, xb + (size_t)(r0 + row) * D + k0 + 4 * q);
      else *reinterpret_cast<float4*>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
  cp_async_commit();
  if (cta < p.NT && p.apply == 3) {               // head constants from their owners' masters (4-aligned groups)
#pragma unroll 1
    for (int i = tid; i < (H * C) >> 2; i += kFsThreads) {
      const long long g = p.off_w2 + 4 * i;
      cp_async_16(s_w2raw + 4 * i, p.masters[fs_owner(s_bounds, g)] + g);
    }
    if (tid < (H >> 2)) {
      const long long g = p.off_b1 + 4 * tid;
      cp_async_16(s_b1 + 4 * tid, p.masters[fs_owner(s_bounds, g)] + g);
    }
    if (tid < kFsCP) {
      const long long g = p.off_b2 + tid;
      if (tid < C) cp_async_4(s_b2 + tid, p.masters[fs_owner(s_bounds, g)] + g);
      else s_b2[tid] = 0.f;
    }
  } else if (cta < p.NT) {
#pragma unroll 1
    for (int i = tid; i < (H * C) >> 2; i += kFsThreads) cp_async_16(s_w2raw + 4 * i, p.w2 + 4 * i);
    if (tid < (H >> 2)) cp_async_16(s_b1 + 4 * tid, p.b1 + 4 * tid);
    if (tid < kFsCP) {
      if (tid < C) cp_async_4(s_b2 + tid, p.b2 + tid);
      else s_b2[tid] = 0.f;
    }
  }
  if (cta < p.NT) {
    if (tid < kFsWarps) {
      const int r = cta * kFsWarps + tid;
      if (r < B) cp_async_4(s_lab + tid, p.labels + row0 + r);
      else s_lab[tid] = 0;
    }
  }
  cp_async_commit();
  const int d0 = cta * kFsDRows;
  if (cta < p.NC) {
#pragma unroll 1
    for (int idx = tid; idx < 2 * B4; idx += kFsThreads) {           // 32-byte runs of the batch rows
      const int b = idx >> 1, q = idx & 1;
      float* dst = s_xc + b * kFsDRows + 4 * q;
      if (b < B) cp_async_16(dst, xb + (size_t)b * D + d0 + 4 * q);
      else *reinterpret_cast<float4*>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    if (p.apply == 1) {
      const int nrow = min(kFsDRows, D - d0);
      #pragma unroll 1
      for (int i = tid; i < (nrow * H) >> 2; i += kFsThreads) cp_async_16(s_w1c + 4 * i, p.w1 + (size_t)d0 * H + 4 * i);
    }
  }
  cp_async_commit();
  FS_STAMP(19);

  // ---------------------------------------------------------------- phase A: Z2 slabs
  cp_async_wait<2>();
  if (in_a) mbar_wait(&s_mbar[0], 0);
  __syncthreads();
  FS_STAMP(2);
  {
    const int rg = tid & 7, cg = tid >> 3;        // 8 row groups (rows rg, rg + 8) x 32 column groups
    if (in_a && cg * 4 < H) {
      float acc[2][4];
#pragma unroll
      for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
      fs_tile_fma<2, 8>(s_xs, p.xpitch, s_ws, H, kw, rg, cg, acc);
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int row = r0 + rg + 8 * i;
        if (row < B)
          *reinterpret_cast<float4*>(p.slabs + ((size_t)sl_a * B + row) * H + cg * 4) =
              make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
      }
    }
  }
  FS_STAMP(3);
  fs_grid_barrier(p.bar, G);
  FS_STAMP(4);

  // ---------------------------------------------------------------- phase B: head, one warp per example
  const int P = H * C + C + H + 2;
  cp_async_wait<1>();
  __syncthreads();
  if (cta < p.NT) {
    for (int i = tid; i < H * kFsCP; i += kFsThreads) {           // W2 -> padded rows of 16 classes (+4)
      const int h = i >> 4, c = i & (kFsCP - 1);
      s_w2[h * kFsCPP + c] = c < C ? s_w2raw[h * C + c] : 0.f;
    }
    __syncthreads();
    float* s_a2 = s_u;                            // [8][H]
    float* s_dz2 = s_a2 + kFsWarps * H;           // [8][H]
    float* s_dz3 = s_dz2 + kFsWarps * H;          // [8][16]
    const int r = cta * kFsWarps + warp;
    const bool row_ok = r < B;
    const int HK = (H + 31) >> 5;
    FS_STAMP(9);
    const int label = s_lab[warp];
    float zk[4], ak[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) zk[k] = 0.f;
    for (int sl0 = 0; sl0 < p.KS; sl0 += 16) {    // up to 16 slabs x 4 units in flight; summed in slice order
      float part[16][4];
#pragma unroll
      for (int u = 0; u < 16; ++u)
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int h = lane + 32 * k;
          part[u][k] = (row_ok && sl0 + u < p.KS && k < HK && h < H)
                           ? __ldcg(p.slabs + ((size_t)(sl0 + u) * B + r) * H + h) : 0.f;
        }
#pragma unroll
      for (int u = 0; u < 16; ++u)
#pragma unroll
        for (int k = 0; k < 4; ++k) zk[k] += part[u][k];
    }
    FS_STAMP(10);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int h = lane + 32 * k;
      const bool ok = row_ok && k < HK && h < H;
      ak[k] = ok ? fs_sigmoid(zk[k] + s_b1[h]) : 0.f;
      if (k < HK && h < H) s_a2[warp * H + h] = ak[k];
    }
    float zc[kFsCP];
#pragma unroll
    for (int c = 0; c < kFsCP; ++c) zc[c] = 0.f;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int h = lane + 32 * k;
      if (k < HK && h < H) {
        const float4* w = reinterpret_cast<const float4*>(s_w2 + h * kFsCPP);
#pragma unroll
        for (int c4 = 0; c4 < kFsCP / 4; ++c4) {
          const float4 wv = w[c4];
          zc[4 * c4] = fmaf(ak[k], wv.x, zc[4 * c4]);
          zc[4 * c4 + 1] = fmaf(ak[k], wv.y, zc[4 * c4 + 1]);
          zc[4 * c4 + 2] = fmaf(ak[k], wv.z, zc[4 * c4 + 2]);
          zc[4 * c4 + 3] = fmaf(ak[k], wv.w, zc[4 * c4 + 3]);
        }
      }
    }
#pragma unroll
    for (int c = 0; c < kFsCP; ++c) zc[c] += __shfl_xor_sync(0xffffffffu, zc[c], 16);
    // recursive halving over the lane bits: lane l ends up owning class l & 15
#pragma unroll
    for (int bit = 8, n = kFsCP / 2; bit >= 1; bit >>= 1, n >>= 1) {
      const bool up = (lane & bit) != 0;
#pragma unroll
      for (int j = 0; j < n; ++j) {
        const float send = up ? zc[j] : zc[j + n];
        const float keep = up ? zc[j + n] : zc[j];
        zc[j] = keep + __shfl_xor_sync(0xffffffffu, send, bit);
      }
    }
    FS_STAMP(11);
    const int myc = lane & (kFsCP - 1);
    const float z = zc[0] + s_b2[myc];
    const bool cls_ok = myc < C;
    float m = cls_ok ? z : -INFINITY;
    int arg = myc;
#pragma unroll
    for (int o = kFsCP / 2; o > 0; o >>= 1) {
      const float om = __shfl_xor_sync(0xffffffffu, m, o);
      const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
      if (om > m || (om == m && oa < arg)) { m = om; arg = oa; }
    }
    const float e = cls_ok ? __expf(z - m) : 0.f;
    float se = e;
#pragma unroll
    for (int o = kFsCP / 2; o > 0; o >>= 1) se += __shfl_xor_sync(0xffffffffu, se, o);
    const float zl = __shfl_sync(0xffffffffu, z, label & (kFsCP - 1));
    const float pr = e / se;
    if (lane == 0) {
      s_loss[warp] = row_ok ? m + __logf(se) - zl : 0.f;
      s_corr[warp] = (row_ok && arg == label) ? 1 : 0;
    }
    const float mine = (row_ok && cls_ok) ? (pr - (myc == label ? 1.f : 0.f)) * p.grad_scale : 0.f;
    FS_STAMP(12);
    if (lane < kFsCP) s_dz3[warp * kFsCP + lane] = mine;
    __syncwarp();
    {
      const float4* dz = reinterpret_cast<const float4*>(s_dz3 + warp * kFsCP);
#pragma unroll
      for (int c4 = 0; c4 < kFsCP / 4; ++c4) {
        const float4 d = dz[c4];
        zc[4 * c4] = d.x; zc[4 * c4 + 1] = d.y; zc[4 * c4 + 2] = d.z; zc[4 * c4 + 3] = d.w;
      }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int h = lane + 32 * k;
      if (k < HK && h < H) {
        const float4* w = reinterpret_cast<const float4*>(s_w2 + h * kFsCPP);
        float d = 0.f;
#pragma unroll
        for (int c4 = 0; c4 < kFsCP / 4; ++c4) {
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
    FS_STAMP(13);
    __syncthreads();
    FS_STAMP(14);
    // this CTA's 8 examples -> partial dW2 | db2 | db1 | loss | correct (rows beyond B hold zeros)
    float* mine_p = p.part + (size_t)cta * P;
    {
      const int c = tid & (kFsCP - 1);            // thread -> class c, hidden units tid/16, tid/16 + 16, ...
      float dz[kFsWarps];
#pragma unroll
      for (int w8 = 0; w8 < kFsWarps; ++w8) dz[w8] = s_dz3[w8 * kFsCP + c];
      if (c < C) {
        for (int h = tid >> 4; h < H; h += kFsThreads / kFsCP) {
          float t = 0.f;
#pragma unroll
          for (int w8 = 0; w8 < kFsWarps; ++w8) t = fmaf(s_a2[w8 * H + h], dz[w8], t);
          mine_p[h * C + c] = t;
        }
      }
    }
    if (tid < C) {
      float t = 0.f;
#pragma unroll
      for (int w8 = 0; w8 < kFsWarps; ++w8) t += s_dz3[w8 * kFsCP + tid];
      mine_p[H * C + tid] = t;
    }
    if (tid < H) {
      float t = 0.f;
#pragma unroll
      for (int w8 = 0; w8 < kFsWarps; ++w8) t += s_dz2[w8 * H + tid];
      mine_p[H * C + C + tid] = t;
    }
    if (tid == 0) {
      float l = 0.f;
      int cr = 0;
#pragma unroll
      for (int w8 = 0; w8 < kFsWarps; ++w8) { l += s_loss[w8]; cr += s_corr[w8]; }
      mine_p[P - 2] = l;
      mine_p[P - 1] = (float)cr;
    }
  }
  fence_proxy_async_all();                        // dZ2 is read through the async proxy (bulk copy) in phase C
  FS_STAMP(5);
  fs_grid_barrier(p.bar + 1, G);
  FS_STAMP(6);
  if (cta < p.NC && tid == 0) {                   // dZ2 is one contiguous run: a single bulk copy, issued first
    fence_proxy_async_all();
    mbar_arrive_expect_tx(&s_mbar[1], (uint32_t)(B * H) * 4u);
    bulk_load_g2s(s_u, p.dz2, (uint32_t)(B * H) * 4u, &s_mbar[1]);
  }
  if (cta == 0 && tid == 0) p.bar[0] = 0;         // everybody is past barrier 1: re-arm it for the next launch

  // ---------------------------------------------------------------- phase C: dW1 (+ apply) and the merge
  // the merge loads go out first so that they fly under the dW1 product
  const int chunk = (P + G - 1) / G;
  const int o_mine = cta * chunk + tid;
  const bool merge_fast = chunk <= kFsThreads && p.NT <= 16;
  float mv[16];
  if (merge_fast) {
#pragma unroll
    for (int u = 0; u < 16; ++u)
      mv[u] = (tid < chunk && o_mine < P && u < p.NT) ? __ldcg(p.part + (size_t)u * P + o_mine) : 0.f;
  }
  if (cta < p.NC) {
    float* s_dz = s_u;                            // [B4][H], rows >= B zero
    for (int i = B * H + tid; i < B4 * H; i += kFsThreads) s_dz[i] = 0.f;
    FS_STAMP(17);
    cp_async_wait<0>();
    mbar_wait(&s_mbar[1], 0);
    FS_STAMP(18);
    __syncthreads();
    FS_STAMP(15);
    const int rg = tid & 7, cg = tid >> 3;
    if (cg * 4 < H) {
      float acc[1][4] = {{0.f, 0.f, 0.f, 0.f}};
      {
        const float* xc = s_xc + rg;              // x[b][d0 + rg], b-major with pitch 8
        const float* dz = s_dz + cg * 4;
#pragma unroll 2
        for (int b = 0; b < B4; b += 4) {
          float xk[4];
          float4 wv[4];
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) {
            xk[kk] = xc[(b + kk) * kFsDRows];
            wv[kk] = *reinterpret_cast<const float4*>(dz + (b + kk) * H);
          }
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) {
            acc[0][0] = fmaf(xk[kk], wv[kk].x, acc[0][0]);
            acc[0][1] = fmaf(xk[kk], wv[kk].y, acc[0][1]);
            acc[0][2] = fmaf(xk[kk], wv[kk].z, acc[0][2]);
            acc[0][3] = fmaf(xk[kk], wv[kk].w, acc[0][3]);
          }
        }
      }
      const int d = d0 + rg;
      if (d < D) {
        const size_t off = (size_t)d * H + cg * 4;
        if (p.apply == 1) {
          float4 w = *reinterpret_cast<const float4*>(s_w1c + rg * H + cg * 4);
          w.x = fmaf(-p.lr, acc[0][0], w.x); w.y = fmaf(-p.lr, acc[0][1], w.y);
          w.z = fmaf(-p.lr, acc[0][2], w.z); w.w = fmaf(-p.lr, acc[0][3], w.w);
          *reinterpret_cast<float4*>(p.w1 + off) = w;
        } else if (p.apply >= 2) {
          // staged: the CTA's 8 x H block is one contiguous run of the flat gradient and leaves below as
          // coalesced packets (32 lanes x 8 bytes per store instruction)
          *reinterpret_cast<float4*>(s_w1c + rg * H + cg * 4) = make_float4(acc[0][0], acc[0][1], acc[0][2], acc[0][3]);
        } else {
          *reinterpret_cast<float4*>(p.gw1 + off) = make_float4(acc[0][0], acc[0][1], acc[0][2], acc[0][3]);
        }
      }
    }
    if (p.apply >= 2) {
      __syncthreads();
      const int n = min(kFsDRows, D - d0) * H;
      const long long base = p.off_w1 + (long long)d0 * H;
      if (p.apply == 2) {
        for (int j = tid; j < n; j += kFsThreads) fs_send1(p, s_bounds, base + j, s_w1c[j], epoch);
      } else {                                    // Hogwild: -lr * dW1 added into the owners' masters, unlocked
        for (int j = tid; j < (n >> 2); j += kFsThreads) {
          const float4 g = *reinterpret_cast<const float4*>(s_w1c + 4 * j);
          const long long i = base + 4 * j;
          red_add_v4f32_sys(p.masters[fs_owner(s_bounds, i)] + i,
                            make_float4(-p.lr * g.x, -p.lr * g.y, -p.lr * g.z, -p.lr * g.w));
        }
      }
    }
  } else {
    cp_async_wait<0>();
  }
  FS_STAMP(7);
  // merge of the head partials: output o is summed over the NT partials in CTA order
  {
    const int o_end = min(P, (cta + 1) * chunk);
    for (int o = o_mine; o < o_end; o += kFsThreads) {
      float t = 0.f;
      if (merge_fast) {
#pragma unroll
        for (int u = 0; u < 16; ++u) t += mv[u];
      } else {
        for (int n = 0; n < p.NT; ++n) t += __ldcg(p.part + (size_t)n * P + o);
      }
      if (o < H * C) {
        if (p.apply == 1) p.w2[o] = fmaf(-p.lr, t, p.w2[o]);
        else if (p.apply == 2) fs_send1(p, s_bounds, p.off_w2 + o, t, epoch);
        else if (p.apply == 3) fs_red1(p, s_bounds, p.off_w2 + o, t);
        else p.gw2[o] = t;
      } else if (o < H * C + C) {
        const int c = o - H * C;
        if (p.apply == 1) p.b2[c] = fmaf(-p.lr, t, p.b2[c]);
        else if (p.apply == 2) fs_send1(p, s_bounds, p.off_b2 + c, t, epoch);
        else if (p.apply == 3) fs_red1(p, s_bounds, p.off_b2 + c, t);
        else p.gb2[c] = t;
      } else if (o < P - 2) {
        const int h = o - H * C - C;
        if (p.apply == 1) p.b1[h] = fmaf(-p.lr, t, p.b1[h]);
        else if (p.apply == 2) fs_send1(p, s_bounds, p.off_b1 + h, t, epoch);
        else if (p.apply == 3) fs_red1(p, s_bounds, p.off_b1 + h, t);
        else p.gb1[h] = t;
      } else if (o == P - 2) {
        if (p.loss) *p.loss = t * p.loss_scale;
      } else {
        if (p.correct) *p.correct = (int)(t + 0.5f);
      }
    }
  }
  FS_STAMP(8);
  if (p.apply == 1) {
    if (cta == 0 && tid == 0) {
      if (p.global_step) *p.global_step += 1ull;
      if (p.cursor_rw) *p.cursor_rw = (*p.cursor_rw + p.cursor_step) % p.cursor_wrap;
    }
    return;
  }
  if (p.apply == 3) {
    if (cta == 0 && tid == 0) {
      if (p.global_steps) atomicAdd_system(p.global_steps[0], 1ull);     // the chief's counter sees every worker's update
      if (p.cursor_rw) *p.cursor_rw = (*p.cursor_rw + p.cursor_step) % p.cursor_wrap;
    }
    return;
  }
  if (p.apply != 2) return;

  // ---------------------------------------------------------------- phases D / E: the in-kernel push
  fs_skew(p, epoch + 77u);
  {
    const int world = p.world, rank = p.rank;
    {   // alignment gaps of the flat layout belong to somebody's range too: they travel as zeros
      const long long total = s_bounds[world];
      const long long w1e = p.off_w1 + (long long)D * H, w2e = p.off_w2 + (long long)H * C;
      const long long b1e = p.off_b1 + H, b2e = p.off_b2 + C;
      const long long first = p.off_w1 == 0 ? w1e : 0;           // W1 is the (large) prefix in every layout we build
      for (long long i = first + (long long)cta * kFsThreads + tid; i < total; i += (long long)G * kFsThreads) {
        const bool covered = (i >= p.off_w1 && i < w1e) || (i >= p.off_w2 && i < w2e) || (i >= p.off_b1 && i < b1e) ||
                             (i >= p.off_b2 && i < b2e);
        if (!covered) fs_send1(p, s_bounds, i, 0.f, epoch);
      }
    }
    float* work = p.masters[rank];
    const long long lo = s_bounds[rank], hi = s_bounds[rank + 1], total = s_bounds[world];
    const long long gtid = (long long)cta * kFsThreads + tid, gstride = (long long)G * kFsThreads;
    const float step = p.lr / (float)world;
    const uint2* mine = p.ll_grads[rank];
    for (long long i = lo + gtid; i < hi; i += gstride) {          // D: reduce in rank order, apply, publish
      float sum = 0.f;
      if (world <= 8) {                                             // all replicas' packets in flight at once
        uint2 pk[8];
#pragma unroll
        for (int r = 0; r < 8; ++r)
          if (r < world) pk[r] = peek_packet(mine + (long long)r * p.max_shard + (i - lo));
#pragma unroll
        for (int r = 0; r < 8; ++r)
          if (r < world)
            sum += pk[r].y == epoch ? __uint_as_float(pk[r].x)
                                    : poll_packet(mine + (long long)r * p.max_shard + (i - lo), epoch);
      } else {
        for (int r = 0; r < world; ++r) sum += poll_packet(mine + (long long)r * p.max_shard + (i - lo), epoch);
      }
      const float w = work[i] - step * sum;
      work[i] = w;
      for (int q = 0; q < world; ++q)
        if (q != rank) st_packet(p.ll_params[q] + i, w, epoch);
    }
    const uint2* fresh = p.ll_params[rank];
    for (long long i0 = gtid; i0 < total; i0 += 4 * gstride) {     // E: the other owners' values, 4 polls in flight
      uint2 pk[4];
      bool need[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const long long i = i0 + u * gstride;
        need[u] = i < total && !(i >= lo && i < hi);
        if (need[u]) pk[u] = peek_packet(fresh + i);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const long long i = i0 + u * gstride;
        if (need[u]) work[i] = pk[u].y == epoch ? __uint_as_float(pk[u].x) : poll_packet(fresh + i, epoch);
      }
    }
    __syncthreads();
    if (tid == 0) {
      uint32_t* pad = p.pads[rank];
      __threadfence();
      if (atomicAdd(pad + kSlotCtaDone, 1u) == (unsigned)G - 1) {
        pad[kSlotCtaDone] = 0;
        pad[kSlotEpoch] = epoch;
        if (p.global_steps) *p.global_steps[rank] += 1ull;
        if (p.cursor_rw) *p.cursor_rw = (*p.cursor_rw + p.cursor_step) % p.cursor_wrap;
      }
    }
  }
}

}  // namespace b200ps

using namespace b200ps;

static long long* g_fs_timing = nullptr;
static int g_fs_skew_ns = 0;
// Debug hook: every following fused-step launch delays each CTA by a pseudo-random time below `ns` nanoseconds at
// entry and before its push phases (0 switches it off).  Used by the multi-GPU skew stress test.
extern "C" int b200ps_debug_set_skew_ns(int32_t ns) {
  g_fs_skew_ns = ns > 0 ? ns : 0;
  return 0;
}
// Debug hook: SM-clock stamps of CTA 0 at the 9 phase boundaries of every following fused-step launch
// are written to `buf` (device memory, 24 x int64); null switches it off.
extern "C" int b200ps_mlp_fused_step_set_timing(void* buf) {
  g_fs_timing = static_cast<long long*>(buf);
  return 0;
}

extern "C" int b200ps_mlp_fused_step(const float* x, const int32_t* labels, const int32_t* cursor, float* w1,
                                     float* b1, float* w2, float* b2, float* gw1, float* gb1, float* gw2,
                                     float* gb2, float* loss, int32_t* correct, uint64_t* global_step,
                                     int32_t* cursor_rw, int32_t cursor_step, int32_t cursor_wrap, int32_t B,
                                     int32_t D, int32_t H, int32_t C, int32_t apply, float lr, float grad_scale,
                                     float loss_scale, const b200ps_fused_push* push, b200ps_stream_t stream) {
  B200PS_REQUIRE(B > 0 && B <= 128 && H > 0 && H <= 128 && H % 4 == 0 && C > 0 && C <= kFsCP && D > 0 && D % 8 == 0,
                 "mlp_fused_step: needs B <= 128, H <= 128 (multiple of 4), C <= 16, D multiple of 8 "
                 "(got B=%d D=%d H=%d C=%d)", B, D, H, C);
  B200PS_REQUIRE(H * C <= kFsMaxPer * kFsThreads, "mlp_fused_step: H*C too large");
  B200PS_REQUIRE(apply >= 0 && apply <= 3, "mlp_fused_step: apply must be 0, 1, 2 or 3");
  B200PS_REQUIRE(apply != 3 || (push && push->masters && push->bounds && push->world >= 1 && push->world <= kMaxWorld &&
                                push->off_w1 % 4 == 0 && push->off_w2 % 4 == 0 && push->off_b1 % 4 == 0),
                 "mlp_fused_step: apply == 3 needs masters/bounds and 4-aligned variable offsets");
  B200PS_REQUIRE(apply != 0 || (gw1 && gb1 && gw2 && gb2), "mlp_fused_step: gradient outputs needed when apply == 0");
  B200PS_REQUIRE(apply != 2 || (push && push->pads && push->masters && push->bounds && push->ll_grads && push->ll_params &&
                                push->world >= 2 && push->world <= kMaxWorld && push->rank >= 0 && push->rank < push->world &&
                                push->off_w1 % 4 == 0),
                 "mlp_fused_step: apply == 2 needs the push tables, 2 <= world <= 32 and a 4-aligned W1 offset");
  B200PS_REQUIRE(((uintptr_t)x | (uintptr_t)w1 | (uintptr_t)w2 | (uintptr_t)b1 | (uintptr_t)gw1) % 16 == 0,
                 "mlp_fused_step: x, w1, w2, b1 and gw1 must be 16-byte aligned");
  FusedParams p;
  p.x = x; p.labels = labels; p.cursor = cursor; p.w1 = w1; p.b1 = b1; p.w2 = w2; p.b2 = b2;
  p.gw1 = gw1; p.gb1 = gb1; p.gw2 = gw2; p.gb2 = gb2; p.loss = loss; p.correct = correct;
  p.global_step = reinterpret_cast<unsigned long long*>(global_step);
  p.cursor_rw = cursor_rw; p.cursor_step = cursor_step; p.cursor_wrap = cursor_wrap > 0 ? cursor_wrap : 1;
  p.timing = g_fs_timing;
  p.skew_ns = g_fs_skew_ns;
  p.pads = nullptr; p.masters = nullptr; p.bounds = nullptr; p.ll_grads = nullptr; p.ll_params = nullptr;
  p.global_steps = nullptr; p.max_shard = 0; p.off_w1 = p.off_b1 = p.off_w2 = p.off_b2 = 0; p.world = 1; p.rank = 0;
  if (apply >= 2) {
    p.pads = push->pads; p.masters = push->masters; p.bounds = reinterpret_cast<const long long*>(push->bounds);
    p.ll_grads = reinterpret_cast<uint2* const*>(push->ll_grads);
    p.ll_params = reinterpret_cast<uint2* const*>(push->ll_params);
    p.global_steps = reinterpret_cast<unsigned long long* const*>(push->global_steps);
    p.max_shard = push->max_shard_elems; p.off_w1 = push->off_w1; p.off_b1 = push->off_b1; p.off_w2 = push->off_w2;
    p.off_b2 = push->off_b2; p.world = push->world; p.rank = push->rank;
  }
  p.B = B; p.D = D; p.H = H; p.C = C; p.apply = apply; p.lr = lr; p.grad_scale = grad_scale; p.loss_scale = loss_scale;
  p.NC = (D + kFsDRows - 1) / kFsDRows;
  p.NT = (B + kFsWarps - 1) / kFsWarps;
  p.RT = (B + kFsRowTile - 1) / kFsRowTile;
  // K slices: enough CTAs to spread the product over the SMs, at most 16 slabs (one batch of loads in the head)
  int ks_target = std::min(16, std::max(1, std::min(sm_count(), 112) / p.RT));
  p.ks = std::max(8, (((D + ks_target - 1) / ks_target) + 7) & ~7);
  p.KS = (D + p.ks - 1) / p.ks;
  p.xpitch = p.ks + 4;                               // ks % 8 == 0  ->  pitch / 4 odd
  const int B4 = (B + 3) & ~3;
  p.cpitch = ((B4 / 4) % 2 == 0) ? B4 + 4 : B4 + 8;  // pitch / 4 odd
  const int G = std::max(std::max(p.RT * p.KS, p.NC), p.NT);
  B200PS_REQUIRE(G <= sm_count(), "mlp_fused_step: needs %d co-resident CTAs, device has %d SMs", G, sm_count());
  const int P = H * C + C + H + 2;
  const size_t slab_f = (size_t)p.KS * B * H, dz_f = (size_t)B * H, part_f = (size_t)p.NT * P;
  Workspace ws;
  int rc = get_workspace(sizeof(float) * (slab_f + dz_f + part_f), &ws, as_stream(stream), true);
  if (rc) return rc;
  p.slabs = reinterpret_cast<float*>(ws.data);
  p.dz2 = p.slabs + slab_f;
  p.part = p.dz2 + dz_f;
  p.bar = ws.counters + 8;     // counters[8], [9]: the two grid barriers (self re-arming)
  const int H4 = (H + 3) & ~3;
  const size_t u_a = (size_t)kFsRowTile * p.xpitch + (size_t)p.ks * H, u_b = (size_t)kFsWarps * (2 * H + kFsCP),
               u_c = (size_t)B4 * H;
  const size_t smem = sizeof(float) * ((size_t)H * kFsCPP + kFsCP + H4 + kFsWarps + (size_t)H * C + (size_t)B4 * kFsDRows +
                                       (size_t)kFsDRows * H + std::max(std::max(u_a, u_b), u_c));
  static bool attr_set[64] = {false};
  int dev = 0;
  B200PS_CUDA_CHECK(cudaGetDevice(&dev));
  if (!attr_set[dev & 63]) {
    B200PS_CUDA_CHECK(cudaFuncSetAttribute(mlp_fused_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    attr_set[dev & 63] = true;
  }
  B200PS_REQUIRE(smem <= 200 * 1024, "mlp_fused_step: shared memory %zu too large", smem);
  // one CTA per SM: ask for more than half of an SM's shared memory so that no two CTAs share an SM's
  // load path and FMA pipes (B200PS_FUSED_SMEM_KB overrides for experiments)
  static const size_t min_smem = [] { const char* e = getenv("B200PS_FUSED_SMEM_KB"); return (size_t)(e ? atoi(e) : 116) * 1024; }();
  const size_t smem_launch = std::max(smem, min_smem);
  // The two grid barriers need all G CTAs on the machine at once.  G <= #SMs was checked above; here the launch
  // itself is checked: one CTA of this size must fit an SM (it does: 1 CTA/SM by construction), and the launch is
  // marked cooperative (measured cost: none, 13.64 vs 13.61 us/step; B200PS_FUSED_COOP=0 switches it off) so that
  // the driver schedules the grid only when every CTA can be resident, instead of letting a co-tenant kernel turn
  // the barrier into a 4-second trap.
  static int occ_ok[64] = {0};
  if (occ_ok[dev & 63] == 0) {
    int per_sm = 0;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, mlp_fused_step_kernel, kFsThreads, smem_launch);
    occ_ok[dev & 63] = (e == cudaSuccess && per_sm >= 1) ? 1 : -1;
    if (e != cudaSuccess) cudaGetLastError();
  }
  B200PS_REQUIRE(occ_ok[dev & 63] == 1, "mlp_fused_step: a CTA of %zu bytes of shared memory does not fit an SM of this device",
                 smem_launch);
  static const bool coop = [] { const char* e = getenv("B200PS_FUSED_COOP"); return !(e && e[0] == '0'); }();
  if (coop) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(G); cfg.blockDim = dim3(kFsThreads); cfg.dynamicSmemBytes = smem_launch; cfg.stream = as_stream(stream);
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeCooperative; at[0].val.cooperative = 1;
    at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl_enabled() ? 2 : 1;
    B200PS_CUDA_CHECK(cudaLaunchKernelEx(&cfg, mlp_fused_step_kernel, p));
  } else {
    B200PS_CUDA_CHECK(launch_pdl(mlp_fused_step_kernel, dim3(G), dim3(kFsThreads), smem_launch, as_stream(stream), p));
  }
  B200PS_LAUNCH_CHECK();
  workspace_mark_used(as_stream(stream));
  return 0;
}
