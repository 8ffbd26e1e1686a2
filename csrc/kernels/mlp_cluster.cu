// b200ps K14 — the narrow sigmoid MLP (example.py:95-118) trained by ONE thread-block cluster that stays
// resident for many steps: parameters live in shared memory across steps, the matrix products run on the
// tensor cores, and every cross-CTA exchange goes through distributed shared memory.
//
// K13 (mlp_fused.cu) runs one step as a 98-CTA grid with two device-wide barriers and re-fetches W1 every
// launch; its step is bounded by those barriers and by "fetch what another SM just wrote" round trips through
// L2 (profiles/EXPERIMENTS.md).  Here the 16 CTAs of one cluster tile the step as 4 K slices x 4 hidden
// column groups, CTA (i, j):
//
//   owns        W1[K_i, H_j] (in shared memory for the whole launch, stored pre-split for the tensor cores),
//               and a replica of W2[H_j, :], b1[H_j], b2
//   forward     P_i[b, H_j] = X[b, K_i] . W1[K_i, H_j]   for all b      (mma.sync m16n8k8 tf32, 3xTF32 split)
//   exchange 1  reduce-scatter over the 4 K slices: rows of group i' go to CTA (i', j)         (DSMEM stores)
//   head a      a2 = sigmoid(sum_i P_i + b1) for rows R_i, columns H_j; z3 partial = a2 . W2[H_j, :]
//   exchange 2  z3 partials to the 4 CTAs of the row group
//   head b      softmax-CE, accuracy, dZ3 (redundantly in the 4 CTAs of a row group), dZ2[R_i, H_j],
//               partial dW2 / db2 / db1 / loss
//   exchange 3  dZ2 block + partials to the 4 CTAs that share H_j
//   backward    dW1[K_i, H_j] = X[:, K_i]^T . dZ2[:, H_j]   (same X tile, still in shared memory)
//   apply       W1 block updated in shared memory; W2 / b1 / b2 replicas updated identically everywhere
//
// Three barrier.cluster per step (~0.2 us each) replace the two grid barriers (~1.2 us each); W1 is never
// re-read; the only global traffic of a step is the X tile (cp.async.bulk, 1-D, one row slice per copy) and,
// on the last step of the launch, the parameter write-back.  All sums run in a fixed order: the result is
// bit-reproducible.  Precision: tf32 tensor-core products with the 3xTF32 error-compensated split
// (a = a_hi + a_lo, a.b ~ a_lo.b_hi + a_hi.b_lo + a_hi.b_hi) and fp32 accumulation — relative error ~2^-21
// per product, i.e. fp32-grade; B200PS_CLUSTER_SPLIT=0 selects plain tf32 (one mma per tile).
//
// Limits: B <= 128, 32 <= D <= 1024 (multiple of 8), H <= 128 (multiple of 4), C <= 16.
#include "../common.h"
#include "ptx.cuh"

namespace b200ps {

constexpr int kClThreads = 256;
constexpr int kClWarps = 8;
constexpr int kClCtas = 16;          // 4 K slices x 4 column groups
constexpr int kClHC = 32;            // padded hidden columns per group
constexpr int kClRG = 32;            // rows per row group
constexpr int kClCP = 16;            // padded classes
constexpr int kClA2P = 36;           // pitch of the a2 block

struct ClusterParams {
  const float* x; const int* labels; const int* cursor;
  float* w1; float* b1; float* w2; float* b2;
  float* loss; int* correct; int stat_stride;
  unsigned long long* global_step; int* cursor_rw; int cursor_step, cursor_wrap;
  int B, D, H, C, steps;
  int MT;                 // 16-row tiles of the batch
  int KSMAX, xp;          // widest K slice, pitch of the X tile (== 4 mod 8)
  float lr, grad_scale, loss_scale;
  long long* timing;
};

// ---- cluster / DSMEM / tensor-core PTX
__device__ __forceinline__ uint32_t cl_rank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cl_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t cl_map(uint32_t addr, uint32_t rank) {
  uint32_t r; asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank)); return r;
}
// One DMA copy of `bytes` (multiple of 16) from this CTA's shared memory into a peer CTA's, completing `bytes` of
// transaction count on the PEER's mbarrier: data and "it has landed" travel together, no cluster-wide barrier.
// (Per-lane st.shared::cluster stores cost ~12 cycles per 128-byte line touched — measured, tools/ubench/
// cluster_probe.cu — so everything that crosses CTAs is staged locally and shipped as a few KB-sized copies.)
__device__ __forceinline__ void cl_bulk_copy(uint32_t dst_cluster_addr, const void* src, uint32_t bytes, uint32_t mbar_cluster_addr) {
  asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst_cluster_addr), "r"(smem_u32(src)), "r"(bytes), "r"(mbar_cluster_addr) : "memory");
}
__device__ __forceinline__ void mma_tf32(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// x = hi + lo exactly: hi = x with the 13 low mantissa bits cleared (what the tensor core would read anyway),
// lo = x - hi (exact in fp32; the tensor core reads its top 10 mantissa bits) -> products good to ~2^-20
__device__ __forceinline__ void split_tf32(float x, uint32_t& hi, uint32_t& lo) {
  hi = __float_as_uint(x) & 0xffffe000u;
  lo = __float_as_uint(x - __uint_as_float(hi));
}
__device__ __forceinline__ float cl_sigmoid(float x) { return 1.f / (1.f + __expf(-x)); }

constexpr int kClRpRows = 37;        // rows of 16: dW2 (32) | db2 | loss,correct | db1 (2 rows)
constexpr int kClZ3Bytes = kClRG * kClCP * 4;                  // one z3-partial block
constexpr int kClDzBytes = 4 * 2 * 8 * 4 * 4 * 4;              // one row group's dZ2 block in fragment layout (4 k-steps)
constexpr int kClRpBytes = kClRpRows * kClCP * 4;              // one partial-gradient block
constexpr int kClZpBytes = kClRG * kClHC * 4;                  // one row group's partial-Z2 chunk

#define CL_STAMP(k) do { if (p.timing && crank == 0 && tid == 0 && s == stamp_step) p.timing[k] = clock64(); } while (0)

template <bool kSplit>
__global__ void __launch_bounds__(kClThreads, 1) mlp_cluster_kernel(const ClusterParams p) {
  extern __shared__ __align__(128) float sm[];
  __shared__ __align__(8) uint64_t s_xbar, s_mb1, s_mb2, s_mb3;
  __shared__ int s_lab[kClRG];
  __shared__ float s_b1[kClHC], s_b2[kClCP];
  __shared__ float s_rowloss[kClRG];
  __shared__ int s_rowcorr[kClRG];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
  const uint32_t crank = cl_rank();
  const int i = crank >> 2, j = crank & 3;
  const int B = p.B, D = p.D, H = p.H, C = p.C, MT = p.MT, xp = p.xp;
  const int stamp_step = p.steps >= 2 ? p.steps - 2 : 0;
  // K slices in units of 8 columns, column groups in units of 8 hidden units
  const int kbt = D >> 3, kbq = kbt >> 2, kbr = kbt & 3;
  const int kb_i = kbq + (i < kbr ? 1 : 0), kb0 = i * kbq + min(i, kbr);
  const int k0 = kb0 * 8, ks = kb_i * 8;
  const int ntt = (H + 7) >> 3, ntq = ntt >> 2, ntr = ntt & 3;
  const int nt_j = ntq + (j < ntr ? 1 : 0), nt0 = j * ntq + min(j, ntr);
  const int h0 = nt0 * 8, hcv = max(0, min(nt_j * 8, H - h0));
  const int KB = p.KSMAX >> 3;

  // ---- shared memory carve-up (floats; every exchanged block is 16-byte aligned)
  float* Xs = sm;                                        // [MT*16][xp]
  float* Wf = Xs + (size_t)MT * 16 * xp;                 // [KB][32][4][4]   {hi(k=t), lo(k=t), hi(k=t+4), lo(k=t+4)}
  float* Zp = Wf + (size_t)KB * kClHC * 16;              // [4 src][32][32]  partial Z2 of my rows / columns (swizzled)
  float* dZf = Zp + 4 * kClRG * kClHC;                   // [16 ksteps][2 mtiles][8 g][4 t][4]  dZ2 of ALL rows, my columns;
                                                         //   block i' doubles as the staging chunk of the partial Z2 sent to (i', j)
  float* Z3p = dZf + 16 * 2 * 8 * 4 * 4;                 // [2 buf][4 src][32][16]
  float* Z3st = Z3p + 2 * 4 * kClRG * kClCP;             // [2 buf][32][16]   my z3 partial (copy source)
  float* Rp = Z3st + 2 * kClRG * kClCP;                  // [4 src][37][16]
  float* A2s = Rp + 4 * kClRpRows * kClCP;               // [32][36]
  float* dZ3s = A2s + kClRG * kClA2P;                    // [32][16]
  float* W2s = dZ3s + kClRG * kClCP;                     // [32][16]
  float* dZl = W2s + kClHC * kClCP;                      // [32][33]  my rows' dZ2 (for db1)

  // ---------------------------------------------------------------- prologue (once per launch)
  const uint32_t n_col_peers = 3;
  if (tid == 0) {
    mbar_init(&s_xbar, 1); mbar_init(&s_mb1, 1); mbar_init(&s_mb2, 1); mbar_init(&s_mb3, 1);
    fence_mbar_init();
    mbar_arrive_expect_tx(&s_mb1, n_col_peers * kClZpBytes);               // arm step 0
    mbar_arrive_expect_tx(&s_mb2, 4 * kClZ3Bytes);
    mbar_arrive_expect_tx(&s_mb3, n_col_peers * (kClDzBytes + kClRpBytes));
  }
  for (int idx = B * xp + tid; idx < MT * 16 * xp; idx += kClThreads) Xs[idx] = 0.f;     // rows beyond the batch
  for (int idx = tid; idx < 16 * 2 * 8 * 4 * 4; idx += kClThreads) dZf[idx] = 0.f;
  for (int idx = tid; idx < 4 * kClRpRows * kClCP; idx += kClThreads) Rp[idx] = 0.f;
  pdl_wait();                                            // parameters / cursor written by the previous launch
  pdl_trigger();
  const int cursor0 = p.cursor ? __ldcg(p.cursor) : 0;
  __syncthreads();
  auto load_x = [&](int s) {                             // warp 0: one bulk copy per batch row
    const int row0 = (int)(((long long)cursor0 + (long long)s * p.cursor_step) % p.cursor_wrap);
    if (lane == 0) mbar_arrive_expect_tx(&s_xbar, (uint32_t)(B * ks) * 4u);
    __syncwarp();
    for (int b = lane; b < B; b += 32)
      bulk_load_g2s(Xs + (size_t)b * xp, p.x + (size_t)(row0 + b) * D + k0, (uint32_t)ks * 4u, &s_xbar);
  };
  if (warp == 0) load_x(0);
  // W1 block -> tensor-core fragment layout, split hi / lo (their sum is the exact fp32 master)
  for (int idx = tid; idx < KB * 8 * kClHC; idx += kClThreads) {
    const int kk = idx / kClHC, n = idx % kClHC;
    float w = 0.f;
    if (kk < ks && n < hcv) w = __ldcg(p.w1 + (size_t)(k0 + kk) * H + h0 + n);
    uint32_t hi, lo;
    split_tf32(w, hi, lo);
    const int kb = kk >> 3, r8 = kk & 7;
    float* dst = Wf + ((size_t)(kb * kClHC + n) * 4 + (r8 & 3)) * 4 + (r8 >> 2) * 2;
    dst[0] = __uint_as_float(hi);
    dst[1] = __uint_as_float(lo);
  }
  for (int idx = tid; idx < kClHC * kClCP; idx += kClThreads) {
    const int n = idx / kClCP, c = idx % kClCP;
    W2s[idx] = (n < hcv && c < C) ? __ldcg(p.w2 + (size_t)(h0 + n) * C + c) : 0.f;
  }
  if (tid < kClHC) s_b1[tid] = tid < hcv ? __ldcg(p.b1 + h0 + tid) : 0.f;
  if (tid < kClCP) s_b2[tid] = tid < C ? __ldcg(p.b2 + tid) : 0.f;
  __syncthreads();
  cl_sync();                                             // every CTA's barriers and buffers exist before any remote copy

  const uint32_t zp_a = smem_u32(Zp), z3_a = smem_u32(Z3p), dzf_a = smem_u32(dZf), rp_a = smem_u32(Rp);
  const uint32_t mb1_a = smem_u32(&s_mb1), mb2_a = smem_u32(&s_mb2), mb3_a = smem_u32(&s_mb3);

  for (int s = 0; s < p.steps; ++s) {
    const int row0 = (int)(((long long)cursor0 + (long long)s * p.cursor_step) % p.cursor_wrap);
    const uint32_t par = s & 1;
    const bool more = s + 1 < p.steps;
    CL_STAMP(0);
    if (tid < kClRG) {
      const int r = i * kClRG + tid;
      s_lab[tid] = r < B ? __ldg(p.labels + row0 + r) : 0;
    }
    mbar_wait(&s_xbar, par);
    CL_STAMP(1);
    // ------------------------------------------------------------ forward: warp w = rows 16w .. 16w+15
    if (warp < MT) {
      float accm[4][4], accs[4][4];                      // main (hi.hi) and correction (lo.hi + hi.lo) sums
#pragma unroll
      for (int n = 0; n < 4; ++n)
#pragma unroll
        for (int e = 0; e < 4; ++e) accm[n][e] = accs[n][e] = 0.f;
      const float* xr0 = Xs + (size_t)(warp * 16 + g) * xp + t;
      const float* xr1 = xr0 + 8 * xp;
      const float4* wf = reinterpret_cast<const float4*>(Wf) + g * 4 + t;
#pragma unroll 2
      for (int kb = 0; kb < kb_i; ++kb) {
        const float av[4] = {xr0[kb * 8], xr1[kb * 8], xr0[kb * 8 + 4], xr1[kb * 8 + 4]};
        float4 bw[4];
#pragma unroll
        for (int n = 0; n < 4; ++n) bw[n] = wf[(size_t)(kb * kClHC + n * 8) * 4];    // pad columns hold zeros
        uint32_t ah[4], al[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          if (kSplit) split_tf32(av[e], ah[e], al[e]);
          else ah[e] = __float_as_uint(av[e]);
        }
        if (kSplit) {
#pragma unroll
          for (int n = 0; n < 4; ++n)
            if (n < nt_j) mma_tf32(accs[n], al, __float_as_uint(bw[n].x), __float_as_uint(bw[n].z));
#pragma unroll
          for (int n = 0; n < 4; ++n)
            if (n < nt_j) mma_tf32(accs[n], ah, __float_as_uint(bw[n].y), __float_as_uint(bw[n].w));
        }
#pragma unroll
        for (int n = 0; n < 4; ++n)
          if (n < nt_j) mma_tf32(accm[n], ah, __float_as_uint(bw[n].x), __float_as_uint(bw[n].z));
      }
      // my tile belongs to row group io = warp/2: my own group goes straight into Zp[src = i], the others into
      // the staging chunk io (swizzled 16-byte chunks: chunk ^= row & 7)
      const int io = warp >> 1;
      float* dst = (io == i) ? Zp + (size_t)i * kClRG * kClHC : dZf + (size_t)io * (kClDzBytes / 4);
      const int rl = (warp & 1) * 16 + g;                // rl & 7 == g
#pragma unroll
      for (int n = 0; n < 4; ++n) {
        if (n < nt_j) {
          const int ch = ((2 * n + (t >> 1)) ^ g) * 4 + ((2 * t) & 3);
          *reinterpret_cast<float2*>(dst + rl * kClHC + ch) = make_float2(accm[n][0] + accs[n][0], accm[n][1] + accs[n][1]);
          *reinterpret_cast<float2*>(dst + (rl + 8) * kClHC + ch) = make_float2(accm[n][2] + accs[n][2], accm[n][3] + accs[n][3]);
        }
      }
    }
    fence_proxy_async_smem();
    __syncthreads();
    CL_STAMP(2);
    if (tid < 4 && tid != i && 2 * tid < MT) {           // exchange 1: chunk i' -> CTA (i', j), slot [src = i]
      const uint32_t peer = (uint32_t)(tid * 4 + j);
      cl_bulk_copy(cl_map(zp_a + (uint32_t)i * kClZpBytes, peer), dZf + (size_t)tid * (kClDzBytes / 4), kClZpBytes,
                   cl_map(mb1_a, peer));
    }
    // ------------------------------------------------------------ head a: a2 and the z3 partial of my block
    const bool group_live = 2 * i < MT;                  // row groups beyond the batch receive (and do) nothing
    if (group_live) mbar_wait(&s_mb1, par);
    CL_STAMP(3);
    {
      const int b = tid >> 3, q = tid & 7;
      const bool row_ok = i * kClRG + b < B;
      float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
      if (group_live) {
#pragma unroll
        for (int src = 0; src < 4; ++src) {
          const float4 v = *reinterpret_cast<const float4*>(Zp + (size_t)(src * kClRG + b) * kClHC + ((q ^ (b & 7)) << 2));
          z.x += v.x; z.y += v.y; z.z += v.z; z.w += v.w;
        }
      }
      float a[4] = {z.x + s_b1[4 * q], z.y + s_b1[4 * q + 1], z.z + s_b1[4 * q + 2], z.w + s_b1[4 * q + 3]};
#pragma unroll
      for (int e = 0; e < 4; ++e) a[e] = (row_ok && 4 * q + e < hcv) ? cl_sigmoid(a[e]) : 0.f;
      *reinterpret_cast<float4*>(A2s + b * kClA2P + 4 * q) = make_float4(a[0], a[1], a[2], a[3]);
    }
    __syncthreads();
    if (tid == 0 && more && group_live) mbar_arrive_expect_tx(&s_mb1, n_col_peers * kClZpBytes);   // everyone is past the wait: re-arm
    float* z3st = Z3st + par * (kClRG * kClCP);
    {
      const int b = tid >> 3, q = tid & 7;
      float z0 = 0.f, z1 = 0.f;
      const float* ar = A2s + b * kClA2P;
#pragma unroll 8
      for (int n = 0; n < kClHC; ++n) {
        const float a = ar[n];
        z0 = fmaf(a, W2s[n * kClCP + q], z0);
        z1 = fmaf(a, W2s[n * kClCP + q + 8], z1);
      }
      z3st[b * kClCP + q] = z0;
      z3st[b * kClCP + q + 8] = z1;
    }
    fence_proxy_async_smem();
    __syncthreads();
    if (tid < 4) {                                       // exchange 2: my z3 partial -> the 4 CTAs of my row group
      const uint32_t peer = (uint32_t)(i * 4 + tid);
      cl_bulk_copy(cl_map(z3_a + (uint32_t)((par * 4 + j) * kClZ3Bytes), peer), z3st, kClZ3Bytes, cl_map(mb2_a, peer));
    }
    CL_STAMP(4);
    mbar_wait(&s_mb2, par);
    CL_STAMP(5);
    // ------------------------------------------------------------ head b: softmax-CE over my row group (8 lanes per row)
    {
      const int b = tid >> 3, q = tid & 7;
      const bool row_ok = i * kClRG + b < B;
      const float* z3 = Z3p + (size_t)par * 4 * kClRG * kClCP;
      float z0 = s_b2[q], z1 = s_b2[q + 8];
#pragma unroll
      for (int src = 0; src < 4; ++src) {
        z0 += z3[(src * kClRG + b) * kClCP + q];
        z1 += z3[(src * kClRG + b) * kClCP + q + 8];
      }
      const bool ok0 = q < C, ok1 = q + 8 < C;
      float m = ok0 ? z0 : -INFINITY;
      int arg = q;
      if (ok1 && z1 > m) { m = z1; arg = q + 8; }
#pragma unroll
      for (int o = 4; o > 0; o >>= 1) {
        const float om = __shfl_xor_sync(0xffffffffu, m, o);
        const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
        if (om > m || (om == m && oa < arg)) { m = om; arg = oa; }
      }
      const float e0 = ok0 ? __expf(z0 - m) : 0.f, e1 = ok1 ? __expf(z1 - m) : 0.f;
      float se = e0 + e1;
#pragma unroll
      for (int o = 4; o > 0; o >>= 1) se += __shfl_xor_sync(0xffffffffu, se, o);
      const int label = s_lab[b];
      const float zsel = (label & 8) ? z1 : z0;
      const float zl = __shfl_sync(0xffffffffu, zsel, (lane & 24) | (label & 7));
      const float inv = 1.f / se;
      dZ3s[b * kClCP + q] = (row_ok && ok0) ? (e0 * inv - (q == label ? 1.f : 0.f)) * p.grad_scale : 0.f;
      dZ3s[b * kClCP + q + 8] = (row_ok && ok1) ? (e1 * inv - (q + 8 == label ? 1.f : 0.f)) * p.grad_scale : 0.f;
      if (q == 0) {
        s_rowloss[b] = row_ok ? m + __logf(se) - zl : 0.f;
        s_rowcorr[b] = (row_ok && arg == label) ? 1 : 0;
      }
    }
    __syncthreads();
    if (tid == 0 && more) mbar_arrive_expect_tx(&s_mb2, 4 * kClZ3Bytes);
    float* rp_mine = Rp + (size_t)i * kClRpRows * kClCP;
    {
      // dZ2[b, n] of my block, written into MY block of the backward A-fragment layout (k-steps 4i .. 4i+3)
      const int b = tid >> 3, q = tid & 7;
      float dz3[kClCP];
#pragma unroll
      for (int c4 = 0; c4 < kClCP / 4; ++c4) {
        const float4 v = *reinterpret_cast<const float4*>(dZ3s + b * kClCP + 4 * c4);
        dz3[4 * c4] = v.x; dz3[4 * c4 + 1] = v.y; dz3[4 * c4 + 2] = v.z; dz3[4 * c4 + 3] = v.w;
      }
      const int brow = i * kClRG + b, kstep = brow >> 3, r8 = brow & 7;
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int n = 4 * q + e;
        float d = 0.f;
#pragma unroll
        for (int c = 0; c < kClCP; ++c) d = fmaf(dz3[c], W2s[n * kClCP + c], d);
        const float a = A2s[b * kClA2P + n];
        d *= a * (1.f - a);
        dZl[b * 33 + n] = d;
        dZf[(((kstep * 2 + (n >> 4)) * 8 + (n & 7)) * 4 + (r8 & 3)) * 4 + (r8 >> 2) * 2 + ((n >> 3) & 1)] = d;
      }
    }
    {
      // partial dW2[n, c] over my 32 rows (thread: hidden unit n = tid / 8, classes q and q + 8)
      const int n = tid >> 3, q = tid & 7;
      float g0 = 0.f, g1 = 0.f;
#pragma unroll 8
      for (int b = 0; b < kClRG; ++b) {
        const float a = A2s[b * kClA2P + n];
        g0 = fmaf(a, dZ3s[b * kClCP + q], g0);
        g1 = fmaf(a, dZ3s[b * kClCP + q + 8], g1);
      }
      rp_mine[n * kClCP + q] = g0;
      rp_mine[n * kClCP + q + 8] = g1;
    }
    __syncthreads();                                     // dZl, s_rowloss complete
    if (tid < kClHC) {                                   // db1 partial of my rows; db2 / loss / correct partials
      float d = 0.f;
      for (int b = 0; b < kClRG; ++b) d += dZl[b * 33 + tid];
      rp_mine[35 * kClCP + tid] = d;
    } else if (tid < kClHC + kClCP) {
      const int c = tid - kClHC;
      float d = 0.f;
      for (int b = 0; b < kClRG; ++b) d += dZ3s[b * kClCP + c];
      rp_mine[32 * kClCP + c] = d;
    } else if (tid == kClHC + kClCP) {
      float l = 0.f;
      int cr = 0;
      for (int b = 0; b < kClRG; ++b) { l += s_rowloss[b]; cr += s_rowcorr[b]; }
      rp_mine[33 * kClCP] = l;
      rp_mine[33 * kClCP + 1] = (float)cr;
    }
    fence_proxy_async_smem();
    __syncthreads();
    if (tid < 4 && tid != i) {                           // exchange 3: dZ2 block + partial gradients -> my column peers
      const uint32_t peer = (uint32_t)(tid * 4 + j);
      const uint32_t mb = cl_map(mb3_a, peer);
      cl_bulk_copy(cl_map(dzf_a + (uint32_t)i * kClDzBytes, peer), dZf + (size_t)i * (kClDzBytes / 4), kClDzBytes, mb);
      cl_bulk_copy(cl_map(rp_a + (uint32_t)i * kClRpBytes, peer), rp_mine, kClRpBytes, mb);
    }
    CL_STAMP(6);
    mbar_wait(&s_mb3, par);
    CL_STAMP(7);
    // ------------------------------------------------------------ backward: dW1[K_i, H_j]^T = dZ2^T . X, warp w owns
    // the 8-column blocks kb = w, w + 8, w + 16, w + 24 of my K slice
    const bool last = s == p.steps - 1;
    {
      float acc[2][4][4];
#pragma unroll
      for (int m = 0; m < 2; ++m)
#pragma unroll
        for (int n = 0; n < 4; ++n) acc[m][n][0] = acc[m][n][1] = acc[m][n][2] = acc[m][n][3] = 0.f;
      const float4* af = reinterpret_cast<const float4*>(dZf) + g * 4 + t;
      const bool two = nt_j > 2;
      int kbo[4];                                        // column offset of my blocks (clamped: loads stay in bounds)
#pragma unroll
      for (int n = 0; n < 4; ++n) kbo[n] = min(warp + 8 * n, kb_i - 1) * 8;
#pragma unroll 2
      for (int kstep = 0; kstep < 2 * MT; ++kstep) {
        const float4 a0 = af[(kstep * 2) * 32], a1 = af[(kstep * 2 + 1) * 32];
        const float* xb0 = Xs + (size_t)(kstep * 8 + t) * xp + g;
        const float* xb1 = xb0 + 4 * xp;
        float bv[4][2];
#pragma unroll
        for (int n = 0; n < 4; ++n) { bv[n][0] = xb0[kbo[n]]; bv[n][1] = xb1[kbo[n]]; }
        const float av[2][4] = {{a0.x, a0.y, a0.z, a0.w}, {a1.x, a1.y, a1.z, a1.w}};
        uint32_t ah[2][4], al[2][4], bh[4][2], bl[4][2];
#pragma unroll
        for (int m = 0; m < 2; ++m)
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            if (kSplit) split_tf32(av[m][e], ah[m][e], al[m][e]);
            else ah[m][e] = __float_as_uint(av[m][e]);
          }
#pragma unroll
        for (int n = 0; n < 4; ++n)
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            if (kSplit) split_tf32(bv[n][e], bh[n][e], bl[n][e]);
            else bh[n][e] = __float_as_uint(bv[n][e]);
          }
        if (kSplit) {
#pragma unroll
          for (int n = 0; n < 4; ++n)
#pragma unroll
            for (int m = 0; m < 2; ++m)
              if (warp + 8 * n < kb_i && (m == 0 || two)) mma_tf32(acc[m][n], al[m], bh[n][0], bh[n][1]);
#pragma unroll
          for (int n = 0; n < 4; ++n)
#pragma unroll
            for (int m = 0; m < 2; ++m)
              if (warp + 8 * n < kb_i && (m == 0 || two)) mma_tf32(acc[m][n], ah[m], bl[n][0], bl[n][1]);
        }
#pragma unroll
        for (int n = 0; n < 4; ++n)
#pragma unroll
          for (int m = 0; m < 2; ++m)
            if (warp + 8 * n < kb_i && (m == 0 || two)) mma_tf32(acc[m][n], ah[m], bh[n][0], bh[n][1]);
      }
      CL_STAMP(8);
      // apply: acc[m][n][e] = dW1[k = kb*8 + 2t + (e&1)][col = m*16 + g + 8*(e>>1)]
#pragma unroll
      for (int n = 0; n < 4; ++n) {
        const int kb = warp + 8 * n;
        if (kb < kb_i) {
#pragma unroll
          for (int m = 0; m < 2; ++m) {
            if (m == 0 || two) {
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const int r8 = 2 * t + (e & 1), col = m * 16 + g + 8 * (e >> 1);
                float2* wp = reinterpret_cast<float2*>(Wf + ((size_t)(kb * kClHC + col) * 4 + (r8 & 3)) * 4 + (r8 >> 2) * 2);
                const float2 hl = *wp;
                const float w = fmaf(-p.lr, acc[m][n][e], hl.x + hl.y);
                uint32_t hi, lo;
                split_tf32(w, hi, lo);
                *wp = make_float2(__uint_as_float(hi), __uint_as_float(lo));
                if (last && col < hcv) p.w1[(size_t)(k0 + kb * 8 + r8) * H + h0 + col] = w;
              }
            }
          }
        }
      }
    }
    // W2 / b1 / b2 replicas: every CTA applies the same sums in the same order
    {
      const int n = tid >> 3, q = tid & 7;
      float g0 = 0.f, g1 = 0.f;
#pragma unroll
      for (int src = 0; src < 4; ++src) {
        g0 += Rp[(src * kClRpRows + n) * kClCP + q];
        g1 += Rp[(src * kClRpRows + n) * kClCP + q + 8];
      }
      const float w0 = fmaf(-p.lr, g0, W2s[n * kClCP + q]), w1v = fmaf(-p.lr, g1, W2s[n * kClCP + q + 8]);
      W2s[n * kClCP + q] = w0;
      W2s[n * kClCP + q + 8] = w1v;
      if (last && i == 0 && n < hcv) {
        if (q < C) p.w2[(size_t)(h0 + n) * C + q] = w0;
        if (q + 8 < C) p.w2[(size_t)(h0 + n) * C + q + 8] = w1v;
      }
    }
    if (tid < kClHC) {
      float d = 0.f;
#pragma unroll
      for (int src = 0; src < 4; ++src) d += Rp[(src * kClRpRows + 35) * kClCP + tid];
      const float v = fmaf(-p.lr, d, s_b1[tid]);
      s_b1[tid] = v;
      if (last && i == 0 && tid < hcv) p.b1[h0 + tid] = v;
    } else if (tid < kClHC + kClCP) {
      const int c = tid - kClHC;
      float d = 0.f;
#pragma unroll
      for (int src = 0; src < 4; ++src) d += Rp[(src * kClRpRows + 32) * kClCP + c];
      const float v = fmaf(-p.lr, d, s_b2[c]);
      s_b2[c] = v;
      if (last && crank == 0 && c < C) p.b2[c] = v;
    } else if (tid == kClHC + kClCP && crank == 0) {
      float l = 0.f, cr = 0.f;
#pragma unroll
      for (int src = 0; src < 4; ++src) {
        l += Rp[(src * kClRpRows + 33) * kClCP];
        cr += Rp[(src * kClRpRows + 33) * kClCP + 1];
      }
      if (p.loss) p.loss[(size_t)s * p.stat_stride] = l * p.loss_scale;
      if (p.correct) p.correct[(size_t)s * p.stat_stride] = (int)(cr + 0.5f);
    }
    __syncthreads();                                     // X tile and parameter replicas are free / updated
    if (tid == 0 && more) mbar_arrive_expect_tx(&s_mb3, n_col_peers * (kClDzBytes + kClRpBytes));
    if (warp == 0 && more) load_x(s + 1);
    CL_STAMP(9);
  }
  if (crank == 0 && tid == 0) {
    if (p.global_step) *p.global_step += (unsigned long long)p.steps;
    if (p.cursor_rw)
      *p.cursor_rw = (int)(((long long)cursor0 + (long long)p.steps * p.cursor_step) % p.cursor_wrap);
  }
  cl_sync();                                             // nobody exits while a peer's copy may still be landing in it
}

}  // namespace b200ps

using namespace b200ps;

static long long* g_cl_timing = nullptr;
extern "C" int b200ps_mlp_cluster_set_timing(void* buf) {
  g_cl_timing = static_cast<long long*>(buf);
  return 0;
}

constexpr size_t kClMaxDynSmem = 227 * 1024 - 1024;    // opt-in maximum minus the kernel's static shared memory

static size_t cluster_smem_bytes(const ClusterParams& p) {
  const size_t KB = p.KSMAX >> 3;
  return sizeof(float) * ((size_t)p.MT * 16 * p.xp + KB * kClHC * 16 + 4 * kClRG * kClHC + 16 * 2 * 8 * 4 * 4 +
                          2 * 4 * kClRG * kClCP + 2 * kClRG * kClCP + 4 * kClRpRows * kClCP + kClRG * kClA2P +
                          kClRG * kClCP + kClHC * kClCP + kClRG * 33) + 128;
}

// Can `steps` consecutive steps of this shape run on the cluster kernel on the current device?  (Shape limits,
// shared memory, and a 16-CTA cluster must be schedulable: checked with cudaOccupancyMaxActiveClusters.)
extern "C" int b200ps_mlp_cluster_supported(int32_t B, int32_t D, int32_t H, int32_t C, int32_t* ok) {
  *ok = 0;
  if (!(B > 0 && B <= 128 && D >= 32 && D <= 1024 && D % 8 == 0 && H > 0 && H <= 128 && H % 4 == 0 && C > 0 && C <= kClCP))
    return 0;
  ClusterParams p{};
  p.MT = (B + 15) / 16;
  p.KSMAX = (((D >> 3) + 3) >> 2) * 8;
  p.xp = p.KSMAX + 4;
  const size_t smem = cluster_smem_bytes(p);
  if (smem > kClMaxDynSmem) return 0;
  static int cached[64] = {0};      // 0 unknown, 1 yes, -1 no (per device; the answer does not depend on the shape)
  int dev = 0;
  B200PS_CUDA_CHECK(cudaGetDevice(&dev));
  if (cached[dev & 63] == 0) {
    auto kern = mlp_cluster_kernel<true>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kClMaxDynSmem);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(mlp_cluster_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kClMaxDynSmem);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(mlp_cluster_kernel<false>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    int n = 0;
    if (e == cudaSuccess) {
      cudaLaunchConfig_t cfg{};
      cfg.gridDim = dim3(kClCtas); cfg.blockDim = dim3(kClThreads); cfg.dynamicSmemBytes = kClMaxDynSmem;
      cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeClusterDimension;
      at[0].val.clusterDim.x = kClCtas; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
      cfg.attrs = at; cfg.numAttrs = 1;
      e = cudaOccupancyMaxActiveClusters(&n, kern, &cfg);
    }
    if (e != cudaSuccess) cudaGetLastError();
    cached[dev & 63] = (e == cudaSuccess && n >= 1) ? 1 : -1;
  }
  *ok = cached[dev & 63] == 1 ? 1 : 0;
  return 0;
}

extern "C" int b200ps_mlp_cluster_steps(const float* x, const int32_t* labels, const int32_t* cursor, float* w1,
                                        float* b1, float* w2, float* b2, float* loss, int32_t* correct,
                                        int32_t stat_stride, uint64_t* global_step, int32_t* cursor_rw,
                                        int32_t cursor_step, int32_t cursor_wrap, int32_t B, int32_t D, int32_t H,
                                        int32_t C, int32_t steps, float lr, float grad_scale, float loss_scale,
                                        int32_t flags, b200ps_stream_t stream) {
  int ok = 0;
  int rc = b200ps_mlp_cluster_supported(B, D, H, C, &ok);
  if (rc) return rc;
  B200PS_REQUIRE(ok, "mlp_cluster_steps: shape B=%d D=%d H=%d C=%d (or this device) is not supported: needs B <= 128, "
                 "32 <= D <= 1024 (multiple of 8), H <= 128 (multiple of 4), C <= 16 and a schedulable 16-CTA cluster",
                 B, D, H, C);
  B200PS_REQUIRE(steps >= 1, "mlp_cluster_steps: steps must be >= 1");
  B200PS_REQUIRE(((uintptr_t)x | (uintptr_t)w1) % 16 == 0, "mlp_cluster_steps: x and w1 must be 16-byte aligned");
  B200PS_REQUIRE(steps == 1 || cursor_wrap >= 1, "mlp_cluster_steps: bad cursor_wrap");
  ClusterParams p{};
  p.x = x; p.labels = labels; p.cursor = cursor; p.w1 = w1; p.b1 = b1; p.w2 = w2; p.b2 = b2;
  p.loss = loss; p.correct = correct; p.stat_stride = stat_stride;
  p.global_step = reinterpret_cast<unsigned long long*>(global_step);
  p.cursor_rw = cursor_rw; p.cursor_step = cursor_step; p.cursor_wrap = cursor_wrap > 0 ? cursor_wrap : 1;
  p.B = B; p.D = D; p.H = H; p.C = C; p.steps = steps;
  p.MT = (B + 15) / 16;
  p.KSMAX = (((D >> 3) + 3) >> 2) * 8;
  p.xp = p.KSMAX + 4;
  p.lr = lr; p.grad_scale = grad_scale; p.loss_scale = loss_scale;
  p.timing = g_cl_timing;
  const size_t smem = cluster_smem_bytes(p);
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(kClCtas); cfg.blockDim = dim3(kClThreads); cfg.dynamicSmemBytes = smem;
  cfg.stream = as_stream(stream);
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = kClCtas; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = pdl_enabled() ? 2 : 1;
  static const bool split = [] { const char* e = getenv("B200PS_CLUSTER_SPLIT"); return !(e && e[0] == '0'); }();
  const bool use_split = (flags & 1) ? false : split;
  if (use_split) B200PS_CUDA_CHECK(cudaLaunchKernelEx(&cfg, mlp_cluster_kernel<true>, p));
  else B200PS_CUDA_CHECK(cudaLaunchKernelEx(&cfg, mlp_cluster_kernel<false>, p));
  B200PS_LAUNCH_CHECK();
  return 0;
}
