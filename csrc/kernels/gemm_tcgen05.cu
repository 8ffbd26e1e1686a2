// b200ps K2/K7 — persistent warp-specialised GEMM on the 5th-gen tensor cores.
//
//   C[M,N] = epilogue(alpha * op(A) * op(B)),  fp32 accumulation in TMEM.
//
// * operands: bf16 (kind::f16) or fp32 storage consumed as tf32 (kind::tf32), each either
//   K-major or MN-major, so forward, dgrad and wgrad of a dense layer all read the SAME
//   row-major tensors with no transpose pass (example.py:95,97 + their gradients);
// * TMA (cp.async.bulk.tensor, 128-byte swizzle) feeds a kStages-deep smem ring; out-of-bounds
//   parts of a box are zero-filled by the hardware, which is how M/N/K tails are handled;
// * one elected thread issues tcgen05.mma (UMMA 128 x BLOCK_N x 32 bytes of K);
// * two TMEM accumulator stages: the epilogue of tile i overlaps the MMAs of tile i+1;
// * epilogue warps read TMEM with tcgen05.ld and apply bias / sigmoid / sigmoid' / SGD
//   directly from registers.
//
// Warp roles (384 threads): w0 TMA producer, w1 MMA issuer, w2 TMEM allocator, w3 K2 gather,
// w4..w11 epilogue (TMEM lane quarter = warp_idx % 4; the two warps of a quarter take alternate
// 32-column chunks).
#include "../common.h"
#include "ptx.cuh"

namespace b200ps {

constexpr int kBlockM = 128;
constexpr int kRowBytes = 128;           // one swizzle-128B row: 64 bf16 or 32 tf32 along K (or MN)
constexpr int kABytes = kBlockM * kRowBytes;  // 16 KiB per stage
constexpr int kGemmThreads = 384;   // 4 role warps + 8 epilogue warps
constexpr int kGatherChunk = 16384;   // bytes per bulk copy of the fused pull; ring of 2 per CTA

struct GemmParams {
  int M, N, K;
  int a_mn, b_mn;
  int num_m_blocks, num_n_blocks, num_k_blocks;
  int split_k, k_blocks_per_split;
  int epilogue, act, c_is_f32, vec_ok;
  float alpha, lr;
  void* C;
  long long ldc;
  const float* bias;
  const void* aux;
  long long ldaux;
  float* master;
  void* lowp;
  const int* a_row_offset;   // device scalar added to A's stored-row coordinate (resident datasets)
  long long c_split_stride;  // EPI_STORE with split_k > 1: split s writes its partial to C + s*stride
  // ---- K2: pull (all-gather of the weight shards) fused into this GEMM.  B is [K,N] row-major
  // bf16, row-sharded over `gather_world` ranks; `gather_src[r]` is rank r's copy (valid in its own
  // row range, peer-mapped), `gather_dst` the local copy the TMA descriptor of B points at.
  const void* const* gather_src;
  void* gather_dst;
  unsigned int* gather_flags;   // [num_k_blocks + 1]: panel-landed flags + CTA arrival counter
  int gather_world, gather_rank, gather_ctas, gather_kb_per_rank;
  float* colsum;   // [N] += column sums of the stored tile (fused bias gradient), or null
};

__device__ __forceinline__ float sigmoidf_fast(float x) { return __fdividef(1.f, 1.f + __expf(-x)); }

template <bool kTf32, int BLOCK_N, int kStages>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA,
                    const __grid_constant__ CUtensorMap tmB, const GemmParams p) {
  constexpr int kElemBytes = kTf32 ? 4 : 2;
  constexpr int BLOCK_K = kRowBytes / kElemBytes;  // K elements per stage: 64 (bf16) / 32 (tf32)
  constexpr int ATOM_MN = kRowBytes / kElemBytes;  // MN elements per swizzle atom row
  constexpr int UMMA_K = 32 / kElemBytes;          // 16 (bf16) / 8 (tf32)
  constexpr int kBBytes = BLOCK_N * kRowBytes;
  constexpr int kStageBytes = kABytes + kBBytes;
  constexpr uint32_t kTmemCols = 2 * BLOCK_N;
  static_assert(BLOCK_N % 32 == 0 && BLOCK_N >= 32 && BLOCK_N <= 256, "BLOCK_N");
  static_assert(kTmemCols >= 32 && kTmemCols <= 512, "TMEM columns");

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // 1024-byte alignment is required by the 128-byte swizzle (pattern repeats every 8 rows).
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kStages * kStageBytes);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tmem_full_bar = empty_bar + kStages;
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;
  uint32_t* tmem_base_smem = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);

  const int warp_idx = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;

  if (warp_idx == 0 && elect_one()) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp_idx == 1 && elect_one()) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full_bar[s], 1);
      mbar_init(&tmem_empty_bar[s], 8);  // one arrive per epilogue warp
      mbar_init(&tmem_empty_bar[3 + s], 1);  // gather ring (K2)
    }
    fence_mbar_init();
  }
  if (warp_idx == 2) tmem_alloc<kTmemCols>(tmem_base_smem);
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_base_smem;

  // Everything above overlapped the previous kernel (programmatic dependent launch); from here
  // on we read what it produced.  Let our own successor start its prologue right away.
  pdl_wait();
  pdl_trigger();
  const int a_row0 = p.a_row_offset ? __ldg(p.a_row_offset) : 0;

  const int tiles_mn = p.num_m_blocks * p.num_n_blocks;
  const int total_tiles = tiles_mn * p.split_k;
  // K2: with a fused pull every CTA's otherwise idle warp 3 streams remote weight panels into the
  // local copy; the k loop starts at this rank's own (already local) panels.
  const int gemm_cta = (int)blockIdx.x;
  const int gemm_ctas = (int)gridDim.x;
  const int k_rot = p.gather_src ? p.gather_rank * p.gather_kb_per_rank : 0;

  if (warp_idx == 3) {
    // ===================== gather role (K2) =====================
    // One thread per CTA moves 16 KiB chunks  peer HBM -> smem ring -> local HBM  with 1-D bulk
    // copies (cp.async.bulk).  Chunks of the remote panels, in ring order after our own range, are
    // dealt round-robin over ALL CTAs of the launch, so ~gridDim x 32 KiB are in flight over NVLink
    // without giving up a single SM.  A panel (BLOCK_K rows of B, contiguous) is published by
    // counting its landed chunks in gather_flags[kb].
    if (p.gather_src && elect_one()) {
      constexpr uint32_t kChunk = kGatherChunk;
      uint8_t* ring = smem + kStages * kStageBytes + 256;
      uint64_t* gbar = tmem_empty_bar + 3;            // two spare mbarriers after the tmem pointer slot
      const size_t panel_bytes = (size_t)BLOCK_K * p.N * kElemBytes;
      const int cpp = (int)(panel_bytes / kChunk);    // chunks per panel (host guarantees divisibility)
      const int nkb = p.num_k_blocks;
      const long long total = (long long)(nkb - p.gather_kb_per_rank) * cpp;   // remote chunks, whole launch
      // software pipeline, two loads in flight: load(n+1) is issued before load(n) is consumed
      auto chunk_kb = [&](long long c) { return (k_rot + p.gather_kb_per_rank + (int)(c / cpp)) % nkb; };
      auto issue_load = [&](long long c, int n) {
        const int slot = n & 1;
        const int kb = chunk_kb(c);
        const size_t off = (size_t)kb * panel_bytes + (size_t)(c % cpp) * kChunk;
        mbar_arrive_expect_tx(&gbar[slot], kChunk);
        bulk_load_g2s(ring + slot * kChunk,
                      reinterpret_cast<const uint8_t*>(p.gather_src[kb / p.gather_kb_per_rank]) + off, kChunk, &gbar[slot]);
      };
      auto publish = [&](long long c) {   // chunk c is completely in local memory
        fence_proxy_async_all();
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(p.gather_flags + chunk_kb(c)) : "memory");
      };
      long long c = blockIdx.x;
      if (c < total) issue_load(c, 0);
      long long prev = -1;   // chunk whose store is committed but not yet known complete
      for (int n = 0; c < total; c += gridDim.x, ++n) {
        const long long cn = c + gridDim.x;
        if (cn < total) {
          if (n >= 1) bulk_wait_read<0>();            // slot (n+1)&1 was read by the store of chunk n-1
          issue_load(cn, n + 1);
        }
        const int slot = n & 1;
        const size_t off = (size_t)chunk_kb(c) * panel_bytes + (size_t)(c % cpp) * kChunk;
        mbar_wait(&gbar[slot], (uint32_t)(n >> 1) & 1u);
        bulk_store_s2g(reinterpret_cast<uint8_t*>(p.gather_dst) + off, ring + slot * kChunk, kChunk);
        bulk_commit();
        if (prev >= 0) { bulk_wait_all<1>(); publish(prev); }
        prev = c;
      }
      if (prev >= 0) { bulk_wait_all<0>(); publish(prev); }
    }
  } else if (warp_idx == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      uint32_t seen[16] = {0};   // panels already observed as landed (<= 512 k-blocks)
      const uint32_t gather_cpp = p.gather_src ? (uint32_t)(((size_t)BLOCK_K * p.N * kElemBytes) / kGatherChunk) : 0u;
      for (int t = gemm_cta; t < total_tiles; t += gemm_ctas) {
        const int split = t % p.split_k;
        const int tmn = t / p.split_k;
        const int m_blk = tmn % p.num_m_blocks;
        const int n_blk = tmn / p.num_m_blocks;
        const int kb0 = split * p.k_blocks_per_split;
        const int kb1 = min(kb0 + p.k_blocks_per_split, p.num_k_blocks);
        for (int ki = kb0; ki < kb1; ++ki) {
          int kb = ki;
          if (p.gather_src) {   // own panels first; a remote panel must have landed locally
            kb = (ki + k_rot) % p.num_k_blocks;
            if (kb / p.gather_kb_per_rank != p.gather_rank && !((seen[kb >> 5] >> (kb & 31)) & 1u)) {
              const uint64_t t0 = globaltimer_ns();
              while (ld_acquire_gpu(p.gather_flags + kb) < gather_cpp) {
                if (globaltimer_ns() - t0 > B200PS_WAIT_TIMEOUT_NS) {
                  printf("b200ps: gather panel %d never landed (block %d)\n", kb, blockIdx.x);
                  __trap();
                }
              }
              // the flag acquire is a generic-proxy read; the panel it guards was written by bulk-async stores
              // and is read next by TMA (async proxy): order the two proxies before the first load of the panel
              fence_proxy_async_all();
              seen[kb >> 5] |= 1u << (kb & 31);
            }
          }
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sA = smem + stage * kStageBytes;
          uint8_t* sB = sA + kABytes;
          mbar_arrive_expect_tx(&full_bar[stage], kStageBytes);
          if (!p.a_mn) {
            tma_load_2d(&tmA, &full_bar[stage], sA, kb * BLOCK_K, a_row0 + m_blk * kBlockM);
          } else {
#pragma unroll
            for (int j = 0; j < kBlockM / ATOM_MN; ++j)
              tma_load_2d(&tmA, &full_bar[stage], sA + j * (BLOCK_K * kRowBytes),
                          m_blk * kBlockM + j * ATOM_MN, a_row0 + kb * BLOCK_K);
          }
          if (!p.b_mn) {
            tma_load_2d(&tmB, &full_bar[stage], sB, kb * BLOCK_K, n_blk * BLOCK_N);
          } else {
#pragma unroll
            for (int j = 0; j < BLOCK_N / ATOM_MN; ++j)
              tma_load_2d(&tmB, &full_bar[stage], sB + j * (BLOCK_K * kRowBytes),
                          n_blk * BLOCK_N + j * ATOM_MN, kb * BLOCK_K);
          }
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp_idx == 1) {
    // ===================== MMA issuer =====================
    const uint32_t idesc = make_idesc(kTf32 ? 2u : 1u, p.a_mn, p.b_mn, kBlockM, BLOCK_N);
    // K-major SW128: 8-row groups 1024 B apart (SBO); LBO unused.  MN-major SW128: atoms of
    // 64 B x 8 k-rows; next MN atom BLOCK_K*128 B away (LBO); next 8 k-rows 1024 B away (SBO).
    const uint32_t a_lbo = p.a_mn ? BLOCK_K * kRowBytes : 16;
    const uint32_t b_lbo = p.b_mn ? BLOCK_K * kRowBytes : 16;
    const uint32_t a_kstep = p.a_mn ? UMMA_K * kRowBytes : 32;  // bytes per UMMA_K advance
    const uint32_t b_kstep = p.b_mn ? UMMA_K * kRowBytes : 32;
    // MN-major tf32 must use the 128B swizzle with 32-byte atoms (4 k-rows per atom => the
    // two 4-row groups of one K=8 MMA are 512 B apart); everything else is plain SWIZZLE_128B.
    const uint32_t a_type = (kTf32 && p.a_mn) ? 1u : 2u, a_sbo = (kTf32 && p.a_mn) ? 512u : 1024u;
    const uint32_t b_type = (kTf32 && p.b_mn) ? 1u : 2u, b_sbo = (kTf32 && p.b_mn) ? 512u : 1024u;
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int t = gemm_cta; t < total_tiles; t += gemm_ctas) {
      const int split = t % p.split_k;
      const int kb0 = split * p.k_blocks_per_split;
      const int kb1 = min(kb0 + p.k_blocks_per_split, p.num_k_blocks);
      mbar_wait(&tmem_empty_bar[acc], acc_phase ^ 1);
      tcgen05_fence_after();
      const uint32_t d_tmem = tmem_base + acc * BLOCK_N;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tcgen05_fence_after();
        if (elect_one()) {
          const uint32_t a_addr = smem_u32(smem + stage * kStageBytes);
          const uint32_t b_addr = a_addr + kABytes;
#pragma unroll
          for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
            const uint64_t adesc = make_smem_desc_sw128(a_addr + k * a_kstep, a_lbo, a_sbo, a_type);
            const uint64_t bdesc = make_smem_desc_sw128(b_addr + k * b_kstep, b_lbo, b_sbo, b_type);
            const uint32_t accumulate = (kb > kb0 || k > 0) ? 1u : 0u;
            if (kTf32) umma_tf32(d_tmem, adesc, bdesc, idesc, accumulate);
            else       umma_f16(d_tmem, adesc, bdesc, idesc, accumulate);
          }
          umma_commit(&empty_bar[stage]);                      // frees the smem slot
          if (kb == kb1 - 1) umma_commit(&tmem_full_bar[acc]);  // accumulator ready
        }
        __syncwarp();
        if (++stage == kStages) { stage = 0; phase ^= 1; }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  } else if (warp_idx >= 4) {
    // ===================== epilogue =====================
    const int quarter = warp_idx & 3;
    const int half = (warp_idx - 4) >> 2;     // which of the quarter's two warps: chunks half, half+2, ...
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int t = gemm_cta; t < total_tiles; t += gemm_ctas) {
      const int tmn = t / p.split_k;
      const int split = t % p.split_k;
      const int m_blk = tmn % p.num_m_blocks;
      const int n_blk = tmn / p.num_m_blocks;
      const int row = m_blk * kBlockM + quarter * 32 + lane;
      const bool row_ok = row < p.M;
      mbar_wait(&tmem_full_bar[acc], acc_phase);
      tcgen05_fence_after();
      const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + acc * BLOCK_N;
      const int n_valid = min(BLOCK_N, p.N - n_blk * BLOCK_N);
      const int n_chunks = (n_valid + 31) / 32;
      if (half >= n_chunks) {   // nothing for this warp in a narrow tile: just hand the stage back
        tcgen05_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tmem_empty_bar[acc]);
      }
      for (int c = half; c < n_chunks; c += 2) {
        uint32_t r[32];
        tmem_ld_32x32(t_addr + c * 32, r);
        tmem_ld_wait();
        if (c + 2 >= n_chunks) {
          // all of this warp's TMEM reads for the tile are done: hand the stage back
          tcgen05_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tmem_empty_bar[acc]);
        }
        const int col0 = n_blk * BLOCK_N + c * 32;
        const int ncols = min(32, p.N - col0);
        if (row_ok) {
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]) * p.alpha;

        if (p.epilogue == B200PS_EPI_STORE) {
          if (p.bias != nullptr) {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (j < ncols) v[j] += __ldg(p.bias + col0 + j);
          }
          if (p.act == B200PS_ACT_SIGMOID) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = sigmoidf_fast(v[j]);
          }
          if (p.aux != nullptr) {
            if (p.c_is_f32) {
              const float* a = reinterpret_cast<const float*>(p.aux) + row * p.ldaux + col0;
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (j < ncols) { const float s = a[j]; v[j] *= s * (1.f - s); }
            } else {
              const __nv_bfloat16* a =
                  reinterpret_cast<const __nv_bfloat16*>(p.aux) + row * p.ldaux + col0;
              if (p.vec_ok && ncols == 32) {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                  const uint4 u = *reinterpret_cast<const uint4*>(a + q * 8);
                  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
                  for (int e = 0; e < 4; ++e) {
                    const float2 f = __bfloat1622float2(h[e]);
                    v[q * 8 + 2 * e] *= f.x * (1.f - f.x);
                    v[q * 8 + 2 * e + 1] *= f.y * (1.f - f.y);
                  }
                }
              } else {
#pragma unroll
                for (int j = 0; j < 32; ++j)
                  if (j < ncols) { const float s = __bfloat162float(a[j]); v[j] *= s * (1.f - s); }
              }
            }
          }
          if (p.colsum != nullptr) {   // bias gradient: column sums of the tile rows, one red per column per warp
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const float cs = warp_sum(v[j]);
              if (lane == 0 && j < ncols) red_add_f32_gpu(p.colsum + col0 + j, cs);
            }
          }
          if (p.c_is_f32) {
            float* cp = reinterpret_cast<float*>(p.C) + split * p.c_split_stride + row * p.ldc + col0;
            if (p.vec_ok && ncols == 32) {
#pragma unroll
              for (int q = 0; q < 8; ++q)
                *reinterpret_cast<float4*>(cp + q * 4) =
                    make_float4(v[q * 4], v[q * 4 + 1], v[q * 4 + 2], v[q * 4 + 3]);
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (j < ncols) cp[j] = v[j];
            }
          } else {
            __nv_bfloat16* cp = reinterpret_cast<__nv_bfloat16*>(p.C) + split * p.c_split_stride + row * p.ldc + col0;
            if (p.vec_ok && ncols == 32) {
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                uint4 u;
                __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
                for (int e = 0; e < 4; ++e)
                  h[e] = __floats2bfloat162_rn(v[q * 8 + 2 * e], v[q * 8 + 2 * e + 1]);
                *reinterpret_cast<uint4*>(cp + q * 8) = u;
              }
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (j < ncols) cp[j] = __float2bfloat16(v[j]);
            }
          }
        } else if (p.epilogue == B200PS_EPI_ATOMIC_ADD) {
          float* cp = reinterpret_cast<float*>(p.C) + row * p.ldc + col0;
          if (p.vec_ok && ncols == 32) {
#pragma unroll
            for (int q = 0; q < 8; ++q)
              asm volatile("red.relaxed.gpu.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(cp + q * 4),
                           "f"(v[q * 4]), "f"(v[q * 4 + 1]), "f"(v[q * 4 + 2]), "f"(v[q * 4 + 3])
                           : "memory");
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (j < ncols) red_add_f32_gpu(cp + j, v[j]);
          }
        } else {  // B200PS_EPI_SGD: master -= lr * grad, refresh the bf16 shadow
          float* mp = p.master + row * p.ldc + col0;
          __nv_bfloat16* lp =
              p.lowp ? reinterpret_cast<__nv_bfloat16*>(p.lowp) + row * p.ldc + col0 : nullptr;
          if (p.vec_ok && ncols == 32) {
            // The master is touched exactly once per step: stream it (evict-first loads and stores)
            // so that 0.67 GB of read-modify-write per layer does not push the GEMM operands out of L2.
            float4 w[8];
#pragma unroll
            for (int q = 0; q < 8; ++q) w[q] = __ldcs(reinterpret_cast<const float4*>(mp + q * 4));
#pragma unroll
            for (int q = 0; q < 8; ++q) {
              w[q].x -= p.lr * v[q * 4]; w[q].y -= p.lr * v[q * 4 + 1];
              w[q].z -= p.lr * v[q * 4 + 2]; w[q].w -= p.lr * v[q * 4 + 3];
              __stcs(reinterpret_cast<float4*>(mp + q * 4), w[q]);
              v[q * 4] = w[q].x; v[q * 4 + 1] = w[q].y; v[q * 4 + 2] = w[q].z; v[q * 4 + 3] = w[q].w;
            }
            if (lp) {
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                uint4 u;
                __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
                for (int e = 0; e < 4; ++e)
                  h[e] = __floats2bfloat162_rn(v[q * 8 + 2 * e], v[q * 8 + 2 * e + 1]);
                *reinterpret_cast<uint4*>(lp + q * 8) = u;
              }
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (j < ncols) {
                const float w = mp[j] - p.lr * v[j];
                mp[j] = w;
                if (lp) lp[j] = __float2bfloat16(w);
              }
          }
        }
        }  // row_ok
        __syncwarp();
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp_idx == 2) {
    tcgen05_fence_after();
    tmem_dealloc<kTmemCols>(tmem_base);
  }
  if (p.gather_src && threadIdx.x == 0) {
    // last CTA out re-arms the panel flags for the next launch (nobody can still be reading them)
    __threadfence();
    if (atomicAdd(p.gather_flags + p.num_k_blocks, 1u) == gridDim.x - 1) {
      for (int kb = 0; kb <= p.num_k_blocks; ++kb) p.gather_flags[kb] = 0u;
    }
  }
}

// ------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------
static int encode_map(CUtensorMap* map, const void* ptr, bool f32, uint64_t inner, uint64_t outer,
                      uint64_t row_stride_elems, uint32_t box_inner, uint32_t box_outer,
                      bool atom32 = false) {
  tensormap_encode_fn enc = get_tensormap_encode();
  if (!enc) return set_error(B200PS_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
  const uint64_t esz = f32 ? 4 : 2;
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {row_stride_elems * esz};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                   const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   atom32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return set_error(B200PS_ERR_CUDA,
                     "cuTensorMapEncodeTiled failed (%d): ptr=%p inner=%llu outer=%llu stride=%llu "
                     "box=%ux%u",
                     (int)r, ptr, (unsigned long long)inner, (unsigned long long)outer,
                     (unsigned long long)(row_stride_elems * esz), box_inner, box_outer);
  return 0;
}

template <bool kTf32, int BLOCK_N, int kStages>
static int launch_gemm(const b200ps_gemm_args& a, cudaStream_t stream) {
  constexpr int esz = kTf32 ? 4 : 2;
  constexpr int BLOCK_K = kRowBytes / esz;
  constexpr int ATOM_MN = kRowBytes / esz;
  constexpr int kStageBytes = kABytes + BLOCK_N * kRowBytes;
  constexpr int kSmem = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/ + 2 * kGatherChunk /*K2 ring*/;
  static_assert(kSmem <= 232448, "shared memory budget");

  CUtensorMap tmA, tmB;
  int rc;
  // a_rows_total widens the stored-row extent of A beyond this call's M (or K): the GEMM then
  // reads rows [*a_row_offset, ...) of a larger resident array; rows past the array are zero-filled
  const uint64_t a_rows = a.a_rows_total > 0 ? (uint64_t)a.a_rows_total : (uint64_t)(a.a_mn_major ? a.K : a.M);
  if (!a.a_mn_major) rc = encode_map(&tmA, a.A, kTf32, a.K, a_rows, a.lda, BLOCK_K, kBlockM);
  else               rc = encode_map(&tmA, a.A, kTf32, a.M, a_rows, a.lda, ATOM_MN, BLOCK_K, kTf32);
  if (rc) return rc;
  if (!a.b_mn_major) rc = encode_map(&tmB, a.B, kTf32, a.K, a.N, a.ldb, BLOCK_K, BLOCK_N);
  else               rc = encode_map(&tmB, a.B, kTf32, a.N, a.K, a.ldb, ATOM_MN, BLOCK_K, kTf32);
  if (rc) return rc;

  GemmParams p{};
  p.M = a.M; p.N = a.N; p.K = a.K;
  p.a_mn = a.a_mn_major ? 1 : 0;
  p.b_mn = a.b_mn_major ? 1 : 0;
  p.num_m_blocks = (a.M + kBlockM - 1) / kBlockM;
  p.num_n_blocks = (a.N + BLOCK_N - 1) / BLOCK_N;
  p.num_k_blocks = (a.K + BLOCK_K - 1) / BLOCK_K;
  int split = a.split_k < 1 ? 1 : a.split_k;
  if (split > p.num_k_blocks) split = p.num_k_blocks;
  p.k_blocks_per_split = (p.num_k_blocks + split - 1) / split;
  p.split_k = (p.num_k_blocks + p.k_blocks_per_split - 1) / p.k_blocks_per_split;
  p.epilogue = a.epilogue;
  p.act = a.act;
  p.c_is_f32 = (a.c_dtype == B200PS_DT_F32) ? 1 : 0;
  p.alpha = a.alpha;
  p.lr = a.sgd_lr;
  p.C = a.C;
  p.ldc = a.ldc;
  p.bias = a.bias;
  p.aux = a.aux;
  p.ldaux = a.ldaux;
  p.master = a.sgd_master;
  p.lowp = a.sgd_lowp;
  p.a_row_offset = a.a_row_offset;
  p.c_split_stride = a.c_split_stride;
  p.colsum = a.colsum_out;
  if (a.colsum_out) {
    B200PS_REQUIRE(a.epilogue == B200PS_EPI_STORE && p.split_k == 1 && a.M % kBlockM == 0,
                   "gemm: colsum_out needs a plain store epilogue and M a multiple of %d (whole warps of valid rows)", kBlockM);
    B200PS_CUDA_CHECK(cudaMemsetAsync(a.colsum_out, 0, sizeof(float) * a.N, stream));
  }
  int gather_ctas = 0;
  if (a.gather_src) {
    const bool is_bf16 = !kTf32;   // (a runtime value keeps the tf32 instantiations free of dead-code warnings)
    B200PS_REQUIRE(is_bf16 && a.b_mn_major && a.ldb == a.N && p.split_k == 1 && a.epilogue == B200PS_EPI_STORE,
                   "gemm: the fused pull needs bf16, B stored [K,N] densely (ldb == N), split_k == 1 and a store epilogue");
    B200PS_REQUIRE(a.gather_world >= 1 && a.gather_rank >= 0 && a.gather_rank < a.gather_world && a.gather_flags,
                   "gemm: bad gather rank/world/flags");
    B200PS_REQUIRE(a.K % (a.gather_world * BLOCK_K) == 0 && p.num_k_blocks <= 512,
                   "gemm: fused pull needs K (%d) divisible by world*%d and at most 512 k-blocks", a.K, BLOCK_K);
    p.gather_src = a.gather_src;
    p.gather_dst = const_cast<void*>(a.B);
    p.gather_flags = a.gather_flags;
    p.gather_world = a.gather_world;
    p.gather_rank = a.gather_rank;
    p.gather_kb_per_rank = p.num_k_blocks / a.gather_world;
    B200PS_REQUIRE(((size_t)BLOCK_K * a.N * 2) % kGatherChunk == 0,
                   "gemm: fused pull needs 64*N*2 bytes (one row panel) to be a multiple of %d", kGatherChunk);
    gather_ctas = 0;
    p.gather_ctas = 0;
    if (a.gather_world == 1) p.gather_src = nullptr;   // nothing to pull: plain GEMM
  }
  // 16-byte vector path: every row start and every 32-column chunk start must be 16 B aligned
  auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
  const int c_esz = (a.epilogue == B200PS_EPI_SGD) ? 2 : (p.c_is_f32 ? 4 : 2);
  bool vec = (a.ldc * c_esz) % 16 == 0;
  if (a.epilogue == B200PS_EPI_SGD) vec = vec && al16(a.sgd_master) && (!a.sgd_lowp || al16(a.sgd_lowp));
  else vec = vec && al16(a.C);
  if (a.aux) vec = vec && al16(a.aux) && (a.ldaux * c_esz) % 16 == 0;
  p.vec_ok = vec ? 1 : 0;

  auto kern = gemm_tcgen05_kernel<kTf32, BLOCK_N, kStages>;
  static bool attr_set[64] = {false};      // per template instantiation AND per device
  int dev = 0;
  B200PS_CUDA_CHECK(cudaGetDevice(&dev));
  if (!attr_set[dev & 63]) {
    B200PS_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
    attr_set[dev & 63] = true;
  }
  const int total_tiles = p.num_m_blocks * p.num_n_blocks * p.split_k;
  int ctas = a.max_ctas > 0 ? a.max_ctas : sm_count();
  if (p.gather_src) ctas = std::min(ctas, sm_count());   // all CTAs co-resident: they wait on each other's gather work
  if (ctas > total_tiles) ctas = total_tiles;
  B200PS_CUDA_CHECK(launch_pdl(kern, dim3(ctas), dim3(kGemmThreads), kSmem, stream, tmA, tmB, p));
  B200PS_LAUNCH_CHECK();
  return 0;
}

}  // namespace b200ps

extern "C" int b200ps_gemm(const b200ps_gemm_args* args, b200ps_stream_t stream) {
  using namespace b200ps;
  B200PS_REQUIRE(args != nullptr, "gemm: null args");
  const b200ps_gemm_args& a = *args;
  B200PS_REQUIRE(a.M > 0 && a.N > 0 && a.K > 0, "gemm: bad shape %dx%dx%d", a.M, a.N, a.K);
  B200PS_REQUIRE(a.ab_dtype == B200PS_DT_BF16 || a.ab_dtype == B200PS_DT_F32, "gemm: bad ab_dtype");
  B200PS_REQUIRE(a.c_dtype == B200PS_DT_BF16 || a.c_dtype == B200PS_DT_F32, "gemm: bad c_dtype");
  const int esz = a.ab_dtype == B200PS_DT_F32 ? 4 : 2;
  B200PS_REQUIRE(a.A && a.B, "gemm: null operand");
  B200PS_REQUIRE((reinterpret_cast<uintptr_t>(a.A) & 15) == 0 && (a.lda * esz) % 16 == 0,
                 "gemm: A must be 16-byte aligned with a 16-byte-multiple row stride (lda=%lld)",
                 (long long)a.lda);
  B200PS_REQUIRE((reinterpret_cast<uintptr_t>(a.B) & 15) == 0 && (a.ldb * esz) % 16 == 0,
                 "gemm: B must be 16-byte aligned with a 16-byte-multiple row stride (ldb=%lld)",
                 (long long)a.ldb);
  B200PS_REQUIRE(a.lda >= (a.a_mn_major ? a.M : a.K), "gemm: lda too small");
  B200PS_REQUIRE(a.ldb >= (a.b_mn_major ? a.N : a.K), "gemm: ldb too small");
  B200PS_REQUIRE(a.ldc >= a.N, "gemm: ldc too small");
  switch (a.epilogue) {
    case B200PS_EPI_STORE:
      B200PS_REQUIRE(a.C != nullptr, "gemm: null C");
      B200PS_REQUIRE(a.split_k <= 1 || (a.c_split_stride >= (int64_t)a.M * a.ldc && !a.bias && !a.aux &&
                                        a.act == B200PS_ACT_NONE),
                     "gemm: split_k > 1 needs B200PS_EPI_ATOMIC_ADD, or a plain store into per-split slabs "
                     "(c_split_stride >= M*ldc)");
      B200PS_REQUIRE(a.split_k <= 1 || (a.c_split_stride * (a.c_dtype == B200PS_DT_F32 ? 4 : 2)) % 16 == 0,
                     "gemm: c_split_stride must keep slabs 16-byte aligned");
      break;
    case B200PS_EPI_ATOMIC_ADD:
      B200PS_REQUIRE(a.C != nullptr && a.c_dtype == B200PS_DT_F32, "gemm: atomic epilogue needs fp32 C");
      B200PS_REQUIRE(!a.bias && !a.aux && a.act == B200PS_ACT_NONE, "gemm: atomic epilogue is plain");
      break;
    case B200PS_EPI_SGD:
      B200PS_REQUIRE(a.sgd_master != nullptr, "gemm: EPI_SGD needs sgd_master");
      B200PS_REQUIRE(a.split_k <= 1, "gemm: EPI_SGD cannot split K");
      break;
    default:
      return set_error(B200PS_ERR_INVALID, "gemm: bad epilogue %d", a.epilogue);
  }
  cudaStream_t s = as_stream(stream);
  if (a.ab_dtype == B200PS_DT_BF16) {
    if (a.N > 128) return launch_gemm<false, 256, 4>(a, s);
    if (a.N > 64) return launch_gemm<false, 128, 6>(a, s);
    return launch_gemm<false, 64, 8>(a, s);
  } else {
    if (a.N > 128) return launch_gemm<true, 256, 4>(a, s);
    if (a.N > 64) return launch_gemm<true, 128, 6>(a, s);
    return launch_gemm<true, 64, 8>(a, s);
  }
}
