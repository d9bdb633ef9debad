// tcgen05 / TMEM GEMM for sm_100a, fed by TMA, persistent and warp-specialised.
//
//   warp 0      : TMA producer (one elected lane) — fills a ring of NUM_STAGES smem stages
//   warp 1      : TMEM allocator + MMA issuer (one elected lane issues tcgen05.mma)
//   warps 2..9  : epilogue — tcgen05.ld the fp32 accumulator out of TMEM, apply
//                 alpha / bias / GELU / residual-add / GELU', convert to bf16, store
//
// The accumulator is double-buffered in TMEM (2 x BLOCK_N columns) so the epilogue of tile
// i overlaps the MMAs of tile i+1.  Operands may be K-major or MN-major ("transposed");
// both use 128-byte-swizzled smem tiles written by TMA and described to the tensor core by
// UMMA shared-memory descriptors, so forward (X·Wᵀ), dgrad (dY·W), wgrad (dYᵀ·X) and the
// batched attention products all run through this one kernel with no explicit transposes.
//
// Replaces the cuBLAS GEMMs implied by model.fit() in the reference
// (scripts/train.py:145; SURVEY.md §2.4 K2/K4/K5/K6 and the K3 matmuls).
#include <mutex>
#include <string>
#include <unordered_map>

#include "../include/hstd_b200.h"
#include "common.cuh"
#include "sm100_ptx.cuh"
#include "tmap.cuh"

namespace hstd {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;  // 64 bf16 = one 128-byte swizzle row
constexpr int UMMA_K = 16;
constexpr int NUM_EPI_WARPS = 8;
constexpr int GEMM_THREADS = (2 + NUM_EPI_WARPS) * 32;

struct GemmKernelParams {
  __nv_bfloat16* C;
  float* Cf32;  // split-K partial output (fp32) when num_splits > 1
  const __nv_bfloat16* bias;
  __nv_bfloat16* aux;
  int64_t ldc, c_bs0, c_bs1, ldaux;
  int M, N, K;
  int nb0, nb1;
  int tiles_m, tiles_n;
  int num_splits, kb_per_split;
  int epilogue;
  float alpha;
  int use_tma_store;  // bf16 output through smem staging + TMA bulk stores (BN >= 128, no split-K)
  long long* trace;   // optional clock64() stamps of CTA 0 (tools/gemm_trace.py); null in production
};
#define HSTD_GTRACE(slot) do { if (p.trace && blockIdx.x == 0 && (slot) < 512) p.trace[slot] = clock64(); } while (0)

// Per-CTA smem: with MC == 2 (cta_group::2) each CTA of the pair stages only half of the B tile.
template <int BN, int MC>
struct GemmCfg {
  static constexpr uint32_t A_BYTES = BLOCK_M * BLOCK_K * 2;
  static constexpr uint32_t B_BYTES = (BN / MC) * BLOCK_K * 2;
  static constexpr int STAGES = (192 * 1024) / int(A_BYTES + B_BYTES) > 8 ? 8 : (192 * 1024) / int(A_BYTES + B_BYTES);
  static constexpr uint32_t STAGING_BYTES = NUM_EPI_WARPS * 4096;  // per warp: 32 rows x 128 B for the TMA store
  static constexpr size_t SMEM = 1024 /*align slack*/ + size_t(STAGES) * (A_BYTES + B_BYTES) + STAGING_BYTES + 256;
};

// MC = CTAs per tile along M.  MC == 2 is the 2-SM form (cta_group::2): a cluster of two CTAs
// computes one 256 x BN tile with a single tcgen05.mma issued by the leader; each CTA stages its
// own 128 rows of A and half of the B tile, so the per-SM shared-memory traffic (TMA writes +
// tensor-core reads) drops from 192 to ~125 B/cycle — below the 128 B/cycle port limit that
// caps the 1-SM form at ~60 % of the tensor peak.
template <int BN, bool A_MN, bool B_MN, int MC>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                 const __grid_constant__ CUtensorMap tmap_c, const __grid_constant__ CUtensorMap tmap_aux,
                 const GemmKernelParams p) {
  constexpr int STAGES = GemmCfg<BN, MC>::STAGES;
  constexpr uint32_t A_BYTES = GemmCfg<BN, MC>::A_BYTES;
  constexpr uint32_t B_BYTES = GemmCfg<BN, MC>::B_BYTES;   // per CTA
  constexpr int BN_LOCAL = BN / MC;                        // B rows staged by this CTA
  // two accumulator buffers; TMEM allocations are powers of two, so the 192-wide tile sits in 2 x 256 columns
  constexpr uint32_t ACC_STRIDE = BN == 192 ? 256 : BN;
  constexpr uint32_t TMEM_COLS = 2 * ACC_STRIDE < 32 ? 32 : 2 * ACC_STRIDE;
  constexpr uint32_t IDESC = make_idesc_bf16(BLOCK_M * MC, BN, A_MN, B_MN);
  constexpr uint16_t PAIR_MASK = (1u << MC) - 1;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + STAGES * A_BYTES;
  uint8_t* smem_stage = smem_b + STAGES * B_BYTES;  // epilogue store staging, 4 KiB per warp
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem_stage + GemmCfg<BN, MC>::STAGING_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full = empty_bar + STAGES;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  if (threadIdx.x == 0) HSTD_GTRACE(0);
  const int cta_rank = MC > 1 ? int(cluster_ctarank()) : 0;
  const bool leader = cta_rank == 0;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    if (p.use_tma_store) tma_prefetch_desc(&tmap_c);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);   // MC == 2: only the leader's copy is used
      mbar_init(&empty_bar[s], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], NUM_EPI_WARPS * MC);  // MC == 2: the leader's copy collects both CTAs' epilogues
    }
    fence_mbar_init();
  }
  if (warp == 1) {
    if constexpr (MC == 1) {
      tmem_alloc(tmem_slot, TMEM_COLS);
      tmem_relinquish();
    } else {
      tmem_alloc_2sm(tmem_slot, TMEM_COLS);
      tmem_relinquish_2sm();
    }
  }
  tc_fence_before();
  __syncthreads();
  if constexpr (MC > 1) cluster_sync_all();  // peer barriers are initialised before anything remote lands
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // PDL: everything above (barrier init, TMEM allocation, descriptor prefetch) overlapped the
  // previous kernel's tail; from here on we read and write global memory.
  pdl_launch_dependents();
  pdl_wait();

  const int tiles_mp = (p.tiles_m + MC - 1) / MC;  // M tiles are handed out in groups of MC
  const int tiles_per_batch = tiles_mp * p.tiles_n * p.num_splits;
  const int total_tiles = tiles_per_batch * p.nb0 * p.nb1;
  const int num_kb_total = (p.K + BLOCK_K - 1) / BLOCK_K;
  const int first_tile = blockIdx.x / MC, tile_stride = gridDim.x / MC;

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer (every CTA)
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = first_tile; tile < total_tiles; tile += tile_stride) {
        const int z = tile / tiles_per_batch;
        int r = tile - z * tiles_per_batch;
        const int split = r / (tiles_mp * p.tiles_n);
        r -= split * (tiles_mp * p.tiles_n);
        const int tmp = r / p.tiles_n, tn = r - tmp * p.tiles_n;
        const int tm = tmp * MC + cta_rank;  // may be one past the last M tile: TMA zero-fills, nothing is stored
        const int z0 = z % p.nb0, z1 = z / p.nb0;
        const int m0 = tm * BLOCK_M, n0 = tn * BN + cta_rank * BN_LOCAL;
        const int kb_begin = split * p.kb_per_split;
        const int kb_end = min(num_kb_total, kb_begin + p.kb_per_split);
        for (int kb = kb_begin; kb < kb_end; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          HSTD_GTRACE(2 + 4 * (kb - kb_begin) + 0);  // producer: slot free
          uint8_t* sa = smem_a + stage * A_BYTES;
          uint8_t* sb = smem_b + stage * B_BYTES;
          const int k0 = kb * BLOCK_K;
          if constexpr (MC == 1) {
            mbar_expect_tx(&full_bar[stage], A_BYTES + B_BYTES);
            if constexpr (!A_MN) {
              tma_load_4d(sa, &tmap_a, &full_bar[stage], k0, m0, z0, z1);
            } else {
#pragma unroll
              for (int j = 0; j < BLOCK_M / 64; ++j)
                tma_load_4d(sa + j * (64 * BLOCK_K * 2), &tmap_a, &full_bar[stage], m0 + 64 * j, k0, z0, z1);
            }
            if constexpr (!B_MN) {
              tma_load_4d(sb, &tmap_b, &full_bar[stage], k0, n0, z0, z1);
            } else {
#pragma unroll
              for (int j = 0; j < BN_LOCAL / 64; ++j)
                tma_load_4d(sb + j * (64 * BLOCK_K * 2), &tmap_b, &full_bar[stage], n0 + 64 * j, k0, z0, z1);
            }
          } else {
            // both CTAs' bytes complete on the leader's barrier
            if (leader) mbar_expect_tx(&full_bar[stage], MC * (A_BYTES + B_BYTES));
            const uint32_t bar = mapa_shared(smem_u32(&full_bar[stage]), 0);
            if constexpr (!A_MN) {
              tma_load_4d_2sm(sa, &tmap_a, bar, k0, m0, z0, z1);
            } else {
#pragma unroll
              for (int j = 0; j < BLOCK_M / 64; ++j)
                tma_load_4d_2sm(sa + j * (64 * BLOCK_K * 2), &tmap_a, bar, m0 + 64 * j, k0, z0, z1);
            }
            if constexpr (!B_MN) {
              tma_load_4d_2sm(sb, &tmap_b, bar, k0, n0, z0, z1);
            } else {
#pragma unroll
              for (int j = 0; j < BN_LOCAL / 64; ++j)
                tma_load_4d_2sm(sb + j * (64 * BLOCK_K * 2), &tmap_b, bar, n0 + 64 * j, k0, z0, z1);
            }
          }
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer (leader CTA only)
    if (leader && elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      const uint64_t desc_a0 = A_MN ? make_smem_desc_sw128(smem_u32(smem_a), 64 * BLOCK_K * 2, 1024)
                                    : make_smem_desc_sw128(smem_u32(smem_a), 16, 1024);
      const uint64_t desc_b0 = B_MN ? make_smem_desc_sw128(smem_u32(smem_b), 64 * BLOCK_K * 2, 1024)
                                    : make_smem_desc_sw128(smem_u32(smem_b), 16, 1024);
      for (int tile = first_tile; tile < total_tiles; tile += tile_stride) {
        int r = tile % tiles_per_batch;
        const int split = r / (tiles_mp * p.tiles_n);
        const int kb_begin = split * p.kb_per_split;
        const int kb_end = min(num_kb_total, kb_begin + p.kb_per_split);
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * ACC_STRIDE;
        for (int kb = kb_begin; kb < kb_end; ++kb) {
          HSTD_GTRACE(2 + 4 * (kb - kb_begin) + 1);  // mma: starts waiting for data
          mbar_wait(&full_bar[stage], phase);
          HSTD_GTRACE(2 + 4 * (kb - kb_begin) + 2);  // mma: data landed
          tc_fence_after();
          // descriptors: stage base + constant per-k-step increments (in 16-byte units); building
          // them from scratch costs ~50 dependent instructions per MMA on the single issuing thread
          const uint64_t da0 = desc_a0 + uint64_t(stage) * (A_BYTES >> 4);
          const uint64_t db0 = desc_b0 + uint64_t(stage) * (B_BYTES >> 4);
#pragma unroll
          for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
            // K-major: advance 16 elements (32 B) inside the swizzled row.
            // MN-major: advance 16 K-rows = two 8-row groups of 1024 B.
            const uint64_t da = da0 + uint64_t(k * (A_MN ? 2048 / 16 : 32 / 16));
            const uint64_t db = db0 + uint64_t(k * (B_MN ? 2048 / 16 : 32 / 16));
            if constexpr (MC == 1) umma_bf16(d_tmem, da, db, IDESC, (kb > kb_begin || k > 0) ? 1u : 0u);
            else umma_bf16_2sm(d_tmem, da, db, IDESC, (kb > kb_begin || k > 0) ? 1u : 0u);
          }
          HSTD_GTRACE(2 + 4 * (kb - kb_begin) + 3);  // mma: k-block issued
          // smem stage free once these MMAs retire (in both CTAs of a pair)
          if constexpr (MC == 1) umma_commit(&empty_bar[stage]);
          else umma_commit_2sm(&empty_bar[stage], PAIR_MASK);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        // accumulator complete: wake the epilogue warps (of both CTAs)
        if constexpr (MC == 1) umma_commit(&tmem_full[acc]);
        else umma_commit_2sm(&tmem_full[acc], PAIR_MASK);
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
  } else {
    // ------------------------------------------------------------ epilogue warps
    const int ew = warp - 2;             // 0..7
    const int quarter = warp & 3;        // TMEM lane quarter this warp may read
    const int col_half = ew >> 2;        // which half of the BN columns
    // 64-column store tiles: the 192-wide tile splits 128 + 64 between the two warps of a lane quarter
    constexpr int COLS_H0 = BN == 192 ? 128 : BN / 2;
    const int col_begin = col_half * COLS_H0;
    const int ncols = col_half ? BN - COLS_H0 : COLS_H0;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = first_tile; tile < total_tiles; tile += tile_stride) {
      const int z = tile / tiles_per_batch;
      int r = tile - z * tiles_per_batch;
      const int split = r / (tiles_mp * p.tiles_n);
      r -= split * (tiles_mp * p.tiles_n);
      const int tmp = r / p.tiles_n, tn = r - tmp * p.tiles_n;
      const int tm = tmp * MC + cta_rank;
      const int z0 = z % p.nb0, z1 = z / p.nb0;
      const int row = tm * BLOCK_M + quarter * 32 + lane;
      const int col_base = tn * BN + col_begin;

      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const uint32_t t_row = tmem_base + acc * ACC_STRIDE + col_begin + (uint32_t(quarter * 32) << 16);
      const int64_t zoff = int64_t(z0) * p.c_bs0 + int64_t(z1) * p.c_bs1;
      if constexpr (BN >= 128) {
        if (p.use_tma_store) {
          // Coalesced path: 32 rows x 64 columns at a time go through a 128B-swizzled smem tile
          // and leave with one TMA bulk store (full 128-byte lines; rows/cols past M/N are clipped
          // by the tensor map).  Direct per-lane stores touch 32 rows x 16 B per instruction and
          // ran at a third of the HBM write bandwidth.
          uint8_t* stage = smem_stage + ew * 4096;
          const int row0 = tm * BLOCK_M + quarter * 32;
#pragma unroll 1
          for (int g = 0; g < ncols; g += 64) {
            const int col0 = col_base + g;
            float f[64];
            {
              uint32_t v0[32], v1[32];
              tmem_ld_32x32(t_row + g, v0);
              tmem_ld_32x32(t_row + g + 32, v1);
              tmem_ld_wait();
#pragma unroll
              for (int j = 0; j < 32; ++j) {
                f[j] = __uint_as_float(v0[j]) * p.alpha;
                f[32 + j] = __uint_as_float(v1[j]) * p.alpha;
              }
            }
            const bool in_range = row < p.M && col0 < p.N;  // (N is a multiple of 8; finer column clipping is the TMA's)
            if (p.bias != nullptr) {
#pragma unroll
              for (int j = 0; j < 64; j += 8) {
                if (col0 + j < p.N) {
                  const uint4 b = __ldg(reinterpret_cast<const uint4*>(p.bias + col0 + j));
                  const float2 b0 = unpack_bf16x2(b.x), b1 = unpack_bf16x2(b.y), b2 = unpack_bf16x2(b.z),
                               b3 = unpack_bf16x2(b.w);
                  f[j] += b0.x; f[j + 1] += b0.y; f[j + 2] += b1.x; f[j + 3] += b1.y;
                  f[j + 4] += b2.x; f[j + 5] += b2.y; f[j + 6] += b3.x; f[j + 7] += b3.y;
                }
              }
            }
            auto flush = [&](const CUtensorMap* tm_out) {
              // (elect.sync picks the same lane for the same member mask, so the thread that waits
              // on the bulk group is the one that committed it)
              if (elect_one()) tma_store_wait_read();  // previous store has drained this staging tile
              __syncwarp();
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const uint4 o = make_uint4(pack_bf16x2(f[8 * j], f[8 * j + 1]), pack_bf16x2(f[8 * j + 2], f[8 * j + 3]),
                                           pack_bf16x2(f[8 * j + 4], f[8 * j + 5]), pack_bf16x2(f[8 * j + 6], f[8 * j + 7]));
                *reinterpret_cast<uint4*>(stage + lane * 128 + ((j ^ (lane & 7)) << 4)) = o;
              }
              fence_proxy_async_smem();
              __syncwarp();
              if (elect_one()) {
                tma_store_4d(tm_out, stage, col0, row0, z0, z1);
                tma_store_commit();
              }
            };
            if (p.epilogue == HSTD_EPI_GELU) {
              if (p.aux) flush(&tmap_aux);  // pre-activation
#pragma unroll
              for (int j = 0; j < 64; ++j) f[j] = gelu_erf(f[j]);
            } else if (p.epilogue == HSTD_EPI_ADD || p.epilogue == HSTD_EPI_GELU_BWD) {
              if (in_range) {
                const __nv_bfloat16* auxrow = p.aux + zoff + int64_t(row) * p.ldaux + col0;
#pragma unroll
                for (int j = 0; j < 64; j += 8) {
                  if (col0 + j < p.N) {
                    const uint4 a = *reinterpret_cast<const uint4*>(auxrow + j);
                    const float2 a0 = unpack_bf16x2(a.x), a1 = unpack_bf16x2(a.y), a2 = unpack_bf16x2(a.z),
                                 a3 = unpack_bf16x2(a.w);
                    if (p.epilogue == HSTD_EPI_ADD) {
                      f[j] += a0.x; f[j + 1] += a0.y; f[j + 2] += a1.x; f[j + 3] += a1.y;
                      f[j + 4] += a2.x; f[j + 5] += a2.y; f[j + 6] += a3.x; f[j + 7] += a3.y;
                    } else {
                      f[j] *= gelu_erf_grad(a0.x); f[j + 1] *= gelu_erf_grad(a0.y);
                      f[j + 2] *= gelu_erf_grad(a1.x); f[j + 3] *= gelu_erf_grad(a1.y);
                      f[j + 4] *= gelu_erf_grad(a2.x); f[j + 5] *= gelu_erf_grad(a2.y);
                      f[j + 6] *= gelu_erf_grad(a3.x); f[j + 7] *= gelu_erf_grad(a3.y);
                    }
                  }
                }
              }
            }
            flush(&tmap_c);
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) {
            if (MC == 1 || leader) mbar_arrive(&tmem_empty[acc]);
            else mbar_arrive_cluster(mapa_shared(smem_u32(&tmem_empty[acc]), 0));
          }
          if (++acc == 2) {
            acc = 0;
            acc_phase ^= 1;
          }
          continue;
        }
      }
#pragma unroll 1
      for (int c = 0; c < ncols; c += 32) {
        uint32_t v[32];
        tmem_ld_32x32(t_row + c, v);
        tmem_ld_wait();
        const int col0 = col_base + c;
        if (p.epilogue >= 100) {  // timing experiments only: 100 = no global stores
          if (v[0] == 0x7fc12345u) p.C[0] = __float2bfloat16(1.f);
          continue;
        }
        if (row < p.M && col0 < p.N) {
          if (p.num_splits > 1) {
            float* dst = p.Cf32 + (int64_t(split) * p.M + row) * p.N + col0;
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              if (col0 + j < p.N) {
                float4 o = make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]), __uint_as_float(v[j + 2]),
                                       __uint_as_float(v[j + 3]));
                *reinterpret_cast<float4*>(dst + j) = o;
              }
            }
          } else {
            float f[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]) * p.alpha;
            if (p.bias != nullptr) {
#pragma unroll
              for (int j = 0; j < 32; j += 8) {
                if (col0 + j < p.N) {
                  const uint4 b = __ldg(reinterpret_cast<const uint4*>(p.bias + col0 + j));
                  const float2 b0 = unpack_bf16x2(b.x), b1 = unpack_bf16x2(b.y), b2 = unpack_bf16x2(b.z),
                               b3 = unpack_bf16x2(b.w);
                  f[j] += b0.x; f[j + 1] += b0.y; f[j + 2] += b1.x; f[j + 3] += b1.y;
                  f[j + 4] += b2.x; f[j + 5] += b2.y; f[j + 6] += b3.x; f[j + 7] += b3.y;
                }
              }
            }
            __nv_bfloat16* crow = p.C + zoff + int64_t(row) * p.ldc + col0;
            __nv_bfloat16* auxrow = p.aux ? p.aux + zoff + int64_t(row) * p.ldaux + col0 : nullptr;
            if (p.epilogue == HSTD_EPI_GELU) {
              if (auxrow) {
#pragma unroll
                for (int j = 0; j < 32; j += 8) {
                  if (col0 + j < p.N) {
                    uint4 o = make_uint4(pack_bf16x2(f[j], f[j + 1]), pack_bf16x2(f[j + 2], f[j + 3]),
                                         pack_bf16x2(f[j + 4], f[j + 5]), pack_bf16x2(f[j + 6], f[j + 7]));
                    *reinterpret_cast<uint4*>(auxrow + j) = o;
                  }
                }
              }
#pragma unroll
              for (int j = 0; j < 32; ++j) f[j] = gelu_erf(f[j]);
            } else if (p.epilogue == HSTD_EPI_ADD) {
#pragma unroll
              for (int j = 0; j < 32; j += 8) {
                if (col0 + j < p.N) {
                  const uint4 a = *reinterpret_cast<const uint4*>(auxrow + j);
                  const float2 a0 = unpack_bf16x2(a.x), a1 = unpack_bf16x2(a.y), a2 = unpack_bf16x2(a.z),
                               a3 = unpack_bf16x2(a.w);
                  f[j] += a0.x; f[j + 1] += a0.y; f[j + 2] += a1.x; f[j + 3] += a1.y;
                  f[j + 4] += a2.x; f[j + 5] += a2.y; f[j + 6] += a3.x; f[j + 7] += a3.y;
                }
              }
            } else if (p.epilogue == HSTD_EPI_GELU_BWD) {
#pragma unroll
              for (int j = 0; j < 32; j += 8) {
                if (col0 + j < p.N) {
                  const uint4 a = *reinterpret_cast<const uint4*>(auxrow + j);
                  const float2 a0 = unpack_bf16x2(a.x), a1 = unpack_bf16x2(a.y), a2 = unpack_bf16x2(a.z),
                               a3 = unpack_bf16x2(a.w);
                  f[j] *= gelu_erf_grad(a0.x); f[j + 1] *= gelu_erf_grad(a0.y);
                  f[j + 2] *= gelu_erf_grad(a1.x); f[j + 3] *= gelu_erf_grad(a1.y);
                  f[j + 4] *= gelu_erf_grad(a2.x); f[j + 5] *= gelu_erf_grad(a2.y);
                  f[j + 6] *= gelu_erf_grad(a3.x); f[j + 7] *= gelu_erf_grad(a3.y);
                }
              }
            }
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
              if (col0 + j < p.N) {
                uint4 o = make_uint4(pack_bf16x2(f[j], f[j + 1]), pack_bf16x2(f[j + 2], f[j + 3]),
                                     pack_bf16x2(f[j + 4], f[j + 5]), pack_bf16x2(f[j + 6], f[j + 7]));
                *reinterpret_cast<uint4*>(crow + j) = o;
              }
            }
          }
        }
      }
      // All TMEM reads of this accumulator buffer are done (wait::ld above): hand it back to
      // the MMA issuer (which lives in the leader CTA).
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (MC == 1 || leader) mbar_arrive(&tmem_empty[acc]);
        else mbar_arrive_cluster(mapa_shared(smem_u32(&tmem_empty[acc]), 0));
      }
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
  }

  if (warp >= 2) {
    __syncwarp();
    if (elect_one()) tma_store_wait_all();  // this thread's bulk stores have landed
  }
  tc_fence_before();
  __syncthreads();
  if constexpr (MC > 1) cluster_sync_all();  // no CTA leaves while its peer may still touch its smem / barriers / TMEM
  if (warp == 1) {
    tc_fence_after();
    if constexpr (MC == 1) tmem_dealloc(tmem_base, TMEM_COLS);
    else tmem_dealloc_2sm(tmem_base, TMEM_COLS);
  }
}

// fp32 split-K partials -> bf16
__global__ void splitk_reduce_kernel(const float* __restrict__ ws, __nv_bfloat16* __restrict__ C, int64_t ldc, int M,
                                     int N, int splits, float alpha) {
  HSTD_PDL_PROLOGUE();
  const int64_t n4 = int64_t(M) * N / 4;
  for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < n4; i += int64_t(gridDim.x) * blockDim.x) {
    float4 s = reinterpret_cast<const float4*>(ws)[i];
    for (int k = 1; k < splits; ++k) {
      const float4 t = reinterpret_cast<const float4*>(ws)[i + k * n4];
      s.x += t.x; s.y += t.y; s.z += t.z; s.w += t.w;
    }
    const int64_t e = i * 4;
    const int row = int(e / N), col = int(e - int64_t(row) * N);
    uint2 o = make_uint2(pack_bf16x2(s.x * alpha, s.y * alpha), pack_bf16x2(s.z * alpha, s.w * alpha));
    *reinterpret_cast<uint2*>(C + int64_t(row) * ldc + col) = o;
  }
}

// ------------------------------------------------------------------ host side
int num_sms() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

template <int BN, bool A_MN, bool B_MN, int MC>
static int launch_gemm_mc(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc, const CUtensorMap& tx,
                          const GemmKernelParams& p, int max_ctas, cudaStream_t stream) {
  auto kern = gemm_bf16_kernel<BN, A_MN, B_MN, MC>;
  constexpr size_t smem = GemmCfg<BN, MC>::SMEM;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
    if (e != cudaSuccess) return set_error(-5, std::string("cudaFuncSetAttribute: ") + cudaGetErrorString(e));
    configured = true;
  }
  const int tiles_mp = (p.tiles_m + MC - 1) / MC;
  const int total = tiles_mp * p.tiles_n * p.num_splits * p.nb0 * p.nb1;  // cluster-tiles
  int clusters = num_sms() / MC;
  if (max_ctas > 0 && max_ctas / MC < clusters) clusters = max_ctas / MC > 0 ? max_ctas / MC : 1;
  if (total < clusters) clusters = total;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(clusters * MC);
  cfg.blockDim = dim3(GEMM_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = MC;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, ta, tb, tc, tx, p);
  if (e != cudaSuccess) return set_error(-6, std::string("gemm launch: ") + cudaGetErrorString(e));
  return 0;
}

template <int BN, bool A_MN, bool B_MN>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc, const CUtensorMap& tx,
                       const GemmKernelParams& p, int max_ctas, int mc, cudaStream_t stream) {
  if (mc == 2) return launch_gemm_mc<BN, A_MN, B_MN, 2>(ta, tb, tc, tx, p, max_ctas, stream);
  return launch_gemm_mc<BN, A_MN, B_MN, 1>(ta, tb, tc, tx, p, max_ctas, stream);
}

static long long* g_gemm_trace = nullptr;

static int pick_block_n(int M, int N, int batches, int splits) {
  // 256-wide tiles halve the B-operand smem traffic per FLOP; use them when they still
  // give every SM a tile, otherwise fall back to 128 (64 for the narrow attention outputs).
  if (N <= 64) return 64;
  const int tm = (M + BLOCK_M - 1) / BLOCK_M;
  if (N >= 256 && tm * ((N + 255) / 256) * batches * splits >= num_sms()) return 256;
  return 128;
}

static int gemm_impl(const hstd_gemm_args* a, int splits, float* workspace, cudaStream_t stream) {
  if (a->M <= 0 || a->N <= 0 || a->K <= 0) return set_error(-1, "gemm: empty problem");
  if (a->N % 8 != 0) return set_error(-1, "gemm: N must be a multiple of 8");
  const int nb0 = a->nb0 > 0 ? a->nb0 : 1, nb1 = a->nb1 > 0 ? a->nb1 : 1;
  int bn = a->block_n ? a->block_n : pick_block_n(a->M, a->N, nb0 * nb1, splits);
  if (bn != 64 && bn != 128 && bn != 192 && bn != 256) return set_error(-1, "gemm: block_n must be 64, 128, 192 or 256");

  CUtensorMap ta, tb;
  int rc;
  if (!a->a_mn_major)
    rc = get_tmap(&ta, a->A, a->K, a->M, nb0, nb1, a->lda, a->a_bs0, a->a_bs1, BLOCK_M);
  else
    rc = get_tmap(&ta, a->A, a->M, a->K, nb0, nb1, a->lda, a->a_bs0, a->a_bs1, 64);
  if (rc) return rc;
  // 2-SM MMA (256 x BN tile per CTA pair) whenever there are at least two M tiles; the
  // 64-wide B tile cannot be split between the pair.
  int mc = (a->M > BLOCK_M && bn >= 128) ? 2 : 1;
  if (a->cluster_m == 1) mc = 1;
  if (bn == 192 && a->b_mn_major) mc = 1;  // half of an MN-major 192-row B tile is not a whole 64-row chunk
  if (!a->b_mn_major)
    rc = get_tmap(&tb, a->B, a->K, a->N, nb0, nb1, a->ldb, a->b_bs0, a->b_bs1, bn / mc);
  else
    rc = get_tmap(&tb, a->B, a->N, a->K, nb0, nb1, a->ldb, a->b_bs0, a->b_bs1, 64);
  if (rc) return rc;

  GemmKernelParams p;
  p.C = static_cast<__nv_bfloat16*>(a->C);
  p.Cf32 = workspace;
  p.bias = static_cast<const __nv_bfloat16*>(a->bias);
  p.aux = static_cast<__nv_bfloat16*>(a->aux);
  p.ldc = a->ldc; p.c_bs0 = a->c_bs0; p.c_bs1 = a->c_bs1; p.ldaux = a->ldaux;
  p.M = a->M; p.N = a->N; p.K = a->K; p.nb0 = nb0; p.nb1 = nb1;
  p.tiles_m = (a->M + BLOCK_M - 1) / BLOCK_M;
  p.tiles_n = (a->N + bn - 1) / bn;
  const int num_kb = (a->K + BLOCK_K - 1) / BLOCK_K;
  if (splits > num_kb) splits = num_kb;
  p.num_splits = splits;
  p.kb_per_split = (num_kb + splits - 1) / splits;
  p.epilogue = a->epilogue;
  p.alpha = a->alpha;
  p.trace = g_gemm_trace;
  if ((a->epilogue == HSTD_EPI_ADD || a->epilogue == HSTD_EPI_GELU_BWD) && a->aux == nullptr)
    return set_error(-1, "gemm: this epilogue needs aux");
  // Output tensor maps (TMA bulk stores of 64-column x 32-row tiles).
  CUtensorMap tc = ta, tx = ta;  // placeholders when unused
  p.use_tma_store = (bn >= 128 && splits == 1 && a->epilogue < 100 && (a->ldc % 8) == 0 &&
                     (reinterpret_cast<uintptr_t>(a->C) & 15) == 0) ? 1 : 0;
  if (p.use_tma_store) {
    rc = get_tmap(&tc, a->C, a->N, a->M, nb0, nb1, a->ldc, a->c_bs0, a->c_bs1, 32);
    if (rc) return rc;
    if (a->epilogue == HSTD_EPI_GELU && a->aux != nullptr) {
      rc = get_tmap(&tx, a->aux, a->N, a->M, nb0, nb1, a->ldaux, a->c_bs0, a->c_bs1, 32);
      if (rc) return rc;
    }
  }

#define HSTD_GEMM_CASE(BN_)                                                                         \
  if (bn == BN_) {                                                                                  \
    if (!a->a_mn_major && !a->b_mn_major) return launch_gemm<BN_, false, false>(ta, tb, tc, tx, p, a->max_ctas, mc, stream); \
    if (!a->a_mn_major && a->b_mn_major) return launch_gemm<BN_, false, true>(ta, tb, tc, tx, p, a->max_ctas, mc, stream);   \
    if (a->a_mn_major && !a->b_mn_major) return launch_gemm<BN_, true, false>(ta, tb, tc, tx, p, a->max_ctas, mc, stream);   \
    return launch_gemm<BN_, true, true>(ta, tb, tc, tx, p, a->max_ctas, mc, stream);                            \
  }
  HSTD_GEMM_CASE(64)
  HSTD_GEMM_CASE(128)
  HSTD_GEMM_CASE(192)
  HSTD_GEMM_CASE(256)
#undef HSTD_GEMM_CASE
  return set_error(-1, "gemm: unreachable");
}

}  // namespace hstd

// Debug aid: device buffer of >= 512 int64 that CTA 0 of the following GEMM launches fills with
// clock64() stamps of its producer / MMA threads per k-block (NULL switches it off).
extern "C" HSTD_API void hstd_gemm_set_trace(void* device_buffer) {
  hstd::g_gemm_trace = static_cast<long long*>(device_buffer);
}

extern "C" HSTD_API int hstd_gemm_bf16(const hstd_gemm_args* args, void* stream) {
  return hstd::gemm_impl(args, 1, nullptr, static_cast<cudaStream_t>(stream));
}

extern "C" HSTD_API int64_t hstd_gemm_splitk_workspace_bytes(int M, int N, int splits) {
  return int64_t(M) * N * splits * 4;
}

extern "C" HSTD_API int hstd_gemm_bf16_splitk(const hstd_gemm_args* args, int splits, void* workspace,
                                     int64_t workspace_bytes, void* stream) {
  using namespace hstd;
  if (splits <= 1) return gemm_impl(args, 1, nullptr, static_cast<cudaStream_t>(stream));
  if (args->nb0 > 1 || args->nb1 > 1) return set_error(-1, "splitk: batched problems are not supported");
  if (args->epilogue != HSTD_EPI_NONE || args->bias) return set_error(-1, "splitk: only EPI_NONE without bias");
  if (args->N % 4 != 0) return set_error(-1, "splitk: N must be a multiple of 4");
  const int num_kb = (args->K + BLOCK_K - 1) / BLOCK_K;
  if (splits > num_kb) splits = num_kb;
  // kb_per_split rounding may leave trailing splits empty; shrink so every split has work.
  const int per = (num_kb + splits - 1) / splits;
  splits = (num_kb + per - 1) / per;
  if (splits <= 1) return gemm_impl(args, 1, nullptr, static_cast<cudaStream_t>(stream));
  if (workspace == nullptr || workspace_bytes < hstd_gemm_splitk_workspace_bytes(args->M, args->N, splits))
    return set_error(-1, "splitk: workspace too small");
  int rc = gemm_impl(args, splits, static_cast<float*>(workspace), static_cast<cudaStream_t>(stream));
  if (rc) return rc;
  const int64_t n4 = int64_t(args->M) * args->N / 4;
  int blocks = int((n4 + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  launch_pdl(splitk_reduce_kernel, dim3(blocks), dim3(256), 0, static_cast<cudaStream_t>(stream),
             static_cast<const float*>(workspace), static_cast<__nv_bfloat16*>(args->C), args->ldc, args->M, args->N, splits,
             args->alpha);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_error(-6, std::string("splitk reduce launch: ") + cudaGetErrorString(e));
  return 0;
}
