// EXPERIMENTAL — NOT part of libhstd_b200.so unless built with --experimental, and NOT yet run on a GPU.
// Attention backward with SIXTEEN softmax warps (four per scheduler instead of two): the control-thread
// schedule, barriers, TMEM map and smem layout are those of attn_bwd_kernel in ../attention_sm100.cu; only the
// softmax side differs.  Warp w owns TMEM lane quarter (w & 3) and the 32-key quarter (w >> 2) of the
// 128-key tile, and walks it in two 16-column tcgen05.ld steps so that the kernel fits the 120 registers
// that 544 threads per CTA leave.  Motivation (tools/protocol_sim.py): a 40 % shorter softmax phase takes
// the modelled CTA from 19.3 k to 14.9 k cycles; the softmax warps of the production kernel run at ~45 %
// issue utilisation with two warps per scheduler.
#include <string>
#include <type_traits>

#include "../attention_common.cuh"
#include "../tmap.cuh"

namespace hstd {

constexpr int W16_SOFTMAX_WARPS = 16;
constexpr int W16_THREADS = (W16_SOFTMAX_WARPS + 1) * 32;
constexpr int W16_ST = W16_SOFTMAX_WARPS * 32;

template <int NQ>
__global__ void __launch_bounds__(W16_THREADS, 1)
attn_bwd_w16_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                const __grid_constant__ CUtensorMap tm_v, const __grid_constant__ CUtensorMap tm_do,
                const __grid_constant__ CUtensorMap tm_dqp, const __grid_constant__ CUtensorMap tm_dk,
                const __grid_constant__ CUtensorMap tm_dv, const AttnParams p) {
  HSTD_PDL_PROLOGUE();
  if (threadIdx.x == 0) HSTD_ATRACE(24);
  constexpr uint32_t TMEM_COLS = 512;
  constexpr uint32_t C_S = 0, C_DP = 128, C_DV = 256, C_DK = 320, C_DQ = 384;  // dQ: two buffers of 64 columns
  constexpr uint32_t IDESC_S = make_idesc_bf16(TILE, TILE, false, false);   // Q·Kᵀ, dO·Vᵀ
  constexpr uint32_t IDESC_T = make_idesc_bf16(TILE, DH, true, true);       // Pdᵀ·dO, dSᵀ·Q
  constexpr uint32_t IDESC_Q = make_idesc_bf16(TILE, DH, false, true);      // dS·K
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sK = smem;
  uint8_t* sV = sK + TILE_BYTES;
  uint8_t* sQ = sV + TILE_BYTES;        // 2 buffers
  uint8_t* sDO = sQ + 2 * TILE_BYTES;   // 2 buffers
  uint8_t* sPd = sDO + 2 * TILE_BYTES;  // 32 KiB
  uint8_t* sDS = sPd + P_BYTES;         // 32 KiB
  uint8_t* sDQ = sDS + P_BYTES;         // 2 x 16 KiB staging of the dQ partial for its TMA store
  uint32_t* sBits = reinterpret_cast<uint32_t*>(sDQ + 2 * TILE_BYTES);  // [4]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sBits + 4);
  uint64_t* bar_kv = bars;
  uint64_t* bar_q = bars + 1;       // [2]
  uint64_t* bar_sdp = bars + 3;
  uint64_t* bar_pds = bars + 4;
  uint64_t* bar_dq = bars + 5;
  uint64_t* bar_dqread = bars + 6;  // dQ of this iteration has left TMEM and sits in its staging tile
  uint64_t* bar_out = bars + 7;     // dK / dV staged in smem
  uint64_t* bar_free = bars + 8;    // [2] every MMA reading Q/dO[buf] (and Pd, dS) of a query tile has completed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 10);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int kt = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int S = p.S;
  const int d = p.H * DH;

  // One synchronisation in the prologue.  The control warp looks at the key mask of this tile on
  // its own (4 loads in flight per lane), so that barriers, TMEM and — above all — the first TMA
  // loads are under way while the softmax warps are still balloting the mask bits.
  bool any_key = true;
  if (warp == W16_SOFTMAX_WARPS) {
    const int32_t* km = p.key_mask + int64_t(b) * S + kt * TILE + lane;
    const int32_t k0 = km[0], k1 = km[32], k2 = km[64], k3 = km[96];
    if (lane == 0) {
      tma_prefetch_desc(&tm_q); tma_prefetch_desc(&tm_k); tma_prefetch_desc(&tm_v); tma_prefetch_desc(&tm_do);
      mbar_init(bar_kv, 1); mbar_init(&bar_q[0], 1); mbar_init(&bar_q[1], 1);
      mbar_init(bar_sdp, 1); mbar_init(bar_pds, W16_ST); mbar_init(bar_dq, 1); mbar_init(bar_dqread, W16_ST);
      mbar_init(bar_out, W16_ST); mbar_init(&bar_free[0], 1); mbar_init(&bar_free[1], 1);
      tma_prefetch_desc(&tm_dqp); tma_prefetch_desc(&tm_dk); tma_prefetch_desc(&tm_dv);
      fence_mbar_init();
    }
    any_key = __any_sync(0xffffffffu, (k0 | k1 | k2 | k3) != 0);
    if (any_key) {
      if (elect_one()) {
        mbar_expect_tx(bar_kv, 2 * TILE_BYTES);
        tma_load_4d(sK, &tm_k, bar_kv, 0, kt * TILE, h, b);
        tma_load_4d(sV, &tm_v, bar_kv, 0, kt * TILE, h, b);
        for (int i = 0; i < 2 && i < NQ; ++i) {
          mbar_expect_tx(&bar_q[i], 2 * TILE_BYTES);
          tma_load_4d(sQ + i * TILE_BYTES, &tm_q, &bar_q[i], 0, i * TILE, h, b);
          tma_load_4d(sDO + i * TILE_BYTES, &tm_do, &bar_q[i], 0, i * TILE, h, b);
        }
      }
      __syncwarp();
      tmem_alloc(tmem_slot, TMEM_COLS);
      tmem_relinquish();
    }
  } else {
    if (warp < 4) {  // four 32-key words
      const uint32_t bb = __ballot_sync(0xffffffffu, p.key_mask[int64_t(b) * S + kt * TILE + warp * 32 + lane] != 0);
      if (lane == 0) sBits[warp] = bb;
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if ((sBits[0] | sBits[1] | sBits[2] | sBits[3]) == 0) {
    // every key of this tile is padding: its dK, dV and its dQ contributions are exactly zero
    // (the control warp reached the same verdict and has neither loaded nor allocated anything)
    const uint4 z = make_uint4(0, 0, 0, 0);
    for (int i = threadIdx.x; i < TILE * 8 * 2; i += W16_THREADS) {  // dK, dV: 128 rows x 128 B each
      const int which = i / (TILE * 8), r = (i / 8) % TILE, c = i % 8;
      *reinterpret_cast<uint4*>(p.dqkv + (int64_t(b) * S + kt * TILE + r) * (3 * int64_t(d)) + (which + 1) * int64_t(d) +
                                h * DH + c * 8) = z;
    }
    for (int i = threadIdx.x; i < S * 8; i += W16_THREADS) {
      const int r = i / 8, c = i % 8;
      *reinterpret_cast<uint4*>(p.dq_part + (int64_t(kt) * p.B * S + int64_t(b) * S + r) * d + h * DH + c * 8) = z;
    }
    return;
  }
  const uint32_t tmem = *tmem_slot;

  if (warp == W16_SOFTMAX_WARPS) {
    if (elect_one()) {
      mbar_wait(bar_kv, 0);
      // descriptors once; per-MMA variation is a constant added to the 16-byte-unit address field
      const uint64_t dK_k = make_smem_desc_sw128(smem_u32(sK), 16, 1024);           // K as K-major B (Q·Kᵀ)
      const uint64_t dK_mn = make_smem_desc_sw128(smem_u32(sK), TILE_BYTES, 1024);  // K as MN-major B (dS·K)
      const uint64_t dV_k = make_smem_desc_sw128(smem_u32(sV), 16, 1024);           // V as K-major B (dO·Vᵀ)
      const uint64_t dPd_mn = make_smem_desc_sw128(smem_u32(sPd), TILE_BYTES, 1024);
      const uint64_t dDS_mn = make_smem_desc_sw128(smem_u32(sDS), TILE_BYTES, 1024);
      const uint64_t dDS_k = make_smem_desc_sw128(smem_u32(sDS), 16, 1024);
      const uint64_t dQ_k0 = make_smem_desc_sw128(smem_u32(sQ), 16, 1024);
      const uint64_t dQ_mn0 = make_smem_desc_sw128(smem_u32(sQ), TILE_BYTES, 1024);
      const uint64_t dDO_k0 = make_smem_desc_sw128(smem_u32(sDO), 16, 1024);
      const uint64_t dDO_mn0 = make_smem_desc_sw128(smem_u32(sDO), TILE_BYTES, 1024);
      // The tensor pipe executes in issue order, so nothing here waits for an MMA group to finish
      // before issuing the next one.  Per query tile i the order is
      //     dQ_i  |  S, dP of tile i+1  |  dV, dK of tile i
      // so the softmax warps get their next scores after ~1200 cycles while dV/dK (which only feed
      // TMEM accumulators) run underneath the next softmax.  dQ is committed on its own barrier, the
      // Pd/dS/Q/dO tiles of tile i are released by bar_free once dV/dK have completed.
      auto issue_scores = [&](int i) {
        const int buf = i & 1;
        const uint64_t boff = uint64_t(buf) * (TILE_BYTES >> 4);
        mbar_wait(&bar_q[buf], (i >> 1) & 1);
        HSTD_ATRACE(32 + 8 * i + 0);  // ctl: Q/dO tile landed
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < DH / 16; ++k) umma_bf16(tmem + C_S, dQ_k0 + boff + k * 2, dK_k + k * 2, IDESC_S, k > 0);
#pragma unroll
        for (int k = 0; k < DH / 16; ++k) umma_bf16(tmem + C_DP, dDO_k0 + boff + k * 2, dV_k + k * 2, IDESC_S, k > 0);
        umma_commit(bar_sdp);
      };
      issue_scores(0);
      for (int i = 0; i < NQ; ++i) {
        const int buf = i & 1;
        const uint64_t boff = uint64_t(buf) * (TILE_BYTES >> 4);
        if (i > 0) {
          // operand tiles of the previous query tile are free once its dV/dK MMAs have completed
          mbar_wait(&bar_free[buf ^ 1], ((i - 1) >> 1) & 1);
          if (i + 1 < NQ) {
            mbar_expect_tx(&bar_q[buf ^ 1], 2 * TILE_BYTES);
            tma_load_4d(sQ + (buf ^ 1) * TILE_BYTES, &tm_q, &bar_q[buf ^ 1], 0, (i + 1) * TILE, h, b);
            tma_load_4d(sDO + (buf ^ 1) * TILE_BYTES, &tm_do, &bar_q[buf ^ 1], 0, (i + 1) * TILE, h, b);
          }
        }
        mbar_wait(bar_pds, i & 1);
        HSTD_ATRACE(32 + 8 * i + 1);  // ctl: Pm / dS' ready
        tc_fence_after();
        if (i + 1 < NQ) {
          issue_scores(i + 1);  // S/dP are free: the softmax warps arrived on bar_pds
          // Measured: MMAs issued behind the score MMAs delay the arrival of their commit by ~800
          // cycles, so everything else is held back until the scores are visible to the softmax warps.
          mbar_wait(bar_sdp, (i + 1) & 1);
        }
        if (i > 0) {
          // dQ partial of the previous query tile: staged by the softmax warps right after their softmax
          // of this tile, stored by TMA from here.  This wait MUST precede the commit of dQ_i below:
          // phase i of bar_dqread needs bar_dq phase i, so the parity wait for phase i-1 cannot be
          // overtaken by two completions (tools/protocol_sim.py reproduces the hang of the other order).
          mbar_wait(bar_dqread, (i - 1) & 1);
          tma_store_4d(&tm_dqp, sDQ + ((i - 1) & 1) * TILE_BYTES, h * DH, (i - 1) * TILE, b, kt);
          tma_store_commit();
          tma_store_wait_read_le1();  // the store of tile i-2 has read the staging tile that dQ_i will use
        }
        // dQ_i = dS · K : reduction over the 128 keys (dS K-major: 4 k-steps per 64-key atom), into
        // the dQ buffer of this parity (the other one is being read out by the softmax warps)
#pragma unroll
        for (int k = 0; k < TILE / 16; ++k)
          umma_bf16(tmem + C_DQ + buf * DH, dDS_k + (k >> 2) * (TILE_BYTES >> 4) + (k & 3) * 2, dK_mn + k * (2048 >> 4),
                    IDESC_Q, k > 0);
        umma_commit(bar_dq);
        // reduction over the 128 query rows of this tile: 8 k-steps of 16 rows = 2048 B each
#pragma unroll
        for (int k = 0; k < TILE / 16; ++k)
          umma_bf16(tmem + C_DV, dPd_mn + k * (2048 >> 4), dDO_mn0 + boff + k * (2048 >> 4), IDESC_T, (i > 0 || k > 0));
#pragma unroll
        for (int k = 0; k < TILE / 16; ++k)
          umma_bf16(tmem + C_DK, dDS_mn + k * (2048 >> 4), dQ_mn0 + boff + k * (2048 >> 4), IDESC_T, (i > 0 || k > 0));
        umma_commit(&bar_free[buf]);
        HSTD_ATRACE(32 + 8 * i + 2);  // ctl: second MMA group issued
      }
      mbar_wait(&bar_free[(NQ - 1) & 1], ((NQ - 1) >> 1) & 1);  // dV / dK complete
      mbar_wait(bar_dqread, (NQ - 1) & 1);
      tma_store_4d(&tm_dqp, sDQ + ((NQ - 1) & 1) * TILE_BYTES, h * DH, (NQ - 1) * TILE, b, kt);
      tma_store_commit();
      // dK / dV of this key tile: staged by the softmax warps in the (now idle) Pd / dS tiles
      mbar_wait(bar_out, 0);
      tma_store_4d(&tm_dk, sDS, 0, kt * TILE, h, b);
      tma_store_4d(&tm_dv, sPd, 0, kt * TILE, h, b);
      tma_store_commit();
      tma_store_wait_read();  // smem is released at CTA exit
      HSTD_ATRACE(26);
    }
  } else {
    const int quarter = warp & 3, colq = warp >> 2;  // TMEM lane quarter, 32-key quarter of the tile
    const uint32_t row = quarter * 32 + lane;
    const uint32_t t_row = tmem + ((uint32_t(quarter) * 32u) << 16);
    const float scale2 = p.scale * LOG2E;
    const Dropout dr = make_dropout(p.st, p.p_drop, p.site);
    const KeepConst kc = make_keep_const(dr);
    // The uniform 1/keep of dropout is factored out of the loop (see attn_bwd_kernel):
    //   dS = ik · (Pm∘dP − (δ/ik)·P) = ik · dS',   dV = ik · Pmᵀ·dO.
    const float ik = dr.active() ? dr.scale : 1.f;
    const float inv_ik = 1.f / ik;
    const uint64_t sc2 = pack2(scale2, scale2);
    // 16 bf16 (eight packed words) of row `row` into a K-major SW128 tile at column col0 (multiple of 16)
    auto store16 = [&](uint8_t* tile, uint32_t col0, const uint32_t (&w)[8]) {
      uint8_t* atom = tile + (col0 >> 6) * TILE_BYTES;
      const uint32_t chunk0 = (col0 & 63u) >> 3;
      *reinterpret_cast<uint4*>(atom + swz(row, chunk0)) = make_uint4(w[0], w[1], w[2], w[3]);
      *reinterpret_cast<uint4*>(atom + swz(row, chunk0 + 1)) = make_uint4(w[4], w[5], w[6], w[7]);
    };
    // 16 fp32 TMEM columns x scale -> staging tile
    auto stage16 = [&](uint32_t taddr, uint8_t* tile, uint32_t col0, float sc) {
      uint32_t v[16];
      tmem_ld_32x16(taddr, v);
      tmem_ld_wait();
      uint32_t w[8];
#pragma unroll
      for (int q = 0; q < 8; ++q) w[q] = pack_bf16x2(__uint_as_float(v[2 * q]) * sc, __uint_as_float(v[2 * q + 1]) * sc);
      store16(tile, col0, w);
    };
    auto stage_dq = [&](int j) {
      mbar_wait(bar_dq, j & 1);
      tc_fence_after();
      stage16(t_row + C_DQ + (j & 1) * DH + colq * 16, sDQ + (j & 1) * TILE_BYTES, colq * 16, p.scale * ik);
      tc_fence_before();
      fence_proxy_async_smem();
      mbar_arrive(bar_dqread);
    };
    const int sb = colq;  // this warp's 32-key sub-block of the tile
    const uint32_t bits = sBits[sb];
    for (int i = 0; i < NQ; ++i) {
      const uint64_t srow = (uint64_t(b) * p.H + h) * S + uint64_t(i) * TILE + row;  // global score row
      const float lse2 = p.lse[srow] * LOG2E;
      const float nd = -p.delta[srow] * inv_ik;
      const uint64_t nl2 = pack2(-lse2, -lse2), nd2 = pack2(nd, nd);
      mbar_wait(bar_sdp, i & 1);
      tc_fence_after();
      if (bits == 0) {  // warp-uniform: every key of this sub-block is padding
        if (i > 0) mbar_wait(&bar_free[(i - 1) & 1], ((i - 1) >> 1) & 1);
        store_row32_zero(sPd, row, sb * 32);
        store_row32_zero(sDS, row, sb * 32);
      } else {
#pragma unroll 1
        for (int g = 0; g < 2; ++g) {
          uint32_t vs[16], vd[16];
          tmem_ld_32x16(t_row + C_S + sb * 32 + g * 16, vs);
          tmem_ld_32x16(t_row + C_DP + sb * 32 + g * 16, vd);
          tmem_ld_wait();
          const uint32_t hbits = (bits >> (16 * g)) & 0xffffu;
          uint32_t pw[8], dw[8];  // Pm and dS' as packed bf16 pairs
          auto body = [&](auto drop_tag, auto full_tag) {
            constexpr bool DROP = decltype(drop_tag)::value, FULL = decltype(full_tag)::value;
            uint32_t fl[4];  // keep flags: MSB of byte b of fl[q] belongs to element 4q + b
            if constexpr (DROP) {
              const uint64_t e0 = srow * S + kt * TILE + sb * 32;
              const uint4 r = dropout_bytes16(dr, (e0 >> 4) + g);
              fl[0] = keep_flags4(r.x, kc);
              fl[1] = keep_flags4(r.y, kc);
              fl[2] = keep_flags4(r.z, kc);
              fl[3] = keep_flags4(r.w, kc);
            }
            if constexpr (FULL) {
#pragma unroll
              for (int j = 0; j < 16; j += 2) {
                float a, b2;
                unpack2(ffma2(pack2(__uint_as_float(vs[j]), __uint_as_float(vs[j + 1])), sc2, nl2), a, b2);
                vs[j] = __float_as_uint(fast_exp2(a));
                vs[j + 1] = __float_as_uint(fast_exp2(b2));
              }
            } else {
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                const float pj = fast_exp2(__uint_as_float(vs[j]) * scale2 - lse2);
                vs[j] = ((hbits >> j) & 1u) ? __float_as_uint(pj) : 0u;
              }
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const int j = 4 * q;
              uint32_t m0 = vs[j], m1 = vs[j + 1], m2 = vs[j + 2], m3 = vs[j + 3];
              if constexpr (DROP) {
                m0 &= prmt_sel<0x8888>(fl[q]);
                m1 &= prmt_sel<0x9999>(fl[q]);
                m2 &= prmt_sel<0xAAAA>(fl[q]);
                m3 &= prmt_sel<0xBBBB>(fl[q]);
              }
              float d0, d1, d2, d3;
              unpack2(ffma2(pack2(__uint_as_float(m0), __uint_as_float(m1)),
                            pack2(__uint_as_float(vd[j]), __uint_as_float(vd[j + 1])),
                            fmul2(pack2(__uint_as_float(vs[j]), __uint_as_float(vs[j + 1])), nd2)), d0, d1);
              unpack2(ffma2(pack2(__uint_as_float(m2), __uint_as_float(m3)),
                            pack2(__uint_as_float(vd[j + 2]), __uint_as_float(vd[j + 3])),
                            fmul2(pack2(__uint_as_float(vs[j + 2]), __uint_as_float(vs[j + 3])), nd2)), d2, d3);
              pw[j >> 1] = pack_bf16x2(__uint_as_float(m0), __uint_as_float(m1));
              pw[(j >> 1) + 1] = pack_bf16x2(__uint_as_float(m2), __uint_as_float(m3));
              dw[j >> 1] = pack_bf16x2(d0, d1);
              dw[(j >> 1) + 1] = pack_bf16x2(d2, d3);
            }
          };
          if (hbits == 0xffffu) {
            if (dr.active()) body(std::true_type{}, std::true_type{});
            else body(std::false_type{}, std::true_type{});
          } else {
            if (dr.active()) body(std::true_type{}, std::false_type{});
            else body(std::false_type{}, std::false_type{});
          }
          // dV/dK of the previous query tile still read Pd / dS: their "free" barrier gates the first overwrite
          if (g == 0 && i > 0) mbar_wait(&bar_free[(i - 1) & 1], ((i - 1) >> 1) & 1);
          store16(sPd, sb * 32 + g * 16, pw);
          store16(sDS, sb * 32 + g * 16, dw);
        }
      }
      tc_fence_before();
      fence_proxy_async_smem();
      mbar_arrive(bar_pds);
      if (i > 0) stage_dq(i - 1);  // dQ of the previous query tile: its MMA ran underneath this softmax
    }
    stage_dq(NQ - 1);
    // dK (x scale) and dV of this key tile (TMEM lane = key row) once the last dV / dK MMAs have completed;
    // the Pd / dS tiles are then free to stage them: 16 of the 64 columns per thread
    mbar_wait(&bar_free[(NQ - 1) & 1], ((NQ - 1) >> 1) & 1);
    tc_fence_after();
    stage16(t_row + C_DK + colq * 16, sDS, colq * 16, p.scale * ik);
    stage16(t_row + C_DV + colq * 16, sPd, colq * 16, ik);
    tc_fence_before();
    fence_proxy_async_smem();
    mbar_arrive(bar_out);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == W16_SOFTMAX_WARPS) {
    tc_fence_after();
    tmem_dealloc(tmem, TMEM_COLS);
  }
}

static size_t bwd_w16_smem() { return 1024 + 8 * size_t(TILE_BYTES) + 2 * P_BYTES + 16 + 256; }

}  // namespace hstd

using namespace hstd;

// Trial entry: runs ONLY the backward kernel.  `workspace` must already hold delta (the first B*H*S floats), e.g.
// from a preceding hstd_attn_bwd call on the same inputs; on return the dQ partials [S/128][B*S][d] sit behind
// delta in the workspace (sum them to get dQ) and dK / dV are in their slots of dqkv.
extern "C" HSTD_API int hstd_attn_bwd_w16_core(const void* qkv, const void* dctx, const float* lse, const int32_t* key_mask,
                                               void* dqkv, void* workspace, int B, int H, int S, float scale, float p_drop,
                                               uint32_t site, const hstd_step_state* st, void* stream) {
  if (S % TILE != 0 || S < TILE || S > 4 * TILE) return set_error(-1, "attn_bwd_w16: S must be 128, 256, 384 or 512");
  const int64_t d = int64_t(H) * DH;
  float* delta = static_cast<float*>(workspace);
  __nv_bfloat16* dq_part = reinterpret_cast<__nv_bfloat16*>(static_cast<uint8_t*>(workspace) + int64_t(B) * H * S * 4);
  const __nv_bfloat16* base = static_cast<const __nv_bfloat16*>(qkv);
  __nv_bfloat16* gbase = static_cast<__nv_bfloat16*>(dqkv);
  CUtensorMap tq, tk, tv, tdo, tdqp, tdk, tdv;
  if (int rc = get_tmap(&tq, base, DH, S, H, B, 3 * d, DH, S * 3 * d, TILE)) return rc;
  if (int rc = get_tmap(&tk, base + d, DH, S, H, B, 3 * d, DH, S * 3 * d, TILE)) return rc;
  if (int rc = get_tmap(&tv, base + 2 * d, DH, S, H, B, 3 * d, DH, S * 3 * d, TILE)) return rc;
  if (int rc = get_tmap(&tdo, dctx, DH, S, H, B, d, DH, S * d, TILE)) return rc;
  if (int rc = get_tmap(&tdqp, dq_part, d, S, B, S / TILE, d, S * d, int64_t(B) * S * d, TILE)) return rc;
  if (int rc = get_tmap(&tdk, gbase + d, DH, S, H, B, 3 * d, DH, S * 3 * d, TILE)) return rc;
  if (int rc = get_tmap(&tdv, gbase + 2 * d, DH, S, H, B, 3 * d, DH, S * 3 * d, TILE)) return rc;
  AttnParams p{};
  p.lse = const_cast<float*>(lse);
  p.delta = delta;
  p.dqkv = gbase;
  p.dq_part = dq_part;
  p.key_mask = key_mask;
  p.st = st; p.p_drop = p_drop; p.site = site;
  p.B = B; p.H = H; p.S = S; p.scale = scale;
  const int nq = S / TILE;
  dim3 grid(nq, H, B);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
#define HSTD_W16(NQ_)                                                                                              \
  case NQ_: {                                                                                                      \
    cudaError_t e = cudaFuncSetAttribute(attn_bwd_w16_kernel<NQ_>, cudaFuncAttributeMaxDynamicSharedMemorySize,    \
                                         int(bwd_w16_smem()));                                                     \
    if (e != cudaSuccess) return set_error(-5, std::string("attn_bwd_w16 smem: ") + cudaGetErrorString(e));        \
    launch_pdl(attn_bwd_w16_kernel<NQ_>, grid, dim3(W16_THREADS), bwd_w16_smem(), s, tq, tk, tv, tdo, tdqp, tdk, tdv, p); \
    break;                                                                                                         \
  }
  switch (nq) {
    HSTD_W16(1) HSTD_W16(2) HSTD_W16(3) HSTD_W16(4)
    default: return set_error(-1, "attn_bwd_w16: unsupported S");
  }
#undef HSTD_W16
  return check_launch("attn_bwd_w16");
}
