// Fused multi-head attention for sm_100a: softmax(QKᵀ/√dh + key mask) (+ dropout) · V,
// forward and backward, on tcgen05 tensor cores with S / dP / O / dQ / dK / dV accumulators
// in TMEM and all operands TMA-loaded straight out of the fused [B*S, 3*d] QKV buffer.
// The [B,H,S,S] probability tensor is never written to memory (the reference's TF graph
// materialises it twice: model.fit at scripts/train.py:145).
//
// head_dim = 64, S ∈ {128, 256, 384, 512}.  One CTA = 9 warps: warps 0-7 are the softmax /
// epilogue warps (two threads per TMEM lane = query row, each owning half of the columns),
// one lane of warp 8 issues every TMA and tcgen05.mma.  Key chunks that are entirely padding
// are skipped (they contribute exactly zero).
//
//  forward  (CTA = 128 query rows of one (batch, head)):
//     S_c = Q·K_cᵀ for all key chunks c (TMEM 128 columns each) → exact row max over all
//     keys → P_c = exp(S_c·scale − max) (·dropout) written as bf16 into a 128B-swizzled smem
//     tile → O += P_c·V_c (V read MN-major) → O / rowsum → ctx ; log-sum-exp saved.
//  backward (CTA = 128 keys of one (batch, head), loop over query tiles i):
//     S = Q_i·Kᵀ, dP = dO_i·Vᵀ → P = exp(S·scale − lse), dS = P∘(dropout'(dP) − δ) →
//     dV += Pdᵀ·dO_i, dK += dSᵀ·Q_i (accumulated in TMEM over i), dQ_i = dS·K written as a
//     per-key-tile partial; a small kernel sums the partials.  The same smem image of a tile
//     serves as K-major A operand and as MN-major B operand (only the descriptors differ).
//     MMA issue order per query tile: scores of tile i+1 | dQ_i (two TMEM buffers, read out under
//     the softmax of tile i+1) | dV, dK; commit barriers release operand tiles, the control thread
//     never waits for an MMA to finish.  tools/protocol_sim.py models this barrier protocol.
//  Softmax loops are written for the pipes that bound them (MUFU ex2 and the half-rate integer
//  ALU): packed FFMA2/FADD2/FMUL2, dropout keep flags four bytes at a time applied by AND, 1/keep
//  folded into the epilogue scales, one straight-line block per (dropout, full sub-block) variant.
//  Every output (O, dQ partials, dK, dV) is staged in swizzled smem and stored by TMA.
#include <string>
#include <type_traits>

#include "common.cuh"
#include "device_utils.cuh"
#include "sm100_ptx.cuh"
#include "tmap.cuh"
#include "attention_common.cuh"

namespace hstd {

// ============================================================================ forward
template <int NC>
__global__ void __launch_bounds__(AT_THREADS, 1)
attn_fwd_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                const __grid_constant__ CUtensorMap tm_v, const __grid_constant__ CUtensorMap tm_ctx, const AttnParams p) {
  pdl_launch_dependents();
  if (threadIdx.x == 0) HSTD_ATRACE(0);
  constexpr uint32_t TMEM_COLS = NC == 1 ? 128 : (NC == 2 ? 256 : 512);
  constexpr uint32_t IDESC_S = make_idesc_bf16(TILE, TILE, false, false);
  constexpr uint32_t IDESC_O = make_idesc_bf16(TILE, DH, false, true);
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + TILE_BYTES;
  uint8_t* sV = sK + NC * TILE_BYTES;
  uint8_t* sP = sV + NC * TILE_BYTES;  // 2 buffers
  float* sRed = reinterpret_cast<float*>(sP + 2 * P_BYTES);  // [2][128] exchange between column halves
  uint32_t* sBits = reinterpret_cast<uint32_t*>(sRed + 2 * TILE);  // [NC*4] key validity, 32 keys per word
  uint64_t* bars = reinterpret_cast<uint64_t*>(sBits + 16);
  uint64_t* bar_qk = bars;
  uint64_t* bar_v = bars + 1;
  uint64_t* bar_o = bars + 2;
  uint64_t* bar_s = bars + 3;            // [NC]
  uint64_t* bar_p = bar_s + NC;          // [2]
  uint64_t* bar_pfree = bar_p + 2;       // [2]
  uint64_t* bar_k2 = bar_pfree + 2;      // K chunks 1..nc-1
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_k2 + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int qt = blockIdx.x, h = blockIdx.y, b = blockIdx.z;

  if (warp == AT_SOFTMAX_WARPS) {
    if (lane == 0) {
      tma_prefetch_desc(&tm_q); tma_prefetch_desc(&tm_k); tma_prefetch_desc(&tm_v); tma_prefetch_desc(&tm_ctx);
      mbar_init(bar_qk, 1); mbar_init(bar_v, 1); mbar_init(bar_o, 1); mbar_init(bar_k2, 1);
      for (int c = 0; c < NC; ++c) mbar_init(&bar_s[c], 1);
      for (int i = 0; i < 2; ++i) { mbar_init(&bar_p[i], AT_ST); mbar_init(&bar_pfree[i], 1); }
      fence_mbar_init();
    }
    __syncwarp();
    pdl_wait();  // the QKV projection has completed
    // Q and the first key chunk are needed whatever the mask says: their loads start before the
    // TMEM allocation and the mask scan instead of after them
    if (elect_one()) {
      mbar_expect_tx(bar_qk, 2 * TILE_BYTES);
      tma_load_4d(sQ, &tm_q, bar_qk, 0, qt * TILE, h, b);
      tma_load_4d(sK, &tm_k, bar_qk, 0, 0, h, b);
    }
    __syncwarp();
    tmem_alloc(tmem_slot, TMEM_COLS);
    tmem_relinquish();
  } else {
    pdl_wait();
    build_mask_bits(sBits, p.key_mask + int64_t(b) * p.S, NC * 4, warp, lane);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  if (threadIdx.x == 0) HSTD_ATRACE(1);
  // key chunks after the last valid key contribute exactly zero: skip them everywhere
  int nc = 1;
#pragma unroll
  for (int w = 0; w < NC * 4; ++w)
    if (sBits[w] != 0) nc = (w >> 2) + 1;

  if (warp == AT_SOFTMAX_WARPS) {
    if (elect_one()) {
      if (nc > 1) {
        mbar_expect_tx(bar_k2, (nc - 1) * TILE_BYTES);
        for (int c = 1; c < nc; ++c) tma_load_4d(sK + c * TILE_BYTES, &tm_k, bar_k2, 0, c * TILE, h, b);
      }
      mbar_expect_tx(bar_v, nc * TILE_BYTES);
      for (int c = 0; c < nc; ++c) tma_load_4d(sV + c * TILE_BYTES, &tm_v, bar_v, 0, c * TILE, h, b);
      // descriptors once; per-MMA variation is a constant added to the 16-byte-unit address field
      // (rebuilding them per instruction cost the single issuing thread ~100 cycles per MMA)
      const uint64_t dq0 = make_smem_desc_sw128(smem_u32(sQ), 16, 1024);
      const uint64_t dk0 = make_smem_desc_sw128(smem_u32(sK), 16, 1024);
      const uint64_t dv0 = make_smem_desc_sw128(smem_u32(sV), TILE_BYTES, 1024);
      const uint64_t dp0 = make_smem_desc_sw128(smem_u32(sP), 16, 1024);
      HSTD_ATRACE(16);
      mbar_wait(bar_qk, 0);
      HSTD_ATRACE(17);
      tc_fence_after();
      for (int c = 0; c < nc; ++c) {
        if (c == 1) {
          mbar_wait(bar_k2, 0);
          tc_fence_after();
        }
        const uint64_t dk = dk0 + uint64_t(c) * (TILE_BYTES >> 4);
#pragma unroll
        for (int k = 0; k < DH / 16; ++k) umma_bf16(tmem + c * TILE, dq0 + k * 2, dk + k * 2, IDESC_S, k > 0);
        umma_commit(&bar_s[c]);
      }
      HSTD_ATRACE(18);
      mbar_wait(bar_v, 0);
      HSTD_ATRACE(19);
      for (int c = 0; c < nc; ++c) {
        const int buf = c & 1;
        mbar_wait(&bar_p[buf], (c >> 1) & 1);
        HSTD_ATRACE(20 + c);
        tc_fence_after();
        const uint64_t dp = dp0 + uint64_t(buf) * (P_BYTES >> 4);
        const uint64_t dv = dv0 + uint64_t(c) * (TILE_BYTES >> 4);
#pragma unroll
        for (int k = 0; k < TILE / 16; ++k)
          umma_bf16(tmem, dp + (k >> 2) * (TILE_BYTES >> 4) + (k & 3) * 2, dv + k * (2048 >> 4), IDESC_O, (c > 0 || k > 0));
        umma_commit(&bar_pfree[buf]);
      }
      umma_commit(bar_o);
    }
  } else {
    const int quarter = warp & 3, half = warp >> 2;
    const uint32_t row = quarter * 32 + lane;               // TMEM lane = query row of the tile
    const uint32_t t_row = tmem + ((uint32_t(quarter) * 32u) << 16);
    const float scale2 = p.scale * LOG2E;
    // TMEM reads are software-pipelined: the tcgen05.ld of the next 32-column sub-block is in flight
    // while the current one is processed (a tcgen05.ld round trip is ~300 cycles and the two warps of
    // a scheduler run in lockstep behind the same barriers, so it was fully exposed).
    uint32_t va[32], vb[32];
    // ---- pass 1: exact row maximum over every valid key (this thread: its column half)
    float mx = -INFINITY, mx1 = -INFINITY;
    auto max_block = [&](const uint32_t (&v)[32], uint32_t bits) {
      if (bits == 0) return;  // warp-uniform
      if (bits == 0xffffffffu) {
        // two independent chains of 3-input maxima (FMNMX3): 16 instructions for 32 columns
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          mx = fmaxf(fmaxf(mx, __uint_as_float(v[j])), __uint_as_float(v[j + 1]));
          mx1 = fmaxf(fmaxf(mx1, __uint_as_float(v[j + 2])), __uint_as_float(v[j + 3]));
        }
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) mx = fmaxf(mx, ((bits >> j) & 1u) ? __uint_as_float(v[j]) : -INFINITY);
      }
    };
    auto load_block = [&](int c, int sbi, uint32_t (&v)[32]) {
      const int sb = half * 2 + sbi;
      if (sBits[c * 4 + sb] != 0) tmem_ld_32x32(t_row + c * TILE + sb * 32, v);
    };
    mbar_wait(&bar_s[0], 0);
    if (threadIdx.x == 0) HSTD_ATRACE(2);
    tc_fence_after();
    load_block(0, 0, va);
    tmem_ld_wait();
    tmem_ld_fence(va);
    for (int c = 0; c < nc; ++c) {
      load_block(c, 1, vb);
      max_block(va, sBits[c * 4 + half * 2]);
      tmem_ld_wait();
      tmem_ld_fence(vb);
      if (c + 1 < nc) {
        mbar_wait(&bar_s[c + 1], 0);
        if (threadIdx.x == 0) HSTD_ATRACE(3 + c);
        tc_fence_after();
        load_block(c + 1, 0, va);
      }
      max_block(vb, sBits[c * 4 + half * 2 + 1]);
      tmem_ld_wait();
      tmem_ld_fence(va);
    }
    mx = fmaxf(mx, mx1);
    if (threadIdx.x == 0) HSTD_ATRACE(6);
    load_block(0, 0, va);  // first sub-block of pass 2: in flight across the exchange of the maxima
    sRed[half * TILE + row] = mx;
    softmax_warps_sync();
    mx = fmaxf(mx, sRed[(half ^ 1) * TILE + row]);
    const float m2 = (mx == -INFINITY) ? 0.f : mx * scale2;
    // ---- pass 2: probabilities -> smem (bf16), running row sum.  The loop is issue-bound, so it is
    // written for instruction count: packed FFMA2 / FADD2, probabilities packed to bf16 first and
    // the dropout mask applied to the packed words with AND (the 1/keep factor is uniform and
    // moves to the epilogue), keep flags for four elements from three integer instructions.
    const Dropout dr = make_dropout(p.st, p.p_drop, p.site);
    const KeepConst kc = make_keep_const(dr);
    const uint64_t row_base = (uint64_t(b) * p.H + h) * p.S + uint64_t(qt) * TILE + row;  // global score row
    const uint64_t sc2 = pack2(scale2, scale2), nm2 = pack2(-m2, -m2);
    uint64_t l2 = pack2(0.f, 0.f);
    auto prob_block = [&](const uint32_t (&v)[32], int c, int sbi) {
      const int sb = half * 2 + sbi;
      const uint32_t bits = sBits[c * 4 + sb];
      uint8_t* pt = sP + (c & 1) * P_BYTES;
      if (bits == 0) {
        store_row32_zero(pt, row, sb * 32);
        return;
      }
      uint32_t pw[16];  // 32 probabilities as packed bf16 pairs
      // Dropout masks (integer work on the half-rate ALU pipe) and exponentials (MUFU) sit in ONE
      // straight-line block per variant, so that ptxas interleaves the two pipes instead of running
      // them back to back: masks first as 16 words of all-ones / all-zeros bf16 lanes, then exp + AND.
      auto body = [&](auto drop_tag, auto full_tag) {
        constexpr bool DROP = decltype(drop_tag)::value, FULL = decltype(full_tag)::value;
        uint32_t mk[16];
        if constexpr (DROP) {
          const uint64_t e0 = row_base * p.S + c * TILE + sb * 32;
#pragma unroll
          for (int g = 0; g < 2; ++g) {
            const uint4 r = dropout_bytes16(dr, (e0 >> 4) + g);
            const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const uint32_t f = keep_flags4(w[i], kc);
              mk[8 * g + 2 * i] = prmt_sel<0x9988>(f);      // elements 4i, 4i+1
              mk[8 * g + 2 * i + 1] = prmt_sel<0xBBAA>(f);  // elements 4i+2, 4i+3
            }
          }
        }
        if constexpr (FULL) {
#pragma unroll
          for (int j = 0; j < 32; j += 2) {
            float a, b2;
            unpack2(ffma2(pack2(__uint_as_float(v[j]), __uint_as_float(v[j + 1])), sc2, nm2), a, b2);
            a = fast_exp2(a);
            b2 = fast_exp2(b2);
            l2 = fadd2(l2, pack2(a, b2));
            pw[j >> 1] = pack_bf16x2(a, b2);
            if constexpr (DROP) pw[j >> 1] &= mk[j >> 1];
          }
        } else {  // the sub-block that straddles the end of the sequence
          float l1 = 0.f;
#pragma unroll
          for (int j = 0; j < 32; j += 2) {
            float a = fast_exp2(__uint_as_float(v[j]) * scale2 - m2), b2 = fast_exp2(__uint_as_float(v[j + 1]) * scale2 - m2);
            a = ((bits >> j) & 1u) ? a : 0.f;
            b2 = ((bits >> (j + 1)) & 1u) ? b2 : 0.f;
            l1 += a + b2;
            pw[j >> 1] = pack_bf16x2(a, b2);
            if constexpr (DROP) pw[j >> 1] &= mk[j >> 1];
          }
          l2 = fadd2(l2, pack2(l1, 0.f));
        }
      };
      if (bits == 0xffffffffu) {
        if (dr.active()) body(std::true_type{}, std::true_type{});
        else body(std::false_type{}, std::true_type{});
      } else {
        if (dr.active()) body(std::true_type{}, std::false_type{});
        else body(std::false_type{}, std::false_type{});
      }
      uint8_t* atom = pt + ((sb * 32) >> 6) * TILE_BYTES;
      const uint32_t chunk0 = ((sb * 32) & 63u) >> 3;
#pragma unroll
      for (int q4 = 0; q4 < 4; ++q4)
        *reinterpret_cast<uint4*>(atom + swz(row, chunk0 + q4)) =
            make_uint4(pw[4 * q4], pw[4 * q4 + 1], pw[4 * q4 + 2], pw[4 * q4 + 3]);
    };
    tmem_ld_wait();
    tmem_ld_fence(va);
    for (int c = 0; c < nc; ++c) {
      const int buf = c & 1;
      if (c >= 2) mbar_wait(&bar_pfree[buf], ((c - 2) >> 1) & 1);
      load_block(c, 1, vb);
      prob_block(va, c, 0);
      tmem_ld_wait();
      tmem_ld_fence(vb);
      if (c + 1 < nc) load_block(c + 1, 0, va);
      prob_block(vb, c, 1);
      tc_fence_before();
      fence_proxy_async_smem();
      mbar_arrive(&bar_p[buf]);
      if (threadIdx.x == 0) HSTD_ATRACE(8 + c);
      tmem_ld_wait();
      tmem_ld_fence(va);
    }
    float l, l_hi;
    unpack2(l2, l, l_hi);
    l += l_hi;
    softmax_warps_sync();  // everyone has read the partner maxima: sRed can be reused
    sRed[half * TILE + row] = l;
    softmax_warps_sync();
    l += sRed[(half ^ 1) * TILE + row];
    // ---- epilogue: O / l -> ctx (this thread: 32 of the 64 output columns), log-sum-exp
    if (threadIdx.x == 0) HSTD_ATRACE(12);
    mbar_wait(bar_o, 0);
    if (threadIdx.x == 0) HSTD_ATRACE(13);
    tc_fence_after();
    const float inv_l = (l > 0.f ? 1.f / l : 0.f) * (dr.active() ? dr.scale : 1.f);  // dropout's 1/keep lands here
    // O leaves through a swizzled smem tile (the first P buffer: every MMA that read it has completed)
    // and ONE TMA bulk store per CTA; per-thread row stores touched 32 different lines per instruction.
    {
      uint32_t v[32];
      tmem_ld_32x32(t_row + half * 32, v);
      tmem_ld_wait();
      float f[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]) * inv_l;
      store_row32(sP, row, half * 32, f);
    }
    fence_proxy_async_smem();
    softmax_warps_sync();
    if (warp == 0 && elect_one()) {
      tma_store_4d(&tm_ctx, sP, 0, qt * TILE, h, b);
      tma_store_commit();
      tma_store_wait_read();  // the tile may not be released (CTA exit) before the engine has read it
    }
    if (half == 0 && p.lse) p.lse[row_base] = (mx == -INFINITY ? 0.f : mx * p.scale) + __logf(fmaxf(l, 1e-30f));
    if (threadIdx.x == 0) HSTD_ATRACE(14);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == AT_SOFTMAX_WARPS) {
    tc_fence_after();
    tmem_dealloc(tmem, TMEM_COLS);
  }
}

// ============================================================================ backward
// delta[b,h,q] = sum_j dO[b,q,h,j] * O[b,q,h,j]; one warp per (b,q) row of width H*64.
__global__ void __launch_bounds__(256) attn_delta_kernel(const __nv_bfloat16* __restrict__ dO,
                                                         const __nv_bfloat16* __restrict__ O, float* __restrict__ delta,
                                                         int B, int S, int H) {
  HSTD_PDL_PROLOGUE();
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= B * S) return;
  const int b = warp / S, q = warp - b * S;
  const int d = H * DH;
  for (int col = lane * 8; col < d; col += 256) {
    float x[8], y[8];
    load8(dO + int64_t(warp) * d + col, x);
    load8(O + int64_t(warp) * d + col, y);
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += x[j] * y[j];
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    s += __shfl_xor_sync(0xffffffffu, s, 4);
    if ((lane & 7) == 0) delta[(int64_t(b) * H + col / DH) * S + q] = s;
  }
}

template <int NQ>
__global__ void __launch_bounds__(AT_THREADS, 1)
attn_bwd_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
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
  if (warp == AT_SOFTMAX_WARPS) {
    const int32_t* km = p.key_mask + int64_t(b) * S + kt * TILE + lane;
    const int32_t k0 = km[0], k1 = km[32], k2 = km[64], k3 = km[96];
    if (lane == 0) {
      tma_prefetch_desc(&tm_q); tma_prefetch_desc(&tm_k); tma_prefetch_desc(&tm_v); tma_prefetch_desc(&tm_do);
      mbar_init(bar_kv, 1); mbar_init(&bar_q[0], 1); mbar_init(&bar_q[1], 1);
      mbar_init(bar_sdp, 1); mbar_init(bar_pds, AT_ST); mbar_init(bar_dq, 1); mbar_init(bar_dqread, AT_ST);
      mbar_init(bar_out, AT_ST); mbar_init(&bar_free[0], 1); mbar_init(&bar_free[1], 1);
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
    build_mask_bits(sBits, p.key_mask + int64_t(b) * S + kt * TILE, 4, warp, lane);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if ((sBits[0] | sBits[1] | sBits[2] | sBits[3]) == 0) {
    // every key of this tile is padding: its dK, dV and its dQ contributions are exactly zero
    // (the control warp reached the same verdict and has neither loaded nor allocated anything)
    const uint4 z = make_uint4(0, 0, 0, 0);
    for (int i = threadIdx.x; i < TILE * 8 * 2; i += AT_THREADS) {  // dK, dV: 128 rows x 128 B each
      const int which = i / (TILE * 8), r = (i / 8) % TILE, c = i % 8;
      *reinterpret_cast<uint4*>(p.dqkv + (int64_t(b) * S + kt * TILE + r) * (3 * int64_t(d)) + (which + 1) * int64_t(d) +
                                h * DH + c * 8) = z;
    }
    for (int i = threadIdx.x; i < S * 8; i += AT_THREADS) {
      const int r = i / 8, c = i % 8;
      *reinterpret_cast<uint4*>(p.dq_part + (int64_t(kt) * p.B * S + int64_t(b) * S + r) * d + h * DH + c * 8) = z;
    }
    return;
  }
  const uint32_t tmem = *tmem_slot;

  if (warp == AT_SOFTMAX_WARPS) {
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
    const int quarter = warp & 3, half = warp >> 2;
    const uint32_t row = quarter * 32 + lane;
    const uint32_t t_row = tmem + ((uint32_t(quarter) * 32u) << 16);
    const float scale2 = p.scale * LOG2E;
    const Dropout dr = make_dropout(p.st, p.p_drop, p.site);
    const KeepConst kc = make_keep_const(dr);
    // The uniform 1/keep of dropout is factored out of the loop: with Pm = mask∘P (unscaled),
    //   dS = ik · (Pm∘dP − (δ/ik)·P) = ik · dS',   dV = ik · Pmᵀ·dO,
    // so Pm and dS' go to the tensor cores and ik is applied when dV / dK / dQ leave TMEM.
    const float ik = dr.active() ? dr.scale : 1.f;
    const float inv_ik = 1.f / ik;
    const uint64_t sc2 = pack2(scale2, scale2);
    auto stage_dq = [&](int j) {
      mbar_wait(bar_dq, j & 1);
      if (threadIdx.x == 0) HSTD_ATRACE(32 + 8 * j + 6);  // dQ visible
      tc_fence_after();
      uint32_t v[32];
      tmem_ld_32x32(t_row + C_DQ + (j & 1) * DH + half * 32, v);
      tmem_ld_wait();
      float f[32];
      const float sc = p.scale * ik;
#pragma unroll
      for (int q = 0; q < 32; ++q) f[q] = __uint_as_float(v[q]) * sc;
      store_row32(sDQ + (j & 1) * TILE_BYTES, row, half * 32, f);
      tc_fence_before();
      fence_proxy_async_smem();
      mbar_arrive(bar_dqread);
      if (threadIdx.x == 0) HSTD_ATRACE(32 + 8 * j + 7);  // dQ staged
    };
    for (int i = 0; i < NQ; ++i) {
      const uint64_t srow = (uint64_t(b) * p.H + h) * S + uint64_t(i) * TILE + row;  // global score row
      const float lse2 = p.lse[srow] * LOG2E;
      const float nd = -p.delta[srow] * inv_ik;
      const uint64_t nl2 = pack2(-lse2, -lse2), nd2 = pack2(nd, nd);
      mbar_wait(bar_sdp, i & 1);
      if (threadIdx.x == 0) HSTD_ATRACE(32 + 8 * i + 4);  // S / dP visible
      tc_fence_after();
#pragma unroll 1
      for (int sbi = 0; sbi < 2; ++sbi) {
        const int sb = half * 2 + sbi;
        const uint32_t bits = sBits[sb];
        if (bits == 0) {
          if (sbi == 0 && i > 0) mbar_wait(&bar_free[(i - 1) & 1], ((i - 1) >> 1) & 1);
          store_row32_zero(sPd, row, sb * 32);
          store_row32_zero(sDS, row, sb * 32);
          continue;
        }
        uint32_t vs[32], vd[32];
        tmem_ld_32x32(t_row + C_S + sb * 32, vs);
        tmem_ld_32x32(t_row + C_DP + sb * 32, vd);
        tmem_ld_wait();
        uint32_t pw[16], dw[16];  // Pm and dS' as packed bf16 pairs
        // One straight-line block per variant: Philox + keep flags (integer, half-rate ALU pipe)
        // are issued first so ptxas can interleave them with the exponentials (MUFU pipe).
        auto body = [&](auto drop_tag, auto full_tag) {
          constexpr bool DROP = decltype(drop_tag)::value, FULL = decltype(full_tag)::value;
          uint32_t fl[8];  // keep flags: MSB of byte b of fl[q] belongs to element 4q + b
          if constexpr (DROP) {
            const uint64_t e0 = srow * S + kt * TILE + sb * 32;
#pragma unroll
            for (int g = 0; g < 2; ++g) {
              const uint4 r = dropout_bytes16(dr, (e0 >> 4) + g);
              fl[4 * g] = keep_flags4(r.x, kc);
              fl[4 * g + 1] = keep_flags4(r.y, kc);
              fl[4 * g + 2] = keep_flags4(r.z, kc);
              fl[4 * g + 3] = keep_flags4(r.w, kc);
            }
          }
          // P in place of S
          if constexpr (FULL) {
#pragma unroll
            for (int j = 0; j < 32; j += 2) {
              float a, b2;
              unpack2(ffma2(pack2(__uint_as_float(vs[j]), __uint_as_float(vs[j + 1])), sc2, nl2), a, b2);
              vs[j] = __float_as_uint(fast_exp2(a));
              vs[j + 1] = __float_as_uint(fast_exp2(b2));
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const float pj = fast_exp2(__uint_as_float(vs[j]) * scale2 - lse2);
              vs[j] = ((bits >> j) & 1u) ? __float_as_uint(pj) : 0u;
            }
          }
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            const int j = 4 * q;
            uint32_t m0 = vs[j], m1 = vs[j + 1], m2 = vs[j + 2], m3 = vs[j + 3];
            if constexpr (DROP) {
              // Pm = P AND (all-ones | zero)
              m0 &= prmt_sel<0x8888>(fl[q]);
              m1 &= prmt_sel<0x9999>(fl[q]);
              m2 &= prmt_sel<0xAAAA>(fl[q]);
              m3 &= prmt_sel<0xBBBB>(fl[q]);
            }
            // dS' = Pm*dP + P*nd, two elements per instruction
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
        if (bits == 0xffffffffu) {
          if (dr.active()) body(std::true_type{}, std::true_type{});
          else body(std::false_type{}, std::true_type{});
        } else {
          if (dr.active()) body(std::true_type{}, std::false_type{});
          else body(std::false_type{}, std::false_type{});
        }
        // dV/dK of the previous query tile run underneath this softmax and still read Pd / dS: its
        // "free" barrier (practically always complete by now) gates the first overwrite
        if (sbi == 0 && i > 0) mbar_wait(&bar_free[(i - 1) & 1], ((i - 1) >> 1) & 1);
        {
          const uint32_t aoff = ((sb * 32) >> 6) * TILE_BYTES;
          const uint32_t chunk0 = ((sb * 32) & 63u) >> 3;
#pragma unroll
          for (int q4 = 0; q4 < 4; ++q4) {
            const uint32_t o = aoff + swz(row, chunk0 + q4);
            *reinterpret_cast<uint4*>(sPd + o) = make_uint4(pw[4 * q4], pw[4 * q4 + 1], pw[4 * q4 + 2], pw[4 * q4 + 3]);
            *reinterpret_cast<uint4*>(sDS + o) = make_uint4(dw[4 * q4], dw[4 * q4 + 1], dw[4 * q4 + 2], dw[4 * q4 + 3]);
          }
        }
      }
      tc_fence_before();
      fence_proxy_async_smem();
      mbar_arrive(bar_pds);
      if (threadIdx.x == 0) HSTD_ATRACE(32 + 8 * i + 5);  // Pm / dS' written
      // dQ partial of the PREVIOUS query tile (its MMA ran underneath the softmax above, into the dQ
      // buffer of its parity): 32 of the 64 columns per thread into the swizzled staging tile; the
      // control thread stores it with one TMA bulk store (per-thread row stores touched 32 different
      // lines per instruction and cost ~900 cycles per query tile)
      if (i > 0) stage_dq(i - 1);
    }
    stage_dq(NQ - 1);
    // dK (x scale) and dV of this key tile (TMEM lane = key row), once the last dV / dK MMAs have
    // completed; the Pd / dS tiles are then free to stage them
    mbar_wait(&bar_free[(NQ - 1) & 1], ((NQ - 1) >> 1) & 1);
    tc_fence_after();
    {
      uint32_t v[32];
      float f[32];
      tmem_ld_32x32(t_row + C_DK + half * 32, v);
      tmem_ld_wait();
      const float sk = p.scale * ik;
#pragma unroll
      for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]) * sk;
      store_row32(sDS, row, half * 32, f);
      tmem_ld_32x32(t_row + C_DV + half * 32, v);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]) * ik;
      store_row32(sPd, row, half * 32, f);
    }
    tc_fence_before();
    fence_proxy_async_smem();
    mbar_arrive(bar_out);
    if (threadIdx.x == 0) HSTD_ATRACE(25);  // dK / dV staged
  }
  tc_fence_before();
  __syncthreads();
  if (warp == AT_SOFTMAX_WARPS) {
    tc_fence_after();
    tmem_dealloc(tmem, TMEM_COLS);
  }
}

// dqkv[:, 0:d] = sum over key tiles of dq_part
__global__ void __launch_bounds__(256) attn_dq_reduce_kernel(const __nv_bfloat16* __restrict__ part,
                                                             __nv_bfloat16* __restrict__ dqkv, int64_t M, int d,
                                                             int nparts) {
  HSTD_PDL_PROLOGUE();
  const int64_t nvec = M * d / 8;
  for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < nvec; i += int64_t(gridDim.x) * blockDim.x) {
    float acc[8];
    load8(part + i * 8, acc);
    for (int k = 1; k < nparts; ++k) {
      float t[8];
      load8(part + (int64_t(k) * M * d) + i * 8, t);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += t[j];
    }
    const int64_t e = i * 8;
    const int64_t row = e / d;
    const int col = int(e - row * d);
    store8(dqkv + row * (3 * int64_t(d)) + col, acc);
  }
}

static size_t fwd_smem(int nc) { return 1024 + size_t(1 + 2 * nc) * TILE_BYTES + 2 * P_BYTES + 2 * TILE * 4 + 64 + 256; }
static size_t bwd_smem() { return 1024 + 8 * size_t(TILE_BYTES) + 2 * P_BYTES + 16 + 256; }

template <typename Kern>
static int set_smem(Kern k, size_t bytes) {
  cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, int(bytes));
  if (e != cudaSuccess) return set_error(-5, std::string("attention cudaFuncSetAttribute: ") + cudaGetErrorString(e));
  return 0;
}

static long long* g_attn_trace = nullptr;

static int check_shape(int B, int H, int S) {
  if (S % TILE != 0 || S < TILE || S > 4 * TILE) return set_error(-1, "attention: S must be 128, 256, 384 or 512");
  if (B <= 0 || H <= 0) return set_error(-1, "attention: empty problem");
  return 0;
}

}  // namespace hstd

using namespace hstd;

// Debug aid: device buffer of >= 64 int64 that CTA (0,0,0) of the following attention launches fills
// with clock64() stamps (NULL switches it off).
extern "C" HSTD_API void hstd_attn_set_trace(void* device_buffer) { g_attn_trace = static_cast<long long*>(device_buffer); }

extern "C" HSTD_API int hstd_attn_fwd(const void* qkv, void* ctx, float* lse, const int32_t* key_mask, int B, int H,
                                      int S, float scale, float p_drop, uint32_t site, const hstd_step_state* st,
                                      void* stream) {
  if (int rc = check_shape(B, H, S)) return rc;
  const int64_t d = int64_t(H) * DH;
  const __nv_bfloat16* base = static_cast<const __nv_bfloat16*>(qkv);
  CUtensorMap tq, tk, tv;
  // dims {dh, S (row stride 3d), H (stride dh), B (stride S*3d)}; box {64, 128}
  if (int rc = get_tmap(&tq, base, DH, S, H, B, 3 * d, DH, S * 3 * d, TILE)) return rc;
  if (int rc = get_tmap(&tk, base + d, DH, S, H, B, 3 * d, DH, S * 3 * d, TILE)) return rc;
  if (int rc = get_tmap(&tv, base + 2 * d, DH, S, H, B, 3 * d, DH, S * 3 * d, TILE)) return rc;
  CUtensorMap tc;  // output [B*S, d] seen as {dh, S, H, B}: one 128-row x 64-column box per CTA
  if (int rc = get_tmap(&tc, ctx, DH, S, H, B, d, DH, S * d, TILE)) return rc;
  AttnParams p{};
  p.ctx = static_cast<__nv_bfloat16*>(ctx);
  p.lse = lse;
  p.key_mask = key_mask;
  p.st = st; p.p_drop = p_drop; p.site = site;
  p.B = B; p.H = H; p.S = S; p.scale = scale;
  p.trace = g_attn_trace;
  const int nc = S / TILE;
  dim3 grid(nc, H, B);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  static bool configured[5] = {false, false, false, false, false};
#define HSTD_ATTN_FWD(NC_)                                                        \
  case NC_: {                                                                     \
    if (!configured[NC_]) {                                                       \
      if (int rc = set_smem(attn_fwd_kernel<NC_>, fwd_smem(NC_))) return rc;      \
      configured[NC_] = true;                                                     \
    }                                                                             \
    launch_pdl(attn_fwd_kernel<NC_>, grid, dim3(AT_THREADS), fwd_smem(NC_), s, tq, tk, tv, tc, p);  \
    break;                                                                        \
  }
  switch (nc) {
    HSTD_ATTN_FWD(1) HSTD_ATTN_FWD(2) HSTD_ATTN_FWD(3) HSTD_ATTN_FWD(4)
    default: return set_error(-1, "attention: unsupported S");
  }
#undef HSTD_ATTN_FWD
  return check_launch("attn_fwd");
}

extern "C" HSTD_API int64_t hstd_attn_bwd_workspace_bytes(int B, int H, int S) {
  // delta [B,H,S] fp32 + dq partials [S/128][B*S][H*64] bf16
  return int64_t(B) * H * S * 4 + int64_t(S / TILE) * B * S * H * DH * 2;
}

extern "C" HSTD_API int hstd_attn_bwd(const void* qkv, const void* ctx, const void* dctx, const float* lse,
                                      const int32_t* key_mask, void* dqkv, void* workspace, int B, int H, int S,
                                      float scale, float p_drop, uint32_t site, const hstd_step_state* st,
                                      void* stream) {
  if (int rc = check_shape(B, H, S)) return rc;
  if (!workspace) return set_error(-1, "attn_bwd: workspace required");
  const int64_t d = int64_t(H) * DH;
  const int64_t M = int64_t(B) * S;
  float* delta = static_cast<float*>(workspace);
  __nv_bfloat16* dq_part = reinterpret_cast<__nv_bfloat16*>(static_cast<uint8_t*>(workspace) + int64_t(B) * H * S * 4);
  const __nv_bfloat16* base = static_cast<const __nv_bfloat16*>(qkv);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  launch_pdl(attn_delta_kernel, dim3(unsigned((M * 32 + 255) / 256)), dim3(256), 0, s,
             static_cast<const __nv_bfloat16*>(dctx), static_cast<const __nv_bfloat16*>(ctx), delta, B, S, H);
  if (int rc = check_launch("attn_delta")) return rc;
  CUtensorMap tq, tk, tv, tdo;
  if (int rc = get_tmap(&tq, base, DH, S, H, B, 3 * d, DH, S * 3 * d, TILE)) return rc;
  if (int rc = get_tmap(&tk, base + d, DH, S, H, B, 3 * d, DH, S * 3 * d, TILE)) return rc;
  if (int rc = get_tmap(&tv, base + 2 * d, DH, S, H, B, 3 * d, DH, S * 3 * d, TILE)) return rc;
  if (int rc = get_tmap(&tdo, dctx, DH, S, H, B, d, DH, S * d, TILE)) return rc;
  // outputs leave by TMA bulk stores: the dQ partials [S/128][B*S][d] seen as {d, S, B, key tile},
  // dK / dV in the K / V slots of dqkv seen like the inputs
  CUtensorMap tdqp, tdk, tdv;
  __nv_bfloat16* gbase = static_cast<__nv_bfloat16*>(dqkv);
  if (int rc = get_tmap(&tdqp, dq_part, d, S, B, S / TILE, d, S * d, int64_t(B) * S * d, TILE)) return rc;
  if (int rc = get_tmap(&tdk, gbase + d, DH, S, H, B, 3 * d, DH, S * 3 * d, TILE)) return rc;
  if (int rc = get_tmap(&tdv, gbase + 2 * d, DH, S, H, B, 3 * d, DH, S * 3 * d, TILE)) return rc;
  AttnParams p{};
  p.lse = const_cast<float*>(lse);
  p.delta = delta;
  p.dqkv = static_cast<__nv_bfloat16*>(dqkv);
  p.dq_part = dq_part;
  p.key_mask = key_mask;
  p.st = st; p.p_drop = p_drop; p.site = site;
  p.B = B; p.H = H; p.S = S; p.scale = scale;
  p.trace = g_attn_trace;
  const int nq = S / TILE;
  dim3 grid(nq, H, B);
  static bool configured[5] = {false, false, false, false, false};
#define HSTD_ATTN_BWD(NQ_)                                                             \
  case NQ_: {                                                                          \
    if (!configured[NQ_]) {                                                            \
      if (int rc = set_smem(attn_bwd_kernel<NQ_>, bwd_smem())) return rc;              \
      configured[NQ_] = true;                                                          \
    }                                                                                  \
    launch_pdl(attn_bwd_kernel<NQ_>, grid, dim3(AT_THREADS), bwd_smem(), s, tq, tk, tv, tdo, tdqp, tdk, tdv, p); \
    break;                                                                             \
  }
  switch (nq) {
    HSTD_ATTN_BWD(1) HSTD_ATTN_BWD(2) HSTD_ATTN_BWD(3) HSTD_ATTN_BWD(4)
    default: return set_error(-1, "attention: unsupported S");
  }
#undef HSTD_ATTN_BWD
  if (int rc = check_launch("attn_bwd")) return rc;
  int blocks = int((M * d / 8 + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  launch_pdl(attn_dq_reduce_kernel, dim3(blocks), dim3(256), 0, s, static_cast<const __nv_bfloat16*>(dq_part),
             static_cast<__nv_bfloat16*>(dqkv), M, int(d), nq);
  return check_launch("attn_dq_reduce");
}
