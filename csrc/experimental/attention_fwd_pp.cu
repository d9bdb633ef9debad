// EXPERIMENTAL — NOT part of libhstd_b200.so (build.py compiles csrc/*.cu only) and NOT yet run on a GPU.
// It compiles for sm_100a (`python tools/build_experimental.py`; `python -m <package>.build --experimental`
// links it into the library as hstd_attn_fwd_pp for trials), and its barrier protocol is the one
// modelled and checked in tools/protocol_sim.py (`attention_fwd_pingpong`): deadlock- and hazard-free
// under random timing, ~14.6 k cycles for two query tiles against 2 x 15.8 k measured for the kernel in
// csrc/attention_sm100.cu.  It is the starting point of the next round's forward kernel.
//
// Fused attention forward, "two query tiles per CTA, ping-pong":
//   * CTA = query tiles A and B (128 rows each) of one (batch, head); keys in chunks of 128 streamed
//     through 2-deep K and V rings shared by both tiles (half the K/V traffic of one tile per CTA);
//   * TMEM: S_A, S_B (one 128-column chunk each) + O_A, O_B (64 columns each) = 384 columns;
//   * online softmax with lazy rescaling: the running row maximum only moves (and O, l are rescaled
//     through tcgen05.ld / tcgen05.st) when it would grow by more than 2^8 — exact, because P and l
//     always use the same reference maximum; the log-sum-exp that is saved does not depend on it;
//   * 4 softmax warps per tile, ONE thread per query row: no cross-thread maximum / sum exchange;
//     warp q of tile A and warp q of tile B sit on the same scheduler, so one tile's exponentials
//     (MUFU) overlap the other tile's waits for its MMAs;
//   * one control thread issues every TMA and tcgen05.mma in the order
//         PV_A(c) | QK_A(c+1) | PV_B(c) | QK_B(c+1)
//     and never waits for an MMA to finish; commit barriers release S buffers, P tiles and ring slots.
// Same dropout stream (Philox counter = element index of the [B,H,S,S] score tensor) and the same
// instruction-lean softmax arithmetic as the kernel in the tree, so the backward kernel is unchanged.
#include <string>
#include <type_traits>

#include "../attention_common.cuh"
#include "../tmap.cuh"

namespace hstd {

constexpr int PP_SOFTMAX_WARPS = 8;  // 4 per query tile
constexpr int PP_THREADS = (PP_SOFTMAX_WARPS + 1) * 32;
constexpr float PP_RESCALE_LOG2 = 8.f;  // move the reference maximum only when it grows by more than 2^8

template <int NC>
__global__ void __launch_bounds__(PP_THREADS, 1)
attn_fwd_pp_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                   const __grid_constant__ CUtensorMap tm_v, const __grid_constant__ CUtensorMap tm_ctx, const AttnParams p) {
  pdl_launch_dependents();
  constexpr uint32_t TMEM_COLS = 512;  // 384 used; allocations are powers of two
  constexpr uint32_t C_S = 0, C_O = 256;
  constexpr uint32_t IDESC_S = make_idesc_bf16(TILE, TILE, false, false);
  constexpr uint32_t IDESC_O = make_idesc_bf16(TILE, DH, false, true);
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sQ = smem;                      // [2 tiles]
  uint8_t* sK = sQ + 2 * TILE_BYTES;       // [2 ring slots]
  uint8_t* sV = sK + 2 * TILE_BYTES;       // [2 ring slots]
  uint8_t* sP = sV + 2 * TILE_BYTES;       // [2 tiles] 32 KiB each
  uint32_t* sBits = reinterpret_cast<uint32_t*>(sP + 2 * P_BYTES);  // [NC*4] key validity, 32 keys per word
  uint64_t* bars = reinterpret_cast<uint64_t*>(sBits + 16);
  uint64_t* bar_q = bars;
  uint64_t* bar_k = bars + 1;    // [2] chunk landed
  uint64_t* bar_v = bars + 3;    // [2]
  uint64_t* bar_kf = bars + 5;   // [2] ring slot free (both tiles' QK of the chunk have completed)
  uint64_t* bar_vf = bars + 7;   // [2]
  uint64_t* bar_s = bars + 9;    // [2 tiles] scores of the current chunk are in TMEM
  uint64_t* bar_p = bars + 11;   // [2 tiles] P written to smem and S consumed (128 arrivals)
  uint64_t* bar_pv = bars + 13;  // [2 tiles] P·V of the current chunk has completed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 15);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int pair = blockIdx.x, h = blockIdx.y, b = blockIdx.z;

  if (warp == PP_SOFTMAX_WARPS) {
    if (lane == 0) {
      tma_prefetch_desc(&tm_q); tma_prefetch_desc(&tm_k); tma_prefetch_desc(&tm_v); tma_prefetch_desc(&tm_ctx);
      mbar_init(bar_q, 1);
      for (int i = 0; i < 2; ++i) {
        mbar_init(&bar_k[i], 1); mbar_init(&bar_v[i], 1); mbar_init(&bar_kf[i], 1); mbar_init(&bar_vf[i], 1);
        mbar_init(&bar_s[i], 1); mbar_init(&bar_p[i], TILE); mbar_init(&bar_pv[i], 1);
      }
      fence_mbar_init();
    }
    __syncwarp();
    pdl_wait();
    if (elect_one()) {  // both query tiles and the first key chunk are needed whatever the mask says
      mbar_expect_tx(bar_q, 2 * TILE_BYTES);
      tma_load_4d(sQ, &tm_q, bar_q, 0, (2 * pair) * TILE, h, b);
      tma_load_4d(sQ + TILE_BYTES, &tm_q, bar_q, 0, (2 * pair + 1) * TILE, h, b);
      mbar_expect_tx(&bar_k[0], TILE_BYTES);
      tma_load_4d(sK, &tm_k, &bar_k[0], 0, 0, h, b);
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
  int nc = 1;  // key chunks after the last valid key contribute exactly zero
#pragma unroll
  for (int w = 0; w < NC * 4; ++w)
    if (sBits[w] != 0) nc = (w >> 2) + 1;

  if (warp == PP_SOFTMAX_WARPS) {
    if (elect_one()) {
      mbar_expect_tx(&bar_v[0], TILE_BYTES);
      tma_load_4d(sV, &tm_v, &bar_v[0], 0, 0, h, b);
      if (nc > 1) {
        mbar_expect_tx(&bar_k[1], TILE_BYTES);
        tma_load_4d(sK + TILE_BYTES, &tm_k, &bar_k[1], 0, TILE, h, b);
        mbar_expect_tx(&bar_v[1], TILE_BYTES);
        tma_load_4d(sV + TILE_BYTES, &tm_v, &bar_v[1], 0, TILE, h, b);
      }
      const uint64_t dq0 = make_smem_desc_sw128(smem_u32(sQ), 16, 1024);
      const uint64_t dk0 = make_smem_desc_sw128(smem_u32(sK), 16, 1024);
      const uint64_t dv0 = make_smem_desc_sw128(smem_u32(sV), TILE_BYTES, 1024);
      const uint64_t dp0 = make_smem_desc_sw128(smem_u32(sP), 16, 1024);
      mbar_wait(bar_q, 0);
      // S_t = Q_t · K_cᵀ.  The S buffer of tile t is free once its softmax warps have arrived on bar_p
      // for the previous chunk (they arrive after their last read of S).
      auto qk = [&](int t, int c) {
        const int r = c & 1;
        mbar_wait(&bar_k[r], (c >> 1) & 1);
        if (c > 0) mbar_wait(&bar_p[t], (c - 1) & 1);
        tc_fence_after();
        const uint64_t da = dq0 + uint64_t(t) * (TILE_BYTES >> 4), db = dk0 + uint64_t(r) * (TILE_BYTES >> 4);
#pragma unroll
        for (int k = 0; k < DH / 16; ++k) umma_bf16(tmem + C_S + t * TILE, da + k * 2, db + k * 2, IDESC_S, k > 0);
        umma_commit(&bar_s[t]);
      };
      // O_t += P_t · V_c (V read MN-major straight from the ring slot)
      auto pv = [&](int t, int c) {
        const int r = c & 1;
        mbar_wait(&bar_p[t], c & 1);
        mbar_wait(&bar_v[r], (c >> 1) & 1);
        tc_fence_after();
        const uint64_t da = dp0 + uint64_t(t) * (P_BYTES >> 4), db = dv0 + uint64_t(r) * (TILE_BYTES >> 4);
#pragma unroll
        for (int k = 0; k < TILE / 16; ++k)
          umma_bf16(tmem + C_O + t * DH, da + (k >> 2) * (TILE_BYTES >> 4) + (k & 3) * 2, db + k * (2048 >> 4), IDESC_O,
                    (c > 0 || k > 0));
        umma_commit(&bar_pv[t]);
      };
      qk(0, 0);
      qk(1, 0);
      umma_commit(&bar_kf[0]);
      for (int c = 0; c < nc; ++c) {
        for (int t = 0; t < 2; ++t) {
          pv(t, c);
          if (t == 1) umma_commit(&bar_vf[c & 1]);
          if (c + 1 < nc) {
            qk(t, c + 1);
            if (t == 1) umma_commit(&bar_kf[(c + 1) & 1]);
          }
        }
        if (c + 2 < nc) {  // refill the ring slots of chunk c with chunk c+2
          const int r = c & 1;
          mbar_wait(&bar_kf[r], (c >> 1) & 1);
          mbar_expect_tx(&bar_k[r], TILE_BYTES);
          tma_load_4d(sK + r * TILE_BYTES, &tm_k, &bar_k[r], 0, (c + 2) * TILE, h, b);
          mbar_wait(&bar_vf[r], (c >> 1) & 1);
          mbar_expect_tx(&bar_v[r], TILE_BYTES);
          tma_load_4d(sV + r * TILE_BYTES, &tm_v, &bar_v[r], 0, (c + 2) * TILE, h, b);
        }
      }
    }
  } else {
    const int t = warp >> 2, quarter = warp & 3;
    const uint32_t row = quarter * 32 + lane;  // TMEM lane = query row of this tile
    const uint32_t t_row = tmem + ((uint32_t(quarter) * 32u) << 16);
    const uint32_t t_s = t_row + C_S + t * TILE, t_o = t_row + C_O + t * DH;
    uint8_t* pt = sP + t * P_BYTES;
    const float scale2 = p.scale * LOG2E;
    const Dropout dr = make_dropout(p.st, p.p_drop, p.site);
    const KeepConst kc = make_keep_const(dr);
    const int qt = 2 * pair + t;
    const uint64_t row_base = (uint64_t(b) * p.H + h) * p.S + uint64_t(qt) * TILE + row;  // global score row
    const uint64_t sc2 = pack2(scale2, scale2);
    float m = -INFINITY;  // reference maximum, in units of log2 of the scaled scores
    uint64_t l2 = pack2(0.f, 0.f);
    for (int c = 0; c < nc; ++c) {
      mbar_wait(&bar_s[t], c & 1);
      tc_fence_after();
      // ---- maximum of this chunk over the valid keys
      float mx = -INFINITY, mx1 = -INFINITY;
#pragma unroll 1
      for (int sb = 0; sb < 4; ++sb) {
        const uint32_t bits = sBits[c * 4 + sb];
        if (bits == 0) continue;  // warp-uniform
        uint32_t v[32];
        tmem_ld_32x32(t_s + sb * 32, v);
        tmem_ld_wait();
        if (bits == 0xffffffffu) {
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            mx = fmaxf(fmaxf(mx, __uint_as_float(v[j])), __uint_as_float(v[j + 1]));
            mx1 = fmaxf(fmaxf(mx1, __uint_as_float(v[j + 2])), __uint_as_float(v[j + 3]));
          }
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) mx = fmaxf(mx, ((bits >> j) & 1u) ? __uint_as_float(v[j]) : -INFINITY);
        }
      }
      const float m_chunk = fmaxf(mx, mx1) * scale2;  // -inf * positive = -inf
      const float m_new = fmaxf(m, m_chunk);
      if (c == 0) {
        m = m_new;
      } else {
        // P tile and O of this tile: free / complete once P·V of the previous chunk has finished
        mbar_wait(&bar_pv[t], (c - 1) & 1);
        tc_fence_after();
        if (__any_sync(0xffffffffu, m_new - m > PP_RESCALE_LOG2)) {  // rare: move the reference maximum
          const float f = (m == -INFINITY) ? 0.f : fast_exp2(m - m_new);
          float la, lb;
          unpack2(l2, la, lb);
          l2 = pack2(la * f, lb * f);
#pragma unroll 1
          for (int g = 0; g < 2; ++g) {
            uint32_t o[32];
            tmem_ld_32x32(t_o + g * 32, o);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) o[j] = __float_as_uint(__uint_as_float(o[j]) * f);
            tmem_st_32x32(t_o + g * 32, o);
          }
          tmem_st_wait();
          m = m_new;
        }
      }
      const float m_ref = (m == -INFINITY) ? 0.f : m;
      const uint64_t nm2 = pack2(-m_ref, -m_ref);
      // ---- probabilities of this chunk -> smem (bf16), running row sum
#pragma unroll 1
      for (int sb = 0; sb < 4; ++sb) {
        const uint32_t bits = sBits[c * 4 + sb];
        if (bits == 0) {
          store_row32_zero(pt, row, sb * 32);
          continue;
        }
        uint32_t v[32];
        tmem_ld_32x32(t_s + sb * 32, v);
        tmem_ld_wait();
        uint32_t pw[16];
        auto body = [&](auto drop_tag, auto full_tag) {
          constexpr bool DROP = decltype(drop_tag)::value, FULL = decltype(full_tag)::value;
          uint32_t mk[16];
          if constexpr (DROP) {
            const uint64_t e0 = row_base * p.S + c * TILE + sb * 32;
#pragma unroll
            for (int g = 0; g < 2; ++g) {
              const uint4 r4 = dropout_bytes16(dr, (e0 >> 4) + g);
              const uint32_t w[4] = {r4.x, r4.y, r4.z, r4.w};
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const uint32_t f = keep_flags4(w[i], kc);
                mk[8 * g + 2 * i] = prmt_sel<0x9988>(f);
                mk[8 * g + 2 * i + 1] = prmt_sel<0xBBAA>(f);
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
          } else {
            float l1 = 0.f;
#pragma unroll
            for (int j = 0; j < 32; j += 2) {
              float a = fast_exp2(__uint_as_float(v[j]) * scale2 - m_ref), b2 = fast_exp2(__uint_as_float(v[j + 1]) * scale2 - m_ref);
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
      }
      tc_fence_before();        // orders the tcgen05.ld / st above before the MMAs the arrival releases
      fence_proxy_async_smem();
      mbar_arrive(&bar_p[t]);
    }
    // ---- epilogue: O / l (and the dropout 1/keep) -> swizzled smem tile -> one TMA store per tile
    float l, l_hi;
    unpack2(l2, l, l_hi);
    l += l_hi;
    mbar_wait(&bar_pv[t], (nc - 1) & 1);
    tc_fence_after();
    const float inv_l = (l > 0.f ? 1.f / l : 0.f) * (dr.active() ? dr.scale : 1.f);
#pragma unroll 1
    for (int g = 0; g < 2; ++g) {
      uint32_t o[32];
      tmem_ld_32x32(t_o + g * 32, o);
      tmem_ld_wait();
      float f[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(o[j]) * inv_l;
      store_row32(pt, row, g * 32, f);  // the P tile is free: every MMA that read it has completed
    }
    fence_proxy_async_smem();
    if (t == 0) asm volatile("bar.sync 1, 128;" ::: "memory");
    else asm volatile("bar.sync 2, 128;" ::: "memory");
    if (quarter == 0 && elect_one()) {
      tma_store_4d(&tm_ctx, pt, 0, qt * TILE, h, b);
      tma_store_commit();
      tma_store_wait_read();
    }
    if (p.lse) p.lse[row_base] = (m == -INFINITY ? 0.f : (m + log2f(fmaxf(l, 1e-30f))) * (1.f / LOG2E));
  }
  tc_fence_before();
  __syncthreads();
  if (warp == PP_SOFTMAX_WARPS) {
    tc_fence_after();
    tmem_dealloc(tmem, TMEM_COLS);
  }
}

static size_t fwd_pp_smem() { return 1024 + 6 * size_t(TILE_BYTES) + 2 * P_BYTES + 64 + 256; }

}  // namespace hstd

using namespace hstd;

// Same contract as hstd_attn_fwd (include/hstd_b200.h); needs S in {256, 512}.
extern "C" HSTD_API int hstd_attn_fwd_pp(const void* qkv, void* ctx, float* lse, const int32_t* key_mask, int B, int H, int S,
                                float scale, float p_drop, uint32_t site, const hstd_step_state* st, void* stream) {
  if (S != 256 && S != 512) return set_error(-1, "attn_fwd_pp: S must be 256 or 512");
  const int64_t d = int64_t(H) * DH;
  const __nv_bfloat16* base = static_cast<const __nv_bfloat16*>(qkv);
  CUtensorMap tq, tk, tv, tc;
  if (int rc = get_tmap(&tq, base, DH, S, H, B, 3 * d, DH, S * 3 * d, TILE)) return rc;
  if (int rc = get_tmap(&tk, base + d, DH, S, H, B, 3 * d, DH, S * 3 * d, TILE)) return rc;
  if (int rc = get_tmap(&tv, base + 2 * d, DH, S, H, B, 3 * d, DH, S * 3 * d, TILE)) return rc;
  if (int rc = get_tmap(&tc, ctx, DH, S, H, B, d, DH, S * d, TILE)) return rc;
  AttnParams p{};
  p.ctx = static_cast<__nv_bfloat16*>(ctx);
  p.lse = lse;
  p.key_mask = key_mask;
  p.st = st; p.p_drop = p_drop; p.site = site;
  p.B = B; p.H = H; p.S = S; p.scale = scale;
  dim3 grid(S / (2 * TILE), H, B);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  cudaError_t e;
  if (S == 512) {
    e = cudaFuncSetAttribute(attn_fwd_pp_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(fwd_pp_smem()));
    if (e != cudaSuccess) return set_error(-5, std::string("attn_fwd_pp smem: ") + cudaGetErrorString(e));
    launch_pdl(attn_fwd_pp_kernel<4>, grid, dim3(PP_THREADS), fwd_pp_smem(), s, tq, tk, tv, tc, p);
  } else {
    e = cudaFuncSetAttribute(attn_fwd_pp_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(fwd_pp_smem()));
    if (e != cudaSuccess) return set_error(-5, std::string("attn_fwd_pp smem: ") + cudaGetErrorString(e));
    launch_pdl(attn_fwd_pp_kernel<2>, grid, dim3(PP_THREADS), fwd_pp_smem(), s, tq, tk, tv, tc, p);
  }
  return check_launch("attn_fwd_pp");
}
