// Memory-bound row kernels: fused (embedding gather | dropout + residual) + LayerNorm
// forward/backward, bias-gradient column sums, masked attention softmax (+dropout)
// forward/backward, embedding scatter-add and position ids.
//
// One warp owns one row; every lane moves 16-byte vectors, statistics are warp-shuffle
// reductions in fp32, dropout masks are regenerated from a Philox counter in backward.
// These replace the TF primitive chains the reference runs for LayerNorm / softmax /
// dropout / gather inside model.fit (scripts/train.py:145).
#include <cub/block/block_scan.cuh>

#include <mutex>
#include <string>

#include "common.cuh"
#include "device_utils.cuh"

namespace hstd {

constexpr int LN_WARPS = 8;
constexpr int LN_BWD_MAX_BLOCKS = 296;  // 2 per SM

// ------------------------------------------------------------------ LayerNorm forward
template <int NV, bool EMBED>
__global__ void __launch_bounds__(LN_WARPS * 32) ln_fwd_kernel(const hstd_ln_args p) {
  HSTD_PDL_PROLOGUE();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * LN_WARPS + warp;
  if (row >= p.M) return;
  const int d = p.d;
  const Dropout din = make_dropout(p.state, p.p_in, p.site_in);
  const Dropout dout = make_dropout(p.state, p.p_out, p.site_out);
  const __nv_bfloat16* gamma = static_cast<const __nv_bfloat16*>(p.gamma);
  const __nv_bfloat16* beta = static_cast<const __nv_bfloat16*>(p.beta);

  float v[NV][8];
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int col = i * 256 + lane * 8;
    if constexpr (EMBED) {
      const int id = p.ids[row], pid = p.pos_ids[row];
      float w[8], q[8];
      load8(static_cast<const __nv_bfloat16*>(p.word) + int64_t(id) * d + col, w);
      load8(static_cast<const __nv_bfloat16*>(p.pos) + int64_t(pid) * d + col, q);
#pragma unroll
      for (int j = 0; j < 8; ++j) v[i][j] = w[j] + q[j];
      if (p.type0) {
        load8(static_cast<const __nv_bfloat16*>(p.type0) + col, q);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[i][j] += q[j];
      }
    } else {
      load8(static_cast<const __nv_bfloat16*>(p.a) + int64_t(row) * d + col, v[i]);
      if (din.active()) {
        float m[8];
        dropout_mask8(din, (uint64_t(row) * d + col) >> 3, m);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[i][j] *= m[j];
      }
      if (p.resid) {
        float r[8];
        load8(static_cast<const __nv_bfloat16*>(p.resid) + int64_t(row) * d + col, r);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[i][j] += r[j];
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      v[i][j] = bf16_round(v[i][j]);
      sum += v[i][j];
    }
    if (p.pre) store8(static_cast<__nv_bfloat16*>(p.pre) + int64_t(row) * d + col, v[i]);
  }
  const float mean = warp_sum(sum) / float(d);
  float sq = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float c = v[i][j] - mean;
      sq += c * c;
    }
  const float rstd = rsqrtf(warp_sum(sq) / float(d) + p.eps);
  if (lane == 0) {
    if (p.mean) p.mean[row] = mean;
    if (p.rstd) p.rstd[row] = rstd;
  }
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int col = i * 256 + lane * 8;
    float g[8], b[8], o[8];
    load8(gamma + col, g);
    load8(beta + col, b);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = (v[i][j] - mean) * rstd * g[j] + b[j];
    if (dout.active()) {
      float m[8];
      dropout_mask8(dout, (uint64_t(row) * d + col) >> 3, m);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] *= m[j];
    }
    store8(static_cast<__nv_bfloat16*>(p.y) + int64_t(row) * d + col, o);
  }
}

// ------------------------------------------------------------------ LayerNorm backward
// ONE kernel: dpre = LN'(dropout_out'(dy)), da = dropout_in'(dpre) row by row (one warp per row, 16-byte
// vectors, warp-shuffle statistics) AND the three column reductions dgamma = sum dy*xhat, dbeta = sum dy,
// dbias = sum da.  A warp cannot keep 3 x d/32 column accumulators in registers next to its row data (that
// version needed 201 registers), so every warp adds its rows' contributions into a private fp32 slice of
// shared memory (conflict-free float4 layout), the block folds its slices, and the block sums go to a small
// fp32 accumulator in global memory with vector reductions (red.global.add.v4.f32).  The last block to finish
// converts the accumulator to the bf16 gradients and leaves it zeroed for the next call — no partials round
// trip, no finalize launch.  (Round 1 ran rows / columns / finalize as three kernels: 21.7 us per LayerNorm,
// reading dy, pre and da twice.)
// `partials` (args): LN_BWD_ACC_FLOATS fp32 + one counter word, zero before the first call.
constexpr int LN_BWD_ACC_FLOATS = 3 * 1024;
constexpr int LN_BWD_ROWS = 2;  // rows per warp

__device__ __forceinline__ void red_add_v4(float* addr, const float4& v) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}

template <int NV, int WARPS>
__global__ void __launch_bounds__(WARPS * 32) ln_bwd_fused_kernel(const hstd_ln_bwd_args p) {
  extern __shared__ float4 ln_sm[];  // [WARPS][3][NV*64] float4, position q = (i*2 + half)*32 + lane
  HSTD_PDL_PROLOGUE();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int d = p.d;
  constexpr int Q = NV * 64;  // float4 positions per quantity
  float4* slice = ln_sm + warp * 3 * Q;
  const Dropout din = make_dropout(p.state, p.p_in, p.site_in);
  const Dropout dout = make_dropout(p.state, p.p_out, p.site_out);
  const __nv_bfloat16* gamma = static_cast<const __nv_bfloat16*>(p.gamma);

#pragma unroll 1
  for (int r = 0; r < LN_BWD_ROWS; ++r) {
    const int row = (blockIdx.x * LN_BWD_ROWS + r) * WARPS + warp;
    if (row >= p.M) {
      if (r == 0) {
        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int q = lane; q < 3 * Q; q += 32) slice[q] = z;
      }
      continue;
    }
    const float mean = p.mean[row], rstd = p.rstd[row];
    float gd[NV][8], xh[NV][8];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int col = i * 256 + lane * 8;
      float g[8];
      load8(static_cast<const __nv_bfloat16*>(p.dy) + int64_t(row) * d + col, gd[i]);
      load8(static_cast<const __nv_bfloat16*>(p.pre) + int64_t(row) * d + col, xh[i]);
      load8(gamma + col, g);
      if (dout.active()) {
        float m[8];
        dropout_mask8(dout, (uint64_t(row) * d + col) >> 3, m);
#pragma unroll
        for (int j = 0; j < 8; ++j) gd[i][j] *= m[j];
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) xh[i][j] = (xh[i][j] - mean) * rstd;
      // column contributions of this row: dgamma += dy * xhat, dbeta += dy
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int q = (i * 2 + h) * 32 + lane;
        float4 cg = make_float4(gd[i][4 * h] * xh[i][4 * h], gd[i][4 * h + 1] * xh[i][4 * h + 1],
                                gd[i][4 * h + 2] * xh[i][4 * h + 2], gd[i][4 * h + 3] * xh[i][4 * h + 3]);
        float4 cb = make_float4(gd[i][4 * h], gd[i][4 * h + 1], gd[i][4 * h + 2], gd[i][4 * h + 3]);
        if (r > 0) {
          const float4 og = slice[q], ob = slice[Q + q];
          cg.x += og.x; cg.y += og.y; cg.z += og.z; cg.w += og.w;
          cb.x += ob.x; cb.y += ob.y; cb.z += ob.z; cb.w += ob.w;
        }
        slice[q] = cg;
        slice[Q + q] = cb;
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        gd[i][j] *= g[j];
        s1 += gd[i][j];
        s2 += gd[i][j] * xh[i][j];
      }
    }
    s1 = warp_sum(s1) / float(d);
    s2 = warp_sum(s2) / float(d);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int col = i * 256 + lane * 8;
      float dp[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) dp[j] = rstd * (gd[i][j] - s1 - xh[i][j] * s2);
      store8(static_cast<__nv_bfloat16*>(p.dpre) + int64_t(row) * d + col, dp);
      if (p.da != nullptr) {
        if (din.active()) {
          float m[8];
          dropout_mask8(din, (uint64_t(row) * d + col) >> 3, m);
#pragma unroll
          for (int j = 0; j < 8; ++j) dp[j] *= m[j];
        }
        store8(static_cast<__nv_bfloat16*>(p.da) + int64_t(row) * d + col, dp);
      }
      // dbias += what the preceding Linear's output gradient is (the value just stored, bf16-rounded)
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int q = (i * 2 + h) * 32 + lane;
        float4 ca = make_float4(bf16_round(dp[4 * h]), bf16_round(dp[4 * h + 1]), bf16_round(dp[4 * h + 2]),
                                bf16_round(dp[4 * h + 3]));
        if (r > 0) {
          const float4 oa = slice[2 * Q + q];
          ca.x += oa.x; ca.y += oa.y; ca.z += oa.z; ca.w += oa.w;
        }
        slice[2 * Q + q] = ca;
      }
    }
  }
  __syncthreads();
  // fold the warps' slices; position q of quantity k -> columns (q/64)*256 + (q%32)*8 + ((q/32)&1)*4 .. +3
  float* acc = p.partials;
  for (int t = threadIdx.x; t < 3 * Q; t += WARPS * 32) {
    float4 s = ln_sm[t];
#pragma unroll
    for (int w = 1; w < WARPS; ++w) {
      const float4 o = ln_sm[w * 3 * Q + t];
      s.x += o.x; s.y += o.y; s.z += o.z; s.w += o.w;
    }
    const int k = t / Q, q = t - k * Q;
    const int col = (q >> 6) * 256 + (q & 31) * 8 + ((q >> 5) & 1) * 4;
    red_add_v4(acc + k * 1024 + col, s);
  }
  __threadfence();
  __syncthreads();
  __shared__ int s_last;
  unsigned int* counter = reinterpret_cast<unsigned int*>(acc + LN_BWD_ACC_FLOATS);
  if (threadIdx.x == 0) s_last = (atomicAdd(counter, 1u) == gridDim.x - 1) ? 1 : 0;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  // every block's sums have landed: bf16 gradients out, accumulator back to zero for the next call
  __nv_bfloat16* outs[3] = {static_cast<__nv_bfloat16*>(p.dgamma), static_cast<__nv_bfloat16*>(p.dbeta),
                            static_cast<__nv_bfloat16*>(p.dbias)};
  for (int t = threadIdx.x; t < 3 * d; t += WARPS * 32) {
    const int k = t / d, c = t - k * d;
    const float v = __ldcg(acc + k * 1024 + c);
    if (outs[k]) outs[k][c] = __float2bfloat16_rn(v);
    __stcg(acc + k * 1024 + c, 0.f);
  }
  if (threadIdx.x == 0) *counter = 0u;
}

// Sum `nparts` partial rows of `width` floats and scatter segments of `seg` columns to up
// to three bf16 destinations (dst[k] may be null).  Block = 32 columns x 8 part-groups:
// coalesced 128-byte row reads, eight independent partial sums per column, smem combine.
constexpr int FIN_COLS = 32, FIN_GROUPS = 8;
__global__ void __launch_bounds__(FIN_COLS * FIN_GROUPS)
partials_finalize_kernel(const float* __restrict__ partials, int nparts, int width, int seg, __nv_bfloat16* d0,
                         __nv_bfloat16* d1, __nv_bfloat16* d2) {
  HSTD_PDL_PROLOGUE();
  const int lc = threadIdx.x & (FIN_COLS - 1), grp = threadIdx.x / FIN_COLS;
  const int c = blockIdx.x * FIN_COLS + lc;
  float s = 0.f;
  if (c < width) {
#pragma unroll 4
    for (int i = grp; i < nparts; i += FIN_GROUPS) s += partials[int64_t(i) * width + c];
  }
  __shared__ float red[FIN_GROUPS][FIN_COLS];
  red[grp][lc] = s;
  __syncthreads();
  if (grp == 0 && c < width) {
#pragma unroll
    for (int g = 1; g < FIN_GROUPS; ++g) s += red[g][lc];
    const int k = c / seg;
    __nv_bfloat16* dst = k == 0 ? d0 : (k == 1 ? d1 : d2);
    if (dst) dst[c - k * seg] = __float2bfloat16_rn(s);
  }
}

// ------------------------------------------------------------------ column sums
constexpr int COLSUM_ROWGROUPS = 32;       // built-in-scratch entry point (512 KiB for N <= 4096)
constexpr int COLSUM_ROWGROUPS_WS = 96;    // caller-workspace entry point: more blocks in flight
__global__ void __launch_bounds__(256) colsum_partial_kernel(const __nv_bfloat16* __restrict__ x, int64_t ld, int M,
                                                             int N, float* __restrict__ partials) {
  HSTD_PDL_PROLOGUE();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int col = blockIdx.x * 256 + lane * 8;
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (col < N) {
    for (int row = blockIdx.y * 8 + warp; row < M; row += gridDim.y * 8) {
      float v[8];
      load8(x + int64_t(row) * ld + col, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += v[j];
    }
  }
  __shared__ float red[8][256];
#pragma unroll
  for (int j = 0; j < 8; ++j) red[warp][lane * 8 + j] = acc[j];
  __syncthreads();
  const int c = threadIdx.x;
  if (blockIdx.x * 256 + c < N) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += red[w][c];
    partials[int64_t(blockIdx.y) * N + blockIdx.x * 256 + c] = s;
  }
}

// ------------------------------------------------------------------ softmax
// NV4 = Sk / 128; lane handles 4 contiguous keys at (i*32 + lane)*4.
template <int NV4>
__global__ void __launch_bounds__(256) softmax_fwd_kernel(__nv_bfloat16* __restrict__ sp, __nv_bfloat16* __restrict__ pd,
                                                          const int32_t* __restrict__ key_mask, int rows, int H, int Sq,
                                                          float p_drop, uint32_t site, const StepState* st) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t row = int64_t(blockIdx.x) * 8 + warp;
  if (row >= rows) return;
  constexpr int Sk = NV4 * 128;
  const int b = int(row / (int64_t(H) * Sq));
  const Dropout dr = make_dropout(st, p_drop, site);
  __nv_bfloat16* src = sp + row * Sk;
  float v[NV4][4];
  float mx = -INFINITY;
#pragma unroll
  for (int i = 0; i < NV4; ++i) {
    const int col = (i * 32 + lane) * 4;
    const uint2 u = *reinterpret_cast<const uint2*>(src + col);
    const int4 km = *reinterpret_cast<const int4*>(key_mask + int64_t(b) * Sk + col);
    const float2 a = unpack_bf16x2_dev(u.x), c = unpack_bf16x2_dev(u.y);
    v[i][0] = km.x ? a.x : -INFINITY;
    v[i][1] = km.y ? a.y : -INFINITY;
    v[i][2] = km.z ? c.x : -INFINITY;
    v[i][3] = km.w ? c.y : -INFINITY;
#pragma unroll
    for (int j = 0; j < 4; ++j) mx = fmaxf(mx, v[i][j]);
  }
  mx = warp_max(mx);
  if (mx == -INFINITY) mx = 0.f;  // fully masked row: all probabilities become 0
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < NV4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      v[i][j] = __expf(v[i][j] - mx);
      sum += v[i][j];
    }
  sum = warp_sum(sum);
  const float inv = sum > 0.f ? 1.f / sum : 0.f;
#pragma unroll
  for (int i = 0; i < NV4; ++i) {
    const int col = (i * 32 + lane) * 4;
    float pr[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) pr[j] = v[i][j] * inv;
    uint2 o;
    o.x = pack_bf16x2_dev(pr[0], pr[1]);
    o.y = pack_bf16x2_dev(pr[2], pr[3]);
    *reinterpret_cast<uint2*>(src + col) = o;
    if (pd != nullptr) {
      if (dr.active()) {
        // one Philox call covers 8 elements: lanes pair up on the same group
        float m[8];
        const uint64_t e = uint64_t(row) * Sk + col;
        dropout_mask8(dr, e >> 3, m);
        const int h = (e & 4) ? 4 : 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) pr[j] *= m[h + j];
      }
      o.x = pack_bf16x2_dev(pr[0], pr[1]);
      o.y = pack_bf16x2_dev(pr[2], pr[3]);
      *reinterpret_cast<uint2*>(pd + row * Sk + col) = o;
    }
  }
}

template <int NV4>
__global__ void __launch_bounds__(256) softmax_bwd_kernel(const __nv_bfloat16* __restrict__ P,
                                                          __nv_bfloat16* __restrict__ dpd, int rows, float p_drop,
                                                          uint32_t site, const StepState* st) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t row = int64_t(blockIdx.x) * 8 + warp;
  if (row >= rows) return;
  constexpr int Sk = NV4 * 128;
  const Dropout dr = make_dropout(st, p_drop, site);
  float pr[NV4][4], dp[NV4][4];
  float dot = 0.f;
#pragma unroll
  for (int i = 0; i < NV4; ++i) {
    const int col = (i * 32 + lane) * 4;
    const uint2 up = *reinterpret_cast<const uint2*>(P + row * Sk + col);
    const uint2 ud = *reinterpret_cast<const uint2*>(dpd + row * Sk + col);
    float2 a = unpack_bf16x2_dev(up.x), c = unpack_bf16x2_dev(up.y);
    pr[i][0] = a.x; pr[i][1] = a.y; pr[i][2] = c.x; pr[i][3] = c.y;
    a = unpack_bf16x2_dev(ud.x); c = unpack_bf16x2_dev(ud.y);
    dp[i][0] = a.x; dp[i][1] = a.y; dp[i][2] = c.x; dp[i][3] = c.y;
    if (dr.active()) {
      float m[8];
      const uint64_t e = uint64_t(row) * Sk + col;
      dropout_mask8(dr, e >> 3, m);
      const int h = (e & 4) ? 4 : 0;
#pragma unroll
      for (int j = 0; j < 4; ++j) dp[i][j] *= m[h + j];
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) dot += dp[i][j] * pr[i][j];
  }
  dot = warp_sum(dot);
#pragma unroll
  for (int i = 0; i < NV4; ++i) {
    const int col = (i * 32 + lane) * 4;
    uint2 o;
    o.x = pack_bf16x2_dev(pr[i][0] * (dp[i][0] - dot), pr[i][1] * (dp[i][1] - dot));
    o.y = pack_bf16x2_dev(pr[i][2] * (dp[i][2] - dot), pr[i][3] * (dp[i][3] - dot));
    *reinterpret_cast<uint2*>(dpd + row * Sk + col) = o;
  }
}

// ------------------------------------------------------------------ embeddings
__global__ void position_ids_kernel(const int32_t* __restrict__ ids, int32_t* __restrict__ pos, int S, int mode,
                                    int pad_id) {
  using Scan = cub::BlockScan<int, 1024>;
  __shared__ typename Scan::TempStorage tmp;
  const int b = blockIdx.x, s = threadIdx.x;
  int nonpad = 0;
  if (s < S) nonpad = (ids[int64_t(b) * S + s] != pad_id) ? 1 : 0;
  int incl;
  Scan(tmp).InclusiveSum(nonpad, incl);
  if (s < S) pos[int64_t(b) * S + s] = mode == 0 ? s : incl * nonpad + pad_id;
}

template <int NV>
__global__ void __launch_bounds__(256) embed_bwd_kernel(const __nv_bfloat16* __restrict__ de,
                                                        const int32_t* __restrict__ ids,
                                                        const int32_t* __restrict__ pos_ids,
                                                        __nv_bfloat16* __restrict__ word_grad,
                                                        __nv_bfloat16* __restrict__ pos_grad, int M, int d) {
  HSTD_PDL_PROLOGUE();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * 8 + warp;
  if (row >= M) return;
  const int id = ids[row], pid = pos_ids[row];
  uint4 u[NV];
  bool nz = false;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    u[i] = *reinterpret_cast<const uint4*>(de + int64_t(row) * d + i * 256 + lane * 8);
    // +0.0 / -0.0 both count as zero
    nz |= ((u[i].x | u[i].y | u[i].z | u[i].w) & 0x7FFF7FFFu) != 0;
  }
  // Padding positions receive an exactly-zero gradient (nothing attends to them and only the
  // CLS row is pooled); skipping them removes the atomic pile-up on the [PAD] embedding row.
  if (!__any_sync(0xffffffffu, nz)) return;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int col = i * 256 + lane * 8;
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u[i]);
    __nv_bfloat162* w = reinterpret_cast<__nv_bfloat162*>(word_grad + int64_t(id) * d + col);
    __nv_bfloat162* q = reinterpret_cast<__nv_bfloat162*>(pos_grad + int64_t(pid) * d + col);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      atomicAdd(w + j, h[j]);
      atomicAdd(q + j, h[j]);
    }
  }
}

static int check_d(int d, const char* who) {
  if (d % 256 != 0 || d < 256 || d > 1024)
    return set_error(-1, std::string(who) + ": hidden size must be a multiple of 256 in [256, 1024]");
  return 0;
}

}  // namespace hstd

using namespace hstd;

#define HSTD_DISPATCH_NV(NV_EXPR, ...)         \
  switch (NV_EXPR) {                           \
    case 1: { constexpr int NV = 1; __VA_ARGS__; break; } \
    case 2: { constexpr int NV = 2; __VA_ARGS__; break; } \
    case 3: { constexpr int NV = 3; __VA_ARGS__; break; } \
    case 4: { constexpr int NV = 4; __VA_ARGS__; break; } \
    default: return set_error(-1, "unsupported vector count"); \
  }

extern "C" HSTD_API int hstd_ln_fwd(const hstd_ln_args* a, void* stream) {
  if (int rc = check_d(a->d, "ln_fwd")) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int blocks = (a->M + LN_WARPS - 1) / LN_WARPS;
  const bool embed = a->ids != nullptr;
  HSTD_DISPATCH_NV(a->d / 256, if (embed) launch_pdl(ln_fwd_kernel<NV, true>, dim3(blocks), dim3(LN_WARPS * 32), 0, s, *a);
                   else launch_pdl(ln_fwd_kernel<NV, false>, dim3(blocks), dim3(LN_WARPS * 32), 0, s, *a));
  return check_launch("ln_fwd");
}

extern "C" HSTD_API int64_t hstd_ln_bwd_partials_floats(int d) { (void)d; return LN_BWD_ACC_FLOATS + 32; }

extern "C" HSTD_API int hstd_ln_bwd(const hstd_ln_bwd_args* a, void* stream) {
  if (int rc = check_d(a->d, "ln_bwd")) return rc;
  if (!a->partials) return set_error(-1, "ln_bwd: partials scratch is required");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  // 16 warps x 2 rows per block: 128 blocks for 4096 tokens, one wave, few block sums to reduce (measured 12.6 us
  // at d = 768 with dropout against 15.3 us for 8-warp blocks; HSTD_LNB_WARPS=8 selects those)
  static const int warps = [] { const char* e = getenv("HSTD_LNB_WARPS"); return e && atoi(e) == 8 ? 8 : 16; }();
  const int rows_per_block = warps * LN_BWD_ROWS;
  const int blocks = (a->M + rows_per_block - 1) / rows_per_block;
  const size_t smem = size_t(warps) * 3 * (a->d / 4) * sizeof(float4);
#define HSTD_LNB_LAUNCH(W_)                                                                                   \
  HSTD_DISPATCH_NV(a->d / 256, {                                                                               \
    static bool configured = false;                                                                            \
    if (!configured) {                                                                                         \
      cudaFuncSetAttribute(ln_bwd_fused_kernel<NV, W_>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)); \
      configured = true;                                                                                       \
    }                                                                                                          \
    launch_pdl(ln_bwd_fused_kernel<NV, W_>, dim3(blocks), dim3(W_ * 32), smem, s, *a);                        \
  })
  if (warps == 16) { HSTD_LNB_LAUNCH(16); } else { HSTD_LNB_LAUNCH(8); }
#undef HSTD_LNB_LAUNCH
  return check_launch("ln_bwd");
}

extern "C" HSTD_API int64_t hstd_colsum_partials_floats(int N) { return int64_t(COLSUM_ROWGROUPS_WS) * N; }

// The column-sum kernels read eight bf16 per 16-byte vector load: a raw-C caller handing in an
// odd pointer or row stride would otherwise get a misaligned-address fault instead of an error.
static int colsum_check_alignment(const void* x, int64_t ld, const char* who) {
  if (reinterpret_cast<uintptr_t>(x) & 15u)
    return set_error(-1, std::string(who) + ": x must be 16-byte aligned");
  if (ld % 8 != 0) return set_error(-1, std::string(who) + ": ld must be a multiple of 8 elements (16 bytes)");
  return 0;
}

extern "C" HSTD_API int hstd_colsum(const void* x, int64_t ld, int M, int N, void* out, float* partials, void* stream) {
  if (N % 8 != 0) return set_error(-1, "colsum: N must be a multiple of 8");
  if (int rc = colsum_check_alignment(x, ld, "colsum")) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  dim3 grid((N + 255) / 256, COLSUM_ROWGROUPS_WS);
  launch_pdl(colsum_partial_kernel, grid, dim3(256), 0, s, static_cast<const __nv_bfloat16*>(x), ld, M, N, partials);
  if (int rc = check_launch("colsum")) return rc;
  launch_pdl(partials_finalize_kernel, dim3((N + FIN_COLS - 1) / FIN_COLS), dim3(FIN_COLS * FIN_GROUPS), 0, s,
             static_cast<const float*>(partials), COLSUM_ROWGROUPS_WS, N, N, static_cast<__nv_bfloat16*>(out),
             static_cast<__nv_bfloat16*>(nullptr), static_cast<__nv_bfloat16*>(nullptr));
  return check_launch("colsum finalize");
}

extern "C" HSTD_API int hstd_softmax_fwd(void* sp, void* pd, const int32_t* key_mask, int B, int H, int Sq, int Sk,
                                         float p_drop, uint32_t site, const hstd_step_state* st, void* stream) {
  if (Sk % 128 != 0 || Sk > 512) return set_error(-1, "softmax: Sk must be 128, 256, 384 or 512");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int64_t rows = int64_t(B) * H * Sq;
  const int blocks = int((rows + 7) / 8);
  HSTD_DISPATCH_NV(Sk / 128, softmax_fwd_kernel<NV><<<blocks, 256, 0, s>>>(
                                 static_cast<__nv_bfloat16*>(sp), static_cast<__nv_bfloat16*>(pd), key_mask, int(rows), H,
                                 Sq, p_drop, site, st));
  return check_launch("softmax_fwd");
}

extern "C" HSTD_API int hstd_softmax_bwd(const void* p, void* dpd, int B, int H, int Sq, int Sk, float p_drop,
                                         uint32_t site, const hstd_step_state* st, void* stream) {
  if (Sk % 128 != 0 || Sk > 512) return set_error(-1, "softmax: Sk must be 128, 256, 384 or 512");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int64_t rows = int64_t(B) * H * Sq;
  const int blocks = int((rows + 7) / 8);
  HSTD_DISPATCH_NV(Sk / 128, softmax_bwd_kernel<NV><<<blocks, 256, 0, s>>>(static_cast<const __nv_bfloat16*>(p),
                                                                           static_cast<__nv_bfloat16*>(dpd), int(rows),
                                                                           p_drop, site, st));
  return check_launch("softmax_bwd");
}

extern "C" HSTD_API int hstd_position_ids(const int32_t* ids, int32_t* pos, int B, int S, int mode, int pad_id,
                                          void* stream) {
  if (S > 1024) return set_error(-1, "position_ids: S must be <= 1024");
  position_ids_kernel<<<B, 1024, 0, static_cast<cudaStream_t>(stream)>>>(ids, pos, S, mode, pad_id);
  return check_launch("position_ids");
}

extern "C" HSTD_API int hstd_embed_bwd(const void* de, const int32_t* ids, const int32_t* pos_ids, void* word_grad,
                                       void* pos_grad, int M, int d, void* stream) {
  if (int rc = check_d(d, "embed_bwd")) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  HSTD_DISPATCH_NV(d / 256, launch_pdl(embed_bwd_kernel<NV>, dim3((M + 7) / 8), dim3(256), 0, s,
                                       static_cast<const __nv_bfloat16*>(de), ids, pos_ids,
                                       static_cast<__nv_bfloat16*>(word_grad), static_cast<__nv_bfloat16*>(pos_grad), M, d));
  return check_launch("embed_bwd");
}

// ------------------------------------------------------------------ column sums without a workspace
// Some users of this library cannot hand us scratch memory and must not cause a device
// allocation: a C caller that has no device allocator of its own to draw a workspace from, or
// a process whose HBM is entirely owned by another pool (the training framework's caching
// allocator, an inference server) so that a stray cudaMalloc would fail or, worse, force an
// implicit device-wide synchronisation in the middle of an overlapped step.  For them the
// per-row-group partial sums live in this fixed array instead of a caller buffer.  It is part
// of the library image: every GPU that loads libhstd_b200.so carries these 512 KiB until the
// process exits.
constexpr int COLSUM_NOALLOC_MAX_N = 4096;  // widest matrix the fixed scratch can serve
static __device__ __align__(256) float g_colsum_noalloc_scratch[COLSUM_ROWGROUPS * COLSUM_NOALLOC_MAX_N];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_noalloc_mu;
float* g_noalloc_scratch_ptr[kMaxDevices] = {};

// The array's name is only a host-side handle; the device address has to be asked for, once
// per device and with that device current (each GPU has its own copy of the array).
int colsum_noalloc_scratch(float** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return set_error(-6, std::string("colsum_noalloc: cudaGetDevice: ") + cudaGetErrorString(e));
  if (dev < 0 || dev >= kMaxDevices) return set_error(-1, "colsum_noalloc: device ordinal out of range");
  std::lock_guard<std::mutex> g(g_noalloc_mu);
  if (g_noalloc_scratch_ptr[dev] == nullptr) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_colsum_noalloc_scratch);
    if (e != cudaSuccess || p == nullptr)
      return set_error(-6, std::string("colsum_noalloc: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
    g_noalloc_scratch_ptr[dev] = static_cast<float*>(p);
  }
  *out = g_noalloc_scratch_ptr[dev];
  return 0;
}
}  // namespace

extern "C" HSTD_API int hstd_colsum_noalloc_max_n(void) { return COLSUM_NOALLOC_MAX_N; }

extern "C" HSTD_API int hstd_colsum_noalloc(const void* x, int64_t ld, int M, int N, void* out, void* stream) {
  if (N % 8 != 0) return set_error(-1, "colsum_noalloc: N must be a multiple of 8");
  if (N > COLSUM_NOALLOC_MAX_N)
    return set_error(-1, "colsum_noalloc: N = " + std::to_string(N) + " exceeds the built-in scratch (max N = " +
                             std::to_string(COLSUM_NOALLOC_MAX_N) + "); use hstd_colsum with a caller workspace");
  if (M <= 0 || N <= 0) return set_error(-1, "colsum_noalloc: empty matrix");
  if (int rc = colsum_check_alignment(x, ld, "colsum_noalloc")) return rc;
  float* scratch = nullptr;
  if (int rc = colsum_noalloc_scratch(&scratch)) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  // Nothing in the scratch has to start at zero: the first kernel overwrites every
  // (row group, column < N) cell the second one reads, so leftovers of an earlier, wider
  // call are never observed.
  dim3 grid((N + 255) / 256, COLSUM_ROWGROUPS);
  colsum_partial_kernel<<<grid, 256, 0, s>>>(static_cast<const __nv_bfloat16*>(x), ld, M, N, scratch);
  if (int rc = check_launch("colsum_noalloc")) return rc;
  partials_finalize_kernel<<<(N + FIN_COLS - 1) / FIN_COLS, FIN_COLS * FIN_GROUPS, 0, s>>>(scratch, COLSUM_ROWGROUPS, N, N,
                                                           static_cast<__nv_bfloat16*>(out), nullptr, nullptr);
  return check_launch("colsum_noalloc finalize");
}
