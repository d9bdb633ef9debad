// Device helpers for the memory-bound kernels: per-step state, counter-based dropout RNG
// (Philox4x32-7, regenerated in backward instead of storing masks — SURVEY.md K13), warp
// reductions and 8-wide bf16 vector load/store.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../include/hstd_b200.h"

namespace hstd {

typedef hstd_step_state StepState;

// ------------------------------------------------------------------ Philox4x32-7
// Counter-based RNG (Salmon et al.); 7 rounds is the smallest count that passes BigCrush.
__device__ __forceinline__ uint4 philox4x32_7(uint4 ctr, uint2 key) {
  constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int i = 0; i < 7; ++i) {
    const uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
    const uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += W0;
    key.y += W1;
  }
  return ctr;
}

// Dropout descriptor resolved once per kernel from the step state.  One Philox call yields
// sixteen 8-bit uniforms, so drop probabilities are quantised to multiples of 1/256 (the
// rescale factor uses the quantised value, keeping the expectation exact).
struct Dropout {
  uint2 key;
  uint32_t step, site;
  uint32_t thresh8;  // drop when r8 < thresh8
  float scale;       // 1 / keep_prob
  __device__ __forceinline__ bool active() const { return thresh8 != 0; }
};
__device__ __forceinline__ Dropout make_dropout(const StepState* st, float p, uint32_t site) {
  Dropout d;
  const uint32_t t = p > 0.f ? min(255u, uint32_t(p * 256.f + 0.5f)) : 0u;
  d.thresh8 = (st != nullptr) ? t : 0u;
  d.scale = 256.f / float(256u - d.thresh8);
  d.key = st ? make_uint2(uint32_t(st->seed), uint32_t(st->seed >> 32)) : make_uint2(0, 0);
  d.step = st ? uint32_t(st->step) : 0u;
  d.site = site;
  return d;
}
// Multipliers (0 or scale) for the 16 consecutive elements whose linear index starts at
// 16 * group16.
__device__ __forceinline__ void dropout_mask16(const Dropout& d, uint64_t group16, float (&m)[16]) {
  const uint4 r = philox4x32_7(make_uint4(uint32_t(group16), uint32_t(group16 >> 32), d.site, d.step), d.key);
  const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int b = 0; b < 4; ++b) m[4 * i + b] = (((w[i] >> (8 * b)) & 0xFFu) >= d.thresh8) ? d.scale : 0.f;
}
// The raw sixteen bytes behind dropout_mask16 (byte b of word i belongs to element 4*i + b), for
// kernels that apply the mask with bit operations instead of float multipliers.
__device__ __forceinline__ uint4 dropout_bytes16(const Dropout& d, uint64_t group16) {
  return philox4x32_7(make_uint4(uint32_t(group16), uint32_t(group16 >> 32), d.site, d.step), d.key);
}
// Byte-parallel "byte >= thresh8": x >= t  <=>  x + (256 - t) carries out of the byte.  The low
// seven bits are added without crossing bytes; the carry out of bit 7 is the majority of the two
// top bits and the carry into them.  Result: the MSB of every byte is the keep flag (3 instructions
// for 4 elements; the other bits are garbage).  Needs 1 <= thresh8 <= 255.
struct KeepConst {
  uint32_t k4, k7;
};
__device__ __forceinline__ KeepConst make_keep_const(const Dropout& d) {
  KeepConst kc;
  kc.k4 = (256u - d.thresh8) * 0x01010101u;
  kc.k7 = kc.k4 & 0x7f7f7f7fu;
  return kc;
}
__device__ __forceinline__ uint32_t keep_flags4(uint32_t x, const KeepConst& kc) {
  const uint32_t s = (x & 0x7f7f7f7fu) + kc.k7;
  uint32_t r;
  asm("lop3.b32 %0, %1, %2, %3, 0xE8;" : "=r"(r) : "r"(x), "r"(kc.k4), "r"(s));  // bitwise majority
  return r;
}
// prmt with sign-replicating selectors: expands flag bytes (MSB = flag) to all-ones / all-zeros lanes
template <uint32_t SEL>
__device__ __forceinline__ uint32_t prmt_sel(uint32_t a) {
  uint32_t r;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(0u), "n"(SEL));
  return r;
}

// ------------------------------------------------------------------ packed fp32x2 arithmetic (sm_100)
// FFMA2 / FADD2 / FMUL2 process two floats per issue slot; the softmax loops of the attention
// kernels are issue-bound, so this is where their time goes.
__device__ __forceinline__ uint64_t pack2(float a, float b) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ void unpack2(uint64_t v, float& a, float& b) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
}
__device__ __forceinline__ uint64_t ffma2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ uint64_t fadd2(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ uint64_t fmul2(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}

// Same stream, addressed in groups of 8 elements (group = element index / 8).
__device__ __forceinline__ void dropout_mask8(const Dropout& d, uint64_t group, float (&m)[8]) {
  const uint64_t g16 = group >> 1;
  const uint4 r = philox4x32_7(make_uint4(uint32_t(g16), uint32_t(g16 >> 32), d.site, d.step), d.key);
  const uint32_t w0 = (group & 1) ? r.z : r.x, w1 = (group & 1) ? r.w : r.y;
#pragma unroll
  for (int b = 0; b < 4; ++b) {
    m[b] = (((w0 >> (8 * b)) & 0xFFu) >= d.thresh8) ? d.scale : 0.f;
    m[4 + b] = (((w1 >> (8 * b)) & 0xFFu) >= d.thresh8) ? d.scale : 0.f;
  }
}

// ------------------------------------------------------------------ vectors / reductions
__device__ __forceinline__ void load8(const __nv_bfloat16* p, float (&v)[8]) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 f = __bfloat1622float2(h[i]);
    v[2 * i] = f.x;
    v[2 * i + 1] = f.y;
  }
}
__device__ __forceinline__ void store8(__nv_bfloat16* p, const float (&v)[8]) {
  uint4 u;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
  *reinterpret_cast<uint4*>(p) = u;
}
__device__ __forceinline__ uint32_t pack_bf16x2_dev(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float2 unpack_bf16x2_dev(uint32_t u) {
  __nv_bfloat162 v = *reinterpret_cast<__nv_bfloat162*>(&u);
  return __bfloat1622float2(v);
}
// Round to bf16 and back, so forward statistics match what backward re-reads.
__device__ __forceinline__ float bf16_round(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

}  // namespace hstd
