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
