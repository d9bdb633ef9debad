// Keras-Adam(W) update shared by the flat optimizer kernel and the fused all-reduce.
#pragma once
#include "device_utils.cuh"

namespace hstd {

struct AdamScalars {
  float lr_t, decay, b1, b2, eps;
};
__device__ __forceinline__ AdamScalars load_adam_scalars(const StepState* st) {
  AdamScalars s;
  s.lr_t = st->lr_t; s.decay = st->decay; s.b1 = st->beta1; s.b2 = st->beta2; s.eps = st->eps;
  return s;
}
// m = b1 m + (1-b1) g ; v = b2 v + (1-b2) g^2 ; w = w*decay - lr_t * m / (sqrt(v) + eps)
__device__ __forceinline__ void adam_update(const AdamScalars& s, float g, float& w, float& m, float& v) {
  m = s.b1 * m + (1.f - s.b1) * g;
  v = s.b2 * v + (1.f - s.b2) * g * g;
  // sqrt.approx / rcp.approx (1-2 ulp): the IEEE sqrt + divide sequences cost ~5x the instructions, which
  // made the update pass of the fused all-reduce (few warps per SM, next to a GEMM CTA) compute-bound
  float r, d;
  asm("sqrt.approx.f32 %0, %1;" : "=f"(r) : "f"(v));
  asm("rcp.approx.f32 %0, %1;" : "=f"(d) : "f"(r + s.eps));
  w = w * s.decay - s.lr_t * m * d;
}

}  // namespace hstd
