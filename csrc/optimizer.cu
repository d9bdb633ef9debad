// Per-step device state (LR schedule, Keras-Adam bias correction, dropout re-keying) and the
// flat Keras-AdamW update (reference scripts/train.py:112-113).  The
// same update is also inlined into the fused all-reduce kernel (allreduce_sm100.cu).
#include "common.cuh"
#include "device_utils.cuh"
#include "optimizer.cuh"

namespace hstd {

__global__ void step_state_init_kernel(StepState* st, uint64_t seed, int64_t step) {
  st->step = step;
  st->seed = seed;
  st->lr = st->lr_t = 0.f;
  st->decay = 1.f;
  st->beta1 = 0.9f; st->beta2 = 0.999f; st->eps = 1e-7f;
  for (int i = 0; i < 4; ++i) { st->metrics[i] = 0.f; st->metrics_acc[i] = 0.0; }
}

__global__ void step_prologue_kernel(StepState* st, float base_lr, int schedule, int64_t total_steps,
                                     int64_t warmup_steps, float beta1, float beta2, float eps, float weight_decay) {
  HSTD_PDL_PROLOGUE();
  const int64_t s0 = st->step;  // steps already taken = 0-based index of this step
  const int64_t t = s0 + 1;
  double lr = base_lr;
  if (schedule == HSTD_SCHED_LINEAR) {
    if (warmup_steps > 0 && s0 < warmup_steps) {
      lr = double(base_lr) * double(s0 + 1) / double(warmup_steps);
    } else {
      const int64_t total = total_steps > warmup_steps + 1 ? total_steps : warmup_steps + 1;
      const double f = double(total - s0) / double(total - warmup_steps);
      lr = double(base_lr) * (f > 0.0 ? f : 0.0);
    }
  }
  const double c1 = 1.0 - pow(double(beta1), double(t));
  const double c2 = 1.0 - pow(double(beta2), double(t));
  st->step = t;
  st->lr = float(lr);
  st->lr_t = float(lr * sqrt(c2) / c1);
  st->decay = float(1.0 - lr * double(weight_decay));
  st->beta1 = beta1; st->beta2 = beta2; st->eps = eps;
}

__global__ void metrics_reset_kernel(StepState* st) {
  for (int i = 0; i < 4; ++i) { st->metrics[i] = 0.f; st->metrics_acc[i] = 0.0; }
}

__global__ void __launch_bounds__(256) adamw_flat_kernel(const StepState* __restrict__ st,
                                                         const __nv_bfloat16* __restrict__ g, float* __restrict__ w,
                                                         float* __restrict__ m, float* __restrict__ v,
                                                         __nv_bfloat16* __restrict__ wb, int64_t n4, float gscale) {
  HSTD_PDL_PROLOGUE();
  const AdamScalars sc = load_adam_scalars(st);
  for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < n4; i += int64_t(gridDim.x) * blockDim.x) {
    const uint2 gu = reinterpret_cast<const uint2*>(g)[i];
    const float2 g01 = unpack_bf16x2_dev(gu.x), g23 = unpack_bf16x2_dev(gu.y);
    float4 wv = reinterpret_cast<float4*>(w)[i], mv = reinterpret_cast<float4*>(m)[i], vv = reinterpret_cast<float4*>(v)[i];
    adam_update(sc, g01.x * gscale, wv.x, mv.x, vv.x);
    adam_update(sc, g01.y * gscale, wv.y, mv.y, vv.y);
    adam_update(sc, g23.x * gscale, wv.z, mv.z, vv.z);
    adam_update(sc, g23.y * gscale, wv.w, mv.w, vv.w);
    reinterpret_cast<float4*>(w)[i] = wv;
    reinterpret_cast<float4*>(m)[i] = mv;
    reinterpret_cast<float4*>(v)[i] = vv;
    uint2 o;
    o.x = pack_bf16x2_dev(wv.x, wv.y);
    o.y = pack_bf16x2_dev(wv.z, wv.w);
    reinterpret_cast<uint2*>(wb)[i] = o;
  }
}

}  // namespace hstd

using namespace hstd;

extern "C" HSTD_API int hstd_step_state_init(hstd_step_state* st, uint64_t seed, int64_t step, void* stream) {
  step_state_init_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(st, seed, step);
  return check_launch("step_state_init");
}
extern "C" HSTD_API int hstd_step_prologue(hstd_step_state* st, float base_lr, int schedule, int64_t total_steps,
                                           int64_t warmup_steps, float beta1, float beta2, float eps,
                                           float weight_decay, void* stream) {
  launch_pdl(step_prologue_kernel, dim3(1), dim3(1), 0, static_cast<cudaStream_t>(stream), st, base_lr, schedule, total_steps,
             warmup_steps, beta1, beta2, eps, weight_decay);
  return check_launch("step_prologue");
}
extern "C" HSTD_API int hstd_metrics_reset(hstd_step_state* st, void* stream) {
  metrics_reset_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(st);
  return check_launch("metrics_reset");
}
extern "C" HSTD_API int hstd_adamw_flat(const hstd_step_state* st, const void* grads, float* master, float* m, float* v,
                                        void* w_bf16, int64_t n, float grad_scale, void* stream) {
  if (n % 4 != 0) return set_error(-1, "adamw_flat: n must be a multiple of 4");
  const int64_t n4 = n / 4;
  int blocks = int((n4 + 255) / 256);
  if (blocks > 148 * 16) blocks = 148 * 16;
  if (blocks < 1) blocks = 1;
  launch_pdl(adamw_flat_kernel, dim3(blocks), dim3(256), 0, static_cast<cudaStream_t>(stream), st,
             static_cast<const __nv_bfloat16*>(grads), master, m, v, static_cast<__nv_bfloat16*>(w_bf16), n4, grad_scale);
  return check_launch("adamw_flat");
}
