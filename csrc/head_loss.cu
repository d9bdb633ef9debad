// Classification head + sparse softmax cross-entropy + accuracy, forward and backward
// (reference scripts/train.py:117-119: the HF seq-cls head, Keras
// SparseCategoricalCrossentropy(from_logits=True) and SparseCategoricalAccuracy).
// Batch x hidden here is tiny (8 x 768), so these are CUDA-core kernels; the only goal is
// to keep the whole head in a handful of launches with no host round trip.
#include "common.cuh"
#include "device_utils.cuh"

namespace hstd {

__device__ __forceinline__ float head_act(int kind, float z) { return kind == 0 ? fmaxf(z, 0.f) : tanhf(z); }
// derivative expressed through the stored activation output h
__device__ __forceinline__ float head_act_grad(int kind, float h) { return kind == 0 ? (h > 0.f ? 1.f : 0.f) : 1.f - h * h; }

__device__ __forceinline__ float drop_mult(const Dropout& d, uint64_t idx) {
  if (!d.active()) return 1.f;
  float m[8];
  dropout_mask8(d, idx >> 3, m);
  return m[idx & 7];
}

constexpr int HB = 8;  // batch rows handled together (weights are read once per HB rows)

// h[b,j] = act(b1[j] + sum_k W1[j,k] * drop0(x[b*S, k]));  one warp per output feature j: the
// W1 row is read once and used for HB batch rows at a time.
__global__ void __launch_bounds__(256) head_dense_fwd_kernel(const hstd_head_args p) {
  HSTD_PDL_PROLOGUE();
  const int j = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (j >= p.d) return;
  const Dropout d0 = make_dropout(p.train ? p.state : nullptr, p.kind == 2 ? p.p_drop : 0.f, p.site0);
  const __nv_bfloat16* x = static_cast<const __nv_bfloat16*>(p.x);
  const __nv_bfloat16* w = static_cast<const __nv_bfloat16*>(p.w1) + int64_t(j) * p.d;
  const float bias = __bfloat162float(static_cast<const __nv_bfloat16*>(p.b1)[j]);
  for (int b0 = 0; b0 < p.B; b0 += HB) {
    float acc[HB];
#pragma unroll
    for (int bb = 0; bb < HB; ++bb) acc[bb] = 0.f;
    for (int k = lane * 8; k < p.d; k += 256) {
      float wv[8];
      load8(w + k, wv);
#pragma unroll
      for (int bb = 0; bb < HB; ++bb) {
        const int b = b0 + bb;
        if (b < p.B) {
          float xv[8];
          load8(x + int64_t(b) * p.S * p.d + k, xv);
          if (d0.active()) {
            float m[8];
            dropout_mask8(d0, (uint64_t(b) * p.d + k) >> 3, m);
#pragma unroll
            for (int i = 0; i < 8; ++i) xv[i] *= m[i];
          }
#pragma unroll
          for (int i = 0; i < 8; ++i) acc[bb] += xv[i] * wv[i];
        }
      }
    }
#pragma unroll
    for (int bb = 0; bb < HB; ++bb) {
      const float v = warp_sum(acc[bb]);
      if (lane == 0 && b0 + bb < p.B) p.h[int64_t(b0 + bb) * p.d + j] = head_act(p.kind, v + bias);
    }
  }
}

// logits, loss, accuracy, dlogits.  Single block; warp w handles samples w, w+nwarps, ...
// In training the dropout of the hidden activation is applied here once, eight elements per
// Philox call: h is overwritten with the dropped activation hd = h*m and, when a dz buffer is
// given, dz receives m * act'(h) — head_bwd reads both instead of regenerating the masks.
__global__ void __launch_bounds__(1024) head_logits_loss_kernel(const hstd_head_args p) {
  HSTD_PDL_PROLOGUE();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const Dropout d1 = make_dropout(p.train ? p.state : nullptr, p.p_drop, p.site1);
  extern __shared__ float sm[];  // [B] loss, [B] correct, [B] valid
  float* s_loss = sm;
  float* s_corr = sm + p.B;
  float* s_valid = sm + 2 * p.B;
  const __nv_bfloat16* w2 = static_cast<const __nv_bfloat16*>(p.w2);
  const __nv_bfloat16* b2 = static_cast<const __nv_bfloat16*>(p.b2);
  for (int b = warp; b < p.B; b += nwarps) {
    float* hrow = p.h + int64_t(b) * p.d;
    for (int c0 = 0; c0 < p.C; c0 += 4) {
      float acc[4] = {0.f, 0.f, 0.f, 0.f};
      for (int j = lane * 8; j < p.d; j += 256) {
        float hv[8];
        *reinterpret_cast<float4*>(&hv[0]) = *reinterpret_cast<const float4*>(hrow + j);
        *reinterpret_cast<float4*>(&hv[4]) = *reinterpret_cast<const float4*>(hrow + j + 4);
        if (c0 == 0 && p.train) {
          float m[8];
          if (d1.active()) {
            dropout_mask8(d1, (uint64_t(b) * p.d + j) >> 3, m);
          } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) m[i] = 1.f;
          }
          if (p.dz != nullptr) {
            float g[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) g[i] = m[i] * head_act_grad(p.kind, hv[i]);
            *reinterpret_cast<float4*>(p.dz + int64_t(b) * p.d + j) = *reinterpret_cast<float4*>(&g[0]);
            *reinterpret_cast<float4*>(p.dz + int64_t(b) * p.d + j + 4) = *reinterpret_cast<float4*>(&g[4]);
          }
          if (d1.active()) {
#pragma unroll
            for (int i = 0; i < 8; ++i) hv[i] *= m[i];
            *reinterpret_cast<float4*>(hrow + j) = *reinterpret_cast<float4*>(&hv[0]);
            *reinterpret_cast<float4*>(hrow + j + 4) = *reinterpret_cast<float4*>(&hv[4]);
          }
        }
#pragma unroll
        for (int cc = 0; cc < 4; ++cc) {
          if (c0 + cc < p.C) {
            float wv[8];
            load8(w2 + int64_t(c0 + cc) * p.d + j, wv);
#pragma unroll
            for (int i = 0; i < 8; ++i) acc[cc] += hv[i] * wv[i];
          }
        }
      }
#pragma unroll
      for (int cc = 0; cc < 4; ++cc) {
        const float v = warp_sum(acc[cc]);
        if (lane == 0 && c0 + cc < p.C) p.logits[b * p.C + c0 + cc] = v + __bfloat162float(b2[c0 + cc]);
      }
      __syncwarp();  // a second pass over the classes re-reads the dropped activations written above
    }
    __syncwarp();
    if (lane == 0) {
      const int label = p.labels[b];
      float mx = -INFINITY;
      int arg = 0;
      for (int c = 0; c < p.C; ++c) {
        const float l = p.logits[b * p.C + c];
        if (l > mx) { mx = l; arg = c; }
      }
      float se = 0.f;
      for (int c = 0; c < p.C; ++c) se += expf(p.logits[b * p.C + c] - mx);
      const float lse = mx + logf(se);
      const bool valid = label >= 0 && label < p.C;
      s_valid[b] = valid ? 1.f : 0.f;
      s_loss[b] = valid ? lse - p.logits[b * p.C + label] : 0.f;
      s_corr[b] = (valid && arg == label) ? 1.f : 0.f;
    }
  }
  __syncthreads();
  __shared__ float s_count;
  if (threadIdx.x == 0) {
    float ls = 0.f, cs = 0.f, n = 0.f;
    for (int b = 0; b < p.B; ++b) { ls += s_loss[b]; cs += s_corr[b]; n += s_valid[b]; }
    s_count = n;
    p.state->metrics[0] = ls; p.state->metrics[1] = cs; p.state->metrics[2] = n;
    p.state->metrics_acc[0] += ls; p.state->metrics_acc[1] += cs; p.state->metrics_acc[2] += n;
  }
  __syncthreads();
  if (p.dlogits != nullptr) {
    const float inv = s_count > 0.f ? 1.f / s_count : 0.f;
    for (int i = threadIdx.x; i < p.B * p.C; i += blockDim.x) {
      const int b = i / p.C, c = i - b * p.C;
      const int label = p.labels[b];
      float g = 0.f;
      if (s_valid[b] > 0.f) {
        float mx = -INFINITY;
        for (int k = 0; k < p.C; ++k) mx = fmaxf(mx, p.logits[b * p.C + k]);
        float se = 0.f;
        for (int k = 0; k < p.C; ++k) se += expf(p.logits[b * p.C + k] - mx);
        g = (expf(p.logits[i] - mx) / se - (c == label ? 1.f : 0.f)) * inv;
      }
      p.dlogits[i] = g;
    }
  }
}

// thread per hidden column j: dW2[c,j] for all c and dz[b,j] for all b; the first C threads also
// write db2.  Reads what the forward left behind: h = dropped activation, dz = m * act'(h).
__global__ void __launch_bounds__(256) head_bwd_dz_kernel(const hstd_head_args p) {
  HSTD_PDL_PROLOGUE();
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  const __nv_bfloat16* w2 = static_cast<const __nv_bfloat16*>(p.w2);
  if (j < p.d) {
    for (int c = 0; c < p.C; ++c) {
      float acc = 0.f;
      for (int b = 0; b < p.B; ++b) acc += p.dlogits[b * p.C + c] * p.h[int64_t(b) * p.d + j];
      static_cast<__nv_bfloat16*>(p.dw2)[int64_t(c) * p.d + j] = __float2bfloat16_rn(acc);
    }
    for (int b = 0; b < p.B; ++b) {
      float dh = 0.f;
      for (int c = 0; c < p.C; ++c) dh += p.dlogits[b * p.C + c] * __bfloat162float(w2[int64_t(c) * p.d + j]);
      p.dz[int64_t(b) * p.d + j] *= dh;
    }
  }
  if (j < p.C) {
    float acc = 0.f;
    for (int b = 0; b < p.B; ++b) acc += p.dlogits[b * p.C + j];
    static_cast<__nv_bfloat16*>(p.db2)[j] = __float2bfloat16_rn(acc);
  }
}

// dW1[j,k] = sum_b dz[b,j] * drop0(x[b,k]);  db1[j] = sum_b dz[b,j].  One thread per (j, 8
// consecutive k): 16-byte loads of x and one 16-byte store.
__global__ void __launch_bounds__(256) head_bwd_dw1_kernel(const hstd_head_args p) {
  HSTD_PDL_PROLOGUE();
  const int kchunks = p.d >> 3;
  const int item = blockIdx.x * blockDim.x + threadIdx.x;
  if (item >= p.d * kchunks) return;
  const int j = item / kchunks, k = (item - j * kchunks) * 8;
  const Dropout d0 = make_dropout(p.state, p.kind == 2 ? p.p_drop : 0.f, p.site0);
  const __nv_bfloat16* x = static_cast<const __nv_bfloat16*>(p.x);
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  float accb = 0.f;
  for (int b = 0; b < p.B; ++b) {
    const float dz = p.dz[int64_t(b) * p.d + j];
    float xv[8];
    load8(x + int64_t(b) * p.S * p.d + k, xv);
    if (d0.active()) {
      float m[8];
      dropout_mask8(d0, (uint64_t(b) * p.d + k) >> 3, m);
#pragma unroll
      for (int i = 0; i < 8; ++i) xv[i] *= m[i];
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] += dz * xv[i];
    accb += dz;
  }
  store8(static_cast<__nv_bfloat16*>(p.dw1) + int64_t(j) * p.d + k, acc);
  if (k == 0) static_cast<__nv_bfloat16*>(p.db1)[j] = __float2bfloat16_rn(accb);
}

// dx[b*S, k] = drop0'( sum_j dz[b,j] * W1[j,k] ).  grid (d/64, ceil(B/8)); block = 64 columns x 8
// j-groups.  dz of the HB batch rows is staged in smem so every W1 element is read once per HB
// rows; the j-groups are combined through the same smem afterwards.
constexpr int HDX_COLS = 64, HDX_GROUPS = 8, HDX_MAX_D = 1024;
static_assert(HDX_GROUPS == HB, "the final stage maps j-group g to batch row b0 + g");
__global__ void __launch_bounds__(HDX_COLS * HDX_GROUPS) head_bwd_dx_kernel(const hstd_head_args p) {
  HSTD_PDL_PROLOGUE();
  __shared__ float sh[HB * HDX_MAX_D];  // dz[HB][d], then red[HDX_GROUPS][HB][HDX_COLS] (16 KiB of it)
  const int lc = threadIdx.x % HDX_COLS, grp = threadIdx.x / HDX_COLS;
  const int k = blockIdx.x * HDX_COLS + lc, b0 = blockIdx.y * HB;
  for (int i = threadIdx.x; i < HB * p.d; i += blockDim.x) {
    const int bb = i / p.d, j = i - bb * p.d;
    sh[i] = (b0 + bb < p.B) ? p.dz[int64_t(b0 + bb) * p.d + j] : 0.f;
  }
  __syncthreads();
  const __nv_bfloat16* w1 = static_cast<const __nv_bfloat16*>(p.w1);
  float acc[HB];
#pragma unroll
  for (int bb = 0; bb < HB; ++bb) acc[bb] = 0.f;
  if (k < p.d) {
#pragma unroll 4
    for (int j = grp; j < p.d; j += HDX_GROUPS) {
      const float w = __bfloat162float(w1[int64_t(j) * p.d + k]);
#pragma unroll
      for (int bb = 0; bb < HB; ++bb) acc[bb] += sh[bb * p.d + j] * w;
    }
  }
  __syncthreads();
#pragma unroll
  for (int bb = 0; bb < HB; ++bb) sh[(grp * HB + bb) * HDX_COLS + lc] = acc[bb];
  __syncthreads();
  // thread (grp, lc) finishes batch row b0 + grp of column k
  if (k < p.d && b0 + grp < p.B) {
    float v = 0.f;
#pragma unroll
    for (int g = 0; g < HDX_GROUPS; ++g) v += sh[(g * HB + grp) * HDX_COLS + lc];
    const int b = b0 + grp;
    const Dropout d0 = make_dropout(p.state, p.kind == 2 ? p.p_drop : 0.f, p.site0);
    v *= drop_mult(d0, uint64_t(b) * p.d + k);
    static_cast<__nv_bfloat16*>(p.dx)[int64_t(b) * p.S * p.d + k] = __float2bfloat16_rn(v);
  }
}

}  // namespace hstd

using namespace hstd;

extern "C" HSTD_API int hstd_head_fwd(const hstd_head_args* a, void* stream) {
  if (a->d % 8 != 0) return set_error(-1, "head: d must be a multiple of 8");
  if (a->B > 2048) return set_error(-1, "head: batch too large");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (a->d > HDX_MAX_D) return set_error(-1, "head: d must be <= 1024");
  const int64_t threads = int64_t(a->d) * 32;
  launch_pdl(head_dense_fwd_kernel, dim3(unsigned((threads + 255) / 256)), dim3(256), 0, s, *a);
  if (int rc = check_launch("head_dense_fwd")) return rc;
  launch_pdl(head_logits_loss_kernel, dim3(1), dim3(1024), 3 * a->B * sizeof(float), s, *a);
  return check_launch("head_logits_loss");
}

extern "C" HSTD_API int hstd_head_bwd(const hstd_head_args* a, void* stream) {
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (!a->dlogits || !a->dz || !a->dx) return set_error(-1, "head_bwd: dlogits, dz and dx are required");
  cudaError_t e = cudaMemsetAsync(a->dx, 0, size_t(a->B) * a->S * a->d * 2, s);
  if (e != cudaSuccess) return set_error(-6, std::string("head_bwd memset: ") + cudaGetErrorString(e));
  const int cols = a->d > a->C ? a->d : a->C;
  launch_pdl(head_bwd_dz_kernel, dim3((cols + 255) / 256), dim3(256), 0, s, *a);
  if (int rc = check_launch("head_bwd_dz")) return rc;
  launch_pdl(head_bwd_dw1_kernel, dim3(unsigned((int64_t(a->d) * (a->d / 8) + 255) / 256)), dim3(256), 0, s, *a);
  if (int rc = check_launch("head_bwd_dw1")) return rc;
  launch_pdl(head_bwd_dx_kernel, dim3((a->d + HDX_COLS - 1) / HDX_COLS, (a->B + HB - 1) / HB), dim3(HDX_COLS * HDX_GROUPS), 0, s,
             *a);
  return check_launch("head_bwd_dx");
}
