// Gradient all-reduce over NVLink 5 / NVSwitch peer memory, fused with the 1/world scale,
// the bf16→fp32 cast, the Keras-AdamW update and the bf16 weight write-back
// (replaces hvd.DistributedOptimizer → ncclAllReduce + postscale and the
// per-variable Adam apply of the reference, scripts/train.py:113-114).
//
// Every rank maps every peer's arena (CUDA IPC, or the multicast window for NVLS).  One
// kernel per gradient bucket, three variants picked by bucket size:
//
//   two-shot : rank r owns slice r of the bucket.  It loads slice r from all peers (reduce-
//              scatter by peer loads), averages, applies AdamW to its fp32 master/m/v slice
//              and stores the refreshed bf16 weights into every peer's weight arena
//              (all-gather by peer stores).  Bandwidth-optimal; optimizer state is sharded.
//   one-shot : every rank loads the whole bucket from all peers and updates a replicated
//              fp32 copy — no remote stores; for small buckets where latency dominates.
//   NVLS     : like two-shot but the reduction happens inside the switch
//              (multimem.ld_reduce) and the weight broadcast is one multimem.st.
//
// Synchronisation is a flag protocol in peer memory with system-scope release/acquire and
// monotonically increasing per-(bucket, block) epochs, so a captured CUDA graph can replay
// the kernels with no host involvement.  A clock-based watchdog turns a dead peer into an
// error word instead of a hang.
#include <cstring>
#include <string>

#include "common.cuh"
#include "device_utils.cuh"
#include "optimizer.cuh"

namespace hstd {

constexpr int AR_MAX_WORLD = 8;
constexpr int AR_MAX_BLOCKS = 128;
constexpr int AR_THREADS = 512;
constexpr long long AR_TIMEOUT_CYCLES = 20ll * 1000 * 1000 * 1000;  // ~10 s at 2 GHz

struct ArKernelParams {
  const __nv_bfloat16* g_peers[AR_MAX_WORLD];  // gradient arena base of every rank (own included)
  __nv_bfloat16* w_peers[AR_MAX_WORLD];        // weight arena base of every rank
  uint32_t* flag_peers[AR_MAX_WORLD];          // flag area of every rank
  const __nv_bfloat16* g_mc;                   // multicast views (NVLS) or null
  __nv_bfloat16* w_mc;
  uint32_t* epoch;                             // local, [AR_MAX_BLOCKS] for this bucket
  uint32_t* error;                             // local error word
  float* master; float* m; float* v;           // owned fp32 state of this bucket (compact)
  const StepState* st;
  int64_t start;  // bucket offset in the arenas (elements)
  int64_t numel;  // bucket size (elements), multiple of 8 * world
  int world, rank;
  int flag_slot;  // first flag slot of this bucket: flags[(flag_slot + block) * AR_MAX_WORLD + rank]
  int mode;       // hstd_ar_mode
  float inv_world;
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ld_peer_16B(const void* p) {
  uint4 v;
  asm volatile("ld.global.relaxed.sys.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void st_peer_16B(void* p, const uint4& v) {
  asm volatile("st.global.relaxed.sys.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}
// In-switch reduction of 8 bf16 over all ranks mapped by the multicast address.
__device__ __forceinline__ uint4 multimem_ld_reduce_bf16x8(const void* mc) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(mc)
               : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_16B(void* mc, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w)
               : "memory");
}

// All ranks' block `blockIdx.x` meet here.  Threads 0..world-1 each signal one peer and wait
// for it.  `val` is monotonically increasing per (bucket, block).
__device__ __forceinline__ void block_barrier(const ArKernelParams& p, uint32_t val) {
  __syncthreads();
  if (threadIdx.x < p.world) {
    const int peer = threadIdx.x;
    const int64_t slot = int64_t(p.flag_slot + blockIdx.x) * AR_MAX_WORLD;
    __threadfence_system();
    st_release_sys(p.flag_peers[peer] + slot + p.rank, val);
    const uint32_t* mine = p.flag_peers[p.rank] + slot + peer;
    const long long t0 = clock64();
    while (int32_t(ld_acquire_sys(mine) - val) < 0) {
      if (clock64() - t0 > AR_TIMEOUT_CYCLES) {
        atomicExch(p.error, 1u + uint32_t(peer));
        break;
      }
    }
  }
  __syncthreads();
}

__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
  float2 a = unpack_bf16x2_dev(u.x), b = unpack_bf16x2_dev(u.y), c = unpack_bf16x2_dev(u.z), d = unpack_bf16x2_dev(u.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  return make_uint4(pack_bf16x2_dev(f[0], f[1]), pack_bf16x2_dev(f[2], f[3]), pack_bf16x2_dev(f[4], f[5]),
                    pack_bf16x2_dev(f[6], f[7]));
}

// AdamW on 8 consecutive owned elements at compact index `o`; returns the new bf16 weights.
__device__ __forceinline__ uint4 adam8(const AdamScalars& sc, const ArKernelParams& p, int64_t o, const float (&g)[8]) {
  float4 w0 = *reinterpret_cast<float4*>(p.master + o), w1 = *reinterpret_cast<float4*>(p.master + o + 4);
  float4 m0 = *reinterpret_cast<float4*>(p.m + o), m1 = *reinterpret_cast<float4*>(p.m + o + 4);
  float4 v0 = *reinterpret_cast<float4*>(p.v + o), v1 = *reinterpret_cast<float4*>(p.v + o + 4);
  adam_update(sc, g[0], w0.x, m0.x, v0.x); adam_update(sc, g[1], w0.y, m0.y, v0.y);
  adam_update(sc, g[2], w0.z, m0.z, v0.z); adam_update(sc, g[3], w0.w, m0.w, v0.w);
  adam_update(sc, g[4], w1.x, m1.x, v1.x); adam_update(sc, g[5], w1.y, m1.y, v1.y);
  adam_update(sc, g[6], w1.z, m1.z, v1.z); adam_update(sc, g[7], w1.w, m1.w, v1.w);
  *reinterpret_cast<float4*>(p.master + o) = w0; *reinterpret_cast<float4*>(p.master + o + 4) = w1;
  *reinterpret_cast<float4*>(p.m + o) = m0; *reinterpret_cast<float4*>(p.m + o + 4) = m1;
  *reinterpret_cast<float4*>(p.v + o) = v0; *reinterpret_cast<float4*>(p.v + o + 4) = v1;
  const float w[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
  return pack8(w);
}

// MODE: 0 two-shot, 1 one-shot, 2 NVLS.  UPDATE: fused AdamW (true) or plain averaged
// all-reduce written back into every rank's gradient arena (false; used by the bandwidth
// sweep and the collective tests).
template <int WORLD, int MODE, bool UPDATE>
__global__ void __launch_bounds__(AR_THREADS) fused_allreduce_kernel(const ArKernelParams p) {
  const uint32_t e = p.epoch[blockIdx.x];
  block_barrier(p, 2 * e + 1);  // every rank's gradients of this bucket are complete

  AdamScalars sc;
  if (UPDATE) sc = load_adam_scalars(p.st);
  const int64_t nvec_total = p.numel / 8;
  const int64_t nvec = MODE == 1 ? nvec_total : nvec_total / WORLD;     // vectors this rank reduces
  const int64_t first = MODE == 1 ? 0 : int64_t(p.rank) * nvec;         // first vector (bucket-relative)
  const int64_t stride = int64_t(gridDim.x) * blockDim.x;

  for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < nvec; i += stride) {
    const int64_t elem = p.start + (first + i) * 8;  // arena element offset
    float g[8];
    if (MODE == 2) {
      unpack8(multimem_ld_reduce_bf16x8(p.g_mc + elem), g);
    } else {
      uint4 raw[WORLD];
#pragma unroll
      for (int r = 0; r < WORLD; ++r) raw[r] = ld_peer_16B(p.g_peers[r] + elem);
#pragma unroll
      for (int j = 0; j < 8; ++j) g[j] = 0.f;
#pragma unroll
      for (int r = 0; r < WORLD; ++r) {  // fixed rank order: every rank would form the same sum
        float t[8];
        unpack8(raw[r], t);
#pragma unroll
        for (int j = 0; j < 8; ++j) g[j] += t[j];
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) g[j] *= p.inv_world;

    if (UPDATE) {
      const uint4 wnew = adam8(sc, p, i * 8, g);
      if (MODE == 1) {
        *reinterpret_cast<uint4*>(p.w_peers[p.rank] + elem) = wnew;
      } else if (MODE == 2) {
        multimem_st_16B(p.w_mc + elem, wnew);
      } else {
#pragma unroll
        for (int r = 0; r < WORLD; ++r) st_peer_16B(p.w_peers[(p.rank + r) % WORLD] + elem, wnew);
      }
    } else {
      const uint4 gout = pack8(g);
      if (MODE == 1) {
        // one-shot plain all-reduce cannot be in place (peers still read our copy): result
        // goes to the weight arena slot of the same offset
        *reinterpret_cast<uint4*>(p.w_peers[p.rank] + elem) = gout;
      } else if (MODE == 2) {
        multimem_st_16B(const_cast<__nv_bfloat16*>(p.g_mc) + elem, gout);
      } else {
#pragma unroll
        for (int r = 0; r < WORLD; ++r)
          st_peer_16B(const_cast<__nv_bfloat16*>(p.g_peers[(p.rank + r) % WORLD]) + elem, gout);
      }
    }
  }

  block_barrier(p, 2 * e + 2);  // all peers finished reading our gradients / writing our weights
  if (threadIdx.x == 0) p.epoch[blockIdx.x] = e + 1;
}

// world == 1 or "my optimizer is local" case on IPC arenas: AdamW over owned vectors only.
__global__ void __launch_bounds__(AR_THREADS) local_update_kernel(const ArKernelParams p) {
  const AdamScalars sc = load_adam_scalars(p.st);
  const int64_t nvec = p.numel / 8;
  for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < nvec; i += int64_t(gridDim.x) * blockDim.x) {
    const int64_t elem = p.start + i * 8;
    float g[8];
    unpack8(*reinterpret_cast<const uint4*>(p.g_peers[p.rank] + elem), g);
    *reinterpret_cast<uint4*>(p.w_peers[p.rank] + elem) = adam8(sc, p, i * 8, g);
  }
}

template <int MODE, bool UPDATE>
static int launch_world(const ArKernelParams& p, int blocks, cudaStream_t s) {
  switch (p.world) {
    case 2: fused_allreduce_kernel<2, MODE, UPDATE><<<blocks, AR_THREADS, 0, s>>>(p); break;
    case 4: fused_allreduce_kernel<4, MODE, UPDATE><<<blocks, AR_THREADS, 0, s>>>(p); break;
    case 8: fused_allreduce_kernel<8, MODE, UPDATE><<<blocks, AR_THREADS, 0, s>>>(p); break;
    default: return set_error(-1, "allreduce: world size must be 2, 4 or 8");
  }
  return check_launch("fused_allreduce");
}

}  // namespace hstd

using namespace hstd;

extern "C" HSTD_API int hstd_allreduce_max_blocks(void) { return AR_MAX_BLOCKS; }
extern "C" HSTD_API int64_t hstd_allreduce_flag_words(int num_buckets) {
  return int64_t(num_buckets) * AR_MAX_BLOCKS * AR_MAX_WORLD;
}

extern "C" HSTD_API int hstd_fused_allreduce(const hstd_allreduce_args* a, void* stream) {
  if (a->world < 1 || a->world > AR_MAX_WORLD) return set_error(-1, "allreduce: bad world size");
  const bool local_only = a->mode == HSTD_AR_LOCAL || a->world == 1;
  if (a->numel % (local_only ? 8 : 8 * a->world) != 0)
    return set_error(-1, "allreduce: bucket size must be a multiple of 8*world (8 for the local update)");
  if (a->start % 8 != 0) return set_error(-1, "allreduce: bucket start must be a multiple of 8");
  ArKernelParams p;
  for (int r = 0; r < AR_MAX_WORLD; ++r) {
    p.g_peers[r] = r < a->world ? static_cast<const __nv_bfloat16*>(a->g_peers[r]) : nullptr;
    p.w_peers[r] = r < a->world ? static_cast<__nv_bfloat16*>(a->w_peers[r]) : nullptr;
    p.flag_peers[r] = r < a->world ? static_cast<uint32_t*>(a->flag_peers[r]) : nullptr;
  }
  p.g_mc = static_cast<const __nv_bfloat16*>(a->g_mc);
  p.w_mc = static_cast<__nv_bfloat16*>(a->w_mc);
  p.epoch = static_cast<uint32_t*>(a->epoch);
  p.error = static_cast<uint32_t*>(a->error);
  p.master = a->master; p.m = a->m; p.v = a->v;
  p.st = a->state;
  p.start = a->start; p.numel = a->numel;
  p.world = a->world; p.rank = a->rank;
  p.flag_slot = a->bucket_index * AR_MAX_BLOCKS;
  p.mode = a->mode;
  p.inv_world = 1.0f / float(a->world);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  int blocks = a->blocks > 0 ? a->blocks : 64;
  if (blocks > AR_MAX_BLOCKS) blocks = AR_MAX_BLOCKS;
  const bool update = a->update != 0;
  if (a->mode == HSTD_AR_LOCAL || a->world == 1) {
    if (!update) return 0;
    const int64_t nvec = a->numel / 8;
    int64_t need = (nvec + AR_THREADS - 1) / AR_THREADS;
    if (need < blocks) blocks = int(need > 0 ? need : 1);
    local_update_kernel<<<blocks, AR_THREADS, 0, s>>>(p);
    return check_launch("local_update");
  }
  if (update && (!a->master || !a->m || !a->v || !a->state)) return set_error(-1, "allreduce: optimizer state missing");
  if (a->mode == HSTD_AR_NVLS && (!a->g_mc || !a->w_mc)) return set_error(-1, "allreduce: NVLS needs multicast pointers");
  switch (a->mode) {
    case HSTD_AR_TWO_SHOT: return update ? launch_world<0, true>(p, blocks, s) : launch_world<0, false>(p, blocks, s);
    case HSTD_AR_ONE_SHOT: return update ? launch_world<1, true>(p, blocks, s) : launch_world<1, false>(p, blocks, s);
    case HSTD_AR_NVLS: return update ? launch_world<2, true>(p, blocks, s) : launch_world<2, false>(p, blocks, s);
    default: return set_error(-1, "allreduce: unknown mode");
  }
}

// ------------------------------------------------------------------ IPC plumbing (C++ side of N4)
extern "C" HSTD_API int hstd_ipc_alloc(int64_t bytes, void** ptr, void* handle64) {
  cudaError_t e = cudaMalloc(ptr, size_t(bytes));
  if (e != cudaSuccess) return set_error(-7, std::string("ipc_alloc: cudaMalloc: ") + cudaGetErrorString(e));
  e = cudaMemset(*ptr, 0, size_t(bytes));
  if (e != cudaSuccess) return set_error(-7, std::string("ipc_alloc: cudaMemset: ") + cudaGetErrorString(e));
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  e = cudaIpcGetMemHandle(static_cast<cudaIpcMemHandle_t*>(handle64), *ptr);
  if (e != cudaSuccess) return set_error(-7, std::string("ipc_alloc: cudaIpcGetMemHandle: ") + cudaGetErrorString(e));
  return 0;
}
extern "C" HSTD_API int hstd_ipc_open(const void* handle64, void** ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, sizeof(h));
  cudaError_t e = cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) return set_error(-7, std::string("ipc_open: cudaIpcOpenMemHandle: ") + cudaGetErrorString(e));
  return 0;
}
extern "C" HSTD_API int hstd_ipc_close(void* ptr) {
  cudaError_t e = cudaIpcCloseMemHandle(ptr);
  return e == cudaSuccess ? 0 : set_error(-7, std::string("ipc_close: ") + cudaGetErrorString(e));
}
extern "C" HSTD_API int hstd_ipc_free(void* ptr) {
  cudaError_t e = cudaFree(ptr);
  return e == cudaSuccess ? 0 : set_error(-7, std::string("ipc_free: ") + cudaGetErrorString(e));
}
// Same-process multi-GPU (MirroredStrategy front-end): plain peer access, no IPC handles.
extern "C" HSTD_API int hstd_enable_peer_access(int dev, int peer) {
  int can = 0;
  cudaDeviceCanAccessPeer(&can, dev, peer);
  if (!can) return set_error(-7, "peer access not possible between these devices");
  int cur = 0;
  cudaGetDevice(&cur);
  cudaSetDevice(dev);
  cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
  cudaSetDevice(cur);
  if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled)
    return set_error(-7, std::string("enable_peer_access: ") + cudaGetErrorString(e));
  cudaGetLastError();
  return 0;
}
