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
// Small blocks with a small register footprint: a tcgen05 GEMM CTA holds ~54 K of the SM's 64 K registers and
// ~200 KB of its shared memory, an attention CTA ~48 K registers.  A 128-thread block at <= 85 registers (and
// no shared memory) fits NEXT to them, so bucket kernels running underneath backward share SMs with the
// compute kernels instead of displacing their (persistent, statically scheduled) CTAs.
constexpr int AR_THREADS = 128;
constexpr int AR_MIN_BLOCKS_PER_SM = 6;
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
  int phase;      // 0: reduce + update in one launch, 1: reduce only, 2: update only (see allreduce_block)
  int sync;       // bit 0: flag barrier before the work, bit 1: flag barrier after it
  float inv_world;
  // Row-sparse buckets (the embedding tables): the bucket is a stack of rows of `row_elems` elements and
  // `row_stamp[row] == stamp` marks the rows that received a gradient on ANY rank in this step.
  const uint32_t* row_stamp;
  int row_elems;
  int64_t row_vec0;  // bucket-relative index of the first 8-element vector of [start, start+numel)
  // sparse-row marking kernel only
  const int32_t* ids_peers[AR_MAX_WORLD];      // this step's token ids of every rank
  int n_ids;                                   // ids per rank
  int dense_rows;                              // leading rows of the bucket that are always live
  int vocab_rows;                              // rows that ids may address (after the dense ones)
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

// All ranks' block `block` meet here.  Threads 0..world-1 each signal one peer and wait for it.
// `val` is monotonically increasing per (bucket, block).  Returns false when a peer never arrived
// (the watchdog has then set the error word): the caller must not touch shared state any more.
__device__ __forceinline__ bool block_barrier(const ArKernelParams& p, int block, uint32_t val) {
  __syncthreads();
  int failed = 0;
  if (threadIdx.x < p.world) {
    const int peer = threadIdx.x;
    const int64_t slot = int64_t(p.flag_slot + block) * AR_MAX_WORLD;
    __threadfence_system();
    st_release_sys(p.flag_peers[peer] + slot + p.rank, val);
    const uint32_t* mine = p.flag_peers[p.rank] + slot + peer;
    const long long t0 = clock64();
    while (int32_t(ld_acquire_sys(mine) - val) < 0) {
      if (clock64() - t0 > AR_TIMEOUT_CYCLES) {
        atomicExch(p.error, 1u + uint32_t(peer));
        failed = 1;
        break;
      }
    }
  }
  return __syncthreads_or(failed) == 0;
}
// Once the watchdog has fired on this rank every later bucket kernel leaves at once (instead of
// waiting another ~10 s per bucket per step) until the host notices the error word.
__device__ __forceinline__ bool already_failed(const ArKernelParams& p) {
  return *reinterpret_cast<volatile const uint32_t*>(p.error) != 0u;
}

__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
  float2 a = unpack_bf16x2_dev(u.x), b = unpack_bf16x2_dev(u.y), c = unpack_bf16x2_dev(u.z), d = unpack_bf16x2_dev(u.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  return make_uint4(pack_bf16x2_dev(f[0], f[1]), pack_bf16x2_dev(f[2], f[3]), pack_bf16x2_dev(f[4], f[5]),
                    pack_bf16x2_dev(f[6], f[7]));
}

// fp32 master weight + Adam moments of 8 consecutive owned elements (compact index `o`), loaded ahead of
// use so that several vectors' worth of HBM requests are in flight per thread.
struct AdamState8 {
  float4 w0, w1, m0, m1, v0, v1;
};
__device__ __forceinline__ AdamState8 adam_load8(const ArKernelParams& p, int64_t o) {
  AdamState8 s;
  // streaming (evict-first) accesses: 24 bytes of state per parameter pass through once per step and must
  // not push the activations and weights of the concurrently running backward out of L2
  s.w0 = __ldcs(reinterpret_cast<const float4*>(p.master + o)); s.w1 = __ldcs(reinterpret_cast<const float4*>(p.master + o + 4));
  s.m0 = __ldcs(reinterpret_cast<const float4*>(p.m + o)); s.m1 = __ldcs(reinterpret_cast<const float4*>(p.m + o + 4));
  s.v0 = __ldcs(reinterpret_cast<const float4*>(p.v + o)); s.v1 = __ldcs(reinterpret_cast<const float4*>(p.v + o + 4));
  return s;
}
// AdamW on those 8 elements; stores the new state and returns the new bf16 weights.
__device__ __forceinline__ uint4 adam8(const AdamScalars& sc, const ArKernelParams& p, int64_t o, AdamState8& s,
                                       const float (&g)[8]) {
  adam_update(sc, g[0], s.w0.x, s.m0.x, s.v0.x); adam_update(sc, g[1], s.w0.y, s.m0.y, s.v0.y);
  adam_update(sc, g[2], s.w0.z, s.m0.z, s.v0.z); adam_update(sc, g[3], s.w0.w, s.m0.w, s.v0.w);
  adam_update(sc, g[4], s.w1.x, s.m1.x, s.v1.x); adam_update(sc, g[5], s.w1.y, s.m1.y, s.v1.y);
  adam_update(sc, g[6], s.w1.z, s.m1.z, s.v1.z); adam_update(sc, g[7], s.w1.w, s.m1.w, s.v1.w);
  __stcs(reinterpret_cast<float4*>(p.master + o), s.w0); __stcs(reinterpret_cast<float4*>(p.master + o + 4), s.w1);
  __stcs(reinterpret_cast<float4*>(p.m + o), s.m0); __stcs(reinterpret_cast<float4*>(p.m + o + 4), s.m1);
  __stcs(reinterpret_cast<float4*>(p.v + o), s.v0); __stcs(reinterpret_cast<float4*>(p.v + o + 4), s.v1);
  const float w[8] = {s.w0.x, s.w0.y, s.w0.z, s.w0.w, s.w1.x, s.w1.y, s.w1.z, s.w1.w};
  return pack8(w);
}

// Vectors per thread whose peer loads are issued before the first one is consumed.  A peer / multimem load
// is a ~2-3 us round trip through the switch: with ONE 16-byte request per thread a bucket kernel moved
// < 100 GB/s, the comm stream fell behind backward and the backlog showed up as exposed communication.
// 32 landing registers per thread: 8 multimem loads (NVLS) or 8 / WORLD vectors x WORLD peers (two-shot).
template <int WORLD, int MODE>
__host__ __device__ constexpr int ar_unroll() {
  return MODE == 2 ? 8 : (MODE == 1 ? 1 : (WORLD <= 2 ? 4 : (WORLD <= 4 ? 2 : 1)));
}
constexpr int AR_UPDATE_UNROLL = 2;  // vectors whose optimizer state (24 floats each) is fetched together

template <int WORLD, int MODE>
__device__ __forceinline__ void load_vec(const ArKernelParams& p, int64_t elem, uint4 (&raw)[MODE == 2 ? 1 : WORLD]) {
  if (MODE == 2) {
    raw[0] = multimem_ld_reduce_bf16x8(p.g_mc + elem);
  } else {
#pragma unroll
    for (int r = 0; r < WORLD; ++r) raw[r] = ld_peer_16B(p.g_peers[r] + elem);
  }
}
// Sum over ranks (fixed rank order: every rank would form the same sum) x 1/world.
template <int WORLD, int MODE>
__device__ __forceinline__ void reduce_vec(const ArKernelParams& p, const uint4 (&raw)[MODE == 2 ? 1 : WORLD], float (&g)[8]) {
  if (MODE == 2) {
    unpack8(raw[0], g);
  } else {
#pragma unroll
    for (int j = 0; j < 8; ++j) g[j] = 0.f;
#pragma unroll
    for (int r = 0; r < WORLD; ++r) {
      float t[8];
      unpack8(raw[r], t);
#pragma unroll
      for (int j = 0; j < 8; ++j) g[j] += t[j];
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) g[j] *= p.inv_world;
}
template <int WORLD, int MODE>
__device__ __forceinline__ void store_all_ranks(const ArKernelParams& p, __nv_bfloat16* const* peers, __nv_bfloat16* mc,
                                                int64_t elem, const uint4& v) {
  if (MODE == 1) {
    *reinterpret_cast<uint4*>(peers[p.rank] + elem) = v;
  } else if (MODE == 2) {
    multimem_st_16B(mc + elem, v);
  } else {
#pragma unroll
    for (int r = 0; r < WORLD; ++r) st_peer_16B(peers[(p.rank + r) % WORLD] + elem, v);
  }
}

// Row filter of a launch over a row-sparse bucket.
//   ROWS_ALL      : dense bucket, every vector is reduced and updated.
//   ROWS_LIVE     : only rows stamped in this step (some rank produced a gradient for them).
//   ROWS_IDLE     : only rows NOT stamped: their gradient is exactly zero on every rank, so nothing is
//                   read from any peer and no rank has to be waited for — m and v decay, the weight takes
//                   its (history-driven) Adam step, the new bf16 row goes to every rank.  No rank reads an
//                   idle embedding row during the step, so this part runs under the forward pass.
//   ROWS_BOTH     : (update-only launches) every row; live rows take the reduced gradient, idle rows zero.
enum { ROWS_ALL = 0, ROWS_LIVE = 1, ROWS_IDLE = 2, ROWS_BOTH = 3 };

// The work of one block of one rank.  `block` / `nblocks` are passed in (not blockIdx / gridDim) so the
// single-GPU emulation below can run several ranks inside one cooperative grid.
// MODE: 0 two-shot, 1 one-shot, 2 NVLS.  UPDATE: fused AdamW (true) or plain averaged all-reduce written
// back into every rank's gradient arena (false; used by the bandwidth sweep and the collective tests).
// Sharded (two-shot / NVLS) optimizer launches come in three forms (p.phase):
//   0  reduce + update of the bucket in one launch (barrier | reduce | AdamW + weight broadcast | barrier);
//   1  reduce only: barrier, then the averaged gradient of this rank's slice is written back (bf16) into the
//      rank's OWN copy of that slice.  Little HBM traffic (2 B/param), so these run underneath backward
//      without fighting the GEMMs for bandwidth;
//   2  update only: AdamW over the slice from those averaged gradients + weight broadcast, launched for all
//      buckets at the end of the step with big blocks at full HBM speed.  (An HBM-bound kernel at full
//      bandwidth needs most of an SM's registers for requests in flight: measured, hiding AdamW underneath
//      backward in small co-resident blocks cost more through contention than it saved.)  By then every
//      rank has finished backward, so no start barrier; the last launch of the step carries the end barrier.
template <int WORLD, int MODE, bool UPDATE, int ROWS>
__device__ __forceinline__ void allreduce_block(const ArKernelParams& p, int block, int nblocks) {
  if (already_failed(p)) return;
  uint32_t e = p.epoch[block];
  if (p.sync & 1) {
    // every rank's gradients of this bucket are complete
    if (!block_barrier(p, block, ++e)) return;  // a peer is gone: leave weights and optimizer state alone
  }
  AdamScalars sc;
  uint32_t stamp = 0;
  if (UPDATE) {
    sc = load_adam_scalars(p.st);
    stamp = uint32_t(p.st->step);
  }
  const int64_t nvec_total = p.numel / 8;
  const int64_t nvec = MODE == 1 ? nvec_total : nvec_total / WORLD;     // vectors this rank reduces
  const int64_t first = MODE == 1 ? 0 : int64_t(p.rank) * nvec;         // first vector (bucket-relative)
  const int64_t stride = int64_t(nblocks) * blockDim.x;
  const uint32_t row_vecs = ROWS != ROWS_ALL ? uint32_t(p.row_elems) / 8u : 1u;

  constexpr int NL = MODE == 2 ? 1 : WORLD;  // 16-byte loads per vector
  const int64_t t0 = block * int64_t(blockDim.x) + threadIdx.x;
  auto row_live = [&](int64_t i) -> bool {
    return p.row_stamp[uint32_t(p.row_vec0 + first + i) / row_vecs] == stamp;
  };
  auto selected = [&](int64_t i) -> bool {
    if (i >= nvec) return false;
    if (ROWS == ROWS_ALL || ROWS == ROWS_BOTH) return true;
    return row_live(i) == (ROWS == ROWS_LIVE);
  };
  __nv_bfloat16* const* w_peers = p.w_peers;
  __nv_bfloat16* const* g_peers_rw = const_cast<__nv_bfloat16* const*>(p.g_peers);

  if (MODE == 1 || !UPDATE) {
    // one-shot (small replicated buckets) and the plain averaging all-reduce: one pass
    constexpr int U = ar_unroll<WORLD, MODE>();
    for (int64_t i0 = t0; i0 < nvec; i0 += stride * U) {
      uint4 raw[U][NL];
      bool act[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        act[u] = selected(i0 + u * stride);
        if (act[u]) load_vec<WORLD, MODE>(p, p.start + (first + i0 + u * stride) * 8, raw[u]);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (!act[u]) continue;
        const int64_t i = i0 + u * stride, elem = p.start + (first + i) * 8;
        float g[8];
        reduce_vec<WORLD, MODE>(p, raw[u], g);
        if (UPDATE) {
          AdamState8 as = adam_load8(p, i * 8);
          store_all_ranks<WORLD, MODE>(p, w_peers, p.w_mc, elem, adam8(sc, p, i * 8, as, g));
        } else {
          // one-shot plain all-reduce cannot be in place (peers still read our copy): its result goes to
          // the weight arena slot of the same offset
          store_all_ranks<WORLD, MODE>(p, MODE == 1 ? w_peers : g_peers_rw,
                                       const_cast<__nv_bfloat16*>(p.g_mc), elem, pack8(g));
        }
      }
    }
  } else {
    // Sharded optimizer (two-shot / NVLS), two passes of the same thread over the same vectors.
    // Pass A, latency-bound on the switch: many peer loads in flight, the averaged gradient goes back (bf16)
    // into this rank's own copy of its slice — no other rank reads that part of our arena.
    // Pass B, HBM-bound: AdamW on the fp32 state, new bf16 weights to every rank.
    // Splitting them keeps the register footprint small enough to co-reside with a GEMM CTA.
    __nv_bfloat16* own_g = g_peers_rw[p.rank];
    if (ROWS != ROWS_IDLE && ROWS != ROWS_BOTH && p.phase != 2) {
      constexpr int U = ar_unroll<WORLD, MODE>();
      for (int64_t i0 = t0; i0 < nvec; i0 += stride * U) {
        uint4 raw[U][NL];
        bool act[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          act[u] = selected(i0 + u * stride);
          if (act[u]) load_vec<WORLD, MODE>(p, p.start + (first + i0 + u * stride) * 8, raw[u]);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          if (!act[u]) continue;
          float g[8];
          reduce_vec<WORLD, MODE>(p, raw[u], g);
          *reinterpret_cast<uint4*>(own_g + p.start + (first + i0 + u * stride) * 8) = pack8(g);
        }
      }
    }
    constexpr int UB = AR_UPDATE_UNROLL;
    for (int64_t i0 = t0; i0 < nvec && p.phase != 1; i0 += stride * UB) {
      uint4 gv[UB];
      AdamState8 as[UB];
      bool act[UB];
#pragma unroll
      for (int u = 0; u < UB; ++u) {
        const int64_t i = i0 + u * stride;
        act[u] = selected(i);
        if (!act[u]) continue;
        gv[u] = make_uint4(0, 0, 0, 0);  // (bf16 zeros) idle rows: the gradient is zero on every rank
        if (ROWS == ROWS_ALL || ROWS == ROWS_LIVE || (ROWS == ROWS_BOTH && row_live(i)))
          gv[u] = *reinterpret_cast<const uint4*>(own_g + p.start + (first + i) * 8);
        as[u] = adam_load8(p, i * 8);
      }
#pragma unroll
      for (int u = 0; u < UB; ++u) {
        if (!act[u]) continue;
        const int64_t i = i0 + u * stride;
        float g[8];
        unpack8(gv[u], g);
        store_all_ranks<WORLD, MODE>(p, w_peers, p.w_mc, p.start + (first + i) * 8, adam8(sc, p, i * 8, as[u], g));
      }
    }
  }

  if (p.sync & 2) block_barrier(p, block, ++e);  // all peers finished reading our gradients / writing our weights
  if ((p.sync & 3) && threadIdx.x == 0) p.epoch[block] = e;
}

// THREADS = AR_THREADS: the co-resident form for buckets that run underneath backward.  THREADS = AR_TAIL_THREADS:
// the trailing buckets of a step have no compute kernel left to share the SM with and both passes are bound by
// the number of requests in flight, so they take four times the threads (and the registers that go with them).
constexpr int AR_TAIL_THREADS = 512;
template <int WORLD, int MODE, bool UPDATE, int ROWS, int THREADS>
__global__ void __launch_bounds__(THREADS, THREADS == AR_THREADS ? AR_MIN_BLOCKS_PER_SM : 1)
fused_allreduce_kernel(const ArKernelParams p) {
  allreduce_block<WORLD, MODE, UPDATE, ROWS>(p, blockIdx.x, gridDim.x);
}

// Single-GPU emulation of a WORLD-rank launch, for tests and profiling on a box with fewer GPUs than
// ranks: ONE cooperative grid of WORLD x blocks CTAs (co-residency is what lets the flag barriers
// complete), rank = blockIdx.x / blocks, every "peer" arena lives on this device.  Runs the very same
// per-block code as the real kernel.  NVLS needs a multicast mapping and cannot be emulated.
struct ArEmuParams {
  ArKernelParams base;
  float* master[AR_MAX_WORLD]; float* m[AR_MAX_WORLD]; float* v[AR_MAX_WORLD];
  uint32_t* epoch[AR_MAX_WORLD];
  uint32_t* error[AR_MAX_WORLD];
  int blocks;
};
template <int WORLD, int MODE, bool UPDATE, int ROWS>
__global__ void __launch_bounds__(AR_THREADS, AR_MIN_BLOCKS_PER_SM) fused_allreduce_emulated_kernel(const __grid_constant__ ArEmuParams ep) {
  const int rank = blockIdx.x / ep.blocks, block = blockIdx.x % ep.blocks;
  ArKernelParams p = ep.base;
  p.rank = rank;
  p.master = ep.master[rank]; p.m = ep.m[rank]; p.v = ep.v[rank];
  p.epoch = ep.epoch[rank];
  p.error = ep.error[rank];
  allreduce_block<WORLD, MODE, UPDATE, ROWS>(p, block, ep.blocks);
}

// Start of a step, row-sparse bucket: learn which embedding rows ANY rank will produce a gradient for.
// The token ids are inputs, so this is known long before backward: every rank reads every rank's ids
// (peer loads, a few KB), stamps the rows with the current step number and clears those rows of its own
// gradient arena (the scatter-add of the embedding backward accumulates into them; rows nobody stamps are
// never read, so the 47 MB table is never zeroed or reduced densely).  The leading `dense_rows` rows of the
// bucket (LayerNorm, position and token-type tables) are always live.
__device__ __forceinline__ void mark_rows_block(const ArKernelParams& p, int block, int nblocks) {
  if (already_failed(p)) return;
  const uint32_t e = p.epoch[block];
  // every rank has staged this step's ids (the host-to-device copy precedes this kernel in stream order)
  if (!block_barrier(p, block, 2 * e + 1)) return;
  const uint32_t stamp = uint32_t(p.st->step);
  uint32_t* stamps = const_cast<uint32_t*>(p.row_stamp);
  for (int r = block * blockDim.x + threadIdx.x; r < p.dense_rows; r += nblocks * blockDim.x) stamps[r] = stamp;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, warps = blockDim.x >> 5;
  const int total = p.world * p.n_ids;
  const int row_vecs = p.row_elems / 8;
  // 32 ids per warp and round trip (one coalesced peer load; a load per id made this kernel latency-bound:
  // 250 us at 4 ranks), then the warp walks through them and clears the rows it was first to stamp
  for (int t0 = (block * warps + warp) * 32; t0 < total; t0 += nblocks * warps * 32) {
    const int t = t0 + lane;
    int id = -1;
    uint32_t old = stamp;
    if (t < total) {
      const int src = t / p.n_ids, k = t - src * p.n_ids;
      id = p.ids_peers[src][k];  // relaxed peer load; ordered after the barrier's acquire
      if (id >= 0 && id < p.vocab_rows) old = atomicExch(stamps + p.dense_rows + id, stamp);
    }
    uint32_t first = __ballot_sync(0xffffffffu, old != stamp);  // lanes whose row still has to be cleared
    while (first) {
      const int src_lane = __ffs(first) - 1;
      first &= first - 1;
      const int rid = __shfl_sync(0xffffffffu, id, src_lane);
      uint4* row = reinterpret_cast<uint4*>(const_cast<__nv_bfloat16*>(p.g_peers[p.rank]) + p.start +
                                            int64_t(p.dense_rows + rid) * p.row_elems);
      for (int c = lane; c < row_vecs; c += 32) row[c] = make_uint4(0, 0, 0, 0);
    }
  }
  // peers have finished reading our ids: the host may stage the next step's batch once this step is over
  block_barrier(p, block, 2 * e + 2);
  if (threadIdx.x == 0) p.epoch[block] = e + 1;
}
__global__ void __launch_bounds__(AR_THREADS, AR_MIN_BLOCKS_PER_SM) mark_rows_kernel(const ArKernelParams p) {
  mark_rows_block(p, blockIdx.x, gridDim.x);
}
__global__ void __launch_bounds__(AR_THREADS, AR_MIN_BLOCKS_PER_SM) mark_rows_emulated_kernel(const __grid_constant__ ArEmuParams ep) {
  const int rank = blockIdx.x / ep.blocks, block = blockIdx.x % ep.blocks;
  ArKernelParams p = ep.base;
  p.rank = rank;
  p.epoch = ep.epoch[rank];
  p.error = ep.error[rank];
  // in the emulation every rank has its own stamp array and gradient arena; they are passed through the
  // optimizer-state slots that this kernel does not otherwise use
  p.row_stamp = reinterpret_cast<const uint32_t*>(ep.master[rank]);
  mark_rows_block(p, block, ep.blocks);
}

// world == 1 or "my optimizer is local" case on IPC arenas: AdamW over owned vectors only.
__global__ void __launch_bounds__(AR_THREADS, AR_MIN_BLOCKS_PER_SM) local_update_kernel(const ArKernelParams p) {
  const AdamScalars sc = load_adam_scalars(p.st);
  const int64_t nvec = p.numel / 8;
  for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < nvec; i += int64_t(gridDim.x) * blockDim.x) {
    const int64_t elem = p.start + i * 8;
    float g[8];
    unpack8(*reinterpret_cast<const uint4*>(p.g_peers[p.rank] + elem), g);
    AdamState8 as = adam_load8(p, i * 8);
    *reinterpret_cast<uint4*>(p.w_peers[p.rank] + elem) = adam8(sc, p, i * 8, as, g);
  }
}

template <int MODE, bool UPDATE, int ROWS, int THREADS>
static int launch_world(const ArKernelParams& p, int blocks, cudaStream_t s) {
  switch (p.world) {
    case 1:  // a rank on its own: the row-sparse optimizer without peers (timing aid, single-GPU sparse update)
      if (MODE != 0 || !UPDATE) return set_error(-1, "allreduce: world 1 exists for the two-shot optimizer form only");
      fused_allreduce_kernel<1, 0, true, ROWS, THREADS><<<blocks, THREADS, 0, s>>>(p);
      break;
    case 2: fused_allreduce_kernel<2, MODE, UPDATE, ROWS, THREADS><<<blocks, THREADS, 0, s>>>(p); break;
    case 4: fused_allreduce_kernel<4, MODE, UPDATE, ROWS, THREADS><<<blocks, THREADS, 0, s>>>(p); break;
    case 8: fused_allreduce_kernel<8, MODE, UPDATE, ROWS, THREADS><<<blocks, THREADS, 0, s>>>(p); break;
    default: return set_error(-1, "allreduce: world size must be 2, 4 or 8");
  }
  return check_launch("fused_allreduce");
}
// sharded optimizer form: row filter x (co-resident | tail) block size
template <int MODE>
static int launch_update(const ArKernelParams& p, int rows, bool tail, int blocks, cudaStream_t s) {
  switch (rows) {
    case ROWS_ALL:
      return tail ? launch_world<MODE, true, ROWS_ALL, AR_TAIL_THREADS>(p, blocks, s)
                  : launch_world<MODE, true, ROWS_ALL, AR_THREADS>(p, blocks, s);
    case ROWS_LIVE:
      return tail ? launch_world<MODE, true, ROWS_LIVE, AR_TAIL_THREADS>(p, blocks, s)
                  : launch_world<MODE, true, ROWS_LIVE, AR_THREADS>(p, blocks, s);
    case ROWS_IDLE: return launch_world<MODE, true, ROWS_IDLE, AR_THREADS>(p, blocks, s);
    case ROWS_BOTH: return launch_world<MODE, true, ROWS_BOTH, AR_TAIL_THREADS>(p, blocks, s);
    default: return set_error(-1, "allreduce: unknown row filter");
  }
}

template <typename Kern>
static int launch_cooperative(Kern kern, const ArEmuParams& ep, int world, cudaStream_t s) {
  int per_sm = 0, dev = 0, sms = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, AR_THREADS, 0);
  if (world * ep.blocks > per_sm * sms)
    return set_error(-1, "allreduce emulation: world x blocks CTAs cannot be co-resident on this GPU");
  void* args[] = {const_cast<ArEmuParams*>(&ep)};
  cudaError_t e = cudaLaunchCooperativeKernel(reinterpret_cast<const void*>(kern), dim3(world * ep.blocks),
                                              dim3(AR_THREADS), args, 0, s);
  if (e != cudaSuccess) return set_error(-6, std::string("allreduce emulation: ") + cudaGetErrorString(e));
  return 0;
}
template <int MODE, bool UPDATE, int ROWS>
static int launch_emulated_world(const ArEmuParams& ep, cudaStream_t s) {
  switch (ep.base.world) {
    case 2: return launch_cooperative(fused_allreduce_emulated_kernel<2, MODE, UPDATE, ROWS>, ep, 2, s);
    case 4: return launch_cooperative(fused_allreduce_emulated_kernel<4, MODE, UPDATE, ROWS>, ep, 4, s);
    case 8: return launch_cooperative(fused_allreduce_emulated_kernel<8, MODE, UPDATE, ROWS>, ep, 8, s);
    default: return set_error(-1, "allreduce: world size must be 2, 4 or 8");
  }
}

}  // namespace hstd

using namespace hstd;

extern "C" HSTD_API int hstd_allreduce_max_blocks(void) { return AR_MAX_BLOCKS; }
extern "C" HSTD_API int64_t hstd_allreduce_flag_words(int num_buckets) {
  return int64_t(num_buckets) * AR_MAX_BLOCKS * AR_MAX_WORLD;
}

static int fill_params(const hstd_allreduce_args* a, ArKernelParams& p) {
  if (a->world < 1 || a->world > AR_MAX_WORLD) return set_error(-1, "allreduce: bad world size");
  for (int r = 0; r < AR_MAX_WORLD; ++r) {
    p.g_peers[r] = r < a->world ? static_cast<const __nv_bfloat16*>(a->g_peers[r]) : nullptr;
    p.w_peers[r] = r < a->world ? static_cast<__nv_bfloat16*>(a->w_peers[r]) : nullptr;
    p.flag_peers[r] = r < a->world ? static_cast<uint32_t*>(a->flag_peers[r]) : nullptr;
    p.ids_peers[r] = r < a->world ? static_cast<const int32_t*>(a->ids_peers[r]) : nullptr;
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
  p.phase = a->phase;
  // barriers: both for the one-launch form (none when only idle rows are touched), before a reduce-only launch,
  // after an update-only launch only where the caller asks for it (the last one of the step)
  p.sync = a->phase == 1 ? 1 : (a->phase == 2 ? (a->end_barrier ? 2 : 0) : (a->rows == HSTD_AR_ROWS_IDLE ? 0 : 3));
  p.inv_world = 1.0f / float(a->world);
  p.row_stamp = static_cast<const uint32_t*>(a->row_stamp);
  p.row_elems = a->row_elems;
  p.row_vec0 = a->row_elem0 / 8;
  p.n_ids = a->n_ids; p.dense_rows = a->dense_rows; p.vocab_rows = a->vocab_rows;
  return 0;
}

static int check_rows(const hstd_allreduce_args* a) {
  if (a->rows == HSTD_AR_ROWS_ALL) return 0;
  if (a->rows != HSTD_AR_ROWS_LIVE && a->rows != HSTD_AR_ROWS_IDLE && a->rows != HSTD_AR_ROWS_BOTH)
    return set_error(-1, "allreduce: unknown row filter");
  if (a->rows == HSTD_AR_ROWS_BOTH && a->phase != 2)
    return set_error(-1, "allreduce: HSTD_AR_ROWS_BOTH belongs to update-only launches (phase 2)");
  if (!a->update) return set_error(-1, "allreduce: row filters exist for the fused optimizer form only");
  if (!a->row_stamp || a->row_elems <= 0 || a->row_elems % 8 != 0)
    return set_error(-1, "allreduce: row filter needs row_stamp and row_elems (multiple of 8)");
  // (a rank on its own — world 1, the exposed-communication timing aid — works on its slice of the bucket, which
  // need not start or end on a row boundary: rows are looked up per vector through row_elem0)
  if (a->world > 1 && a->numel % a->row_elems != 0)
    return set_error(-1, "allreduce: a row-sparse bucket must be whole rows");
  if (a->mode == HSTD_AR_ONE_SHOT) return set_error(-1, "allreduce: row filters need the sharded (two-shot / NVLS) form");
  return 0;
}

static int clamp_blocks(int blocks) {
  if (blocks <= 0) blocks = 64;
  return blocks > AR_MAX_BLOCKS ? AR_MAX_BLOCKS : blocks;
}

extern "C" HSTD_API int hstd_fused_allreduce(const hstd_allreduce_args* a, void* stream) {
  ArKernelParams p;
  if (int rc = fill_params(a, p)) return rc;
  const bool local_only = a->mode == HSTD_AR_LOCAL || (a->world == 1 && a->rows == HSTD_AR_ROWS_ALL);
  if (a->row_elem0 % 8 != 0) return set_error(-1, "allreduce: row_elem0 must be a multiple of 8");
  if (a->numel % (local_only ? 8 : 8 * a->world) != 0)
    return set_error(-1, "allreduce: bucket size must be a multiple of 8*world (8 for the local update)");
  if (a->start % 8 != 0) return set_error(-1, "allreduce: bucket start must be a multiple of 8");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  int blocks = clamp_blocks(a->blocks);
  const bool update = a->update != 0;
  if (local_only) {
    if (!update) return 0;
    const int64_t nvec = a->numel / 8;
    int64_t need = (nvec + AR_THREADS - 1) / AR_THREADS;
    if (need < blocks) blocks = int(need > 0 ? need : 1);
    local_update_kernel<<<blocks, AR_THREADS, 0, s>>>(p);
    return check_launch("local_update");
  }
  if (update && (!a->master || !a->m || !a->v || !a->state)) return set_error(-1, "allreduce: optimizer state missing");
  if (a->mode == HSTD_AR_NVLS && (!a->g_mc || !a->w_mc)) return set_error(-1, "allreduce: NVLS needs multicast pointers");
  if (int rc = check_rows(a)) return rc;
  if (a->phase < 0 || a->phase > 2) return set_error(-1, "allreduce: phase must be 0, 1 or 2");
  if (a->phase != 0 && (!update || a->mode == HSTD_AR_ONE_SHOT))
    return set_error(-1, "allreduce: split reduce / update launches exist for the sharded optimizer form only");
  const bool tail = a->tail != 0;
  switch (a->mode) {
    case HSTD_AR_TWO_SHOT:
      if (update) return launch_update<0>(p, a->rows, tail, blocks, s);
      return tail ? launch_world<0, false, ROWS_ALL, AR_TAIL_THREADS>(p, blocks, s)
                  : launch_world<0, false, ROWS_ALL, AR_THREADS>(p, blocks, s);
    case HSTD_AR_ONE_SHOT:
      return update ? launch_world<1, true, ROWS_ALL, AR_THREADS>(p, blocks, s)
                    : launch_world<1, false, ROWS_ALL, AR_THREADS>(p, blocks, s);
    case HSTD_AR_NVLS:
      if (update) return launch_update<2>(p, a->rows, tail, blocks, s);
      return tail ? launch_world<2, false, ROWS_ALL, AR_TAIL_THREADS>(p, blocks, s)
                  : launch_world<2, false, ROWS_ALL, AR_THREADS>(p, blocks, s);
    default: return set_error(-1, "allreduce: unknown mode");
  }
}

extern "C" HSTD_API int hstd_sparse_rows_mark(const hstd_allreduce_args* a, void* stream) {
  ArKernelParams p;
  if (int rc = fill_params(a, p)) return rc;
  if (!a->row_stamp || !a->state || a->row_elems <= 0 || a->row_elems % 8 != 0 || a->n_ids <= 0 || a->dense_rows < 0 ||
      a->vocab_rows <= 0)
    return set_error(-1, "sparse_rows_mark: row_stamp, state, row_elems (multiple of 8), n_ids, vocab_rows required");
  for (int r = 0; r < a->world; ++r)
    if (!a->ids_peers[r]) return set_error(-1, "sparse_rows_mark: ids pointer of every rank required");
  mark_rows_kernel<<<clamp_blocks(a->blocks), AR_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(p);
  return check_launch("sparse_rows_mark");
}

// ------------------------------------------------------------------ single-GPU emulation (tests, profiling)
static int fill_emu(const hstd_allreduce_args* per_rank, int world, ArEmuParams& ep) {
  if (world != 2 && world != 4 && world != 8) return set_error(-1, "allreduce emulation: world must be 2, 4 or 8");
  if (int rc = fill_params(&per_rank[0], ep.base)) return rc;
  ep.base.world = world;
  ep.base.inv_world = 1.0f / float(world);
  for (int r = 0; r < world; ++r) {
    const hstd_allreduce_args& a = per_rank[r];
    if (a.start != per_rank[0].start || a.numel != per_rank[0].numel || a.mode != per_rank[0].mode ||
        a.bucket_index != per_rank[0].bucket_index)
      return set_error(-1, "allreduce emulation: every rank must describe the same bucket");
    ep.master[r] = a.master; ep.m[r] = a.m; ep.v[r] = a.v;
    ep.epoch[r] = static_cast<uint32_t*>(a.epoch);
    ep.error[r] = static_cast<uint32_t*>(a.error);
  }
  ep.blocks = clamp_blocks(per_rank[0].blocks);
  return 0;
}

extern "C" HSTD_API int hstd_fused_allreduce_emulated(const hstd_allreduce_args* per_rank, int world, void* stream) {
  ArEmuParams ep;
  if (int rc = fill_emu(per_rank, world, ep)) return rc;
  const hstd_allreduce_args& a = per_rank[0];
  if (a.numel % (8 * world) != 0 || a.start % 8 != 0) return set_error(-1, "allreduce emulation: bad bucket bounds");
  if (int rc = check_rows(&a)) return rc;
  const bool update = a.update != 0;
  if (update && !a.state) return set_error(-1, "allreduce emulation: step state missing");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (a.mode == HSTD_AR_TWO_SHOT) {
    if (!update) return launch_emulated_world<0, false, ROWS_ALL>(ep, s);
    switch (a.rows) {
      case HSTD_AR_ROWS_LIVE: return launch_emulated_world<0, true, ROWS_LIVE>(ep, s);
      case HSTD_AR_ROWS_IDLE: return launch_emulated_world<0, true, ROWS_IDLE>(ep, s);
      case HSTD_AR_ROWS_BOTH: return launch_emulated_world<0, true, ROWS_BOTH>(ep, s);
      default: return launch_emulated_world<0, true, ROWS_ALL>(ep, s);
    }
  }
  if (a.mode == HSTD_AR_ONE_SHOT)
    return update ? launch_emulated_world<1, true, ROWS_ALL>(ep, s) : launch_emulated_world<1, false, ROWS_ALL>(ep, s);
  return set_error(-1, "allreduce emulation: two-shot and one-shot only (NVLS needs a multicast mapping)");
}

// per_rank[r].master carries rank r's row_stamp array here (each emulated rank keeps its own).
extern "C" HSTD_API int hstd_sparse_rows_mark_emulated(const hstd_allreduce_args* per_rank, int world, void* stream) {
  ArEmuParams ep;
  if (int rc = fill_emu(per_rank, world, ep)) return rc;
  for (int r = 0; r < world; ++r) ep.master[r] = reinterpret_cast<float*>(per_rank[r].row_stamp);
  return launch_cooperative(mark_rows_emulated_kernel, ep, world, static_cast<cudaStream_t>(stream));
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
