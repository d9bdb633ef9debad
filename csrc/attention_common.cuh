// Shared pieces of the fused attention kernels (csrc/attention_sm100.cu and the experimental variants
// under csrc/experimental/): tile geometry, kernel parameters, the 128B-swizzle address function, the
// helpers that write bf16 rows into K-major SW128 tiles, and the key-mask ballot.
#pragma once
#include "common.cuh"
#include "device_utils.cuh"
#include "sm100_ptx.cuh"

namespace hstd {

constexpr int AT_SOFTMAX_WARPS = 8;               // two threads per query row (column halves)
constexpr int AT_THREADS = (AT_SOFTMAX_WARPS + 1) * 32;
constexpr int AT_ST = AT_SOFTMAX_WARPS * 32;      // softmax / epilogue threads
constexpr int TILE = 128;        // query rows / keys per tile
constexpr int DH = 64;           // head dim
constexpr uint32_t TILE_BYTES = TILE * DH * 2;   // 16 KiB: [128 rows x 128 B]
constexpr uint32_t P_BYTES = TILE * TILE * 2;    // 32 KiB: two 64-column swizzle atoms
constexpr float LOG2E = 1.4426950408889634f;

struct AttnParams {
  __nv_bfloat16* ctx;          // fwd out [B,S,H*64]
  float* lse;                  // [B,H,S] natural-log log-sum-exp of the scaled scores
  const float* delta;          // bwd [B,H,S] rowsum(dO ∘ O)
  __nv_bfloat16* dqkv;         // bwd out [B*S, 3*d] (K and V slots written here)
  __nv_bfloat16* dq_part;      // bwd out [S/128][B*S][d] partial dQ per key tile
  const int32_t* key_mask;     // [B,S] 1 = attend
  const StepState* st;
  float p_drop;
  uint32_t site;
  int B, H, S;
  float scale;
  long long* trace;  // optional clock64() stamps of CTA (0,0,0) (tools/attn_trace.py); null in production
};
#define HSTD_ATRACE(slot) do { if (p.trace && (blockIdx.x | blockIdx.y | blockIdx.z) == 0) p.trace[slot] = clock64(); } while (0)

// byte offset of 16-byte chunk `chunk` (0..7) of row `row` inside a [128 x 128B] swizzled atom
__device__ __forceinline__ uint32_t swz(uint32_t row, uint32_t chunk) { return row * 128u + ((chunk ^ (row & 7u)) << 4); }

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void softmax_warps_sync() { asm volatile("bar.sync 1, %0;" ::"n"(AT_ST) : "memory"); }

// Writes 32 consecutive columns [col0, col0+32) of row `row` (bf16) into a K-major SW128
// tile made of 64-column atoms of 16 KiB each.
__device__ __forceinline__ void store_row32(uint8_t* tile, uint32_t row, uint32_t col0, const float (&v)[32]) {
  uint8_t* atom = tile + (col0 >> 6) * TILE_BYTES;
  const uint32_t chunk0 = (col0 & 63u) >> 3;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    uint4 o = make_uint4(pack_bf16x2(v[8 * c], v[8 * c + 1]), pack_bf16x2(v[8 * c + 2], v[8 * c + 3]),
                         pack_bf16x2(v[8 * c + 4], v[8 * c + 5]), pack_bf16x2(v[8 * c + 6], v[8 * c + 7]));
    *reinterpret_cast<uint4*>(atom + swz(row, chunk0 + c)) = o;
  }
}
__device__ __forceinline__ void store_row32_zero(uint8_t* tile, uint32_t row, uint32_t col0) {
  uint8_t* atom = tile + (col0 >> 6) * TILE_BYTES;
  const uint32_t chunk0 = (col0 & 63u) >> 3;
#pragma unroll
  for (int c = 0; c < 4; ++c) *reinterpret_cast<uint4*>(atom + swz(row, chunk0 + c)) = make_uint4(0, 0, 0, 0);
}
// 32 fp32 TMEM values scaled and stored as 32 bf16 (64 contiguous bytes)
__device__ __forceinline__ void store_out32(__nv_bfloat16* out, const uint32_t (&v)[32], float sc) {
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    uint4 o = make_uint4(pack_bf16x2(__uint_as_float(v[8 * c]) * sc, __uint_as_float(v[8 * c + 1]) * sc),
                         pack_bf16x2(__uint_as_float(v[8 * c + 2]) * sc, __uint_as_float(v[8 * c + 3]) * sc),
                         pack_bf16x2(__uint_as_float(v[8 * c + 4]) * sc, __uint_as_float(v[8 * c + 5]) * sc),
                         pack_bf16x2(__uint_as_float(v[8 * c + 6]) * sc, __uint_as_float(v[8 * c + 7]) * sc));
    *reinterpret_cast<uint4*>(out + c * 8) = o;
  }
}
// Ballot the validity of `nwords*32` keys starting at key0 into smem words (warps 0..7).
__device__ __forceinline__ void build_mask_bits(uint32_t* bits, const int32_t* key_mask, int nwords, int warp, int lane) {
  for (int w = warp; w < nwords; w += AT_SOFTMAX_WARPS) {
    const uint32_t b = __ballot_sync(0xffffffffu, key_mask[w * 32 + lane] != 0);
    if (lane == 0) bits[w] = b;
  }
}

}  // namespace hstd
