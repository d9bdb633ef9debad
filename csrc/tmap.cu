// TMA tensor-map cache shared by the GEMM and attention kernels.
#include <cuda_runtime.h>

#include <mutex>
#include <string>
#include <unordered_map>

#include "common.cuh"
#include "tmap.cuh"

namespace hstd {

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode_fn() {
  static PFN_encodeTiled fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_encodeTiled>(ptr);
  });
  return fn;
}

struct TmapKey {
  const void* ptr;
  uint64_t d0, d1, d2, d3, s1, s2, s3;
  uint32_t b0, b1;
  bool operator==(const TmapKey& o) const {
    return ptr == o.ptr && d0 == o.d0 && d1 == o.d1 && d2 == o.d2 && d3 == o.d3 && s1 == o.s1 && s2 == o.s2 &&
           s3 == o.s3 && b0 == o.b0 && b1 == o.b1;
  }
};
struct TmapKeyHash {
  size_t operator()(const TmapKey& k) const {
    uint64_t h = reinterpret_cast<uint64_t>(k.ptr);
    auto mix = [&h](uint64_t v) { h ^= v + 0x9e3779b97f4a7c15ull + (h << 6) + (h >> 2); };
    mix(k.d0); mix(k.d1); mix(k.d2); mix(k.d3); mix(k.s1); mix(k.s2); mix(k.s3); mix(k.b0); mix(k.b1);
    return size_t(h);
  }
};

// bf16 4-D tensor map {d0 (contiguous), d1, d2, d3} with element strides s1..s3 and a
// 128B-swizzled box of {64, box_rows, 1, 1}.
int get_tmap(CUtensorMap* out, const void* ptr, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t d3,
                    uint64_t s1, uint64_t s2, uint64_t s3, uint32_t box_rows) {
  static std::mutex mu;
  static std::unordered_map<TmapKey, CUtensorMap, TmapKeyHash> cache;
  TmapKey key{ptr, d0, d1, d2, d3, s1, s2, s3, 64u, box_rows};
  {
    std::lock_guard<std::mutex> g(mu);
    auto it = cache.find(key);
    if (it != cache.end()) {
      *out = it->second;
      return 0;
    }
  }
  PFN_encodeTiled enc = get_encode_fn();
  if (!enc) return set_error(-2, "cuTensorMapEncodeTiled not available from the driver");
  // A batch dimension of extent 1 never advances, but its stride must still be legal.
  if (d2 <= 1) s2 = s1 * d1;
  if (d3 <= 1) s3 = s2 * (d2 ? d2 : 1);
  cuuint64_t dims[4] = {d0, d1, d2 ? d2 : 1, d3 ? d3 : 1};
  cuuint64_t strides[3] = {s1 * 2, s2 * 2, s3 * 2};
  cuuint32_t box[4] = {64, box_rows, 1, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) || (strides[0] & 15) || (strides[1] & 15) || (strides[2] & 15))
    return set_error(-3, "gemm operand pointer/strides must be 16-byte aligned");
  CUtensorMap m;
  CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error(-4, "cuTensorMapEncodeTiled failed (code " + std::to_string(int(r)) + ")");
  {
    std::lock_guard<std::mutex> g(mu);
    // a training job touches a few hundred distinct (pointer, shape) pairs; callers that stream fresh
    // tensors through the C ABI must not grow the table without bound (128 B per entry)
    if (cache.size() >= 8192) cache.clear();
    cache.emplace(key, m);
  }
  *out = m;
  return 0;
}


}  // namespace hstd
