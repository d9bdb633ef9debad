// Cached creation of 4-D bf16 TMA tensor maps (128B swizzle, box {64, box_rows, 1, 1}).
#pragma once
#include <cuda.h>
#include <stdint.h>

namespace hstd {
// dims {d0 (contiguous), d1, d2, d3}, element strides s1..s3.
int get_tmap(CUtensorMap* out, const void* ptr, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t d3, uint64_t s1,
             uint64_t s2, uint64_t s3, uint32_t box_rows);
int num_sms();
}  // namespace hstd
