// Error plumbing and small helpers shared by every translation unit of libhstd_b200.so.
#pragma once
#include <cuda_runtime.h>

#include <string>

namespace hstd {

// Records `msg` as the calling thread's last error and returns `code` (negative).
int set_error(int code, const std::string& msg);

inline int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_error(-6, std::string(what) + ": " + cudaGetErrorString(e));
  return 0;
}

}  // namespace hstd
