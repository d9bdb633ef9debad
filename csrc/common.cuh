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

// ---------------------------------------------------------------- programmatic dependent launch
// Every kernel starts with HSTD_PDL_PROLOGUE(): it lets the *next* kernel of the stream be
// scheduled right away (its launch latency and prologue overlap our execution) and then waits
// until the *previous* kernel has completed and flushed, before touching global memory.
// launch_pdl() sets the matching launch attribute (HSTD_PDL=0 in the environment turns it off).
#if defined(__CUDACC__)
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
#define HSTD_PDL_PROLOGUE() do { hstd::pdl_launch_dependents(); hstd::pdl_wait(); } while (0)
#endif

bool pdl_enabled();

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

}  // namespace hstd
