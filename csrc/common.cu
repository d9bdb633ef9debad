#include "common.cuh"

#include <cstdlib>

#include "../include/hstd_b200.h"

namespace hstd {
static thread_local std::string g_last_error;
int set_error(int code, const std::string& msg) {
  g_last_error = msg;
  return code;
}
bool pdl_enabled() {
  static const bool on = [] {
    const char* e = getenv("HSTD_PDL");
    return !(e && e[0] == '0');
  }();
  return on;
}
}  // namespace hstd

extern "C" HSTD_API const char* hstd_last_error(void) { return hstd::g_last_error.c_str(); }
extern "C" HSTD_API int hstd_version(void) { return 1; }
