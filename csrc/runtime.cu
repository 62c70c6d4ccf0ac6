// Library runtime: error string, device queries, version.  Everything else in csrc/ is a
// kernel family with its extern "C" launchers (declared in include/mamba_b200.h).
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "../include/mamba_b200.h"

namespace mb {
static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static unsigned long long g_launches = 0;

int check_launch(const char* what) {
  ++g_launches;  // one per kernel launch issued by this library (host-side counter)
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: launch failed: %s", what, cudaGetErrorString(e));
    return 1;
  }
  return 0;
}

bool pdl_enabled() {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("MAMBA_B200_PDL");
    on = (e != nullptr && e[0] == '0') ? 0 : 1;
  }
  return on == 1;
}

int num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (cached[dev] == 0) {
    int n = 0;
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    cached[dev] = n > 0 ? n : 148;
  }
  return cached[dev];
}
}  // namespace mb

extern "C" const char* mb_last_error(void) { return mb::g_err; }
extern "C" unsigned long long mb_launch_count(void) { return mb::g_launches; }
extern "C" int mb_version(void) { return 100; }
extern "C" int mb_num_sms(void) { return mb::num_sms(); }
