// Library-level plumbing: last-error string, version, device queries.
#include "common.cuh"
#include <stdarg.h>
#include <string.h>

static thread_local char g_err[512] = "";
unsigned long long g_clsm_launches = 0;
thread_local void* g_clsm_marks[2] = {nullptr, nullptr};      // clsm_debug_mark_conv_bwd: events for the next clsm_conv_pool_bwd

void clsm_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int clsm_sm_count() {
  static int cached[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

extern "C" {
int clsm_version(void) { return 100; }
const char* clsm_last_error(void) { return g_err; }
int clsm_num_sms(void) { return clsm_sm_count(); }
unsigned long long clsm_launch_count(void) { return g_clsm_launches; }
void clsm_debug_mark_conv_bwd(void* ev_after_dw, void* ev_before_dx) { g_clsm_marks[0] = ev_after_dw; g_clsm_marks[1] = ev_before_dx; }
}
