// Library-wide plumbing: error string, launch counter, device check.
#include <stdarg.h>

#include <atomic>

#include "common.cuh"

namespace tnt {

static thread_local char g_err[1024] = "";
static std::atomic<uint64_t> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add(static_cast<uint64_t>(n), std::memory_order_relaxed); }

}  // namespace tnt

extern "C" int tnt_version(void) { return 100; }
extern "C" const char* tnt_last_error(void) { return tnt::g_err; }
extern "C" uint64_t tnt_launch_count(void) { return tnt::g_launches.load(); }

extern "C" int tnt_device_check(void) {
  int dev = 0, major = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) {
    tnt::set_error("no CUDA device");
    return TNT_ERR_NODEVICE;
  }
  cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
  if (major != 10) {
    tnt::set_error("tnt_b200 is built for sm_100a only; device has compute capability %d.x", major);
    return TNT_ERR_NODEVICE;
  }
  return TNT_OK;
}
