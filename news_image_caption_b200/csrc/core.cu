// Library-wide plumbing: error string, launch counter, device check.
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <mutex>
#include <string>
#include <vector>

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
static int g_pdl = -1;      // -1: read TNT_PDL from the environment on first use (default on)
bool pdl_enabled() {
  if (g_pdl < 0) {
    const char* e = getenv("TNT_PDL");
    g_pdl = (e && e[0] == '0') ? 0 : 1;
  }
  return g_pdl != 0;
}
static std::atomic<uint64_t> g_variants[V_COUNT];
static const char* const kVariantNames[V_COUNT] = {
    "gemm.tc.1cta", "gemm.tc.pair", "gemm.splitk", "gemm.simt",
    "xattn.tc", "xattn.tc.tileskip", "xattn.tc.weights", "xattn.simt", "xattn.step", "xattn.short",
    "xattn_bwd.tc", "xattn_bwd.simt",
    "dynconv.fwd.tc", "dynconv.fwd.simt", "dynconv.bwd.tc", "dynconv.bwd.simt", "dynconv.step.fast", "dynconv.step.simt",
    "gemm.ln_fused", "kv.beside"};
void variant_hit(Variant v) { g_variants[v].fetch_add(1, std::memory_order_relaxed); }
void count_launch(int n) { g_launches.fetch_add(static_cast<uint64_t>(n), std::memory_order_relaxed); }

// ---- named device-time regions (CUDA events on the launching stream), off by default.
// bench.py uses them to time the dominant kernel INSIDE the timed step loop.
struct ProfRec { std::string name; cudaEvent_t e0, e1; };
static bool g_prof_on = false;
static std::vector<ProfRec> g_prof;
static std::mutex g_prof_mu;

bool prof_enabled() { return g_prof_on; }
int prof_begin(const char* name, cudaStream_t st) {
  if (!g_prof_on) return -1;
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(st, &cap) != cudaSuccess || cap != cudaStreamCaptureStatusNone) return -1;   // graph capture: no timing events
  std::lock_guard<std::mutex> lk(g_prof_mu);
  ProfRec r;
  r.name = name;
  if (cudaEventCreate(&r.e0) != cudaSuccess || cudaEventCreate(&r.e1) != cudaSuccess) return -1;
  cudaEventRecord(r.e0, st);
  g_prof.push_back(r);
  return static_cast<int>(g_prof.size()) - 1;
}
void prof_end(int id, cudaStream_t st) {
  if (id < 0) return;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  if (id < static_cast<int>(g_prof.size())) cudaEventRecord(g_prof[id].e1, st);
}

}  // namespace tnt

extern "C" int tnt_version(void) { return 100; }
extern "C" const char* tnt_last_error(void) { return tnt::g_err; }
extern "C" uint64_t tnt_launch_count(void) { return tnt::g_launches.load(); }

// launches of dispatcher variant `name` since load (-1: unknown name); name == NULL resets every counter
extern "C" int64_t tnt_variant_count(const char* name) {
  if (!name) {
    for (auto& v : tnt::g_variants) v.store(0);
    return 0;
  }
  for (int i = 0; i < tnt::V_COUNT; ++i)
    if (!strcmp(name, tnt::kVariantNames[i])) return static_cast<int64_t>(tnt::g_variants[i].load());
  return -1;
}

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

extern "C" void tnt_debug_pdl(int on) { tnt::g_pdl = on ? 1 : 0; }

extern "C" void tnt_prof_enable(int on) {
  std::lock_guard<std::mutex> lk(tnt::g_prof_mu);
  tnt::g_prof_on = on != 0;
}
extern "C" void tnt_prof_reset(void) {
  std::lock_guard<std::mutex> lk(tnt::g_prof_mu);
  for (auto& r : tnt::g_prof) {
    cudaEventDestroy(r.e0);
    cudaEventDestroy(r.e1);
  }
  tnt::g_prof.clear();
}
extern "C" int tnt_prof_read(const char* name, double* ms_total, int* count) {
  std::lock_guard<std::mutex> lk(tnt::g_prof_mu);
  double tot = 0;
  int n = 0;
  for (auto& r : tnt::g_prof) {
    if (name && r.name != name) continue;
    if (cudaEventSynchronize(r.e1) != cudaSuccess) continue;
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, r.e0, r.e1) == cudaSuccess) {
      tot += ms;
      ++n;
    }
  }
  if (ms_total) *ms_total = tot;
  if (count) *count = n;
  return TNT_OK;
}
