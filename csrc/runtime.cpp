// Error reporting, launch accounting, version.
#include "common.cuh"
#include "../include/kubeml_b200.h"
#include <cstring>
#include <mutex>
#include <unordered_set>

namespace kml {
static thread_local char g_err[1024] = "";
std::atomic<unsigned long long> g_launches{0};

void note_kernel(const void* fn) {
  // keyed by (kernel, device): the attribute is per device (see PerDeviceOnce in common.cuh)
  static std::mutex mu;
  static std::unordered_set<unsigned long long> seen;
  const unsigned long long key = reinterpret_cast<unsigned long long>(fn) ^
                                 (static_cast<unsigned long long>(current_device() + 1) << 56);
  std::lock_guard<std::mutex> lock(mu);
  if (seen.insert(key).second) {
    const char* e = getenv("KUBEML_UNIFORM_CARVEOUT");
    if (!(e && e[0] == '0'))
      cudaFuncSetAttribute(fn, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  }
}

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
}  // namespace kml

extern "C" {
const char* kml_last_error(void) { return kml::g_err; }
int kml_version(void) { return 100; }
unsigned long long kml_launch_count(void) { return kml::g_launches.load(); }
void kml_reset_launch_count(void) { kml::g_launches.store(0); }
int kml_num_sms(void) { return kml::num_sms(); }
}
