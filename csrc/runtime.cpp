// Error reporting, launch accounting, version.
#include "common.cuh"
#include "../include/kubeml_b200.h"
#include <cstring>

namespace kml {
static thread_local char g_err[1024] = "";
std::atomic<unsigned long long> g_launches{0};

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
}
