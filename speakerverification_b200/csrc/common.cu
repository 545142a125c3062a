// Error reporting, device queries and launch accounting for libsvb.
#include <stdarg.h>
#include <atomic>

#include "svb_internal.h"

namespace svb {

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

const char* last_error() { return g_err; }

int num_sms() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

// Kernel attributes (cudaFuncSetAttribute) and device allocations belong to ONE device: code that does something "once"
// asks per device ordinal, so that a process that drives several GPUs (one engine per device) works on all of them.
bool first_on_device(unsigned long long* seen) {
  int dev = 0;
  cudaGetDevice(&dev);
  const unsigned long long bit = 1ull << (dev & 63);
  if (__atomic_fetch_or(seen, bit, __ATOMIC_ACQ_REL) & bit) return false;
  return true;
}

int count_launch(int n) {
  g_launches += n;
  return 0;
}
long long launches() { return g_launches.load(); }
void reset_launches() { g_launches = 0; }

}  // namespace svb
