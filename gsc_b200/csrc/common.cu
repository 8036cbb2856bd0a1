// Error reporting shared by every entry point of libgsc_b200.
#include "common.cuh"
#include <atomic>

namespace gsc {

char* err_buf() {
  static thread_local char buf[512] = {0};
  return buf;
}

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(err_buf(), 512, fmt, ap);
  va_end(ap);
  return code;
}

}  // namespace gsc

extern "C" int gsc_version(void) { return 200; }   // 200: round 2 (gsc_gemm_args grew by the dot_* fields)

namespace gsc {
static std::atomic<int> g_sm_reserve{0};
int sm_reserve() { return g_sm_reserve.load(std::memory_order_relaxed); }
}  // namespace gsc

extern "C" int gsc_set_sm_reserve(int num_sms) {
  if (num_sms < 0 || num_sms > gsc::kNumSMs / 2) return gsc::fail(GSC_EINVAL, "gsc_set_sm_reserve: %d outside 0..%d", num_sms, gsc::kNumSMs / 2);
  gsc::g_sm_reserve.store(num_sms, std::memory_order_relaxed);
  return GSC_OK;
}
extern "C" const char* gsc_last_error(void) { return gsc::err_buf(); }
