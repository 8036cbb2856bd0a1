// Error reporting shared by every entry point of libgsc_b200.
#include "common.cuh"

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

extern "C" int gsc_version(void) { return 100; }
extern "C" const char* gsc_last_error(void) { return gsc::err_buf(); }
