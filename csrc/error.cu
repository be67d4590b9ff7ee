// Error text storage for the C ABI (thread-local), version query.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

namespace b200dp {
namespace {
thread_local char g_err[1024] = "";
}

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int fail(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return B200DP_ERR;
}
}  // namespace b200dp

extern "C" const char* b200dp_last_error(void) { return b200dp::g_err; }
extern "C" int b200dp_version(void) { return 100; }
