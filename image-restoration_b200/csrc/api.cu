// api.cu — version / error plumbing of the C ABI (include/b200ir.h).
#include <stdarg.h>

#include "common.cuh"

namespace b200ir {
static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
}  // namespace b200ir

extern "C" int b200ir_version(void) { return 100; }
extern "C" const char* b200ir_last_error_string(void) { return b200ir::g_err; }
