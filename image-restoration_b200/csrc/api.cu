// api.cu — version / error plumbing of the C ABI (include/b200ir.h).
#include <stdarg.h>

#include <map>
#include <mutex>
#include <utility>

#include "common.cuh"

namespace b200ir {
static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int num_sms() {
  static std::mutex mu;
  static int cache[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  std::lock_guard<std::mutex> lk(mu);
  if (cache[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n < 1) n = 148;
    cache[dev] = n;
  }
  return cache[dev];
}

void ensure_dyn_smem(const void* func, int bytes) {
  static std::mutex mu;
  static std::map<std::pair<const void*, int>, int> done;
  int dev = 0;
  cudaGetDevice(&dev);
  std::lock_guard<std::mutex> lk(mu);
  int& cur = done[std::make_pair(func, dev)];
  if (cur >= bytes) return;
  cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  cur = bytes;
}
}  // namespace b200ir

extern "C" int b200ir_version(void) { return 100; }
extern "C" const char* b200ir_last_error_string(void) { return b200ir::g_err; }
