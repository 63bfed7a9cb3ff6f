// common.cuh — shared helpers for libb200ir (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/b200ir.h"

namespace b200ir {

typedef __nv_bfloat16 bf16;

// thread-local error message (b200ir_last_error_string)
void set_error(const char* fmt, ...);

#define B200IR_CHECK_ARG(cond, ...)                  \
  do {                                               \
    if (!(cond)) {                                   \
      ::b200ir::set_error(__VA_ARGS__);              \
      return B200IR_E_BADARG;                        \
    }                                                \
  } while (0)

#define B200IR_UNSUPPORTED(cond, ...)                \
  do {                                               \
    if (!(cond)) {                                   \
      ::b200ir::set_error(__VA_ARGS__);              \
      return B200IR_E_UNSUPPORTED;                   \
    }                                                \
  } while (0)

inline int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: %s", what, cudaGetErrorString(e));
    return (int)e;
  }
  return 0;
}

// ---- dtype-generic element access (dtype is kernel-uniform: no divergence) ----
template <typename T> __device__ __forceinline__ float to_f(T v);
template <> __device__ __forceinline__ float to_f<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f<bf16>(bf16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }
template <> __device__ __forceinline__ bf16 from_f<bf16>(float v) { return __float2bfloat16_rn(v); }

__device__ __forceinline__ float ld_any(const void* p, int dtype, int64_t i) {
  return dtype == B200IR_BF16 ? __bfloat162float(((const bf16*)p)[i]) : ((const float*)p)[i];
}
__device__ __forceinline__ void st_any(void* p, int dtype, int64_t i, float v) {
  if (dtype == B200IR_BF16) ((bf16*)p)[i] = __float2bfloat16_rn(v);
  else ((float*)p)[i] = v;
}

__device__ __forceinline__ float gelu_erf(float x) {
  // F.gelu default (approximate='none'): 0.5*x*(1+erf(x/sqrt(2)))   (Restormer.py:91)
  return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f));
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// SM count of the CURRENT device (one cudaDeviceGetAttribute per device, cached): grids are sized in multiples of it
int num_sms();
// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) once per (function, device): the attribute is per device, a
// process-wide std::call_once would leave a second GPU of the same process at the 48 KB default
void ensure_dyn_smem(const void* func, int bytes);

inline int dtype_size(int dt) { return dt == B200IR_BF16 ? 2 : 4; }
inline bool dtype_ok(int dt) { return dt == B200IR_F32 || dt == B200IR_BF16; }

}  // namespace b200ir
