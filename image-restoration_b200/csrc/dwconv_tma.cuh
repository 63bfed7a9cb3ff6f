// dwconv_tma.cuh — shared pieces of the TMA-staged depthwise 3x3 kernels (forward: dwconv_tma.cu, backward:
// dw_backward_tma.cu): work-item geometry, packed fp32x2 FMA, bf16 window loads from the staged halo tile.
#pragma once
#include <map>
#include <mutex>

#include "tc_common.cuh"

namespace b200ir {

// CTAs of `kernel` resident per SM at this dynamic shared-memory size (cached: the occupancy query is a host call)
template <typename K>
inline int resident_ctas(K kernel, int threads, size_t smem) {
  static std::mutex mu;
  static std::map<size_t, int> cache;
  std::lock_guard<std::mutex> lk(mu);
  auto it = cache.find(smem);
  if (it != cache.end()) return it->second;
  int n = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, threads, smem) != cudaSuccess || n < 1) {
    cudaGetLastError();
    n = 1;
  }
  cache[smem] = n;
  return n;
}

struct DwTmaParams {
  bf16* y;
  int64_t y_bs;
  const float* w;
  const float* bias;
  int Cout, H, W, TW, TH, nxg, nstrips, tiles_x, tiles_y, S;
  int nitems;
  int flip;  // 1: taps read in reverse order (= input gradient of the depthwise conv)
  int ncg;   // channels per work item (> 1 for planes narrower than 64 pixels: one TMA box carries ncg channels)
};

constexpr int DW_RS = 8;   // output rows per thread
constexpr int DW_XP = 8;   // left/right padding of the staged tile in pixels (keeps TMA coordinates 16B aligned)

#ifndef B200IR_GELU_TANH
#define B200IR_GELU_TANH 1
#endif
__device__ __forceinline__ float gelu_fast(float x) {
  // x * sigmoid(x * q(min(x^2, 64))), q = degree-2 fit in x^2 (max abs err 2.5e-5 vs erf GELU); clamping x^2 keeps the
  // odd polynomial monotone outside the fitted range (|x| > 8 -> exact 0 / x in fp32 anyway).
  const float x2 = fminf(x * x, 64.f);
#if B200IR_GELU_TANH
  // sigmoid(z) = 0.5 + 0.5 tanh(z/2): ONE MUFU op (tanh.approx) instead of ex2 + rcp -- the XU pipe was the busiest
  // pipe of the gate kernel (ncu r1j: 33.7 %).  Coefficients are q/2.
  float t = fmaf(-3.5151753375e-4f, x2, 3.7005650955e-2f);
  t = fmaf(t, x2, 0.79750787845f);
  float th;
  asm("tanh.approx.f32 %0, %1;" : "=f"(th) : "f"(t * x));
  const float hx = 0.5f * x;
  return fmaf(hx, th, hx);
#else
  float t = fmaf(1.0142652e-3f, x2, -1.0677574e-1f);   // -log2e * (-7.030350675e-4, 7.401130191e-2)
  t = fmaf(t, x2, -2.3011213f);                        // -log2e * 1.5950157569
  float e, r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(t * x));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.0f + e));
  return x * r;
#endif
}

// 2 * gelu(x) = x + x * tanh(x * q(x^2)/2) (same fit as gelu_fast; the caller folds the 0.5 into the other factor)
__device__ __forceinline__ float gelu2_fast(float x) {
  const float x2 = fminf(x * x, 64.f);
  float t = fmaf(-3.5151753375e-4f, x2, 3.7005650955e-2f);
  t = fmaf(t, x2, 0.79750787845f);
  float th;
  asm("tanh.approx.f32 %0, %1;" : "=f"(th) : "f"(t * x));
  return fmaf(x, th, x);
}

// packed fp32x2 FMA (sm_100): d = a * b + c on both halves, one issue slot
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
  uint64_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;"
      : "=l"(d)
      : "l"(*reinterpret_cast<uint64_t*>(&a)), "l"(*reinterpret_cast<uint64_t*>(&b)),
        "l"(*reinterpret_cast<uint64_t*>(&c)));
  return *reinterpret_cast<float2*>(&d);
}


__device__ __forceinline__ float bf_lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf_hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }

// gated variant: both planes at once, window element = (plane A value, plane B value) for packed FMA
__device__ __forceinline__ void dw_load_row2(uint32_t sa, uint32_t sb, float2* o) {
  uint32_t l0, r0, a0, b0, c0, d0, l1, r1, a1, b1, c1, d1;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(l0) : "r"(sa - 4));
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(a0), "=r"(b0), "=r"(c0), "=r"(d0) : "r"(sa));
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(r0) : "r"(sa + 16));
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(l1) : "r"(sb - 4));
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(a1), "=r"(b1), "=r"(c1), "=r"(d1) : "r"(sb));
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(r1) : "r"(sb + 16));
  o[0] = make_float2(bf_hi(l0), bf_hi(l1));
  o[1] = make_float2(bf_lo(a0), bf_lo(a1)); o[2] = make_float2(bf_hi(a0), bf_hi(a1));
  o[3] = make_float2(bf_lo(b0), bf_lo(b1)); o[4] = make_float2(bf_hi(b0), bf_hi(b1));
  o[5] = make_float2(bf_lo(c0), bf_lo(c1)); o[6] = make_float2(bf_hi(c0), bf_hi(c1));
  o[7] = make_float2(bf_lo(d0), bf_lo(d1)); o[8] = make_float2(bf_hi(d0), bf_hi(d1));
  o[9] = make_float2(bf_lo(r0), bf_lo(r1));
}

// load the 10-pixel window row (x-1 .. x+8) of one staged plane
__device__ __forceinline__ void dw_load_row(uint32_t saddr, float* o) {
  uint32_t l, r, a, b, c, d;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(l) : "r"(saddr - 4));
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(saddr));
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(r) : "r"(saddr + 16));
  o[0] = bf_hi(l);
  o[1] = bf_lo(a); o[2] = bf_hi(a); o[3] = bf_lo(b); o[4] = bf_hi(b);
  o[5] = bf_lo(c); o[6] = bf_hi(c); o[7] = bf_lo(d); o[8] = bf_hi(d);
  o[9] = bf_lo(r);
}


}  // namespace b200ir
