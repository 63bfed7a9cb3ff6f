// moce_ops.cu — the MoCE-IR decoder pieces that are not a Restormer block: the FFT expert's per-patch
// spectral product, the small elementwise gates around it, and the global average pools of the router / frequency
// embedding.  Reference: /root/reference/moce_ir.py FFTAttention :373-422, ModExpert.process :548-560,
// RoutingFunction.gate (AdaptiveAvgPool2d) :703-707, FrequencyEmbedding.forward :1068-1075.
//
// FFTAttention computes, per p x p patch and channel, irfft2(rfft2(q) * rfft2(k)): by the convolution theorem that is
// the CIRCULAR CONVOLUTION of the two real patches,
//     out[i, j] = sum_{a, b} q[a, b] * k[(i - a) mod p, (j - b) mod p],
// evaluated here directly from shared memory in fp32 (p in {4, 8, 16, 32}: at most 1024 x 1024 MACs per patch, no FFT
// library, no complex arithmetic, exact zero padding of ragged planes like F.pad at :391).  The two gradients are
// circular cross-correlations and run on the same kernel (mode 1): dq = corr(dout, k), dk = corr(dout, q).
#include "common.cuh"
#include <cooperative_groups.h>

namespace b200ir {

// mode 0: y = a (*) b (circular convolution);  mode 1: y[s] = sum_i a[i] * b[(i - s) mod p]  (correlation of a with b)
template <typename T>
__global__ void __launch_bounds__(256)
circconv_patch_kernel(const T* __restrict__ a, int64_t a_bs, const T* __restrict__ b, int64_t b_bs, T* __restrict__ y,
                      int64_t y_bs, int C, int H, int W, int p, int npx, int mode) {
  extern __shared__ float sm[];
  float* sa = sm;
  float* sb = sm + p * p;
  const int patch = blockIdx.x, c = blockIdx.y, n = blockIdx.z;
  const int py = patch / npx, px = patch - py * npx;
  const int y0 = py * p, x0 = px * p, pp = p * p, msk = p - 1;
  const int lg = 31 - __clz(p);
  const int64_t plane = (int64_t)c * H * W;
  for (int i = threadIdx.x; i < pp; i += blockDim.x) {
    const int r = i >> lg, q = i & msk;
    const bool in = (y0 + r) < H && (x0 + q) < W;
    const int64_t off = plane + (int64_t)(y0 + r) * W + (x0 + q);
    sa[i] = in ? to_f<T>(a[(int64_t)n * a_bs + off]) : 0.f;
    sb[i] = in ? to_f<T>(b[(int64_t)n * b_bs + off]) : 0.f;
  }
  __syncthreads();
  for (int o = threadIdx.x; o < pp; o += blockDim.x) {
    const int oi = o >> lg, oj = o & msk;
    if (y0 + oi >= H || x0 + oj >= W) continue;
    float acc = 0.f;
    if (mode == 0) {
      for (int r = 0; r < p; ++r) {
        const float* ar = sa + r * p;
        const float* br = sb + ((oi - r) & msk) * p;
#pragma unroll 4
        for (int q = 0; q < p; ++q) acc = fmaf(ar[q], br[(oj - q) & msk], acc);
      }
    } else {
      for (int r = 0; r < p; ++r) {
        const float* ar = sa + r * p;
        const float* br = sb + ((r - oi) & msk) * p;
#pragma unroll 4
        for (int q = 0; q < p; ++q) acc = fmaf(ar[q], br[(q - oj) & msk], acc);
      }
    }
    y[(int64_t)n * y_bs + plane + (int64_t)(y0 + oi) * W + (x0 + oj)] = from_f<T>(acc);
  }
}

// Small patches (p = 4 / 8: the two cheapest experts, 6-24 channels at full resolution): the kernel above launches one
// 16- / 64-thread CTA per patch (165 us for 12 MB at cfg3).  Here a 256-thread CTA takes 256 / p^2 patches that are neighbours in
// x (one p-row band, 64 B - 128 B contiguous per row), a thread owns one output pixel.
template <int P, typename T>
__global__ void __launch_bounds__(256)
circconv_small_kernel(const T* __restrict__ a, int64_t a_bs, const T* __restrict__ b, int64_t b_bs, T* __restrict__ y,
                      int64_t y_bs, int C, int H, int W, int ngx, int mode) {
  constexpr int NP = 256 / (P * P), TW = NP * P, M = P - 1;
  __shared__ float sa[P][TW + 1], sb[P][TW + 1];
  const int grp = blockIdx.x % ngx, py = blockIdx.x / ngx, c = blockIdx.y, n = blockIdx.z;
  const int y0 = py * P, x0 = grp * TW;
  const int64_t plane = (int64_t)c * H * W;
  {
    const int r = threadIdx.x / TW, q = threadIdx.x % TW;   // 256 = P * TW
    const bool in = (y0 + r) < H && (x0 + q) < W;
    const int64_t off = plane + (int64_t)(y0 + r) * W + (x0 + q);
    sa[r][q] = in ? to_f<T>(a[(int64_t)n * a_bs + off]) : 0.f;
    sb[r][q] = in ? to_f<T>(b[(int64_t)n * b_bs + off]) : 0.f;
  }
  __syncthreads();
  const int oi = threadIdx.x / TW, col = threadIdx.x % TW, pb = (col / P) * P, oj = col % P;
  if (y0 + oi >= H || x0 + col >= W) return;
  float acc = 0.f;
#pragma unroll
  for (int r = 0; r < P; ++r) {
    const int br = mode == 0 ? ((oi - r) & M) : ((r - oi) & M);
#pragma unroll
    for (int q = 0; q < P; ++q) {
      const int bq = mode == 0 ? ((oj - q) & M) : ((q - oj) & M);
      acc = fmaf(sa[r][pb + q], sb[br][pb + bq], acc);
    }
  }
  y[(int64_t)n * y_bs + plane + (int64_t)(y0 + oi) * W + (x0 + col)] = from_f<T>(acc);
}

// Register-tiled form for p = 16 / 32 (the two large experts; measured on B200 the shared-memory-bound kernel above ran
// the p = 32 expert of cfg3 at 3.7 TMAC/s = 10 % of the FP32 peak and was half of the whole MoCE-IR forward): one WARP per
// (patch, channel), a thread owns one output row (p = 32) or half of one (p = 16) in registers.  Per row a of the first
// operand it loads the matching row of the second operand once (row stride p+1: conflict-free) and then runs p x NO
// fully unrolled FMAs whose register indices (j -+ b) mod p are compile-time: ~40 shared loads per 1024 FMAs.
template <int P, int MODE, typename T>
__global__ void __launch_bounds__(128)
circconv_warp_kernel(const T* __restrict__ a, int64_t a_bs, const T* __restrict__ b, int64_t b_bs, T* __restrict__ y,
                     int64_t y_bs, int C, int H, int W, int npx, int npatch, int64_t total) {
  constexpr int SEG = 32 / P, NO = P / SEG, M = P - 1;
  __shared__ float sa[4][P * P];
  __shared__ float sb[4][P * (P + 1)];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t item = (int64_t)blockIdx.x * 4 + warp;
  if (item >= total) return;                       // whole warp
  const int patch = (int)(item % npatch);
  const int64_t pc = item / npatch;
  const int c = (int)(pc % C), n = (int)(pc / C);
  const int py = patch / npx, px = patch - py * npx;
  const int y0 = py * P, x0 = px * P;
  const int64_t plane = (int64_t)c * H * W;
  for (int i = lane; i < P * P; i += 32) {
    const int r = i / P, q = i % P;
    const bool in = (y0 + r) < H && (x0 + q) < W;
    const int64_t off = plane + (int64_t)(y0 + r) * W + (x0 + q);
    sa[warp][i] = in ? to_f<T>(a[(int64_t)n * a_bs + off]) : 0.f;
    sb[warp][r * (P + 1) + q] = in ? to_f<T>(b[(int64_t)n * b_bs + off]) : 0.f;
  }
  __syncwarp();
  const int row = lane % P, seg = lane / P;
  float acc[NO];
#pragma unroll
  for (int j = 0; j < NO; ++j) acc[j] = 0.f;
  for (int r = 0; r < P; ++r) {
    // second-operand row this output row meets at first-operand row r, rotated so that column indices are compile-time
    const int br = MODE == 0 ? ((row - r) & M) : ((r - row) & M);
    const int rot = MODE == 0 ? seg * NO : (P - seg * NO) & M;
    float kr[P];
#pragma unroll
    for (int q = 0; q < P; ++q) kr[q] = sb[warp][br * (P + 1) + ((q + rot) & M)];
    const float* ar = sa[warp] + r * P;
#pragma unroll
    for (int q = 0; q < P; ++q) {
      const float av = ar[q];
#pragma unroll
      for (int j = 0; j < NO; ++j) acc[j] = fmaf(av, MODE == 0 ? kr[(j - q) & M] : kr[(q - j) & M], acc[j]);
    }
  }
  if (y0 + row < H) {
    T* yp = y + (int64_t)n * y_bs + plane + (int64_t)(y0 + row) * W + x0 + seg * NO;
#pragma unroll
    for (int j = 0; j < NO; ++j)
      if (x0 + seg * NO + j < W) yp[j] = from_f<T>(acc[j]);
  }
}

// elementwise over NCHW views (batch stride free): op 0 y = a*b, 1 y = a*silu(b), 2 y = a+b, 3 y = gelu_erf(a),
// 4 y = a * (1 - ...) unused; 5: y = b * (sigmoid(a) * (1 + a * (1 - sigmoid(a))))  (d silu(a)/da times b)
template <typename T>
__global__ void ew_kernel(const T* __restrict__ a, int64_t a_bs, const T* __restrict__ b, int64_t b_bs, T* __restrict__ y,
                          int64_t y_bs, int64_t per, int op) {
  const int n = blockIdx.y;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < per; i += (int64_t)gridDim.x * blockDim.x) {
    const float av = to_f<T>(a[(int64_t)n * a_bs + i]);
    const float bv = b ? to_f<T>(b[(int64_t)n * b_bs + i]) : 0.f;
    float r;
    if (op == 0) r = av * bv;
    else if (op == 1) r = av * (bv / (1.f + __expf(-bv)));
    else if (op == 2) r = av + bv;
    else if (op == 3) r = gelu_erf(av);
    else {
      const float s = 1.f / (1.f + __expf(-av));
      r = bv * (s * (1.f + av * (1.f - s)));
    }
    y[(int64_t)n * y_bs + i] = from_f<T>(r);
  }
}

// y[b] = (accumulate ? y[b] : 0) + g[b] * x[b]   (per-sample gate of an expert output, AdapterLayer.forward :672-676)
template <typename T>
__global__ void scale_rows_kernel(const T* __restrict__ x, int64_t x_bs, const float* __restrict__ g, T* __restrict__ y,
                                  int64_t y_bs, int64_t per, int accumulate) {
  const int n = blockIdx.y;
  const float gv = g[n];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < per; i += (int64_t)gridDim.x * blockDim.x) {
    float r = gv * to_f<T>(x[(int64_t)n * x_bs + i]);
    if (accumulate) r += to_f<T>(y[(int64_t)n * y_bs + i]);
    y[(int64_t)n * y_bs + i] = from_f<T>(r);
  }
}

// out[b, c] = mean over HW of x[b, c, :]  (fp32 accumulation, one warp per plane, fixed order)
template <typename T>
__global__ void spatial_mean_kernel(const T* __restrict__ x, int64_t x_bs, float* __restrict__ out, int B, int C,
                                    int64_t HW, int gelu) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int c = warp % C, n = warp / C;
  if (n >= B) return;
  const T* p = x + (int64_t)n * x_bs + (int64_t)c * HW;
  float s = 0.f;
  for (int64_t i = lane; i < HW; i += 32) {
    const float v = to_f<T>(p[i]);
    s += gelu ? gelu_erf(v) : v;
  }
  s = warp_sum(s);
  if (lane == 0) out[(int64_t)n * C + c] = s / (float)HW;
}

// Large planes: one CTA per plane, 16-byte loads (8 bf16 / 4 fp32 per lane and step), fixed-order fold (lane sums ->
// warp butterfly -> 8 warp sums added in order): one warp per 65 536-pixel plane ran at 0.27 TB/s (184 us at cfg3).
template <typename T>
__global__ void __launch_bounds__(256)
spatial_mean_cta_kernel(const T* __restrict__ x, int64_t x_bs, float* __restrict__ out, int C, int64_t HW, int gelu) {
  constexpr int V = 16 / (int)sizeof(T);
  __shared__ float red[8];
  const int c = blockIdx.x, n = blockIdx.y;
  const T* p = x + (int64_t)n * x_bs + (int64_t)c * HW;
  float s = 0.f;
  for (int64_t i = (int64_t)threadIdx.x * V; i < HW; i += 256 * V) {
    const uint4 q = __ldg(reinterpret_cast<const uint4*>(p + i));
    const T* e = reinterpret_cast<const T*>(&q);
#pragma unroll
    for (int j = 0; j < V; ++j) {
      const float v = to_f<T>(e[j]);
      s += gelu ? gelu_erf(v) : v;
    }
  }
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += red[w];
    out[(int64_t)n * C + c] = t / (float)HW;
  }
}

}  // namespace b200ir

namespace b200ir {
bool circconv_dft_launch(const void* a, int64_t a_bs, const void* b, int64_t b_bs, void* y, int64_t y_bs, int B, int C, int H,
                         int W, int patch, int mode, cudaStream_t st, int* rc);   // circconv_dft.cu
}
using namespace b200ir;

extern "C" int b200ir_circconv_patch(const void* a, int64_t a_bstride, const void* b, int64_t b_bstride, void* y,
                                     int64_t y_bstride, int dtype, int B, int C, int H, int W, int patch, int mode,
                                     b200ir_stream_t stream) {
  B200IR_CHECK_ARG(a && b && y && B > 0 && C > 0 && H > 0 && W > 0, "circconv_patch: bad args");
  B200IR_CHECK_ARG(dtype_ok(dtype) && (mode == 0 || mode == 1), "circconv_patch: bad dtype / mode");
  B200IR_UNSUPPORTED(patch >= 2 && patch <= 32 && (patch & (patch - 1)) == 0, "circconv_patch: patch %d unsupported (2..32, power of 2)", patch);
  B200IR_UNSUPPORTED(B <= 65535 && C <= 65535, "circconv_patch: B, C <= 65535");
  const int npx = (W + patch - 1) / patch, npy = (H + patch - 1) / patch;
  if (patch >= 16 && dtype == B200IR_BF16) {   // tensor-core DFT (circconv_dft.cu); falls through for unaligned planes
    int rc = 0;
    if (circconv_dft_launch(a, a_bstride, b, b_bstride, y, y_bstride, B, C, H, W, patch, mode, (cudaStream_t)stream, &rc))
      return rc;
  }
  if (patch >= 16) {
    const int64_t total = (int64_t)B * C * npx * npy;
    B200IR_UNSUPPORTED((total + 3) / 4 < (1ll << 31), "circconv_patch: too many patches");
    const unsigned g = (unsigned)((total + 3) / 4);
#define B200IR_CC(P_, T_)                                                                                            \
  do {                                                                                                               \
    if (mode == 0)                                                                                                   \
      circconv_warp_kernel<P_, 0, T_><<<g, 128, 0, (cudaStream_t)stream>>>((const T_*)a, a_bstride, (const T_*)b,     \
                                                                           b_bstride, (T_*)y, y_bstride, C, H, W, npx, \
                                                                           npx * npy, total);                        \
    else                                                                                                             \
      circconv_warp_kernel<P_, 1, T_><<<g, 128, 0, (cudaStream_t)stream>>>((const T_*)a, a_bstride, (const T_*)b,     \
                                                                           b_bstride, (T_*)y, y_bstride, C, H, W, npx, \
                                                                           npx * npy, total);                        \
  } while (0)
    if (patch == 32) {
      if (dtype == B200IR_BF16) B200IR_CC(32, bf16); else B200IR_CC(32, float);
    } else {
      if (dtype == B200IR_BF16) B200IR_CC(16, bf16); else B200IR_CC(16, float);
    }
#undef B200IR_CC
    return check_launch("circconv_patch_warp");
  }
  if (patch == 4 || patch == 8) {
    const int tw = 256 / patch, ngx = (W + tw - 1) / tw;
    const dim3 g2((unsigned)(ngx * npy), (unsigned)C, (unsigned)B);
#define B200IR_CS(P_, T_)                                                                                              \
  circconv_small_kernel<P_, T_><<<g2, 256, 0, (cudaStream_t)stream>>>((const T_*)a, a_bstride, (const T_*)b, b_bstride, \
                                                                      (T_*)y, y_bstride, C, H, W, ngx, mode)
    if (patch == 4) {
      if (dtype == B200IR_BF16) B200IR_CS(4, bf16); else B200IR_CS(4, float);
    } else {
      if (dtype == B200IR_BF16) B200IR_CS(8, bf16); else B200IR_CS(8, float);
    }
#undef B200IR_CS
    return check_launch("circconv_small");
  }
  const dim3 grid((unsigned)(npx * npy), (unsigned)C, (unsigned)B);
  const size_t smem = (size_t)2 * patch * patch * sizeof(float);
  const int threads = patch * patch >= 256 ? 256 : (patch * patch < 32 ? 32 : patch * patch);
  if (dtype == B200IR_BF16)
    circconv_patch_kernel<bf16><<<grid, threads, smem, (cudaStream_t)stream>>>((const bf16*)a, a_bstride, (const bf16*)b,
                                                                              b_bstride, (bf16*)y, y_bstride, C, H, W,
                                                                              patch, npx, mode);
  else
    circconv_patch_kernel<float><<<grid, threads, smem, (cudaStream_t)stream>>>((const float*)a, a_bstride,
                                                                               (const float*)b, b_bstride, (float*)y,
                                                                               y_bstride, C, H, W, patch, npx, mode);
  return check_launch("circconv_patch");
}

extern "C" int b200ir_ew(const void* a, int64_t a_bstride, const void* b, int64_t b_bstride, void* y, int64_t y_bstride,
                         int dtype, int B, int64_t per_sample, int op, b200ir_stream_t stream) {
  B200IR_CHECK_ARG(a && y && B > 0 && B <= 65535 && per_sample > 0 && op >= 0 && op <= 5, "ew: bad args");
  B200IR_CHECK_ARG(b || op == 3, "ew: second operand missing");
  B200IR_CHECK_ARG(dtype_ok(dtype), "ew: bad dtype");
  int64_t gx = (per_sample + 255) / 256;
  if (gx > num_sms() * 8) gx = num_sms() * 8;
  const dim3 grid((unsigned)gx, (unsigned)B);
  if (dtype == B200IR_BF16)
    ew_kernel<bf16><<<grid, 256, 0, (cudaStream_t)stream>>>((const bf16*)a, a_bstride, (const bf16*)b, b_bstride, (bf16*)y,
                                                            y_bstride, per_sample, op);
  else
    ew_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>((const float*)a, a_bstride, (const float*)b, b_bstride,
                                                             (float*)y, y_bstride, per_sample, op);
  return check_launch("ew");
}

extern "C" int b200ir_scale_rows(const void* x, int64_t x_bstride, const float* g, void* y, int64_t y_bstride, int dtype,
                                 int B, int64_t per_sample, int accumulate, b200ir_stream_t stream) {
  B200IR_CHECK_ARG(x && g && y && B > 0 && B <= 65535 && per_sample > 0, "scale_rows: bad args");
  B200IR_CHECK_ARG(dtype_ok(dtype), "scale_rows: bad dtype");
  int64_t gx = (per_sample + 255) / 256;
  if (gx > num_sms() * 8) gx = num_sms() * 8;
  const dim3 grid((unsigned)gx, (unsigned)B);
  if (dtype == B200IR_BF16)
    scale_rows_kernel<bf16><<<grid, 256, 0, (cudaStream_t)stream>>>((const bf16*)x, x_bstride, g, (bf16*)y, y_bstride,
                                                                    per_sample, accumulate);
  else
    scale_rows_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>((const float*)x, x_bstride, g, (float*)y, y_bstride,
                                                                     per_sample, accumulate);
  return check_launch("scale_rows");
}

extern "C" int b200ir_spatial_mean(const void* x, int64_t x_bstride, int dtype, float* out, int B, int C, int64_t HW,
                                   int gelu, b200ir_stream_t stream) {
  B200IR_CHECK_ARG(x && out && B > 0 && B <= 65535 && C > 0 && HW > 0, "spatial_mean: bad args");
  B200IR_CHECK_ARG(dtype_ok(dtype), "spatial_mean: bad dtype");
  const int esz = dtype == B200IR_BF16 ? 2 : 4;
  if (HW >= 4096 && C <= 65535 && (HW * esz) % 16 == 0 && (x_bstride * esz) % 16 == 0 && ((uintptr_t)x & 15) == 0) {
    const dim3 g2((unsigned)C, (unsigned)B);
    if (dtype == B200IR_BF16)
      spatial_mean_cta_kernel<bf16><<<g2, 256, 0, (cudaStream_t)stream>>>((const bf16*)x, x_bstride, out, C, HW, gelu);
    else
      spatial_mean_cta_kernel<float><<<g2, 256, 0, (cudaStream_t)stream>>>((const float*)x, x_bstride, out, C, HW, gelu);
    return check_launch("spatial_mean");
  }
  const unsigned grid = (unsigned)(((int64_t)B * C + 7) / 8);   // one warp per (sample, channel) plane, 8 warps per CTA
  if (dtype == B200IR_BF16)
    spatial_mean_kernel<bf16><<<grid, 256, 0, (cudaStream_t)stream>>>((const bf16*)x, x_bstride, out, B, C, HW, gelu);
  else
    spatial_mean_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>((const float*)x, x_bstride, out, B, C, HW, gelu);
  return check_launch("spatial_mean");
}

// =====================================================================================================
// Training mode of the MoCE-IR adapter (SparseDispatcher, moce_ir.py:71-141, 659-667): samples are gathered per expert,
// the expert outputs are scattered back weighted by their gates, and the global pools need gradients.
// =====================================================================================================
namespace b200ir {

// y[dst(j)] (+)= (scale ? scale[j] : 1) * x[src(j)],  j < n rows of `per` elements (row = one sample of an NCHW tensor);
// src / dst index arrays may be NULL (identity).  dispatch = gather by src, combine = scatter-accumulate by dst; the
// indices of one call are distinct (a sample meets an expert once), so there is no write conflict.
template <typename T>
__global__ void rows_copy_kernel(const T* __restrict__ x, int64_t x_rs, T* __restrict__ y, int64_t y_rs,
                                 const int* __restrict__ src, const int* __restrict__ dst, const float* __restrict__ scale,
                                 int64_t per, int accumulate) {
  const int j = blockIdx.y;
  const int64_t so = (int64_t)(src ? src[j] : j) * x_rs, dof = (int64_t)(dst ? dst[j] : j) * y_rs;
  const float g = scale ? scale[j] : 1.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < per; i += (int64_t)gridDim.x * blockDim.x) {
    float r = g * to_f<T>(x[so + i]);
    if (accumulate) r += to_f<T>(y[dof + i]);
    y[dof + i] = from_f<T>(r);
  }
}

// out[j] = sum_i a[ia(j), i] * b[ib(j), i]  (fp32): the gate gradient of the weighted combine.  A row is one sample of
// an NCHW tensor (MBs) and a call has 1-3 rows, so one row is spread over a cluster of ROWDOT_CL CTAs; the CTA partials
// are folded by rank 0 over distributed shared memory in rank order (fixed order, no workspace, no atomics).
constexpr int ROWDOT_CL = 8, ROWDOT_T = 512;
template <typename T, bool VEC>
__global__ void __launch_bounds__(ROWDOT_T)
rowdot_kernel(const T* __restrict__ a, int64_t a_rs, const int* __restrict__ ia, const T* __restrict__ b, int64_t b_rs,
              const int* __restrict__ ib, float* __restrict__ out, int64_t per) {
  namespace cg = cooperative_groups;
  cg::cluster_group cl = cg::this_cluster();
  __shared__ float red[ROWDOT_T / 32];
  __shared__ float part;
  const int j = blockIdx.y, r = (int)cl.block_rank();
  const T* pa = a + (int64_t)(ia ? ia[j] : j) * a_rs;
  const T* pb = b + (int64_t)(ib ? ib[j] : j) * b_rs;
  float s = 0.f;
  if constexpr (VEC) {
    constexpr int V = 16 / (int)sizeof(T);
    const int64_t nv = per / V;
    const uint4* va = reinterpret_cast<const uint4*>(pa);
    const uint4* vb = reinterpret_cast<const uint4*>(pb);
    float s1 = 0.f;
    for (int64_t i = (int64_t)r * ROWDOT_T + threadIdx.x; i < nv; i += (int64_t)ROWDOT_CL * ROWDOT_T) {
      const uint4 qa = __ldg(va + i), qb = __ldg(vb + i);
      const T* ea = reinterpret_cast<const T*>(&qa);
      const T* eb = reinterpret_cast<const T*>(&qb);
#pragma unroll
      for (int e = 0; e < V; e += 2) {
        s = fmaf(to_f<T>(ea[e]), to_f<T>(eb[e]), s);
        s1 = fmaf(to_f<T>(ea[e + 1]), to_f<T>(eb[e + 1]), s1);
      }
    }
    s += s1;
  } else {
    for (int64_t i = (int64_t)r * ROWDOT_T + threadIdx.x; i < per; i += (int64_t)ROWDOT_CL * ROWDOT_T)
      s = fmaf(to_f<T>(pa[i]), to_f<T>(pb[i]), s);
  }
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < ROWDOT_T / 32; ++w) t += red[w];
    part = t;
  }
  cl.sync();
  if (r == 0 && threadIdx.x == 0) {
    float t = 0.f;
    for (int q = 0; q < ROWDOT_CL; ++q) t += *cl.map_shared_rank(&part, q);
    out[j] = t;
  }
  cl.sync();   // a CTA's shared memory must outlive rank 0's reads
}

// backward of spatial_mean: dx[b, c, i] = g[b, c] / HW * (gelu ? gelu'(x[b, c, i]) : 1)
template <typename T>
__global__ void spatial_mean_bwd_kernel(const float* __restrict__ g, const T* __restrict__ x, int64_t x_bs,
                                        T* __restrict__ dx, int64_t dx_bs, int C, int64_t HW, int gelu) {
  const int c = blockIdx.y, n = blockIdx.z;
  const float gv = g[(int64_t)n * C + c] / (float)HW;
  const T* xp = x ? x + (int64_t)n * x_bs + (int64_t)c * HW : nullptr;
  T* dp = dx + (int64_t)n * dx_bs + (int64_t)c * HW;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < HW; i += (int64_t)gridDim.x * blockDim.x) {
    float r = gv;
    if (gelu) {
      const float v = to_f<T>(xp[i]);
      // d/dv [v * Phi(v)] = Phi(v) + v * phi(v)
      r *= 0.5f * (1.f + erff(v * 0.70710678118654752f)) + v * 0.3989422804014327f * __expf(-0.5f * v * v);
    }
    dp[i] = from_f<T>(r);
  }
}

}  // namespace b200ir

static inline unsigned moce_grid(int64_t n, int threads, int cap) {
  int64_t g = (n + threads - 1) / threads;
  if (g > cap) g = cap;
  return (unsigned)(g < 1 ? 1 : g);
}

extern "C" int b200ir_rows_copy(const void* x, int64_t x_rowstride, void* y, int64_t y_rowstride, const int* src_idx,
                                const int* dst_idx, const float* scale, int n, int64_t per, int accumulate, int dtype,
                                b200ir_stream_t stream) {
  B200IR_CHECK_ARG(x && y && n > 0 && n <= 65535 && per > 0, "rows_copy: bad args");
  B200IR_CHECK_ARG(dtype_ok(dtype), "rows_copy: bad dtype");
  const dim3 grid(moce_grid(per, 256, num_sms() * 4), (unsigned)n);
  if (dtype == B200IR_BF16)
    rows_copy_kernel<bf16><<<grid, 256, 0, (cudaStream_t)stream>>>((const bf16*)x, x_rowstride, (bf16*)y, y_rowstride, src_idx,
                                                                    dst_idx, scale, per, accumulate);
  else
    rows_copy_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>((const float*)x, x_rowstride, (float*)y, y_rowstride,
                                                                     src_idx, dst_idx, scale, per, accumulate);
  return check_launch("rows_copy");
}

extern "C" int b200ir_rowdot(const void* a, int64_t a_rowstride, const int* a_idx, const void* b, int64_t b_rowstride,
                             const int* b_idx, float* out, int n, int64_t per, int dtype, b200ir_stream_t stream) {
  B200IR_CHECK_ARG(a && b && out && n > 0 && per > 0, "rowdot: bad args");
  B200IR_CHECK_ARG(dtype_ok(dtype), "rowdot: bad dtype");
  B200IR_CHECK_ARG(n <= 65535, "rowdot: more than 65535 rows");
  const int esz = dtype == B200IR_BF16 ? 2 : 4, V = 16 / esz;
  const bool vec = per % V == 0 && (a_rowstride * esz) % 16 == 0 && (b_rowstride * esz) % 16 == 0 &&
                   ((uintptr_t)a & 15) == 0 && ((uintptr_t)b & 15) == 0;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(ROWDOT_CL, (unsigned)n);
  cfg.blockDim = dim3(ROWDOT_T);
  cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = ROWDOT_CL;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  cudaError_t e;
  if (dtype == B200IR_BF16) {
    const bf16 *pa = (const bf16*)a, *pb = (const bf16*)b;
    e = vec ? cudaLaunchKernelEx(&cfg, rowdot_kernel<bf16, true>, pa, a_rowstride, a_idx, pb, b_rowstride, b_idx, out, per)
            : cudaLaunchKernelEx(&cfg, rowdot_kernel<bf16, false>, pa, a_rowstride, a_idx, pb, b_rowstride, b_idx, out, per);
  } else {
    const float *pa = (const float*)a, *pb = (const float*)b;
    e = vec ? cudaLaunchKernelEx(&cfg, rowdot_kernel<float, true>, pa, a_rowstride, a_idx, pb, b_rowstride, b_idx, out, per)
            : cudaLaunchKernelEx(&cfg, rowdot_kernel<float, false>, pa, a_rowstride, a_idx, pb, b_rowstride, b_idx, out, per);
  }
  if (e != cudaSuccess) {
    set_error("rowdot: %s", cudaGetErrorString(e));
    return (int)e;
  }
  return check_launch("rowdot");
}

extern "C" int b200ir_spatial_mean_bwd(const float* g, const void* x, int64_t x_bstride, void* dx, int64_t dx_bstride,
                                       int dtype, int B, int C, int64_t HW, int gelu, b200ir_stream_t stream) {
  B200IR_CHECK_ARG(g && dx && (x || !gelu) && B > 0 && B <= 65535 && C > 0 && C <= 65535 && HW > 0, "spatial_mean_bwd: bad args");
  B200IR_CHECK_ARG(dtype_ok(dtype), "spatial_mean_bwd: bad dtype");
  const dim3 grid(moce_grid(HW, 256, 64), (unsigned)C, (unsigned)B);
  if (dtype == B200IR_BF16)
    spatial_mean_bwd_kernel<bf16><<<grid, 256, 0, (cudaStream_t)stream>>>(g, (const bf16*)x, x_bstride, (bf16*)dx, dx_bstride,
                                                                           C, HW, gelu);
  else
    spatial_mean_bwd_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>(g, (const float*)x, x_bstride, (float*)dx,
                                                                            dx_bstride, C, HW, gelu);
  return check_launch("spatial_mean_bwd");
}
