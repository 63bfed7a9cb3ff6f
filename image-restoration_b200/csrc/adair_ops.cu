// adair_ops.cu — AdaIR's frequency mining / modulation pieces that are not a Restormer block:
// FreModule.fft (fft2 -> centred rectangular mask -> ifft2 -> abs), F.interpolate(bilinear), SpatialGate, ChannelGate,
// FreRefine's gating and the final per-channel affine.  Reference: /root/reference/AdaIR-main/net/model.py:234-372.
//
// FreModule.fft keeps, per sample, the spectrum bins u in [-h_, h_), v in [-w_, w_) (h_ = int(H/128 * t0) <= H/128:
// a handful of bins) as `low` and everything else as `high`, then takes |ifft2|.  With so few bins the low-pass part is
// synthesised directly:  lowc[y,x] = sum_{u,v} F[u,v] e^{+2 pi i (u y/H + v x/W)},  F[u,v] = (1/HW) sum x e^{-2 pi i (...)},
// low = |lowc|, high = |x - lowc|  (ifft2 is linear and fft_high = fft - fft_low).  No FFT library, exact for any H, W.
#include "common.cuh"

namespace b200ir {

constexpr int kMaxModes = 64;   // (2 h_) * (2 w_) <= 64: images up to 512 x 512 at n = 128 (model.py:341)

// coefficients: grid (C, B), 256 threads; coef[b][c][mode] = (re, im)
template <typename T>
__global__ void __launch_bounds__(256)
lowfreq_coef_kernel(const T* __restrict__ x, int64_t x_bs, const int* __restrict__ hw_, float2* __restrict__ coef, int C,
                    int H, int W) {
  const int c = blockIdx.x, b = blockIdx.y;
  const int h_ = hw_[2 * b], w_ = hw_[2 * b + 1];
  const int nm = 4 * h_ * w_;
  __shared__ float2 red[8];
  const T* p = x + (int64_t)b * x_bs + (int64_t)c * H * W;
  const float inv = 1.0f / ((float)H * (float)W);
  for (int m = 0; m < nm; ++m) {
    const int u = m / (2 * w_) - h_, v = m % (2 * w_) - w_;
    float re = 0.f, im = 0.f;
    for (int i = threadIdx.x; i < H * W; i += 256) {
      const int yy = i / W, xx = i - yy * W;
      // phase = -2 pi (u y / H + v x / W); reduce the arguments mod 1 in integers to keep sincospif exact
      const float ph = -2.0f * ((float)((u * yy) % H) / (float)H + (float)((v * xx) % W) / (float)W);
      float s, co;
      sincospif(ph, &s, &co);
      const float val = to_f<T>(p[i]);
      re = fmaf(val, co, re);
      im = fmaf(val, s, im);
    }
    re = warp_sum(re);
    im = warp_sum(im);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = make_float2(re, im);
    __syncthreads();
    if (threadIdx.x == 0) {
      float r2 = 0.f, i2 = 0.f;
      for (int k = 0; k < 8; ++k) r2 += red[k].x, i2 += red[k].y;
      coef[((int64_t)b * C + c) * kMaxModes + m] = make_float2(r2 * inv, i2 * inv);
    }
    __syncthreads();
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
lowfreq_apply_kernel(const T* __restrict__ x, int64_t x_bs, const int* __restrict__ hw_, const float2* __restrict__ coef,
                     T* __restrict__ high, int64_t h_bs, T* __restrict__ low, int64_t l_bs, int C, int H, int W) {
  const int c = blockIdx.y, b = blockIdx.z;
  const int h_ = hw_[2 * b], w_ = hw_[2 * b + 1];
  const int nm = 4 * h_ * w_;
  __shared__ float2 sc[kMaxModes];
  for (int m = threadIdx.x; m < nm; m += 256) sc[m] = coef[((int64_t)b * C + c) * kMaxModes + m];
  __syncthreads();
  const int64_t plane = (int64_t)c * H * W;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < H * W; i += gridDim.x * 256) {
    const int yy = i / W, xx = i - yy * W;
    float re = 0.f, im = 0.f;
    for (int m = 0; m < nm; ++m) {
      const int u = m / (2 * w_) - h_, v = m % (2 * w_) - w_;
      const float ph = 2.0f * ((float)((u * yy) % H) / (float)H + (float)((v * xx) % W) / (float)W);
      float s, co;
      sincospif(ph, &s, &co);
      re += sc[m].x * co - sc[m].y * s;
      im += sc[m].x * s + sc[m].y * co;
    }
    const float val = to_f<T>(x[(int64_t)b * x_bs + plane + i]);
    low[(int64_t)b * l_bs + plane + i] = from_f<T>(sqrtf(re * re + im * im));
    high[(int64_t)b * h_bs + plane + i] = from_f<T>(sqrtf((val - re) * (val - re) + im * im));
  }
}

// F.interpolate(mode='bilinear', align_corners=False)
template <typename T>
__global__ void bilinear_kernel(const T* __restrict__ x, int64_t x_bs, T* __restrict__ y, int64_t y_bs, int C, int Hi, int Wi,
                                int Ho, int Wo) {
  const int b = blockIdx.y;
  const float sh = (float)Hi / (float)Ho, sw = (float)Wi / (float)Wo;
  const int64_t per = (int64_t)C * Ho * Wo;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < per; i += (int64_t)gridDim.x * blockDim.x) {
    const int xo = (int)(i % Wo), yo = (int)((i / Wo) % Ho), c = (int)(i / ((int64_t)Wo * Ho));
    const float fy = fmaxf(((float)yo + 0.5f) * sh - 0.5f, 0.f), fx = fmaxf(((float)xo + 0.5f) * sw - 0.5f, 0.f);
    const int y0 = min((int)fy, Hi - 1), x0 = min((int)fx, Wi - 1);
    const int y1 = min(y0 + 1, Hi - 1), x1 = min(x0 + 1, Wi - 1);
    const float ly = fy - (float)y0, lx = fx - (float)x0;
    const T* p = x + (int64_t)b * x_bs + (int64_t)c * Hi * Wi;
    const float v = (1.f - ly) * ((1.f - lx) * to_f<T>(p[y0 * Wi + x0]) + lx * to_f<T>(p[y0 * Wi + x1])) +
                    ly * ((1.f - lx) * to_f<T>(p[y1 * Wi + x0]) + lx * to_f<T>(p[y1 * Wi + x1]));
    y[(int64_t)b * y_bs + i] = from_f<T>(v);
  }
}

// SpatialGate (model.py:234-246): sigmoid(conv7x7([max_c x, mean_c x])) -> [B,1,H,W] fp32
template <typename T>
__global__ void __launch_bounds__(256)
chan_maxmean_kernel(const T* __restrict__ x, int64_t x_bs, float* __restrict__ mm, int C, int64_t HW) {
  const int b = blockIdx.y;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < HW; i += (int64_t)gridDim.x * 256) {
    float mx = -INFINITY, s = 0.f;
    for (int c = 0; c < C; ++c) {
      const float v = to_f<T>(x[(int64_t)b * x_bs + (int64_t)c * HW + i]);
      mx = fmaxf(mx, v);
      s += v;
    }
    mm[((int64_t)b * 2) * HW + i] = mx;
    mm[((int64_t)b * 2 + 1) * HW + i] = s / (float)C;
  }
}
__global__ void __launch_bounds__(256)
conv7x7_2to1_sigmoid_kernel(const float* __restrict__ mm, const float* __restrict__ w, float* __restrict__ y, int H, int W) {
  const int b = blockIdx.y;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < H * W; i += gridDim.x * 256) {
    const int yy = i / W, xx = i - yy * W;
    float a = 0.f;
    for (int k = 0; k < 2; ++k)
      for (int r = -3; r <= 3; ++r)
        for (int s = -3; s <= 3; ++s) {
          const int yr = yy + r, xs = xx + s;
          if (yr >= 0 && yr < H && xs >= 0 && xs < W)
            a = fmaf(w[k * 49 + (r + 3) * 7 + (s + 3)], mm[((int64_t)b * 2 + k) * H * W + (int64_t)yr * W + xs], a);
        }
    y[(int64_t)b * H * W + i] = 1.f / (1.f + __expf(-a));
  }
}

// AdaptiveMaxPool2d(1): one warp per (b, c)
template <typename T>
__global__ void spatial_max_kernel(const T* __restrict__ x, int64_t x_bs, float* __restrict__ out, int B, int C, int64_t HW) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int c = warp % C, n = warp / C;
  if (n >= B) return;
  const T* p = x + (int64_t)n * x_bs + (int64_t)c * HW;
  float m = -INFINITY;
  for (int64_t i = lane; i < HW; i += 32) m = fmaxf(m, to_f<T>(p[i]));
  m = warp_max(m);
  if (lane == 0) out[(int64_t)n * C + c] = m;
}

// FreRefine (model.py:275-287): y = low * spatial[b,0,p] + high * channel[b,c]
template <typename T>
__global__ void fre_combine_kernel(const T* __restrict__ low, int64_t l_bs, const T* __restrict__ high, int64_t h_bs,
                                   const float* __restrict__ sw, const float* __restrict__ cw, T* __restrict__ y, int64_t y_bs,
                                   int C, int64_t HW) {
  const int b = blockIdx.y;
  const int64_t per = (int64_t)C * HW;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < per; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i / HW);
    const int64_t p = i - (int64_t)c * HW;
    const float v = to_f<T>(low[(int64_t)b * l_bs + i]) * sw[(int64_t)b * HW + p] +
                    to_f<T>(high[(int64_t)b * h_bs + i]) * cw[(int64_t)b * C + c];
    y[(int64_t)b * y_bs + i] = from_f<T>(v);
  }
}

// out * para1[c] + y * para2[c]   (model.py:324)
template <typename T>
__global__ void affine2_kernel(const T* __restrict__ a, int64_t a_bs, const T* __restrict__ b2, int64_t b_bs,
                               const float* __restrict__ p1, const float* __restrict__ p2, T* __restrict__ y, int64_t y_bs,
                               int C, int64_t HW) {
  const int b = blockIdx.y;
  const int64_t per = (int64_t)C * HW;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < per; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i / HW);
    y[(int64_t)b * y_bs + i] = from_f<T>(to_f<T>(a[(int64_t)b * a_bs + i]) * p1[c] + to_f<T>(b2[(int64_t)b * b_bs + i]) * p2[c]);
  }
}

static inline unsigned ew_grid(int64_t n) {
  int64_t g = (n + 255) / 256;
  const int64_t cap = (int64_t)num_sms() * 8;
  return (unsigned)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace b200ir

using namespace b200ir;

#define B200IR_BY_DTYPE(KERNEL, GRID, BLOCK, ...)                                             \
  do {                                                                                        \
    if (dtype == B200IR_BF16) KERNEL<bf16><<<GRID, BLOCK, 0, (cudaStream_t)stream>>>(__VA_ARGS__); \
    else KERNEL<float><<<GRID, BLOCK, 0, (cudaStream_t)stream>>>(__VA_ARGS__);                \
  } while (0)

extern "C" int b200ir_lowfreq_split(const void* x, int64_t x_bstride, const int* hw_counts, float* coef_ws, void* high,
                                    int64_t h_bstride, void* low, int64_t l_bstride, int dtype, int B, int C, int H, int W,
                                    int max_modes, b200ir_stream_t stream) {
  B200IR_CHECK_ARG(x && hw_counts && coef_ws && high && low && B > 0 && B <= 65535 && C > 0 && C <= 65535 && H > 0 && W > 0,
                   "lowfreq_split: bad args");
  B200IR_CHECK_ARG(dtype_ok(dtype), "lowfreq_split: bad dtype");
  B200IR_UNSUPPORTED(max_modes <= kMaxModes, "lowfreq_split: more than %d spectrum bins kept (image larger than 512x512?)", kMaxModes);
  if (dtype == B200IR_BF16) {
    lowfreq_coef_kernel<bf16><<<dim3((unsigned)C, (unsigned)B), 256, 0, (cudaStream_t)stream>>>(
        (const bf16*)x, x_bstride, hw_counts, (float2*)coef_ws, C, H, W);
    lowfreq_apply_kernel<bf16><<<dim3(ew_grid((int64_t)H * W), (unsigned)C, (unsigned)B), 256, 0, (cudaStream_t)stream>>>(
        (const bf16*)x, x_bstride, hw_counts, (const float2*)coef_ws, (bf16*)high, h_bstride, (bf16*)low, l_bstride, C, H, W);
  } else {
    lowfreq_coef_kernel<float><<<dim3((unsigned)C, (unsigned)B), 256, 0, (cudaStream_t)stream>>>(
        (const float*)x, x_bstride, hw_counts, (float2*)coef_ws, C, H, W);
    lowfreq_apply_kernel<float><<<dim3(ew_grid((int64_t)H * W), (unsigned)C, (unsigned)B), 256, 0, (cudaStream_t)stream>>>(
        (const float*)x, x_bstride, hw_counts, (const float2*)coef_ws, (float*)high, h_bstride, (float*)low, l_bstride, C, H, W);
  }
  return check_launch("lowfreq_split");
}

extern "C" int b200ir_bilinear_resize(const void* x, int64_t x_bstride, void* y, int64_t y_bstride, int dtype, int B, int C,
                                      int Hin, int Win, int Hout, int Wout, b200ir_stream_t stream) {
  B200IR_CHECK_ARG(x && y && B > 0 && B <= 65535 && C > 0 && Hin > 0 && Win > 0 && Hout > 0 && Wout > 0, "bilinear_resize: bad args");
  B200IR_CHECK_ARG(dtype_ok(dtype), "bilinear_resize: bad dtype");
  const dim3 grid(ew_grid((int64_t)C * Hout * Wout), (unsigned)B);
  if (dtype == B200IR_BF16)
    bilinear_kernel<bf16><<<grid, 256, 0, (cudaStream_t)stream>>>((const bf16*)x, x_bstride, (bf16*)y, y_bstride, C, Hin, Win, Hout, Wout);
  else
    bilinear_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>((const float*)x, x_bstride, (float*)y, y_bstride, C, Hin, Win, Hout, Wout);
  return check_launch("bilinear_resize");
}

extern "C" int b200ir_spatial_gate(const void* x, int64_t x_bstride, int dtype, const float* w7, float* mm_ws, float* gate,
                                   int B, int C, int H, int W, b200ir_stream_t stream) {
  B200IR_CHECK_ARG(x && w7 && mm_ws && gate && B > 0 && B <= 65535 && C > 0 && H > 0 && W > 0, "spatial_gate: bad args");
  B200IR_CHECK_ARG(dtype_ok(dtype), "spatial_gate: bad dtype");
  const dim3 grid(ew_grid((int64_t)H * W), (unsigned)B);
  if (dtype == B200IR_BF16)
    chan_maxmean_kernel<bf16><<<grid, 256, 0, (cudaStream_t)stream>>>((const bf16*)x, x_bstride, mm_ws, C, (int64_t)H * W);
  else
    chan_maxmean_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>((const float*)x, x_bstride, mm_ws, C, (int64_t)H * W);
  conv7x7_2to1_sigmoid_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(mm_ws, w7, gate, H, W);
  return check_launch("spatial_gate");
}

extern "C" int b200ir_spatial_max(const void* x, int64_t x_bstride, int dtype, float* out, int B, int C, int64_t HW,
                                  b200ir_stream_t stream) {
  B200IR_CHECK_ARG(x && out && B > 0 && C > 0 && HW > 0, "spatial_max: bad args");
  B200IR_CHECK_ARG(dtype_ok(dtype), "spatial_max: bad dtype");
  const unsigned grid = (unsigned)(((int64_t)B * C + 7) / 8);
  if (dtype == B200IR_BF16)
    spatial_max_kernel<bf16><<<grid, 256, 0, (cudaStream_t)stream>>>((const bf16*)x, x_bstride, out, B, C, HW);
  else
    spatial_max_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>((const float*)x, x_bstride, out, B, C, HW);
  return check_launch("spatial_max");
}

extern "C" int b200ir_fre_combine(const void* low, int64_t l_bstride, const void* high, int64_t h_bstride, const float* sw,
                                  const float* cw, void* y, int64_t y_bstride, int dtype, int B, int C, int64_t HW,
                                  b200ir_stream_t stream) {
  B200IR_CHECK_ARG(low && high && sw && cw && y && B > 0 && B <= 65535 && C > 0 && HW > 0, "fre_combine: bad args");
  B200IR_CHECK_ARG(dtype_ok(dtype), "fre_combine: bad dtype");
  const dim3 grid(ew_grid((int64_t)C * HW), (unsigned)B);
  if (dtype == B200IR_BF16)
    fre_combine_kernel<bf16><<<grid, 256, 0, (cudaStream_t)stream>>>((const bf16*)low, l_bstride, (const bf16*)high, h_bstride, sw, cw, (bf16*)y, y_bstride, C, HW);
  else
    fre_combine_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>((const float*)low, l_bstride, (const float*)high, h_bstride, sw, cw, (float*)y, y_bstride, C, HW);
  return check_launch("fre_combine");
}

extern "C" int b200ir_affine2(const void* a, int64_t a_bstride, const void* b, int64_t b_bstride, const float* p1,
                              const float* p2, void* y, int64_t y_bstride, int dtype, int B, int C, int64_t HW,
                              b200ir_stream_t stream) {
  B200IR_CHECK_ARG(a && b && p1 && p2 && y && B > 0 && B <= 65535 && C > 0 && HW > 0, "affine2: bad args");
  B200IR_CHECK_ARG(dtype_ok(dtype), "affine2: bad dtype");
  const dim3 grid(ew_grid((int64_t)C * HW), (unsigned)B);
  if (dtype == B200IR_BF16)
    affine2_kernel<bf16><<<grid, 256, 0, (cudaStream_t)stream>>>((const bf16*)a, a_bstride, (const bf16*)b, b_bstride, p1, p2, (bf16*)y, y_bstride, C, HW);
  else
    affine2_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>((const float*)a, a_bstride, (const float*)b, b_bstride, p1, p2, (float*)y, y_bstride, C, HW);
  return check_launch("affine2");
}
