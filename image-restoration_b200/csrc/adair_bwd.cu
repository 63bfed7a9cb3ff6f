// adair_bwd.cu — backward of AdaIR's frequency mining / modulation pieces, the adjoints of adair_ops.cu.
// Reference: /root/reference/AdaIR-main/net/model.py:234-372 under torch autograd.
//
// FreModule.fft:  low = |P x|, high = |x - P x| with P = ifft2 . mask . fft2, an orthogonal projection onto the kept bins
// (Hermitian: P^H = P).  With z = P x, h = x - z, u_l = g_low z/|z|, u_h = g_high h/|h| (0 where the modulus is 0, as
// torch's sgn):   dx = Re(u_h) + Re(P (u_l - u_h)).
// The thresholds reach the mask through .int() in the reference, so nothing flows into rate_conv.
// Every reduction is per CTA in a fixed order: bitwise deterministic.
#include "common.cuh"

namespace b200ir {

constexpr int kMaxModesB = 64;   // = kMaxModes of adair_ops.cu

__device__ __forceinline__ void mode_uv(int m, int h_, int w_, int& u, int& v) {
  u = m / (2 * w_) - h_;
  v = m % (2 * w_) - w_;
}
__device__ __forceinline__ void twiddle(int u, int v, int yy, int xx, int H, int W, float& s, float& co) {
  const float ph = 2.0f * ((float)((u * yy) % H) / (float)H + (float)((v * xx) % W) / (float)W);
  sincospif(ph, &s, &co);
}

// coefficients of w = u_l - u_h:  coefw[b][c][m] = (1/HW) sum_p w_p e^{-i phi_m(p)};  grid (C, B), 256 threads
template <typename T>
__global__ void __launch_bounds__(256)
lowfreq_bwd_coef_kernel(const T* __restrict__ x, int64_t x_bs, const int* __restrict__ hw_, const float2* __restrict__ coef,
                        const T* __restrict__ gh, int64_t gh_bs, const T* __restrict__ gl, int64_t gl_bs,
                        float2* __restrict__ coefw, int C, int H, int W) {
  const int c = blockIdx.x, b = blockIdx.y;
  const int h_ = hw_[2 * b], w_ = hw_[2 * b + 1];
  const int nm = 4 * h_ * w_;
  if (nm == 0) return;
  __shared__ float2 sc[kMaxModesB];
  __shared__ float2 acc[8][kMaxModesB];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int m = threadIdx.x; m < nm; m += 256) sc[m] = coef[((int64_t)b * C + c) * kMaxModesB + m];
  for (int m = threadIdx.x; m < 8 * kMaxModesB; m += 256) acc[m / kMaxModesB][m % kMaxModesB] = make_float2(0.f, 0.f);
  __syncthreads();
  const int64_t plane = (int64_t)c * H * W;
  const T* xp = x + (int64_t)b * x_bs + plane;
  const T* ghp = gh + (int64_t)b * gh_bs + plane;
  const T* glp = gl + (int64_t)b * gl_bs + plane;
  const int HW = H * W;
  constexpr int K = 4;
  for (int base = 0; base < HW; base += 256 * K) {
    float2 wv[K];
    int yy[K], xx[K];
#pragma unroll
    for (int k = 0; k < K; ++k) {
      const int i = base + k * 256 + threadIdx.x;
      wv[k] = make_float2(0.f, 0.f);
      yy[k] = xx[k] = 0;
      if (i < HW) {
        yy[k] = i / W;
        xx[k] = i - yy[k] * W;
        float re = 0.f, im = 0.f;
        for (int m = 0; m < nm; ++m) {
          int u, v;
          mode_uv(m, h_, w_, u, v);
          float s, co;
          twiddle(u, v, yy[k], xx[k], H, W, s, co);
          re += sc[m].x * co - sc[m].y * s;
          im += sc[m].x * s + sc[m].y * co;
        }
        const float val = to_f<T>(xp[i]), g1 = to_f<T>(ghp[i]), g0 = to_f<T>(glp[i]);
        const float la = sqrtf(re * re + im * im);
        const float hr = val - re, hi = -im, ha = sqrtf(hr * hr + hi * hi);
        const float fl = la > 0.f ? g0 / la : 0.f, fh = ha > 0.f ? g1 / ha : 0.f;
        wv[k] = make_float2(fl * re - fh * hr, fl * im - fh * hi);
      }
    }
    for (int m = 0; m < nm; ++m) {
      int u, v;
      mode_uv(m, h_, w_, u, v);
      float re = 0.f, im = 0.f;
#pragma unroll
      for (int k = 0; k < K; ++k) {
        float s, co;
        twiddle(u, v, yy[k], xx[k], H, W, s, co);
        re += wv[k].x * co + wv[k].y * s;      // w * (co - i s)
        im += wv[k].y * co - wv[k].x * s;
      }
      re = warp_sum(re);
      im = warp_sum(im);
      if (lane == 0) {
        acc[warp][m].x += re;
        acc[warp][m].y += im;
      }
    }
  }
  __syncthreads();
  const float inv = 1.0f / ((float)H * (float)W);
  for (int m = threadIdx.x; m < nm; m += 256) {
    float re = 0.f, im = 0.f;
    for (int q = 0; q < 8; ++q) re += acc[q][m].x, im += acc[q][m].y;
    coefw[((int64_t)b * C + c) * kMaxModesB + m] = make_float2(re * inv, im * inv);
  }
}

// dx = Re(u_h) + Re(sum_m coefw[m] e^{+i phi_m});  grid (chunks, C, B)
template <typename T>
__global__ void __launch_bounds__(256)
lowfreq_bwd_apply_kernel(const T* __restrict__ x, int64_t x_bs, const int* __restrict__ hw_, const float2* __restrict__ coef,
                         const float2* __restrict__ coefw, const T* __restrict__ gh, int64_t gh_bs, T* __restrict__ dx,
                         int64_t dx_bs, int C, int H, int W) {
  const int c = blockIdx.y, b = blockIdx.z;
  const int h_ = hw_[2 * b], w_ = hw_[2 * b + 1];
  const int nm = 4 * h_ * w_;
  __shared__ float2 sc[kMaxModesB], sw[kMaxModesB];
  for (int m = threadIdx.x; m < nm; m += 256) {
    sc[m] = coef[((int64_t)b * C + c) * kMaxModesB + m];
    sw[m] = coefw[((int64_t)b * C + c) * kMaxModesB + m];
  }
  __syncthreads();
  const int64_t plane = (int64_t)c * H * W;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < H * W; i += gridDim.x * 256) {
    const int yy = i / W, xx = i - yy * W;
    float re = 0.f, im = 0.f, pr = 0.f;
    for (int m = 0; m < nm; ++m) {
      int u, v;
      mode_uv(m, h_, w_, u, v);
      float s, co;
      twiddle(u, v, yy, xx, H, W, s, co);
      re += sc[m].x * co - sc[m].y * s;
      im += sc[m].x * s + sc[m].y * co;
      pr += sw[m].x * co - sw[m].y * s;
    }
    const float val = to_f<T>(x[(int64_t)b * x_bs + plane + i]), g1 = to_f<T>(gh[(int64_t)b * gh_bs + plane + i]);
    const float hr = val - re, ha = sqrtf(hr * hr + im * im);
    dx[(int64_t)b * dx_bs + plane + i] = from_f<T>((ha > 0.f ? g1 * hr / ha : 0.f) + pr);
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// SpatialGate backward.  gate = sigmoid(a), a = conv7x7(mm), mm = [max_c x, mean_c x]
// ---------------------------------------------------------------------------------------------------------------------
__global__ void sg_ds_kernel(const float* __restrict__ dgate, const float* __restrict__ gate, float* __restrict__ ds, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float g = gate[i];
    ds[i] = dgate[i] * g * (1.f - g);
  }
}

// dw7[k][r][s] = sum_b sum_p ds[b, p] * mm[b, k, p + (r - 3, s - 3)];  grid 98
__global__ void __launch_bounds__(256)
sg_dw_kernel(const float* __restrict__ ds, const float* __restrict__ mm, float* __restrict__ dw, int B, int H, int W) {
  const int k = blockIdx.x / 49, r = (blockIdx.x % 49) / 7 - 3, s = blockIdx.x % 7 - 3;
  __shared__ float red[8];
  float a = 0.f;
  for (int b = 0; b < B; ++b) {
    const float* dp = ds + (int64_t)b * H * W;
    const float* mp = mm + ((int64_t)b * 2 + k) * H * W;
    for (int i = threadIdx.x; i < H * W; i += 256) {
      const int yy = i / W, xx = i - yy * W, yr = yy + r, xs = xx + s;
      if (yr >= 0 && yr < H && xs >= 0 && xs < W) a = fmaf(dp[i], mp[yr * W + xs], a);
    }
  }
  a = warp_sum(a);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = a;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int q = 0; q < 8; ++q) t += red[q];
    dw[blockIdx.x] = t;
  }
}

// dx[b, c, q] (+)= dmean(q) / C + [c == first argmax_c x[b, :, q]] * dmax(q),  dmm[k](q) = sum_{r,s} w[k][r][s] ds(q - (r,s))
template <typename T>
__global__ void __launch_bounds__(256)
sg_dx_kernel(const T* __restrict__ x, int64_t x_bs, const float* __restrict__ ds, const float* __restrict__ w,
             T* __restrict__ dx, int64_t dx_bs, int C, int H, int W, int accumulate) {
  const int b = blockIdx.y;
  __shared__ float sw[98];
  if (threadIdx.x < 98) sw[threadIdx.x] = w[threadIdx.x];
  __syncthreads();
  const int64_t HW = (int64_t)H * W;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < H * W; i += gridDim.x * 256) {
    const int yy = i / W, xx = i - yy * W;
    float dmax = 0.f, dmean = 0.f;
    for (int r = -3; r <= 3; ++r)
      for (int s = -3; s <= 3; ++s) {
        const int yr = yy - r, xs = xx - s;
        if (yr >= 0 && yr < H && xs >= 0 && xs < W) {
          const float d = ds[(int64_t)b * HW + (int64_t)yr * W + xs];
          dmax = fmaf(sw[(r + 3) * 7 + (s + 3)], d, dmax);
          dmean = fmaf(sw[49 + (r + 3) * 7 + (s + 3)], d, dmean);
        }
      }
    dmean /= (float)C;
    const T* xp = x + (int64_t)b * x_bs + i;
    float mx = -INFINITY;
    int arg = 0;
    for (int c = 0; c < C; ++c) {
      const float v = to_f<T>(xp[(int64_t)c * HW]);
      if (v > mx) mx = v, arg = c;
    }
    T* dp = dx + (int64_t)b * dx_bs + i;
    for (int c = 0; c < C; ++c) {
      float v = dmean + (c == arg ? dmax : 0.f);
      if (accumulate) v += to_f<T>(dp[(int64_t)c * HW]);
      dp[(int64_t)c * HW] = from_f<T>(v);
    }
  }
}

// AdaptiveMaxPool2d(1) backward: dx[b, c, :] = 0 except g[b, c] at the first maximum;  grid (C, B)
template <typename T>
__global__ void __launch_bounds__(256)
spatial_max_bwd_kernel(const T* __restrict__ x, int64_t x_bs, const float* __restrict__ g, T* __restrict__ dx, int64_t dx_bs,
                       int C, int64_t HW) {
  const int c = blockIdx.x, b = blockIdx.y;
  const T* xp = x + (int64_t)b * x_bs + (int64_t)c * HW;
  T* dp = dx + (int64_t)b * dx_bs + (int64_t)c * HW;
  float mx = -INFINITY;
  long long arg = HW;
  for (int64_t i = threadIdx.x; i < HW; i += 256) {
    const float v = to_f<T>(xp[i]);
    if (v > mx) mx = v, arg = i;
  }
  __shared__ float smx[8];
  __shared__ long long sarg[8];
  __shared__ long long s_arg;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, mx, o);
    const long long oa = __shfl_xor_sync(0xffffffffu, arg, o);
    if (ov > mx || (ov == mx && oa < arg)) mx = ov, arg = oa;
  }
  if ((threadIdx.x & 31) == 0) smx[threadIdx.x >> 5] = mx, sarg[threadIdx.x >> 5] = arg;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int q = 1; q < 8; ++q)
      if (smx[q] > mx || (smx[q] == mx && sarg[q] < arg)) mx = smx[q], arg = sarg[q];
    s_arg = arg;
  }
  __syncthreads();
  const long long a = s_arg;
  const float gv = g[(int64_t)b * C + c];
  for (int64_t i = threadIdx.x; i < HW; i += 256) dp[i] = from_f<T>(i == a ? gv : 0.f);
}

// ---------------------------------------------------------------------------------------------------------------------
// FreRefine combine backward: y = low * sw[b, p] + high * cw[b, c]
// ---------------------------------------------------------------------------------------------------------------------
// per plane: dhigh = g * cw[b, c], dcw[b, c] = sum_p g * high;  grid (C, B)
template <typename T>
__global__ void __launch_bounds__(256)
fre_combine_bwd_plane_kernel(const T* __restrict__ g, int64_t g_bs, const T* __restrict__ high, int64_t h_bs,
                             const float* __restrict__ cw, T* __restrict__ dhigh, int64_t dh_bs, float* __restrict__ dcw, int C,
                             int64_t HW) {
  const int c = blockIdx.x, b = blockIdx.y;
  const float wv = cw[(int64_t)b * C + c];
  const T* gp = g + (int64_t)b * g_bs + (int64_t)c * HW;
  const T* hp = high + (int64_t)b * h_bs + (int64_t)c * HW;
  T* dp = dhigh + (int64_t)b * dh_bs + (int64_t)c * HW;
  __shared__ float red[8];
  float a = 0.f;
  for (int64_t i = threadIdx.x; i < HW; i += 256) {
    const float gv = to_f<T>(gp[i]);
    a = fmaf(gv, to_f<T>(hp[i]), a);
    dp[i] = from_f<T>(gv * wv);
  }
  a = warp_sum(a);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = a;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int q = 0; q < 8; ++q) t += red[q];
    dcw[(int64_t)b * C + c] = t;
  }
}
// per pixel: dlow = g * sw[b, p], dsw[b, p] = sum_c g * low;  grid (chunks, B)
template <typename T>
__global__ void __launch_bounds__(256)
fre_combine_bwd_pixel_kernel(const T* __restrict__ g, int64_t g_bs, const T* __restrict__ low, int64_t l_bs,
                             const float* __restrict__ sw, T* __restrict__ dlow, int64_t dl_bs, float* __restrict__ dsw, int C,
                             int64_t HW) {
  const int b = blockIdx.y;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < HW; i += (int64_t)gridDim.x * 256) {
    const float sv = sw[(int64_t)b * HW + i];
    float a = 0.f;
    for (int c = 0; c < C; ++c) {
      const float gv = to_f<T>(g[(int64_t)b * g_bs + (int64_t)c * HW + i]);
      a = fmaf(gv, to_f<T>(low[(int64_t)b * l_bs + (int64_t)c * HW + i]), a);
      dlow[(int64_t)b * dl_bs + (int64_t)c * HW + i] = from_f<T>(gv * sv);
    }
    dsw[(int64_t)b * HW + i] = a;
  }
}

// y = a * p1[c] + b2 * p2[c]:  da = g p1, db = g p2, parts[b][c] = (sum_p g a, sum_p g b2);  grid (C, B)
template <typename T>
__global__ void __launch_bounds__(256)
affine2_bwd_kernel(const T* __restrict__ g, int64_t g_bs, const T* __restrict__ a, int64_t a_bs, const T* __restrict__ b2,
                   int64_t b_bs, const float* __restrict__ p1, const float* __restrict__ p2, T* __restrict__ da, int64_t da_bs,
                   T* __restrict__ db, int64_t db_bs, float2* __restrict__ parts, int C, int64_t HW) {
  const int c = blockIdx.x, b = blockIdx.y;
  const int64_t off = (int64_t)c * HW;
  const float q1 = p1[c], q2 = p2[c];
  __shared__ float2 red[8];
  float s1 = 0.f, s2 = 0.f;
  for (int64_t i = threadIdx.x; i < HW; i += 256) {
    const float gv = to_f<T>(g[(int64_t)b * g_bs + off + i]);
    s1 = fmaf(gv, to_f<T>(a[(int64_t)b * a_bs + off + i]), s1);
    s2 = fmaf(gv, to_f<T>(b2[(int64_t)b * b_bs + off + i]), s2);
    da[(int64_t)b * da_bs + off + i] = from_f<T>(gv * q1);
    db[(int64_t)b * db_bs + off + i] = from_f<T>(gv * q2);
  }
  s1 = warp_sum(s1);
  s2 = warp_sum(s2);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = make_float2(s1, s2);
  __syncthreads();
  if (threadIdx.x == 0) {
    float t1 = 0.f, t2 = 0.f;
    for (int q = 0; q < 8; ++q) t1 += red[q].x, t2 += red[q].y;
    parts[(int64_t)b * C + c] = make_float2(t1, t2);
  }
}

static inline unsigned bw_grid(int64_t n) {
  int64_t g = (n + 255) / 256;
  const int64_t cap = (int64_t)num_sms() * 8;
  return (unsigned)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace b200ir

using namespace b200ir;

#define B200IR_PLANE_ARGS(B, C) ((B) > 0 && (B) <= 65535 && (C) > 0 && (C) <= 65535)

extern "C" int b200ir_lowfreq_split_bwd(const void* x, int64_t x_bstride, const int* hw_counts, const float* coef,
                                        const void* g_high, int64_t gh_bstride, const void* g_low, int64_t gl_bstride,
                                        float* coefw_ws, void* dx, int64_t dx_bstride, int dtype, int B, int C, int H, int W,
                                        int max_modes, b200ir_stream_t stream) {
  B200IR_CHECK_ARG(x && hw_counts && coef && g_high && g_low && coefw_ws && dx && B200IR_PLANE_ARGS(B, C) && H > 0 && W > 0,
                   "lowfreq_split_bwd: bad args");
  B200IR_CHECK_ARG(dtype_ok(dtype), "lowfreq_split_bwd: bad dtype");
  B200IR_UNSUPPORTED(max_modes <= kMaxModesB, "lowfreq_split_bwd: more than %d spectrum bins kept", kMaxModesB);
  const cudaStream_t st = (cudaStream_t)stream;
  const dim3 gc((unsigned)C, (unsigned)B), ga(bw_grid((int64_t)H * W), (unsigned)C, (unsigned)B);
  if (dtype == B200IR_BF16) {
    lowfreq_bwd_coef_kernel<bf16><<<gc, 256, 0, st>>>((const bf16*)x, x_bstride, hw_counts, (const float2*)coef,
                                                      (const bf16*)g_high, gh_bstride, (const bf16*)g_low, gl_bstride,
                                                      (float2*)coefw_ws, C, H, W);
    lowfreq_bwd_apply_kernel<bf16><<<ga, 256, 0, st>>>((const bf16*)x, x_bstride, hw_counts, (const float2*)coef,
                                                       (const float2*)coefw_ws, (const bf16*)g_high, gh_bstride, (bf16*)dx,
                                                       dx_bstride, C, H, W);
  } else {
    lowfreq_bwd_coef_kernel<float><<<gc, 256, 0, st>>>((const float*)x, x_bstride, hw_counts, (const float2*)coef,
                                                       (const float*)g_high, gh_bstride, (const float*)g_low, gl_bstride,
                                                       (float2*)coefw_ws, C, H, W);
    lowfreq_bwd_apply_kernel<float><<<ga, 256, 0, st>>>((const float*)x, x_bstride, hw_counts, (const float2*)coef,
                                                        (const float2*)coefw_ws, (const float*)g_high, gh_bstride, (float*)dx,
                                                        dx_bstride, C, H, W);
  }
  return check_launch("lowfreq_split_bwd");
}

extern "C" int b200ir_spatial_gate_bwd(const void* x, int64_t x_bstride, int dtype, const float* w7, const float* mm,
                                       const float* gate, const float* dgate, float* ds_ws, void* dx, int64_t dx_bstride,
                                       float* dw7, int B, int C, int H, int W, int accumulate, b200ir_stream_t stream) {
  B200IR_CHECK_ARG(x && w7 && mm && gate && dgate && ds_ws && dx && dw7 && B > 0 && B <= 65535 && C > 0 && H > 0 && W > 0,
                   "spatial_gate_bwd: bad args");
  B200IR_CHECK_ARG(dtype_ok(dtype), "spatial_gate_bwd: bad dtype");
  const cudaStream_t st = (cudaStream_t)stream;
  const int64_t n = (int64_t)B * H * W;
  sg_ds_kernel<<<bw_grid(n), 256, 0, st>>>(dgate, gate, ds_ws, n);
  sg_dw_kernel<<<98, 256, 0, st>>>(ds_ws, mm, dw7, B, H, W);
  const dim3 grid(bw_grid((int64_t)H * W), (unsigned)B);
  if (dtype == B200IR_BF16)
    sg_dx_kernel<bf16><<<grid, 256, 0, st>>>((const bf16*)x, x_bstride, ds_ws, w7, (bf16*)dx, dx_bstride, C, H, W, accumulate);
  else
    sg_dx_kernel<float><<<grid, 256, 0, st>>>((const float*)x, x_bstride, ds_ws, w7, (float*)dx, dx_bstride, C, H, W, accumulate);
  return check_launch("spatial_gate_bwd");
}

extern "C" int b200ir_spatial_max_bwd(const void* x, int64_t x_bstride, const float* g, void* dx, int64_t dx_bstride, int dtype,
                                      int B, int C, int64_t HW, b200ir_stream_t stream) {
  B200IR_CHECK_ARG(x && g && dx && B200IR_PLANE_ARGS(B, C) && HW > 0, "spatial_max_bwd: bad args");
  B200IR_CHECK_ARG(dtype_ok(dtype), "spatial_max_bwd: bad dtype");
  const dim3 grid((unsigned)C, (unsigned)B);
  if (dtype == B200IR_BF16)
    spatial_max_bwd_kernel<bf16><<<grid, 256, 0, (cudaStream_t)stream>>>((const bf16*)x, x_bstride, g, (bf16*)dx, dx_bstride, C, HW);
  else
    spatial_max_bwd_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>((const float*)x, x_bstride, g, (float*)dx, dx_bstride, C, HW);
  return check_launch("spatial_max_bwd");
}

extern "C" int b200ir_fre_combine_bwd(const void* g, int64_t g_bstride, const void* low, int64_t l_bstride, const void* high,
                                      int64_t h_bstride, const float* sw, const float* cw, void* dlow, int64_t dl_bstride,
                                      void* dhigh, int64_t dh_bstride, float* dsw, float* dcw, int dtype, int B, int C,
                                      int64_t HW, b200ir_stream_t stream) {
  B200IR_CHECK_ARG(g && low && high && sw && cw && dlow && dhigh && dsw && dcw && B200IR_PLANE_ARGS(B, C) && HW > 0,
                   "fre_combine_bwd: bad args");
  B200IR_CHECK_ARG(dtype_ok(dtype), "fre_combine_bwd: bad dtype");
  const cudaStream_t st = (cudaStream_t)stream;
  const dim3 gp((unsigned)C, (unsigned)B), gx(bw_grid(HW), (unsigned)B);
  if (dtype == B200IR_BF16) {
    fre_combine_bwd_plane_kernel<bf16><<<gp, 256, 0, st>>>((const bf16*)g, g_bstride, (const bf16*)high, h_bstride, cw,
                                                           (bf16*)dhigh, dh_bstride, dcw, C, HW);
    fre_combine_bwd_pixel_kernel<bf16><<<gx, 256, 0, st>>>((const bf16*)g, g_bstride, (const bf16*)low, l_bstride, sw,
                                                           (bf16*)dlow, dl_bstride, dsw, C, HW);
  } else {
    fre_combine_bwd_plane_kernel<float><<<gp, 256, 0, st>>>((const float*)g, g_bstride, (const float*)high, h_bstride, cw,
                                                            (float*)dhigh, dh_bstride, dcw, C, HW);
    fre_combine_bwd_pixel_kernel<float><<<gx, 256, 0, st>>>((const float*)g, g_bstride, (const float*)low, l_bstride, sw,
                                                            (float*)dlow, dl_bstride, dsw, C, HW);
  }
  return check_launch("fre_combine_bwd");
}

extern "C" int b200ir_affine2_bwd(const void* g, int64_t g_bstride, const void* a, int64_t a_bstride, const void* b,
                                  int64_t b_bstride, const float* p1, const float* p2, void* da, int64_t da_bstride, void* db,
                                  int64_t db_bstride, float* parts, int dtype, int B, int C, int64_t HW, b200ir_stream_t stream) {
  B200IR_CHECK_ARG(g && a && b && p1 && p2 && da && db && parts && B200IR_PLANE_ARGS(B, C) && HW > 0, "affine2_bwd: bad args");
  B200IR_CHECK_ARG(dtype_ok(dtype), "affine2_bwd: bad dtype");
  const dim3 grid((unsigned)C, (unsigned)B);
  if (dtype == B200IR_BF16)
    affine2_bwd_kernel<bf16><<<grid, 256, 0, (cudaStream_t)stream>>>((const bf16*)g, g_bstride, (const bf16*)a, a_bstride,
                                                                     (const bf16*)b, b_bstride, p1, p2, (bf16*)da, da_bstride,
                                                                     (bf16*)db, db_bstride, (float2*)parts, C, HW);
  else
    affine2_bwd_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>((const float*)g, g_bstride, (const float*)a, a_bstride,
                                                                      (const float*)b, b_bstride, p1, p2, (float*)da,
                                                                      da_bstride, (float*)db, db_bstride, (float2*)parts, C, HW);
  return check_launch("affine2_bwd");
}
