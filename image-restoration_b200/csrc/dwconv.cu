// dwconv.cu — depthwise 3x3 (stride 1, zero pad 1) bandwidth kernels, optional fused GELU gate.
// Reference: Attention.qkv_dwconv Restormer.py:106,114 and FeedForward.dwconv + chunk(2) +
// F.gelu(x1)*x2 Restormer.py:84,90-91.
//
// v1 ("register window"): every thread owns a 16-byte-wide column strip (8 bf16 / 4 fp32 pixels)
// and slides a 3-row window down RS output rows, so each input row is loaded once per strip
// with one 128-bit load + two neighbour scalars (L1 hits) and each output row leaves as one
// 128-bit store.  Algorithmic bytes: read C*H*W (2*Hd*H*W gated) + write C*H*W (Hd*H*W).
#include "common.cuh"

namespace b200ir {

constexpr int RS = 8;  // output rows per thread strip

template <typename T> struct Vec;
template <> struct Vec<float> {
  static constexpr int N = 4;
  typedef float4 raw;
  static __device__ __forceinline__ void unpack(const raw& r, float* f) { f[0] = r.x; f[1] = r.y; f[2] = r.z; f[3] = r.w; }
  static __device__ __forceinline__ raw pack(const float* f) { return make_float4(f[0], f[1], f[2], f[3]); }
};
template <> struct Vec<bf16> {
  static constexpr int N = 8;
  typedef uint4 raw;
  static __device__ __forceinline__ void unpack(const raw& r, float* f) {
    const uint32_t u[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      f[2 * i] = __uint_as_float(u[i] << 16);
      f[2 * i + 1] = __uint_as_float(u[i] & 0xffff0000u);
    }
  }
  static __device__ __forceinline__ raw pack(const float* f) {
    uint32_t u[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      __nv_bfloat162 h = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
      u[i] = *reinterpret_cast<uint32_t*>(&h);
    }
    return make_uint4(u[0], u[1], u[2], u[3]);
  }
};

template <typename T>
__device__ __forceinline__ void load_row(const T* __restrict__ plane, int r, int H, int W, int x0, float* out /*N+2*/) {
  constexpr int N = Vec<T>::N;
  if (r < 0 || r >= H) {
#pragma unroll
    for (int i = 0; i < N + 2; ++i) out[i] = 0.f;
    return;
  }
  const T* row = plane + (int64_t)r * W + x0;
  typename Vec<T>::raw v = __ldg(reinterpret_cast<const typename Vec<T>::raw*>(row));
  Vec<T>::unpack(v, out + 1);
  out[0] = (x0 > 0) ? to_f<T>(__ldg(row - 1)) : 0.f;
  out[N + 1] = (x0 + N < W) ? to_f<T>(__ldg(row + N)) : 0.f;
}

template <typename T, int GATE>
__global__ void __launch_bounds__(256)
dwconv3x3_vec_kernel(const T* __restrict__ x, int64_t x_bs, const float* __restrict__ w,
                     const float* __restrict__ bias, T* __restrict__ y, int64_t y_bs, int Cout, int H, int W,
                     int nxg, int nstrip, int64_t total, int flip) {
  constexpr int N = Vec<T>::N;
  constexpr int NIN = GATE ? 2 : 1;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int xg = (int)(idx % nxg);
  int64_t t = idx / nxg;
  const int strip = (int)(t % nstrip);
  t /= nstrip;
  const int c = (int)(t % Cout);
  const int b = (int)(t / Cout);
  const int x0 = xg * N, y0 = strip * RS;
  const int y1 = min(H, y0 + RS);
  const int64_t HWp = (int64_t)H * W;

  float wk[NIN][9], bs[NIN];
  const T* plane[NIN];
#pragma unroll
  for (int k = 0; k < NIN; ++k) {
    const int cin = c + k * Cout;
#pragma unroll
    for (int i = 0; i < 9; ++i) wk[k][i] = __ldg(w + cin * 9 + (flip ? 8 - i : i));
    bs[k] = bias ? __ldg(bias + cin) : 0.f;
    plane[k] = x + (int64_t)b * x_bs + (int64_t)cin * HWp;
  }
  T* oplane = y + (int64_t)b * y_bs + (int64_t)c * HWp;

  float win[NIN][3][N + 2];
#pragma unroll
  for (int k = 0; k < NIN; ++k) {
    load_row<T>(plane[k], y0 - 1, H, W, x0, win[k][0]);
    load_row<T>(plane[k], y0, H, W, x0, win[k][1]);
  }
  for (int yy = y0; yy < y1; ++yy) {
    float o[NIN][N];
#pragma unroll
    for (int k = 0; k < NIN; ++k) {
      load_row<T>(plane[k], yy + 1, H, W, x0, win[k][2]);
#pragma unroll
      for (int i = 0; i < N; ++i) {
        float a = bs[k];
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
          for (int s = 0; s < 3; ++s) a = fmaf(wk[k][r * 3 + s], win[k][r][i + s], a);
        o[k][i] = a;
      }
#pragma unroll
      for (int i = 0; i < N + 2; ++i) {
        win[k][0][i] = win[k][1][i];
        win[k][1][i] = win[k][2][i];
      }
    }
    float outv[N];
#pragma unroll
    for (int i = 0; i < N; ++i)
      outv[i] = GATE == 2 ? gelu_erf(o[0][i]) / (1.f + __expf(-o[NIN - 1][i]))
                          : (GATE ? gelu_erf(o[0][i]) * o[NIN - 1][i] : o[0][i]);
    *reinterpret_cast<typename Vec<T>::raw*>(oplane + (int64_t)yy * W + x0) = Vec<T>::pack(outv);
  }
}

// generic fallback: any W / alignment, one output pixel per thread
template <typename T, int GATE>
__global__ void __launch_bounds__(256)
dwconv3x3_scalar_kernel(const T* __restrict__ x, int64_t x_bs, const float* __restrict__ w,
                        const float* __restrict__ bias, T* __restrict__ y, int64_t y_bs, int Cout, int H, int W,
                        int64_t total, int flip) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int xx = (int)(idx % W);
  int64_t t = idx / W;
  const int yy = (int)(t % H);
  t /= H;
  const int c = (int)(t % Cout);
  const int b = (int)(t / Cout);
  const int64_t HWp = (int64_t)H * W;
  float o[2] = {0.f, 0.f};
#pragma unroll
  for (int k = 0; k < (GATE ? 2 : 1); ++k) {
    const int cin = c + k * Cout;
    const T* plane = x + (int64_t)b * x_bs + (int64_t)cin * HWp;
    float a = bias ? bias[cin] : 0.f;
#pragma unroll
    for (int r = -1; r <= 1; ++r)
#pragma unroll
      for (int s = -1; s <= 1; ++s) {
        const int yr = yy + r, xs = xx + s;
        if (yr >= 0 && yr < H && xs >= 0 && xs < W)
          a = fmaf(w[cin * 9 + (flip ? 8 - ((r + 1) * 3 + (s + 1)) : (r + 1) * 3 + (s + 1))],
                   to_f<T>(plane[(int64_t)yr * W + xs]), a);
      }
    o[k] = a;
  }
  y[(int64_t)b * y_bs + (int64_t)c * HWp + (int64_t)yy * W + xx] =
      from_f<T>(GATE == 2 ? gelu_erf(o[0]) / (1.f + __expf(-o[1])) : (GATE ? gelu_erf(o[0]) * o[1] : o[0]));
}

template <typename T, int GATE>
static int launch_dwconv(const void* x, int64_t x_bs, const float* w, const float* bias, void* y, int64_t y_bs,
                         int B, int Cout, int H, int W, cudaStream_t st, int flip = 0) {
  constexpr int N = Vec<T>::N;
  const bool vec_ok = (W % N == 0) && (x_bs % N == 0) && (y_bs % N == 0) && ((uintptr_t)x % 16 == 0) &&
                      ((uintptr_t)y % 16 == 0);
  if (vec_ok) {
    const int nxg = W / N, nstrip = (H + RS - 1) / RS;
    const int64_t total = (int64_t)B * Cout * nstrip * nxg;
    const int64_t blocks = (total + 255) / 256;
    dwconv3x3_vec_kernel<T, GATE><<<(unsigned)blocks, 256, 0, st>>>((const T*)x, x_bs, w, bias, (T*)y, y_bs, Cout, H,
                                                                    W, nxg, nstrip, total, flip);
  } else {
    const int64_t total = (int64_t)B * Cout * H * W;
    const int64_t blocks = (total + 255) / 256;
    dwconv3x3_scalar_kernel<T, GATE><<<(unsigned)blocks, 256, 0, st>>>((const T*)x, x_bs, w, bias, (T*)y, y_bs, Cout,
                                                                       H, W, total, flip);
  }
  return check_launch(GATE ? "dwconv3x3_gelu_gate" : "dwconv3x3");
}

// dwconv_tma.cu: TMA-staged bf16 path
bool dwconv_tma_supported(int dtype, int H, int W, int64_t x_bs, int64_t y_bs, const void* x, const void* y);
template <bool GATE>
int dwconv_tma_launch(const void* x, int64_t x_bs, const float* w, const float* bias, void* y, int64_t y_bs, int B,
                      int Cout, int H, int W, cudaStream_t st, int flip = 0);

}  // namespace b200ir

using namespace b200ir;

extern "C" int b200ir_dwconv3x3_fwd(const void* x, int64_t x_bstride, const float* w, const float* bias, void* y,
                                    int64_t y_bstride, int dtype, int B, int C, int H, int W,
                                    b200ir_stream_t stream) {
  B200IR_CHECK_ARG(x && w && y && B > 0 && C > 0 && H > 0 && W > 0, "dwconv3x3: bad args");
  B200IR_CHECK_ARG(dtype_ok(dtype), "dwconv3x3: bad dtype %d", dtype);
  if (dwconv_tma_supported(dtype, H, W, x_bstride, y_bstride, x, y))
    return dwconv_tma_launch<false>(x, x_bstride, w, bias, y, y_bstride, B, C, H, W, (cudaStream_t)stream);
  if (dtype == B200IR_BF16)
    return launch_dwconv<bf16, 0>(x, x_bstride, w, bias, y, y_bstride, B, C, H, W, (cudaStream_t)stream);
  return launch_dwconv<float, 0>(x, x_bstride, w, bias, y, y_bstride, B, C, H, W, (cudaStream_t)stream);
}

extern "C" int b200ir_dwconv3x3_bwd_input(const void* dy, int64_t dy_bstride, const float* w, void* dx,
                                          int64_t dx_bstride, int dtype, int B, int C, int H, int W,
                                          b200ir_stream_t stream) {
  B200IR_CHECK_ARG(dy && w && dx && B > 0 && C > 0 && H > 0 && W > 0, "dwconv3x3_bwd_input: bad args");
  B200IR_CHECK_ARG(dtype_ok(dtype), "dwconv3x3_bwd_input: bad dtype %d", dtype);
  if (dwconv_tma_supported(dtype, H, W, dy_bstride, dx_bstride, dy, dx))
    return dwconv_tma_launch<false>(dy, dy_bstride, w, nullptr, dx, dx_bstride, B, C, H, W, (cudaStream_t)stream, 1);
  if (dtype == B200IR_BF16)
    return launch_dwconv<bf16, 0>(dy, dy_bstride, w, nullptr, dx, dx_bstride, B, C, H, W, (cudaStream_t)stream, 1);
  return launch_dwconv<float, 0>(dy, dy_bstride, w, nullptr, dx, dx_bstride, B, C, H, W, (cudaStream_t)stream, 1);
}

extern "C" int b200ir_dwconv3x3_gelu_gate_fwd(const void* x, int64_t x_bstride, const float* w, const float* bias,
                                              void* y, int64_t y_bstride, int dtype, int B, int Hd, int H, int W,
                                              b200ir_stream_t stream) {
  B200IR_CHECK_ARG(x && w && y && B > 0 && Hd > 0 && H > 0 && W > 0, "dwconv3x3_gelu_gate: bad args");
  B200IR_CHECK_ARG(dtype_ok(dtype), "dwconv3x3_gelu_gate: bad dtype %d", dtype);
  if (dwconv_tma_supported(dtype, H, W, x_bstride, y_bstride, x, y))
    return dwconv_tma_launch<true>(x, x_bstride, w, bias, y, y_bstride, B, Hd, H, W, (cudaStream_t)stream);
  if (dtype == B200IR_BF16)
    return launch_dwconv<bf16, 1>(x, x_bstride, w, bias, y, y_bstride, B, Hd, H, W, (cudaStream_t)stream);
  return launch_dwconv<float, 1>(x, x_bstride, w, bias, y, y_bstride, B, Hd, H, W, (cudaStream_t)stream);
}

// Split-gate FFN of the reference's RectiFormer experiment (experiment/RectiFormer.py:231-266):
// out[c] = gelu(dw_content(t)[c]) * sigmoid(dw_gate(t)[c + Hd]); same kernel, gate function 2 (CUDA-core path).
extern "C" int b200ir_dwconv3x3_gelu_sigmoid_gate_fwd(const void* x, int64_t x_bstride, const float* w, const float* bias,
                                                      void* y, int64_t y_bstride, int dtype, int B, int Hd, int H, int W,
                                                      b200ir_stream_t stream) {
  B200IR_CHECK_ARG(x && w && y && B > 0 && Hd > 0 && H > 0 && W > 0, "dwconv3x3_gelu_sigmoid_gate: bad args");
  B200IR_CHECK_ARG(dtype_ok(dtype), "dwconv3x3_gelu_sigmoid_gate: bad dtype %d", dtype);
  if (dtype == B200IR_BF16)
    return launch_dwconv<bf16, 2>(x, x_bstride, w, bias, y, y_bstride, B, Hd, H, W, (cudaStream_t)stream);
  return launch_dwconv<float, 2>(x, x_bstride, w, bias, y, y_bstride, B, Hd, H, W, (cudaStream_t)stream);
}
