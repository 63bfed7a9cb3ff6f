// dwconv_kxk.cu — depthwise K x K convolution (stride 1, zero padding K/2, odd K <= 7), the generalisation of the 3x3
// depthwise kernel that MoCE-IR's CrossAttention needs for its 7x7 kv_dwconv (moce_ir.py:341-342).
// One CTA = one 32 x 32 output tile of one (sample, channel) plane: the (32+K-1)^2 halo tile is staged in shared memory
// as fp32 (zero padding applied while staging), a thread owns one column and four consecutive rows, so that every
// staged value it loads is used for up to 4 x K products; the K*K taps live in registers.  fp32 accumulation.
#include "common.cuh"

namespace b200ir {

constexpr int KT = 32;   // tile edge

template <int K>
__global__ void __launch_bounds__(256)
dwconv_kxk_kernel(const void* __restrict__ x, int64_t x_bs, const float* __restrict__ w, const float* __restrict__ bias,
                  void* __restrict__ y, int64_t y_bs, int dt, int C, int H, int W, int tiles_x, int tiles, int flip) {
  constexpr int R = K / 2, SW = KT + K - 1;
  __shared__ float s[SW][SW + 1];
  const int tid = threadIdx.x, tx = tid & 31, ty = tid >> 5;
  const int tile = blockIdx.x % tiles, pc = blockIdx.x / tiles;
  const int c = pc % C, b = pc / C;
  const int x0 = (tile % tiles_x) * KT, y0 = (tile / tiles_x) * KT;
  const int64_t HWp = (int64_t)H * W;
  const int64_t xo = (int64_t)b * x_bs + (int64_t)c * HWp, yo = (int64_t)b * y_bs + (int64_t)c * HWp;
  for (int e = tid; e < SW * SW; e += 256) {
    const int r = e / SW, cc = e - r * SW;
    const int gy = y0 + r - R, gx = x0 + cc - R;
    s[r][cc] = (gy >= 0 && gy < H && gx >= 0 && gx < W) ? ld_any(x, dt, xo + (int64_t)gy * W + gx) : 0.f;
  }
  float wk[K * K];
#pragma unroll
  for (int i = 0; i < K * K; ++i) wk[i] = __ldg(w + (size_t)c * K * K + (flip ? K * K - 1 - i : i));
  const float bv = bias ? __ldg(bias + c) : 0.f;
  __syncthreads();
  float acc[4] = {bv, bv, bv, bv};
#pragma unroll
  for (int kc = 0; kc < K; ++kc) {
    float col[4 + K - 1];
#pragma unroll
    for (int r = 0; r < 4 + K - 1; ++r) col[r] = s[ty * 4 + r][tx + kc];
#pragma unroll
    for (int kr = 0; kr < K; ++kr)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[j] = fmaf(wk[kr * K + kc], col[j + kr], acc[j]);
  }
  const int gx = x0 + tx;
  if (gx >= W) return;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int gy = y0 + ty * 4 + j;
    if (gy < H) st_any(y, dt, yo + (int64_t)gy * W + gx, acc[j]);
  }
}

// Weight / bias gradient: parts[(b * tiles + tile)][c][K*K + 1] = sum over the tile of dy * shifted x | dy.
// Same staging as the forward kernel plus the dy tile; a thread keeps the K*K + 1 sums of its 4 pixels in registers,
// the CTA folds them in fixed order (warp butterfly, then 8 warp sums): deterministic.
template <int K>
__global__ void __launch_bounds__(256)
dwconv_kxk_wgrad_kernel(const void* __restrict__ x, int64_t x_bs, const void* __restrict__ dy, int64_t dy_bs, int dt,
                        float* __restrict__ parts, int C, int H, int W, int tiles_x, int tiles) {
  constexpr int R = K / 2, SW = KT + K - 1, NV = K * K + 1;
  __shared__ float s[SW][SW + 1];
  __shared__ float red[8][NV];
  const int tid = threadIdx.x, tx = tid & 31, ty = tid >> 5, warp = ty, lane = tx;
  const int tile = blockIdx.x % tiles, pc = blockIdx.x / tiles;
  const int c = pc % C, b = pc / C;
  const int x0 = (tile % tiles_x) * KT, y0 = (tile / tiles_x) * KT;
  const int64_t HWp = (int64_t)H * W;
  const int64_t xo = (int64_t)b * x_bs + (int64_t)c * HWp, go = (int64_t)b * dy_bs + (int64_t)c * HWp;
  for (int e = tid; e < SW * SW; e += 256) {
    const int r = e / SW, cc = e - r * SW;
    const int gy = y0 + r - R, gx = x0 + cc - R;
    s[r][cc] = (gy >= 0 && gy < H && gx >= 0 && gx < W) ? ld_any(x, dt, xo + (int64_t)gy * W + gx) : 0.f;
  }
  float g[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int gy = y0 + ty * 4 + j, gx = x0 + tx;
    g[j] = (gy < H && gx < W) ? ld_any(dy, dt, go + (int64_t)gy * W + gx) : 0.f;
  }
  __syncthreads();
  float acc[NV];
#pragma unroll
  for (int kc = 0; kc < K; ++kc) {
    float col[4 + K - 1];
#pragma unroll
    for (int r = 0; r < 4 + K - 1; ++r) col[r] = s[ty * 4 + r][tx + kc];
#pragma unroll
    for (int kr = 0; kr < K; ++kr) {
      float a = 0.f;
#pragma unroll
      for (int j = 0; j < 4; ++j) a = fmaf(g[j], col[j + kr], a);
      acc[kr * K + kc] = a;
    }
  }
  acc[K * K] = (g[0] + g[1]) + (g[2] + g[3]);
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const float v = warp_sum(acc[i]);
    if (lane == 0) red[warp][i] = v;
  }
  __syncthreads();
  if (tid < NV) {
    float t = 0.f;
#pragma unroll
    for (int wv = 0; wv < 8; ++wv) t += red[wv][tid];
    parts[(((size_t)b * tiles + tile) * C + c) * NV + tid] = t;
  }
}

}  // namespace b200ir

using namespace b200ir;

static int dwconv_kxk_launch(const void* x, int64_t x_bstride, const float* w, const float* bias, void* y,
                             int64_t y_bstride, int dtype, int B, int C, int H, int W, int K, int flip,
                             b200ir_stream_t stream) {
  B200IR_CHECK_ARG(x && w && y && B > 0 && C > 0 && H > 0 && W > 0, "dwconv_kxk: bad args");
  B200IR_CHECK_ARG(dtype_ok(dtype), "dwconv_kxk: bad dtype %d", dtype);
  B200IR_UNSUPPORTED(K == 3 || K == 5 || K == 7, "dwconv_kxk: kernel size %d unsupported (3, 5, 7)", K);
  const int tiles_x = (W + KT - 1) / KT, tiles_y = (H + KT - 1) / KT;
  const int64_t nblocks = (int64_t)B * C * tiles_x * tiles_y;
  B200IR_UNSUPPORTED(nblocks < (1ll << 31), "dwconv_kxk: grid too large");
  cudaStream_t st = (cudaStream_t)stream;
#define B200IR_KXK(KK)                                                                                       \
  dwconv_kxk_kernel<KK><<<(unsigned)nblocks, 256, 0, st>>>(x, x_bstride, w, bias, y, y_bstride, dtype, C, H, W, \
                                                            tiles_x, tiles_x * tiles_y, flip)
  if (K == 3) B200IR_KXK(3);
  else if (K == 5) B200IR_KXK(5);
  else B200IR_KXK(7);
#undef B200IR_KXK
  return check_launch("dwconv_kxk");
}

extern "C" int b200ir_dwconv_kxk_fwd(const void* x, int64_t x_bstride, const float* w, const float* bias, void* y,
                                     int64_t y_bstride, int dtype, int B, int C, int H, int W, int K,
                                     b200ir_stream_t stream) {
  return dwconv_kxk_launch(x, x_bstride, w, bias, y, y_bstride, dtype, B, C, H, W, K, 0, stream);
}

extern "C" int b200ir_dwconv_kxk_bwd_input(const void* dy, int64_t dy_bstride, const float* w, void* dx,
                                           int64_t dx_bstride, int dtype, int B, int C, int H, int W, int K,
                                           b200ir_stream_t stream) {
  return dwconv_kxk_launch(dy, dy_bstride, w, nullptr, dx, dx_bstride, dtype, B, C, H, W, K, 1, stream);
}

extern "C" int b200ir_dwconv_kxk_wgrad_nparts(int B, int H, int W) {
  if (B <= 0 || H <= 0 || W <= 0) return B200IR_E_BADARG;
  return B * ((W + KT - 1) / KT) * ((H + KT - 1) / KT);
}

extern "C" int b200ir_dwconv_kxk_wgrad(const void* dy, int64_t dy_bstride, const void* x, int64_t x_bstride, int dtype,
                                       float* parts, int nparts, int B, int C, int H, int W, int K,
                                       b200ir_stream_t stream) {
  B200IR_CHECK_ARG(dy && x && parts && B > 0 && C > 0 && H > 0 && W > 0, "dwconv_kxk_wgrad: bad args");
  B200IR_CHECK_ARG(dtype_ok(dtype), "dwconv_kxk_wgrad: bad dtype %d", dtype);
  B200IR_UNSUPPORTED(K == 3 || K == 5 || K == 7, "dwconv_kxk_wgrad: kernel size %d unsupported (3, 5, 7)", K);
  B200IR_CHECK_ARG(nparts == b200ir_dwconv_kxk_wgrad_nparts(B, H, W), "dwconv_kxk_wgrad: nparts=%d, expected %d", nparts,
                   b200ir_dwconv_kxk_wgrad_nparts(B, H, W));
  const int tiles_x = (W + KT - 1) / KT, tiles_y = (H + KT - 1) / KT;
  const int64_t nblocks = (int64_t)B * C * tiles_x * tiles_y;
  B200IR_UNSUPPORTED(nblocks < (1ll << 31), "dwconv_kxk_wgrad: grid too large");
  cudaStream_t st = (cudaStream_t)stream;
#define B200IR_KXKW(KK)                                                                                          \
  dwconv_kxk_wgrad_kernel<KK><<<(unsigned)nblocks, 256, 0, st>>>(x, x_bstride, dy, dy_bstride, dtype, parts, C, H, W, \
                                                                  tiles_x, tiles_x * tiles_y)
  if (K == 3) B200IR_KXKW(3);
  else if (K == 5) B200IR_KXKW(5);
  else B200IR_KXKW(7);
#undef B200IR_KXKW
  return check_launch("dwconv_kxk_wgrad");
}
