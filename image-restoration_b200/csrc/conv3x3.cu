// conv3x3.cu — dense 3x3 conv (pad 1) on CUDA cores with PixelUnshuffle(2)/PixelShuffle(2) folded
// into the store addressing, optional bias and residual.
// Reference: OverlapPatchEmbed Restormer.py:162-165; Downsample :171-179 (conv + nn.PixelUnshuffle(2):
// out[c*4+i*2+j, y, x] = t[c, 2y+i, 2x+j]); Upsample :181-189 (conv + nn.PixelShuffle(2), the inverse);
// output conv + inp_img :281.
#include "common.cuh"

namespace b200ir {

constexpr int CT = 16;   // spatial tile 16x16, one pixel per thread
constexpr int COT = 8;   // output channels per thread
constexpr int CIK = 8;   // input channels per smem chunk

__global__ void __launch_bounds__(256)
conv3x3_simt_kernel(const void* __restrict__ x, int64_t x_bs, int x_dt, const float* __restrict__ w,
                    const float* __restrict__ bias, const void* __restrict__ res, int64_t res_bs, int res_dt,
                    void* __restrict__ y, int64_t y_bs, int y_dt, int Cin, int Cout, int H, int W, int mode,
                    int tiles_x) {
  __shared__ float xs[CIK][CT + 2][CT + 3];
  __shared__ __align__(16) float ws[CIK][9][COT];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int tile = blockIdx.x, b = blockIdx.z;
  const int x0 = (tile % tiles_x) * CT, y0 = (tile / tiles_x) * CT;
  const int co0 = blockIdx.y * COT;
  const int64_t HWp = (int64_t)H * W;

  float acc[COT];
#pragma unroll
  for (int i = 0; i < COT; ++i) acc[i] = 0.f;

  for (int c0 = 0; c0 < Cin; c0 += CIK) {
    for (int e = tid; e < CIK * (CT + 2) * (CT + 2); e += 256) {
      const int ci = e / ((CT + 2) * (CT + 2));
      const int r = (e / (CT + 2)) % (CT + 2), cc = e % (CT + 2);
      const int gy = y0 + r - 1, gx = x0 + cc - 1, c = c0 + ci;
      float v = 0.f;
      if (c < Cin && gy >= 0 && gy < H && gx >= 0 && gx < W)
        v = ld_any(x, x_dt, (int64_t)b * x_bs + (int64_t)c * HWp + (int64_t)gy * W + gx);
      xs[ci][r][cc] = v;
    }
    for (int e = tid; e < CIK * 9 * COT; e += 256) {
      const int co = e / (CIK * 9), rem = e % (CIK * 9);
      const int ci = rem / 9, t = rem % 9;
      float v = 0.f;
      if (co0 + co < Cout && c0 + ci < Cin) v = w[((size_t)(co0 + co) * Cin + c0 + ci) * 9 + t];
      ws[ci][t][co] = v;
    }
    __syncthreads();
#pragma unroll 2
    for (int ci = 0; ci < CIK; ++ci) {
#pragma unroll
      for (int t = 0; t < 9; ++t) {
        const float xv = xs[ci][ty + t / 3][tx + t % 3];
        const float4 w0 = *reinterpret_cast<const float4*>(&ws[ci][t][0]);
        const float4 w1 = *reinterpret_cast<const float4*>(&ws[ci][t][4]);
        acc[0] = fmaf(w0.x, xv, acc[0]);
        acc[1] = fmaf(w0.y, xv, acc[1]);
        acc[2] = fmaf(w0.z, xv, acc[2]);
        acc[3] = fmaf(w0.w, xv, acc[3]);
        acc[4] = fmaf(w1.x, xv, acc[4]);
        acc[5] = fmaf(w1.y, xv, acc[5]);
        acc[6] = fmaf(w1.z, xv, acc[6]);
        acc[7] = fmaf(w1.w, xv, acc[7]);
      }
    }
    __syncthreads();
  }

  const int gy = y0 + ty, gx = x0 + tx;
  if (gy >= H || gx >= W) return;
#pragma unroll
  for (int i = 0; i < COT; ++i) {
    const int co = co0 + i;
    if (co >= Cout) break;
    float v = acc[i] + (bias ? bias[co] : 0.f);
    int64_t off;
    if (mode == B200IR_UNSHUFFLE2) {
      const int oc = co * 4 + (gy & 1) * 2 + (gx & 1);
      off = (int64_t)oc * (HWp / 4) + (int64_t)(gy >> 1) * (W / 2) + (gx >> 1);
    } else if (mode == B200IR_SHUFFLE2) {
      const int oc = co >> 2, si = (co >> 1) & 1, sj = co & 1;
      off = (int64_t)oc * (HWp * 4) + (int64_t)(2 * gy + si) * (2 * W) + (2 * gx + sj);
    } else {
      off = (int64_t)co * HWp + (int64_t)gy * W + gx;
    }
    if (res) v += ld_any(res, res_dt, (int64_t)b * res_bs + off);
    st_any(y, y_dt, (int64_t)b * y_bs + off, v);
  }
}

int conv3x3_thin_launch(const void* x, int64_t x_bs, int x_dt, const float* w, const float* bias, const void* res,
                        int64_t res_bs, int res_dt, void* y, int64_t y_bs, int y_dt, int B, int Cin, int Cout, int H,
                        int W, cudaStream_t st);  // conv3x3_thin.cu

}  // namespace b200ir

using namespace b200ir;

extern "C" int b200ir_conv3x3_fwd(const void* x, int64_t x_bstride, int x_dtype, const float* w, const float* bias,
                                  const void* res, int64_t res_bstride, int res_dtype, void* y, int64_t y_bstride,
                                  int y_dtype, int B, int Cin, int Cout, int H, int W, int shuffle_mode,
                                  b200ir_stream_t stream) {
  B200IR_CHECK_ARG(x && w && y && B > 0 && Cin > 0 && Cout > 0 && H > 0 && W > 0, "conv3x3: bad args");
  B200IR_CHECK_ARG(dtype_ok(x_dtype) && dtype_ok(y_dtype) && (!res || dtype_ok(res_dtype)), "conv3x3: bad dtype");
  B200IR_CHECK_ARG(shuffle_mode >= 0 && shuffle_mode <= 2, "conv3x3: bad shuffle_mode %d", shuffle_mode);
  B200IR_CHECK_ARG(shuffle_mode != B200IR_UNSHUFFLE2 || (H % 2 == 0 && W % 2 == 0),
                   "conv3x3: PixelUnshuffle(2) needs even H,W (got %dx%d)", H, W);
  B200IR_CHECK_ARG(shuffle_mode != B200IR_SHUFFLE2 || Cout % 4 == 0, "conv3x3: PixelShuffle(2) needs Cout%%4==0");
  B200IR_CHECK_ARG(B <= 65535 && (Cout + COT - 1) / COT <= 65535, "conv3x3: grid too large");
  if (shuffle_mode == B200IR_SHUFFLE_NONE) {  // patch-embed / output conv: one side of the channel pair is tiny
    const int rc = conv3x3_thin_launch(x, x_bstride, x_dtype, w, bias, res, res_bstride, res_dtype, y, y_bstride,
                                       y_dtype, B, Cin, Cout, H, W, (cudaStream_t)stream);
    if (rc >= 0) return rc;   // -1: not a thin shape
  }
  const int tiles_x = (W + CT - 1) / CT, tiles_y = (H + CT - 1) / CT;
  dim3 grid((unsigned)(tiles_x * tiles_y), (unsigned)((Cout + COT - 1) / COT), (unsigned)B);
  conv3x3_simt_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x, x_bstride, x_dtype, w, bias, res, res_bstride,
                                                              res_dtype, y, y_bstride, y_dtype, Cin, Cout, H, W,
                                                              shuffle_mode, tiles_x);
  return check_launch("conv3x3_simt");
}
