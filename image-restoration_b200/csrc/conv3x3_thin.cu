// conv3x3_thin.cu — the two "thin" dense 3x3 convolutions at the ends of the U-Net (Restormer.py:220 patch_embed
// OverlapPatchEmbed 3 -> dim, Restormer.py:257 output dim*2 -> 3 with the `+ inp_img` residual of :306).  One side of
// the channel pair is tiny, so neither tensor cores nor the generic 8-outputs-per-thread SIMT kernel fit: these are
// HBM-bound streaming kernels with exact fp32 accumulation.
//   thin-in  (Cin <= 4): thread = 2 horizontally adjacent pixels x ALL output channels (8 at a time); the 3x4xCin input
//                        window lives in registers, weights come as broadcast LDS.128; stores are pixel pairs.
//   thin-out (Cout <= 4): thread = 4 pixels in a row x all outputs; input channels are staged 8 at a time as fp32 halo
//                        tiles (128 x 8 pixels) in shared memory.
#include "common.cuh"

namespace b200ir {

// ------------------------------------------------------------------------------------------------
// thin-in
// ------------------------------------------------------------------------------------------------
constexpr int TI_TW = 64, TI_TH = 8;  // CTA tile in pixels (256 threads x 2 px)

template <int CIN>
__global__ void __launch_bounds__(256)
conv3x3_thin_in_kernel(const void* __restrict__ x, int64_t x_bs, int x_dt, const float* __restrict__ w,
                       const float* __restrict__ bias, const void* __restrict__ res, int64_t res_bs, int res_dt,
                       void* __restrict__ y, int64_t y_bs, int y_dt, int Cout, int CoutPad, int H, int W,
                       int tiles_x) {
  extern __shared__ __align__(16) float sm_in[];
  float* wsm = sm_in;                                      // [CIN*9][CoutPad]
  float* xs = wsm + CIN * 9 * CoutPad;                     // [CIN][TI_TH+2][TI_TW+2]
  const int tid = threadIdx.x, tx = tid & 31, ty = tid >> 5;
  const int tile = blockIdx.x, b = blockIdx.y;
  const int x0 = (tile % tiles_x) * TI_TW, y0 = (tile / tiles_x) * TI_TH;
  const int64_t HWp = (int64_t)H * W;
  for (int e = tid; e < CIN * 9 * CoutPad; e += 256) {
    const int k = e / CoutPad, co = e - k * CoutPad;       // k = ci*9 + tap
    wsm[e] = co < Cout ? w[(size_t)co * CIN * 9 + k] : 0.f;
  }
  constexpr int XW = TI_TW + 2, XH = TI_TH + 2;
  for (int e = tid; e < CIN * XH * XW; e += 256) {
    const int ci = e / (XH * XW), r = (e / XW) % XH, cc = e % XW;
    const int gy = y0 + r - 1, gx = x0 + cc - 1;
    float v = 0.f;
    if (gy >= 0 && gy < H && gx >= 0 && gx < W) v = ld_any(x, x_dt, (int64_t)b * x_bs + ci * HWp + (int64_t)gy * W + gx);
    xs[e] = v;
  }
  __syncthreads();
  float win[CIN][3][4];
#pragma unroll
  for (int ci = 0; ci < CIN; ++ci)
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int c = 0; c < 4; ++c) win[ci][r][c] = xs[(ci * XH + ty + r) * XW + 2 * tx + c];
  const int gy = y0 + ty, gx = x0 + 2 * tx;
  if (gy >= H || gx >= W) return;
  const bool two = gx + 1 < W;
  const int64_t pix = (int64_t)gy * W + gx;
  const bool vec = two && ((W & 1) == 0) && ((y_bs & 1) == 0);   // pixel pair is 4-/8-byte aligned in every channel
  for (int c0 = 0; c0 < Cout; c0 += 8) {
    float a0[8], a1[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) a0[j] = a1[j] = 0.f;
#pragma unroll
    for (int ci = 0; ci < CIN; ++ci)
#pragma unroll
      for (int t = 0; t < 9; ++t) {
        const float4 w0 = *reinterpret_cast<const float4*>(&wsm[(ci * 9 + t) * CoutPad + c0]);
        const float4 w1 = *reinterpret_cast<const float4*>(&wsm[(ci * 9 + t) * CoutPad + c0 + 4]);
        const float wv[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
        const float xa = win[ci][t / 3][t % 3], xb = win[ci][t / 3][t % 3 + 1];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          a0[j] = fmaf(wv[j], xa, a0[j]);
          a1[j] = fmaf(wv[j], xb, a1[j]);
        }
      }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int co = c0 + j;
      if (co >= Cout) break;
      const float bv = bias ? __ldg(bias + co) : 0.f;
      float v0 = a0[j] + bv, v1 = a1[j] + bv;
      const int64_t off = (int64_t)co * HWp + pix;
      if (res) {
        v0 += ld_any(res, res_dt, (int64_t)b * res_bs + off);
        if (two) v1 += ld_any(res, res_dt, (int64_t)b * res_bs + off + 1);
      }
      const int64_t yo = (int64_t)b * y_bs + off;
      if (vec) {
        if (y_dt == B200IR_BF16) *reinterpret_cast<__nv_bfloat162*>((bf16*)y + yo) = __floats2bfloat162_rn(v0, v1);
        else *reinterpret_cast<float2*>((float*)y + yo) = make_float2(v0, v1);
      } else {
        st_any(y, y_dt, yo, v0);
        if (two) st_any(y, y_dt, yo + 1, v1);
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------
// thin-out
// ------------------------------------------------------------------------------------------------
constexpr int TO_TW = 128, TO_TH = 8, TO_CK = 8;
constexpr int TO_XW = TO_TW + 8;  // interior starts at column 4 (16-byte aligned), halos at 3 and 4+TO_TW

template <int COUT>
__global__ void __launch_bounds__(256)
conv3x3_thin_out_kernel(const void* __restrict__ x, int64_t x_bs, int x_dt, const float* __restrict__ w,
                        const float* __restrict__ bias, const void* __restrict__ res, int64_t res_bs, int res_dt,
                        void* __restrict__ y, int64_t y_bs, int y_dt, int Cin, int H, int W, int tiles_x) {
  __shared__ __align__(16) float xs[TO_CK][TO_TH + 2][TO_XW];
  __shared__ float wsm[TO_CK][9 * COUT + 1];
  const int tid = threadIdx.x, tx = tid & 31, ty = tid >> 5;
  const int tile = blockIdx.x, b = blockIdx.y;
  const int x0 = (tile % tiles_x) * TO_TW, y0 = (tile / tiles_x) * TO_TH;
  const int64_t HWp = (int64_t)H * W;
  const bool fast = x_dt == B200IR_BF16 && (W & 7) == 0 && (x_bs & 7) == 0 && x0 + TO_TW <= W &&
                    (reinterpret_cast<uintptr_t>(x) & 15) == 0;
  float acc[COUT][4];
#pragma unroll
  for (int o = 0; o < COUT; ++o)
#pragma unroll
    for (int p = 0; p < 4; ++p) acc[o][p] = 0.f;

  for (int c0 = 0; c0 < Cin; c0 += TO_CK) {
    __syncthreads();
    if (fast) {
      // interior: 16-byte loads of 8 bf16 pixels -> two STS.128
      for (int e = tid; e < TO_CK * (TO_TH + 2) * (TO_TW / 8); e += 256) {
        const int v8 = e % (TO_TW / 8), r = (e / (TO_TW / 8)) % (TO_TH + 2), ci = e / ((TO_TW / 8) * (TO_TH + 2));
        const int gy = y0 + r - 1, c = c0 + ci;
        float f[8];
        if (c < Cin && gy >= 0 && gy < H) {
          const uint4 u = __ldg(reinterpret_cast<const uint4*>((const bf16*)x + (int64_t)b * x_bs + c * HWp +
                                                               (int64_t)gy * W + x0 + v8 * 8));
          const uint32_t uu[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            f[2 * k] = __uint_as_float(uu[k] << 16);
            f[2 * k + 1] = __uint_as_float(uu[k] & 0xffff0000u);
          }
        } else {
#pragma unroll
          for (int k = 0; k < 8; ++k) f[k] = 0.f;
        }
        float4* d = reinterpret_cast<float4*>(&xs[ci][r][4 + v8 * 8]);
        d[0] = make_float4(f[0], f[1], f[2], f[3]);
        d[1] = make_float4(f[4], f[5], f[6], f[7]);
      }
      for (int e = tid; e < TO_CK * (TO_TH + 2) * 2; e += 256) {
        const int side = e & 1, r = (e >> 1) % (TO_TH + 2), ci = (e >> 1) / (TO_TH + 2);
        const int gy = y0 + r - 1, gx = side ? x0 + TO_TW : x0 - 1, c = c0 + ci;
        float v = 0.f;
        if (c < Cin && gy >= 0 && gy < H && gx >= 0 && gx < W)
          v = ld_any(x, x_dt, (int64_t)b * x_bs + c * HWp + (int64_t)gy * W + gx);
        xs[ci][r][side ? 4 + TO_TW : 3] = v;
      }
    } else {
      for (int e = tid; e < TO_CK * (TO_TH + 2) * (TO_TW + 2); e += 256) {
        const int cc = e % (TO_TW + 2), r = (e / (TO_TW + 2)) % (TO_TH + 2), ci = e / ((TO_TW + 2) * (TO_TH + 2));
        const int gy = y0 + r - 1, gx = x0 + cc - 1, c = c0 + ci;
        float v = 0.f;
        if (c < Cin && gy >= 0 && gy < H && gx >= 0 && gx < W)
          v = ld_any(x, x_dt, (int64_t)b * x_bs + c * HWp + (int64_t)gy * W + gx);
        xs[ci][r][3 + cc] = v;
      }
    }
    for (int e = tid; e < TO_CK * 9 * COUT; e += 256) {
      const int ci = e / (9 * COUT), k = e - ci * 9 * COUT;   // k = tap*COUT + o
      const int t = k / COUT, o = k - t * COUT;
      wsm[ci][k] = c0 + ci < Cin ? w[((size_t)o * Cin + c0 + ci) * 9 + t] : 0.f;
    }
    __syncthreads();
#pragma unroll 2
    for (int ci = 0; ci < TO_CK; ++ci) {
      float win[3][6];
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const float* row = &xs[ci][ty + r][4 * tx];
        const float4 m = *reinterpret_cast<const float4*>(row + 4);
        win[r][0] = row[3];
        win[r][1] = m.x; win[r][2] = m.y; win[r][3] = m.z; win[r][4] = m.w;
        win[r][5] = row[8];
      }
#pragma unroll
      for (int t = 0; t < 9; ++t)
#pragma unroll
        for (int o = 0; o < COUT; ++o) {
          const float wv = wsm[ci][t * COUT + o];
#pragma unroll
          for (int p = 0; p < 4; ++p) acc[o][p] = fmaf(wv, win[t / 3][t % 3 + p], acc[o][p]);
        }
    }
  }
  const int gy = y0 + ty, gx = x0 + 4 * tx;
  if (gy >= H) return;
#pragma unroll
  for (int o = 0; o < COUT; ++o) {
    const float bv = bias ? __ldg(bias + o) : 0.f;
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      if (gx + p >= W) break;
      const int64_t off = (int64_t)o * HWp + (int64_t)gy * W + gx + p;
      float v = acc[o][p] + bv;
      if (res) v += ld_any(res, res_dt, (int64_t)b * res_bs + off);
      st_any(y, y_dt, (int64_t)b * y_bs + off, v);
    }
  }
}

// returns -1 if this shape is not a thin conv, else the launch status (0 = ok)
int conv3x3_thin_launch(const void* x, int64_t x_bs, int x_dt, const float* w, const float* bias, const void* res,
                        int64_t res_bs, int res_dt, void* y, int64_t y_bs, int y_dt, int B, int Cin, int Cout, int H,
                        int W, cudaStream_t st) {
  if (B > 65535) return -1;
  if ((Cin == 1 || Cin == 3) && Cout >= 8 && Cout <= 256) {
    const int CoutPad = (Cout + 7) / 8 * 8;
    const int tiles_x = (W + TI_TW - 1) / TI_TW, tiles_y = (H + TI_TH - 1) / TI_TH;
    const size_t smem = ((size_t)Cin * 9 * CoutPad + (size_t)Cin * (TI_TH + 2) * (TI_TW + 2)) * sizeof(float);
    dim3 grid((unsigned)(tiles_x * tiles_y), (unsigned)B);
    if (Cin == 3)
      conv3x3_thin_in_kernel<3><<<grid, 256, smem, st>>>(x, x_bs, x_dt, w, bias, res, res_bs, res_dt, y, y_bs, y_dt,
                                                         Cout, CoutPad, H, W, tiles_x);
    else
      conv3x3_thin_in_kernel<1><<<grid, 256, smem, st>>>(x, x_bs, x_dt, w, bias, res, res_bs, res_dt, y, y_bs, y_dt,
                                                         Cout, CoutPad, H, W, tiles_x);
    return check_launch("conv3x3_thin_in");
  }
  if ((Cout == 1 || Cout == 3) && Cin >= 8) {
    const int tiles_x = (W + TO_TW - 1) / TO_TW, tiles_y = (H + TO_TH - 1) / TO_TH;
    dim3 grid((unsigned)(tiles_x * tiles_y), (unsigned)B);
    if (Cout == 3)
      conv3x3_thin_out_kernel<3><<<grid, 256, 0, st>>>(x, x_bs, x_dt, w, bias, res, res_bs, res_dt, y, y_bs, y_dt, Cin,
                                                       H, W, tiles_x);
    else
      conv3x3_thin_out_kernel<1><<<grid, 256, 0, st>>>(x, x_bs, x_dt, w, bias, res, res_bs, res_dt, y, y_bs, y_dt, Cin,
                                                       H, W, tiles_x);
    return check_launch("conv3x3_thin_out");
  }
  return -1;
}

}  // namespace b200ir
