// pipeline_ops.cu — the training input pipeline and the evaluation metrics of the reference on the device
// (SURVEY.md §8 f4): at ~40 MP/s per GPU the PIL / numpy loader and the skimage metrics are the next bottleneck of
// train.py / test.py.
//   crop + flip/rot90 augmentation + Gaussian-noise synthesis + clamp + ToTensor:
//     MoCE-IR-main/src/data/dataset_utils.py:122-133,160-186 (_crop_patch, random_augmentation, ToTensor),
//     MoCE-IR-main/src/utils/image_utils.py:136-166 (data_augmentation modes 0..7),
//     MoCE-IR-main/src/data/degradation_utils.py:21-24 (clip(clean + noise * sigma, 0, 255).astype(uint8)).
//   PSNR / SSIM of clamp(restored, 0, 1) against the clean image: MoCE-IR-main/src/test.py:46-52,104-112
//     (skimage peak_signal_noise_ratio / structural_similarity(gaussian_weights=True, data_range=1, channel_axis=2):
//      11 x 11 Gaussian window sigma 1.5, sample covariance (x 121/120), K1 = 0.01, K2 = 0.03, mean over the image
//      without a 5-pixel border, then over channels).
#include "common.cuh"

namespace b200ir {

// source pixel of augmented output pixel (y, x) of a P x P patch (numpy flipud / rot90 composition, image_utils.py:136-166)
__device__ __forceinline__ void aug_src(int mode, int P, int y, int x, int& sy, int& sx) {
  switch (mode) {
    case 1: sy = P - 1 - y; sx = x; break;              // flipud
    case 2: sy = x; sx = P - 1 - y; break;              // rot90
    case 3: sy = x; sx = y; break;                      // rot90 + flipud (transpose)
    case 4: sy = P - 1 - y; sx = P - 1 - x; break;      // rot180
    case 5: sy = y; sx = P - 1 - x; break;              // rot180 + flipud (fliplr)
    case 6: sy = P - 1 - x; sx = y; break;              // rot270
    case 7: sy = P - 1 - x; sx = P - 1 - y; break;      // rot270 + flipud (anti-transpose)
    default: sy = y; sx = x; break;
  }
}

// src uint8 [B][Hs][Ws][3] (HWC like PIL / numpy), noise fp32 [B][P][P][3] (HWC, the layout np.random.randn(*patch.shape)
// has in the reference) or NULL, out clean / degraded fp32 [B][3][P][P] in [0, 1].
__global__ void __launch_bounds__(256)
crop_augment_degrade_kernel(const uint8_t* __restrict__ src, int Hs, int Ws, const int* __restrict__ offs,
                            const int* __restrict__ modes, const float* __restrict__ sigma, const float* __restrict__ noise,
                            float* __restrict__ clean, float* __restrict__ degraded, int P) {
  const int b = blockIdx.y;
  const int oy = offs[2 * b], ox = offs[2 * b + 1], mode = modes[b];
  const float sg = sigma ? sigma[b] : 0.f;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < P * P * 3; i += gridDim.x * 256) {
    const int ch = i % 3, x = (i / 3) % P, y = i / (3 * P);
    int sy, sx;
    aug_src(mode, P, y, x, sy, sx);
    const float v = (float)src[(((int64_t)b * Hs + oy + sy) * Ws + ox + sx) * 3 + ch];
    const int64_t o = (((int64_t)b * 3 + ch) * P + y) * P + x;
    clean[o] = v / 255.0f;                             // ToTensor divides (v * (1/255) differs by one ulp)
    if (degraded) {
      float d = v;
      if (noise && sg > 0.f) {
        d = fminf(fmaxf(v + noise[(int64_t)b * P * P * 3 + i] * sg, 0.f), 255.f);
        d = truncf(d);                                   // .astype(np.uint8) truncates
      }
      degraded[o] = d / 255.0f;
    }
  }
}

// per-image squared error of clamp(restored) vs clamp(clean): parts[b][block]
__global__ void __launch_bounds__(256)
sqerr_kernel(const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ parts, int64_t per, int nblk) {
  const int n = blockIdx.y;
  double s = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < per; i += (int64_t)nblk * 256) {
    const float x = fminf(fmaxf(a[(int64_t)n * per + i], 0.f), 1.f), y = fminf(fmaxf(b[(int64_t)n * per + i], 0.f), 1.f);
    s += (double)((x - y) * (x - y));
  }
  float f = warp_sum((float)s);
  __shared__ float red[8];
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = f;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int k = 0; k < 8; ++k) t += red[k];
    parts[(int64_t)n * nblk + blockIdx.x] = t;
  }
}

// SSIM map summed over the valid region: parts[b][c][block]; one thread per valid output pixel
__global__ void __launch_bounds__(256)
ssim_kernel(const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ parts, int H, int W, int nblk) {
  __shared__ float g[11];
  __shared__ float red[8];
  if (threadIdx.x < 11) {
    float s = 0.f;
    for (int k = -5; k <= 5; ++k) s += expf(-(float)(k * k) / (2.f * 1.5f * 1.5f));
    const int k = (int)threadIdx.x - 5;
    g[threadIdx.x] = expf(-(float)(k * k) / (2.f * 1.5f * 1.5f)) / s;
  }
  __syncthreads();
  const int c = blockIdx.y, n = blockIdx.z, C = gridDim.y;
  const float* pa = a + ((int64_t)n * C + c) * H * W;
  const float* pb = b + ((int64_t)n * C + c) * H * W;
  const int vh = H - 10, vw = W - 10;
  const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f, cov_norm = 121.f / 120.f;
  float acc = 0.f;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < vh * vw; i += nblk * 256) {
    const int y = i / vw + 5, x = i % vw + 5;
    float ux = 0.f, uy = 0.f, uxx = 0.f, uyy = 0.f, uxy = 0.f;
    for (int r = -5; r <= 5; ++r) {
      float rx = 0.f, ry = 0.f, rxx = 0.f, ryy = 0.f, rxy = 0.f;
      for (int s = -5; s <= 5; ++s) {
        const float va = fminf(fmaxf(pa[(int64_t)(y + r) * W + x + s], 0.f), 1.f);
        const float vb = fminf(fmaxf(pb[(int64_t)(y + r) * W + x + s], 0.f), 1.f);
        const float w = g[s + 5];
        rx = fmaf(w, va, rx); ry = fmaf(w, vb, ry);
        rxx = fmaf(w, va * va, rxx); ryy = fmaf(w, vb * vb, ryy); rxy = fmaf(w, va * vb, rxy);
      }
      const float w = g[r + 5];
      ux = fmaf(w, rx, ux); uy = fmaf(w, ry, uy); uxx = fmaf(w, rxx, uxx); uyy = fmaf(w, ryy, uyy); uxy = fmaf(w, rxy, uxy);
    }
    const float vx = cov_norm * (uxx - ux * ux), vy = cov_norm * (uyy - uy * uy), vxy = cov_norm * (uxy - ux * uy);
    acc += ((2.f * ux * uy + C1) * (2.f * vxy + C2)) / ((ux * ux + uy * uy + C1) * (vx + vy + C2));
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int k = 0; k < 8; ++k) t += red[k];
    parts[(((int64_t)n * C + c) * nblk) + blockIdx.x] = t;
  }
}

// out[b] = (psnr, ssim) from the partial sums (fixed order)
__global__ void metrics_finish_kernel(const float* __restrict__ se, const float* __restrict__ ss, float* __restrict__ out, int B,
                                      int C, int64_t per, int nblk_se, int nblk_ss, int valid) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= B) return;
  double e = 0.0;
  for (int k = 0; k < nblk_se; ++k) e += (double)se[(int64_t)n * nblk_se + k];
  const double mse = e / (double)per;
  out[2 * n] = mse > 0.0 ? (float)(10.0 * log10(1.0 / mse)) : INFINITY;
  double s = 0.0;
  for (int k = 0; k < C * nblk_ss; ++k) s += (double)ss[(int64_t)n * C * nblk_ss + k];
  out[2 * n + 1] = (float)(s / ((double)C * (double)valid));
}

}  // namespace b200ir

using namespace b200ir;

// ---- full-frame inference in tiles (the E / W recipe of the reference's tile helper, e.g.
// 20260104_CG_IR/src/data/dataset_utils.py:513-534: E[.., h:h+t, w:w+t] += out_patch, W += 1, restored = E / W) ----
namespace b200ir {
// patches[(ti * B + b)][c][y][x] = src[b][c][h0(ti) + y][w0(ti) + x]
__global__ void tile_gather_kernel(const float* __restrict__ src, const int* __restrict__ origins, float* __restrict__ dst,
                                   int B, int C, int H, int W, int t) {
  const int ti = blockIdx.z, b = blockIdx.y;
  const int h0 = origins[2 * ti], w0 = origins[2 * ti + 1];
  const int64_t per = (int64_t)C * t * t;
  float* d = dst + ((int64_t)ti * B + b) * per;
  const float* s = src + (int64_t)b * C * H * W;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < per; i += (int64_t)gridDim.x * blockDim.x) {
    const int x = (int)(i % t), y = (int)((i / t) % t), c = (int)(i / ((int64_t)t * t));
    d[i] = s[((int64_t)c * H + h0 + y) * W + w0 + x];
  }
}
// out[b][c][y][x] = clamp01( (sum over the tiles covering (y, x), in tile order, of their pixel) / (number of such tiles) ):
// pull form of the accumulate -- no atomics, and the same order of additions as the reference's loop over h_idx, w_idx
__global__ void tile_blend_kernel(const float* __restrict__ patches, const int* __restrict__ hs, int nh,
                                  const int* __restrict__ ws, int nw, float* __restrict__ out, int B, int C, int H, int W, int t,
                                  int clamp01) {
  const int b = blockIdx.y;
  const int64_t per = (int64_t)C * H * W, pp = (int64_t)C * t * t;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < per; i += (int64_t)gridDim.x * blockDim.x) {
    const int x = (int)(i % W), y = (int)((i / W) % H), c = (int)(i / ((int64_t)W * H));
    float e = 0.f, wsum = 0.f;
    for (int a = 0; a < nh; ++a) {
      const int dy = y - hs[a];
      if (dy < 0 || dy >= t) continue;
      for (int q = 0; q < nw; ++q) {
        const int dx = x - ws[q];
        if (dx < 0 || dx >= t) continue;
        e += patches[((int64_t)(a * nw + q) * B + b) * pp + ((int64_t)c * t + dy) * t + dx];
        wsum += 1.f;
      }
    }
    float v = e / wsum;
    if (clamp01) v = fminf(fmaxf(v, 0.f), 1.f);
    out[(int64_t)b * per + i] = v;
  }
}
}  // namespace b200ir

extern "C" int b200ir_tile_gather(const float* src, const int* origins, float* patches, int ntiles, int B, int C, int H, int W,
                                  int tile, b200ir_stream_t stream) {
  B200IR_CHECK_ARG(src && origins && patches && ntiles > 0 && ntiles <= 65535 && B > 0 && B <= 65535 && C > 0 && tile > 0 &&
                   tile <= H && tile <= W, "tile_gather: bad args");
  int gx = (int)(((int64_t)C * tile * tile + 255) / 256);
  if (gx > 64) gx = 64;
  b200ir::tile_gather_kernel<<<dim3((unsigned)gx, (unsigned)B, (unsigned)ntiles), 256, 0, (cudaStream_t)stream>>>(
      src, origins, patches, B, C, H, W, tile);
  return b200ir::check_launch("tile_gather");
}

extern "C" int b200ir_tile_blend(const float* patches, const int* h_idx, int nh, const int* w_idx, int nw, float* out, int B,
                                 int C, int H, int W, int tile, int clamp01, b200ir_stream_t stream) {
  B200IR_CHECK_ARG(patches && h_idx && w_idx && out && nh > 0 && nw > 0 && B > 0 && B <= 65535 && C > 0 && tile > 0 &&
                   tile <= H && tile <= W, "tile_blend: bad args");
  int64_t gx = ((int64_t)C * H * W + 255) / 256;
  if (gx > (int64_t)b200ir::num_sms() * 8) gx = (int64_t)b200ir::num_sms() * 8;
  b200ir::tile_blend_kernel<<<dim3((unsigned)gx, (unsigned)B), 256, 0, (cudaStream_t)stream>>>(patches, h_idx, nh, w_idx, nw, out,
                                                                                              B, C, H, W, tile, clamp01);
  return b200ir::check_launch("tile_blend");
}

extern "C" int b200ir_crop_augment_degrade(const uint8_t* src, int Hs, int Ws, const int* offsets, const int* modes,
                                           const float* sigma, const float* noise, float* clean, float* degraded, int B,
                                           int P, b200ir_stream_t stream) {
  B200IR_CHECK_ARG(src && offsets && modes && clean && B > 0 && B <= 65535 && P > 0 && P <= Hs && P <= Ws,
                   "crop_augment_degrade: bad args");
  int gx = (P * P * 3 + 255) / 256;
  if (gx > num_sms() * 4) gx = num_sms() * 4;
  crop_augment_degrade_kernel<<<dim3((unsigned)gx, (unsigned)B), 256, 0, (cudaStream_t)stream>>>(src, Hs, Ws, offsets, modes, sigma,
                                                                                                  noise, clean, degraded, P);
  return check_launch("crop_augment_degrade");
}

extern "C" int b200ir_psnr_ssim_nblocks(int H, int W) {
  int64_t n = ((int64_t)H * W + 255) / 256;
  return (int)(n > 64 ? 64 : (n < 1 ? 1 : n));
}

extern "C" int b200ir_psnr_ssim(const float* restored, const float* clean, float* ws, float* out, int B, int C, int H, int W,
                                b200ir_stream_t stream) {
  B200IR_CHECK_ARG(restored && clean && ws && out && B > 0 && B <= 65535 && C > 0 && C <= 65535, "psnr_ssim: bad args");
  B200IR_UNSUPPORTED(H >= 11 && W >= 11, "psnr_ssim: images smaller than the 11x11 SSIM window");
  const int nblk = b200ir_psnr_ssim_nblocks(H, W);
  const int64_t per = (int64_t)C * H * W;
  float* se = ws;                       // [B][nblk]
  float* ss = ws + (int64_t)B * nblk;   // [B][C][nblk]
  sqerr_kernel<<<dim3((unsigned)nblk, (unsigned)B), 256, 0, (cudaStream_t)stream>>>(restored, clean, se, per, nblk);
  ssim_kernel<<<dim3((unsigned)nblk, (unsigned)C, (unsigned)B), 256, 0, (cudaStream_t)stream>>>(restored, clean, ss, H, W, nblk);
  metrics_finish_kernel<<<(B + 127) / 128, 128, 0, (cudaStream_t)stream>>>(se, ss, out, B, C, per, nblk, nblk, (H - 10) * (W - 10));
  return check_launch("psnr_ssim");
}
