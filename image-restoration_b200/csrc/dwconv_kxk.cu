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

// bf16 fast path (round 2; the 7x7 kv convolutions of CrossAttention were 4.4 ms of the 31.7 ms MoCE-IR forward on the kernel
// above: 1 024 outputs per CTA behind one latency-bound scalar load phase).  Tile = 64 x 32 outputs of one (sample, channel)
// plane; the halo tile is fetched with 16-byte loads over the 8-pixel-aligned span [x0 - 8, x0 + 72) (requires W % 8 == 0 and
// 16-byte aligned planes: vectors are then entirely inside or entirely outside the image) and kept as fp32 in shared memory;
// a thread owns 2 adjacent pixels x 4 rows: per input row it reads its 8 + ... columns with 64-bit shared loads once and
// feeds every (output row, tap column) that uses them -> 392 FMAs per 40 shared loads, bf16x2 stores (128 B per warp).
template <int K>
__global__ void __launch_bounds__(256)
dwconv_kxk_bf16_kernel(const bf16* __restrict__ x, int64_t x_bs, const float* __restrict__ w, const float* __restrict__ bias,
                       bf16* __restrict__ y, int64_t y_bs, int C, int H, int W, int tiles_x, int tiles, int flip) {
  constexpr int R = K / 2, TX = 64, TY = 32, SH = TY + K - 1, SWD = TX + 16;   // smem columns <-> gx = x0 - 8 + col
  constexpr int NV = 2 + K - 1;                                                // columns a thread needs per input row
  __shared__ __align__(16) float s[SH][SWD];
  const int tid = threadIdx.x, tx = tid & 31, ty = tid >> 5;
  const int tile = blockIdx.x % tiles, pc = blockIdx.x / tiles;
  const int c = pc % C, b = pc / C;
  const int x0 = (tile % tiles_x) * TX, y0 = (tile / tiles_x) * TY;
  const int64_t HWp = (int64_t)H * W;
  const bf16* xp = x + (int64_t)b * x_bs + (int64_t)c * HWp;
  bf16* yp = y + (int64_t)b * y_bs + (int64_t)c * HWp;
  for (int e = tid; e < SH * (SWD / 8); e += 256) {
    const int r = e / (SWD / 8), v = e - r * (SWD / 8);
    const int gy = y0 + r - R, gx = x0 - 8 + v * 8;
    uint4 q = make_uint4(0u, 0u, 0u, 0u);
    if (gy >= 0 && gy < H && gx >= 0 && gx < W) q = __ldg(reinterpret_cast<const uint4*>(xp + (int64_t)gy * W + gx));
    float4 lo, hi;
    lo.x = __uint_as_float(q.x << 16); lo.y = __uint_as_float(q.x & 0xffff0000u);
    lo.z = __uint_as_float(q.y << 16); lo.w = __uint_as_float(q.y & 0xffff0000u);
    hi.x = __uint_as_float(q.z << 16); hi.y = __uint_as_float(q.z & 0xffff0000u);
    hi.z = __uint_as_float(q.w << 16); hi.w = __uint_as_float(q.w & 0xffff0000u);
    *reinterpret_cast<float4*>(&s[r][v * 8]) = lo;
    *reinterpret_cast<float4*>(&s[r][v * 8 + 4]) = hi;
  }
  float wk[K * K];
#pragma unroll
  for (int i = 0; i < K * K; ++i) wk[i] = __ldg(w + (size_t)c * K * K + (flip ? K * K - 1 - i : i));
  const float bv = bias ? __ldg(bias + c) : 0.f;
  __syncthreads();
  float acc[4][2];
#pragma unroll
  for (int j = 0; j < 4; ++j) acc[j][0] = acc[j][1] = bv;
  // output pixel px (0/1) at tile column 2*tx + px needs smem columns 2*tx + px + kc + 8 - R, kc = 0..K-1
  constexpr int OFF = 8 - R;                       // 5 / 6 / 7: first needed column relative to 2*tx
  constexpr int A0 = OFF & ~1;                     // even start -> 8-byte aligned shared loads
  constexpr int NL = (OFF - A0 + NV + 1) / 2;      // float2 loads per input row
#pragma unroll
  for (int r = 0; r < 4 + K - 1; ++r) {
    float v[2 * NL];
    const float2* row = reinterpret_cast<const float2*>(&s[ty * 4 + r][2 * tx + A0]);
#pragma unroll
    for (int i = 0; i < NL; ++i) {
      const float2 t = row[i];
      v[2 * i] = t.x;
      v[2 * i + 1] = t.y;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int kr = r - j;
      if (kr < 0 || kr >= K) continue;
#pragma unroll
      for (int kc = 0; kc < K; ++kc) {
        acc[j][0] = fmaf(wk[kr * K + kc], v[OFF - A0 + kc], acc[j][0]);
        acc[j][1] = fmaf(wk[kr * K + kc], v[OFF - A0 + kc + 1], acc[j][1]);
      }
    }
  }
  const int gx = x0 + 2 * tx;
  if (gx >= W) return;                             // W is even: both pixels of the pair are inside or outside
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int gy = y0 + ty * 4 + j;
    if (gy < H) *reinterpret_cast<__nv_bfloat162*>(yp + (int64_t)gy * W + gx) = __floats2bfloat162_rn(acc[j][0], acc[j][1]);
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
  if (dtype == B200IR_BF16 && W % 8 == 0 && x_bstride % 8 == 0 && y_bstride % 8 == 0 &&
      ((uintptr_t)x & 15) == 0 && ((uintptr_t)y & 3) == 0) {
    const int tx2 = (W + 63) / 64, ty2 = (H + 31) / 32;
    const int64_t nb2 = (int64_t)B * C * tx2 * ty2;
#define B200IR_KXK2(KK)                                                                                             \
  dwconv_kxk_bf16_kernel<KK><<<(unsigned)nb2, 256, 0, st>>>((const bf16*)x, x_bstride, w, bias, (bf16*)y, y_bstride, C, H, \
                                                             W, tx2, tx2 * ty2, flip)
    if (K == 3) B200IR_KXK2(3);
    else if (K == 5) B200IR_KXK2(5);
    else B200IR_KXK2(7);
#undef B200IR_KXK2
    return check_launch("dwconv_kxk_bf16");
  }
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
