// backward.cu — backward (training) kernels of the Restormer block path, dtype-generic CUDA-core versions.
// These are the gradients of the forward kernels in gemm_simt.cu / dwconv.cu / mdta.cu / conv3x3.cu, i.e. of
// Restormer.py:37-39,54-57 (LayerNorm), :89-92 (GDFN), :114-131 (MDTA), :162-189 (3x3 convs + pixel shuffles), as the
// reference gets them from torch autograd (MoCE-IR-main/src/train.py:59-72: loss.backward inside Lightning).
// Input-gradient GEMMs / depthwise convs reuse the forward kernels with transposed / flipped weights (host side);
// what lives here are the reductions over pixels (weight / bias / LayerNorm-parameter gradients), the LayerNorm and
// GELU-gate input gradients, and the small per-(sample, head) softmax / normalisation backward of MDTA.
// Every reduction is split into fixed-order partials + b200ir_reduce_partials: bitwise deterministic.
#include <math.h>
#include <mutex>

#include "common.cuh"

namespace b200ir {

// =====================================================================================================
// LayerNorm kernels.  CTA = 32 pixels x 8 warps; lane = pixel (coalesced along HW), warp w owns channels w, w+8, ...
// Per-pixel sums over C are combined across the 8 warps through shared memory in fixed order.
// =====================================================================================================
constexpr int LN_PX = 32;

// DT: compile-time activation dtype of ALL tensors of the launch (B200IR_F32 / B200IR_BF16), or -1 = per-tensor runtime
// dtype (then every load is a branch and the loads of an unrolled channel loop cannot be issued together).
template <int DT> __device__ __forceinline__ float ldt(const void* p, int dt, int64_t i) {
  if (DT == B200IR_BF16) return __bfloat162float(((const bf16*)p)[i]);
  if (DT == B200IR_F32) return ((const float*)p)[i];
  return ld_any(p, dt, i);
}
template <int DT> __device__ __forceinline__ void stt(void* p, int dt, int64_t i, float v) {
  if (DT == B200IR_BF16) ((bf16*)p)[i] = __float2bfloat16_rn(v);
  else if (DT == B200IR_F32) ((float*)p)[i] = v;
  else st_any(p, dt, i, v);
}

// y = LN(x) materialised (operand of the weight-gradient GEMM of the fused LN + 1x1 conv)
template <int DT>
__global__ void __launch_bounds__(256)
ln_fwd_kernel(const void* __restrict__ x, int64_t x_bs, int x_dt, int ln_type, const float* __restrict__ lw,
              const float* __restrict__ lb, void* __restrict__ y, int64_t y_bs, int y_dt, int C, int64_t HW, int tps,
              int64_t ntiles) {
  __shared__ float red2[8][2][LN_PX];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const float invC = 1.0f / (float)C;
  for (int64_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
    const int b = (int)(t / tps);
    const int64_t p = (t % tps) * LN_PX + lane;
    const bool valid = p < HW;
    const int64_t xo = (int64_t)b * x_bs + p, yo = (int64_t)b * y_bs + p;
    const float x0 = valid ? ldt<DT>(x, x_dt, xo) : 0.f;   // shift: keeps E[d^2] - E[d]^2 well conditioned
    float S1 = 0.f, S2 = 0.f;
    if (valid) {
#pragma unroll 4
      for (int c = warp; c < C; c += 8) {
        const float d = ldt<DT>(x, x_dt, xo + (int64_t)c * HW) - x0;
        S1 += d;
        S2 = fmaf(d, d, S2);
      }
    }
    __syncthreads();
    red2[warp][0][lane] = S1;
    red2[warp][1][lane] = S2;
    __syncthreads();
    S1 = S2 = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) {
      S1 += red2[w][0][lane];
      S2 += red2[w][1][lane];
    }
    const float md = S1 * invC, mu = x0 + md;
    const float rstd = 1.0f / sqrtf(fmaxf(S2 * invC - md * md, 0.f) + 1e-5f);
    if (valid) {
#pragma unroll 4
      for (int c = warp; c < C; c += 8) {
        const float xv = ldt<DT>(x, x_dt, xo + (int64_t)c * HW);
        float n;
        if (ln_type == B200IR_LN_WITHBIAS) n = (xv - mu) * rstd * __ldg(lw + c) + __ldg(lb + c);   // Restormer.py:57
        else n = xv * rstd * __ldg(lw + c);                                                       // Restormer.py:39
        stt<DT>(y, y_dt, yo + (int64_t)c * HW, n);
      }
    }
  }
}

// LayerNorm backward.  g = dL/dLN(x), h = g*gamma, xh = (x-mu)*rstd:
//   WithBias: dx = rstd*(h - mean(h) - xh*mean(h*xh)),  dgamma = sum_p g*xh,  dbeta = sum_p g
//   BiasFree: dx = rstd*h - rstd^3*(x-mu)*sum_c(h*x)/C,  dgamma = sum_p g*x*rstd
// dx (+)= res (gradient arriving through the residual path of TransformerBlock, Restormer.py:147-148).
// parts: fp32 [gridDim.x][2][C] partial dgamma / dbeta; a channel is only ever touched by one warp of the CTA, which
// accumulates it tile after tile in shared memory: fixed order, deterministic.
template <int DT, int NW>
__global__ void __launch_bounds__(32 * NW)
ln_bwd_kernel(const void* __restrict__ x, int64_t x_bs, int x_dt, const void* __restrict__ g, int64_t g_bs, int g_dt,
              int ln_type, const float* __restrict__ lw, const float* __restrict__ lb, const void* __restrict__ res,
              int64_t res_bs, int res_dt, void* __restrict__ dx, int64_t dx_bs, int dx_dt, void* __restrict__ xn,
              int64_t xn_bs, float* __restrict__ parts, int C, int64_t HW, int tps, int64_t ntiles) {
  extern __shared__ float acc[];  // [2][C]
  __shared__ float red4[NW][4][LN_PX];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 2 * C; i += 32 * NW) acc[i] = 0.f;
  __syncthreads();
  const float invC = 1.0f / (float)C;
  for (int64_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
    const int b = (int)(t / tps);
    const int64_t p = (t % tps) * LN_PX + lane;
    const bool valid = p < HW;
    const int64_t xo = (int64_t)b * x_bs + p, go = (int64_t)b * g_bs + p;
    // pass A (one trip over x and g): shifted sums S1 = sum(x-x0), S2 = sum((x-x0)^2), H1 = sum(h), H2 = sum(h*(x-x0))
    const float x0 = valid ? ldt<DT>(x, x_dt, xo) : 0.f;
    float S1 = 0.f, S2 = 0.f, H1 = 0.f, H2 = 0.f;
    if (valid) {
#pragma unroll 4
      for (int c = warp; c < C; c += NW) {
        const float d = ldt<DT>(x, x_dt, xo + (int64_t)c * HW) - x0;
        const float h = ldt<DT>(g, g_dt, go + (int64_t)c * HW) * __ldg(lw + c);
        S1 += d;
        S2 = fmaf(d, d, S2);
        H1 += h;
        H2 = fmaf(h, d, H2);
      }
    }
    __syncthreads();   // previous tile's readers of red4 are done
    red4[warp][0][lane] = S1;
    red4[warp][1][lane] = S2;
    red4[warp][2][lane] = H1;
    red4[warp][3][lane] = H2;
    __syncthreads();
    S1 = S2 = H1 = H2 = 0.f;
#pragma unroll
    for (int w = 0; w < NW; ++w) {
      S1 += red4[w][0][lane];
      S2 += red4[w][1][lane];
      H1 += red4[w][2][lane];
      H2 += red4[w][3][lane];
    }
    const float md = S1 * invC;                       // mu - x0
    const float mu = x0 + md;
    const float rstd = 1.0f / sqrtf(fmaxf(S2 * invC - md * md, 0.f) + 1e-5f);
    const float s1 = H1 * invC;                       // mean(h)
    // WithBias: s2 = mean(h*xh) = rstd*(mean(h*(x-x0)) - md*mean(h));  BiasFree: s2 = mean(h*x)
    const float s2 = ln_type == B200IR_LN_WITHBIAS ? rstd * (H2 * invC - md * s1) : (H2 * invC + x0 * s1);
    // pass B: dx, and this tile's contribution to dgamma / dbeta
    const int64_t dxo = (int64_t)b * dx_bs + p, ro = (int64_t)b * res_bs + p, no = (int64_t)b * xn_bs + p;
#pragma unroll 2
    for (int c = warp; c < C; c += NW) {
      float xv = 0.f, gv = 0.f, rv = 0.f;
      if (valid) {
        xv = ldt<DT>(x, x_dt, xo + (int64_t)c * HW);
        gv = ldt<DT>(g, g_dt, go + (int64_t)c * HW);
        if (res) rv = ldt<DT>(res, res_dt, ro + (int64_t)c * HW);
      }
      const float gam = __ldg(lw + c);
      const float h = gv * gam;
      float d, dgam, dbet = gv, nv;
      if (ln_type == B200IR_LN_WITHBIAS) {
        const float xh = (xv - mu) * rstd;
        d = rstd * (h - s1 - xh * s2);
        dgam = gv * xh;
        nv = xn ? fmaf(xh, gam, __ldg(lb + c)) : 0.f;
      } else {
        d = rstd * h - rstd * rstd * rstd * (xv - mu) * s2;
        dgam = gv * xv * rstd;
        nv = xv * rstd * gam;
      }
      if (valid) stt<DT>(dx, dx_dt, dxo + (int64_t)c * HW, d + rv);
      if (valid && xn) stt<DT>(xn, dx_dt, no + (int64_t)c * HW, nv);
      dgam = warp_sum(dgam);
      dbet = warp_sum(dbet);
      if (lane == 0) {
        acc[c] += dgam;
        acc[C + c] += dbet;
      }
    }
  }
  __syncthreads();
  float* out = parts + (size_t)blockIdx.x * 2 * C;
  for (int i = threadIdx.x; i < 2 * C; i += 32 * NW) out[i] = acc[i];
}

// bf16 fast path of ln_bwd_kernel: a lane owns TWO adjacent pixels (bf16x2 accesses: 128 B per warp instruction, half the
// instructions and half the dgamma/dbeta shuffles per byte), channel pointers advance by 8*HW instead of being
// recomputed.  Needs HW even, even batch strides and 4-byte aligned bases.  Same two-pass math as above.
__global__ void __launch_bounds__(256)
ln_bwd_bf16x2_kernel(const bf16* __restrict__ x, int64_t x_bs, const bf16* __restrict__ g, int64_t g_bs, int ln_type,
                     const float* __restrict__ lw, const float* __restrict__ lb, const bf16* __restrict__ res,
                     int64_t res_bs, bf16* __restrict__ dx, int64_t dx_bs, bf16* __restrict__ xn, int64_t xn_bs,
                     float* __restrict__ parts, int C, int64_t HW, int tps, int64_t ntiles) {
  extern __shared__ float acc[];  // [2][C]
  __shared__ float2 red4[8][4][32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 2 * C; i += 256) acc[i] = 0.f;
  __syncthreads();
  const float invC = 1.0f / (float)C;
  const int64_t cstep = 8 * HW;
  auto ld2 = [](const bf16* p) {
    return __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(p));
  };
  for (int64_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
    const int b = (int)(t / tps);
    const int64_t p = (t % tps) * 64 + 2 * lane;
    const bool valid = p < HW;
    const bf16* xw = x + (int64_t)b * x_bs + p + (int64_t)warp * HW;
    const bf16* gw = g + (int64_t)b * g_bs + p + (int64_t)warp * HW;
    const float2 x0 = valid ? ld2(x + (int64_t)b * x_bs + p) : make_float2(0.f, 0.f);
    float2 S1 = {0.f, 0.f}, S2 = {0.f, 0.f}, H1 = {0.f, 0.f}, H2 = {0.f, 0.f};
    if (valid) {
      const bf16 *xp = xw, *gp = gw;
#pragma unroll 4
      for (int c = warp; c < C; c += 8, xp += cstep, gp += cstep) {
        const float2 xv = ld2(xp), gv = ld2(gp);
        const float gam = __ldg(lw + c);
        const float d0 = xv.x - x0.x, d1 = xv.y - x0.y, h0 = gv.x * gam, h1 = gv.y * gam;
        S1.x += d0; S1.y += d1;
        S2.x = fmaf(d0, d0, S2.x); S2.y = fmaf(d1, d1, S2.y);
        H1.x += h0; H1.y += h1;
        H2.x = fmaf(h0, d0, H2.x); H2.y = fmaf(h1, d1, H2.y);
      }
    }
    __syncthreads();
    red4[warp][0][lane] = S1;
    red4[warp][1][lane] = S2;
    red4[warp][2][lane] = H1;
    red4[warp][3][lane] = H2;
    __syncthreads();
    S1 = S2 = H1 = H2 = make_float2(0.f, 0.f);
#pragma unroll
    for (int w = 0; w < 8; ++w) {
      const float2 a = red4[w][0][lane], bq = red4[w][1][lane], cq = red4[w][2][lane], dq = red4[w][3][lane];
      S1.x += a.x; S1.y += a.y; S2.x += bq.x; S2.y += bq.y;
      H1.x += cq.x; H1.y += cq.y; H2.x += dq.x; H2.y += dq.y;
    }
    float mu[2], rstd[2], s1[2], s2[2];
    {
      const float S1v[2] = {S1.x, S1.y}, S2v[2] = {S2.x, S2.y}, H1v[2] = {H1.x, H1.y}, H2v[2] = {H2.x, H2.y};
      const float x0v[2] = {x0.x, x0.y};
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const float md = S1v[k] * invC;
        mu[k] = x0v[k] + md;
        rstd[k] = 1.0f / sqrtf(fmaxf(S2v[k] * invC - md * md, 0.f) + 1e-5f);
        s1[k] = H1v[k] * invC;
        s2[k] = ln_type == B200IR_LN_WITHBIAS ? rstd[k] * (H2v[k] * invC - md * s1[k]) : (H2v[k] * invC + x0v[k] * s1[k]);
      }
    }
    const bf16 *xp = xw, *gp = gw;
    const bf16* rp = res ? res + (int64_t)b * res_bs + p + (int64_t)warp * HW : nullptr;
    bf16* dp = dx + (int64_t)b * dx_bs + p + (int64_t)warp * HW;
    bf16* np = xn ? xn + (int64_t)b * xn_bs + p + (int64_t)warp * HW : nullptr;
#pragma unroll 2
    for (int c = warp; c < C; c += 8, xp += cstep, gp += cstep, dp += cstep) {
      float2 xv = {0.f, 0.f}, gv = {0.f, 0.f}, rv = {0.f, 0.f};
      if (valid) {
        xv = ld2(xp);
        gv = ld2(gp);
        if (rp) rv = ld2(rp);
      }
      if (rp) rp += cstep;
      const float gam = __ldg(lw + c);
      const float xs[2] = {xv.x, xv.y}, gs[2] = {gv.x, gv.y};
      float d[2], nv[2], dgam = 0.f;
      const float bet = (np && ln_type == B200IR_LN_WITHBIAS) ? __ldg(lb + c) : 0.f;
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const float h = gs[k] * gam;
        if (ln_type == B200IR_LN_WITHBIAS) {
          const float xh = (xs[k] - mu[k]) * rstd[k];
          d[k] = rstd[k] * (h - s1[k] - xh * s2[k]);
          dgam = fmaf(gs[k], xh, dgam);
          nv[k] = fmaf(xh, gam, bet);
        } else {
          d[k] = rstd[k] * h - rstd[k] * rstd[k] * rstd[k] * (xs[k] - mu[k]) * s2[k];
          dgam = fmaf(gs[k] * xs[k], rstd[k], dgam);
          nv[k] = xs[k] * rstd[k] * gam;
        }
      }
      if (valid) *reinterpret_cast<__nv_bfloat162*>(dp) = __floats2bfloat162_rn(d[0] + rv.x, d[1] + rv.y);
      if (np) {
        if (valid) *reinterpret_cast<__nv_bfloat162*>(np) = __floats2bfloat162_rn(nv[0], nv[1]);
        np += cstep;
      }
      dgam = warp_sum(dgam);
      const float dbet = warp_sum(gv.x + gv.y);
      if (lane == 0) {
        acc[c] += dgam;
        acc[C + c] += dbet;
      }
    }
  }
  __syncthreads();
  float* out = parts + (size_t)blockIdx.x * 2 * C;
  for (int i = threadIdx.x; i < 2 * C; i += 256) out[i] = acc[i];
}

// Register-resident form of the kernel above for C = 8 * CPW, CPW <= 12 (C = 48 / 96: 24 of the 44 blocks of the default
// net; CPW = 24 needs 247 registers).  Measured in a CUDA graph at B8 C48 128x128 the two-pass kernel took 32 us for 38 MB (0.2 of the HBM
// roof): both passes load, and every channel of every tile pays two 5-step warp reductions for dgamma / dbeta.  Here a
// warp's CPW channels of the tile stay in registers between the passes (one trip over x and g, all loads of a tile in
// flight together), and dgamma / dbeta are accumulated PER LANE over all tiles of the CTA and folded across lanes once
// at the end (fixed order: tile loop, then the shuffle tree) -- 2 * CPW reductions per CTA instead of per tile.
template <int CPW>
__global__ void __launch_bounds__(256)
ln_bwd_bf16_reg_kernel(const bf16* __restrict__ x, int64_t x_bs, const bf16* __restrict__ g, int64_t g_bs, int ln_type,
                       const float* __restrict__ lw, const float* __restrict__ lb, const bf16* __restrict__ res,
                       int64_t res_bs, bf16* __restrict__ dx, int64_t dx_bs, bf16* __restrict__ xn, int64_t xn_bs,
                       float* __restrict__ parts, int64_t HW, int tps, int64_t ntiles) {
  constexpr int C = 8 * CPW;
  __shared__ float2 red4[8][4][32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const float invC = 1.0f / (float)C;
  const int64_t cstep = 8 * HW;
  float gam[CPW], bet[CPW], dgam[CPW], dbet[CPW];
#pragma unroll
  for (int j = 0; j < CPW; ++j) {
    gam[j] = __ldg(lw + warp + 8 * j);
    bet[j] = (xn && ln_type == B200IR_LN_WITHBIAS) ? __ldg(lb + warp + 8 * j) : 0.f;
    dgam[j] = dbet[j] = 0.f;
  }
  for (int64_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
    const int b = (int)(t / tps);
    const int64_t p = (t % tps) * 64 + 2 * lane;
    const bool valid = p < HW;
    const int64_t off = p + (int64_t)warp * HW;
    uint32_t xr[CPW], gr[CPW];
    uint32_t x0r = 0u;
    if (valid) {
      const bf16* xp = x + (int64_t)b * x_bs + off;
      const bf16* gp = g + (int64_t)b * g_bs + off;
      x0r = *reinterpret_cast<const uint32_t*>(x + (int64_t)b * x_bs + p);
#pragma unroll
      for (int j = 0; j < CPW; ++j) {
        xr[j] = *reinterpret_cast<const uint32_t*>(xp + j * cstep);
        gr[j] = *reinterpret_cast<const uint32_t*>(gp + j * cstep);
      }
    } else {
#pragma unroll
      for (int j = 0; j < CPW; ++j) xr[j] = gr[j] = 0u;
    }
    const float2 x0 = make_float2(__uint_as_float(x0r << 16), __uint_as_float(x0r & 0xffff0000u));
    float2 S1 = {0.f, 0.f}, S2 = {0.f, 0.f}, H1 = {0.f, 0.f}, H2 = {0.f, 0.f};
#pragma unroll
    for (int j = 0; j < CPW; ++j) {
      const float d0 = __uint_as_float(xr[j] << 16) - x0.x, d1 = __uint_as_float(xr[j] & 0xffff0000u) - x0.y;
      const float h0 = __uint_as_float(gr[j] << 16) * gam[j], h1 = __uint_as_float(gr[j] & 0xffff0000u) * gam[j];
      S1.x += d0; S1.y += d1;
      S2.x = fmaf(d0, d0, S2.x); S2.y = fmaf(d1, d1, S2.y);
      H1.x += h0; H1.y += h1;
      H2.x = fmaf(h0, d0, H2.x); H2.y = fmaf(h1, d1, H2.y);
    }
    __syncthreads();          // the previous tile's readers of red4 are done
    red4[warp][0][lane] = S1;
    red4[warp][1][lane] = S2;
    red4[warp][2][lane] = H1;
    red4[warp][3][lane] = H2;
    __syncthreads();
    S1 = S2 = H1 = H2 = make_float2(0.f, 0.f);
#pragma unroll
    for (int w = 0; w < 8; ++w) {
      const float2 a = red4[w][0][lane], bq = red4[w][1][lane], cq = red4[w][2][lane], dq = red4[w][3][lane];
      S1.x += a.x; S1.y += a.y; S2.x += bq.x; S2.y += bq.y;
      H1.x += cq.x; H1.y += cq.y; H2.x += dq.x; H2.y += dq.y;
    }
    float mu[2], rstd[2], s1[2], s2[2];
    {
      const float S1v[2] = {S1.x, S1.y}, S2v[2] = {S2.x, S2.y}, H1v[2] = {H1.x, H1.y}, H2v[2] = {H2.x, H2.y};
      const float x0v[2] = {x0.x, x0.y};
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const float md = S1v[k] * invC;
        mu[k] = x0v[k] + md;
        rstd[k] = 1.0f / sqrtf(fmaxf(S2v[k] * invC - md * md, 0.f) + 1e-5f);
        s1[k] = H1v[k] * invC;
        s2[k] = ln_type == B200IR_LN_WITHBIAS ? rstd[k] * (H2v[k] * invC - md * s1[k]) : (H2v[k] * invC + x0v[k] * s1[k]);
      }
    }
    if (!valid) continue;     // warp-uniform per lane only in the ragged last tile; no barrier below this point in the body
    const bf16* rp = res ? res + (int64_t)b * res_bs + off : nullptr;
    bf16* dp = dx + (int64_t)b * dx_bs + off;
    bf16* np = xn ? xn + (int64_t)b * xn_bs + off : nullptr;
#pragma unroll
    for (int j = 0; j < CPW; ++j) {
      const float xs[2] = {__uint_as_float(xr[j] << 16), __uint_as_float(xr[j] & 0xffff0000u)};
      const float gs[2] = {__uint_as_float(gr[j] << 16), __uint_as_float(gr[j] & 0xffff0000u)};
      float2 rv = {0.f, 0.f};
      if (rp) rv = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(rp + j * cstep));
      float d[2], nv[2];
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const float h = gs[k] * gam[j];
        if (ln_type == B200IR_LN_WITHBIAS) {
          const float xh = (xs[k] - mu[k]) * rstd[k];
          d[k] = rstd[k] * (h - s1[k] - xh * s2[k]);
          dgam[j] = fmaf(gs[k], xh, dgam[j]);
          nv[k] = fmaf(xh, gam[j], bet[j]);
        } else {
          d[k] = rstd[k] * h - rstd[k] * rstd[k] * rstd[k] * (xs[k] - mu[k]) * s2[k];
          dgam[j] = fmaf(gs[k] * xs[k], rstd[k], dgam[j]);
          nv[k] = xs[k] * rstd[k] * gam[j];
        }
      }
      dbet[j] += gs[0] + gs[1];
      *reinterpret_cast<__nv_bfloat162*>(dp + j * cstep) = __floats2bfloat162_rn(d[0] + rv.x, d[1] + rv.y);
      if (np) *reinterpret_cast<__nv_bfloat162*>(np + j * cstep) = __floats2bfloat162_rn(nv[0], nv[1]);
    }
  }
  float* out = parts + (size_t)blockIdx.x * 2 * C;
#pragma unroll
  for (int j = 0; j < CPW; ++j) {
    const float a = warp_sum(dgam[j]), bq = warp_sum(dbet[j]);
    if (lane == 0) {
      out[warp + 8 * j] = a;
      out[C + warp + 8 * j] = bq;
    }
  }
}

// out[k] = sum_{i<n} in[i*stride + k].  CTA = 32 outputs x SL slices of the partial index; each slice is summed
// sequentially, the SL slice sums are combined in fixed order (deterministic).
template <int SL>
__global__ void __launch_bounds__(32 * SL)
reduce_partials_kernel(const float* __restrict__ in, float* __restrict__ out, int n, int64_t K, int64_t stride) {
  __shared__ float red[SL][32];
  const int lane = threadIdx.x & 31, sl = threadIdx.x >> 5;
  for (int64_t k0 = (int64_t)blockIdx.x * 32; k0 < K; k0 += (int64_t)gridDim.x * 32) {
    const int64_t k = k0 + lane;
    float s = 0.f;
    if (k < K)
      for (int i = sl; i < n; i += SL) s += in[(size_t)i * stride + k];
    __syncthreads();
    red[sl][lane] = s;
    __syncthreads();
    if (sl == 0 && k < K) {
      float t = 0.f;
#pragma unroll
      for (int w = 0; w < SL; ++w) t += red[w][lane];
      out[k] = t;
    }
  }
}

// The same fold for the depthwise-conv partial sums parts[i][c][nv] (nv - 1 taps + the bias sum per channel), written as two
// dense tensors: dw[c][nv - 1] and db[c] -- the parameter gradients themselves, not strided views into a [C][nv] buffer.
__global__ void __launch_bounds__(256)
reduce_partials_taps_kernel(const float* __restrict__ in, float* __restrict__ dw, float* __restrict__ db, int n, int C, int nv) {
  __shared__ float red[8][32];
  const int lane = threadIdx.x & 31, sl = threadIdx.x >> 5;
  const int64_t K = (int64_t)C * nv;
  for (int64_t k0 = (int64_t)blockIdx.x * 32; k0 < K; k0 += (int64_t)gridDim.x * 32) {
    const int64_t k = k0 + lane;
    float s = 0.f;
    if (k < K)
      for (int i = sl; i < n; i += 8) s += in[(size_t)i * K + k];
    __syncthreads();
    red[sl][lane] = s;
    __syncthreads();
    if (sl == 0 && k < K) {
      float t = 0.f;
#pragma unroll
      for (int w = 0; w < 8; ++w) t += red[w][lane];
      const int c = (int)(k / nv), q = (int)(k - (int64_t)c * nv);
      if (q < nv - 1) dw[(size_t)c * (nv - 1) + q] = t;
      else if (db) db[c] = t;
    }
  }
}

// =====================================================================================================
// Weight gradient of a 1x1 conv (taps=1) or a dense 3x3 conv (taps=9): reduction over pixels
//   dW[g][o][c*taps+t] = sum_{b in g} sum_p dy[b,o,p] * x[b,c,p+shift(t)],   db[g][o] = sum dy[b,o,p]
// g = sample (per_sample, MDTA dM_b = dout_b . v_b^T) or the whole batch.  Split-K over 64-pixel chunks.
// CUDA-core fp32 FMA version (exact path); grid (n-tiles, m-tiles, G*nsplit).
// =====================================================================================================
constexpr int WM = 64, WN = 64, WK = 64;   // WK = pixels per split-K chunk (same chunking as wgrad_tc.cu)

__global__ void __launch_bounds__(256)
wgrad_simt_kernel(const void* __restrict__ A, int64_t a_bs, int a_dt, const void* __restrict__ X, int64_t x_bs, int x_dt,
                  float* __restrict__ pw, float* __restrict__ pb, int M, int Cin, int taps, int H, int W,
                  int nsplit, int chunks_per_sample, int chunks_per_split, int chunks_total, int per_sample) {
  __shared__ __align__(16) float As[WK][WM + 4];
  __shared__ __align__(16) float Bs[WK][WN + 4];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int n0 = blockIdx.x * WN, o0 = blockIdx.y * WM;
  const int g = blockIdx.z / nsplit, split = blockIdx.z % nsplit;
  const int N = Cin * taps;
  const int64_t HW = (int64_t)H * W;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  float bsum = 0.f;

  const int cbeg = split * chunks_per_split;
  const int cend = min(chunks_total, cbeg + chunks_per_split);
  for (int ch = cbeg; ch < cend; ++ch) {
    const int b = per_sample ? g : ch / chunks_per_sample;
    const int64_t p0 = (int64_t)(ch % chunks_per_sample) * WK;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int e = tid + i * 256;
      const int r = e >> 6, cpx = e & 63;
      const int64_t p = p0 + cpx;
      float av = 0.f, xv = 0.f;
      if (p < HW) {
        if (o0 + r < M) av = ld_any(A, a_dt, (int64_t)b * a_bs + (int64_t)(o0 + r) * HW + p);
        const int n = n0 + r;
        if (n < N) {
          if (taps == 1) {
            xv = ld_any(X, x_dt, (int64_t)b * x_bs + (int64_t)n * HW + p);
          } else {
            const int c = n / 9, t = n - c * 9;
            const int yy = (int)(p / W) + t / 3 - 1, xx = (int)(p % W) + t % 3 - 1;
            if (yy >= 0 && yy < H && xx >= 0 && xx < W)
              xv = ld_any(X, x_dt, (int64_t)b * x_bs + (int64_t)c * HW + (int64_t)yy * W + xx);
          }
        }
      }
      As[cpx][r] = av;
      Bs[cpx][r] = xv;
    }
    __syncthreads();
#pragma unroll 8
    for (int k = 0; k < WK; ++k) {
      const float4 a = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
      const float4 bb = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w};
      const float bv[4] = {bb.x, bb.y, bb.z, bb.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    if (pb && blockIdx.x == 0 && tid < WM) {
#pragma unroll 8
      for (int k = 0; k < WK; ++k) bsum += As[k][tid];
    }
    __syncthreads();
  }
  const size_t slot = (size_t)split * (gridDim.z / nsplit) + g;  // partials laid out [split][g][M][N]
  float* out = pw + slot * M * N;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int o = o0 + ty * 4 + i;
    if (o >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n < N) out[(size_t)o * N + n] = acc[i][j];
    }
  }
  if (pb && blockIdx.x == 0 && tid < WM && o0 + tid < M) pb[slot * M + o0 + tid] = bsum;
}

// =====================================================================================================
// MDTA small backward, one CTA per (head, sample).  Inputs: Gram partials (forward), attn, temperature, W_out and
// dM_b = dout_b . v_b^T  [B,C,C].  With G^ = G/(a b^T) (a_i = max(|q_i|,eps), b_j likewise), S = temp*G^,
// attn = softmax(S), M_b = W_out . blockdiag(attn):
//   dattn = W_out[:,h]^T . dM[:,h],  dW_out[:,h] (per sample) = dM[:,h] . attn^T
//   dS = attn o (dattn - rowsum(dattn o attn)),  dtemp = sum dS o G^,  dG^ = temp*dS
//   P = dG^/(a b^T),  r_i = sum_j dG^_ij G^_ij / a_i^2,  s_j = sum_i dG^_ij G^_ij / b_j^2
//   dq = P.k - diag(r).q,  dk = P^T.q - diag(s).k      -> mqk[b] (2C x 2C) so that d[q;k] = mqk[b] . [q;k]
//   mt[b] = M_b^T                                      -> dv = mt[b] . dout
// =====================================================================================================
// Kernel 1 (chain): grid (heads, B), 1024 threads: dattn, softmax / temperature / normalise backward -> mqk, dtemp.
// Kernel 2 (fold):  grid (ceil(C/32), heads, B), 256 threads: per-sample dW_out and mt for 32 output channels.
constexpr int MB_THREADS = 1024;

__global__ void __launch_bounds__(MB_THREADS)
mdta_bwd_chain_kernel(const float* __restrict__ partials, int nslab, const float* __restrict__ attn,
                      const float* __restrict__ temperature, const float* __restrict__ w_out,
                      const float* __restrict__ dM, float* __restrict__ mqk, float* __restrict__ dtemp_part, int C,
                      int d, int temp_per_sample, int CO, int S) {
  extern __shared__ float sm[];
  const int D1 = d + 1;
  float* at = sm;                        // attn  [d][d+1]
  float* gh = at + (size_t)d * D1;       // G^    [d][d+1]
  float* da = gh + (size_t)d * D1;       // dattn -> dG^ [d][d+1]
  float* na = da + (size_t)d * D1;       // a [d]
  float* nb = na + d;                    // b [d]
  float* rs = nb + d;                    // rowsum / r [d]
  float* cs = rs + d;                    // s [d]
  __shared__ float red[MB_THREADS / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int NW = MB_THREADS / 32;
  const int h = blockIdx.x, b = blockIdx.y, heads = gridDim.x;
  const size_t per = (size_t)d * d + 2 * d;
  const float* base = partials + ((size_t)b * heads + h) * nslab * per;
  const float* a_g = attn + ((size_t)b * heads + h) * d * d;
  const float* dMb = dM + (size_t)b * C * C;
  const float temp = temperature[temp_per_sample ? (size_t)b * heads + h : h];   // [B][heads]: dynamic temperature

  // sum over the Gram slabs: eight loads in flight, added in slab order (the order of the plain loop)
  auto slab_sum = [&](size_t off) {
    float s = 0.f;
    int sl = 0;
    for (; sl + 8 <= nslab; sl += 8) {
      float v[8];
#pragma unroll
      for (int q = 0; q < 8; ++q) v[q] = base[(size_t)(sl + q) * per + off];
#pragma unroll
      for (int q = 0; q < 8; ++q) s += v[q];
    }
    for (; sl < nslab; ++sl) s += base[(size_t)sl * per + off];
    return s;
  };
  for (int j = tid; j < 2 * d; j += MB_THREADS)
    na[j] = fmaxf(sqrtf(slab_sum((size_t)d * d + j)), 1e-12f);  // na[d..2d) aliases nb
  __syncthreads();
  for (int e = tid; e < d * d; e += MB_THREADS) {
    const int i = e / d, j = e % d;
    const float s = slab_sum((size_t)e);
    gh[i * D1 + j] = s / (na[i] * nb[j]);
    at[i * D1 + j] = a_g[e];
    da[i * D1 + j] = 0.f;
  }
  // stage this head's column blocks W_out[:, hd:hd+d] and dM_b[:, hd:hd+d] in shared memory TRANSPOSED ([i][o], row stride S
  // with S % 32 == 4: a lane's float4 along o is conflict-free across the j of a warp), CO rows of o at a time (CO = C
  // unless the two blocks do not fit: AdaIR's FreModule runs C = 384 with 4 heads).  An element's fma chain over o runs in
  // the same order whatever CO is, through da[] between chunks, so the result does not depend on CO; the zero padding of
  // o up to a multiple of 4 adds fma(0, 0, acc) = acc.
  float* wsm = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(cs + d) + 15) & ~(uintptr_t)15);   // [d][S], 16-B aligned
  float* dsm = wsm + (size_t)d * S;      // [d][S]
  for (int o0 = 0; o0 < C; o0 += CO) {
    const int co = min(CO, C - o0), co4 = (co + 3) & ~3;
    __syncthreads();
    for (int e = tid; e < co4 * d; e += MB_THREADS) {
      const int o = e / d, i = e - o * d;
      const bool in = o < co;
      wsm[i * S + o] = in ? w_out[(size_t)(o0 + o) * C + h * d + i] : 0.f;
      dsm[i * S + o] = in ? dMb[(size_t)(o0 + o) * C + h * d + i] : 0.f;
    }
    __syncthreads();
    for (int e = tid; e < d * d; e += MB_THREADS) {
      const int i = e / d, j = e % d;
      // dattn[i][j] = sum_o W[o][hd+i] * dM[o][hd+j]
      float acc = da[i * D1 + j];
      const float4* wp = reinterpret_cast<const float4*>(wsm + i * S);
      const float4* dp = reinterpret_cast<const float4*>(dsm + j * S);
#pragma unroll 4
      for (int o = 0; o < co4 / 4; ++o) {
        const float4 a = wp[o], b = dp[o];
        acc = fmaf(a.x, b.x, acc);
        acc = fmaf(a.y, b.y, acc);
        acc = fmaf(a.z, b.z, acc);
        acc = fmaf(a.w, b.w, acc);
      }
      da[i * D1 + j] = acc;
    }
  }
  __syncthreads();
  for (int i = warp; i < d; i += NW) {  // row sums of dattn o attn
    float s = 0.f;
    for (int j = lane; j < d; j += 32) s += da[i * D1 + j] * at[i * D1 + j];
    s = warp_sum(s);
    if (lane == 0) rs[i] = s;
  }
  __syncthreads();
  float tsum = 0.f;
  for (int e = tid; e < d * d; e += MB_THREADS) {
    const int i = e / d, j = e % d;
    const float dS = at[i * D1 + j] * (da[i * D1 + j] - rs[i]);
    tsum += dS * gh[i * D1 + j];
    da[i * D1 + j] = temp * dS;  // dG^
  }
  tsum = warp_sum(tsum);
  if (lane == 0) red[warp] = tsum;
  __syncthreads();
  if (tid == 0) {
    float s = 0.f;
    for (int w = 0; w < NW; ++w) s += red[w];
    dtemp_part[(size_t)b * heads + h] = s;
  }
  for (int i = warp; i < d; i += NW) {  // r_i, s_j
    float s = 0.f, s2 = 0.f;
    for (int j = lane; j < d; j += 32) {
      s += da[i * D1 + j] * gh[i * D1 + j];
      s2 += da[j * D1 + i] * gh[j * D1 + i];
    }
    s = warp_sum(s);
    s2 = warp_sum(s2);
    if (lane == 0) {
      rs[i] = s / (na[i] * na[i]);
      cs[i] = s2 / (nb[i] * nb[i]);
    }
  }
  __syncthreads();
  // mqk rows of this head: [hd+i] (dq) and [C+hd+i] (dk), full 2C columns
  const int C2 = 2 * C;
  float* mq = mqk + (size_t)b * C2 * C2;
  for (int e = tid; e < d * C2; e += MB_THREADS) {
    const int i = e / C2, col = e % C2;
    float vq = 0.f, vk = 0.f;
    if (col < C) {
      const int jj = col - h * d;
      if (jj == i) vq = -rs[i];                                              // dq_i <- -r_i q_i
      if (jj >= 0 && jj < d) vk = da[jj * D1 + i] / (na[jj] * nb[i]);        // dk_i <- sum_jj P[jj][i] q_jj
    } else {
      const int jj = col - C - h * d;
      if (jj >= 0 && jj < d) vq = da[i * D1 + jj] / (na[i] * nb[jj]);        // dq_i <- sum_jj P[i][jj] k_jj
      if (jj == i) vk = -cs[i];                                              // dk_i <- -s_i k_i
    }
    mq[(size_t)(h * d + i) * C2 + col] = vq;
    mq[(size_t)(C + h * d + i) * C2 + col] = vk;
  }
}

// per-sample dW_out[o][hd+i] = sum_j dM[o][hd+j]*attn[i][j];  mt[b][hd+j][o] = sum_i W[o][hd+i]*attn[i][j]
__global__ void __launch_bounds__(256)
mdta_bwd_fold_kernel(const float* __restrict__ attn, const float* __restrict__ w_out, const float* __restrict__ dM,
                     float* __restrict__ mt, float* __restrict__ dw_out_part, int C, int d) {
  extern __shared__ float sm[];
  const int D1 = d + 1;
  float* at = sm;                          // attn [d][d+1]
  float* ws = at + (size_t)d * D1;         // W_out[o0..o0+32)[hd..hd+d)   [32][d+1]
  float* ds = ws + (size_t)32 * D1;        // dM   [o0..o0+32)[hd..hd+d)   [32][d+1]
  const int tid = threadIdx.x;
  const int o0 = blockIdx.x * 32, h = blockIdx.y, b = blockIdx.z, heads = gridDim.y;
  const float* a_g = attn + ((size_t)b * heads + h) * d * d;
  const float* dMb = dM + (size_t)b * C * C;
  for (int e = tid; e < d * d; e += 256) at[(e / d) * D1 + e % d] = a_g[e];
  for (int e = tid; e < 32 * d; e += 256) {
    const int ol = e / d, i = e % d, o = o0 + ol;
    ws[ol * D1 + i] = o < C ? w_out[(size_t)o * C + h * d + i] : 0.f;
    ds[ol * D1 + i] = o < C ? dMb[(size_t)o * C + h * d + i] : 0.f;
  }
  __syncthreads();
  for (int e = tid; e < 32 * d; e += 256) {   // dW_out: threads along i (contiguous in the output row)
    const int ol = e / d, i = e % d, o = o0 + ol;
    if (o >= C) continue;
    float s = 0.f;
    for (int j = 0; j < d; ++j) s = fmaf(ds[ol * D1 + j], at[i * D1 + j], s);
    dw_out_part[((size_t)b * C + o) * C + h * d + i] = s;
  }
  for (int e = tid; e < 32 * d; e += 256) {   // mt: threads along o (contiguous in the output row)
    const int j = e / 32, ol = e % 32, o = o0 + ol;
    if (o >= C) continue;
    float s = 0.f;
    for (int i = 0; i < d; ++i) s = fmaf(ws[ol * D1 + i], at[i * D1 + j], s);
    mt[((size_t)b * C + h * d + j) * C + o] = s;
  }
}

// =====================================================================================================
// im2col for the dense 3x3 weight gradient on tensor cores: y[b][c*9+t][p] = x[b][c][p + shift(t)] (zero padded),
// bf16 output, so that dW[o][c*9+t] = sum_p dy[o,p]*y[c*9+t,p] is a plain K-major GEMM (wgrad_tc.cu).
// grid (ceil(HW/256), Cin, B)
// =====================================================================================================
__global__ void __launch_bounds__(256)
im2col3x3_kernel(const void* __restrict__ x, int64_t x_bs, int x_dt, bf16* __restrict__ y, int Cin, int H, int W) {
  const int c = blockIdx.y, b = blockIdx.z;
  const int64_t HW = (int64_t)H * W;
  const int64_t p = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (p >= HW) return;
  const int yy = (int)(p / W), xx = (int)(p % W);
  const int64_t xo = (int64_t)b * x_bs + (int64_t)c * HW;
  bf16* yo = y + ((int64_t)b * Cin + c) * 9 * HW + p;
#pragma unroll
  for (int t = 0; t < 9; ++t) {
    const int y2 = yy + t / 3 - 1, x2 = xx + t % 3 - 1;
    float v = 0.f;
    if (y2 >= 0 && y2 < H && x2 >= 0 && x2 < W) v = ld_any(x, x_dt, xo + (int64_t)y2 * W + x2);
    yo[(int64_t)t * HW] = __float2bfloat16_rn(v);
  }
}

// =====================================================================================================
// PixelShuffle(2) / PixelUnshuffle(2) as a copy (gradients of Upsample / Downsample, Restormer.py:171-189).
// x [B,C,H,W].  SHUFFLE2: y [B,C/4,2H,2W], y[c,2h+i,2w+j] = x[4c+2i+j,h,w].  UNSHUFFLE2: the inverse,
// y [B,4C,H/2,W/2].  One thread per input element.
// =====================================================================================================
__global__ void pixel_shuffle_kernel(const void* __restrict__ x, int64_t x_bs, void* __restrict__ y, int64_t y_bs,
                                     int dt, int C, int H, int W, int mode) {
  const int b = blockIdx.y;
  const int64_t HW = (int64_t)H * W, n = (int64_t)C * HW;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(e / HW);
    const int64_t p = e % HW;
    const int hh = (int)(p / W), ww = (int)(p % W);
    int64_t off;
    if (mode == B200IR_SHUFFLE2) {
      const int oc = c >> 2, i = (c >> 1) & 1, j = c & 1;
      off = (int64_t)oc * (HW * 4) + (int64_t)(2 * hh + i) * (2 * W) + (2 * ww + j);
    } else {
      const int oc = c * 4 + (hh & 1) * 2 + (ww & 1);
      off = (int64_t)oc * (HW / 4) + (int64_t)(hh >> 1) * (W / 2) + (ww >> 1);
    }
    st_any(y, dt, (int64_t)b * y_bs + off, ld_any(x, dt, (int64_t)b * x_bs + e));
  }
}

// =====================================================================================================
// L1 loss (nn.L1Loss, MoCE-IR-main/src/train.py:51,65): parts[blk] = sum |pred-target| ; loss = sum(parts)/n.
// Backward: grad = sign(pred-target) * gscale[0] / n.
// =====================================================================================================
__global__ void __launch_bounds__(256)
l1_loss_fwd_kernel(const void* __restrict__ pred, int p_dt, const void* __restrict__ target, int t_dt, int64_t n,
                   float* __restrict__ parts) {
  float s = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256)
    s += fabsf(ld_any(pred, p_dt, i) - ld_any(target, t_dt, i));
  __shared__ float red[8];
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < 8; ++w) t += red[w];
    parts[blockIdx.x] = t;
  }
}
__global__ void l1_loss_finish_kernel(const float* __restrict__ parts, int nparts, float inv_n, float* __restrict__ loss) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    float s = 0.f;
    for (int i = 0; i < nparts; ++i) s += parts[i];
    loss[0] = s * inv_n;
  }
}
__global__ void l1_loss_bwd_kernel(const void* __restrict__ pred, int p_dt, const void* __restrict__ target, int t_dt,
                                   int64_t n, const float* __restrict__ gscale, float inv_n, void* __restrict__ grad,
                                   int g_dt) {
  const float gs = gscale[0] * inv_n;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float d = ld_any(pred, p_dt, i) - ld_any(target, t_dt, i);
    st_any(grad, g_dt, i, d > 0.f ? gs : (d < 0.f ? -gs : 0.f));
  }
}

// =====================================================================================================
// AdamW (torch.optim.AdamW semantics, MoCE-IR-main/src/train.py:83: decoupled weight decay, bias-corrected
// moments, eps added after the sqrt / sqrt(bias_correction2)); one launch per flat fp32 buffer.
// =====================================================================================================
__global__ void adamw_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                             float* __restrict__ v, int64_t n, float lr, float beta1, float beta2, float eps, float wd,
                             float bc1, float bc2_sqrt, float gscale, const int* __restrict__ step_dev,
                             const float* __restrict__ hyper_dev) {
  if (hyper_dev) {  // learning rate / weight decay live on the device: a scheduler can change them under a captured graph
    lr = hyper_dev[0];
    wd = hyper_dev[1];
  }
  if (step_dev) {  // CUDA-graph replay: the step count lives on the device (b200ir_adamw_step, step_dev != NULL)
    const double st = (double)step_dev[0];
    bc1 = (float)(1.0 - pow((double)beta1, st));
    bc2_sqrt = (float)sqrt(1.0 - pow((double)beta2, st));
  }
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float gi = g[i] * gscale;
    float pi = p[i];
    pi *= 1.0f - lr * wd;
    const float mi = m[i] + (1.0f - beta1) * (gi - m[i]);       // lerp, as torch's _single_tensor_adamw
    const float vi = beta2 * v[i] + (1.0f - beta2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    const float denom = sqrtf(vi) / bc2_sqrt + eps;
    p[i] = pi - (lr / bc1) * (mi / denom);
  }
}
__global__ void counter_inc_kernel(int* c) {
  if (threadIdx.x == 0 && blockIdx.x == 0) c[0] += 1;
}

static inline int grid_1d(int64_t n, int threads, int cap) {
  int64_t g = (n + threads - 1) / threads;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (int)g;
}

}  // namespace b200ir

using namespace b200ir;

extern "C" int b200ir_ln_fwd(const void* x, int64_t x_bstride, int x_dtype, int ln_type, const float* ln_w,
                             const float* ln_b, void* y, int64_t y_bstride, int y_dtype, int B, int C, int64_t HW,
                             b200ir_stream_t stream) {
  B200IR_CHECK_ARG(x && y && ln_w && B > 0 && C > 0 && HW > 0, "ln_fwd: bad args");
  B200IR_CHECK_ARG(ln_type == B200IR_LN_WITHBIAS || ln_type == B200IR_LN_BIASFREE, "ln_fwd: bad ln_type %d", ln_type);
  B200IR_CHECK_ARG(ln_type != B200IR_LN_WITHBIAS || ln_b, "ln_fwd: ln_b missing");
  B200IR_CHECK_ARG(dtype_ok(x_dtype) && dtype_ok(y_dtype), "ln_fwd: bad dtype");
  const int tps = (int)((HW + LN_PX - 1) / LN_PX);
  const int64_t ntiles = (int64_t)B * tps;
#define B200IR_LN_FWD(DT)                                                                                        \
  ln_fwd_kernel<DT><<<grid_1d(ntiles, 1, num_sms() * 16), 256, 0, (cudaStream_t)stream>>>(x, x_bstride, x_dtype, ln_type, ln_w, \
                                                                                   ln_b, y, y_bstride, y_dtype, C, HW,  \
                                                                                   tps, ntiles)
  if (x_dtype == y_dtype && x_dtype == B200IR_BF16) B200IR_LN_FWD(B200IR_BF16);
  else if (x_dtype == y_dtype) B200IR_LN_FWD(B200IR_F32);
  else B200IR_LN_FWD(-1);
#undef B200IR_LN_FWD
  return check_launch("ln_fwd");
}

extern "C" int b200ir_ln_bwd_nparts(int B, int64_t HW) {
  if (B <= 0 || HW <= 0) return B200IR_E_BADARG;
  const int64_t ntiles = (int64_t)B * ((HW + LN_PX - 1) / LN_PX);
  return grid_1d(ntiles, 1, num_sms() * 4);
}

extern "C" int b200ir_ln_bwd(const void* x, int64_t x_bstride, int x_dtype, const void* g, int64_t g_bstride,
                             int g_dtype, int ln_type, const float* ln_w, const float* ln_b, const void* res,
                             int64_t res_bstride, int res_dtype, void* dx, int64_t dx_bstride, int dx_dtype, void* xn,
                             int64_t xn_bstride, float* parts, int nparts, int B, int C, int64_t HW,
                             b200ir_stream_t stream) {
  B200IR_CHECK_ARG(!xn || ln_type != B200IR_LN_WITHBIAS || ln_b, "ln_bwd: xn output needs ln_b");
  B200IR_CHECK_ARG(x && g && dx && ln_w && parts && B > 0 && C > 0 && HW > 0, "ln_bwd: bad args");
  B200IR_CHECK_ARG(ln_type == B200IR_LN_WITHBIAS || ln_type == B200IR_LN_BIASFREE, "ln_bwd: bad ln_type %d", ln_type);
  B200IR_CHECK_ARG(dtype_ok(x_dtype) && dtype_ok(g_dtype) && dtype_ok(dx_dtype) && (!res || dtype_ok(res_dtype)),
                   "ln_bwd: bad dtype");
  B200IR_CHECK_ARG(nparts == b200ir_ln_bwd_nparts(B, HW), "ln_bwd: nparts=%d, expected %d", nparts,
                   b200ir_ln_bwd_nparts(B, HW));
  B200IR_UNSUPPORTED(C <= 4096, "ln_bwd: C=%d > 4096 unsupported", C);
  const int tps = (int)((HW + LN_PX - 1) / LN_PX);
  const int64_t ntiles = (int64_t)B * tps;
  // few tiles (the deep U-Net levels: 64 / 256 CTAs with one 32-pixel tile each): 32 warps per CTA, so that a warp walks
  // C / 32 channels instead of C / 8 -- these launches are latency chains (C = 384 at 16 x 16 x 8: 47 -> see DESIGN 4b)
  const bool wide = ntiles <= 2 * (int64_t)num_sms() && C >= 128;
#define B200IR_LN_BWD(DT)                                                                                   \
  do {                                                                                                      \
    if (wide)                                                                                               \
      ln_bwd_kernel<DT, 32><<<nparts, 1024, (size_t)2 * C * sizeof(float), (cudaStream_t)stream>>>(                 \
          x, x_bstride, x_dtype, g, g_bstride, g_dtype, ln_type, ln_w, ln_b, res, res_bstride, res_dtype, dx,       \
          dx_bstride, dx_dtype, xn, xn_bstride, parts, C, HW, tps, ntiles);                                         \
    else                                                                                                    \
      ln_bwd_kernel<DT, 8><<<nparts, 256, (size_t)2 * C * sizeof(float), (cudaStream_t)stream>>>(                   \
          x, x_bstride, x_dtype, g, g_bstride, g_dtype, ln_type, ln_w, ln_b, res, res_bstride, res_dtype, dx,       \
          dx_bstride, dx_dtype, xn, xn_bstride, parts, C, HW, tps, ntiles);                                         \
  } while (0)
  const bool same = x_dtype == g_dtype && x_dtype == dx_dtype && (!res || res_dtype == x_dtype);
  auto even = [](const void* q, int64_t bs) { return !q || (((uintptr_t)q & 3) == 0 && (bs & 1) == 0); };
  // (tiny planes keep the 32-pixel tiles: twice as many CTAs to spread the channel loop over)
  if (same && x_dtype == B200IR_BF16 && (HW & 1) == 0 && (int64_t)B * ((HW + 63) / 64) >= 2 * num_sms() && even(x, x_bstride) && even(g, g_bstride) &&
      even(res, res_bstride) && even(dx, dx_bstride) && even(xn, xn_bstride)) {
    const int tps2 = (int)((HW + 63) / 64);
#define B200IR_LN_BWD_REG(CPW_)                                                                                        \
  {                                                                                                                    \
    ln_bwd_bf16_reg_kernel<CPW_><<<nparts, 256, 0, (cudaStream_t)stream>>>(                                            \
        (const bf16*)x, x_bstride, (const bf16*)g, g_bstride, ln_type, ln_w, ln_b, (const bf16*)res, res_bstride,      \
        (bf16*)dx, dx_bstride, (bf16*)xn, xn_bstride, parts, HW, tps2, (int64_t)B * tps2);                             \
    return check_launch("ln_bwd_bf16_reg");                                                                            \
  }
    if (C == 48) B200IR_LN_BWD_REG(6)
    if (C == 96) B200IR_LN_BWD_REG(12)
#undef B200IR_LN_BWD_REG
    ln_bwd_bf16x2_kernel<<<nparts, 256, (size_t)2 * C * sizeof(float), (cudaStream_t)stream>>>(
        (const bf16*)x, x_bstride, (const bf16*)g, g_bstride, ln_type, ln_w, ln_b, (const bf16*)res, res_bstride,
        (bf16*)dx, dx_bstride, (bf16*)xn, xn_bstride, parts, C, HW, tps2, (int64_t)B * tps2);
    return check_launch("ln_bwd_bf16x2");
  }
  if (same && x_dtype == B200IR_BF16) B200IR_LN_BWD(B200IR_BF16);
  else if (same) B200IR_LN_BWD(B200IR_F32);
  else B200IR_LN_BWD(-1);
#undef B200IR_LN_BWD
  return check_launch("ln_bwd");
}

extern "C" int b200ir_reduce_partials(const float* in, float* out, int n, int64_t K, int64_t stride,
                                      b200ir_stream_t stream) {
  B200IR_CHECK_ARG(in && out && n > 0 && K > 0 && stride >= K, "reduce_partials: bad args");
  // (SL = 32 for the few-output / many-partial folds was measured: 26.4 vs 26.0-26.1 ms per training step -- the 1024-thread
  // CTAs on the side stream take residency from the main-stream kernels; not adopted)
  reduce_partials_kernel<8><<<grid_1d(K, 32, num_sms() * 16), 256, 0, (cudaStream_t)stream>>>(in, out, n, K, stride);
  return check_launch("reduce_partials");
}

extern "C" int b200ir_reduce_partials_taps(const float* in, float* dw, float* db, int n, int C, int nv, b200ir_stream_t stream) {
  B200IR_CHECK_ARG(in && dw && n > 0 && C > 0 && nv > 1, "reduce_partials_taps: bad args");
  reduce_partials_taps_kernel<<<grid_1d((int64_t)C * nv, 32, num_sms() * 16), 256, 0, (cudaStream_t)stream>>>(in, dw, db, n, C, nv);
  return check_launch("reduce_partials_taps");
}


namespace b200ir {
bool wgrad_tc_supported(const void* dy, int64_t dy_bs, int dy_dt, const void* x, int64_t x_bs, int x_dt, int taps,
                        int64_t HW);                                                    // wgrad_tc.cu
void wgrad_tc_tiles(int M, int N, int* mtiles, int* ntiles, int* Ntile);
int wgrad_tc_launch(const void* dy, int64_t dy_bs, int dy_dt, const void* x, int64_t x_bs, float* pw, float* pb, int B,
                    int M, int N, int64_t HW, int per_sample, int nsplit, int chunks_per_sample, int chunks_per_split,
                    int chunks_total, cudaStream_t st);

// split-K plan shared by the CUDA-core and the tcgen05 kernels (depends on the shape only, so the caller can size the
// partial buffers before knowing which kernel runs): ~2 waves of 148 SMs worth of 128 x <=256 tiles, <= 64 splits.
static void wgrad_split(int B, int M, int N, int64_t HW, int per_sample, int* nsplit, int* cps, int* cpsplit,
                        int* ctotal) {
  const int chunks_per_sample = (int)((HW + WK - 1) / WK);
  const int total = per_sample ? chunks_per_sample : B * chunks_per_sample;
  int mt, nt, ntile;
  wgrad_tc_tiles(M, N, &mt, &nt, &ntile);
  const int64_t tiles = (int64_t)mt * nt * (per_sample ? B : 1);
  int64_t want = (2 * num_sms() + tiles - 1) / tiles;
  if (want > 64) want = 64;
  if (want > total) want = total;
  if (want < 1) want = 1;
  const int per = (int)((total + want - 1) / want);
  *nsplit = (total + per - 1) / per;
  *cps = chunks_per_sample;
  *cpsplit = per;
  *ctotal = total;
}
}  // namespace b200ir

extern "C" int b200ir_conv_wgrad_nsplit(int B, int Cout, int Cin, int taps, int64_t HW, int per_sample) {
  if (B <= 0 || Cout <= 0 || Cin <= 0 || HW <= 0 || (taps != 1 && taps != 9)) return B200IR_E_BADARG;
  int ns, a, b, c;
  wgrad_split(B, Cout, Cin * taps, HW, per_sample, &ns, &a, &b, &c);
  return ns;
}

extern "C" int b200ir_conv_wgrad(const void* dy, int64_t dy_bstride, int dy_dtype, const void* x, int64_t x_bstride,
                                 int x_dtype, float* parts_w, float* parts_b, int nsplit, int B, int Cout, int Cin,
                                 int H, int W, int taps, int per_sample, int impl, b200ir_stream_t stream) {
  B200IR_CHECK_ARG(dy && x && parts_w && B > 0 && Cout > 0 && Cin > 0 && H > 0 && W > 0, "conv_wgrad: bad args");
  B200IR_CHECK_ARG(taps == 1 || taps == 9, "conv_wgrad: taps must be 1 or 9");
  B200IR_CHECK_ARG(dtype_ok(dy_dtype) && dtype_ok(x_dtype), "conv_wgrad: bad dtype");
  int ns, cps, cpsplit, ctotal;
  wgrad_split(B, Cout, Cin * taps, (int64_t)H * W, per_sample, &ns, &cps, &cpsplit, &ctotal);
  B200IR_CHECK_ARG(ns == nsplit, "conv_wgrad: nsplit=%d, expected %d (b200ir_conv_wgrad_nsplit)", nsplit, ns);
  const int G = per_sample ? B : 1;
  B200IR_CHECK_ARG((int64_t)G * ns <= 65535, "conv_wgrad: grid too large");
  B200IR_CHECK_ARG(impl >= 0 && impl <= 2, "conv_wgrad: bad impl %d", impl);
  if (impl != B200IR_IMPL_SIMT &&
      wgrad_tc_supported(dy, dy_bstride, dy_dtype, x, x_bstride, x_dtype, taps, (int64_t)H * W))
    return wgrad_tc_launch(dy, dy_bstride, dy_dtype, x, x_bstride, parts_w, parts_b, B, Cout, Cin, (int64_t)H * W,
                           per_sample, ns, cps, cpsplit, ctotal, (cudaStream_t)stream);
  B200IR_UNSUPPORTED(impl != B200IR_IMPL_TC, "conv_wgrad: tcgen05 path unsupported for this shape/dtype");
  dim3 grid((unsigned)((Cin * taps + WN - 1) / WN), (unsigned)((Cout + WM - 1) / WM), (unsigned)(G * ns));
  wgrad_simt_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(dy, dy_bstride, dy_dtype, x, x_bstride, x_dtype, parts_w,
                                                            parts_b, Cout, Cin, taps, H, W, ns, cps, cpsplit, ctotal,
                                                            per_sample);
  return check_launch("conv_wgrad_simt");
}

extern "C" int b200ir_mdta_bwd(const float* partials, int nslab, const float* attn, const float* temperature,
                               const float* w_out, const float* dM, float* mt, float* mqk, float* dw_out_part,
                               float* dtemp_part, int B, int C, int heads, b200ir_stream_t stream) {
  return b200ir_mdta_bwd_ex(partials, nslab, attn, temperature, 0, w_out, dM, mt, mqk, dw_out_part, dtemp_part, B, C, heads,
                            stream);
}

extern "C" int b200ir_mdta_bwd_ex(const float* partials, int nslab, const float* attn, const float* temperature,
                                  int temp_per_sample, const float* w_out, const float* dM, float* mt, float* mqk,
                                  float* dw_out_part, float* dtemp_part, int B, int C, int heads, b200ir_stream_t stream) {
  return b200ir_mdta_bwd_parts(partials, nslab, attn, temperature, temp_per_sample, w_out, dM, mt, mqk, dw_out_part,
                               dtemp_part, B, C, heads, B200IR_MDTA_BWD_CHAIN | B200IR_MDTA_BWD_FOLD, stream);
}

extern "C" int b200ir_mdta_bwd_parts(const float* partials, int nslab, const float* attn, const float* temperature,
                                     int temp_per_sample, const float* w_out, const float* dM, float* mt, float* mqk,
                                     float* dw_out_part, float* dtemp_part, int B, int C, int heads, int parts,
                                     b200ir_stream_t stream) {
  const bool chain = parts & B200IR_MDTA_BWD_CHAIN, fold = parts & B200IR_MDTA_BWD_FOLD;
  B200IR_CHECK_ARG(chain || fold, "mdta_bwd: no part selected");
  B200IR_CHECK_ARG(attn && w_out && dM && (!chain || (partials && temperature && mqk && dtemp_part)) &&
                   (!fold || (mt && dw_out_part)), "mdta_bwd: null pointer");
  B200IR_CHECK_ARG(nslab > 0 && B > 0 && C > 0 && heads > 0 && C % heads == 0 && B <= 65535, "mdta_bwd: bad shape");
  const int d = C / heads;
  B200IR_UNSUPPORTED(d <= 128, "mdta_bwd: head dim %d > 128 unsupported", d);
  // fixed part (three d x (d+1) matrices + four vectors) + the staged W_out / dM rows, CO of C at a time
  const size_t fixed = (3 * (size_t)d * (d + 1) + 4 * (size_t)d) * sizeof(float), limit = 220 * 1024;
  B200IR_UNSUPPORTED(fixed + 2 * 72 * (size_t)d * sizeof(float) <= limit, "mdta_bwd: head dim %d needs %zu B of shared memory",
                     d, fixed);
  // rows of o staged at a time (multiple of 4) and the padded row stride S of the transposed blocks (S % 32 == 4)
  int CO = (int)((limit - fixed - 16) / (2 * (size_t)d * sizeof(float))) - 32;
  if (CO > C) CO = (C + 3) & ~3;
  CO &= ~3;
  const int S = CO + ((36 - CO % 32) % 32);
  const size_t smem = fixed + 16 + 2 * (size_t)d * S * sizeof(float);
  const size_t smem2 = ((size_t)d * (d + 1) + 64 * (size_t)(d + 1)) * sizeof(float);
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(mdta_bwd_chain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    cudaFuncSetAttribute(mdta_bwd_fold_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
    attr_set = true;
  }
  if (chain) {
    mdta_bwd_chain_kernel<<<dim3((unsigned)heads, (unsigned)B), MB_THREADS, smem, (cudaStream_t)stream>>>(
        partials, nslab, attn, temperature, w_out, dM, mqk, dtemp_part, C, d, temp_per_sample ? 1 : 0, CO, S);
    const int rc = check_launch("mdta_bwd_chain");
    if (rc) return rc;
  }
  if (fold)
    mdta_bwd_fold_kernel<<<dim3((unsigned)((C + 31) / 32), (unsigned)heads, (unsigned)B), 256, smem2,
                           (cudaStream_t)stream>>>(attn, w_out, dM, mt, dw_out_part, C, d);
  return check_launch("mdta_bwd");
}

extern "C" int b200ir_im2col3x3(const void* x, int64_t x_bstride, int x_dtype, void* y, int B, int Cin, int H, int W,
                                b200ir_stream_t stream) {
  B200IR_CHECK_ARG(x && y && B > 0 && Cin > 0 && H > 0 && W > 0 && B <= 65535 && Cin <= 65535, "im2col3x3: bad args");
  B200IR_CHECK_ARG(dtype_ok(x_dtype), "im2col3x3: bad dtype");
  dim3 grid((unsigned)(((int64_t)H * W + 255) / 256), (unsigned)Cin, (unsigned)B);
  im2col3x3_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x, x_bstride, x_dtype, (bf16*)y, Cin, H, W);
  return check_launch("im2col3x3");
}

extern "C" int b200ir_pixel_shuffle(const void* x, int64_t x_bstride, void* y, int64_t y_bstride, int dtype, int B,
                                    int C, int H, int W, int mode, b200ir_stream_t stream) {
  B200IR_CHECK_ARG(x && y && B > 0 && C > 0 && H > 0 && W > 0 && B <= 65535, "pixel_shuffle: bad args");
  B200IR_CHECK_ARG(dtype_ok(dtype), "pixel_shuffle: bad dtype");
  B200IR_CHECK_ARG(mode == B200IR_SHUFFLE2 || mode == B200IR_UNSHUFFLE2, "pixel_shuffle: bad mode %d", mode);
  B200IR_CHECK_ARG(mode != B200IR_SHUFFLE2 || C % 4 == 0, "pixel_shuffle: PixelShuffle(2) needs C%%4==0");
  B200IR_CHECK_ARG(mode != B200IR_UNSHUFFLE2 || (H % 2 == 0 && W % 2 == 0), "pixel_shuffle: needs even H,W");
  const int64_t n = (int64_t)C * H * W;
  pixel_shuffle_kernel<<<dim3((unsigned)grid_1d(n, 256, num_sms() * 8), (unsigned)B), 256, 0, (cudaStream_t)stream>>>(
      x, x_bstride, y, y_bstride, dtype, C, H, W, mode);
  return check_launch("pixel_shuffle");
}

extern "C" int b200ir_l1_loss_fwd(const void* pred, int pred_dtype, const void* target, int target_dtype, int64_t n,
                                  float* parts, int nparts, float* loss, b200ir_stream_t stream) {
  B200IR_CHECK_ARG(pred && target && parts && loss && n > 0 && nparts > 0 && nparts <= 4096, "l1_loss_fwd: bad args");
  B200IR_CHECK_ARG(dtype_ok(pred_dtype) && dtype_ok(target_dtype), "l1_loss_fwd: bad dtype");
  l1_loss_fwd_kernel<<<nparts, 256, 0, (cudaStream_t)stream>>>(pred, pred_dtype, target, target_dtype, n, parts);
  int rc = check_launch("l1_loss_fwd");
  if (rc) return rc;
  l1_loss_finish_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(parts, nparts, 1.0f / (float)n, loss);
  return check_launch("l1_loss_finish");
}

// ---- allocation-free L1 loss (no counterpart in the reference) ----------
// The per-block partial sums live in a file-scope device array instead of a caller-provided buffer, so the call needs no
// cudaMalloc (callers under CUDA graph capture, tight serving loops).  16 KB per GPU that loads the library, held until
// exit.  In host code `g_l1_noalloc_parts` is only a stand-in for the symbol: its device address is looked up once per
// device (with that device current) and passed to the kernels.  Every call on a device shares the array: calls that
// overlap on the GPU (two streams) race.  Nothing has to start at zero: the first kernel writes parts[0, nparts) before
// the finishing kernel (same stream) reads exactly that range.
static __device__ __align__(256) float g_l1_noalloc_parts[B200IR_L1_NOALLOC_MAX_PARTS];

static int l1_noalloc_scratch(float** out) {
  constexpr int kMaxDev = 64;
  static std::mutex mu;
  static float* addr[kMaxDev] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxDev) {
    set_error("l1_loss_fwd_noalloc: cudaGetDevice: %s", e != cudaSuccess ? cudaGetErrorString(e) : "device index >= 64");
    return B200IR_E_DRIVER;
  }
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[dev]) {
    float* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, g_l1_noalloc_parts);   // the current device's copy of the array
    if (e != cudaSuccess || !p) {
      set_error("l1_loss_fwd_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
      return B200IR_E_DRIVER;
    }
    addr[dev] = p;
  }
  *out = addr[dev];
  return 0;
}

extern "C" int b200ir_l1_loss_fwd_noalloc(const void* pred, int pred_dtype, const void* target, int target_dtype,
                                          int64_t n, float* loss, b200ir_stream_t stream) {
  B200IR_CHECK_ARG(pred && target && loss && n > 0, "l1_loss_fwd_noalloc: bad args");
  B200IR_CHECK_ARG(dtype_ok(pred_dtype) && dtype_ok(target_dtype), "l1_loss_fwd_noalloc: bad dtype");
  if (n > (int64_t)B200IR_L1_NOALLOC_MAX_N) {
    set_error("l1_loss_fwd_noalloc: n = %lld exceeds the built-in scratch (max %lld elements); use b200ir_l1_loss_fwd "
              "with a caller-provided parts buffer", (long long)n, (long long)B200IR_L1_NOALLOC_MAX_N);
    return B200IR_E_WORKSPACE;
  }
  float* parts = nullptr;
  int rc = l1_noalloc_scratch(&parts);
  if (rc) return rc;
  const int nparts = (int)((n + B200IR_L1_NOALLOC_ELEMS_PER_PART - 1) / B200IR_L1_NOALLOC_ELEMS_PER_PART);
  l1_loss_fwd_kernel<<<nparts, 256, 0, (cudaStream_t)stream>>>(pred, pred_dtype, target, target_dtype, n, parts);
  rc = check_launch("l1_loss_fwd_noalloc");
  if (rc) return rc;
  l1_loss_finish_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(parts, nparts, 1.0f / (float)n, loss);
  return check_launch("l1_loss_finish_noalloc");
}

extern "C" int b200ir_l1_loss_bwd(const void* pred, int pred_dtype, const void* target, int target_dtype, int64_t n,
                                  const float* gscale, void* grad, int grad_dtype, b200ir_stream_t stream) {
  B200IR_CHECK_ARG(pred && target && gscale && grad && n > 0, "l1_loss_bwd: bad args");
  B200IR_CHECK_ARG(dtype_ok(pred_dtype) && dtype_ok(target_dtype) && dtype_ok(grad_dtype), "l1_loss_bwd: bad dtype");
  l1_loss_bwd_kernel<<<grid_1d(n, 256, num_sms() * 8), 256, 0, (cudaStream_t)stream>>>(pred, pred_dtype, target, target_dtype,
                                                                                n, gscale, 1.0f / (float)n, grad,
                                                                                grad_dtype);
  return check_launch("l1_loss_bwd");
}

extern "C" int b200ir_adamw_step(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1,
                                 float beta2, float eps, float weight_decay, int step, int* step_dev, float grad_scale,
                                 b200ir_stream_t stream) {
  return b200ir_adamw_step_ex(p, g, m, v, n, lr, beta1, beta2, eps, weight_decay, step, step_dev, 1, nullptr, grad_scale,
                              stream);
}

extern "C" int b200ir_adamw_step_ex(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1,
                                    float beta2, float eps, float weight_decay, int step, int* step_dev, int tick,
                                    const float* hyper_dev, float grad_scale, b200ir_stream_t stream) {
  B200IR_CHECK_ARG(p && g && m && v && n > 0 && (step >= 1 || step_dev), "adamw_step: bad args");
  double bc1 = 1.0, bc2 = 1.0;
  if (step_dev) {  // ++*step_dev on the stream, then the update reads it: replayable inside a CUDA graph
    if (tick) {
      counter_inc_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(step_dev);
      int rc = check_launch("adamw_tick");
      if (rc) return rc;
    }
  } else {
    bc1 = 1.0 - pow((double)beta1, (double)step);
    bc2 = 1.0 - pow((double)beta2, (double)step);
  }
  adamw_kernel<<<grid_1d(n, 256, num_sms() * 8), 256, 0, (cudaStream_t)stream>>>(
      p, g, m, v, n, lr, beta1, beta2, eps, weight_decay, (float)bc1, (float)sqrt(bc2), grad_scale, step_dev, hyper_dev);
  return check_launch("adamw_step");
}
