// gemm_simt.cu — LayerNorm + 1x1 conv (+bias, +residual) with fp32 FMA on CUDA cores.
// This is the fp32-exact path (max-abs <= 1e-4 vs the CPU reference rules out bf16/TF32 MMA,
// SURVEY.md §7.3-2) and the yard-stick the tcgen05 path (gemm_tc.cu) is checked against.
// Reference: LayerNorm.forward Restormer.py:37-39,54-57,68-70 followed by a 1x1 nn.Conv2d
// (:114 qkv, :89 project_in, :92/:131 project_out, :261/:266 reduce_chan) and the residual
// add of TransformerBlock.forward (:147-148).
#include "common.cuh"

namespace b200ir {

constexpr int TM = 64;   // out-channel tile
constexpr int TN = 64;   // pixel tile
constexpr int TK = 16;   // in-channel chunk

__global__ void __launch_bounds__(256)
ln_conv1x1_simt_kernel(const void* __restrict__ x, int64_t x_bs, int x_dt,
                       const float* __restrict__ w, int64_t w_bs, const float* __restrict__ bias,
                       int ln_type, const float* __restrict__ ln_w, const float* __restrict__ ln_b,
                       const void* __restrict__ res, int64_t res_bs, int res_dt,
                       void* __restrict__ y, int64_t y_bs, int y_dt,
                       int Cin, int Cout, int64_t HW) {
  __shared__ __align__(16) float Xs[TK][TN];
  __shared__ __align__(16) float Ws[TK][TM + 4];
  __shared__ float red[4][TN];
  __shared__ float s_mu[TN], s_rstd[TN];

  const int tid = threadIdx.x;
  const int b = blockIdx.z;
  const int64_t p0 = (int64_t)blockIdx.x * TN;
  const int o0 = blockIdx.y * TM;
  const char* xb = (const char*)x + (size_t)b * x_bs * (x_dt == B200IR_BF16 ? 2 : 4);
  const float* wb = w + (size_t)b * w_bs;

  // ---- LayerNorm statistics over channels for the 64 pixels of this tile (two-pass) ----
  if (ln_type != B200IR_LN_NONE) {
    const int px = tid & 63, q = tid >> 6;
    const int64_t p = p0 + px;
    float s = 0.f;
    if (p < HW)
      for (int c = q; c < Cin; c += 4) s += ld_any(xb, x_dt, (int64_t)c * HW + p);
    red[q][px] = s;
    __syncthreads();
    const float mu = (red[0][px] + red[1][px] + red[2][px] + red[3][px]) / (float)Cin;
    __syncthreads();
    float v = 0.f;
    if (p < HW)
      for (int c = q; c < Cin; c += 4) {
        float d = ld_any(xb, x_dt, (int64_t)c * HW + p) - mu;
        v += d * d;
      }
    red[q][px] = v;
    __syncthreads();
    if (q == 0) {
      float var = (red[0][px] + red[1][px] + red[2][px] + red[3][px]) / (float)Cin;
      s_mu[px] = (ln_type == B200IR_LN_WITHBIAS) ? mu : 0.f;  // BiasFree: numerator not centred (:39)
      s_rstd[px] = 1.0f / sqrtf(var + 1e-5f);
    }
    __syncthreads();
  }

  const int tx = tid & 15, ty = tid >> 4;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int k0 = 0; k0 < Cin; k0 += TK) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int e = tid + i * 256;
      {  // X tile: row = channel, col = pixel (coalesced over pixels)
        const int r = e >> 6, cpx = e & 63;
        const int c = k0 + r;
        const int64_t p = p0 + cpx;
        float v = 0.f;
        if (c < Cin && p < HW) {
          v = ld_any(xb, x_dt, (int64_t)c * HW + p);
          if (ln_type != B200IR_LN_NONE) {
            v = (v - s_mu[cpx]) * s_rstd[cpx] * ln_w[c];
            if (ln_type == B200IR_LN_WITHBIAS) v += ln_b[c];
          }
        }
        Xs[r][cpx] = v;
      }
      {  // W tile, stored transposed [k][o]
        const int o = e >> 4, k = e & 15;
        float v = 0.f;
        if (o0 + o < Cout && k0 + k < Cin) v = wb[(size_t)(o0 + o) * Cin + k0 + k];
        Ws[k][o] = v;
      }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < TK; ++k) {
      const float4 a = *reinterpret_cast<const float4*>(&Ws[k][ty * 4]);
      const float4 bb = *reinterpret_cast<const float4*>(&Xs[k][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w};
      const float bv[4] = {bb.x, bb.y, bb.z, bb.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }

  // ---- epilogue: + bias, + residual, store ----
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int o = o0 + ty * 4 + i;
    if (o >= Cout) continue;
    const float bo = bias ? bias[o] : 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t p = p0 + tx * 4 + j;
      if (p >= HW) continue;
      float v = acc[i][j] + bo;
      if (res) v += ld_any(res, res_dt, (int64_t)b * res_bs + (int64_t)o * HW + p);
      st_any(y, y_dt, (int64_t)b * y_bs + (int64_t)o * HW + p, v);
    }
  }
}

// ---- strided copy / cast ----
__global__ void copy_cast_kernel(const void* __restrict__ x, int64_t x_bs, int x_dt, void* __restrict__ y,
                                 int64_t y_bs, int y_dt, int64_t n) {
  const int b = blockIdx.y;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    st_any(y, y_dt, (int64_t)b * y_bs + i, ld_any(x, x_dt, (int64_t)b * x_bs + i));
}

}  // namespace b200ir

using namespace b200ir;

extern "C" int b200ir_ln_conv1x1_fwd(const void* x, int64_t x_bstride, int x_dtype, const float* w,
                                     int64_t w_bstride, const float* bias, int ln_type, const float* ln_w,
                                     const float* ln_b, const void* res, int64_t res_bstride, int res_dtype,
                                     void* y, int64_t y_bstride, int y_dtype, int B, int Cin, int Cout,
                                     int64_t HW, b200ir_stream_t stream) {
  B200IR_CHECK_ARG(x && w && y, "ln_conv1x1: null pointer");
  B200IR_CHECK_ARG(B > 0 && Cin > 0 && Cout > 0 && HW > 0, "ln_conv1x1: bad shape B=%d Cin=%d Cout=%d HW=%lld",
                   B, Cin, Cout, (long long)HW);
  B200IR_CHECK_ARG(dtype_ok(x_dtype) && dtype_ok(y_dtype) && (!res || dtype_ok(res_dtype)), "ln_conv1x1: bad dtype");
  B200IR_CHECK_ARG(ln_type >= 0 && ln_type <= 2, "ln_conv1x1: bad ln_type %d", ln_type);
  B200IR_CHECK_ARG(ln_type == B200IR_LN_NONE || ln_w, "ln_conv1x1: ln_w missing");
  B200IR_CHECK_ARG(ln_type != B200IR_LN_WITHBIAS || ln_b, "ln_conv1x1: ln_b missing for WithBias");
  B200IR_CHECK_ARG(B <= 65535, "ln_conv1x1: B too large");
  dim3 grid((unsigned)((HW + TN - 1) / TN), (unsigned)((Cout + TM - 1) / TM), (unsigned)B);
  ln_conv1x1_simt_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x, x_bstride, x_dtype, w, w_bstride, bias, ln_type,
                                                                 ln_w, ln_b, res, res_bstride, res_dtype, y,
                                                                 y_bstride, y_dtype, Cin, Cout, HW);
  return check_launch("ln_conv1x1_simt");
}

extern "C" int b200ir_copy_cast(const void* x, int64_t x_bstride, int x_dtype, void* y, int64_t y_bstride,
                                int y_dtype, int B, int64_t CHW, b200ir_stream_t stream) {
  B200IR_CHECK_ARG(x && y && B > 0 && CHW > 0 && B <= 65535, "copy_cast: bad args");
  B200IR_CHECK_ARG(dtype_ok(x_dtype) && dtype_ok(y_dtype), "copy_cast: bad dtype");
  int64_t blocks = (CHW + 255) / 256;
  if (blocks > num_sms() * 16) blocks = num_sms() * 16;
  copy_cast_kernel<<<dim3((unsigned)blocks, (unsigned)B), 256, 0, (cudaStream_t)stream>>>(x, x_bstride, x_dtype, y,
                                                                                          y_bstride, y_dtype, CHW);
  return check_launch("copy_cast");
}
