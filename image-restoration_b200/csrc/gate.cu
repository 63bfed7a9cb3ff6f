// gate.cu — multiplicative sigmoid gate on the attention output, y = u * sigmoid(g), used by the gated MDTA variants of
// the reference's experiment models (experiment/baseline_restormer_new_train/
// 20251116_use_headwise_gate_model/.../moce_ir.py:155-166 GatedMDTA, experiment/RectiFormer.py:316 ElementwiseGated-
// Attention).  g has G channels, each gating C/G consecutive channels of u: G = C is the element-wise gate, G = heads
// the head-wise one (one scalar per head and pixel).  HBM-bound streaming kernels: a thread owns one gate element
// (sample, gate channel, pixel) and walks its C/G channels of u, so sigmoid(g) is evaluated once per gate element and
// the backward's reduction over the gated channels stays inside the thread (fixed order, deterministic).
#include "common.cuh"

namespace b200ir {

__device__ __forceinline__ float sigmoidf_(float v) { return 1.0f / (1.0f + __expf(-v)); }

__global__ void __launch_bounds__(256)
sigmoid_gate_fwd_kernel(const void* __restrict__ u, int64_t u_bs, const void* __restrict__ g, int64_t g_bs, int dt,
                        void* __restrict__ y, int64_t y_bs, int G, int cpg, int64_t HW, int64_t total) {
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (int64_t)gridDim.x * 256) {
    const int64_t p = i % HW;
    const int64_t t = i / HW;
    const int gc = (int)(t % G), b = (int)(t / G);
    const float s = sigmoidf_(ld_any(g, dt, (int64_t)b * g_bs + (int64_t)gc * HW + p));
    for (int k = 0; k < cpg; ++k) {
      const int64_t off = (int64_t)(gc * cpg + k) * HW + p;
      st_any(y, dt, (int64_t)b * y_bs + off, ld_any(u, dt, (int64_t)b * u_bs + off) * s);
    }
  }
}

// du = dy * sigmoid(g);  dg = sigmoid(g) (1 - sigmoid(g)) * sum_{c in group} dy[c] * u[c]
__global__ void __launch_bounds__(256)
sigmoid_gate_bwd_kernel(const void* __restrict__ u, int64_t u_bs, const void* __restrict__ g, int64_t g_bs,
                        const void* __restrict__ dy, int64_t dy_bs, int dt, void* __restrict__ du, int64_t du_bs,
                        void* __restrict__ dg, int64_t dg_bs, int G, int cpg, int64_t HW, int64_t total) {
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (int64_t)gridDim.x * 256) {
    const int64_t p = i % HW;
    const int64_t t = i / HW;
    const int gc = (int)(t % G), b = (int)(t / G);
    const float s = sigmoidf_(ld_any(g, dt, (int64_t)b * g_bs + (int64_t)gc * HW + p));
    float acc = 0.f;
    for (int k = 0; k < cpg; ++k) {
      const int64_t off = (int64_t)(gc * cpg + k) * HW + p;
      const float d = ld_any(dy, dt, (int64_t)b * dy_bs + off);
      acc = fmaf(d, ld_any(u, dt, (int64_t)b * u_bs + off), acc);
      st_any(du, dt, (int64_t)b * du_bs + off, d * s);
    }
    st_any(dg, dt, (int64_t)b * dg_bs + (int64_t)gc * HW + p, acc * s * (1.0f - s));
  }
}

static inline int gate_grid(int64_t n) {
  int64_t b = (n + 255) / 256;
  return (int)(b < num_sms() * 16 ? (b < 1 ? 1 : b) : num_sms() * 16);
}

}  // namespace b200ir

using namespace b200ir;

extern "C" int b200ir_sigmoid_gate_fwd(const void* u, int64_t u_bstride, const void* g, int64_t g_bstride, int dtype,
                                       void* y, int64_t y_bstride, int B, int C, int G, int64_t HW,
                                       b200ir_stream_t stream) {
  B200IR_CHECK_ARG(u && g && y && B > 0 && C > 0 && G > 0 && HW > 0, "sigmoid_gate: bad args");
  B200IR_CHECK_ARG(C % G == 0, "sigmoid_gate: C=%d not a multiple of the gate channels G=%d", C, G);
  B200IR_CHECK_ARG(dtype_ok(dtype), "sigmoid_gate: bad dtype %d", dtype);
  const int64_t total = (int64_t)B * G * HW;
  sigmoid_gate_fwd_kernel<<<gate_grid(total), 256, 0, (cudaStream_t)stream>>>(
      u, u_bstride, g, g_bstride, dtype, y, y_bstride, G, C / G, HW, total);
  return check_launch("sigmoid_gate_fwd");
}

extern "C" int b200ir_sigmoid_gate_bwd(const void* u, int64_t u_bstride, const void* g, int64_t g_bstride,
                                       const void* dy, int64_t dy_bstride, int dtype, void* du, int64_t du_bstride,
                                       void* dg, int64_t dg_bstride, int B, int C, int G, int64_t HW,
                                       b200ir_stream_t stream) {
  B200IR_CHECK_ARG(u && g && dy && du && dg && B > 0 && C > 0 && G > 0 && HW > 0, "sigmoid_gate_bwd: bad args");
  B200IR_CHECK_ARG(C % G == 0, "sigmoid_gate_bwd: C=%d not a multiple of the gate channels G=%d", C, G);
  B200IR_CHECK_ARG(dtype_ok(dtype), "sigmoid_gate_bwd: bad dtype %d", dtype);
  const int64_t total = (int64_t)B * G * HW;
  sigmoid_gate_bwd_kernel<<<gate_grid(total), 256, 0, (cudaStream_t)stream>>>(
      u, u_bstride, g, g_bstride, dy, dy_bstride, dtype, du, du_bstride, dg, dg_bstride, G, C / G, HW, total);
  return check_launch("sigmoid_gate_bwd");
}
