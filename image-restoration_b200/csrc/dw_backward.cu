// dw_backward.cu — backward of the depthwise 3x3 convs (Attention.qkv_dwconv Restormer.py:106,114 and
// FeedForward.dwconv + gelu gate :84,90-91): weight / bias gradients and the gate's input gradient.
// (The input gradient of a depthwise conv is the forward kernel with flipped taps, launched from the host side.)
//
// Both kernels stage a (TH+2) x (TW+2) halo tile of one plane in shared memory (1024 output pixels per tile, the
// tile shape follows the image width), every thread owns 4 consecutive pixels of a row, and the 9 tap products + the
// bias sum are accumulated in registers across all tiles a CTA visits, then reduced once per CTA in fixed order:
//   parts[part][c][10],  part = b * nsplit + split   ->  b200ir_reduce_partials folds parts (deterministic).
#include "common.cuh"

namespace b200ir {

constexpr int DWB_PIX = 1024;              // pixels per tile
constexpr int DWB_SM = 1200;               // >= (TH+2)*(TW+2) for (TW,TH) in {(64,16),(32,32),(16,64)}

struct DwTile {
  int TW, TH, tiles_x, tiles_y, ntiles;
};

__host__ __device__ inline DwTile dw_tile(int H, int W) {
  DwTile t;
  t.TW = W > 32 ? 64 : (W > 16 ? 32 : 16);
  t.TH = DWB_PIX / t.TW;
  t.tiles_x = (W + t.TW - 1) / t.TW;
  t.tiles_y = (H + t.TH - 1) / t.TH;
  t.ntiles = t.tiles_x * t.tiles_y;
  return t;
}

__device__ __forceinline__ void load_halo(float* __restrict__ s, const void* __restrict__ x, int dt, int64_t base,
                                          int H, int W, int y0, int x0, int TW, int TH) {
  const int SW = TW + 2, n = (TH + 2) * SW;
  for (int e = threadIdx.x; e < n; e += 256) {
    const int r = e / SW, cc = e - r * SW;
    const int gy = y0 + r - 1, gx = x0 + cc - 1;
    float v = 0.f;
    if (gy >= 0 && gy < H && gx >= 0 && gx < W) v = ld_any(x, dt, base + (int64_t)gy * W + gx);
    s[e] = v;
  }
}

// reduce NV per-thread accumulators over the CTA (fixed order) and write them to out[0..NV)
template <int NV>
__device__ __forceinline__ void block_reduce_store(float (&acc)[NV], float* __restrict__ out) {
  __shared__ float red[8][NV];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const float v = warp_sum(acc[i]);
    if (lane == 0) red[warp][i] = v;
  }
  __syncthreads();
  if (threadIdx.x < NV) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += red[w][threadIdx.x];
    out[threadIdx.x] = s;
  }
}

// ---- depthwise weight/bias gradient: grid (nsplit, C, B) ----
__global__ void __launch_bounds__(256)
dwconv_wgrad_tile_kernel(const void* __restrict__ dy, int64_t dy_bs, const void* __restrict__ x, int64_t x_bs, int dt,
                         float* __restrict__ parts, int C, int H, int W) {
  __shared__ float xs[DWB_SM];
  const DwTile T = dw_tile(H, W);
  const int c = blockIdx.y, b = blockIdx.z, nsplit = gridDim.x;
  const int64_t HW = (int64_t)H * W;
  const int64_t dyo = (int64_t)b * dy_bs + (int64_t)c * HW, xo = (int64_t)b * x_bs + (int64_t)c * HW;
  const int tpr = T.TW >> 2;                       // threads per tile row
  const int tx = threadIdx.x % tpr, ty = threadIdx.x / tpr;
  const int SW = T.TW + 2;
  float acc[10];
#pragma unroll
  for (int i = 0; i < 10; ++i) acc[i] = 0.f;
  for (int tile = blockIdx.x; tile < T.ntiles; tile += nsplit) {
    const int y0 = (tile / T.tiles_x) * T.TH, x0 = (tile % T.tiles_x) * T.TW;
    __syncthreads();
    load_halo(xs, x, dt, xo, H, W, y0, x0, T.TW, T.TH);
    __syncthreads();
    const int gy = y0 + ty;
    if (gy < H) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int lx = tx * 4 + j, gx = x0 + lx;
        if (gx < W) {
          const float g = ld_any(dy, dt, dyo + (int64_t)gy * W + gx);
          const float* s = xs + ty * SW + lx;
#pragma unroll
          for (int k = 0; k < 9; ++k) acc[k] = fmaf(g, s[(k / 3) * SW + (k % 3)], acc[k]);
          acc[9] += g;
        }
      }
    }
  }
  block_reduce_store<10>(acc, parts + (((size_t)b * nsplit + blockIdx.x) * C + c) * 10);
}

// ---- GDFN gate backward + depthwise weight/bias gradient: grid (nsplit, Hd, B) ----
//   a = dw(t)+bias (recomputed), da[c] = dg[c]*a[c+Hd]*gelu'(a[c]), da[c+Hd] = dg[c]*gelu(a[c])
//   parts[part][c][.] and parts[part][c+Hd][.] = partial sum_p da * shifted t | sum_p da
__global__ void __launch_bounds__(256)
gelu_gate_bwd_tile_kernel(const void* __restrict__ t, int64_t t_bs, const float* __restrict__ w,
                          const float* __restrict__ bias, const void* __restrict__ dg, int64_t dg_bs,
                          void* __restrict__ da, int64_t da_bs, int dt, float* __restrict__ parts, int Hd, int H, int W,
                          int gate_fn) {
  __shared__ float s1[DWB_SM], s2[DWB_SM];
  const DwTile T = dw_tile(H, W);
  const int c = blockIdx.y, b = blockIdx.z, nsplit = gridDim.x;
  const int64_t HW = (int64_t)H * W;
  const int64_t t1o = (int64_t)b * t_bs + (int64_t)c * HW, t2o = t1o + (int64_t)Hd * HW;
  const int64_t go = (int64_t)b * dg_bs + (int64_t)c * HW;
  const int64_t a1o = (int64_t)b * da_bs + (int64_t)c * HW, a2o = a1o + (int64_t)Hd * HW;
  const int tpr = T.TW >> 2;
  const int tx = threadIdx.x % tpr, ty = threadIdx.x / tpr;
  const int SW = T.TW + 2;
  float w1[9], w2[9];
#pragma unroll
  for (int k = 0; k < 9; ++k) {
    w1[k] = w[c * 9 + k];
    w2[k] = w[(c + Hd) * 9 + k];
  }
  const float b1 = bias ? bias[c] : 0.f, b2 = bias ? bias[c + Hd] : 0.f;
  float acc[20];
#pragma unroll
  for (int i = 0; i < 20; ++i) acc[i] = 0.f;
  for (int tile = blockIdx.x; tile < T.ntiles; tile += nsplit) {
    const int y0 = (tile / T.tiles_x) * T.TH, x0 = (tile % T.tiles_x) * T.TW;
    __syncthreads();
    load_halo(s1, t, dt, t1o, H, W, y0, x0, T.TW, T.TH);
    load_halo(s2, t, dt, t2o, H, W, y0, x0, T.TW, T.TH);
    __syncthreads();
    const int gy = y0 + ty;
    if (gy < H) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int lx = tx * 4 + j, gx = x0 + lx;
        if (gx < W) {
          const float* p1 = s1 + ty * SW + lx;
          const float* p2 = s2 + ty * SW + lx;
          float v1[9], v2[9], a1 = b1, a2 = b2;
#pragma unroll
          for (int k = 0; k < 9; ++k) {
            v1[k] = p1[(k / 3) * SW + (k % 3)];
            v2[k] = p2[(k / 3) * SW + (k % 3)];
            a1 = fmaf(w1[k], v1[k], a1);
            a2 = fmaf(w2[k], v2[k], a2);
          }
          const int64_t off = (int64_t)gy * W + gx;
          const float g = ld_any(dg, dt, go + off);
          const float cdf = 0.5f * (1.0f + erff(a1 * 0.70710678118654752440f));
          const float pdf = 0.39894228040143267794f * expf(-0.5f * a1 * a1);
          float d1, d2;
          if (gate_fn == 0) {                      // y = gelu(a1) * a2               (Restormer.py:90-91)
            d1 = g * a2 * (cdf + a1 * pdf);
            d2 = g * a1 * cdf;
          } else {                                 // y = gelu(a1) * sigmoid(a2)      (experiment/RectiFormer.py:258-262)
            const float sg = 1.0f / (1.0f + expf(-a2));
            d1 = g * sg * (cdf + a1 * pdf);
            d2 = g * a1 * cdf * sg * (1.0f - sg);
          }
          st_any(da, dt, a1o + off, d1);
          st_any(da, dt, a2o + off, d2);
#pragma unroll
          for (int k = 0; k < 9; ++k) {
            acc[k] = fmaf(d1, v1[k], acc[k]);
            acc[10 + k] = fmaf(d2, v2[k], acc[10 + k]);
          }
          acc[9] += d1;
          acc[19] += d2;
        }
      }
    }
  }
  // two planes: [c] <- acc[0..10), [c+Hd] <- acc[10..20)
  __shared__ float red[8][20];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int i = 0; i < 20; ++i) {
    const float v = warp_sum(acc[i]);
    if (lane == 0) red[warp][i] = v;
  }
  __syncthreads();
  if (threadIdx.x < 20) {
    float s = 0.f;
#pragma unroll
    for (int wv = 0; wv < 8; ++wv) s += red[wv][threadIdx.x];
    const int plane = threadIdx.x < 10 ? c : c + Hd;
    parts[(((size_t)b * nsplit + blockIdx.x) * (2 * Hd) + plane) * 10 + threadIdx.x % 10] = s;
  }
}

static int dw_nsplit(int B, int C, int H, int W) {
  const DwTile T = dw_tile(H, W);
  int64_t want = (8 * num_sms() + (int64_t)B * C - 1) / ((int64_t)B * C);
  if (want > T.ntiles) want = T.ntiles;
  if (want > 256) want = 256;
  if (want < 1) want = 1;
  return (int)want;
}

// dw_backward_tma.cu (bf16 fast path)
bool dw_bwd_tma_shape_ok(int H, int W);
int dw_bwd_tma_tiles_per_plane(int H, int W);
template <bool GATE>
int dw_bwd_tma_launch(const void* x, int64_t x_bs, const void* g, int64_t g_bs, void* da, int64_t da_bs, const float* w,
                      const float* bias, float* parts, int B, int Cout, int H, int W, cudaStream_t st);

static bool dw_tma_ok(int dtype, int H, int W, const void* a, int64_t a_bs, const void* b, int64_t b_bs, const void* c,
                      int64_t c_bs) {
  auto al = [](const void* p, int64_t bs) { return !p || (((uintptr_t)p % 16) == 0 && bs % 8 == 0); };
  return dtype == B200IR_BF16 && dw_bwd_tma_shape_ok(H, W) && al(a, a_bs) && al(b, b_bs) && al(c, c_bs);
}

// parts beyond the ones the chosen kernel writes are zeroed, so the caller always folds `nparts` parts
static int zero_tail(float* parts, int nparts, int used, int64_t per_part, cudaStream_t st) {
  if (nparts > used) {
    cudaError_t e = cudaMemsetAsync(parts + (size_t)used * per_part, 0, (size_t)(nparts - used) * per_part * sizeof(float), st);
    if (e != cudaSuccess) {
      set_error("dw backward: cudaMemsetAsync: %s", cudaGetErrorString(e));
      return (int)e;
    }
  }
  return 0;
}

}  // namespace b200ir

using namespace b200ir;

extern "C" int b200ir_dwconv3x3_bwd_nparts(int B, int C, int H, int W) {
  if (B <= 0 || C <= 0 || H <= 0 || W <= 0) return B200IR_E_BADARG;
  int n = dw_nsplit(B, C, H, W);
  if (dw_bwd_tma_shape_ok(H, W)) {
    const int t = dw_bwd_tma_tiles_per_plane(H, W);
    if (t > n) n = t;
  }
  return B * n;   // capacity: whichever kernel runs writes <= this many parts and zero-fills the rest
}

extern "C" int b200ir_dwconv3x3_wgrad(const void* dy, int64_t dy_bstride, const void* x, int64_t x_bstride, int dtype,
                                      float* parts, int nparts, int B, int C, int H, int W, b200ir_stream_t stream) {
  B200IR_CHECK_ARG(dy && x && parts && B > 0 && C > 0 && H > 0 && W > 0 && B <= 65535 && C <= 65535,
                   "dwconv3x3_wgrad: bad args");
  B200IR_CHECK_ARG(dtype_ok(dtype), "dwconv3x3_wgrad: bad dtype");
  B200IR_CHECK_ARG(nparts == b200ir_dwconv3x3_bwd_nparts(B, C, H, W),
                   "dwconv3x3_wgrad: nparts=%d, expected %d (b200ir_dwconv3x3_bwd_nparts)", nparts,
                   b200ir_dwconv3x3_bwd_nparts(B, C, H, W));
  cudaStream_t st = (cudaStream_t)stream;
  if (dw_tma_ok(dtype, H, W, dy, dy_bstride, x, x_bstride, nullptr, 0)) {
    int rc = zero_tail(parts, nparts, B * dw_bwd_tma_tiles_per_plane(H, W), (int64_t)C * 10, st);
    if (rc) return rc;
    return dw_bwd_tma_launch<false>(x, x_bstride, dy, dy_bstride, nullptr, 0, nullptr, nullptr, parts, B, C, H, W, st);
  }
  const int ns = dw_nsplit(B, C, H, W);
  int rc = zero_tail(parts, nparts, B * ns, (int64_t)C * 10, st);
  if (rc) return rc;
  dwconv_wgrad_tile_kernel<<<dim3((unsigned)ns, (unsigned)C, (unsigned)B), 256, 0, st>>>(dy, dy_bstride, x, x_bstride,
                                                                                         dtype, parts, C, H, W);
  return check_launch("dwconv3x3_wgrad");
}

static int gate_bwd_impl(const void* t, int64_t t_bstride, const float* w, const float* bias, const void* dg,
                         int64_t dg_bstride, void* da, int64_t da_bstride, int dtype, float* parts, int nparts, int B,
                         int Hd, int H, int W, int gate_fn, b200ir_stream_t stream);

extern "C" int b200ir_dwconv3x3_gelu_gate_bwd(const void* t, int64_t t_bstride, const float* w, const float* bias,
                                              const void* dg, int64_t dg_bstride, void* da, int64_t da_bstride,
                                              int dtype, float* parts, int nparts, int B, int Hd, int H, int W,
                                              b200ir_stream_t stream) {
  return gate_bwd_impl(t, t_bstride, w, bias, dg, dg_bstride, da, da_bstride, dtype, parts, nparts, B, Hd, H, W, 0, stream);
}

extern "C" int b200ir_dwconv3x3_gelu_sigmoid_gate_bwd(const void* t, int64_t t_bstride, const float* w, const float* bias,
                                                      const void* dg, int64_t dg_bstride, void* da, int64_t da_bstride,
                                                      int dtype, float* parts, int nparts, int B, int Hd, int H, int W,
                                                      b200ir_stream_t stream) {
  return gate_bwd_impl(t, t_bstride, w, bias, dg, dg_bstride, da, da_bstride, dtype, parts, nparts, B, Hd, H, W, 1, stream);
}

static int gate_bwd_impl(const void* t, int64_t t_bstride, const float* w, const float* bias, const void* dg,
                         int64_t dg_bstride, void* da, int64_t da_bstride, int dtype, float* parts, int nparts, int B,
                         int Hd, int H, int W, int gate_fn, b200ir_stream_t stream) {
  B200IR_CHECK_ARG(t && w && dg && da && parts && B > 0 && Hd > 0 && H > 0 && W > 0, "gelu_gate_bwd: bad args");
  B200IR_CHECK_ARG(dtype_ok(dtype) && B <= 65535 && Hd <= 65535, "gelu_gate_bwd: bad dtype / grid");
  B200IR_CHECK_ARG(nparts == b200ir_dwconv3x3_bwd_nparts(B, Hd, H, W),
                   "gelu_gate_bwd: nparts=%d, expected %d (b200ir_dwconv3x3_bwd_nparts with C=Hd)", nparts,
                   b200ir_dwconv3x3_bwd_nparts(B, Hd, H, W));
  cudaStream_t st = (cudaStream_t)stream;
  if (gate_fn == 0 && dw_tma_ok(dtype, H, W, t, t_bstride, dg, dg_bstride, da, da_bstride)) {   // (sigmoid gate: tile kernel)
    int rc = zero_tail(parts, nparts, B * dw_bwd_tma_tiles_per_plane(H, W), (int64_t)Hd * 20, st);
    if (rc) return rc;
    return dw_bwd_tma_launch<true>(t, t_bstride, dg, dg_bstride, da, da_bstride, w, bias, parts, B, Hd, H, W, st);
  }
  const int ns = dw_nsplit(B, Hd, H, W);
  int rc = zero_tail(parts, nparts, B * ns, (int64_t)Hd * 20, st);
  if (rc) return rc;
  gelu_gate_bwd_tile_kernel<<<dim3((unsigned)ns, (unsigned)Hd, (unsigned)B), 256, 0, st>>>(
      t, t_bstride, w, bias, dg, dg_bstride, da, da_bstride, dtype, parts, Hd, H, W, gate_fn);
  return check_launch("gelu_gate_bwd");
}
