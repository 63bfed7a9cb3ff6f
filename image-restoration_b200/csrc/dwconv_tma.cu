// dwconv_tma.cu — depthwise 3x3 (stride 1, zero pad 1) for bf16 NCHW planes as a shared-memory-staged, halo-tiled,
// vectorised bandwidth kernel, optionally with the GDFN gate gelu(x1)*x2 fused.
// Reference: Attention.qkv_dwconv Restormer.py:106,114; FeedForward.dwconv + chunk(2) + F.gelu(x1)*x2 :84,90-91.
//
// One work item = (sample, channel [and channel+Hd], TH x TW output tile).  A persistent CTA streams items through a
// ring of shared-memory stages: ONE 4D TMA box per plane brings the (TH+2) x (TW+16) halo tile (the x origin is moved
// 8 pixels left so that the innermost TMA coordinate stays 16-byte aligned; zero padding = TMA out-of-bounds fill, so
// the inner loops carry no boundary predicates).  A thread owns an 8-pixel-wide column strip and slides a 3-row fp32
// register window down 8 output rows: per row one LDS.128 + two LDS.32 (left/right neighbour), 72 FFMA per plane,
// one 128-bit coalesced global store.  fp32 accumulation, bf16 storage.
// The gate uses x*sigmoid(p(x)), p an odd degree-5 polynomial fitted to the exact erf GELU (max abs error 2.5e-5,
// i.e. below bf16 resolution); the fp32 path (dwconv.cu) keeps erff.
#include <mutex>

#include "dwconv_tma.cuh"

namespace b200ir {

// forward geometry: FW_THREADS threads, each sliding its window down RS output rows (tile = TW x (threads/(TW/8))*RS).
// Measured on B200 (tools/bench_ops.py, C=96 level): plain dwconv is fastest with RS=8 (fewer window refills, 87 % of
// HBM); the gated kernel is issue/latency-bound and gains 22 % from RS=4 (smaller tiles -> twice the resident CTAs).
#ifndef B200IR_DW_GATE_RS
#define B200IR_DW_GATE_RS 4
#endif
#ifndef B200IR_DW_GATE_STAGES
#define B200IR_DW_GATE_STAGES 2
#endif
#ifndef B200IR_DW_FW_THREADS
#define B200IR_DW_FW_THREADS 128
#endif
constexpr int FW_THREADS = B200IR_DW_FW_THREADS;

template <bool GATE>
__global__ void __launch_bounds__(FW_THREADS)
dwconv_tma_kernel(const __grid_constant__ CUtensorMap tmX, const DwTmaParams P) {
  using namespace tc;
  constexpr int FW_RS = GATE ? B200IR_DW_GATE_RS : 8;
  constexpr int NIN = GATE ? 2 : 1;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 127) & ~(uintptr_t)127);
  __shared__ uint64_t bar_full[4];
  const int tid = threadIdx.x;
  const int tpc = P.nxg * P.nstrips;                                   // threads per channel plane tile
  const int nthreads = tpc * P.ncg;
  const uint32_t pitch = (uint32_t)(P.TW + 2 * DW_XP) * 2u;            // bytes per staged row
  const uint32_t cplane = (uint32_t)(P.TH + 2) * pitch;                // one channel of a box (dense)
  const uint32_t plane_bytes = ((uint32_t)P.ncg * cplane + 127u) & ~127u;   // one TMA box = ncg channels
  const uint32_t stage_bytes = plane_bytes * NIN;

  if (tid == 0) {
    for (int i = 0; i < P.S; ++i) mbar_init(&bar_full[i], 1);
    fence_barrier_init();
    tma_prefetch_desc(&tmX);
  }
  __syncthreads();

  const int tiles = P.tiles_x * P.tiles_y;
  const uint32_t cgroups = (uint32_t)((P.Cout + P.ncg - 1) / P.ncg);
  auto issue = [&](uint32_t item, int stage) {   // thread 0
    const uint32_t t = item % (uint32_t)tiles, pc = item / (uint32_t)tiles;
    const int c = (int)(pc % cgroups) * P.ncg, b = (int)(pc / cgroups);
    const int x0 = (t % P.tiles_x) * P.TW, y0 = (t / P.tiles_x) * P.TH;
    mbar_arrive_expect_tx(&bar_full[stage], cplane * (uint32_t)P.ncg * NIN);
#pragma unroll
    for (int k = 0; k < NIN; ++k)
      tma_load_4d(smem + (size_t)stage * stage_bytes + (size_t)k * plane_bytes, &tmX, &bar_full[stage], x0 - DW_XP,
                  y0 - 1, c + k * P.Cout, b);
  };

  const uint32_t first = blockIdx.x, step = gridDim.x;
  if (tid == 0)
    for (int s = 0; s < P.S; ++s)
      if (first + s * step < (uint32_t)P.nitems) issue(first + s * step, s);

  const int cs = tid / tpc, rr = tid - cs * tpc;       // channel within the group, thread within the plane tile
  const int xg = rr % P.nxg, strip = rr / P.nxg;
  const int64_t HWp = (int64_t)P.H * P.W;
  uint32_t it = 0;
  for (uint32_t item = first; item < (uint32_t)P.nitems; item += step, ++it) {
    const uint32_t stage = it % P.S, ph = (it / P.S) & 1;
    const uint32_t t = item % (uint32_t)tiles, pc = item / (uint32_t)tiles;
    const int c = (int)(pc % cgroups) * P.ncg + cs, b = (int)(pc / cgroups);
    const int x0 = (t % P.tiles_x) * P.TW, y0 = (t / P.tiles_x) * P.TH;

    if (tid < nthreads && c < P.Cout) {
      const int gx = x0 + xg * 8;
      bf16* orow = P.y + (int64_t)b * P.y_bs + (int64_t)c * HWp + (int64_t)(y0 + strip * FW_RS) * P.W + gx;
      const bool xok = gx < P.W;
      const int yrem = P.H - (y0 + strip * FW_RS);
      // staged row r <-> image row y0-1+r ; staged col 8+i <-> image col x0+i
      const uint32_t base = smem_u32(smem) + stage * stage_bytes + (uint32_t)cs * cplane +
                            (uint32_t)(strip * FW_RS) * pitch + (uint32_t)(DW_XP + xg * 8) * 2u;
      if (GATE) {
        float2 wk[9];
#pragma unroll
        for (int i = 0; i < 9; ++i) wk[i] = make_float2(__ldg(P.w + c * 9 + i), __ldg(P.w + (c + P.Cout) * 9 + i));
        const float2 bs = P.bias ? make_float2(__ldg(P.bias + c), __ldg(P.bias + c + P.Cout)) : make_float2(0.f, 0.f);
        mbar_wait(&bar_full[stage], ph);
        float2 win[3][10];
        dw_load_row2(base, base + plane_bytes, win[0]);
        dw_load_row2(base + pitch, base + plane_bytes + pitch, win[1]);
#pragma unroll
        for (int j = 0; j < FW_RS; ++j) {
          const uint32_t ra = base + (uint32_t)(j + 2) * pitch;
          dw_load_row2(ra, ra + plane_bytes, win[(j + 2) % 3]);
          uint32_t pk[4];
#pragma unroll
          for (int i = 0; i < 8; i += 2) {
            float2 a0 = bs, a1 = bs;
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
              for (int s2 = 0; s2 < 3; ++s2) {
                a0 = ffma2(wk[r * 3 + s2], win[(j + r) % 3][i + s2], a0);
                a1 = ffma2(wk[r * 3 + s2], win[(j + r) % 3][i + 1 + s2], a1);
              }
            __nv_bfloat162 h = __floats2bfloat162_rn(gelu_fast(a0.x) * a0.y, gelu_fast(a1.x) * a1.y);
            pk[i >> 1] = *reinterpret_cast<uint32_t*>(&h);
          }
          if (xok && j < yrem) *reinterpret_cast<uint4*>(orow + (int64_t)j * P.W) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
        }
      } else {
        float wk[9];
#pragma unroll
        for (int i = 0; i < 9; ++i) wk[i] = __ldg(P.w + c * 9 + (P.flip ? 8 - i : i));
        const float bs = P.bias ? __ldg(P.bias + c) : 0.f;
        mbar_wait(&bar_full[stage], ph);
        float win[3][10];
        dw_load_row(base, win[0]);
        dw_load_row(base + pitch, win[1]);
#pragma unroll
        for (int j = 0; j < FW_RS; ++j) {
          dw_load_row(base + (uint32_t)(j + 2) * pitch, win[(j + 2) % 3]);
          float o[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            float a = bs;
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
              for (int s2 = 0; s2 < 3; ++s2) a = fmaf(wk[r * 3 + s2], win[(j + r) % 3][i + s2], a);
            o[i] = a;
          }
          uint32_t pk[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            __nv_bfloat162 h = __floats2bfloat162_rn(o[2 * i], o[2 * i + 1]);
            pk[i] = *reinterpret_cast<uint32_t*>(&h);
          }
          if (xok && j < yrem) *reinterpret_cast<uint4*>(orow + (int64_t)j * P.W) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
        }
      }
    }
    __syncthreads();   // every thread is done reading this stage
    if (tid == 0) {
      const uint32_t nxt = item + (uint32_t)P.S * step;
      if (nxt < (uint32_t)P.nitems) {
        fence_proxy_async();
        issue(nxt, stage);
      }
    }
  }
}

bool dwconv_tma_supported(int dtype, int H, int W, int64_t x_bs, int64_t y_bs, const void* x, const void* y) {
  return dtype == B200IR_BF16 && W % 8 == 0 && W >= 16 && H >= 8 && x_bs % 8 == 0 && y_bs % 8 == 0 &&
         ((uintptr_t)x % 16) == 0 && ((uintptr_t)y % 16) == 0;
}

template <bool GATE>
int dwconv_tma_launch(const void* x, int64_t x_bs, const float* w, const float* bias, void* y, int64_t y_bs, int B,
                      int Cout, int H, int W, cudaStream_t st, int flip) {
  constexpr int FW_RS = GATE ? B200IR_DW_GATE_RS : 8;
  DwTmaParams P;
  P.y = (bf16*)y; P.y_bs = y_bs; P.w = w; P.bias = bias; P.Cout = Cout; P.H = H; P.W = W; P.flip = flip;
  P.TW = W >= 128 ? 128 : (W >= 64 ? 64 : W);           // planes narrower than 64 pixels: the whole width
  P.nxg = P.TW / 8;
  P.nstrips = FW_THREADS / P.nxg;                       // 8 (TW=128) or 16 (TW=64)
  const int maxstrips = (H + FW_RS - 1) / FW_RS;
  if (P.nstrips > maxstrips) P.nstrips = maxstrips;
  // narrow planes would leave most threads idle: carry several channels per item (one TMA box with ncg channels)
  P.ncg = W < 64 ? FW_THREADS / (P.nxg * P.nstrips) : 1;
  if (P.ncg > Cout) P.ncg = Cout;
  P.TH = P.nstrips * FW_RS;
  P.tiles_x = (W + P.TW - 1) / P.TW;
  P.tiles_y = (H + P.TH - 1) / P.TH;
  const int64_t nitems64 = (int64_t)B * ((Cout + P.ncg - 1) / P.ncg) * P.tiles_x * P.tiles_y;
  if (nitems64 >= (1ll << 30)) {
    set_error("dwconv_tma: too many work items");
    return B200IR_E_UNSUPPORTED;
  }
  P.nitems = (int)nitems64;
  const int Ctot = GATE ? 2 * Cout : Cout;
  CUtensorMap tm;
  const uint64_t dims[4] = {(uint64_t)W, (uint64_t)H, (uint64_t)Ctot, (uint64_t)B};
  const uint64_t str[4] = {2, (uint64_t)W * 2, (uint64_t)H * W * 2, (uint64_t)x_bs * 2};
  const uint32_t box[4] = {(uint32_t)(P.TW + 2 * DW_XP), (uint32_t)(P.TH + 2), (uint32_t)P.ncg, 1};
  int rc = make_tmap_bf16(&tm, x, 4, dims, str, box, false);
  if (rc) return rc;
  const size_t pitch = (size_t)(P.TW + 2 * DW_XP) * 2;
  const size_t plane = ((size_t)P.ncg * (P.TH + 2) * pitch + 127) & ~(size_t)127;
  const size_t stage = plane * (GATE ? 2 : 1);
  P.S = GATE ? B200IR_DW_GATE_STAGES : 3;
  const size_t smem = stage * P.S + 256;
  static std::once_flag once;
  std::call_once(once, [] {
    cudaFuncSetAttribute(dwconv_tma_kernel<GATE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  });
  int ctas_per_sm = (int)((220 * 1024) / (smem + 1024));
  if (ctas_per_sm > 8) ctas_per_sm = 8;
  if (ctas_per_sm < 1) ctas_per_sm = 1;
  int64_t grid = 148ll * ctas_per_sm;
  if (grid > P.nitems) grid = P.nitems;
  dwconv_tma_kernel<GATE><<<(unsigned)grid, FW_THREADS, smem, st>>>(tm, P);
  return check_launch(GATE ? "dwconv3x3_gelu_gate_tma" : "dwconv3x3_tma");
}

template int dwconv_tma_launch<false>(const void*, int64_t, const float*, const float*, void*, int64_t, int, int, int,
                                      int, cudaStream_t, int);
template int dwconv_tma_launch<true>(const void*, int64_t, const float*, const float*, void*, int64_t, int, int, int,
                                     int, cudaStream_t, int);

}  // namespace b200ir
