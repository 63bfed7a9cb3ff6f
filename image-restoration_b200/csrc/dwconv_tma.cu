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
#ifndef B200IR_DW_FOLD
#define B200IR_DW_FOLD 1
#endif
#ifndef B200IR_DW_HOIST
#define B200IR_DW_HOIST 0
#endif
#ifndef B200IR_DW_STAGEDIV
#define B200IR_DW_STAGEDIV 2   // stage / phase from the item counter with the compile-time ring depth (measured fastest)
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

  // A CTA owns a CONTIGUOUS range of work items (tile index fastest, then channel group, then sample): consecutive
  // items are neighbouring tiles of one plane, so the position advances with a few compares instead of divisions and
  // the 3x3 taps are re-read only when the channel changes (per-item overhead was ~27 % of the issued instructions).
  const uint32_t cgroups = (uint32_t)((P.Cout + P.ncg - 1) / P.ncg);
  struct Pos {
    int tx, ty, cg, b;
  };
  auto decode = [&](uint32_t item) {
    const uint32_t tiles = (uint32_t)(P.tiles_x * P.tiles_y);
    const uint32_t t = item % tiles, pc = item / tiles;
    Pos q;
    q.tx = (int)(t % (uint32_t)P.tiles_x); q.ty = (int)(t / (uint32_t)P.tiles_x);
    q.cg = (int)(pc % cgroups); q.b = (int)(pc / cgroups);
    return q;
  };
  auto advance = [&](Pos& q) {
    if (++q.tx == P.tiles_x) {
      q.tx = 0;
      if (++q.ty == P.tiles_y) {
        q.ty = 0;
        if (++q.cg == (int)cgroups) { q.cg = 0; ++q.b; }
      }
    }
  };
  auto issue = [&](const Pos& q, int stage) {   // thread 0
    mbar_arrive_expect_tx(&bar_full[stage], cplane * (uint32_t)P.ncg * NIN);
#pragma unroll
    for (int k = 0; k < NIN; ++k)
      tma_load_4d(smem + (size_t)stage * stage_bytes + (size_t)k * plane_bytes, &tmX, &bar_full[stage],
                  q.tx * P.TW - DW_XP, q.ty * P.TH - 1, q.cg * P.ncg + k * P.Cout, q.b);
  };

  const uint32_t per = (uint32_t)P.nitems / gridDim.x, rem = (uint32_t)P.nitems % gridDim.x;
  const uint32_t begin = blockIdx.x * per + (blockIdx.x < rem ? blockIdx.x : rem);
  const uint32_t count = per + (blockIdx.x < rem ? 1u : 0u);
  Pos cur = decode(begin), nxt = cur;     // nxt: the next item to stage (thread 0 runs P.S items ahead)
  uint32_t staged = 0;
  if (tid == 0)
    for (; staged < (uint32_t)P.S && staged < count; ++staged) {
      issue(nxt, (int)staged);
      advance(nxt);
    }

  const int cs = tid / tpc, rr = tid - cs * tpc;       // channel within the group, thread within the plane tile
  const int xg = rr % P.nxg, strip = rr / P.nxg;
  const int64_t HWp = (int64_t)P.H * P.W;
  // staged row r <-> image row y0-1+r ; staged col 8+i <-> image col x0+i
  const uint32_t thr_base = smem_u32(smem) + (uint32_t)cs * cplane + (uint32_t)(strip * FW_RS) * pitch +
                            (uint32_t)(DW_XP + xg * 8) * 2u;
  int wc = -1;                            // channel whose taps are in registers
  float2 wk2[9], bs2 = make_float2(0.f, 0.f);
  float wk1[9], bs1 = 0.f;
  uint32_t stage = 0, ph = 0;
  for (uint32_t it = 0; it < count; ++it) {
#if B200IR_DW_STAGEDIV == 1
    stage = it % P.S; ph = (it / P.S) & 1;
#elif B200IR_DW_STAGEDIV == 2
    if (GATE) { stage = it % B200IR_DW_GATE_STAGES; ph = (it / B200IR_DW_GATE_STAGES) & 1; }
    else { stage = it % 3u; ph = (it / 3u) & 1; }
#endif
    const int c = cur.cg * P.ncg + cs, b = cur.b;
    const int x0 = cur.tx * P.TW, y0 = cur.ty * P.TH;

    if (tid < nthreads && c < P.Cout) {
      const int gx = x0 + xg * 8;
      bf16* orow = P.y + (int64_t)b * P.y_bs + (int64_t)c * HWp + (int64_t)(y0 + strip * FW_RS) * P.W + gx;
      const bool xok = gx < P.W;
      const int yrem = P.H - (y0 + strip * FW_RS);
#if B200IR_DW_HOIST
      const uint32_t base = thr_base + stage * stage_bytes;
#else
      const uint32_t base = smem_u32(smem) + stage * stage_bytes + (uint32_t)cs * cplane +
                            (uint32_t)(strip * FW_RS) * pitch + (uint32_t)(DW_XP + xg * 8) * 2u;
#endif
      if (GATE) {
        if (c != wc) {
          wc = c;
#pragma unroll
          constexpr float kh = B200IR_DW_FOLD ? 0.5f : 1.0f;
          for (int i = 0; i < 9; ++i)   // the 0.5 of gelu(x) = 0.5 x (1 + tanh(..)) rides on the gate plane's taps
            wk2[i] = make_float2(__ldg(P.w + c * 9 + i), kh * __ldg(P.w + (c + P.Cout) * 9 + i));
          if (P.bias) bs2 = make_float2(__ldg(P.bias + c), kh * __ldg(P.bias + c + P.Cout));
        }
        const float2* wk = wk2;
        const float2 bs = bs2;
        mbar_wait(&bar_full[stage], ph);
        float2 win[3][10];
#if B200IR_DW_HOIST
        uint32_t ra = base, rb = base + plane_bytes;
        dw_load_row2(ra, rb, win[0]);
        ra += pitch; rb += pitch;
        dw_load_row2(ra, rb, win[1]);
#else
        dw_load_row2(base, base + plane_bytes, win[0]);
        dw_load_row2(base + pitch, base + plane_bytes + pitch, win[1]);
#endif
#pragma unroll
        for (int j = 0; j < FW_RS; ++j) {
#if B200IR_DW_HOIST
          ra += pitch; rb += pitch;
#else
          const uint32_t ra = base + (uint32_t)(j + 2) * pitch, rb = ra + plane_bytes;
#endif
          dw_load_row2(ra, rb, win[(j + 2) % 3]);
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
#if B200IR_DW_FOLD
            __nv_bfloat162 h = __floats2bfloat162_rn(gelu2_fast(a0.x) * a0.y, gelu2_fast(a1.x) * a1.y);
#else
            __nv_bfloat162 h = __floats2bfloat162_rn(gelu_fast(a0.x) * a0.y, gelu_fast(a1.x) * a1.y);
#endif
            pk[i >> 1] = *reinterpret_cast<uint32_t*>(&h);
          }
          if (xok && j < yrem) *reinterpret_cast<uint4*>(orow + (int64_t)j * P.W) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
        }
      } else {
        if (c != wc) {
          wc = c;
#pragma unroll
          for (int i = 0; i < 9; ++i) wk1[i] = __ldg(P.w + c * 9 + (P.flip ? 8 - i : i));
          if (P.bias) bs1 = __ldg(P.bias + c);
        }
        const float* wk = wk1;
        const float bs = bs1;
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
    if (tid == 0 && staged < count) {
      fence_proxy_async();
      issue(nxt, (int)stage);
      advance(nxt);
      ++staged;
    }
    advance(cur);
#if !B200IR_DW_STAGEDIV
    if (++stage == (uint32_t)P.S) { stage = 0; ph ^= 1u; }
#endif
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
  ensure_dyn_smem((const void*)dwconv_tma_kernel<GATE>, 200 * 1024);
  // static contiguous ranges: the grid must be exactly what is resident at once (registers limit the gated kernel to
  // 4 CTAs per SM although 5 would fit in shared memory -- a fifth "wave" of 148 CTAs ran alone at quarter occupancy)
  int ctas_per_sm = resident_ctas(dwconv_tma_kernel<GATE>, FW_THREADS, smem);
  if (ctas_per_sm > 8) ctas_per_sm = 8;
  if (ctas_per_sm < 1) ctas_per_sm = 1;
  int64_t grid = (int64_t)num_sms() * ctas_per_sm;
  if (grid > P.nitems) grid = P.nitems;
  dwconv_tma_kernel<GATE><<<(unsigned)grid, FW_THREADS, smem, st>>>(tm, P);
  return check_launch(GATE ? "dwconv3x3_gelu_gate_tma" : "dwconv3x3_tma");
}

template int dwconv_tma_launch<false>(const void*, int64_t, const float*, const float*, void*, int64_t, int, int, int,
                                      int, cudaStream_t, int);
template int dwconv_tma_launch<true>(const void*, int64_t, const float*, const float*, void*, int64_t, int, int, int,
                                     int, cudaStream_t, int);

}  // namespace b200ir
