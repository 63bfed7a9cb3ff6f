// dw_backward_tma.cu — bf16 fast path of the depthwise backward kernels (see dw_backward.cu for the math), built on
// the same TMA-staged halo-tile pipeline as the forward kernel (dwconv_tma.cu): a persistent CTA streams
// (sample, channel, tile) items through a ring of shared-memory stages, one 4D TMA box per plane, zero padding from
// the TMA out-of-bounds fill, 8-pixel column strips with a 3-row sliding register window.
//   GATE = true : GDFN gate backward fused with the depthwise weight/bias gradient.  Stages t[c] and t[c+Hd],
//                 recomputes a = dw(t)+bias with packed fp32x2 FMAs, reads dg (one 128-bit load per row), writes
//                 da[c], da[c+Hd] (128-bit stores) and accumulates sum da*shifted t / sum da in registers.
//   GATE = false: depthwise weight/bias gradient only: stages x[c], reads dy, accumulates sum dy*shifted x / sum dy.
// Per item the 10 (20) accumulators are reduced over the CTA in fixed order and written to
//   parts[b * tiles + tile][channel][10]   (b200ir_reduce_partials folds the parts: deterministic).
// Reference: autograd of FeedForward.forward Restormer.py:90-91 and Attention.forward :114.
#include <cstdlib>
#include <mutex>

#include "dwconv_tma.cuh"

namespace b200ir {

#ifndef B200IR_DW_BW_RS
#define B200IR_DW_BW_RS 4   // measured: 64x64 planes 82 -> 49 us, 128x128 unchanged (more resident CTAs)
#endif
constexpr int BW_RS = B200IR_DW_BW_RS;   // output rows per thread

struct DwBwdParams {
  const bf16* g;     // dg [B,Hd,H,W] (GATE) or dy [B,C,H,W]
  int64_t g_bs;
  bf16* da;          // GATE only: [B,2Hd,H,W]
  int64_t da_bs;
  const float* w;    // GATE only
  const float* bias;
  float* parts;
  int Cout, Ctot, H, W, TW, TH, nxg, nstrips, tiles_x, tiles_y, S, nitems;
  int ncg;   // channels per work item: 1, or 128 / (nxg*nstrips) for small planes (W = 16 | 32), see dw_bwd_tma_tiles
};

__device__ __forceinline__ void unpack8(const uint4 v, float* o) {
  o[0] = bf_lo(v.x); o[1] = bf_hi(v.x); o[2] = bf_lo(v.y); o[3] = bf_hi(v.y);
  o[4] = bf_lo(v.z); o[5] = bf_hi(v.z); o[6] = bf_lo(v.w); o[7] = bf_hi(v.w);
}

template <bool GATE>
__global__ void __launch_bounds__(128)
dw_bwd_tma_kernel(const __grid_constant__ CUtensorMap tmX, const DwBwdParams P) {
  using namespace tc;
  constexpr int NIN = GATE ? 2 : 1;
  constexpr int NV = GATE ? 20 : 10;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 127) & ~(uintptr_t)127);
  __shared__ uint64_t bar_full[4];
  __shared__ float red[2][4][NV];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int tpc = P.nxg * P.nstrips;          // threads per channel plane
  const int nthreads = tpc * P.ncg;
  const uint32_t pitch = (uint32_t)(P.TW + 2 * DW_XP) * 2u;
  const uint32_t cplane = (uint32_t)(P.TH + 2) * pitch;                        // one channel of a box (dense)
  const uint32_t plane_bytes = ((uint32_t)P.ncg * cplane + 127u) & ~127u;      // one TMA box = ncg channels
  const uint32_t stage_bytes = plane_bytes * NIN;

  if (tid == 0) {
    for (int i = 0; i < P.S; ++i) mbar_init(&bar_full[i], 1);
    fence_barrier_init();
    tma_prefetch_desc(&tmX);
  }
  __syncthreads();

  // A CTA owns a CONTIGUOUS range of work items (tile fastest, then channel group, then sample), as in the forward
  // kernel (dwconv_tma.cu): the position advances without divisions, the taps are re-read only when the channel changes,
  // and the weight-gradient sums keep accumulating in registers over the consecutive tiles of one plane -- the CTA-wide
  // reduction runs once per plane segment (the other tiles' partial slots are written as zeros).
  const int tiles = P.tiles_x * P.tiles_y;
  const uint32_t cgroups = (uint32_t)((P.Cout + P.ncg - 1) / P.ncg);
  struct Pos {
    int tx, ty, cg, b;
  };
  auto decode = [&](uint32_t item) {
    const uint32_t t = item % (uint32_t)tiles, pc = item / (uint32_t)tiles;
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

  const int cs = tid / tpc, rr = tid - cs * tpc;      // channel of the group, thread within the channel plane
  const int xg = rr % P.nxg, strip = rr / P.nxg;
  const int64_t HWp = (int64_t)P.H * P.W;
  const uint32_t thr_base = smem_u32(smem) + (uint32_t)cs * cplane + (uint32_t)(strip * BW_RS) * pitch +
                            (uint32_t)(DW_XP + xg * 8) * 2u;
  int wc = -1;                            // channel whose taps are in registers (GATE)
  float2 wk[9], bs = make_float2(0.f, 0.f);
  float acc[NV];                          // weight / bias gradient sums of the current plane segment
#pragma unroll
  for (int i = 0; i < NV; ++i) acc[i] = 0.f;
  constexpr uint32_t NS = GATE ? 2 : 3;   // ring depth (= P.S; compile-time so that stage / phase are bit ops of `it`)
  for (uint32_t it = 0; it < count; ++it) {
    const uint32_t stage = it % NS, ph = (it / NS) & 1;
    const int cb = cur.cg * P.ncg, c = cb + cs, b = cur.b;
    const int x0 = cur.tx * P.TW, y0 = cur.ty * P.TH;
    const int t = cur.ty * P.tiles_x + cur.tx;
    Pos after = cur;
    advance(after);
    const bool flush = it + 1 == count || after.cg != cur.cg || after.b != cur.b;   // last tile of the plane segment

    if (tid < nthreads && c < P.Cout) {
      const int gx = x0 + xg * 8;
      const int64_t poff = (int64_t)(y0 + strip * BW_RS) * P.W + gx;
      const bf16* grow = P.g + (int64_t)b * P.g_bs + (int64_t)c * HWp + poff;
      const bool xok = gx < P.W;
      const int yrem = P.H - (y0 + strip * BW_RS);
      const uint32_t base = thr_base + stage * stage_bytes;
      if (GATE) {
        bf16* a1row = P.da + (int64_t)b * P.da_bs + (int64_t)c * HWp + poff;
        bf16* a2row = a1row + (int64_t)P.Cout * HWp;
        if (c != wc) {
          wc = c;
#pragma unroll
          for (int i = 0; i < 9; ++i) wk[i] = make_float2(__ldg(P.w + c * 9 + i), __ldg(P.w + (c + P.Cout) * 9 + i));
          if (P.bias) bs = make_float2(__ldg(P.bias + c), __ldg(P.bias + c + P.Cout));
        }
        float2 ac2[10];
#pragma unroll
        for (int i = 0; i < 10; ++i) ac2[i] = make_float2(0.f, 0.f);
        mbar_wait(&bar_full[stage], ph);
        float2 win[3][10];
        dw_load_row2(base, base + plane_bytes, win[0]);
        dw_load_row2(base + pitch, base + plane_bytes + pitch, win[1]);
#pragma unroll
        for (int j = 0; j < BW_RS; ++j) {
          const uint32_t ra = base + (uint32_t)(j + 2) * pitch;
          dw_load_row2(ra, ra + plane_bytes, win[(j + 2) % 3]);
          const bool ok = xok && j < yrem;
          float gv[8];
          unpack8(ok ? __ldg(reinterpret_cast<const uint4*>(grow + (int64_t)j * P.W)) : make_uint4(0, 0, 0, 0), gv);
          float d1[8], d2[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            float2 a = bs;
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
              for (int s2 = 0; s2 < 3; ++s2) a = ffma2(wk[r * 3 + s2], win[(j + r) % 3][i + s2], a);
            // cdf = Phi(a1) from the same fitted sigmoid form as gelu_fast; pdf = exp(-a1^2/2)/sqrt(2 pi)
            const float x = a.x, x2 = fminf(x * x, 64.f);
            float q = fmaf(1.0142652e-3f, x2, -1.0677574e-1f);
            q = fmaf(q, x2, -2.3011213f);
            float e, cdf, pe;
            asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(q * x));
            asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(cdf) : "f"(1.0f + e));
            asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(pe) : "f"(-0.72134752044448170368f * x * x));
            const float pdf = 0.39894228040143267794f * pe;
            d1[i] = gv[i] * a.y * fmaf(x, pdf, cdf);
            d2[i] = gv[i] * x * cdf;
            const float2 dd = make_float2(d1[i], d2[i]);
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
              for (int s2 = 0; s2 < 3; ++s2) ac2[r * 3 + s2] = ffma2(dd, win[(j + r) % 3][i + s2], ac2[r * 3 + s2]);
            ac2[9].x += d1[i];
            ac2[9].y += d2[i];
          }
          if (ok) {
            uint32_t p1[4], p2[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              __nv_bfloat162 h1 = __floats2bfloat162_rn(d1[2 * i], d1[2 * i + 1]);
              __nv_bfloat162 h2 = __floats2bfloat162_rn(d2[2 * i], d2[2 * i + 1]);
              p1[i] = *reinterpret_cast<uint32_t*>(&h1);
              p2[i] = *reinterpret_cast<uint32_t*>(&h2);
            }
            *reinterpret_cast<uint4*>(a1row + (int64_t)j * P.W) = make_uint4(p1[0], p1[1], p1[2], p1[3]);
            *reinterpret_cast<uint4*>(a2row + (int64_t)j * P.W) = make_uint4(p2[0], p2[1], p2[2], p2[3]);
          }
        }
#pragma unroll
        for (int i = 0; i < 10; ++i) {
          acc[i] += ac2[i].x;
          acc[(NV - 10) + i] += ac2[i].y;
        }
      } else {
        mbar_wait(&bar_full[stage], ph);
        float win[3][10];
        dw_load_row(base, win[0]);
        dw_load_row(base + pitch, win[1]);
#pragma unroll
        for (int j = 0; j < BW_RS; ++j) {
          dw_load_row(base + (uint32_t)(j + 2) * pitch, win[(j + 2) % 3]);
          const bool ok = xok && j < yrem;
          float gv[8];
          unpack8(ok ? __ldg(reinterpret_cast<const uint4*>(grow + (int64_t)j * P.W)) : make_uint4(0, 0, 0, 0), gv);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
              for (int s2 = 0; s2 < 3; ++s2) acc[r * 3 + s2] = fmaf(gv[i], win[(j + r) % 3][i + s2], acc[r * 3 + s2]);
            acc[9] += gv[i];
          }
        }
      }
    } else {
      mbar_wait(&bar_full[stage], ph);   // keep every thread's view of the barrier phase in step
    }
    const size_t slot = ((size_t)b * tiles + t) * P.Ctot;
    if (P.ncg > 1) {
      // small planes: the tpc (power of two <= 32) threads of a channel are consecutive lanes of one warp: butterfly
      // within that lane group (fixed order), its first lane writes the channel's sums
      const bool writer = tid < nthreads && c < P.Cout && rr == 0;
      if (flush) {
#pragma unroll
        for (int i = 0; i < NV; ++i) {
          float v = acc[i];
          for (int o = tpc >> 1; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
          acc[i] = v;
        }
      }
      if (writer) {
#pragma unroll
        for (int i = 0; i < NV; ++i) {
          const int plane = (GATE && i >= 10) ? c + P.Cout : c;
          P.parts[(slot + plane) * 10 + (i % 10)] = flush ? acc[i] : 0.f;
        }
      }
      if (flush) {
#pragma unroll
        for (int i = 0; i < NV; ++i) acc[i] = 0.f;
      }
      __syncthreads();   // every thread is done reading this stage
    } else {
      // fixed-order CTA reduction of the plane segment's sums
      float (*rb)[NV] = red[it & 1];
      if (flush) {
#pragma unroll
        for (int i = 0; i < NV; ++i) {
          const float v = warp_sum(acc[i]);
          if (lane == 0) rb[warp][i] = v;
          acc[i] = 0.f;
        }
      }
      __syncthreads();   // every thread is done reading this stage; rb complete
      if (tid < NV) {
        const float s = flush ? rb[0][tid] + rb[1][tid] + rb[2][tid] + rb[3][tid] : 0.f;
        const int plane = (GATE && tid >= 10) ? cb + P.Cout : cb;   // ncg == 1 here
        P.parts[(slot + plane) * 10 + (tid % 10)] = s;
      }
    }
    if (tid == 0 && staged < count) {
      fence_proxy_async();
      issue(nxt, (int)stage);
      advance(nxt);
      ++staged;
    }
    cur = after;
  }
}

// Small planes (W = 16 or 32: the two deep levels at 128x128 training patches) would leave most of the 128 threads
// idle with one plane per item, so an item then carries ncg = 128 / tpc channels (one TMA box with ncg channels), where
// tpc = nxg * nstrips threads (a power of two <= 32) cover one channel plane tile.
void dw_bwd_tma_tiles(int H, int W, int* TW, int* TH, int* nxg, int* nstrips, int* tiles_x, int* tiles_y, int* ncg) {
  const int maxstrips = (H + BW_RS - 1) / BW_RS;
  if (W < 64) {
    *TW = W;
    *nxg = W / 8;
    int ns = 1;
    while (ns < maxstrips && ns * 2 * *nxg <= 32) ns *= 2;
    *nstrips = ns;
    *ncg = 128 / (*nxg * ns);
  } else {
    *TW = W >= 128 ? 128 : 64;
    *nxg = *TW / 8;
    *nstrips = 128 / *nxg;
    if (*nstrips > maxstrips) *nstrips = maxstrips;
    *ncg = 1;
  }
  *TH = *nstrips * BW_RS;
  *tiles_x = (W + *TW - 1) / *TW;
  *tiles_y = (H + *TH - 1) / *TH;
}

bool dw_bwd_tma_shape_ok(int H, int W) {
#ifdef B200IR_AB
  static const bool small_ok = !(getenv("B200IR_DW_SMALL") && getenv("B200IR_DW_SMALL")[0] == '0');   // A/B switch
#else
  const bool small_ok = true;
#endif
  return (((W == 16 || W == 32) && small_ok) || (W % 8 == 0 && W >= 64)) && H >= 8;
}

int dw_bwd_tma_tiles_per_plane(int H, int W) {
  int TW, TH, nxg, ns, tx, ty, ncg;
  dw_bwd_tma_tiles(H, W, &TW, &TH, &nxg, &ns, &tx, &ty, &ncg);
  return tx * ty;
}

// x: staged tensor (t for GATE, x otherwise) [B,Ctot,H,W]; g: dg / dy
template <bool GATE>
int dw_bwd_tma_launch(const void* x, int64_t x_bs, const void* g, int64_t g_bs, void* da, int64_t da_bs, const float* w,
                      const float* bias, float* parts, int B, int Cout, int H, int W, cudaStream_t st) {
  DwBwdParams P;
  P.g = (const bf16*)g; P.g_bs = g_bs; P.da = (bf16*)da; P.da_bs = da_bs; P.w = w; P.bias = bias; P.parts = parts;
  P.Cout = Cout; P.Ctot = GATE ? 2 * Cout : Cout; P.H = H; P.W = W;
  dw_bwd_tma_tiles(H, W, &P.TW, &P.TH, &P.nxg, &P.nstrips, &P.tiles_x, &P.tiles_y, &P.ncg);
  const int64_t nitems64 = (int64_t)B * ((Cout + P.ncg - 1) / P.ncg) * P.tiles_x * P.tiles_y;
  if (nitems64 >= (1ll << 30)) {
    set_error("dw_bwd_tma: too many work items");
    return B200IR_E_UNSUPPORTED;
  }
  P.nitems = (int)nitems64;
  CUtensorMap tm;
  const uint64_t dims[4] = {(uint64_t)W, (uint64_t)H, (uint64_t)P.Ctot, (uint64_t)B};
  const uint64_t str[4] = {2, (uint64_t)W * 2, (uint64_t)H * W * 2, (uint64_t)x_bs * 2};
  const uint32_t box[4] = {(uint32_t)(P.TW + 2 * DW_XP), (uint32_t)(P.TH + 2), (uint32_t)P.ncg, 1};
  int rc = make_tmap_bf16(&tm, x, 4, dims, str, box, false);
  if (rc) return rc;
  const size_t pitch = (size_t)(P.TW + 2 * DW_XP) * 2;
  const size_t plane = ((size_t)P.ncg * (P.TH + 2) * pitch + 127) & ~(size_t)127;
  const size_t stage = plane * (GATE ? 2 : 1);
  P.S = GATE ? 2 : 3;
  const size_t smem = stage * P.S + 256;
  ensure_dyn_smem((const void*)dw_bwd_tma_kernel<GATE>, 200 * 1024);
  int ctas_per_sm = resident_ctas(dw_bwd_tma_kernel<GATE>, 128, smem);   // exactly the resident CTAs (static contiguous ranges; registers are the limit here)
  if (ctas_per_sm > 8) ctas_per_sm = 8;
  if (ctas_per_sm < 1) ctas_per_sm = 1;
  int64_t grid = (int64_t)num_sms() * ctas_per_sm;
  if (grid > P.nitems) grid = P.nitems;
  dw_bwd_tma_kernel<GATE><<<(unsigned)grid, 128, smem, st>>>(tm, P);
  return check_launch(GATE ? "gelu_gate_bwd_tma" : "dwconv3x3_wgrad_tma");
}

template int dw_bwd_tma_launch<false>(const void*, int64_t, const void*, int64_t, void*, int64_t, const float*,
                                      const float*, float*, int, int, int, int, cudaStream_t);
template int dw_bwd_tma_launch<true>(const void*, int64_t, const void*, int64_t, void*, int64_t, const float*,
                                     const float*, float*, int, int, int, int, cudaStream_t);

}  // namespace b200ir
