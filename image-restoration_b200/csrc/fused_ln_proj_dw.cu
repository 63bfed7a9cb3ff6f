// fused_ln_proj_dw.cu — LayerNorm + 1x1 conv + depthwise 3x3 (+ GELU gate) as ONE kernel: the 2*Hd-channel (GDFN) or
// 3*C-channel (MDTA) intermediate of the reference never reaches HBM.
//
// Reference ops replaced (all of them, for one call):
//   GDFN : norm2 -> project_in (1x1) -> dwconv 3x3 -> chunk(2) -> gelu(x1)*x2        Restormer.py:54-57,88-91
//   MDTA : norm1 -> qkv (1x1) -> qkv_dwconv 3x3                                       Restormer.py:54-57,114
// Algorithmic HBM bytes per pixel: (C + Hd)*2 instead of (C + 2Hd)*2 + (2Hd + Hd)*2 for the two-kernel path.
//
// Orientation: D[out-channel (TMEM lane)][pixel (TMEM column)] = W[128 x K] . Xhat[K x 64]   (tcgen05, fp32 in TMEM)
//   * the depthwise conv is then THREAD-LOCAL: a thread owns one channel (pair) = one TMEM lane and reads the 3x3
//     neighbourhood of its pixels straight from TMEM columns with tcgen05.ld -- no shared-memory staging of the
//     intermediate, no bf16 round trip, no unpacking; packed fma.rn.f32x2 over pixel pairs (the odd-aligned pairs of
//     the +-1 taps come from a second tcgen05.ld shifted by one column).
//   * LayerNorm is applied to the ACTIVATION tile (C values per pixel, 5.3x fewer than the 2Hd outputs): four warps
//     compute the per-pixel statistics from the TMA-staged tile, rewrite it in place as bf16 xhat = (x-mu)*rstd and set
//     two constant-one K rows whose weight columns hold t = bias + W.beta (hi + lo bf16 parts), so the accumulator IS
//     the conv input: no epilogue arithmetic on the 2Hd-channel tensor at all.  Rows / columns outside the image are
//     zero in x (TMA out-of-bounds fill) and get a zero in the constant rows: exactly the zero padding of the reference's
//     depthwise conv, which pads the OUTPUT of the 1x1 conv.
//   * a CTA walks a 64-pixel-wide column strip of one image downwards, one image row per MMA group (N = 64); TMEM holds
//     a ring of 4 row slots per accumulator (A = channels c, B = channels c + Hd of the same gate pairs: 2 x 256 columns),
//     so the MMAs of row r+2 run while the conv warps read rows r-1, r, r+1.
// Work item = (sample, band of RB rows, strip, channel block of 128 pairs); strips are 64 loaded / 48 written columns
// (three 16-pixel windows, one per conv warp of a lane quadrant; 16-pixel alignment makes the output a TMA store).
// Items are dealt to the CTAs round-robin with the channel block fastest and the strip next, so the CTAs running at the
// same time read neighbouring 128-byte pieces of the same image rows (DRAM page locality: a CTA walking alone through a
// strip fetches one 128-byte piece per channel plane per row, ~1.6 TB/s measured) and the second channel block of a
// (band, strip) finds the activations in L2.  148 % ncb == 0 keeps one weight block resident per CTA.
#include "dwconv_tma.cuh"

namespace b200ir {

struct FusedParams {
  bf16* y;
  int64_t y_bs;
  const float* dw_w;   // [Ctot][9]
  const float* dw_b;   // [Ctot] or NULL
  int ln_type, C, Cp, Cout, H, W, B;
  int KS, Kbox, kf64;
  int ncb, nstrips, nbands, RB, total_items, S;
  float inv_c;
};

#ifndef B200IR_FU_NCW
#define B200IR_FU_NCW 12
#endif
// timing experiments (-DB200IR_FU_DBG=<bits> A/B builds only, results WRONG): 1 read each TMEM row once (no re-read for the
// other two tap rows), 2 no conv arithmetic, 4 no LayerNorm work, 8 no output stores, 16 no MMAs, 32 no activation loads,
// 64 one TMEM read per row but the FMAs of all three tap rows
#ifndef B200IR_FU_DBG
#define B200IR_FU_DBG 0
#endif
#ifdef B200IR_TIMELINE
// debug builds only: clock64 timeline of CTA 0, one private region per role (producer, LN warp 0, MMA, conv warp 0)
__device__ unsigned long long g_ftl[4 * 4096 * 2];
#define FTL(slot, tag, id)                                                                            \
  do {                                                                                                \
    if (blockIdx.x == 0 && (threadIdx.x & 31) == 0 && tl_cnt < 4096) {                                \
      g_ftl[((slot) * 4096 + tl_cnt) * 2] = ((unsigned long long)(tag) << 32) | (unsigned)(id);       \
      g_ftl[((slot) * 4096 + tl_cnt) * 2 + 1] = (unsigned long long)clock64();                        \
      ++tl_cnt;                                                                                       \
    }                                                                                                 \
  } while (0)
#define FTL_DECL int tl_cnt = 0;
#else
#define FTL(slot, tag, id)
#define FTL_DECL
#endif
constexpr int kFuStripStep = 48;   // image columns between strips; strip k loads [48k-8, 48k+56), writes [48k, 48k+48)
constexpr int kFuLnWarps = 6;     // three teams of two warps; a team normalises every third activation row
constexpr int kFuLnTeams = kFuLnWarps / 2;

__device__ __forceinline__ uint64_t globaltimer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
// mbarrier wait bounded in TIME (2 s): a protocol bug traps instead of hanging the GPU box
__device__ __forceinline__ void mbar_wait_to(uint64_t* bar, uint32_t parity) {
  using namespace tc;
  if (mbar_try_wait(bar, parity)) return;
  const uint64_t t0 = globaltimer_ns();
  while (!mbar_try_wait(bar, parity)) {
    if (globaltimer_ns() - t0 > 2000000000ull) {
      printf("b200ir fused: mbarrier timeout (block %d warp %d lane %d bar %p parity %u)\n", blockIdx.x, threadIdx.x >> 5,
             threadIdx.x & 31, (void*)bar, parity);
      __trap();
    }
  }
}

__device__ __forceinline__ void tmem_ld_32x2(uint32_t taddr, uint32_t* r) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0, %1}, [%2];" : "=r"(r[0]), "=r"(r[1]) : "r"(taddr) : "memory");
}

__device__ __forceinline__ float2 f2(uint32_t a, uint32_t b) { return make_float2(__uint_as_float(a), __uint_as_float(b)); }
__device__ __forceinline__ float2 fmul2(float2 a, float2 b) {
  uint64_t d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(*reinterpret_cast<uint64_t*>(&a)), "l"(*reinterpret_cast<uint64_t*>(&b)));
  return *reinterpret_cast<float2*>(&d);
}
__device__ __forceinline__ float2 fadd2(float2 a, float2 b) {
  uint64_t d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(*reinterpret_cast<uint64_t*>(&a)), "l"(*reinterpret_cast<uint64_t*>(&b)));
  return *reinterpret_cast<float2*>(&d);
}
// a pixel pair stored as bf16x2 -> (lo, hi) fp32
__device__ __forceinline__ float2 bf2_unpack(uint32_t w) { return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u)); }
__device__ __forceinline__ void lds128(uint32_t addr, uint32_t* w) {
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "r"(addr));
}
// 2*gelu on a pixel pair (same fit as gelu2_fast, packed fp32x2 arithmetic): x + x*tanh(x*q(min(x^2,64)))
__device__ __forceinline__ float2 gelu2_fast2(float2 x) {
  float2 x2 = fmul2(x, x);
  x2.x = fminf(x2.x, 64.f);
  x2.y = fminf(x2.y, 64.f);
  float2 t = ffma2(make_float2(-3.5151753375e-4f, -3.5151753375e-4f), x2, make_float2(3.7005650955e-2f, 3.7005650955e-2f));
  t = ffma2(t, x2, make_float2(0.79750787845f, 0.79750787845f));
  const float2 u = fmul2(t, x);
  float2 th;
  asm("tanh.approx.f32 %0, %1;" : "=f"(th.x) : "f"(u.x));
  asm("tanh.approx.f32 %0, %1;" : "=f"(th.y) : "f"(u.y));
  return ffma2(x, th, x);
}

struct FuItem {
  int cb, b, y0, nr, xs, lo, hi;
};
__device__ __forceinline__ FuItem fu_decode(int t, const FusedParams& P) {
  FuItem it;
  it.cb = t % P.ncb;
  int r = t / P.ncb;
  const int strip = r % P.nstrips;
  r /= P.nstrips;
  const int band = r % P.nbands;
  it.b = r / P.nbands;
  it.y0 = band * P.RB;
  it.nr = min(P.RB, P.H - it.y0);
  it.xs = strip * kFuStripStep - 8;
  it.lo = it.xs + 8;
  it.hi = min(it.xs + 56, P.W);
  return it;
}

// GATE: out[c] = gelu(dw(t)[c]) * dw(t)[c + Cout], c < Cout (Cout = hidden_features); conv channels = 2*Cout.
// !GATE: out[c] = dw(t)[c], c < Cout; the two accumulators hold two independent 128-channel blocks.
template <bool GATE, int NCW, int LNB>
__global__ void __launch_bounds__((NCW + kFuLnWarps + 2) * 32, 1)
fused_ln_proj_dw_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW,
                        const __grid_constant__ CUtensorMap tmY, const FusedParams P) {
  using namespace tc;
  constexpr int kWarpLn0 = NCW, kWarpProducer = NCW + kFuLnWarps, kWarpMma = NCW + kFuLnWarps + 1;
  static_assert(NCW == 12, "three 16-pixel windows per 48-pixel strip: one conv warp per (lane quadrant, window)");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t w_bytes = (uint32_t)P.kf64 * 256u * 128u;
  const uint32_t stage_bytes = (uint32_t)P.Kbox * 128u;
  uint8_t* Wsm = smem;
  uint8_t* Xsm = smem + w_bytes;
  uint8_t* Ysm = Xsm + (size_t)P.S * stage_bytes;   // [NCW warps][2 buffers][32 channels][32 B] output staging
  __shared__ uint64_t bar_xfull[8], bar_xnorm[8], bar_xempty[8], bar_rowfull[4], bar_rowempty[4], bar_wfull[1], bar_wempty[1];
  __shared__ uint32_t tmem_slot[1];
  // -rstd*mu and rstd of the 8 pixels of a group, per LN warp: two separate arrays so that one LDS.128 returns pixel
  // PAIRS in aligned register pairs (an interleaved (nm, rstd) layout cost ~110 register moves per normalise batch)
  __shared__ __align__(16) float s_nm[kFuLnWarps][4][8], s_rs[kFuLnWarps][4][8];
  __shared__ uint32_t s_adesc[2 * 16];                       // low words of the weight descriptors, (acc, ks)

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int i = 0; i < P.S; ++i) {
      mbar_init(&bar_xfull[i], 1);
      mbar_init(&bar_xnorm[i], 2);   // the two warps of the team that owns the row
      mbar_init(&bar_xempty[i], 1);
    }
    for (int i = 0; i < 4; ++i) {
      mbar_init(&bar_rowfull[i], 1);
      mbar_init(&bar_rowempty[i], NCW);
    }
    mbar_init(bar_wfull, 1);
    mbar_init(bar_wempty, 1);
    fence_barrier_init();
  }
  if (warp == kWarpProducer) {
    if (lane == 0) {
      tma_prefetch_desc(&tmX);
      tma_prefetch_desc(&tmW);
      tma_prefetch_desc(&tmY);
    }
    tmem_alloc(tmem_slot, 512);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int t_begin = (int)blockIdx.x, t_end = P.total_items, t_step = (int)gridDim.x;   // round-robin deal

  if (warp == kWarpProducer) {
    // ================= TMA producer =================
    if (lane == 0) {
      FTL_DECL
      uint32_t it = 0, wload = 0;
      int cur_cb = -1;
      for (int t = t_begin; t < t_end; t += t_step) {
        const FuItem I = fu_decode(t, P);
        if (I.cb != cur_cb) {
          if (wload > 0) mbar_wait_to(bar_wempty, (wload - 1) & 1);
          mbar_arrive_expect_tx(bar_wfull, w_bytes);
          for (int k64 = 0; k64 < P.kf64; ++k64)
            tma_load_3d(Wsm + (size_t)k64 * 32768, &tmW, bar_wfull, k64 * 64, I.cb * 256, 0);
          ++wload;
          cur_cb = I.cb;
        }
        for (int i = 0; i < I.nr + 2; ++i, ++it) {
          const uint32_t stage = it % (uint32_t)P.S, ph = (it / (uint32_t)P.S) & 1;
          mbar_wait_to(&bar_xempty[stage], ph ^ 1);
          FTL(0, 1, it);
          if (B200IR_FU_DBG & 32) { mbar_arrive(&bar_xfull[stage]); continue; }
          mbar_arrive_expect_tx(&bar_xfull[stage], stage_bytes);
          tma_load_4d(Xsm + (size_t)stage * stage_bytes, &tmX, &bar_xfull[stage], I.xs, I.y0 - 1 + i, 0, I.b);
        }
      }
    }
  } else if (warp == kWarpMma) {
    // ================= MMA issuer (whole warp, one elected lane per instruction) =================
    FTL_DECL
    uint32_t it = 0, n = 0, wload = 0;
    int cur_cb = -1;
    const uint32_t x_addr = smem_u32(Xsm), w_addr = smem_u32(Wsm);
    const uint32_t idesc = make_idesc_bf16(128, 64, 0, 1);   // A (weights) K-major, B (activations) MN-major
    const uint32_t desc_hi = (uint32_t)(make_smem_desc_sw128(0u, 16u, 1024u) >> 32);
    // the weight descriptors do not depend on the row: tabulate their low words once ((acc, ks) -> s_adesc), the
    // activation descriptor advances by a constant per K step -> one MMA issue is a table load, an add and the instruction
    for (int e = lane; e < 2 * P.KS; e += 32) {
      const int acc = e / P.KS, ks = e - acc * P.KS, kglob = ks * 16;
      const uint32_t a_addr = w_addr + (uint32_t)(kglob >> 6) * 32768u + (uint32_t)acc * 16384u + (uint32_t)(kglob & 63) * 2u;
      s_adesc[acc * 16 + ks] = ((a_addr >> 4) & 0x3fffu) | (1u << 16);
    }
    __syncwarp();
    for (int t = t_begin; t < t_end; t += t_step) {
      const FuItem I = fu_decode(t, P);
      if (I.cb != cur_cb) {
        if (cur_cb >= 0) umma_commit_elect(bar_wempty);   // every MMA that read the old weights has been issued
        mbar_wait_to(bar_wfull, wload & 1);
        ++wload;
        cur_cb = I.cb;
        tc_fence_after();
      }
      const int nacc = (GATE || (2 * I.cb + 1) * 128 < P.Cout) ? 2 : 1;
      for (int i = 0; i < I.nr + 2; ++i, ++it, ++n) {
        const uint32_t stage = it % (uint32_t)P.S, ph = (it / (uint32_t)P.S) & 1;
        const uint32_t slot = n & 3, rph = (n >> 2) & 1;
        FTL(2, 10, it);
        mbar_wait_to(&bar_xnorm[stage], ph);
        FTL(2, 11, it);
        mbar_wait_to(&bar_rowempty[slot], rph ^ 1);
        FTL(2, 12, it);
        tc_fence_after();
        const uint32_t b_lo0 = (((x_addr + stage * stage_bytes) >> 4) & 0x3fffu) | ((((uint32_t)P.Kbox * 128u) >> 4) << 16);
        for (int acc = 0; acc < ((B200IR_FU_DBG & 16) ? 0 : nacc); ++acc) {
          const uint32_t d_tmem = tmem_base + (uint32_t)acc * 256u + slot * 64u;
          uint32_t b_lo = b_lo0;
          const uint32_t* atab = s_adesc + acc * 16;
          for (int ks = 0; ks < P.KS; ++ks) {
            const uint32_t a_lo = atab[ks];
            umma_bf16_elect(d_tmem, ((uint64_t)desc_hi << 32) | a_lo, ((uint64_t)desc_hi << 32) | b_lo, idesc,
                            ks != 0 ? 1u : 0u);
            b_lo += 2048u >> 4;
          }
          FTL(2, 14, it);
        }
        umma_commit_elect(&bar_xempty[stage]);
        umma_commit_elect(&bar_rowfull[slot]);
        FTL(2, 13, it);
      }
    }
  } else if (warp >= kWarpLn0) {
    // ================= LayerNorm: statistics + in-place normalisation of the staged activation row =================
    // a team = 2 warps = 64 threads per activation row: G = m>>3 = 8-pixel group (16-byte chunk of every 128-byte channel
    // row), sub = m&7 = row subset (channels == sub mod 8); the teams take the rows round-robin, so three rows are being
    // normalised at any time (one warp alone is latency-bound: ~3500 cycles per row measured with one 4-warp team)
    const int lw = warp - kWarpLn0;
    const int team = lw >> 1;
    const int m = (lw & 1) * 32 + lane;
    const uint32_t G = (uint32_t)m >> 3, sub = (uint32_t)m & 7u;
    const uint32_t x_addr = smem_u32(Xsm);
    const bool withbias = P.ln_type == B200IR_LN_WITHBIAS;
    FTL_DECL
    uint32_t it = 0;
    for (int t = t_begin; t < t_end; t += t_step) {
      const FuItem I = fu_decode(t, P);
      const int gx0 = I.xs + (int)G * 8;
      for (int i = 0; i < I.nr + 2; ++i, ++it) {
        const uint32_t stage = it % (uint32_t)P.S, ph = (it / (uint32_t)P.S) & 1;
        if ((int)(it % (uint32_t)kFuLnTeams) != team) continue;
        const int yy = I.y0 - 1 + i;
        const bool inimg = yy >= 0 && yy < P.H && gx0 >= 0 && gx0 < P.W;
        if (lw == 0) FTL(1, 20, it);
        mbar_wait_to(&bar_xfull[stage], ph);
        if (lw == 0) FTL(1, 21, it);
        if (B200IR_FU_DBG & 4) {
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) mbar_arrive(&bar_xnorm[stage]);
          continue;
        }
        const uint32_t tile = x_addr + stage * stage_bytes;
        // the thread's rows are channels sub, sub+8, ...: r & 7 == sub, so every row sits at base + q * 1024 (8 rows of
        // 128 B) in the same swizzled 16-byte chunk; shifted sums in packed fp32x2 over pixel pairs
        const uint32_t base = tile + sub * 128u + ((G ^ sub) << 4);
        uint32_t w0[4];
        lds128(tile + (G << 4), w0);           // row 0 (chunk G ^ 0): the shift of the one-trip variance
        float2 c0[4], sm[4], sq[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          c0[j] = bf2_unpack(w0[j]);
          c0[j].x = -c0[j].x; c0[j].y = -c0[j].y;
          sm[j] = sq[j] = make_float2(0.f, 0.f);
        }
        const int nrow = P.C >> 3;             // C % (8 * LNB) == 0 (launcher): no row predicates
        for (int qb = 0; qb < nrow; qb += LNB) {
          uint32_t xw[LNB][4];
#pragma unroll
          for (int q = 0; q < LNB; ++q) lds128(base + (uint32_t)(qb + q) * 1024u, xw[q]);
#pragma unroll
          for (int q = 0; q < LNB; ++q) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const float2 d = fadd2(bf2_unpack(xw[q][j]), c0[j]);
              sm[j] = fadd2(sm[j], d);
              sq[j] = ffma2(d, d, sq[j]);
            }
          }
        }
        if (lw == 0) FTL(1, 23, it);
        float x0[8], sum[8], sqv[8];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          x0[2 * j] = -c0[j].x; x0[2 * j + 1] = -c0[j].y;
          sum[2 * j] = sm[j].x; sum[2 * j + 1] = sm[j].y;
          sqv[2 * j] = sq[j].x; sqv[2 * j + 1] = sq[j].y;
        }
        // reduce-scatter over the 8 row subsets (lane bits 2..0): 8 -> 4 -> 2 -> 1 pixels kept; lane sub ends with pixel sub
        const bool h3 = (sub & 4u) != 0, h2 = (sub & 2u) != 0, h1 = (sub & 1u) != 0;
        float s4[4], q4[4], c4[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float ks = h3 ? sum[4 + j] : sum[j], ss = h3 ? sum[j] : sum[4 + j];
          const float kq = h3 ? sqv[4 + j] : sqv[j], sq_ = h3 ? sqv[j] : sqv[4 + j];
          s4[j] = ks + __shfl_xor_sync(0xffffffffu, ss, 4);
          q4[j] = kq + __shfl_xor_sync(0xffffffffu, sq_, 4);
          c4[j] = h3 ? x0[4 + j] : x0[j];
        }
        float s2[2], q2[2], c2[2];
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const float ks = h2 ? s4[2 + j] : s4[j], ss = h2 ? s4[j] : s4[2 + j];
          const float kq = h2 ? q4[2 + j] : q4[j], sq_ = h2 ? q4[j] : q4[2 + j];
          s2[j] = ks + __shfl_xor_sync(0xffffffffu, ss, 2);
          q2[j] = kq + __shfl_xor_sync(0xffffffffu, sq_, 2);
          c2[j] = h2 ? c4[2 + j] : c4[j];
        }
        const float s1 = (h1 ? s2[1] : s2[0]) + __shfl_xor_sync(0xffffffffu, h1 ? s2[0] : s2[1], 1);
        const float q1 = (h1 ? q2[1] : q2[0]) + __shfl_xor_sync(0xffffffffu, h1 ? q2[0] : q2[1], 1);
        const float c1 = h1 ? c2[1] : c2[0];
        const float md = s1 * P.inv_c;
        const float var = fmaxf(q1 * P.inv_c - md * md, 0.f);
        const float rstd = rsqrtf(var + 1e-5f);
        const float nmv = withbias ? -rstd * (c1 + md) : 0.f;
        s_nm[lw][lane >> 3][sub] = nmv;
        s_rs[lw][lane >> 3][sub] = rstd;
        __syncwarp();
        if (lw == 0) FTL(1, 24, it);
        float2 nm2[4], rs2[4];
        {
          const float4* pn = reinterpret_cast<const float4*>(&s_nm[lw][lane >> 3][0]);
          const float4* pr = reinterpret_cast<const float4*>(&s_rs[lw][lane >> 3][0]);
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            const float4 u = pn[j], v = pr[j];
            nm2[2 * j] = make_float2(u.x, u.y);
            nm2[2 * j + 1] = make_float2(u.z, u.w);
            rs2[2 * j] = make_float2(v.x, v.y);
            rs2[2 * j + 1] = make_float2(v.z, v.w);
          }
        }
        for (int qb = 0; qb < nrow; qb += LNB) {
          uint32_t xw[LNB][4];
#pragma unroll
          for (int q = 0; q < LNB; ++q) lds128(base + (uint32_t)(qb + q) * 1024u, xw[q]);
#pragma unroll
          for (int q = 0; q < LNB; ++q) {
            uint32_t pk[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const float2 o = ffma2(bf2_unpack(xw[q][j]), rs2[j], nm2[j]);
              __nv_bfloat162 h = __floats2bfloat162_rn(o.x, o.y);
              pk[j] = *reinterpret_cast<uint32_t*>(&h);
            }
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(base + (uint32_t)(qb + q) * 1024u), "r"(pk[0]),
                         "r"(pk[1]), "r"(pk[2]), "r"(pk[3])
                         : "memory");
          }
        }
        if (lw == 0) FTL(1, 25, it);
        if (sub == 0) {   // the two constant-one K rows (weight columns Cp, Cp+1 hold t = bias + W.beta, hi + lo parts)
          const uint32_t one = inimg ? 0x3f803f80u : 0u;
#pragma unroll
          for (int rr = 0; rr < 2; ++rr) {
            const uint32_t r = (uint32_t)P.Cp + rr;
            asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};" ::"r"(tile + r * 128u + ((G ^ (r & 7u)) << 4)), "r"(one)
                         : "memory");
          }
        }
        fence_proxy_async();
        if (lw == 0) FTL(1, 26, it);
        __syncwarp();
        if (lane == 0) mbar_arrive(&bar_xnorm[stage]);
        if (lw == 0) FTL(1, 22, it);
      }
    }
  } else {
    // ================= depthwise 3x3 (+ gate) straight out of TMEM: thread = channel (pair), warp = 16-pixel window ===
    const int quad = warp & 3, jpart = warp >> 2;
    const uint32_t tlane = tmem_base + ((uint32_t)(quad * 32) << 16);
    uint8_t* stg = Ysm + (size_t)warp * 2048;   // two 1 KB buffers [32 channels][16 px] for the TMA stores
    FTL_DECL
    uint32_t n = 0, st_it = 0;
    int cur_cb = -1;
    float2 wA[9], wB[9];
    float bA = 0.f, bB = 0.f;
    const int e0 = 8 + 16 * jpart;              // strip-local column of the window (image column xs + e0, 16-aligned)
    for (int t = t_begin; t < t_end; t += t_step) {
      const FuItem I = fu_decode(t, P);
      if (I.cb != cur_cb) {
        cur_cb = I.cb;
        int cA, cB;
        bool okA, okB;
        if (GATE) {
          cA = I.cb * 128 + quad * 32 + lane;
          cB = P.Cout + cA;
          okA = okB = cA < P.Cout;
        } else {
          cA = (2 * I.cb) * 128 + quad * 32 + lane;
          cB = cA + 128;
          okA = cA < P.Cout;
          okB = cB < P.Cout;
        }
        const float kh = GATE ? 0.5f : 1.0f;   // the 0.5 of gelu(x) = 0.5 x (1 + tanh(..)) rides on the gate channel's taps
#pragma unroll
        for (int i = 0; i < 9; ++i) {
          const float a = okA ? __ldg(P.dw_w + (int64_t)cA * 9 + i) : 0.f;
          const float b = okB ? kh * __ldg(P.dw_w + (int64_t)cB * 9 + i) : 0.f;
          wA[i] = make_float2(a, a);
          wB[i] = make_float2(b, b);
        }
        bA = (okA && P.dw_b) ? __ldg(P.dw_b + cA) : 0.f;
        bB = (okB && P.dw_b) ? kh * __ldg(P.dw_b + cB) : 0.f;
      }
      const bool haveB = GATE || (2 * I.cb + 1) * 128 < P.Cout;   // warp-uniform
      const int x0 = I.xs + e0;
      const bool active = x0 < P.W;                               // warp-uniform: the window lies inside the image
      const int chA = GATE ? I.cb * 128 + quad * 32 : (2 * I.cb) * 128 + quad * 32;   // first channel of this warp's rows
      for (int i = 0; i < I.nr; ++i) {
        const uint32_t nn = n + (uint32_t)i + 2u;
        if (warp == 0) FTL(3, 30, nn);
        mbar_wait_to(&bar_rowfull[nn & 3], (nn >> 2) & 1);
        if (warp == 0) FTL(3, 31, nn);
        tc_fence_after();
        if (active && chA < P.Cout) {
          float2 accA[8], accB[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            accA[k] = make_float2(bA, bA);
            accB[k] = make_float2(bB, bB);
          }
#pragma unroll
          for (int r = 0; r < ((B200IR_FU_DBG & 1) ? 1 : 3); ++r) {
            const uint32_t slot = (n + (uint32_t)i + (uint32_t)r) & 3u;
            const uint32_t ta = tlane + slot * 64u + (uint32_t)e0;
            if ((B200IR_FU_DBG & 64) && r > 0) {   // experiment: arithmetic of all three tap rows, TMEM read of one
#pragma unroll
              for (int k = 0; k < 8; ++k) {
                accA[k] = ffma2(wA[r * 3 + 0], accB[(k + 1) & 7], accA[k]);
                accA[k] = ffma2(wA[r * 3 + 2], accB[(k + 2) & 7], accA[k]);
                accA[k].x = fmaf(wA[r * 3 + 1].x, accB[(k + 3) & 7].x, accA[k].x);
                accA[k].y = fmaf(wA[r * 3 + 1].x, accB[(k + 3) & 7].y, accA[k].y);
                accB[k] = ffma2(wB[r * 3 + 0], accA[(k + 1) & 7], accB[k]);
                accB[k] = ffma2(wB[r * 3 + 2], accA[(k + 2) & 7], accB[k]);
                accB[k].x = fmaf(wB[r * 3 + 1].x, accA[(k + 3) & 7].x, accB[k].x);
                accB[k].y = fmaf(wB[r * 3 + 1].x, accA[(k + 3) & 7].y, accB[k].y);
              }
              continue;
            }
            {
              // 18 columns e0-1 .. e0+16: the +-1 taps read even-aligned register pairs (packed FMA); the centre tap is
              // odd-aligned -> scalar FMAs on the same registers (a second, shifted tcgen05.ld doubled the TMEM traffic,
              // which the MMAs' accumulator updates compete with)
              uint32_t O[18];
              tmem_ld_32x16(ta - 1, O);
              tmem_ld_32x2(ta + 15, O + 16);
              tmem_ld_wait();
              if (B200IR_FU_DBG & 2) { accA[0].x += __uint_as_float(O[0] ^ O[17] ^ O[3]); } else
#pragma unroll
              for (int k = 0; k < 8; ++k) {
                accA[k] = ffma2(wA[r * 3 + 0], f2(O[2 * k], O[2 * k + 1]), accA[k]);
                accA[k] = ffma2(wA[r * 3 + 2], f2(O[2 * k + 2], O[2 * k + 3]), accA[k]);
#ifndef B200IR_FU_CENTER2
                accA[k].x = fmaf(wA[r * 3 + 1].x, __uint_as_float(O[2 * k + 1]), accA[k].x);
                accA[k].y = fmaf(wA[r * 3 + 1].x, __uint_as_float(O[2 * k + 2]), accA[k].y);
#endif
              }
#ifdef B200IR_FU_CENTER2
              {   // A/B: centre tap packed too, from a second (even-aligned) tcgen05.ld of the same 16 columns
                uint32_t Q[16];
                tmem_ld_32x16(ta, Q);
                tmem_ld_wait();
#pragma unroll
                for (int k = 0; k < 8; ++k) accA[k] = ffma2(wA[r * 3 + 1], f2(Q[2 * k], Q[2 * k + 1]), accA[k]);
              }
#endif
            }
            if (haveB) {
              uint32_t O[18];
              tmem_ld_32x16(ta + 256u - 1, O);
              tmem_ld_32x2(ta + 256u + 15, O + 16);
              tmem_ld_wait();
              if (B200IR_FU_DBG & 2) { accB[0].x += __uint_as_float(O[0] ^ O[17] ^ O[3]); } else
#pragma unroll
              for (int k = 0; k < 8; ++k) {
                accB[k] = ffma2(wB[r * 3 + 0], f2(O[2 * k], O[2 * k + 1]), accB[k]);
                accB[k] = ffma2(wB[r * 3 + 2], f2(O[2 * k + 2], O[2 * k + 3]), accB[k]);
#ifndef B200IR_FU_CENTER2
                accB[k].x = fmaf(wB[r * 3 + 1].x, __uint_as_float(O[2 * k + 1]), accB[k].x);
                accB[k].y = fmaf(wB[r * 3 + 1].x, __uint_as_float(O[2 * k + 2]), accB[k].y);
#endif
              }
#ifdef B200IR_FU_CENTER2
              {   // A/B: centre tap packed too, from a second (even-aligned) tcgen05.ld of the same 16 columns
                uint32_t Q[16];
                tmem_ld_32x16(ta + 256u, Q);
                tmem_ld_wait();
#pragma unroll
                for (int k = 0; k < 8; ++k) accB[k] = ffma2(wB[r * 3 + 1], f2(Q[2 * k], Q[2 * k + 1]), accB[k]);
              }
#endif
            }
          }
          // results -> bf16 -> this warp's staging tile [32 channels][16 px] -> ONE TMA store box {16 px, 1 row, 32 ch}
          // (channels >= Cout and pixels >= W are clipped by TMA)
          const int npass = GATE ? 1 : (haveB ? 2 : 1);
          for (int pass = 0; pass < npass; ++pass) {
            uint32_t pk[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              __nv_bfloat162 h;
              if (GATE) {
                const float2 gv = fmul2(gelu2_fast2(accA[k]), accB[k]);
                h = __floats2bfloat162_rn(gv.x, gv.y);
              }
              else h = pass == 0 ? __floats2bfloat162_rn(accA[k].x, accA[k].y) : __floats2bfloat162_rn(accB[k].x, accB[k].y);
              pk[k] = *reinterpret_cast<uint32_t*>(&h);
            }
            uint8_t* sbuf = stg + (st_it & 1) * 1024;
            if (lane == 0) bulk_wait_read<1>();   // the store that read this buffer two stores ago is done with it
            __syncwarp();
            const uint32_t sa = smem_u32(sbuf) + (uint32_t)lane * 32u;
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(sa), "r"(pk[0]), "r"(pk[1]), "r"(pk[2]), "r"(pk[3]) : "memory");
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(sa + 16u), "r"(pk[4]), "r"(pk[5]), "r"(pk[6]), "r"(pk[7]) : "memory");
            fence_proxy_async();
            __syncwarp();
            if (lane == 0 && !(B200IR_FU_DBG & 8)) {
              asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.tile.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(&tmY),
                           "r"(smem_u32(sbuf)), "r"(x0), "r"(I.y0 + i), "r"(chA + pass * 128), "r"(I.b)
                           : "memory");
              bulk_commit();
            }
            ++st_it;
          }
        }
        if (warp == 0) FTL(3, 32, nn);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bar_rowempty[(n + (uint32_t)i) & 3]);   // input row i of the band is no longer needed
      }
      __syncwarp();
      if (lane == 0) {   // the last two input rows of the band
        mbar_arrive(&bar_rowempty[(n + (uint32_t)I.nr) & 3]);
        mbar_arrive(&bar_rowempty[(n + (uint32_t)I.nr + 1u) & 3]);
      }
      n += (uint32_t)I.nr + 2u;
    }
    if (lane == 0) bulk_wait<0>();   // all output stores have left shared memory and are complete
  }

  tc_fence_before();
  __syncthreads();
  if (warp == kWarpProducer) tmem_dealloc(tmem_base, 512);
}

// ---------------------------------------------------------------------------------------------
// weight packer: wf[(cb*256 + acc*128 + l)][Kf] bf16, K-major; columns [0,C) = W*gamma, Cp / Cp+1 = hi / lo parts of
// t = bias + W.beta.  One warp per row.
// ---------------------------------------------------------------------------------------------
__global__ void pack_fused_kernel(const float* __restrict__ w, const float* __restrict__ bias, int ln_type,
                                  const float* __restrict__ ln_w, const float* __restrict__ ln_b, int C, int Cp, int Cout,
                                  int gate, int Kf, int nrows, bf16* __restrict__ wf) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= nrows) return;
  const int cb = row >> 8, acc = (row >> 7) & 1, l = row & 127;
  int ch;
  bool ok;
  if (gate) {
    const int c = cb * 128 + l;
    ok = c < Cout;
    ch = acc ? Cout + c : c;
  } else {
    ch = (2 * cb + acc) * 128 + l;
    ok = ch < Cout;
  }
  bf16* dst = wf + (size_t)row * Kf;
  float t = 0.f;
  for (int k = lane; k < Kf; k += 32) {
    float v = 0.f;
    if (ok && k < C) {
      const float wv = w[(size_t)ch * C + k];
      v = ln_type != B200IR_LN_NONE ? wv * ln_w[k] : wv;
      if (ln_type == B200IR_LN_WITHBIAS) t = fmaf(wv, ln_b[k], t);
    }
    if (k != Cp && k != Cp + 1) dst[k] = __float2bfloat16_rn(v);
  }
  t = warp_sum(t);
  if (lane == 0) {
    const float tt = ok ? t + (bias ? bias[ch] : 0.f) : 0.f;
    const bf16 hi = __float2bfloat16_rn(tt);
    dst[Cp] = hi;
    dst[Cp + 1] = __float2bfloat16_rn(tt - __bfloat162float(hi));
  }
}

struct FusedGeom {
  int Cp, Kbox, KS, Kf, kf64, ncb, nrows, S;
  size_t smem;
};

static bool fused_geometry(int C, int Cout, int gate, FusedGeom* g) {
  if (C < 16 || C % 16 || Cout < 1) return false;   // LayerNorm rows are dealt 8 at a time, K steps are 16 channels
  g->Cp = C;
  g->Kbox = g->Cp + 16;
  if (g->Kbox > 256) return false;
  g->KS = g->Kbox / 16;
  g->Kf = (g->Kbox + 63) / 64 * 64;
  g->kf64 = g->Kf / 64;
  g->ncb = gate ? (Cout + 127) / 128 : (Cout + 255) / 256;
  g->nrows = g->ncb * 256;
  const size_t wbytes = (size_t)g->kf64 * 32768, stage = (size_t)g->Kbox * 128;
  const size_t ybytes = (size_t)B200IR_FU_NCW * 2048;   // output staging of the conv warps
  const long avail = 224 * 1024 - 1024 - (long)wbytes - (long)ybytes;
  int S = avail > 0 ? (int)(avail / (long)stage) : 0;
  if (S > 6) S = 6;
  if (S < 3) return false;
  g->S = S;
  g->smem = wbytes + (size_t)S * stage + ybytes + 1024;
  return true;
}

}  // namespace b200ir

using namespace b200ir;

#ifdef B200IR_TIMELINE
extern "C" int b200ir_debug_fused_timeline(unsigned long long* out, int max_pairs) {
  cudaDeviceSynchronize();
  const int n = 4 * 4096;
  if (out && max_pairs >= n) cudaMemcpyFromSymbol(out, b200ir::g_ftl, (size_t)n * 16);
  void* p = nullptr;
  cudaGetSymbolAddress(&p, b200ir::g_ftl);
  cudaMemset(p, 0, (size_t)n * 16);
  return n;
}
#endif

extern "C" int b200ir_fused_proj_dw_geometry(int C, int Cout, int gate, int* kf, int* nrows) {
  FusedGeom g;
  if (!fused_geometry(C, Cout, gate, &g)) {
    set_error("fused_proj_dw: unsupported C=%d Cout=%d", C, Cout);
    return B200IR_E_UNSUPPORTED;
  }
  if (kf) *kf = g.Kf;
  if (nrows) *nrows = g.nrows;
  return 0;
}

extern "C" int b200ir_pack_fused_proj_dw(const float* w, const float* bias, int ln_type, const float* ln_w,
                                         const float* ln_b, int C, int Cout, int gate, void* wf,
                                         b200ir_stream_t stream) {
  B200IR_CHECK_ARG(w && wf, "pack_fused_proj_dw: bad args");
  B200IR_CHECK_ARG(ln_type == B200IR_LN_NONE || ln_w, "pack_fused_proj_dw: ln_w missing");
  B200IR_CHECK_ARG(ln_type != B200IR_LN_WITHBIAS || ln_b, "pack_fused_proj_dw: ln_b missing");
  FusedGeom g;
  B200IR_UNSUPPORTED(fused_geometry(C, Cout, gate, &g), "pack_fused_proj_dw: unsupported C=%d Cout=%d", C, Cout);
  pack_fused_kernel<<<(g.nrows + 7) / 8, 256, 0, (cudaStream_t)stream>>>(w, bias, ln_type, ln_w, ln_b, C, g.Cp, Cout,
                                                                         gate ? 1 : 0, g.Kf, g.nrows, (bf16*)wf);
  return check_launch("pack_fused_proj_dw");
}

extern "C" int b200ir_ln_proj_dw_supported(int C, int Cout, int H, int W, int gate) {
  FusedGeom g;
  return (fused_geometry(C, Cout, gate, &g) && W % 8 == 0 && W >= 16 && H >= 1) ? 1 : 0;
}

template <bool GATE>
static int fused_launch(const void* x, int64_t x_bs, const void* wf, const float* dw_w, const float* dw_b, int ln_type,
                        void* y, int64_t y_bs, int B, int C, int Cout, int H, int W, cudaStream_t st) {
  constexpr int NCW = B200IR_FU_NCW;   // conv warps (8: 128 registers per thread, no spills; 16 spills ~55 registers)
  FusedGeom g;
  B200IR_UNSUPPORTED(fused_geometry(C, Cout, GATE ? 1 : 0, &g), "ln_proj_dw: unsupported C=%d Cout=%d", C, Cout);
  FusedParams P;
  P.y = (bf16*)y; P.y_bs = y_bs; P.dw_w = dw_w; P.dw_b = dw_b;
  P.ln_type = ln_type; P.C = C; P.Cp = g.Cp; P.Cout = Cout; P.H = H; P.W = W; P.B = B;
  P.KS = g.KS; P.Kbox = g.Kbox; P.kf64 = g.kf64; P.ncb = g.ncb; P.S = g.S;
  P.inv_c = 1.0f / (float)C;
  P.nstrips = (W + kFuStripStep - 1) / kFuStripStep;
  const int sms = num_sms();
  // band height: fewest (rows incl. the 2 halo rows) on the busiest CTA of the static contiguous split
  int best_rb = 8;
  double best_cost = 1e30;
  for (int rb = 8; rb <= 64; rb *= 2) {
    const int nb = (H + rb - 1) / rb;
    const int64_t items = (int64_t)g.ncb * B * nb * P.nstrips;
    const int64_t grid = items < sms ? items : sms;
    const double per = (double)((items + grid - 1) / grid) * (double)((rb < H ? rb : H) + 2);
    if (per < best_cost - 1e-9) best_cost = per, best_rb = rb;
  }
  P.RB = best_rb;
  P.nbands = (H + P.RB - 1) / P.RB;
  const int64_t items = (int64_t)g.ncb * B * P.nbands * P.nstrips;
  B200IR_UNSUPPORTED(items < (1ll << 30), "ln_proj_dw: too many work items");
  P.total_items = (int)items;
  CUtensorMap tmX, tmW;
  {
    const uint64_t dims[4] = {(uint64_t)W, (uint64_t)H, (uint64_t)C, (uint64_t)B};
    const uint64_t str[4] = {2, (uint64_t)W * 2, (uint64_t)H * W * 2, (uint64_t)x_bs * 2};
    const uint32_t box[4] = {64, 1, (uint32_t)g.Kbox, 1};
    int rc = make_tmap_bf16(&tmX, x, 4, dims, str, box);
    if (rc) return rc;
  }
  {
    const uint64_t dims[3] = {(uint64_t)g.Kf, (uint64_t)g.nrows, 1};
    const uint64_t str[3] = {2, (uint64_t)g.Kf * 2, (uint64_t)g.nrows * g.Kf * 2};
    const uint32_t box[3] = {64, 256, 1};
    int rc = make_tmap_bf16(&tmW, wf, 3, dims, str, box);
    if (rc) return rc;
  }
  CUtensorMap tmY;
  {
    const uint64_t dims[4] = {(uint64_t)W, (uint64_t)H, (uint64_t)Cout, (uint64_t)B};
    const uint64_t str[4] = {2, (uint64_t)W * 2, (uint64_t)H * W * 2, (uint64_t)y_bs * 2};
    const uint32_t box[4] = {16, 1, 32, 1};
    int rc = make_tmap_bf16(&tmY, y, 4, dims, str, box, false);
    if (rc) return rc;
  }
  int grid = sms < P.total_items ? sms : P.total_items;
  // LayerNorm rows per thread = C / 8, processed in batches of LNB loads (compile-time: no row predicates)
  const int nrow = C / 8;
#define B200IR_FU_LAUNCH(LNB_)                                                                              \
  do {                                                                                                      \
    ensure_dyn_smem((const void*)fused_ln_proj_dw_kernel<GATE, NCW, LNB_>, 224 * 1024); /* + 2 KB static */  \
    fused_ln_proj_dw_kernel<GATE, NCW, LNB_><<<grid, (NCW + kFuLnWarps + 2) * 32, g.smem, st>>>(tmX, tmW, tmY, P); \
  } while (0)
  if (nrow % 6 == 0) B200IR_FU_LAUNCH(6);
  else if (nrow % 4 == 0) B200IR_FU_LAUNCH(4);
  else B200IR_FU_LAUNCH(2);
#undef B200IR_FU_LAUNCH
  return check_launch(GATE ? "ln_proj_dwgate_fused" : "ln_proj_dw_fused");
}

extern "C" int b200ir_ln_proj_dw_fwd(const void* x, int64_t x_bstride, const void* wf, const float* dw_w,
                                     const float* dw_b, int ln_type, void* y, int64_t y_bstride, int B, int C, int Cout,
                                     int H, int W, int gate, b200ir_stream_t stream) {
  B200IR_CHECK_ARG(x && wf && dw_w && y && B > 0 && C > 0 && Cout > 0 && H > 0 && W > 0, "ln_proj_dw: bad args");
  B200IR_CHECK_ARG(ln_type == B200IR_LN_WITHBIAS || ln_type == B200IR_LN_BIASFREE, "ln_proj_dw: bad ln_type %d", ln_type);
  B200IR_UNSUPPORTED(W % 8 == 0 && W >= 16 && x_bstride % 8 == 0 && y_bstride % 8 == 0 && ((uintptr_t)x % 16) == 0 &&
                         ((uintptr_t)y % 16) == 0,
                     "ln_proj_dw: needs W%%8==0, W>=16, 16B-aligned bf16 activations");
  if (gate)
    return fused_launch<true>(x, x_bstride, wf, dw_w, dw_b, ln_type, y, y_bstride, B, C, Cout, H, W, (cudaStream_t)stream);
  return fused_launch<false>(x, x_bstride, wf, dw_w, dw_b, ln_type, y, y_bstride, B, C, Cout, H, W, (cudaStream_t)stream);
}
