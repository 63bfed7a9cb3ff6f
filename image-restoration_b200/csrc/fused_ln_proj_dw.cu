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
// Work item = (channel block of 128 pairs, sample, band of RB rows, strip); strips are 64 loaded / <= 56 written columns.
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
#define B200IR_FU_NCW 8
#endif
constexpr int kFuStripStep = 56;   // image columns between strips; strip k loads [56k-8, 56k+56), writes [56k-4, 56k+52)
constexpr int kFuLnWarps = 4;

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

struct FuItem {
  int cb, b, y0, nr, xs, lo, hi;
};
__device__ __forceinline__ FuItem fu_decode(int t, const FusedParams& P) {
  FuItem it;
  const int per_b = P.nbands * P.nstrips, per_cb = P.B * per_b;
  it.cb = t / per_cb;
  int r = t - it.cb * per_cb;
  it.b = r / per_b;
  r -= it.b * per_b;
  const int band = r / P.nstrips, strip = r - band * P.nstrips;
  it.y0 = band * P.RB;
  it.nr = min(P.RB, P.H - it.y0);
  it.xs = strip * kFuStripStep - 8;
  it.lo = max(it.xs + 4, 0);
  it.hi = min(it.xs + 60, P.W);
  return it;
}

// GATE: out[c] = gelu(dw(t)[c]) * dw(t)[c + Cout], c < Cout (Cout = hidden_features); conv channels = 2*Cout.
// !GATE: out[c] = dw(t)[c], c < Cout; the two accumulators hold two independent 128-channel blocks.
template <bool GATE, int NCW>
__global__ void __launch_bounds__((NCW + kFuLnWarps + 2) * 32, 1)
fused_ln_proj_dw_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW,
                        const FusedParams P) {
  using namespace tc;
  constexpr int kWarpLn0 = NCW, kWarpProducer = NCW + kFuLnWarps, kWarpMma = NCW + kFuLnWarps + 1;
  constexpr int NJ = NCW / 4;   // column parts per TMEM lane quadrant
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t w_bytes = (uint32_t)P.kf64 * 256u * 128u;
  const uint32_t stage_bytes = (uint32_t)P.Kbox * 128u;
  uint8_t* Wsm = smem;
  uint8_t* Xsm = smem + w_bytes;
  __shared__ uint64_t bar_xfull[8], bar_xnorm[8], bar_xempty[8], bar_rowfull[4], bar_rowempty[4], bar_wfull[1], bar_wempty[1];
  __shared__ uint32_t tmem_slot[1];
  __shared__ __align__(16) float2 s_nr[kFuLnWarps][2][8];   // (-rstd*mu, rstd) of the 8 pixels of a group, per LN warp

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int i = 0; i < P.S; ++i) {
      mbar_init(&bar_xfull[i], 1);
      mbar_init(&bar_xnorm[i], kFuLnWarps);
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
    }
    tmem_alloc(tmem_slot, 512);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int gx = gridDim.x;
  const int t_begin = (int)(((int64_t)P.total_items * blockIdx.x) / gx);
  const int t_end = (int)(((int64_t)P.total_items * (blockIdx.x + 1)) / gx);

  if (warp == kWarpProducer) {
    // ================= TMA producer =================
    if (lane == 0) {
      uint32_t it = 0, wload = 0;
      int cur_cb = -1;
      for (int t = t_begin; t < t_end; ++t) {
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
          mbar_arrive_expect_tx(&bar_xfull[stage], stage_bytes);
          tma_load_4d(Xsm + (size_t)stage * stage_bytes, &tmX, &bar_xfull[stage], I.xs, I.y0 - 1 + i, 0, I.b);
        }
      }
    }
  } else if (warp == kWarpMma) {
    // ================= MMA issuer (whole warp, one elected lane per instruction) =================
    uint32_t it = 0, n = 0, wload = 0;
    int cur_cb = -1;
    const uint32_t x_addr = smem_u32(Xsm), w_addr = smem_u32(Wsm);
    const uint32_t idesc = make_idesc_bf16(128, 64, 0, 1);   // A (weights) K-major, B (activations) MN-major
    const uint32_t desc_hi = (uint32_t)(make_smem_desc_sw128(0u, 16u, 1024u) >> 32);
    for (int t = t_begin; t < t_end; ++t) {
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
        mbar_wait_to(&bar_xnorm[stage], ph);
        mbar_wait_to(&bar_rowempty[slot], rph ^ 1);
        tc_fence_after();
        const uint32_t b_lo0 = (((x_addr + stage * stage_bytes) >> 4) & 0x3fffu) | ((((uint32_t)P.Kbox * 128u) >> 4) << 16);
        for (int acc = 0; acc < nacc; ++acc) {
          const uint32_t d_tmem = tmem_base + (uint32_t)acc * 256u + slot * 64u;
          uint32_t b_lo = b_lo0;
          for (int ks = 0; ks < P.KS; ++ks) {
            const int kglob = ks * 16;
            const uint32_t a_addr = w_addr + (uint32_t)(kglob >> 6) * 32768u + (uint32_t)acc * 16384u + (uint32_t)(kglob & 63) * 2u;
            const uint32_t a_lo = ((a_addr >> 4) & 0x3fffu) | (1u << 16);
            umma_bf16_elect(d_tmem, ((uint64_t)desc_hi << 32) | a_lo, ((uint64_t)desc_hi << 32) | b_lo, idesc,
                            ks != 0 ? 1u : 0u);
            b_lo += 2048u >> 4;
          }
        }
        umma_commit_elect(&bar_xempty[stage]);
        umma_commit_elect(&bar_rowfull[slot]);
      }
    }
  } else if (warp >= kWarpLn0) {
    // ================= LayerNorm: statistics + in-place normalisation of the staged activation row =================
    // 128 threads: G = m>>4 = 8-pixel group (16-byte chunk of every 128-byte channel row), sub = m&15 = row subset
    const int lw = warp - kWarpLn0;
    const int m = lw * 32 + lane;
    const uint32_t G = (uint32_t)m >> 4, sub = (uint32_t)m & 15u;
    const uint32_t x_addr = smem_u32(Xsm);
    const bool withbias = P.ln_type == B200IR_LN_WITHBIAS;
    uint32_t it = 0;
    for (int t = t_begin; t < t_end; ++t) {
      const FuItem I = fu_decode(t, P);
      const int gx0 = I.xs + (int)G * 8;
      for (int i = 0; i < I.nr + 2; ++i, ++it) {
        const uint32_t stage = it % (uint32_t)P.S, ph = (it / (uint32_t)P.S) & 1;
        const int yy = I.y0 - 1 + i;
        const bool inimg = yy >= 0 && yy < P.H && gx0 >= 0 && gx0 < P.W;
        mbar_wait_to(&bar_xfull[stage], ph);
        const uint32_t tile = x_addr + stage * stage_bytes;
        float x0[8], sum[8], sq[8];
        ln_stat_unpack(tile + (G << 4), x0);   // row 0 (chunk G ^ 0): the shift of the one-trip variance
#pragma unroll
        for (int j = 0; j < 8; ++j) sum[j] = sq[j] = 0.f;
        for (int r = (int)sub; r < P.C; r += 16) {
          float v[8];
          ln_stat_unpack(tile + (uint32_t)r * 128u + ((G ^ ((uint32_t)r & 7u)) << 4), v);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float d = v[j] - x0[j];
            sum[j] += d;
            sq[j] = fmaf(d, d, sq[j]);
          }
        }
        // reduce-scatter over the 16 row subsets (lane bits 3..0): 8 -> 4 -> 2 -> 1 pixels kept, then one all-reduce
        const bool h3 = (sub & 8u) != 0, h2 = (sub & 4u) != 0, h1 = (sub & 2u) != 0;
        float s4[4], q4[4], c4[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float ks = h3 ? sum[4 + j] : sum[j], ss = h3 ? sum[j] : sum[4 + j];
          const float kq = h3 ? sq[4 + j] : sq[j], sq_ = h3 ? sq[j] : sq[4 + j];
          s4[j] = ks + __shfl_xor_sync(0xffffffffu, ss, 8);
          q4[j] = kq + __shfl_xor_sync(0xffffffffu, sq_, 8);
          c4[j] = h3 ? x0[4 + j] : x0[j];
        }
        float s2[2], q2[2], c2[2];
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const float ks = h2 ? s4[2 + j] : s4[j], ss = h2 ? s4[j] : s4[2 + j];
          const float kq = h2 ? q4[2 + j] : q4[j], sq_ = h2 ? q4[j] : q4[2 + j];
          s2[j] = ks + __shfl_xor_sync(0xffffffffu, ss, 4);
          q2[j] = kq + __shfl_xor_sync(0xffffffffu, sq_, 4);
          c2[j] = h2 ? c4[2 + j] : c4[j];
        }
        float s1 = (h1 ? s2[1] : s2[0]) + __shfl_xor_sync(0xffffffffu, h1 ? s2[0] : s2[1], 2);
        float q1 = (h1 ? q2[1] : q2[0]) + __shfl_xor_sync(0xffffffffu, h1 ? q2[0] : q2[1], 2);
        const float c1 = h1 ? c2[1] : c2[0];
        s1 += __shfl_xor_sync(0xffffffffu, s1, 1);
        q1 += __shfl_xor_sync(0xffffffffu, q1, 1);
        const int jpx = (h3 ? 4 : 0) + (h2 ? 2 : 0) + (h1 ? 1 : 0);   // the pixel of the group this lane finished
        const float md = s1 * P.inv_c;
        const float var = fmaxf(q1 * P.inv_c - md * md, 0.f);
        const float rstd = 1.0f / sqrtf(var + 1e-5f);
        const float nmv = withbias ? -rstd * (c1 + md) : 0.f;
        if ((sub & 1u) == 0) s_nr[lw][lane >> 4][jpx] = make_float2(nmv, rstd);
        __syncwarp();
        float nm[8], rs[8];
        {
          const float4* p4 = reinterpret_cast<const float4*>(&s_nr[lw][lane >> 4][0]);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float4 u = p4[j];
            nm[2 * j] = u.x; rs[2 * j] = u.y; nm[2 * j + 1] = u.z; rs[2 * j + 1] = u.w;
          }
        }
        for (int r = (int)sub; r < P.C; r += 16) {
          const uint32_t addr = tile + (uint32_t)r * 128u + ((G ^ ((uint32_t)r & 7u)) << 4);
          float v[8];
          ln_stat_unpack(addr, v);
          uint32_t pk[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            __nv_bfloat162 h = __floats2bfloat162_rn(fmaf(v[2 * j], rs[2 * j], nm[2 * j]),
                                                     fmaf(v[2 * j + 1], rs[2 * j + 1], nm[2 * j + 1]));
            pk[j] = *reinterpret_cast<uint32_t*>(&h);
          }
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(pk[0]), "r"(pk[1]), "r"(pk[2]), "r"(pk[3])
                       : "memory");
        }
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
        __syncwarp();
        if (lane == 0) mbar_arrive(&bar_xnorm[stage]);
      }
    }
  } else {
    // ================= depthwise 3x3 (+ gate) straight out of TMEM: thread = channel (pair), warp = column part ======
    const int quad = warp & 3, jpart = warp >> 2;
    const uint32_t tlane = tmem_base + ((uint32_t)(quad * 32) << 16);
    const int64_t HWp = (int64_t)P.H * P.W;
    uint32_t n = 0;
    int cur_cb = -1;
    float2 wA[9], wB[9];
    float bA = 0.f, bB = 0.f;
    int cA = 0, cB = 0;
    bool okA = false, okB = false;
    for (int t = t_begin; t < t_end; ++t) {
      const FuItem I = fu_decode(t, P);
      if (I.cb != cur_cb) {
        cur_cb = I.cb;
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
      const int Pn = (I.hi - I.lo) >> 1;
      const int p0 = (Pn * jpart) / NJ, p1 = (Pn * (jpart + 1)) / NJ;
      const int lo_l = I.lo - I.xs;
      for (int i = 0; i < I.nr; ++i) {
        const uint32_t nn = n + (uint32_t)i + 2u;
        mbar_wait_to(&bar_rowfull[nn & 3], (nn >> 2) & 1);
        tc_fence_after();
        bf16* orow = P.y + (int64_t)I.b * P.y_bs + (int64_t)(I.y0 + i) * P.W + I.xs;
        for (int ps = p0; ps < p1; ps += 8) {
          int e0 = lo_l + 2 * ps;
          if (e0 > 46) e0 = 46;                       // the 16-pixel window [e0, e0+16) (+1 halo each side) stays in the slot
          const int koff = (lo_l + 2 * ps - e0) >> 1; // first pair of the window that belongs to this sub-segment
          const int cnt = min(8, p1 - ps);
          float2 accA[8], accB[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            accA[k] = make_float2(bA, bA);
            accB[k] = make_float2(bB, bB);
          }
#pragma unroll
          for (int r = 0; r < 3; ++r) {
            const uint32_t slot = (n + (uint32_t)i + (uint32_t)r) & 3u;
            const uint32_t ta = tlane + slot * 64u + (uint32_t)e0;
            {
              uint32_t O[18], E[16];
              tmem_ld_32x16(ta - 1, O);
              tmem_ld_32x2(ta + 15, O + 16);
              tmem_ld_32x16(ta, E);
              tmem_ld_wait();
#pragma unroll
              for (int k = 0; k < 8; ++k) {
                accA[k] = ffma2(wA[r * 3 + 0], f2(O[2 * k], O[2 * k + 1]), accA[k]);
                accA[k] = ffma2(wA[r * 3 + 1], f2(E[2 * k], E[2 * k + 1]), accA[k]);
                accA[k] = ffma2(wA[r * 3 + 2], f2(O[2 * k + 2], O[2 * k + 3]), accA[k]);
              }
            }
            if (haveB) {
              uint32_t O[18], E[16];
              tmem_ld_32x16(ta + 256u - 1, O);
              tmem_ld_32x2(ta + 256u + 15, O + 16);
              tmem_ld_32x16(ta + 256u, E);
              tmem_ld_wait();
#pragma unroll
              for (int k = 0; k < 8; ++k) {
                accB[k] = ffma2(wB[r * 3 + 0], f2(O[2 * k], O[2 * k + 1]), accB[k]);
                accB[k] = ffma2(wB[r * 3 + 1], f2(E[2 * k], E[2 * k + 1]), accB[k]);
                accB[k] = ffma2(wB[r * 3 + 2], f2(O[2 * k + 2], O[2 * k + 3]), accB[k]);
              }
            }
          }
          if (GATE) {
            if (okA) {
              bf16* op = orow + (int64_t)cA * HWp + e0;
#pragma unroll
              for (int k = 0; k < 8; ++k) {
                if (k >= koff && k < koff + cnt) {
                  const __nv_bfloat162 h = __floats2bfloat162_rn(gelu2_fast(accA[k].x) * accB[k].x,
                                                                 gelu2_fast(accA[k].y) * accB[k].y);
                  *reinterpret_cast<__nv_bfloat162*>(op + 2 * k) = h;
                }
              }
            }
          } else {
            bf16* opA = orow + (int64_t)cA * HWp + e0;
            bf16* opB = orow + (int64_t)cB * HWp + e0;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              if (k >= koff && k < koff + cnt) {
                if (okA) *reinterpret_cast<__nv_bfloat162*>(opA + 2 * k) = __floats2bfloat162_rn(accA[k].x, accA[k].y);
                if (haveB && okB)
                  *reinterpret_cast<__nv_bfloat162*>(opB + 2 * k) = __floats2bfloat162_rn(accB[k].x, accB[k].y);
              }
            }
          }
        }
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
  if (C < 16 || Cout < 1) return false;
  g->Cp = (C + 15) / 16 * 16;
  g->Kbox = g->Cp + 16;
  if (g->Kbox > 256) return false;
  g->KS = g->Kbox / 16;
  g->Kf = (g->Kbox + 63) / 64 * 64;
  g->kf64 = g->Kf / 64;
  g->ncb = gate ? (Cout + 127) / 128 : (Cout + 255) / 256;
  g->nrows = g->ncb * 256;
  const size_t wbytes = (size_t)g->kf64 * 32768, stage = (size_t)g->Kbox * 128;
  const long avail = 227 * 1024 - 4 * 1024 - 1024 - (long)wbytes;
  int S = avail > 0 ? (int)(avail / (long)stage) : 0;
  if (S > 6) S = 6;
  if (S < 3) return false;
  g->S = S;
  g->smem = wbytes + (size_t)S * stage + 1024;
  return true;
}

}  // namespace b200ir

using namespace b200ir;

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
  P.nstrips = (W + 4 + kFuStripStep - 1) / kFuStripStep;
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
  int grid = sms < P.total_items ? sms : P.total_items;
  ensure_dyn_smem((const void*)fused_ln_proj_dw_kernel<GATE, NCW>, 227 * 1024);
  fused_ln_proj_dw_kernel<GATE, NCW><<<grid, (NCW + kFuLnWarps + 2) * 32, g.smem, st>>>(tmX, tmW, P);
  return check_launch(GATE ? "ln_proj_dwgate_fused" : "ln_proj_dw_fused");
}

extern "C" int b200ir_ln_proj_dw_fwd(const void* x, int64_t x_bstride, const void* wf, const float* dw_w,
                                     const float* dw_b, int ln_type, void* y, int64_t y_bstride, int B, int C, int Cout,
                                     int H, int W, int gate, b200ir_stream_t stream) {
  B200IR_CHECK_ARG(x && wf && dw_w && y && B > 0 && C > 0 && Cout > 0 && H > 0 && W > 0, "ln_proj_dw: bad args");
  B200IR_CHECK_ARG(ln_type == B200IR_LN_WITHBIAS || ln_type == B200IR_LN_BIASFREE, "ln_proj_dw: bad ln_type %d", ln_type);
  B200IR_UNSUPPORTED(W % 8 == 0 && W >= 16 && x_bstride % 8 == 0 && y_bstride % 2 == 0 && ((uintptr_t)x % 16) == 0 &&
                         ((uintptr_t)y % 4) == 0,
                     "ln_proj_dw: needs W%%8==0, W>=16, 16B-aligned bf16 activations");
  if (gate)
    return fused_launch<true>(x, x_bstride, wf, dw_w, dw_b, ln_type, y, y_bstride, B, C, Cout, H, W, (cudaStream_t)stream);
  return fused_launch<false>(x, x_bstride, wf, dw_w, dw_b, ln_type, y, y_bstride, B, C, Cout, H, W, (cudaStream_t)stream);
}
