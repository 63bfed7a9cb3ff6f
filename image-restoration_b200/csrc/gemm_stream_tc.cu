// gemm_stream_tc.cu — 1x1 conv (+ folded LayerNorm, bias, residual) on tcgen05 for the DEEP U-Net levels, where the
// packed weight matrix (e.g. 2042 x 384 at level 4, Restormer.py:80 project_in) does not fit in shared memory next to
// the activation ring: gemm_tc.cu then cuts it in up to 15 resident slices of ~144 rows, every slice re-reads all
// activations and issues N=144 MMAs.  Here BOTH operands are streamed: a work item is 128 pixels x Ntile (<= 256) output
// channels, K runs over 64-channel chunks, each pipeline stage holds
//   A: the NCHW activation tile  [64 ch][128 px]  (MN-major, two 64-pixel SWIZZLE_128B atoms, 3D TMA)
//   B: the packed weight rows    [Ntile][64 k]    (K-major, from b200ir_pack_conv1x1; <= 4 MB, L2-resident)
// Persistent CTAs (one per SM) walk the items with the out-channel tile fastest, so the CTAs running at the same time
// share activation tiles through L2.  Two 256-column TMEM accumulators let the epilogue of one item overlap the MMAs of
// the next; four extra warps compute the per-pixel LayerNorm statistics from the staged activation tiles.
// Same math as gemm_tc.cu MODE 0: y = acc*rstd + (-rstd*mu)*s[o] + t[o] (+ res).
#include <mutex>
#include <type_traits>

#include "tc_common.cuh"

namespace b200ir {

struct GemmStreamParams {
  void* y; int64_t y_bs; int y_dt;
  const void* res; int64_t res_bs; int res_dt;
  const float* s; const float* t;
  int ln_type, Cin, Cout, Nrows, Ntile, n_tiles, kchunks, S, per_sample_w;
  int tiles_per_sample, total_items;
  int64_t HW;
  float inv_cin;
};

constexpr int kGsEpiWarps = 8;
constexpr int kGsWarpProducer = 8, kGsWarpMma = 9, kGsWarpStats0 = 10;
constexpr int kGsThreads = (kGsWarpStats0 + 4) * 32;  // 448

template <typename T> __device__ __forceinline__ float ldf(const T* p);
template <> __device__ __forceinline__ float ldf<bf16>(const bf16* p) { return __bfloat162float(*p); }
template <> __device__ __forceinline__ float ldf<float>(const float* p) { return *p; }
template <typename T> __device__ __forceinline__ void stf(T* p, float v);
template <> __device__ __forceinline__ void stf<bf16>(bf16* p, float v) { *p = __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ void stf<float>(float* p, float v) { *p = v; }

// TRES: void (no residual) | bf16 | float ; TOUT: bf16 | float
template <typename TRES, typename TOUT>
__global__ void __launch_bounds__(kGsThreads, 1)
gemm_stream_tc_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW,
                      const __grid_constant__ CUtensorMap tmY, const GemmStreamParams P) {
  using namespace tc;
  constexpr bool HAS_RES = !std::is_same<TRES, void>::value;
  typedef typename std::conditional<HAS_RES, TRES, float>::type TR;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* ring = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  // bf16 output without residual leaves through shared memory + TMA stores; with a residual the 16-column steps make
  // the store boxes too small to hide the TMA latency (measured: 1021->384 60 us vs 39 us direct), so it stores directly
  constexpr bool OUT16 = std::is_same<TOUT, bf16>::value && !HAS_RES;
  const uint32_t a_bytes = 16384u, w_bytes = (uint32_t)P.Ntile * 128u, stage_bytes = a_bytes + w_bytes;
  __shared__ uint64_t bar_full[8], bar_empty[8], bar_accfull[2], bar_accempty[2], bar_stfull[2], bar_stempty[2];
  __shared__ uint32_t tmem_slot[1];
  __shared__ __align__(16) float2 s_st[512];   // [2][256]: (s, t) of the item's out-channel tile
  __shared__ float s_nm[256], s_rstd[256];   // [2][128]: (-rstd*mu, rstd) per pixel of the item, double buffered

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool has_ln = P.ln_type != B200IR_LN_NONE;

  if (threadIdx.x == 0) {
    for (int i = 0; i < P.S; ++i) {
      mbar_init(&bar_full[i], 1);
      mbar_init(&bar_empty[i], has_ln ? 5 : 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bar_accfull[i], 1);
      mbar_init(&bar_accempty[i], kGsEpiWarps);
      mbar_init(&bar_stfull[i], 128);
      mbar_init(&bar_stempty[i], kGsEpiWarps);
    }
    fence_barrier_init();
  }
  if (warp == kGsWarpProducer) {
    if (lane == 0) {
      tma_prefetch_desc(&tmX);
      tma_prefetch_desc(&tmW);
      if (OUT16) tma_prefetch_desc(&tmY);
    }
    tmem_alloc(tmem_slot, 512);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == kGsWarpProducer) {
    if (lane == 0) {
      uint32_t it = 0;
      for (int item = blockIdx.x; item < P.total_items; item += gridDim.x) {
        const int nt = item % P.n_tiles, pt = item / P.n_tiles;
        const int b = pt / P.tiles_per_sample, p0 = (pt - b * P.tiles_per_sample) * 128;
        for (int kc = 0; kc < P.kchunks; ++kc, ++it) {
          const uint32_t stage = it % P.S, ph = (it / P.S) & 1;
          mbar_wait(&bar_empty[stage], ph ^ 1);
          mbar_arrive_expect_tx(&bar_full[stage], stage_bytes);
          uint8_t* dst = ring + (size_t)stage * stage_bytes;
          tma_load_3d(dst, &tmX, &bar_full[stage], p0, kc * 64, b);
          tma_load_3d(dst + 8192, &tmX, &bar_full[stage], p0 + 64, kc * 64, b);
          tma_load_3d(dst + a_bytes, &tmW, &bar_full[stage], kc * 64, nt * P.Ntile, P.per_sample_w ? b : 0);
        }
      }
    }
  } else if (warp == kGsWarpMma) {
    {   // whole warp, convergent; the issuing lane is elected per instruction (tc_common.cuh)
      const uint32_t idesc = make_idesc_bf16(128, P.Ntile, 1, 0);
      const uint32_t ring_addr = smem_u32(ring);
      uint32_t it = 0, acc_it = 0;
      for (int item = blockIdx.x; item < P.total_items; item += gridDim.x, ++acc_it) {
        const uint32_t as = acc_it & 1, aph = (acc_it >> 1) & 1;
        mbar_wait(&bar_accempty[as], aph ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * 256;
        for (int kc = 0; kc < P.kchunks; ++kc, ++it) {
          const uint32_t stage = it % P.S, ph = (it / P.S) & 1;
          mbar_wait(&bar_full[stage], ph);
          tc_fence_after();
          const uint32_t aa = ring_addr + stage * stage_bytes, wa = aa + a_bytes;
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) {
            const uint64_t da = make_smem_desc_sw128(aa + ks * 2048, 8192u, 1024u);
            const uint64_t dw = make_smem_desc_sw128(wa + ks * 32, 16u, 1024u);
            umma_bf16_elect(d_tmem, da, dw, idesc, (kc | ks) != 0 ? 1u : 0u);
          }
          umma_commit_elect(&bar_empty[stage]);
        }
        umma_commit_elect(&bar_accfull[as]);
      }
    }
  } else if (warp >= kGsWarpStats0) {
    // LayerNorm statistics of the 128 pixels of each item, from the staged activation tiles (shifted sums)
    if (has_ln) {
      const int m = (warp - kGsWarpStats0) * 32 + lane;
      const uint32_t ring_addr = smem_u32(ring);
      uint32_t it = 0, icount = 0;
      for (int item = blockIdx.x; item < P.total_items; item += gridDim.x, ++icount) {
        LnStatAcc acc;
        for (int kc = 0; kc < P.kchunks; ++kc, ++it) {
          const uint32_t stage = it % P.S, ph = (it / P.S) & 1;
          mbar_wait(&bar_full[stage], ph);
          const int rows = min(64, P.Cin - kc * 64);
          ln_stat_chunk(ring_addr + stage * stage_bytes, 8192u, m, rows, kc == 0, acc);
          __syncwarp();
          if (lane == 0) mbar_arrive(&bar_empty[stage]);
        }
        float x0, sum, sq;
        ln_stat_finish(acc, m, x0, sum, sq);
        const float md = sum * P.inv_cin;
        const float var = fmaxf(sq * P.inv_cin - md * md, 0.f);
        const uint32_t sp = icount & 1, sph = (icount >> 1) & 1;
        mbar_wait(&bar_stempty[sp], sph ^ 1);
        const float rstd = 1.0f / sqrtf(var + 1e-5f);
        s_nm[sp * 128 + m] = (P.ln_type == B200IR_LN_WITHBIAS) ? -rstd * (x0 + md) : 0.f;
        s_rstd[sp * 128 + m] = rstd;
        mbar_arrive(&bar_stfull[sp]);
      }
    }
  } else {
    // epilogue: warp w <-> TMEM lane quadrant w&3 (32 pixels = 64 contiguous bytes per output channel), 16/32-column
    // groups g with g&1 == w>>2.  (s,t) of the item's out-channel tile are staged in shared memory by the epilogue
    // warps themselves; residual values of a group are all loaded before the first store (y may alias nothing here,
    // but the compiler cannot know), so their latencies overlap.
    const int quad = warp & 3, half = warp >> 2;
    const int row = quad * 32 + lane;
    constexpr int GW = HAS_RES ? 16 : 32;   // columns per step (residual values double the live registers)
    const int groups = (P.Ntile + GW - 1) / GW;
    TOUT* yout = reinterpret_cast<TOUT*>(P.y);
    const TR* rin = reinterpret_cast<const TR*>(P.res);
    // staging of one warp: two [GW ch][32 px] bf16 tiles, sent by TMA store boxes {32 px, GW ch} (clipped at HW / Cout)
    uint8_t* stg = ring + (size_t)P.S * stage_bytes + (size_t)warp * 4096;
    uint32_t acc_it = 0, st_it = 0;
    for (int item = blockIdx.x; item < P.total_items; item += gridDim.x, ++acc_it) {
      const int nt = item % P.n_tiles, pt = item / P.n_tiles;
      const int b = pt / P.tiles_per_sample;
      const int pw0 = (pt - b * P.tiles_per_sample) * 128 + quad * 32;   // first pixel of this warp
      const int64_t p = (int64_t)pw0 + lane;
      const bool pok = p < P.HW;
      const int n0 = nt * P.Ntile;
      const uint32_t as = acc_it & 1, aph = (acc_it >> 1) & 1;
      {  // every warp passed the named barrier of item-1, i.e. finished reading s_st[as] for item-2: free to overwrite
        const int e = threadIdx.x;   // 0..255
        const int co = n0 + e;
        s_st[as * 256 + e] = (e < P.Ntile && co < P.Nrows) ? make_float2(__ldg(P.s + co), __ldg(P.t + co))
                                                            : make_float2(0.f, 0.f);
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");
      float nm = 0.f, rs = 1.f;
      if (has_ln) {
        mbar_wait(&bar_stfull[as], aph);
        nm = s_nm[as * 128 + row];
        rs = s_rstd[as * 128 + row];
      }
      mbar_wait(&bar_accfull[as], aph);
      tc_fence_after();
      const uint32_t tb = tmem_base + as * 256 + ((uint32_t)(quad * 32) << 16);
      TOUT* yp = yout + (int64_t)b * P.y_bs + p;
      const TR* rp = HAS_RES ? rin + (int64_t)b * P.res_bs + p : nullptr;
      const float2* stv = &s_st[as * 256];
      for (int g = half; g < groups; g += 2) {
        const int c0 = g * GW;
        if (n0 + c0 >= P.Cout) break;
        uint32_t r[GW];
        tmem_ld_32x16(tb + c0, r);
        if (GW == 32 && c0 + 16 < P.Ntile) tmem_ld_32x16(tb + c0 + 16, r + GW - 16);
        float rv[GW];
        if (HAS_RES) {
#pragma unroll
          for (int j = 0; j < GW; ++j) {
            const int co = n0 + c0 + j;
            rv[j] = (pok && co < P.Cout && c0 + j < P.Ntile) ? ldf<TR>(rp + (int64_t)co * P.HW) : 0.f;
          }
        }
        tmem_ld_wait();
        bf16* sbuf = reinterpret_cast<bf16*>(stg + (st_it & 1) * 2048);
        if (OUT16) {
          if (lane == 0) bulk_wait_read<1>();   // the store issued from this buffer two groups ago has read it
          __syncwarp();
        }
#pragma unroll
        for (int j = 0; j < GW; ++j) {
          const int co = n0 + c0 + j;
          const float2 st2 = stv[c0 + j];
          float v = has_ln ? fmaf(__uint_as_float(r[j]), rs, fmaf(nm, st2.x, st2.y)) : __uint_as_float(r[j]) + st2.y;
          if (HAS_RES) v += rv[j];
          if (OUT16) sbuf[j * 32 + lane] = __float2bfloat16_rn(v);
          else if (pok && co < P.Cout && c0 + j < P.Ntile) stf<TOUT>(yp + (int64_t)co * P.HW, v);
        }
        if (OUT16) {
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) {
            tma_store_3d(&tmY, sbuf, pw0, n0 + c0, b);
            bulk_commit();
          }
          ++st_it;
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&bar_accempty[as]);
        if (has_ln) mbar_arrive(&bar_stempty[as]);
      }
    }
  }
  if (OUT16 && warp < kGsEpiWarps && lane == 0) bulk_wait_read<0>();   // staging must outlive its last store
  tc_fence_before();
  __syncthreads();
  if (warp == kGsWarpProducer) tmem_dealloc(tmem_base, 512);
}

// wp: packed [Bw][Nrows][Kpad] bf16 (b200ir_pack_conv1x1), st: [2][Nrows]
int gemm_stream_launch(const void* x, int64_t x_bs, const void* wp, int Bw, const float* st, int ln_type,
                       const void* res, int64_t res_bs, int res_dt, void* y, int64_t y_bs, int y_dt, int B, int Cin,
                       int Cout, int64_t HW, int Kpad, int Nrows, cudaStream_t stream) {
  const int n_tiles = (Nrows + 255) / 256;
  // multiple of 32: a 32-channel store box never reaches into the next out-channel tile
  const int Ntile = ((Nrows + n_tiles - 1) / n_tiles + 31) / 32 * 32;
  CUtensorMap tmX, tmW;
  {
    const uint64_t dims[3] = {(uint64_t)HW, (uint64_t)Cin, (uint64_t)B};
    const uint64_t str[3] = {2, (uint64_t)HW * 2, (uint64_t)x_bs * 2};
    const uint32_t box[3] = {64, 64, 1};
    int rc = make_tmap_bf16(&tmX, x, 3, dims, str, box);
    if (rc) return rc;
  }
  {
    const uint64_t dims[3] = {(uint64_t)Kpad, (uint64_t)Nrows, (uint64_t)Bw};
    const uint64_t str[3] = {2, (uint64_t)Kpad * 2, (uint64_t)Nrows * Kpad * 2};
    const uint32_t box[3] = {64, (uint32_t)Ntile, 1};
    int rc = make_tmap_bf16(&tmW, wp, 3, dims, str, box);
    if (rc) return rc;
  }
  CUtensorMap tmY = tmX;
  if (y_dt == B200IR_BF16 && !res) {
    const uint64_t dims[3] = {(uint64_t)HW, (uint64_t)Cout, (uint64_t)B};
    const uint64_t str[3] = {2, (uint64_t)HW * 2, (uint64_t)y_bs * 2};
    const uint32_t box[3] = {32, 32, 1};
    int rc = make_tmap_bf16(&tmY, y, 3, dims, str, box, false);
    if (rc) return rc;
  }
  GemmStreamParams P;
  P.y = y; P.y_bs = y_bs; P.y_dt = y_dt; P.res = res; P.res_bs = res_bs; P.res_dt = res_dt;
  P.s = st; P.t = st + Nrows;
  P.ln_type = ln_type; P.Cin = Cin; P.Cout = Cout; P.Nrows = Nrows; P.Ntile = Ntile; P.n_tiles = n_tiles;
  P.kchunks = Kpad / 64; P.per_sample_w = Bw > 1 ? 1 : 0;
  P.tiles_per_sample = (int)((HW + 127) / 128);
  const int64_t items = (int64_t)P.tiles_per_sample * B * n_tiles;
  if (items >= (1ll << 31)) {
    set_error("gemm_stream: too many work items");
    return B200IR_E_UNSUPPORTED;
  }
  P.total_items = (int)items;
  P.HW = HW;
  P.inv_cin = 1.0f / (float)Cin;
  const size_t stage = 16384 + (size_t)Ntile * 128;
  int S = (int)((172 * 1024) / stage);   // + 32 KB of epilogue staging
  if (S > 8) S = 8;
  P.S = S;
  const size_t smem = (size_t)S * stage + kGsEpiWarps * 4096 + 1024;
  int grid = num_sms();
  if (grid > P.total_items) grid = P.total_items;
#define B200IR_LAUNCH_GS(TRES, TOUT)                                                                             \
  do {                                                                                                          \
    ensure_dyn_smem((const void*)gemm_stream_tc_kernel<TRES, TOUT>, 216 * 1024);                                                                                                         \
    gemm_stream_tc_kernel<TRES, TOUT><<<grid, kGsThreads, smem, stream>>>(tmX, tmW, tmY, P);                         \
  } while (0)
  const bool ob = y_dt == B200IR_BF16;
  if (!res) {
    if (ob) B200IR_LAUNCH_GS(void, bf16); else B200IR_LAUNCH_GS(void, float);
  } else if (res_dt == B200IR_BF16) {
    if (ob) B200IR_LAUNCH_GS(bf16, bf16); else B200IR_LAUNCH_GS(bf16, float);
  } else {
    if (ob) B200IR_LAUNCH_GS(float, bf16); else B200IR_LAUNCH_GS(float, float);
  }
#undef B200IR_LAUNCH_GS
  return check_launch("gemm_stream_tc");
}

}  // namespace b200ir
