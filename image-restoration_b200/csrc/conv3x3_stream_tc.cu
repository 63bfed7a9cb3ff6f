// conv3x3_stream_tc.cu — dense 3x3 conv (pad 1) on tcgen05 for WIDE layers, where the 9*Cin x Cout weight matrix does
// not fit next to the activation ring in shared memory (Upsample 384->768 / 192->384 / 96->192, Downsample 192->96,
// Restormer.py:171-189): both operands are streamed.  One CTA = 128 output pixels (a TH x TW patch of one image) x up
// to 256 output channels; K runs over 9 taps x ceil(Cin/64) chunks; per chunk TMA brings
//   A: the tap-shifted NHWC activation patch  [128 px][64 ch]  (4D box, zero padding = TMA out-of-bounds fill)
//   B: the packed weight rows                 [Ntile rows][64 k] (K-major, from b200ir_pack_conv3x3)
// into one SWIZZLE_128B stage; fp32 accumulation in TMEM; the epilogue adds the bias and stores with
// PixelUnshuffle(2) / PixelShuffle(2) folded into the addressing.  Weights (<= 5.3 MB) stay L2-resident, the
// activation patch is shared by the CTAs of the same pixel tile (adjacent blockIdx.x) through L2.
#include <mutex>

#include "tc_common.cuh"

namespace b200ir {

struct ConvStreamParams {
  void* y;
  int64_t y_bs;
  int y_dt;
  const float* t;   // bias per output channel (st[1] of the packed weights)
  int Cout, H, W, TW, TWs, TH, tiles_x, tiles_y, kpt, kchunks, Ntile, S, mode;
};

constexpr int kCsThreads = 192;  // warps 0-3 epilogue, 4 TMA producer, 5 MMA issuer

__global__ void __launch_bounds__(kCsThreads)
conv3x3_stream_tc_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW,
                         const ConvStreamParams P) {
  using namespace tc;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t a_bytes = 128u * 128u;
  const uint32_t w_bytes = (uint32_t)P.Ntile * 128u;
  const uint32_t stage_bytes = a_bytes + w_bytes;
  uint8_t* ring = smem;
  uint8_t* tail = ring + (size_t)P.S * stage_bytes;
  uint64_t* bar_full = (uint64_t*)tail;   // [8]
  uint64_t* bar_empty = bar_full + 8;     // [8]
  uint64_t* bar_acc = bar_empty + 8;      // [1]
  uint32_t* tmem_slot = (uint32_t*)(bar_acc + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.x * P.Ntile;
  const int tile = blockIdx.y;
  const int tpi = P.tiles_x * P.tiles_y;
  const int b = tile / tpi, r = tile - b * tpi;
  const int cy0 = (r / P.tiles_x) * P.TH, cx0 = (r % P.tiles_x) * P.TW;
  const uint32_t ncols = P.Ntile <= 32 ? 32u : (P.Ntile <= 64 ? 64u : (P.Ntile <= 128 ? 128u : 256u));

  if (threadIdx.x == 0) {
    for (int i = 0; i < P.S; ++i) {
      mbar_init(&bar_full[i], 1);
      mbar_init(&bar_empty[i], 1);
    }
    mbar_init(bar_acc, 1);
    fence_barrier_init();
  }
  if (warp == 4) {
    if (lane == 0) {
      tma_prefetch_desc(&tmX);
      tma_prefetch_desc(&tmW);
    }
    tmem_alloc(tmem_slot, ncols);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 4) {
    if (lane == 0) {
      for (int kc = 0; kc < P.kchunks; ++kc) {
        const uint32_t stage = kc % P.S, ph = (kc / P.S) & 1;
        const int tap = kc / P.kpt, cc = kc - tap * P.kpt;
        const int dy = tap / 3 - 1, dx = tap - (tap / 3) * 3 - 1;
        mbar_wait(&bar_empty[stage], ph ^ 1);
        mbar_arrive_expect_tx(&bar_full[stage], stage_bytes);
        uint8_t* dst = ring + (size_t)stage * stage_bytes;
        tma_load_4d(dst, &tmX, &bar_full[stage], cc * 64, cx0 + dx, cy0 + dy, b);
        tma_load_3d(dst + a_bytes, &tmW, &bar_full[stage], kc * 64, n0, 0);
      }
    }
  } else if (warp == 5) {
    {   // whole warp, convergent; the issuing lane is elected per instruction (tc_common.cuh)
      const uint32_t idesc = make_idesc_bf16(128, P.Ntile, 0, 0);
      const uint32_t ring_addr = smem_u32(ring);
      for (int kc = 0; kc < P.kchunks; ++kc) {
        const uint32_t stage = kc % P.S, ph = (kc / P.S) & 1;
        mbar_wait(&bar_full[stage], ph);
        tc_fence_after();
        const uint32_t aa = ring_addr + stage * stage_bytes, wa = aa + a_bytes;
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          const uint64_t da = make_smem_desc_sw128(aa + ks * 32, 16u, 1024u);
          const uint64_t dw = make_smem_desc_sw128(wa + ks * 32, 16u, 1024u);
          umma_bf16_elect(tmem_base, da, dw, idesc, (kc | ks) != 0 ? 1u : 0u);
        }
        umma_commit_elect(&bar_empty[stage]);
      }
      umma_commit_elect(bar_acc);
    }
  } else {
    // epilogue: thread <-> TMEM lane = pixel of the patch
    const int row = warp * 32 + lane;
    const int gy = cy0 + (row >> P.TWs), gx = cx0 + (row & (P.TW - 1));
    const bool valid = gy < P.H && gx < P.W;
    const int64_t HWp = (int64_t)P.H * P.W;
    const uint32_t lane_base = tmem_base + ((uint32_t)(warp * 32) << 16);
    mbar_wait(bar_acc, 0);
    tc_fence_after();
    for (int c0 = 0; c0 < P.Ntile && n0 + c0 < P.Cout; c0 += 16) {
      uint32_t rr[16];
      tmem_ld_32x16(lane_base + c0, rr);
      tmem_ld_wait();
      if (valid) {
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const int co = n0 + c0 + j;
          if (co >= P.Cout) break;
          const float v = __uint_as_float(rr[j]) + __ldg(P.t + co);
          int64_t off;
          if (P.mode == B200IR_UNSHUFFLE2) {
            const int oc = co * 4 + (gy & 1) * 2 + (gx & 1);
            off = (int64_t)oc * (HWp / 4) + (int64_t)(gy >> 1) * (P.W / 2) + (gx >> 1);
          } else if (P.mode == B200IR_SHUFFLE2) {
            const int oc = co >> 2, si = (co >> 1) & 1, sj = co & 1;
            off = (int64_t)oc * (HWp * 4) + (int64_t)(2 * gy + si) * (2 * P.W) + (2 * gx + sj);
          } else {
            off = (int64_t)co * HWp + (int64_t)gy * P.W + gx;
          }
          st_any(P.y, P.y_dt, (int64_t)b * P.y_bs + off, v);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) tmem_dealloc(tmem_base, ncols);
}

// x: NHWC bf16 [B][H][W][Cin]; wp: packed [Nrows][Kpad] (k = (tap*kpt + chunk)*64 + c); st: [2][Nrows]
int conv3x3_stream_launch(const void* x, int64_t x_bs, const void* wp, const float* st, int Nrows, int Kpad, int kpt,
                          void* y, int64_t y_bs, int y_dt, int B, int Cin, int Cout, int H, int W, int mode,
                          cudaStream_t stream) {
  int TWs = 0;
  while ((1 << TWs) < W && TWs < 7) ++TWs;
  const int TW = 1 << TWs, TH = 128 / TW;
  const int nt = (Nrows + 255) / 256;
  const int Ntile = ((Nrows + nt - 1) / nt + 15) / 16 * 16;
  CUtensorMap tmX, tmW;
  {
    const uint64_t dims[4] = {(uint64_t)Cin, (uint64_t)W, (uint64_t)H, (uint64_t)B};
    const uint64_t str[4] = {2, (uint64_t)Cin * 2, (uint64_t)W * Cin * 2, (uint64_t)x_bs * 2};
    const uint32_t box[4] = {64, (uint32_t)TW, (uint32_t)TH, 1};
    int rc = make_tmap_bf16(&tmX, x, 4, dims, str, box);
    if (rc) return rc;
  }
  {
    const uint64_t dims[3] = {(uint64_t)Kpad, (uint64_t)Nrows, 1};
    const uint64_t str[3] = {2, (uint64_t)Kpad * 2, (uint64_t)Nrows * Kpad * 2};
    const uint32_t box[3] = {64, (uint32_t)Ntile, 1};
    int rc = make_tmap_bf16(&tmW, wp, 3, dims, str, box);
    if (rc) return rc;
  }
  ConvStreamParams P;
  P.y = y; P.y_bs = y_bs; P.y_dt = y_dt; P.t = st + Nrows; P.Cout = Cout; P.H = H; P.W = W;
  P.TW = TW; P.TWs = TWs; P.TH = TH; P.tiles_x = (W + TW - 1) / TW; P.tiles_y = (H + TH - 1) / TH;
  P.kpt = kpt; P.kchunks = 9 * kpt; P.Ntile = Ntile; P.mode = mode;
  const size_t stage = 128 * 128 + (size_t)Ntile * 128;
  int S = (int)((200 * 1024) / stage);
  if (S > 6) S = 6;
  if (S < 2) S = 2;
  P.S = S;
  const size_t smem = (size_t)S * stage + 1024 + 256;
  const int64_t tiles = (int64_t)B * P.tiles_x * P.tiles_y;
  if (tiles > 65535) {
    set_error("conv3x3_stream: too many pixel tiles (%lld)", (long long)tiles);
    return B200IR_E_UNSUPPORTED;
  }
  ensure_dyn_smem((const void*)conv3x3_stream_tc_kernel, 227 * 1024);
  conv3x3_stream_tc_kernel<<<dim3((unsigned)nt, (unsigned)tiles), kCsThreads, smem, stream>>>(tmX, tmW, P);
  return check_launch("conv3x3_stream_tc");
}

}  // namespace b200ir
