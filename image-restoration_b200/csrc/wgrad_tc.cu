// wgrad_tc.cu — weight gradient of a 1x1 conv on tcgen05: dW[o][c] = sum_b sum_p dy[b,o,p] * x[b,c,p].
// (autograd of the 1x1 nn.Conv2d calls of Restormer.py:89,92,114,131,261,266 and, per sample, the gradient of
// the folded MDTA matrix dM_b = dout_b . v_b^T, Restormer.py:127-131.)
//
// In NCHW both operands are [channels, HW] with HW contiguous = K-major for a GEMM whose K is the pixel index, so
// the tiles come straight from TMA (SWIZZLE_128B, 64 pixels = 128 B per row) with no transposition.  One CTA owns a
// 128-row slab of dy (MMA M), up to 256 rows of x (MMA N) and one split of the pixel range; fp32 accumulation in
// TMEM over the whole split; the epilogue writes the fp32 partial tile, splits are folded in fixed order by
// b200ir_reduce_partials (deterministic).  Out-of-range channels / pixels are zero-filled by TMA.
#include <mutex>

#include "tc_common.cuh"

namespace b200ir {

struct WgradTcParams {
  float* pw;
  int M, N, Ntile, G, nsplit;
  int chunks_per_sample, chunks_per_split, chunks_total, per_sample;
  int S;
};

constexpr int kWgThreads = 192;  // warps 0-3 epilogue, 4 TMA producer, 5 MMA issuer

__global__ void __launch_bounds__(kWgThreads)
wgrad_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmX,
                const WgradTcParams P) {
  using namespace tc;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t a_bytes = 128u * 128u;
  const uint32_t x_bytes = (uint32_t)P.Ntile * 128u;
  const uint32_t stage_bytes = a_bytes + x_bytes;   // both multiples of 1024 (Ntile % 16 == 0 -> x_bytes % 2048 == 0)
  uint8_t* ring = smem;
  uint8_t* tail = ring + (size_t)P.S * stage_bytes;
  uint64_t* bar_full = (uint64_t*)tail;   // [8]
  uint64_t* bar_empty = bar_full + 8;     // [8]
  uint64_t* bar_acc = bar_empty + 8;      // [1]
  uint32_t* tmem_slot = (uint32_t*)(bar_acc + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.x * P.Ntile, o0 = blockIdx.y * 128;
  const int g = blockIdx.z / P.nsplit, split = blockIdx.z % P.nsplit;
  const int cbeg = split * P.chunks_per_split;
  const int cend = min(P.chunks_total, cbeg + P.chunks_per_split);
  const int nchunks = cend - cbeg;
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
      tma_prefetch_desc(&tmA);
      tma_prefetch_desc(&tmX);
    }
    tmem_alloc(tmem_slot, ncols);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 4) {
    if (lane == 0) {
      for (int i = 0; i < nchunks; ++i) {
        const uint32_t stage = i % P.S, ph = (i / P.S) & 1;
        const int ch = cbeg + i;
        const int b = P.per_sample ? g : ch / P.chunks_per_sample;
        const int px = (ch % P.chunks_per_sample) * 64;
        mbar_wait(&bar_empty[stage], ph ^ 1);
        mbar_arrive_expect_tx(&bar_full[stage], stage_bytes);
        uint8_t* dst = ring + (size_t)stage * stage_bytes;
        tma_load_3d(dst, &tmA, &bar_full[stage], px, o0, b);
        tma_load_3d(dst + a_bytes, &tmX, &bar_full[stage], px, n0, b);
      }
    }
  } else if (warp == 5) {
    {   // whole warp, convergent; the issuing lane is elected per instruction (tc_common.cuh)
      const uint32_t idesc = make_idesc_bf16(128, P.Ntile, 0, 0);
      const uint32_t ring_addr = smem_u32(ring);
      for (int i = 0; i < nchunks; ++i) {
        const uint32_t stage = i % P.S, ph = (i / P.S) & 1;
        mbar_wait(&bar_full[stage], ph);
        tc_fence_after();
        const uint32_t aa = ring_addr + stage * stage_bytes, xa = aa + a_bytes;
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          const uint64_t da = make_smem_desc_sw128(aa + ks * 32, 16u, 1024u);
          const uint64_t dx = make_smem_desc_sw128(xa + ks * 32, 16u, 1024u);
          umma_bf16_elect(tmem_base, da, dx, idesc, (i | ks) != 0 ? 1u : 0u);
        }
        umma_commit_elect(&bar_empty[stage]);
      }
      umma_commit_elect(bar_acc);
    }
  } else {
    // epilogue: thread <-> TMEM lane = row of the dy slab
    const int row = warp * 32 + lane, o = o0 + row;
    const size_t slot = (size_t)split * P.G + g;  // partials laid out [split][g][M][N]
    float* out = P.pw + (slot * P.M + (size_t)o) * P.N + n0;
    const uint32_t lane_base = tmem_base + ((uint32_t)(warp * 32) << 16);
    if (nchunks > 0) {
      mbar_wait(bar_acc, 0);
      tc_fence_after();
    }
    const bool vec = (P.N % 4 == 0);
    if (o0 + warp * 32 < P.M) {  // warp-uniform: skip slabs entirely past M
      for (int c0 = 0; c0 < P.Ntile && n0 + c0 < P.N; c0 += 16) {
        uint32_t r[16];
        if (nchunks > 0) {
          tmem_ld_32x16(lane_base + c0, r);
          tmem_ld_wait();
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j) r[j] = 0u;
        }
        if (o < P.M) {
          if (vec && n0 + c0 + 16 <= P.N) {
            float4* o4 = reinterpret_cast<float4*>(out + c0);
#pragma unroll
            for (int j = 0; j < 4; ++j)
              o4[j] = make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]),
                                  __uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3]));
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (n0 + c0 + j < P.N) out[c0 + j] = __uint_as_float(r[j]);
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) tmem_dealloc(tmem_base, ncols);
}

// db partial: pb[0][g][o] = sum over the group's samples and pixels of dy[b,o,p]; pb[s>0][g][o] = 0.
__global__ void __launch_bounds__(256)
rowsum_kernel(const void* __restrict__ A, int64_t a_bs, int a_dt, float* __restrict__ pb, int M, int64_t HW, int B,
              int per_sample, int nsplit) {
  const int o = blockIdx.x, g = blockIdx.y, G = gridDim.y;
  float s = 0.f;
  const int b0 = per_sample ? g : 0, b1 = per_sample ? g + 1 : B;
  for (int b = b0; b < b1; ++b)
    for (int64_t p = threadIdx.x; p < HW; p += 256) s += ld_any(A, a_dt, (int64_t)b * a_bs + (int64_t)o * HW + p);
  __shared__ float red[8];
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < 8; ++w) t += red[w];
    pb[(size_t)g * M + o] = t;
    for (int sp = 1; sp < nsplit; ++sp) pb[((size_t)sp * G + g) * M + o] = 0.f;
  }
}

bool wgrad_tc_supported(const void* dy, int64_t dy_bs, int dy_dt, const void* x, int64_t x_bs, int x_dt, int taps,
                        int64_t HW) {
  return taps == 1 && dy_dt == B200IR_BF16 && x_dt == B200IR_BF16 && HW % 8 == 0 && HW >= 64 && HW < (1ll << 31) &&
         dy_bs % 8 == 0 && x_bs % 8 == 0 && ((uintptr_t)dy % 16) == 0 && ((uintptr_t)x % 16) == 0;
}

void wgrad_tc_tiles(int M, int N, int* mtiles, int* ntiles, int* Ntile) {
  const int n16 = (N + 15) / 16 * 16;
  const int nt = (n16 + 255) / 256;
  *Ntile = ((n16 + nt - 1) / nt + 15) / 16 * 16;
  *ntiles = nt;
  *mtiles = (M + 127) / 128;
}

int wgrad_tc_launch(const void* dy, int64_t dy_bs, int dy_dt, const void* x, int64_t x_bs, float* pw, float* pb, int B,
                    int M, int N, int64_t HW, int per_sample, int nsplit, int chunks_per_sample, int chunks_per_split,
                    int chunks_total, cudaStream_t st) {
  int mtiles, ntiles, Ntile;
  wgrad_tc_tiles(M, N, &mtiles, &ntiles, &Ntile);
  CUtensorMap tmA, tmX;
  {
    const uint64_t dims[3] = {(uint64_t)HW, (uint64_t)M, (uint64_t)B};
    const uint64_t str[3] = {2, (uint64_t)HW * 2, (uint64_t)dy_bs * 2};
    const uint32_t box[3] = {64, 128, 1};
    int rc = make_tmap_bf16(&tmA, dy, 3, dims, str, box);
    if (rc) return rc;
  }
  {
    const uint64_t dims[3] = {(uint64_t)HW, (uint64_t)N, (uint64_t)B};
    const uint64_t str[3] = {2, (uint64_t)HW * 2, (uint64_t)x_bs * 2};
    const uint32_t box[3] = {64, (uint32_t)Ntile, 1};
    int rc = make_tmap_bf16(&tmX, x, 3, dims, str, box);
    if (rc) return rc;
  }
  WgradTcParams P;
  P.pw = pw; P.M = M; P.N = N; P.Ntile = Ntile; P.G = per_sample ? B : 1; P.nsplit = nsplit;
  P.chunks_per_sample = chunks_per_sample; P.chunks_per_split = chunks_per_split; P.chunks_total = chunks_total;
  P.per_sample = per_sample;
  const size_t stage = 128 * 128 + (size_t)Ntile * 128;
  int S = (int)((100 * 1024) / stage);   // <= ~100 KB of ring: two CTAs per SM
  if (S > 8) S = 8;
  if (S < 2) S = 2;
  P.S = S;
  const size_t smem = (size_t)S * stage + 1024 + 256;
  ensure_dyn_smem((const void*)wgrad_tc_kernel, 227 * 1024);
  dim3 grid((unsigned)ntiles, (unsigned)mtiles, (unsigned)(P.G * nsplit));
  wgrad_tc_kernel<<<grid, kWgThreads, smem, st>>>(tmA, tmX, P);
  int rc = check_launch("conv_wgrad_tc");
  if (rc) return rc;
  if (pb) {
    rowsum_kernel<<<dim3((unsigned)M, (unsigned)P.G), 256, 0, st>>>(dy, dy_bs, dy_dt, pb, M, HW, B, per_sample, nsplit);
    rc = check_launch("conv_wgrad_rowsum");
  }
  return rc;
}

}  // namespace b200ir
