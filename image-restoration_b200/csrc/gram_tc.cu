// gram_tc.cu — MDTA stage 1 on tcgen05: split-K Gram q.k^T over HW slabs with the squared L2 norms of q
// and k produced by the same tensor-core pass from the same shared-memory tiles.
// Reference: Attention.forward Restormer.py:117-124 (rearrange to heads, F.normalize, q @ k^T).
//
// In NCHW, q[h] and k[h] are [d, HW] with HW contiguous = K-major operands for a GEMM whose K is HW.
// One CTA = one (slab, head, sample).  Per 64-pixel chunk TMA brings q (d rows x 128 B) and k (d rows x 128 B)
// into one SWIZZLE_128B stage, q first then k.  The MMA A operand always spans 128 rows starting at a tile
// start (rows past the tile are whatever follows in shared memory; output rows are independent, so those
// rows of D are simply never read):
//     D1 = A(q|k|..) x B(k)^T : rows [0,d)  = q.k^T        rows [d,2d) = k.k^T (only if 2d <= 128)
//     D2 = A(q|k|..) x B(q)^T : rows [0,d)  = q.q^T  (diagonal = sum q^2)
//     D3 = A(k|....) x B(k)^T : rows [0,d)  = k.k^T  (only when 2d > 128)
// fp32 accumulation in TMEM over the whole slab; the epilogue writes the d x d partial and the two diagonals
// to partials[b][h][slab][d*d + 2d]; slabs are reduced in fixed order by the softmax kernel (deterministic).
#include <mutex>

#include "tc_common.cuh"

namespace b200ir {

struct GramTcParams {
  float* partials;
  int C, d, nslab;
  int64_t HW, slab_len;
  int S;          // pipeline stages
  int three;      // 1: needs D3 (2d > 128)
};

constexpr int kGramThreads = 192;  // warps 0-3 epilogue, 4 producer, 5 MMA

__global__ void __launch_bounds__(kGramThreads)
gram_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const GramTcParams P) {
  using namespace tc;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const int d = P.d;
  const uint32_t stage_bytes = 2u * d * 128u;
  uint8_t* ring = smem;
  // ring is followed by 16 KB of slack so that a 128-row A operand starting in the last stage stays in-bounds
  uint8_t* tail = ring + (size_t)P.S * stage_bytes + 16384;
  uint64_t* bar_full = (uint64_t*)tail;   // [8]
  uint64_t* bar_empty = bar_full + 8;     // [8]
  uint64_t* bar_acc = bar_empty + 8;      // [1]
  uint32_t* tmem_slot = (uint32_t*)(bar_acc + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int slab = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int64_t pbeg = (int64_t)slab * P.slab_len;
  const int64_t pend = min(P.HW, pbeg + P.slab_len);
  const int nchunks = (int)((pend - pbeg + 63) >> 6);
  const uint32_t ncols_alloc = P.three ? 512u : (d <= 64 ? 128u : 256u);
  // accumulator column offsets
  const uint32_t colD1 = 0, colD2 = P.three ? 128u : (d <= 64 ? 64u : 128u), colD3 = 256u;

  if (threadIdx.x == 0) {
    for (int i = 0; i < P.S; ++i) {
      mbar_init(&bar_full[i], 1);
      mbar_init(&bar_empty[i], 1);
    }
    mbar_init(bar_acc, 1);
    fence_barrier_init();
  }
  if (warp == 4) {
    if (lane == 0) tma_prefetch_desc(&tmQ);
    tmem_alloc(tmem_slot, ncols_alloc);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 4) {
    if (lane == 0) {
      for (int i = 0; i < nchunks; ++i) {
        const uint32_t stage = i % P.S, ph = (i / P.S) & 1;
        mbar_wait(&bar_empty[stage], ph ^ 1);
        mbar_arrive_expect_tx(&bar_full[stage], stage_bytes);
        uint8_t* dst = ring + (size_t)stage * stage_bytes;
        const int px = (int)(pbeg + (int64_t)i * 64);
        tma_load_3d(dst, &tmQ, &bar_full[stage], px, h * d, b);                    // q[h]
        tma_load_3d(dst + d * 128, &tmQ, &bar_full[stage], px, P.C + h * d, b);    // k[h]
      }
    }
  } else if (warp == 5) {
    {   // whole warp, convergent; the issuing lane is elected per instruction (tc_common.cuh)
      const uint32_t idesc = make_idesc_bf16(128, d, 0, 0);
      const uint32_t ring_addr = smem_u32(ring);
      for (int i = 0; i < nchunks; ++i) {
        const uint32_t stage = i % P.S, ph = (i / P.S) & 1;
        mbar_wait(&bar_full[stage], ph);
        tc_fence_after();
        const uint32_t qa = ring_addr + stage * stage_bytes, ka = qa + d * 128;
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          const uint64_t dq = make_smem_desc_sw128(qa + ks * 32, 16u, 1024u);
          const uint64_t dk = make_smem_desc_sw128(ka + ks * 32, 16u, 1024u);
          const uint32_t acc = (i | ks) != 0 ? 1u : 0u;
          umma_bf16_elect(tmem_base + colD1, dq, dk, idesc, acc);
          umma_bf16_elect(tmem_base + colD2, dq, dq, idesc, acc);
          if (P.three) umma_bf16_elect(tmem_base + colD3, dk, dk, idesc, acc);
        }
        umma_commit_elect(&bar_empty[stage]);
      }
      umma_commit_elect(bar_acc);
    }
  } else {
    // epilogue: thread <-> TMEM lane (row of D)
    const int row = warp * 32 + lane;
    float* out = P.partials + (((size_t)b * gridDim.y + h) * P.nslab + slab) * ((size_t)d * d + 2 * d);
    mbar_wait(bar_acc, 0);
    tc_fence_after();
    const uint32_t lane_base = tmem_base + ((uint32_t)(warp * 32) << 16);
    float nq = 0.f, nk = 0.f;
    // warps whose 32 rows are all >= the rows that matter skip the loads (warp-uniform conditions)
    const int rows_needed = P.three ? d : 2 * d;
    if (warp * 32 < rows_needed) {
      for (int c0 = 0; c0 < d; c0 += 16) {
        uint32_t r1[16], r2[16];
        tmem_ld_32x16(lane_base + colD1 + c0, r1);
        tmem_ld_32x16(lane_base + colD2 + c0, r2);
        tmem_ld_wait();
        if (row < d) {
          float4* o4 = reinterpret_cast<float4*>(out + (size_t)row * d + c0);  // d % 16 == 0 -> 16B aligned rows
#pragma unroll
          for (int j = 0; j < 4; ++j)
            o4[j] = make_float4(__uint_as_float(r1[4 * j]), __uint_as_float(r1[4 * j + 1]),
                                __uint_as_float(r1[4 * j + 2]), __uint_as_float(r1[4 * j + 3]));
#pragma unroll
          for (int j = 0; j < 16; ++j)
            if (c0 + j == row) nq = __uint_as_float(r2[j]);
        } else if (!P.three && row < 2 * d) {
#pragma unroll
          for (int j = 0; j < 16; ++j)
            if (c0 + j == row - d) nk = __uint_as_float(r1[j]);
        }
        if (P.three) {
          uint32_t r3[16];
          tmem_ld_32x16(lane_base + colD3 + c0, r3);
          tmem_ld_wait();
          if (row < d) {
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (c0 + j == row) nk = __uint_as_float(r3[j]);
          }
        }
      }
      if (row < d) out[(size_t)d * d + row] = nq;
      if (P.three) {
        if (row < d) out[(size_t)d * d + d + row] = nk;
      } else if (row >= d && row < 2 * d) {
        out[(size_t)d * d + d + (row - d)] = nk;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) tmem_dealloc(tmem_base, ncols_alloc);
}

bool gram_tc_supported(int dtype, int C, int heads, int64_t HW, int64_t bstride, const void* qkv) {
  const int d = C / heads;
  return dtype == B200IR_BF16 && d % 16 == 0 && d >= 16 && d <= 128 && HW % 8 == 0 && HW >= 64 && bstride % 8 == 0 &&
         ((uintptr_t)qkv % 16) == 0 && HW < (1ll << 31);
}

int gram_tc_launch(const void* qkv, int64_t bstride, int B, int C, int heads, int64_t HW, float* partials, int nslab,
                   int64_t slab_len, cudaStream_t st) {
  const int d = C / heads;
  CUtensorMap tm;
  const uint64_t dims[3] = {(uint64_t)HW, (uint64_t)(3 * C), (uint64_t)B};
  const uint64_t str[3] = {2, (uint64_t)HW * 2, (uint64_t)bstride * 2};
  const uint32_t box[3] = {64, (uint32_t)d, 1};
  int rc = make_tmap_bf16(&tm, qkv, 3, dims, str, box);
  if (rc) return rc;
  GramTcParams P;
  P.partials = partials;
  P.C = C; P.d = d; P.nslab = nslab; P.HW = HW; P.slab_len = slab_len;
  P.three = (2 * d > 128) ? 1 : 0;
  const size_t stage = 2 * (size_t)d * 128;
  // d <= 64: keep two CTAs per SM (<= ~100 KB each); larger heads take the whole SM
  const size_t budget = (d <= 64) ? 96 * 1024 : 200 * 1024;
  int S = (int)((budget - 16384 - 2048) / stage);
  if (S > 8) S = 8;
  if (S < 2) S = 2;
  P.S = S;
  const size_t smem = (size_t)S * stage + 16384 + 1024 + 1024;
  ensure_dyn_smem((const void*)gram_tc_kernel, 227 * 1024);
  gram_tc_kernel<<<dim3((unsigned)nslab, (unsigned)heads, (unsigned)B), kGramThreads, smem, st>>>(tm, P);
  return check_launch("mdta_gram_tc");
}

}  // namespace b200ir
