// gemm_tc.cu — LayerNorm + 1x1 conv (+bias, +residual) as a persistent tcgen05 GEMM, bf16 in HBM.
//
//   D[pixel, out_ch] = X^T[pixel, in_ch] . W^T[in_ch, out_ch]        (fp32 accumulators in TMEM)
//
// Reference ops replaced: LayerNorm.forward (Restormer.py:37-39,54-57,68-70) + the 1x1 nn.Conv2d that
// follows it (:114 qkv, :89 project_in), plain 1x1 convs (:92 project_out, :261/:266 reduce_chan), the
// folded attn.v+project_out (:127-131) with a per-sample weight, and the residual adds (:147-148).
//
// Layout / roles (320 threads, 1 CTA per SM, grid = (persistent CTAs, out-channel slices)):
//   A operand = activation tile, 128 pixels x BK channels, straight from NCHW by TMA: box {64 px, BK ch}
//               twice -> two MN-major SWIZZLE_128B blocks (pixels contiguous), ring of S stages.
//   B operand = the CTA's weight slice [Nslice x Kpad] bf16 (LayerNorm gamma pre-multiplied), K-major
//               SWIZZLE_128B, loaded ONCE by TMA and kept resident in shared memory (re-loaded only when
//               the sample changes for per-sample weights).
//   warp 4    = TMA producer; warp 5 = MMA issuer (one thread, tcgen05.mma M=128, N<=256, K=16);
//   warps 6-9 = LayerNorm statistics: read the activation tile from the swizzled smem stage while the MMA
//               consumes it (shifted single pass, fp32) -> (mu, rstd) per pixel;
//   warps 0-3 = epilogue: tcgen05.ld 32 lanes x 32 columns, y = rstd*(acc - mu*s[o]) + t[o] (+res), coalesced
//               NCHW stores (a warp writes 32 consecutive pixels of one channel).
//   LayerNorm is folded algebraically: W.LN(x) = rstd*(W.diag(g).x - mu*s) + t with s[o]=sum_c bf16(W g)[o,c],
//   t[o] = bias[o] + sum_c W[o,c] b[c], so the tensor core consumes the raw bf16 activations.
//   TMEM: 2 accumulator stages x 256 columns; the epilogue of chunk i overlaps the MMAs of chunk i+1.
#include <stdlib.h>

#include <mutex>
#include <type_traits>

#include "tc_common.cuh"

namespace b200ir {

// ---------------------------------------------------------------------------------------------
// host: driver entry point + tensor maps
// ---------------------------------------------------------------------------------------------
PFN_encodeTiled get_encode_tiled() {
  static PFN_encodeTiled fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (PFN_encodeTiled)p;
  });
  return fn;
}

int make_tmap_bf16(CUtensorMap* m, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                   const uint32_t* box, bool swizzle128) {
  return make_tmap_bf16_sw(m, base, rank, dims, strides_bytes, box, swizzle128 ? 128 : 0);
}

int make_tmap_bf16_sw(CUtensorMap* m, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                      const uint32_t* box, int swizzle_bytes) {
  PFN_encodeTiled enc = get_encode_tiled();
  if (!enc) {
    set_error("cuTensorMapEncodeTiled entry point unavailable");
    return B200IR_E_DRIVER;
  }
  cuuint64_t gd[5], gs[5];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) {
    gd[i] = dims[i];
    bx[i] = box[i];
    es[i] = 1;
    if (i > 0) gs[i - 1] = strides_bytes[i];
  }
  const CUtensorMapSwizzle sw = swizzle_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                                                    : (swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_NONE);
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), gd, gs, bx, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r == CUDA_ERROR_INVALID_CONTEXT || r == CUDA_ERROR_NOT_INITIALIZED) {
    // a driver-API call in a thread that has made no runtime-API call yet (the first node of a backward pass running in
    // an autograd worker thread): bind the primary context to this thread and try again
    cudaFree(nullptr);
    r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), gd, gs, bx, es,
            CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  }
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d): rank %d dims [%llu,%llu,%llu] box [%u,%u,%u]", (int)r, rank,
              (unsigned long long)dims[0], (unsigned long long)dims[1], (unsigned long long)(rank > 2 ? dims[2] : 0),
              box[0], box[1], rank > 2 ? box[2] : 0);
    return B200IR_E_DRIVER;
  }
  return 0;
}

// ---------------------------------------------------------------------------------------------
// geometry shared by the packer and the launcher
// ---------------------------------------------------------------------------------------------
struct GemmGeom {
  int Kpad, BK, kchunks, kpt, Nslice, nslices, Nrows, NC, nchunks, nbox, S;
  int gpw;              // 16-column groups per epilogue warp (TMA epilogue staging)
  size_t epi_warp_bytes;
  size_t smem;
};

static inline int ceil_to(int v, int m) { return (v + m - 1) / m * m; }

// Packing (Kpad, Nrows) depends only on (Cin, Cout, taps); the out-channel slicing may additionally depend on
// whether a residual tile is staged in shared memory (has_res) and whether the TMA-store epilogue is used.
bool gemm_tc_geometry(int Cin, int Cout, GemmGeom* g, int taps = 1, bool has_res = false, bool tma_epi = false) {
  if (Cin < 16 || Cout < 1) return false;
  // BK: channels per pipeline stage (multiple of 16, <= 64), least zero-padding first, then larger
  int best = 64;
  double beste = -1;
  for (int bk = 64; bk >= 16; bk -= 16) {
    if (bk > ceil_to(Cin, 16)) continue;
    double e = (double)Cin / (double)(ceil_to(Cin, bk));
    if (e > beste + 1e-9) beste = e, best = bk;
  }
  if (taps != 1) best = 64;  // conv: K-major activation rows are 64 channels (128 B) wide
  g->BK = best;
  g->kpt = (Cin + g->BK - 1) / g->BK;
  g->kchunks = taps * g->kpt;
  g->Kpad = taps == 1 ? ceil_to(Cin, 64) : ceil_to(taps * g->kpt * g->BK, 64);
  const int Npad = ceil_to(Cout, 16);
  g->Nrows = Npad;
  // shrink the resident weight slice until the activation ring has enough stages next to the epilogue staging
  for (int wkb = 112; wkb >= 16; wkb -= 16) {
    int nmax = (wkb * 1024) / (g->Kpad * 2) / 16 * 16;
    if (nmax > (taps == 1 ? 512 : 256)) nmax = (taps == 1 ? 512 : 256);
    if (tma_epi && has_res && nmax > 128) nmax = 128;
    if (nmax < 16) continue;
    g->nslices = (Npad + nmax - 1) / nmax;
    g->Nslice = ceil_to((Npad + g->nslices - 1) / g->nslices, 16);
    g->nchunks = (g->Nslice + 255) / 256;
    g->NC = ceil_to((g->Nslice + g->nchunks - 1) / g->nchunks, 16);
    // the 16-column groups of a chunk are dealt to 4 epilogue warps per TMEM quadrant: prefer chunks whose group count
    // is a multiple of 4 (288 = 192 + 96 -> busiest warp handles 5 groups per tile instead of 6 with 144 + 144)
    if (g->nchunks > 1 && ceil_to(g->NC, 64) <= 256 && ceil_to(g->NC, 64) * (g->nchunks - 1) < g->Nslice)
      g->NC = ceil_to(g->NC, 64);
    g->nbox = 16;
    for (int nb = 256; nb >= 16; nb -= 16)
      if (g->Nslice % nb == 0) {
        g->nbox = nb;
        break;
      }
    const size_t wbytes = (size_t)(g->Kpad / 64) * g->Nslice * 128;
    g->gpw = ((g->NC >> 4) + 3) / 4;
    g->epi_warp_bytes = tma_epi ? (size_t)g->gpw * 1024 * (has_res ? 3 : 1) : 0;
    const size_t fixed = 1024 + 16 * g->epi_warp_bytes;  // alignment slack + per-warp epilogue staging
    // static shared (barriers, LayerNorm stats, s/t vectors) is ~7 KB and counts against the 227 KB too
    const long avail = 227 * 1024 - 8 * 1024 - (long)wbytes - (long)fixed;
    int S = avail > 0 ? (int)(avail / 16384) : 0;
    if (S > 8) S = 8;
    g->S = S;
    g->smem = wbytes + (size_t)S * 16384 + fixed;
    const int need = g->nchunks > 1 ? g->kchunks + 1 : 3;
    if (S >= need) return true;
  }
  if (taps != 1) {
    // no resident weight slice fits (e.g. 768 -> 384: one 16-row slice is already 221 KB): the 3x3 path then streams
    // both operands (conv3x3_stream_tc.cu), which only needs the packing geometry (Kpad, Nrows, kpt)
    g->Nslice = 16;
    g->nslices = Npad / 16 > 2 ? Npad / 16 : 2;
    g->NC = 16; g->nchunks = 1; g->nbox = 16; g->S = 0; g->gpw = 1; g->epi_warp_bytes = 0; g->smem = 0;
    return true;
  }
  return false;
}

// geometry of the channels-on-lanes variant (MODE 4): same packing / slicing, but the out-channel slice of a CTA is cut
// in 128-row MMA tiles and 32-row store boxes (so slices must be multiples of 32 rows when there are several), and the
// epilogue staging is two 2 KB tiles per warp (bf16 output) instead of per-group buffers.
bool gemm_tc_geometry_t(int Cin, int Cout, bool out16, GemmGeom* g) {
  if (!gemm_tc_geometry(Cin, Cout, g, 1, false, false)) return false;
  if (g->nslices > 1 && g->Nslice % 32) {
    g->Nslice = ceil_to(g->Nslice, 32);
    g->nslices = (g->Nrows + g->Nslice - 1) / g->Nslice;
    g->nbox = 16;
    for (int nb = 256; nb >= 16; nb -= 16)
      if (g->Nslice % nb == 0) {
        g->nbox = nb;
        break;
      }
  }
  if (g->Nslice > 512) return false;
  const size_t wbytes = (size_t)(g->Kpad / 64) * g->Nslice * 128;
  g->epi_warp_bytes = out16 ? 4096 : 0;
  const size_t fixed = 1024 + 16 * g->epi_warp_bytes;
  const long avail = 227 * 1024 - 8 * 1024 - (long)wbytes - (long)fixed;
  int S = avail > 0 ? (int)(avail / 16384) : 0;
  if (S > 8) S = 8;
  if (S < 3) return false;
  g->S = S;
  g->smem = wbytes + (size_t)S * 16384 + fixed;
  return true;
}

// ---------------------------------------------------------------------------------------------
// weight packer: wp[b][Nrows][Kpad] bf16 = w[b][o][c] * ln_w[c];  st[0][o] = sum_c bf16(.) ; st[1][o] = bias + W.ln_b
// one warp per output row
// ---------------------------------------------------------------------------------------------
__global__ void pack_conv1x1_kernel(const float* __restrict__ w, int64_t w_bs, const float* __restrict__ bias,
                                    int ln_type, const float* __restrict__ ln_w, const float* __restrict__ ln_b,
                                    int Cin, int Cout, int Kpad, int Nrows, bf16* __restrict__ wp,
                                    float* __restrict__ st) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int b = blockIdx.y;
  if (row >= Nrows) return;
  bf16* dst = wp + ((size_t)b * Nrows + row) * Kpad;
  float s = 0.f, t = 0.f;
  if (row < Cout) {
    const float* src = w + (size_t)b * w_bs + (size_t)row * Cin;
    for (int c = lane; c < Kpad; c += 32) {
      float v = 0.f;
      if (c < Cin) {
        const float wv = src[c];
        v = ln_type != B200IR_LN_NONE ? wv * ln_w[c] : wv;
        if (ln_type == B200IR_LN_WITHBIAS) t = fmaf(wv, ln_b[c], t);
      }
      const bf16 q = __float2bfloat16_rn(v);
      s += __bfloat162float(q);
      dst[c] = q;
    }
  } else {
    for (int c = lane; c < Kpad; c += 32) dst[c] = __float2bfloat16_rn(0.f);
  }
  s = warp_sum(s);
  t = warp_sum(t);
  if (lane == 0 && b == 0) {
    st[row] = s;
    st[Nrows + row] = (row < Cout) ? t + (bias ? bias[row] : 0.f) : 0.f;
  }
}

// wp[c][o] = bf16(w[o][c]) for c < Cin (rows), o < Cout (k), zero padded to [Nrows][Kpad]; st = 0.
// 32x32 tiles through shared memory: coalesced fp32 reads along c, coalesced bf16 writes along o.
__global__ void __launch_bounds__(256)
pack_conv1x1_t_kernel(const float* __restrict__ w, int Cin, int Cout, int Kpad, int Nrows, bf16* __restrict__ wp,
                      float* __restrict__ st) {
  __shared__ float tile[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int k0 = blockIdx.x * 32, r0 = blockIdx.y * 32;   // k = o (source row), r = c (source column)
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int o = k0 + ty + 8 * i, c = r0 + tx;
    tile[ty + 8 * i][tx] = (o < Cout && c < Cin) ? w[(size_t)o * Cin + c] : 0.f;
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int r = r0 + ty + 8 * i, k = k0 + tx;
    if (r < Nrows && k < Kpad) wp[(size_t)r * Kpad + k] = __float2bfloat16_rn(tile[tx][ty + 8 * i]);
  }
  if (blockIdx.x == 0 && threadIdx.x < 32 && r0 + threadIdx.x < Nrows) {
    st[r0 + threadIdx.x] = 0.f;
    st[Nrows + r0 + threadIdx.x] = 0.f;
  }
}

// Every pack of a model in one launch: CTA -> descriptor by binary search over first_block, then the same per-row /
// per-tile work as pack_conv1x1_kernel / pack_conv1x1_t_kernel (bitwise the same results).
__global__ void __launch_bounds__(256)
pack_multi_kernel(const b200ir_pack_desc_t* __restrict__ descs, int ndesc) {
  __shared__ float tile[32][33];
  int lo = 0, hi = ndesc - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (descs[mid].first_block <= (int)blockIdx.x) lo = mid; else hi = mid - 1;
  }
  const b200ir_pack_desc_t d = descs[lo];
  const int blk = (int)blockIdx.x - d.first_block;
  bf16* wp = (bf16*)d.wp;
  if (d.transposed) {
    const int tiles_k = (d.Kpad + 31) / 32;
    const int k0 = (blk % tiles_k) * 32, r0 = (blk / tiles_k) * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int o = k0 + ty + 8 * i, c = r0 + tx;
      tile[ty + 8 * i][tx] = (o < d.Cout && c < d.Cin) ? d.w[(size_t)o * d.Cin + c] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int r = r0 + ty + 8 * i, k = k0 + tx;
      if (r < d.Nrows && k < d.Kpad) wp[(size_t)r * d.Kpad + k] = __float2bfloat16_rn(tile[tx][ty + 8 * i]);
    }
    if (k0 == 0 && threadIdx.x < 32 && r0 + threadIdx.x < d.Nrows) {
      d.st[r0 + threadIdx.x] = 0.f;
      d.st[d.Nrows + r0 + threadIdx.x] = 0.f;
    }
    return;
  }
  const int lane = threadIdx.x & 31;
  const int row = blk * 8 + (threadIdx.x >> 5);
  if (row >= d.Nrows) return;
  bf16* dst = wp + (size_t)row * d.Kpad;
  float s = 0.f, t = 0.f;
  if (row < d.Cout) {
    const float* src = d.w + (size_t)row * d.Cin;
    for (int c = lane; c < d.Kpad; c += 32) {
      float v = 0.f;
      if (c < d.Cin) {
        const float wv = src[c];
        v = d.ln_type != B200IR_LN_NONE ? wv * d.ln_w[c] : wv;
        if (d.ln_type == B200IR_LN_WITHBIAS) t = fmaf(wv, d.ln_b[c], t);
      }
      const bf16 q = __float2bfloat16_rn(v);
      s += __bfloat162float(q);
      dst[c] = q;
    }
  } else {
    for (int c = lane; c < d.Kpad; c += 32) dst[c] = __float2bfloat16_rn(0.f);
  }
  s = warp_sum(s);
  t = warp_sum(t);
  if (lane == 0) {
    d.st[row] = s;
    d.st[d.Nrows + row] = (row < d.Cout) ? t + (d.bias ? d.bias[row] : 0.f) : 0.f;
  }
}

}  // namespace b200ir

namespace b200ir {

struct GemmTcParams {
  void* y; int64_t y_bs; int y_dt;
  const void* res; int64_t res_bs; int res_dt;
  const float* s; const float* t;
  int ln_type, Cin, Cout;
  int64_t HW;
  int tiles_per_sample, total_tiles;
  int BK, kchunks, Nslice, NC, nchunks, nbox, kc64, S;
  int per_sample_w;
  float inv_cin;
  // 3x3 conv mode (MODE > 0): K loop = 9 taps x kpt channel chunks, A tiles come from tap-shifted 4D TMA boxes
  int Nrows;               // rows of the packed weight / length of s,t
  int gpw, epi_warp_bytes; // TMA epilogue: 16-col groups per warp, bytes of staging per warp
  int kpt;                 // channel chunks per tap (== kchunks for 1x1)
  int H, W, TW, TWs, TH, tiles_x, tiles_y;   // conv tile = TW x TH pixels (TW*TH = 128, TW = 1<<TWs)
  int dbg;                 // B200IR_GEMM_DBG (timing experiments only, wrong results): 1 no TMA stores, 2 no statistics, 4 no epilogue math
};

#ifdef B200IR_TIMELINE
// debug builds only (tools/ab): per-role clock64 timeline of CTA (0,0); one private region per role (plain stores)
__device__ unsigned long long g_tl[4 * 4096 * 2];
#define TL_DECL int tl_cnt = 0; const int tl_slot = (blockIdx.x | blockIdx.y) ? -1 : (warp == kWarpProducer ? 0 : warp == kWarpMma ? 1 : warp == kWarpStats0 ? 2 : warp == 0 ? 3 : -1);
#define TL(tag, id)                                                                                    \
  do {                                                                                                 \
    if (tl_slot >= 0 && tl_cnt < 4096) {                                                               \
      g_tl[(tl_slot * 4096 + tl_cnt) * 2] = ((unsigned long long)(tag) << 32) | (unsigned)(id);        \
      g_tl[(tl_slot * 4096 + tl_cnt) * 2 + 1] = (unsigned long long)clock64();                         \
      ++tl_cnt;                                                                                        \
    }                                                                                                  \
  } while (0)
#else
#define TL_DECL
#define TL(tag, id)
#endif

// timing experiments (tools/ab builds with -DB200IR_AB only; results with a bit set are WRONG by design): 1 no TMA stores,
// 2 no statistics, 4 no epilogue math, 8 no MMAs, 16 no activation loads.  The default library compiles them out.
#ifdef B200IR_AB
#define B200IR_DBG(bit) (P.dbg & (bit))
#else
#define B200IR_DBG(bit) false
#endif
constexpr int kStageBytes = 16384;  // 128 px x 64 ch x 2 B
constexpr int kEpiWarps = 16;       // 4 per TMEM lane quadrant
constexpr int kWarpProducer = 16, kWarpMma = 17, kWarpStats0 = 18;
constexpr int kThreads = (kWarpStats0 + 4) * 32;  // 704
constexpr int kFixedSmem = 8192;    // barriers, stats, s/t vectors

template <typename T> __device__ __forceinline__ float ldf(const T* p);
template <> __device__ __forceinline__ float ldf<bf16>(const bf16* p) { return __bfloat162float(*p); }
template <> __device__ __forceinline__ float ldf<float>(const float* p) { return *p; }
template <typename T> __device__ __forceinline__ void stf(T* p, float v);
template <> __device__ __forceinline__ void stf<bf16>(bf16* p, float v) { *p = __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ void stf<float>(float* p, float v) { *p = v; }

// TRES: void (no residual) | bf16 | float ; TOUT: bf16 | float
// MODE 0: 1x1 conv over flattened HW (A = NCHW activations, MN-major).
// MODE 1/2/3: 3x3 conv (pad 1) as 9 tap-shifted GEMMs accumulated in TMEM, store = plain / PixelUnshuffle(2) /
// PixelShuffle(2) (Restormer.py:171-189).  The activation operand is the NHWC copy of the input (channels
// contiguous = K-major): an M tile is a TW x TH pixel patch fetched by ONE 4D TMA box {64 ch, TW, TH, 1} at
// (x0+dx, y0+dy) -- only non-innermost coordinates are shifted (the innermost TMA coordinate must stay 16-byte
// aligned, measured with tools/tma4d_probe.cu) and the zero padding comes from TMA out-of-bounds fill.
template <typename TRES, typename TOUT, int MODE>
__global__ void __launch_bounds__(kThreads, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW,
               const __grid_constant__ CUtensorMap tmY, const __grid_constant__ CUtensorMap tmR,
               const GemmTcParams P) {
  using namespace tc;
  constexpr bool HAS_RES = !std::is_same<TRES, void>::value;
  constexpr bool ONE = MODE == 0 || MODE == 4;   // 1x1 conv over flattened HW (MODE 4: channels-on-lanes epilogue)
  typedef typename std::conditional<HAS_RES, TRES, float>::type TR;
  // bf16 NCHW output (and bf16 residual) of the 1x1 path goes through shared memory + TMA (see epilogue)
  constexpr bool TMA_EPI = MODE == 0 && std::is_same<TOUT, bf16>::value &&
                           (!HAS_RES || std::is_same<TRES, bf16>::value);
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t wbytes = (uint32_t)P.kc64 * P.Nslice * 128;
  uint8_t* Wsm = smem;
  uint8_t* ring = smem + wbytes;
  uint8_t* epi_smem = ring + (size_t)P.S * kStageBytes;   // [16 warps][epi_warp_bytes]
  __shared__ uint64_t bar_full[8], bar_empty[8], bar_wfull[1], bar_wempty[1], bar_accfull[4], bar_accempty[4],
      bar_stfull[2], bar_stempty[2], bar_res[2 * kEpiWarps];
  __shared__ uint32_t tmem_slot[1];
  __shared__ __align__(16) uint32_t s_bdesc[256];            // low words of the weight descriptors, (ch, kc, ks)
  __shared__ __align__(16) float s_mu[256], s_rstd[256];   // [2][128]: (-rstd*mu, rstd) per pixel, double buffered
  __shared__ __align__(16) float2 s_st[512];               // (s, t) of this CTA's out-channel slice

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool has_ln = P.ln_type != B200IR_LN_NONE;
  const int n_lo = blockIdx.y * P.Nslice;

  if (threadIdx.x == 0) {
    for (int i = 0; i < P.S; ++i) {
      mbar_init(&bar_full[i], 1);
      mbar_init(&bar_empty[i], has_ln ? 5 : 1);
    }
    mbar_init(bar_wfull, 1);
    mbar_init(bar_wempty, 1);
    for (int i = 0; i < 2 * kEpiWarps; ++i) mbar_init(&bar_res[i], 1);
    for (int i = 0; i < 4; ++i) {
      mbar_init(&bar_accfull[i], 1);
      mbar_init(&bar_accempty[i], kEpiWarps);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bar_stfull[i], 128);
      mbar_init(&bar_stempty[i], kEpiWarps);
    }
    fence_barrier_init();
  }
  for (int i = threadIdx.x; i < P.Nslice; i += kThreads)
    s_st[i] = (n_lo + i < P.Nrows) ? make_float2(P.s[n_lo + i], P.t[n_lo + i]) : make_float2(0.f, 0.f);
  if (warp == kWarpProducer) {
    if (lane == 0) {
      tma_prefetch_desc(&tmX);
      tma_prefetch_desc(&tmW);
      if (TMA_EPI || MODE == 4) tma_prefetch_desc(&tmY);
      if (TMA_EPI && HAS_RES) tma_prefetch_desc(&tmR);
    }
    tmem_alloc(tmem_slot, 512);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  TL_DECL

  // contiguous tile range of this CTA
  const int gx = gridDim.x;
  const int t_begin = (int)(((int64_t)P.total_tiles * blockIdx.x) / gx);
  const int t_end = (int)(((int64_t)P.total_tiles * (blockIdx.x + 1)) / gx);

  if (warp == kWarpProducer) {
    // ================= TMA producer =================
    if (lane == 0) {
      uint32_t it = 0, wload = 0;
      int cur = -1;
      for (int tile = t_begin; tile < t_end; ++tile) {
        const int b = ONE ? tile / P.tiles_per_sample : 0;
        const int p0 = (tile - b * P.tiles_per_sample) * 128;
        if (cur < 0 || (P.per_sample_w && b != cur)) {
          if (wload > 0) mbar_wait(bar_wempty, (wload - 1) & 1);
          mbar_arrive_expect_tx(bar_wfull, wbytes);
          for (int k64 = 0; k64 < P.kc64; ++k64)
            for (int r0 = 0; r0 < P.Nslice; r0 += P.nbox)
              tma_load_3d(Wsm + ((size_t)k64 * P.Nslice + r0) * 128, &tmW, bar_wfull, k64 * 64, n_lo + r0,
                          P.per_sample_w ? b : 0);
          ++wload;
          cur = b;
        }
        int cx0 = 0, cy0 = 0, cb = 0;
        if (!ONE) {
          const int tpi = P.tiles_x * P.tiles_y;
          cb = tile / tpi;
          const int r = tile - cb * tpi;
          cy0 = (r / P.tiles_x) * P.TH;
          cx0 = (r % P.tiles_x) * P.TW;
        }
        for (int kc = 0; kc < P.kchunks; ++kc, ++it) {
          const uint32_t stage = it % P.S, ph = (it / P.S) & 1;
          mbar_wait(&bar_empty[stage], ph ^ 1);
          TL(1, tile * 8 + kc);
          if (B200IR_DBG(16)) { mbar_arrive(&bar_full[stage]); continue; }
          mbar_arrive_expect_tx(&bar_full[stage], ONE ? 2u * P.BK * 128u : (uint32_t)kStageBytes);
          uint8_t* dst = ring + (size_t)stage * kStageBytes;
          if (ONE) {
            tma_load_3d(dst, &tmX, &bar_full[stage], p0, kc * P.BK, b);
            tma_load_3d(dst + P.BK * 128, &tmX, &bar_full[stage], p0 + 64, kc * P.BK, b);
            TL(2, tile * 8 + kc);
          } else {
            const int tap = kc / P.kpt, cc = kc - tap * P.kpt;
            const int dy = tap / 3 - 1, dx = tap - (tap / 3) * 3 - 1;
            tma_load_4d(dst, &tmX, &bar_full[stage], cc * 64, cx0 + dx, cy0 + dy, cb);
          }
        }
      }
    }
  } else if (warp == kWarpMma) {
    // ================= MMA issuer =================
    {   // whole warp, convergent: the issuing lane is elected per instruction (tc_common.cuh)
      uint32_t it = 0, acc_it = 0, wload = 0;
      int cur = -1;
      const uint32_t ring_addr = smem_u32(ring), w_addr = smem_u32(Wsm);
      if (MODE == 4) {
        // D[128 out-channels (lanes)][128 pixels (columns)] = W[128 x K] (K-major, resident) . X[K x 128 px] (MN-major
        // activation tile): four 128-column accumulators rotate over (pixel tile, channel tile) items.
        const int mtiles = (P.Nslice + 127) >> 7;
        const uint32_t idesc = make_idesc_bf16(128, 128, 0, 1);
        // same issue-loop form as below: weight descriptor low words tabulated once per CTA ((mt, kc, ks) -> s_bdesc), the
        // activation descriptor of a stage advances by a constant per K step, stage / phase are counters
        const int nent = mtiles * P.kchunks * 4;     // <= 256 (launcher)
        for (int e = lane; e < nent; e += 32) {
          const int ks = e & 3, kc = (e >> 2) % P.kchunks, mt = (e >> 2) / P.kchunks;
          const int kglob = kc * P.BK + ks * 16;
          s_bdesc[e] = (uint32_t)make_smem_desc_sw128(
              w_addr + (uint32_t)(kglob >> 6) * P.Nslice * 128u + (uint32_t)mt * 16384u + (kglob & 63) * 2u, 16u, 1024u);
        }
        __syncwarp();
        const uint32_t desc_hi = (uint32_t)(make_smem_desc_sw128(0u, 16u, 1024u) >> 32);
        const uint32_t x_lbo = (((uint32_t)P.BK * 128u) >> 4) << 16;
        uint32_t st0 = 0, ph0 = 0;
        for (int tile = t_begin; tile < t_end; ++tile) {
          const int b = tile / P.tiles_per_sample;
          if (cur < 0 || (P.per_sample_w && b != cur)) {
            mbar_wait(bar_wfull, wload & 1);
            ++wload;
            cur = b;
          }
          uint32_t stage = st0, ph = ph0;
          for (int mt = 0; mt < mtiles; ++mt, ++acc_it) {
            const uint32_t as = acc_it & 3, aph = (acc_it >> 2) & 1;
            mbar_wait(&bar_accempty[as], aph ^ 1);
            tc_fence_after();
            const uint32_t d_tmem = tmem_base + as * 128;
            const uint32_t* wtab = s_bdesc + mt * P.kchunks * 4;
            stage = st0; ph = ph0;
            for (int kc = 0; kc < P.kchunks; ++kc) {
              if (mt == 0) {
                mbar_wait(&bar_full[stage], ph);
                tc_fence_after();
              }
              const int rows = min(P.BK, P.Cin - kc * P.BK);
              const int ksteps = (rows + 15) >> 4;
              const uint4 wl = *reinterpret_cast<const uint4*>(wtab + kc * 4);
              const uint32_t wlo[4] = {wl.x, wl.y, wl.z, wl.w};
              uint32_t xlo = (((ring_addr + stage * kStageBytes) >> 4) & 0x3fffu) | x_lbo;
#pragma unroll
              for (int ks = 0; ks < 4; ++ks) {
                if (ks < ksteps)
                  umma_bf16_elect(d_tmem, ((uint64_t)desc_hi << 32) | wlo[ks], ((uint64_t)desc_hi << 32) | xlo, idesc,
                                  (kc | ks) != 0 ? 1u : 0u);
                xlo += 2048u >> 4;
              }
              if (mt == mtiles - 1) umma_commit_elect(&bar_empty[stage]);
              if (++stage == (uint32_t)P.S) { stage = 0; ph ^= 1u; }
            }
            umma_commit_elect(&bar_accfull[as]);
          }
          st0 = stage; ph0 = ph;
          if (P.per_sample_w && tile + 1 < t_end && (tile + 1) / P.tiles_per_sample != b) umma_commit_elect(bar_wempty);
        }
      } else
      {
        // Weight descriptors do not depend on the tile: their low words are tabulated once ((ch, kc, ks) -> s_bdesc) and
        // the activation descriptor of a stage advances by a constant per K step, so that one MMA issue is a table
        // load, two adds and the instruction (the issue loop was ~59 instructions per MMA = ~450 cycles of one warp's
        // dependent issue, tools/gemm_timeline.py, against 72..128 cycles of tensor-pipe time per MMA).
        const int nent = P.nchunks * P.kchunks * 4;
        for (int e = lane; e < nent; e += 32) {
          const int ks = e & 3, kc = (e >> 2) % P.kchunks, ch = (e >> 2) / P.kchunks;
          const int kglob = kc * P.BK + ks * 16;
          const uint32_t baddr =
              w_addr + (uint32_t)(kglob >> 6) * P.Nslice * 128u + (uint32_t)(ch * P.NC) * 128u + (kglob & 63) * 2u;
          s_bdesc[e] = (uint32_t)make_smem_desc_sw128(baddr, 16u, 1024u);
        }
        __syncwarp();
        const uint32_t desc_hi = (uint32_t)(make_smem_desc_sw128(0u, 16u, 1024u) >> 32);   // SBO, version, swizzle
        const uint32_t a_lbo = MODE == 0 ? (((uint32_t)P.BK * 128u) >> 4) << 16 : 1u << 16;
        const uint32_t a_step = MODE == 0 ? 2048u >> 4 : 32u >> 4;
        uint32_t st0 = 0, ph0 = 0;   // ring position of the tile's first K chunk
        for (int tile = t_begin; tile < t_end; ++tile) {
          const int b = MODE == 0 ? tile / P.tiles_per_sample : 0;
          if (cur < 0 || (P.per_sample_w && b != cur)) {
            mbar_wait(bar_wfull, wload & 1);
            ++wload;
            cur = b;
          }
          uint32_t stage = st0, ph = ph0;
          for (int ch = 0; ch < P.nchunks; ++ch, ++acc_it) {
            const uint32_t as = acc_it & 1, aph = (acc_it >> 1) & 1;
            mbar_wait(&bar_accempty[as], aph ^ 1);
            TL(10, tile * 8 + ch);
            tc_fence_after();
            const uint32_t d_tmem = tmem_base + as * 256;
            const int ncols = min(P.NC, P.Nslice - ch * P.NC);
            const uint32_t idesc = make_idesc_bf16(128, ncols, MODE == 0 ? 1 : 0, 0);
            const uint32_t* btab = s_bdesc + ch * P.kchunks * 4;
            stage = st0; ph = ph0;
            int kcc = 0;   // channel chunk within the tap (MODE > 0), == kc for 1x1
            for (int kc = 0; kc < P.kchunks; ++kc) {
              if (ch == 0) {
                mbar_wait(&bar_full[stage], ph);
                TL(11, tile * 8 + kc);
                tc_fence_after();
              }
              const int rows = min(P.BK, P.Cin - kcc * P.BK);
              const int ksteps = (rows + 15) >> 4;
              const uint4 bl = *reinterpret_cast<const uint4*>(btab + kc * 4);
              const uint32_t blo[4] = {bl.x, bl.y, bl.z, bl.w};
              uint32_t alo = (((ring_addr + stage * kStageBytes) >> 4) & 0x3fffu) | a_lbo;
#pragma unroll
              for (int ks = 0; ks < 4; ++ks) {
                if (ks < ksteps && !B200IR_DBG(8))
                  umma_bf16_elect(d_tmem, ((uint64_t)desc_hi << 32) | alo, ((uint64_t)desc_hi << 32) | blo[ks], idesc,
                                  (kc | ks) != 0 ? 1u : 0u);
                alo += a_step;
              }
              if (ch == P.nchunks - 1) umma_commit_elect(&bar_empty[stage]);
              if (++stage == (uint32_t)P.S) { stage = 0; ph ^= 1u; }
              if (MODE != 0 && ++kcc == P.kpt) kcc = 0;
              if (MODE == 0) ++kcc;
            }
            umma_commit_elect(&bar_accfull[as]);
            TL(12, tile * 8 + ch);
          }
          st0 = stage; ph0 = ph;
          if (MODE == 0 && P.per_sample_w && tile + 1 < t_end && (tile + 1) / P.tiles_per_sample != b)
            umma_commit_elect(bar_wempty);
        }
      }
    }
  } else if (warp >= kWarpStats0) {
    // ================= LayerNorm statistics =================
    if (has_ln) {
      const int m = (warp - kWarpStats0) * 32 + lane;
      const uint32_t ring_addr = smem_u32(ring);
      uint32_t it = 0, tcount = 0;
      for (int tile = t_begin; tile < t_end; ++tile, ++tcount) {
        LnStatAcc acc;
        for (int kc = 0; kc < P.kchunks; ++kc, ++it) {
          const uint32_t stage = it % P.S, ph = (it / P.S) & 1;
          mbar_wait(&bar_full[stage], ph);
          if (warp == kWarpStats0 && lane == 0) TL(20, tile * 8 + kc);
          const int rows = min(P.BK, P.Cin - kc * P.BK);
          if (!B200IR_DBG(2) || kc == 0) ln_stat_chunk(ring_addr + stage * kStageBytes, (uint32_t)P.BK * 128u, m, B200IR_DBG(2) ? 0 : rows, kc == 0, acc);
          __syncwarp();
          if (lane == 0) mbar_arrive(&bar_empty[stage]);
        }
        float x0, sum, sq;
        ln_stat_finish(acc, m, x0, sum, sq);
        const float md = sum * P.inv_cin;
        const float var = fmaxf(sq * P.inv_cin - md * md, 0.f);
        const uint32_t sp = tcount & 1, sph = (tcount >> 1) & 1;
        if (warp == kWarpStats0 && lane == 0) TL(21, tile * 8);
        mbar_wait(&bar_stempty[sp], sph ^ 1);
        if (warp == kWarpStats0 && lane == 0) TL(22, tile * 8);
        const float rstd = 1.0f / sqrtf(var + 1e-5f);
        // epilogue form: y = acc*rstd + (-rstd*mu)*s[o] + t[o]
        s_mu[sp * 128 + m] = (P.ln_type == B200IR_LN_WITHBIAS) ? -rstd * (x0 + md) : 0.f;
        s_rstd[sp * 128 + m] = rstd;
        mbar_arrive(&bar_stfull[sp]);
      }
    }
  } else if (MODE == 4) {
    // ===== epilogue, channels on lanes: warp w <-> TMEM lane quadrant w%4 (32 out-channels) x pixel columns
    // [32*(w/4), +32).  A thread owns 32 consecutive pixels of ONE output channel: (s,t) live in registers, the per-pixel
    // (-rstd*mu, rstd) come as broadcast LDS.128, results are packed pairwise to bf16 and written as four 16-byte
    // chunks into the warp's private [32 ch][64 B] staging tile (SWIZZLE_64B pattern: conflict-free), which one TMA
    // store box {32 px, 32 ch} sends to the NCHW output (rows / pixels past the tensor are clipped by TMA).
    const int quad = warp & 3, wq = warp >> 2;
    const int mtiles = (P.Nslice + 127) >> 7;
    constexpr bool OUT16 = std::is_same<TOUT, bf16>::value;
    TOUT* yout = reinterpret_cast<TOUT*>(P.y);
    const TR* rin = reinterpret_cast<const TR*>(P.res);
    uint8_t* stg = epi_smem + (size_t)warp * P.epi_warp_bytes;   // [2][32 ch][64 B]
    uint32_t acc_it = 0, tcount = 0, st_it = 0;
    for (int tile = t_begin; tile < t_end; ++tile, ++tcount) {
      const int b = tile / P.tiles_per_sample;
      const int64_t p0 = (int64_t)(tile - b * P.tiles_per_sample) * 128 + wq * 32;
      const uint32_t sp = tcount & 1, sph = (tcount >> 1) & 1;
      if (has_ln) mbar_wait(&bar_stfull[sp], sph);
      const float4* nm4 = reinterpret_cast<const float4*>(&s_mu[sp * 128 + wq * 32]);
      const float4* rs4 = reinterpret_cast<const float4*>(&s_rstd[sp * 128 + wq * 32]);
      for (int mt = 0; mt < mtiles; ++mt, ++acc_it) {
        const uint32_t as = acc_it & 3, aph = (acc_it >> 2) & 1;
        const int rowbase = mt * 128 + quad * 32;              // rows of this CTA's out-channel slice
        const int row = rowbase + lane, co = n_lo + row;
        const bool wok = rowbase < P.Nslice && n_lo + rowbase < P.Cout;   // warp-uniform
        const bool cok = row < P.Nslice && co < P.Cout;
        const float2 stv = cok ? s_st[row] : make_float2(0.f, 0.f);
        mbar_wait(&bar_accfull[as], aph);
        tc_fence_after();
        const uint32_t tb = tmem_base + as * 128 + wq * 32 + ((uint32_t)(quad * 32) << 16);
        uint8_t* sbuf = stg + (st_it & 1) * 2048;
        if (OUT16 && wok) {
          if (lane == 0) bulk_wait_read<1>();   // the store that used this buffer two items ago has read it
          __syncwarp();
        }
        if (wok) {
          uint32_t r[32];
          tmem_ld_32x16(tb, r);             // both halves in flight before the single wait
          tmem_ld_32x16(tb + 16, r + 16);
          tmem_ld_wait();
          tc_fence_before();                 // accumulator is in registers: hand the TMEM slot back right away
          __syncwarp();
          if (lane == 0) mbar_arrive(&bar_accempty[as]);
#pragma unroll
          for (int g8 = 0; g8 < 4; ++g8) {   // 8 pixels at a time: few live registers
            float v[8];
            if (has_ln) {
              const float4 a0 = nm4[g8 * 2], a1 = nm4[g8 * 2 + 1], c0 = rs4[g8 * 2], c1 = rs4[g8 * 2 + 1];
              const float nmv[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
              const float rsv[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
#pragma unroll
              for (int j = 0; j < 8; ++j)
                v[j] = fmaf(__uint_as_float(r[g8 * 8 + j]), rsv[j], fmaf(nmv[j], stv.x, stv.y));
            } else {
#pragma unroll
              for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(r[g8 * 8 + j]) + stv.y;
            }
            const int64_t pp = p0 + g8 * 8;
            const bool pok = pp < P.HW;          // HW % 8 == 0: an 8-pixel group is entirely in or out
            if (HAS_RES && cok && pok) {
              const TR* rp = rin + (int64_t)b * P.res_bs + (int64_t)co * P.HW + pp;
              if (std::is_same<TR, bf16>::value) {
                const uint4 u = __ldg(reinterpret_cast<const uint4*>(rp));
                const uint32_t uu[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                  v[2 * k] += __uint_as_float(uu[k] << 16);
                  v[2 * k + 1] += __uint_as_float(uu[k] & 0xffff0000u);
                }
              } else {
#pragma unroll
                for (int k = 0; k < 8; ++k) v[k] += ldf<TR>(rp + k);
              }
            }
            if (OUT16) {
              uint32_t pk[4];
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * k], v[2 * k + 1]);
                pk[k] = *reinterpret_cast<uint32_t*>(&h);
              }
              const uint32_t chunk = (uint32_t)g8 ^ (((uint32_t)lane >> 1) & 3u);
              *reinterpret_cast<uint4*>(sbuf + lane * 64 + chunk * 16) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
            } else if (cok && pok) {
              float4* d4 = reinterpret_cast<float4*>(yout + (int64_t)b * P.y_bs + (int64_t)co * P.HW + pp);
              d4[0] = make_float4(v[0], v[1], v[2], v[3]);
              d4[1] = make_float4(v[4], v[5], v[6], v[7]);
            }
          }
        }
        if (OUT16 && wok) fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
          if (!wok) mbar_arrive(&bar_accempty[as]);
          if (OUT16 && wok) {
            tma_store_3d(&tmY, sbuf, (int)p0, n_lo + rowbase, b);
            bulk_commit();
          }
        }
        if (OUT16 && wok) ++st_it;
      }
      if (has_ln) {
        __syncwarp();
        if (lane == 0) mbar_arrive(&bar_stempty[sp]);
      }
    }
    if (lane == 0) bulk_wait<0>();
  } else if (TMA_EPI) {
    // ===== epilogue (bf16 NCHW via TMA): warp w <-> TMEM lane quadrant w%4 and a contiguous quarter of the chunk's
    // 16-column groups.  Each warp is self-contained: tcgen05.ld -> fp32 math -> bf16 into its private staging
    // [group][16 ch][32 px] -> one TMA store box {32 px, 16 ch} per group.  The residual tile of the NEXT
    // (tile, chunk) is TMA-loaded into the warp's private double buffer while the current one is processed.
    const int quad = warp & 3, wq = warp >> 2;
    const int m = quad * 32 + lane;
    uint8_t* stg = epi_smem + (size_t)warp * P.epi_warp_bytes;          // [gpw][16][64 B]
    uint8_t* resb = stg + (size_t)P.gpw * 1024;                          // [2][gpw][16][64 B]
    uint64_t* rbar = &bar_res[warp * 2];
    auto groups_of = [&](int ch, int& g_lo, int& g_hi) {
      const int G = min(P.NC, P.Nslice - ch * P.NC) >> 4;
      g_lo = (wq * G) >> 2;
      g_hi = ((wq + 1) * G) >> 2;
    };
    auto issue_res = [&](int tile, int ch, uint32_t buf) {   // lane 0 only
      int g_lo, g_hi;
      groups_of(ch, g_lo, g_hi);
      if (g_hi <= g_lo) return;
      const int b = tile / P.tiles_per_sample;
      const int p0 = (tile - b * P.tiles_per_sample) * 128 + quad * 32;
      mbar_arrive_expect_tx(&rbar[buf], (uint32_t)(g_hi - g_lo) * 1024u);
      for (int g = g_lo; g < g_hi; ++g)
        tma_load_3d(resb + ((size_t)buf * P.gpw + (g - g_lo)) * 1024, &tmR, &rbar[buf], p0,
                    n_lo + ch * P.NC + g * 16, b);
    };
    uint32_t acc_it = 0, tcount = 0;
    if (HAS_RES && lane == 0 && t_begin < t_end) issue_res(t_begin, 0, 0);
    for (int tile = t_begin; tile < t_end; ++tile, ++tcount) {
      const int b = tile / P.tiles_per_sample;
      const int p0 = (tile - b * P.tiles_per_sample) * 128 + quad * 32;
      float nmr = 0.f, rstd = 1.f;  // (-rstd*mu, rstd)
      for (int ch = 0; ch < P.nchunks; ++ch, ++acc_it) {
        int g_lo, g_hi;
        groups_of(ch, g_lo, g_hi);
        const int nbase = ch * P.NC;
        if (HAS_RES && lane == 0) {  // prefetch the next item's residual
          int ntile = tile, nch = ch + 1;
          if (nch == P.nchunks) nch = 0, ++ntile;
          if (ntile < t_end) {
            fence_proxy_async();
            issue_res(ntile, nch, (acc_it + 1) & 1);
          }
        }
        if (ch == 0 && has_ln) {
          const uint32_t sp = tcount & 1, sph = (tcount >> 1) & 1;
          if (warp == 0 && lane == 0) TL(29, tile * 8 + ch);
          mbar_wait(&bar_stfull[sp], sph);
          if (warp == 0 && lane == 0) TL(30, tile * 8 + ch);
          nmr = s_mu[sp * 128 + m];
          rstd = s_rstd[sp * 128 + m];
          __syncwarp();
          if (lane == 0) mbar_arrive(&bar_stempty[sp]);
        }
        const uint32_t as = acc_it & 1, aph = (acc_it >> 1) & 1;
        mbar_wait(&bar_accfull[as], aph);
        if (warp == 0 && lane == 0) TL(31, tile * 8 + ch);
        tc_fence_after();
        if (elect_one()) bulk_wait_read<0>();   // the stores of the previous item have drained the staging buffer
        __syncwarp();
        if (warp == 0 && lane == 0) TL(32, tile * 8 + ch);
        if (HAS_RES && g_hi > g_lo) mbar_wait(&rbar[acc_it & 1], (acc_it >> 1) & 1);
        const uint32_t tbase = tmem_base + as * 256 + ((uint32_t)(quad * 32) << 16);
        const uint32_t rsrc = smem_u32(resb) + (uint32_t)(acc_it & 1) * P.gpw * 1024u + lane * 2u;
        const uint32_t sdst = smem_u32(stg) + lane * 2u;
        for (int g = g_lo; g < g_hi && !B200IR_DBG(4); ++g) {
          uint32_t r[16];
#ifdef B200IR_EPI_NO_TMEMLD   // A/B timing builds only (wrong results): which part of the epilogue slows the tensor pipe
#pragma unroll
          for (int j = 0; j < 16; ++j) r[j] = (uint32_t)(g + j);
#else
          tmem_ld_32x16(tbase + g * 16, r);
#endif
          const float4* stp = reinterpret_cast<const float4*>(&s_st[nbase + g * 16]);  // (s,t) pairs of 2 columns
          float4 st4[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) st4[j] = stp[j];
          float rr[16];
          if (HAS_RES) {
#pragma unroll
            for (int j = 0; j < 16; ++j) rr[j] = ld_shared_bf16(rsrc + (uint32_t)(g - g_lo) * 1024u + j * 64u);
          }
          tmem_ld_wait();
          const uint32_t d0 = sdst + (uint32_t)(g - g_lo) * 1024u;
          // two copies of the loop instead of a per-element select: with the select the compiler issued BOTH arithmetic
          // forms under opposite predicates, and these kernels are issue-bound (ncu r1n: 74-79 % issue slots busy)
#define B200IR_EPI_ELEMS(EXPR)                                                       \
  _Pragma("unroll") for (int j = 0; j < 16; ++j) {                                   \
    const float sx = (j & 1) ? st4[j >> 1].z : st4[j >> 1].x;                         \
    const float ty = (j & 1) ? st4[j >> 1].w : st4[j >> 1].y;                         \
    float v = (EXPR);                                                                 \
    (void)sx;                                                                         \
    if (HAS_RES) v += rr[j];                                                          \
    B200IR_EPI_STORE(d0 + j * 64u, v);                                                \
  }
#ifdef B200IR_EPI_NO_STS
#define B200IR_EPI_STORE(a, v) if ((v) == 1.2345e30f) st_shared_bf16((a), (v))
#else
#define B200IR_EPI_STORE(a, v) st_shared_bf16((a), (v))
#endif
          if (HAS_RES) {   // residual variants: the split costs registers (spills at 704 threads), keep the select
            B200IR_EPI_ELEMS(has_ln ? fmaf(__uint_as_float(r[j]), rstd, fmaf(nmr, sx, ty)) : __uint_as_float(r[j]) + ty)
          } else if (has_ln) {
            B200IR_EPI_ELEMS(fmaf(__uint_as_float(r[j]), rstd, fmaf(nmr, sx, ty)))
          } else {
            B200IR_EPI_ELEMS(__uint_as_float(r[j]) + ty)
          }
#undef B200IR_EPI_ELEMS
#undef B200IR_EPI_STORE
        }
        if (warp == 0 && lane == 0) TL(34, tile * 8 + ch);
        tc_fence_before();
        fence_proxy_async();
        __syncwarp();
        // convergent issue, one elected lane per instruction (the same lane every time: bulk groups are per thread);
        // under `if (lane == 0)` every UTMASTG sat in an R2UR / branch waterfall (~200 cycles per store)
        if (elect_one()) mbar_arrive(&bar_accempty[as]);
        if (warp == 0) TL(35, tile * 8 + ch);
        if (!B200IR_DBG(1))
          for (int g = g_lo; g < g_hi; ++g)
            if (elect_one()) tma_store_3d(&tmY, stg + (size_t)(g - g_lo) * 1024, p0, n_lo + nbase + g * 16, b);
        if (elect_one()) bulk_commit();
        if (warp == 0) TL(36, tile * 8 + ch);
      }
    }
    if (elect_one()) bulk_wait<0>();
  } else {
    // ===== epilogue: warp w <-> TMEM lane quadrant w%4, 16-column groups g = w/4, w/4+4, ... =====
    const int quad = warp & 3, wq = warp >> 2;
    const int m = quad * 32 + lane;
    uint32_t acc_it = 0, tcount = 0;
    for (int tile = t_begin; tile < t_end; ++tile, ++tcount) {
      int b;
      int64_t p;
      bool valid;
      int py = 0, px = 0;
      if (MODE == 0) {
        b = tile / P.tiles_per_sample;
        p = (int64_t)(tile - b * P.tiles_per_sample) * 128 + m;
        valid = p < P.HW;
      } else {
        const int tpi = P.tiles_x * P.tiles_y;
        b = tile / tpi;
        const int r = tile - b * tpi;
        py = (r / P.tiles_x) * P.TH + (m >> P.TWs);
        px = (r % P.tiles_x) * P.TW + (m & (P.TW - 1));
        valid = py < P.H && px < P.W;
        p = (int64_t)py * P.W + px;
      }
      TOUT* yb = (TOUT*)P.y + (int64_t)b * P.y_bs + (MODE <= 1 ? p : 0);
      const TR* rb = HAS_RES ? (const TR*)P.res + (int64_t)b * P.res_bs + p : nullptr;
      float nmr = 0.f, rstd = 1.f;  // (-rstd*mu, rstd)
      for (int ch = 0; ch < P.nchunks; ++ch, ++acc_it) {
        const int ncols = min(P.NC, P.Nslice - ch * P.NC);
        const int nbase = ch * P.NC;  // within the slice
        const int ngroups = ncols >> 4;
        // residual prefetch for this warp's first group (independent of the MMA)
        float rv[16];
        if (HAS_RES) {
          const int c0 = wq * 16;
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int n = n_lo + nbase + c0 + j;
            rv[j] = (valid && wq < ngroups && n < P.Cout) ? ldf<TR>(rb + (int64_t)n * P.HW) : 0.f;
          }
        }
        if (ch == 0 && has_ln) {
          const uint32_t sp = tcount & 1, sph = (tcount >> 1) & 1;
          mbar_wait(&bar_stfull[sp], sph);
          nmr = s_mu[sp * 128 + m];
          rstd = s_rstd[sp * 128 + m];
          __syncwarp();
          if (lane == 0) mbar_arrive(&bar_stempty[sp]);
        }
        const uint32_t as = acc_it & 1, aph = (acc_it >> 1) & 1;
        mbar_wait(&bar_accfull[as], aph);
        tc_fence_after();
        const uint32_t tbase = tmem_base + as * 256 + ((uint32_t)(quad * 32) << 16);
        for (int g = wq; g < ngroups; g += 4) {
          const int c0 = g * 16;
          uint32_t r[16];
          tmem_ld_32x16(tbase + c0, r);
          float rn[16];
          if (HAS_RES) {  // prefetch the next group's residuals while this one is processed
            const int c1 = c0 + 64;
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const int n = n_lo + nbase + c1 + j;
              rn[j] = (valid && g + 4 < ngroups && n < P.Cout) ? ldf<TR>(rb + (int64_t)n * P.HW) : 0.f;
            }
          }
          tmem_ld_wait();
          if (valid) {
            const int n0 = n_lo + nbase + c0;
            const int nlim = P.Cout - n0;
            if (MODE <= 1) {
              TOUT* yp = yb + (int64_t)n0 * P.HW;
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                if (j < nlim) {
                  const float2 st = s_st[nbase + c0 + j];
                  float v = fmaf(__uint_as_float(r[j]), rstd, fmaf(nmr, st.x, st.y));
                  if (HAS_RES) v += rv[j];
                  stf<TOUT>(yp + (int64_t)j * P.HW, v);
                }
              }
            } else if (MODE == 2) {
              // PixelUnshuffle(2): out[n*4 + (y&1)*2 + (x&1), y/2, x/2]
              const int64_t hwq = P.HW >> 2;
              TOUT* yp = yb + (int64_t)(n0 * 4 + (py & 1) * 2 + (px & 1)) * hwq + (int64_t)(py >> 1) * (P.W >> 1) + (px >> 1);
#pragma unroll
              for (int j = 0; j < 16; ++j)
                if (j < nlim) stf<TOUT>(yp + (int64_t)j * 4 * hwq, __uint_as_float(r[j]) + s_st[nbase + c0 + j].y);
            } else {
              // PixelShuffle(2): out[n/4, 2y + (n/2)%2, 2x + n%2]; channels (n, n+1) are horizontal neighbours
              const int64_t hw4 = P.HW << 2;
              const int W2 = P.W * 2;
#pragma unroll
              for (int j = 0; j < 16; j += 2) {
                if (j < nlim) {
                  const int n = n0 + j;
                  TOUT* yp = yb + (int64_t)(n >> 2) * hw4 + (int64_t)(2 * py + ((n >> 1) & 1)) * W2 + 2 * px;
                  const float v0 = __uint_as_float(r[j]) + s_st[nbase + c0 + j].y;
                  const float v1 = __uint_as_float(r[j + 1]) + s_st[nbase + c0 + j + 1].y;
                  if (std::is_same<TOUT, bf16>::value) {
                    *reinterpret_cast<__nv_bfloat162*>(yp) = __floats2bfloat162_rn(v0, v1);
                  } else {
                    *reinterpret_cast<float2*>(yp) = make_float2(v0, v1);
                  }
                }
              }
            }
          }
          if (HAS_RES) {
#pragma unroll
            for (int j = 0; j < 16; ++j) rv[j] = rn[j];
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bar_accempty[as]);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == kWarpProducer) tmem_dealloc(tmem_base, 512);
}

}  // namespace b200ir

using namespace b200ir;

extern "C" int b200ir_gemm_tc_geometry(int Cin, int Cout, int* kpad, int* nrows) {
  GemmGeom g;
  if (!gemm_tc_geometry(Cin, Cout, &g)) {
    set_error("gemm_tc: unsupported Cin=%d Cout=%d", Cin, Cout);
    return B200IR_E_UNSUPPORTED;
  }
  if (kpad) *kpad = g.Kpad;
  if (nrows) *nrows = g.Nrows;
  return 0;
}

extern "C" int b200ir_pack_conv1x1(const float* w, int64_t w_bstride, int Bw, const float* bias, int ln_type,
                                   const float* ln_w, const float* ln_b, int Cin, int Cout, void* wp, float* st,
                                   b200ir_stream_t stream) {
  B200IR_CHECK_ARG(w && wp && st && Bw > 0 && Bw <= 65535, "pack_conv1x1: bad args");
  B200IR_CHECK_ARG(ln_type == B200IR_LN_NONE || ln_w, "pack_conv1x1: ln_w missing");
  B200IR_CHECK_ARG(ln_type != B200IR_LN_WITHBIAS || ln_b, "pack_conv1x1: ln_b missing");
  GemmGeom g;
  B200IR_UNSUPPORTED(gemm_tc_geometry(Cin, Cout, &g), "pack_conv1x1: unsupported Cin=%d Cout=%d", Cin, Cout);
  dim3 grid((unsigned)((g.Nrows + 7) / 8), (unsigned)Bw);
  pack_conv1x1_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(w, w_bstride, bias, ln_type, ln_w, ln_b, Cin, Cout,
                                                              g.Kpad, g.Nrows, (bf16*)wp, st);
  return check_launch("pack_conv1x1");
}

extern "C" int b200ir_pack_conv1x1_multi(const b200ir_pack_desc_t* descs, int ndesc, int total_blocks,
                                         b200ir_stream_t stream) {
  B200IR_CHECK_ARG(descs && ndesc > 0 && total_blocks > 0, "pack_conv1x1_multi: bad args");
  pack_multi_kernel<<<(unsigned)total_blocks, 256, 0, (cudaStream_t)stream>>>(descs, ndesc);
  return check_launch("pack_conv1x1_multi");
}

extern "C" int b200ir_pack_conv1x1_t(const float* w, int Cin, int Cout, void* wp, float* st, b200ir_stream_t stream) {
  B200IR_CHECK_ARG(w && wp && st && Cin > 0 && Cout > 0, "pack_conv1x1_t: bad args");
  GemmGeom g;
  B200IR_UNSUPPORTED(gemm_tc_geometry(Cout, Cin, &g), "pack_conv1x1_t: unsupported Cin=%d Cout=%d", Cin, Cout);
  dim3 grid((unsigned)((g.Kpad + 31) / 32), (unsigned)((g.Nrows + 31) / 32));
  pack_conv1x1_t_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(w, Cin, Cout, g.Kpad, g.Nrows, (bf16*)wp, st);
  return check_launch("pack_conv1x1_t");
}

#ifdef B200IR_TIMELINE
extern "C" int b200ir_debug_timeline(unsigned long long* out, int max_pairs) {
  cudaDeviceSynchronize();
  const int n = 4 * 4096;
  if (out && max_pairs >= n) cudaMemcpyFromSymbol(out, b200ir::g_tl, (size_t)n * 16);
  void* p = nullptr;
  cudaGetSymbolAddress(&p, b200ir::g_tl);
  cudaMemset(p, 0, (size_t)n * 16);
  return n;
}
#endif

namespace b200ir {
int gemm_stream_launch(const void* x, int64_t x_bs, const void* wp, int Bw, const float* st, int ln_type,
                       const void* res, int64_t res_bs, int res_dt, void* y, int64_t y_bs, int y_dt, int B, int Cin,
                       int Cout, int64_t HW, int Kpad, int Nrows, cudaStream_t stream);  // gemm_stream_tc.cu
}

extern "C" int b200ir_gemm_tc_fwd(const void* x, int64_t x_bstride, const void* wp, int Bw, const float* st,
                                  int ln_type, const void* res, int64_t res_bstride, int res_dtype, void* y,
                                  int64_t y_bstride, int y_dtype, int B, int Cin, int Cout, int64_t HW,
                                  b200ir_stream_t stream) {
  return b200ir_gemm_tc_fwd_ex(x, x_bstride, wp, Bw, st, ln_type, res, res_bstride, res_dtype, y, y_bstride, y_dtype, B,
                               Cin, Cout, HW, B200IR_GEMM_AUTO, stream);
}

extern "C" int b200ir_gemm_tc_fwd_ex(const void* x, int64_t x_bstride, const void* wp, int Bw, const float* st,
                                     int ln_type, const void* res, int64_t res_bstride, int res_dtype, void* y,
                                     int64_t y_bstride, int y_dtype, int B, int Cin, int Cout, int64_t HW, int variant,
                                     b200ir_stream_t stream) {
  B200IR_CHECK_ARG(variant >= B200IR_GEMM_AUTO && variant <= B200IR_GEMM_RESIDENT_CH, "gemm_tc: bad variant");
  B200IR_CHECK_ARG(x && wp && st && y && B > 0 && HW > 0, "gemm_tc: bad args");
  B200IR_CHECK_ARG(Bw == 1 || Bw == B, "gemm_tc: Bw must be 1 or B");
  B200IR_CHECK_ARG(dtype_ok(y_dtype) && (!res || dtype_ok(res_dtype)), "gemm_tc: bad dtype");
  GemmGeom g;
  const bool tma_epi = y_dtype == B200IR_BF16 && (!res || res_dtype == B200IR_BF16) && y_bstride % 8 == 0 &&
                       ((uintptr_t)y % 16) == 0 && (!res || (res_bstride % 8 == 0 && ((uintptr_t)res % 16) == 0));
  B200IR_UNSUPPORTED(tma_epi || y_dtype == B200IR_F32,
                     "gemm_tc: bf16 output / residual must be 16B-aligned with batch strides %% 8 == 0");
  // Two epilogues.  Default: pixels on the TMEM lanes (MODE 0).  variant = B200IR_GEMM_RESIDENT_CH selects the channels-on-lanes
  // variant (MODE 4: D = W.X, a thread owns 32 pixels of one channel, 2.4x fewer instructions).  Measured on B200 it is
  // not faster (K4 352 vs 334 us, K6 191 vs 193 us): these GEMMs sit at the HBM WRITE roof (3.95 TB/s write-only vs
  // 6.6 TB/s copy, tools/bw_probe.py), not at the issue limit, so MODE 0 stays the default.
  const bool force_px = variant != B200IR_GEMM_RESIDENT_CH;
  const bool al16 = ((uintptr_t)y % 16) == 0 && y_bstride % 8 == 0 &&
                    (!res || (((uintptr_t)res % 16) == 0 && res_bstride % 8 == 0));
  bool mode4 = !force_px && al16 && gemm_tc_geometry_t(Cin, Cout, y_dtype == B200IR_BF16, &g);
  B200IR_UNSUPPORTED(mode4 || gemm_tc_geometry(Cin, Cout, &g, 1, res != nullptr, tma_epi),
                     "gemm_tc: unsupported Cin=%d Cout=%d", Cin, Cout);
  B200IR_UNSUPPORTED(HW % 8 == 0 && HW >= 64 && x_bstride % 8 == 0 && ((uintptr_t)x % 16) == 0,
                     "gemm_tc: needs HW%%8==0, HW>=64 and 16B-aligned activations (HW=%lld)", (long long)HW);
  B200IR_UNSUPPORTED(HW < (1ll << 31) && (int64_t)B * ((HW + 127) / 128) < (1ll << 31), "gemm_tc: too many tiles");

  // Deep levels (Cin >= 384: level 4 and the level-3/4 project_out): the weight matrix needs 3..15 resident slices, every
  // slice re-reads the activations and issues narrow MMAs -> stream both operands instead (gemm_stream_tc.cu; measured
  // 384->2042 103 -> 62 us, 1021->384 84 -> 39 us; three resident slices still win: 510->192 at 64x64 57 vs 82 us).  With short K the streamed weights cost more L2 traffic than the
  // output is worth, so everything else stays on the resident-weight kernel.  variant = B200IR_GEMM_RESIDENT / B200IR_GEMM_STREAM (b200ir_gemm_tc_fwd_ex) forces never / always
  // (tests, A/B runs); no environment variable is read.
  {
    bool want = g.Kpad >= 384 && Bw == 1 && (g.nslices >= 4 || (!res && g.nslices >= 2));
    if (variant != B200IR_GEMM_AUTO) want = variant == B200IR_GEMM_STREAM;
    if (!mode4 && want && Cin >= 64)
      return gemm_stream_launch(x, x_bstride, wp, Bw, st, ln_type, res, res_bstride, res_dtype, y, y_bstride, y_dtype,
                                B, Cin, Cout, HW, g.Kpad, g.Nrows, (cudaStream_t)stream);
  }
  CUtensorMap tmX, tmW;
  {
    const uint64_t dims[3] = {(uint64_t)HW, (uint64_t)Cin, (uint64_t)B};
    const uint64_t str[3] = {2, (uint64_t)HW * 2, (uint64_t)x_bstride * 2};
    const uint32_t box[3] = {64, (uint32_t)g.BK, 1};
    int rc = make_tmap_bf16(&tmX, x, 3, dims, str, box);
    if (rc) return rc;
  }
  {
    const uint64_t dims[3] = {(uint64_t)g.Kpad, (uint64_t)g.Nrows, (uint64_t)Bw};
    const uint64_t str[3] = {2, (uint64_t)g.Kpad * 2, (uint64_t)g.Nrows * g.Kpad * 2};
    const uint32_t box[3] = {64, (uint32_t)g.nbox, 1};
    int rc = make_tmap_bf16(&tmW, wp, 3, dims, str, box);
    if (rc) return rc;
  }
  CUtensorMap tmY = tmX, tmR = tmX;
  if (mode4) {
    if (y_dtype == B200IR_BF16) {   // one {32 px, 32 ch} box per warp item, SWIZZLE_64B staging
      const uint64_t dims[3] = {(uint64_t)HW, (uint64_t)Cout, (uint64_t)B};
      const uint32_t box[3] = {32, 32, 1};
      const uint64_t stry[3] = {2, (uint64_t)HW * 2, (uint64_t)y_bstride * 2};
      int rc = make_tmap_bf16_sw(&tmY, y, 3, dims, stry, box, 64);
      if (rc) return rc;
    }
  } else if (tma_epi) {
    const uint64_t dims[3] = {(uint64_t)HW, (uint64_t)Cout, (uint64_t)B};
    const uint32_t box[3] = {32, 16, 1};
    const uint64_t stry[3] = {2, (uint64_t)HW * 2, (uint64_t)y_bstride * 2};
    int rc = make_tmap_bf16(&tmY, y, 3, dims, stry, box, false);
    if (rc) return rc;
    if (res) {
      const uint64_t strr[3] = {2, (uint64_t)HW * 2, (uint64_t)res_bstride * 2};
      rc = make_tmap_bf16(&tmR, res, 3, dims, strr, box, false);
      if (rc) return rc;
    }
  }
  GemmTcParams P;
  P.y = y; P.y_bs = y_bstride; P.y_dt = y_dtype;
  P.res = res; P.res_bs = res_bstride; P.res_dt = res_dtype;
  P.s = st; P.t = st + g.Nrows;
  P.Nrows = g.Nrows; P.gpw = g.gpw; P.epi_warp_bytes = (int)g.epi_warp_bytes;
  P.ln_type = ln_type; P.Cin = Cin; P.Cout = Cout; P.HW = HW;
  P.tiles_per_sample = (int)((HW + 127) / 128);
  P.total_tiles = P.tiles_per_sample * B;
  P.BK = g.BK; P.kchunks = g.kchunks; P.Nslice = g.Nslice; P.NC = g.NC; P.nchunks = g.nchunks; P.nbox = g.nbox;
  P.kc64 = g.Kpad / 64; P.S = g.S;
  P.per_sample_w = (Bw > 1) ? 1 : 0;
  P.inv_cin = 1.0f / (float)Cin;
  B200IR_UNSUPPORTED(mode4 || g.nchunks * g.kchunks * 4 <= 256, "gemm_tc: descriptor table too small (Cin=%d Cout=%d)", Cin, Cout);
  B200IR_UNSUPPORTED(!mode4 || ((g.Nslice + 127) / 128) * g.kchunks * 4 <= 256, "gemm_tc: descriptor table too small (Cin=%d Cout=%d)", Cin, Cout);
  P.kpt = g.kchunks; P.H = P.W = P.TW = P.TWs = P.TH = P.tiles_x = P.tiles_y = 0;
  P.dbg = 0;
#ifdef B200IR_AB
  if (const char* de = getenv("B200IR_GEMM_DBG")) P.dbg = atoi(de);
#endif

  int sms = num_sms();
  int gx = sms / g.nslices;
  if (gx < 1) gx = 1;
  if (gx > P.total_tiles) gx = P.total_tiles;
  const dim3 grid((unsigned)gx, (unsigned)g.nslices);
  cudaStream_t st_ = (cudaStream_t)stream;
#define B200IR_LAUNCH_GEMM(TRES, TOUT)                                                                          \
  do {                                                                                                          \
    ensure_dyn_smem((const void*)gemm_tc_kernel<TRES, TOUT, 0>, 219 * 1024);                                                                                                         \
    ensure_dyn_smem((const void*)gemm_tc_kernel<TRES, TOUT, 4>, 219 * 1024);                                                                                                         \
    if (mode4) gemm_tc_kernel<TRES, TOUT, 4><<<grid, kThreads, g.smem, st_>>>(tmX, tmW, tmY, tmR, P);           \
    else gemm_tc_kernel<TRES, TOUT, 0><<<grid, kThreads, g.smem, st_>>>(tmX, tmW, tmY, tmR, P);                 \
  } while (0)
  const bool ob = y_dtype == B200IR_BF16;
  if (!res) {
    if (ob) B200IR_LAUNCH_GEMM(void, bf16); else B200IR_LAUNCH_GEMM(void, float);
  } else if (res_dtype == B200IR_BF16) {
    if (ob) B200IR_LAUNCH_GEMM(bf16, bf16); else B200IR_LAUNCH_GEMM(bf16, float);
  } else {
    if (ob) B200IR_LAUNCH_GEMM(float, bf16); else B200IR_LAUNCH_GEMM(float, float);
  }
#undef B200IR_LAUNCH_GEMM
  return check_launch("gemm_tc");
}


// ---------------------------------------------------------------------------------------------
// 3x3 dense conv on the same kernel (MODE 1..3)
// ---------------------------------------------------------------------------------------------
namespace b200ir {
// wp[n][tap*Kt + c] = w[n][c][tap] (Kt = kpt*BK), zero padded; st[1][n] = bias
__global__ void pack_conv3x3_kernel(const float* __restrict__ w, const float* __restrict__ bias, int Cin, int Cout,
                                    int Kt, int Kpad, int Nrows, bf16* __restrict__ wp, float* __restrict__ st) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= Nrows) return;
  bf16* dst = wp + (size_t)row * Kpad;
  for (int k = lane; k < Kpad; k += 32) {
    const int tap = k / Kt, c = k - tap * Kt;
    float v = 0.f;
    if (row < Cout && tap < 9 && c < Cin) v = w[((size_t)row * Cin + c) * 9 + tap];
    dst[k] = __float2bfloat16_rn(v);
  }
  if (lane == 0) {
    st[row] = 0.f;
    st[Nrows + row] = (row < Cout && bias) ? bias[row] : 0.f;
  }
}
}  // namespace b200ir

extern "C" int b200ir_conv3x3_tc_geometry(int Cin, int Cout, int* kpad, int* nrows) {
  GemmGeom g;
  if (!gemm_tc_geometry(Cin, Cout, &g, 9)) {
    set_error("conv3x3_tc: unsupported Cin=%d Cout=%d", Cin, Cout);
    return B200IR_E_UNSUPPORTED;
  }
  if (kpad) *kpad = g.Kpad;
  if (nrows) *nrows = g.Nrows;
  return 0;
}

extern "C" int b200ir_pack_conv3x3(const float* w, const float* bias, int Cin, int Cout, void* wp, float* st,
                                   b200ir_stream_t stream) {
  B200IR_CHECK_ARG(w && wp && st, "pack_conv3x3: bad args");
  GemmGeom g;
  B200IR_UNSUPPORTED(gemm_tc_geometry(Cin, Cout, &g, 9), "pack_conv3x3: unsupported Cin=%d Cout=%d", Cin, Cout);
  pack_conv3x3_kernel<<<(g.Nrows + 7) / 8, 256, 0, (cudaStream_t)stream>>>(w, bias, Cin, Cout, g.kpt * g.BK, g.Kpad,
                                                                            g.Nrows, (bf16*)wp, st);
  return check_launch("pack_conv3x3");
}

namespace b200ir {
int conv3x3_stream_launch(const void* x, int64_t x_bs, const void* wp, const float* st, int Nrows, int Kpad, int kpt,
                          void* y, int64_t y_bs, int y_dt, int B, int Cin, int Cout, int H, int W, int mode,
                          cudaStream_t stream);  // conv3x3_stream_tc.cu
}

extern "C" int b200ir_conv3x3_tc_fwd(const void* x, int64_t x_bstride, const void* wp, const float* st, void* y,
                                     int64_t y_bstride, int y_dtype, int B, int Cin, int Cout, int H, int W,
                                     int shuffle_mode, b200ir_stream_t stream) {
  B200IR_CHECK_ARG(x && wp && st && y && B > 0 && H > 0 && W > 0, "conv3x3_tc: bad args");
  B200IR_CHECK_ARG(shuffle_mode >= 0 && shuffle_mode <= 2, "conv3x3_tc: bad shuffle mode");
  B200IR_CHECK_ARG(dtype_ok(y_dtype), "conv3x3_tc: bad dtype");
  GemmGeom g;
  B200IR_UNSUPPORTED(gemm_tc_geometry(Cin, Cout, &g, 9), "conv3x3_tc: unsupported Cin=%d Cout=%d", Cin, Cout);
  B200IR_UNSUPPORTED(Cin % 8 == 0 && x_bstride % 8 == 0 && ((uintptr_t)x % 16) == 0,
                     "conv3x3_tc: needs Cin%%8==0 and 16B-aligned NHWC activations");
  B200IR_CHECK_ARG(shuffle_mode != B200IR_UNSHUFFLE2 || (H % 2 == 0 && W % 2 == 0), "conv3x3_tc: odd H/W");
  B200IR_CHECK_ARG(shuffle_mode != B200IR_SHUFFLE2 || Cout % 4 == 0, "conv3x3_tc: Cout%%4 != 0");
  if (g.nslices >= 2)  // the weight matrix does not fit in shared memory: stream both operands (conv3x3_stream_tc.cu)
    return conv3x3_stream_launch(x, x_bstride, wp, st, g.Nrows, g.Kpad, g.kpt, y, y_bstride, y_dtype, B, Cin, Cout, H,
                                 W, shuffle_mode, (cudaStream_t)stream);
  int TWs = 0;
  while ((1 << TWs) < W && TWs < 7) ++TWs;
  const int TW = 1 << TWs, TH = 128 / TW;
  CUtensorMap tmX, tmW;
  {
    // x is the NHWC copy: [B][H][W][Cin], Cin contiguous
    const uint64_t dims[4] = {(uint64_t)Cin, (uint64_t)W, (uint64_t)H, (uint64_t)B};
    const uint64_t str[4] = {2, (uint64_t)Cin * 2, (uint64_t)W * Cin * 2, (uint64_t)x_bstride * 2};
    const uint32_t box[4] = {64, (uint32_t)TW, (uint32_t)TH, 1};
    int rc = make_tmap_bf16(&tmX, x, 4, dims, str, box);
    if (rc) return rc;
  }
  {
    const uint64_t dims[3] = {(uint64_t)g.Kpad, (uint64_t)g.Nrows, 1};
    const uint64_t str[3] = {2, (uint64_t)g.Kpad * 2, (uint64_t)g.Nrows * g.Kpad * 2};
    const uint32_t box[3] = {64, (uint32_t)g.nbox, 1};
    int rc = make_tmap_bf16(&tmW, wp, 3, dims, str, box);
    if (rc) return rc;
  }
  GemmTcParams P;
  P.y = y; P.y_bs = y_bstride; P.y_dt = y_dtype;
  P.res = nullptr; P.res_bs = 0; P.res_dt = 0;
  P.s = st; P.t = st + g.Nrows;
  P.Nrows = g.Nrows; P.gpw = 0; P.epi_warp_bytes = 0;
  P.ln_type = B200IR_LN_NONE; P.Cin = Cin; P.Cout = Cout; P.HW = (int64_t)H * W;
  P.H = H; P.W = W; P.TW = TW; P.TWs = TWs; P.TH = TH;
  P.tiles_x = (W + TW - 1) / TW; P.tiles_y = (H + TH - 1) / TH;
  P.tiles_per_sample = 1 << 30;
  P.total_tiles = B * P.tiles_x * P.tiles_y;
  P.BK = g.BK; P.kchunks = g.kchunks; P.kpt = g.kpt; P.Nslice = g.Nslice; P.NC = g.NC; P.nchunks = g.nchunks;
  P.nbox = g.nbox; P.kc64 = g.Kpad / 64; P.S = g.S;
  B200IR_UNSUPPORTED(g.nchunks * g.kchunks * 4 <= 256, "conv3x3_tc: descriptor table too small (Cin=%d Cout=%d)", Cin, Cout);
  P.per_sample_w = 0;
  P.inv_cin = 0.f; P.dbg = 0;
  int gx = num_sms() / g.nslices;
  if (gx < 1) gx = 1;
  if (gx > P.total_tiles) gx = P.total_tiles;
  const dim3 grid((unsigned)gx, (unsigned)g.nslices);
  cudaStream_t st_ = (cudaStream_t)stream;
#define B200IR_LAUNCH_CONV(TOUT, MODE)                                                                           \
  do {                                                                                                           \
    ensure_dyn_smem((const void*)gemm_tc_kernel<void, TOUT, MODE>, 219 * 1024);                                                                                                          \
    gemm_tc_kernel<void, TOUT, MODE><<<grid, kThreads, g.smem, st_>>>(tmX, tmW, tmX, tmX, P);                    \
  } while (0)
  const int mode = shuffle_mode + 1;
  if (y_dtype == B200IR_BF16) {
    if (mode == 1) B200IR_LAUNCH_CONV(bf16, 1); else if (mode == 2) B200IR_LAUNCH_CONV(bf16, 2); else B200IR_LAUNCH_CONV(bf16, 3);
  } else {
    if (mode == 1) B200IR_LAUNCH_CONV(float, 1); else if (mode == 2) B200IR_LAUNCH_CONV(float, 2); else B200IR_LAUNCH_CONV(float, 3);
  }
#undef B200IR_LAUNCH_CONV
  return check_launch("conv3x3_tc");
}


// ---------------------------------------------------------------------------------------------
// NCHW -> NHWC (bf16) staging copy for the conv activation operand
// ---------------------------------------------------------------------------------------------
namespace b200ir {
__global__ void __launch_bounds__(256)
nchw_to_nhwc_kernel(const bf16* __restrict__ x, int64_t x_bs, bf16* __restrict__ y, int C, int64_t HW) {
  __shared__ bf16 t[64][66];
  const int b = blockIdx.z;
  const int64_t p0 = (int64_t)blockIdx.x * 64;
  const int c0 = blockIdx.y * 64;
  const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;  // 64 x 4
  for (int i = ty; i < 64; i += 4) {
    const int c = c0 + i;
    const int64_t p = p0 + tx;
    t[i][tx] = (c < C && p < HW) ? x[(int64_t)b * x_bs + (int64_t)c * HW + p] : __float2bfloat16_rn(0.f);
  }
  __syncthreads();
  for (int i = ty; i < 64; i += 4) {
    const int64_t p = p0 + i;
    const int c = c0 + tx;
    if (c < C && p < HW) y[((int64_t)b * HW + p) * C + c] = t[tx][i];
  }
}
}  // namespace b200ir

extern "C" int b200ir_nchw_to_nhwc(const void* x, int64_t x_bstride, void* y, int B, int C, int64_t HW,
                                   b200ir_stream_t stream) {
  B200IR_CHECK_ARG(x && y && B > 0 && C > 0 && HW > 0 && B <= 65535, "nchw_to_nhwc: bad args");
  dim3 grid((unsigned)((HW + 63) / 64), (unsigned)((C + 63) / 64), (unsigned)B);
  nchw_to_nhwc_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const bf16*)x, x_bstride, (bf16*)y, C, HW);
  return check_launch("nchw_to_nhwc");
}
