// mdta.cu — MDTA transposed attention, stages 1 (Gram + norms, fp32 CUDA-core path) and 2
// (slab reduce + normalise + temperature softmax + project_out fold).
// Reference: Attention.forward Restormer.py:111-132.
//   :117-119  q,k,v = heads of contiguous channel groups of the q|k|v chunks (index math only here)
//   :121-122  F.normalize(dim=-1): x / max(||x||_2, 1e-12)  -> applied to the d x d Gram, never to q/k
//   :124      (q @ k^T) * temperature[h]
//   :125      softmax over the last dim (rows of length d = C/heads)
//   :127-131  attn @ v, then project_out  -> folded into m[b] = W_out . blockdiag(attn[b,h])
#include <mutex>

#include "common.cuh"

namespace b200ir {

constexpr int GPC = 32;  // pixels per smem chunk in the SIMT Gram

// grid (nslab, heads, B), 256 threads as 16x16, thread (ty,tx) owns G[ty+16*ii][tx+16*jj]
template <int TI>
__global__ void __launch_bounds__(256)
gram_simt_kernel(const void* __restrict__ qkv, int64_t bs, int dt, int C, int d, int64_t HW, int64_t slab_len,
                 float* __restrict__ partials, int nslab) {
  extern __shared__ float sm[];
  float* qs = sm;                       // [d][GPC+1]
  float* ks = sm + (size_t)d * (GPC + 1);
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int slab = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int64_t pbeg = (int64_t)slab * slab_len;
  const int64_t pend = min(HW, pbeg + slab_len);
  const int64_t qoff = (int64_t)b * bs + (int64_t)(h * d) * HW;
  const int64_t koff = qoff + (int64_t)C * HW;

  float acc[TI][TI];
#pragma unroll
  for (int i = 0; i < TI; ++i)
#pragma unroll
    for (int j = 0; j < TI; ++j) acc[i][j] = 0.f;
  float nrm = 0.f;  // thread t<d: sum q[t]^2 ; d<=t<2d: sum k[t-d]^2

  for (int64_t p0 = pbeg; p0 < pend; p0 += GPC) {
    for (int e = tid; e < d * GPC; e += 256) {
      const int r = e / GPC, cpx = e % GPC;
      const int64_t p = p0 + cpx;
      float qv = 0.f, kv = 0.f;
      if (p < pend) {
        qv = ld_any(qkv, dt, qoff + (int64_t)r * HW + p);
        kv = ld_any(qkv, dt, koff + (int64_t)r * HW + p);
      }
      qs[r * (GPC + 1) + cpx] = qv;
      ks[r * (GPC + 1) + cpx] = kv;
    }
    __syncthreads();
#pragma unroll 4
    for (int k = 0; k < GPC; ++k) {
      float a[TI], bb[TI];
#pragma unroll
      for (int i = 0; i < TI; ++i) {
        const int r = ty + 16 * i;
        a[i] = r < d ? qs[r * (GPC + 1) + k] : 0.f;
        const int c2 = tx + 16 * i;
        bb[i] = c2 < d ? ks[c2 * (GPC + 1) + k] : 0.f;
      }
#pragma unroll
      for (int i = 0; i < TI; ++i)
#pragma unroll
        for (int j = 0; j < TI; ++j) acc[i][j] = fmaf(a[i], bb[j], acc[i][j]);
    }
    if (tid < 2 * d) {
      const float* src = tid < d ? qs + tid * (GPC + 1) : ks + (tid - d) * (GPC + 1);
#pragma unroll 8
      for (int k = 0; k < GPC; ++k) nrm = fmaf(src[k], src[k], nrm);
    }
    __syncthreads();
  }
  float* out = partials + (((size_t)b * gridDim.y + h) * nslab + slab) * ((size_t)d * d + 2 * d);
#pragma unroll
  for (int i = 0; i < TI; ++i)
#pragma unroll
    for (int j = 0; j < TI; ++j) {
      const int r = ty + 16 * i, c2 = tx + 16 * j;
      if (r < d && c2 < d) out[r * d + c2] = acc[i][j];
    }
  if (tid < 2 * d) out[d * d + tid] = nrm;
}

// sum of n values spaced `stride` apart, in index order (bitwise reproducible), with the loads of each group of 8 issued
// together so the chain is not one DRAM/L2 latency per term
__device__ __forceinline__ float ordered_sum(const float* __restrict__ p, size_t stride, int n) {
  float s = 0.f;
  int i = 0;
  for (; i + 8 <= n; i += 8) {
    float v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = __ldg(p + (size_t)(i + k) * stride);
#pragma unroll
    for (int k = 0; k < 8; ++k) s += v[k];
  }
  float v[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) v[k] = i + k < n ? __ldg(p + (size_t)(i + k) * stride) : 0.f;
#pragma unroll
  for (int k = 0; k < 8; ++k)
    if (i + k < n) s += v[k];
  return s;
}

// ---- stage 2a: slab reduce + normalise + temperature + row softmax.  grid (ceil(d/8), heads, B), 256 threads:
// one warp per attention row (8 rows per block), so B*heads*d/8 blocks share the slab reduction.
__global__ void __launch_bounds__(256)
mdta_softmax_kernel(const float* __restrict__ partials, int nslab, const float* __restrict__ temperature, int temp_bs,
                    float* __restrict__ attn, int d) {
  extern __shared__ float sm[];
  float* nk = sm;             // [d]  max(||k_j||, eps)
  float* rowbuf = sm + d;     // [8][d]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int h = blockIdx.y, b = blockIdx.z, heads = gridDim.y;
  const size_t per = (size_t)d * d + 2 * d;
  const float* base = partials + ((size_t)b * heads + h) * nslab * per;
  for (int j = tid; j < d; j += 256) {
    const float s = ordered_sum(base + (size_t)d * d + d + j, per, nslab);  // fixed order
    nk[j] = fmaxf(sqrtf(s), 1e-12f);                                                          // (:122)
  }
  __syncthreads();
  const int i = blockIdx.x * 8 + warp;
  if (i >= d) return;
  float nq = 0.f;
  if (lane == 0) {
    const float s = ordered_sum(base + (size_t)d * d + i, per, nslab);
    nq = fmaxf(sqrtf(s), 1e-12f);                                                             // (:121)
  }
  nq = __shfl_sync(0xffffffffu, nq, 0);
  const float temp = temperature[(size_t)b * temp_bs + h];   // temp_bs = heads: per-sample (dynamic) temperature
  float* row = rowbuf + warp * d;
  float mx = -INFINITY;
  for (int j = lane; j < d; j += 32) {
    const float s = ordered_sum(base + (size_t)i * d + j, per, nslab);  // fixed order
    const float v = s / (nq * nk[j]) * temp;                                                  // (:124)
    row[j] = v;
    mx = fmaxf(mx, v);
  }
  mx = warp_max(mx);
  float sum = 0.f;
  for (int j = lane; j < d; j += 32) {
    const float e = expf(row[j] - mx);
    row[j] = e;
    sum += e;
  }
  sum = warp_sum(sum);
  const float inv = 1.0f / sum;
  float* out = attn + (((size_t)b * heads + h) * d + i) * d;
  for (int j = lane; j < d; j += 32) out[j] = row[j] * inv;                                   // (:125)
}

// ---- stage 2b: fold project_out into the attention matrix: m[b][o][h*d+j] = sum_i w_out[o][h*d+i] attn[b,h][i][j]
// grid (ceil(C/kFoldRows), heads, B), 256 threads; kFoldRows = 32 or 8 (launcher).  Writes fp32 m [B,C,C] and/or the bf16 operand of the tcgen05 GEMM
// mp [B][nrows][kpad] (zero padded) so that no separate packing launch is needed.
// Register-tiled: the 256 threads form a 16 x 16 grid, thread (ty, tx) owns output rows ty*RT .. ty*RT+RT-1 and columns
// tx + 16 k (k < KC, KC = ceil(d/16)): per reduction step RT + KC shared loads feed RT*KC FMAs (the one-output-per-thread
// form issued 2 LDS per FMA and was LDS-bound: 40 us at C=384).  The i-sum runs in increasing i in every variant, so
// the result does not depend on RT / the row-block size.
template <int RT, int KC>
__global__ void __launch_bounds__(256)
mdta_fold_kernel(const float* __restrict__ attn, const float* __restrict__ w_out, float* __restrict__ m,
                 bf16* __restrict__ mp, int kpad, int nrows, int C, int d) {
  constexpr int kFoldRows = 16 * RT;
  extern __shared__ float sm[];
  float* at = sm;                    // [d][d]      (column reads are consecutive over tx: conflict-free unpadded)
  float* ws = sm + (size_t)d * d;    // [kFoldRows][d+1]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int tx = tid & 15, ty = tid >> 4;
  const int h = blockIdx.y, b = blockIdx.z, heads = gridDim.y;
  const int o0 = blockIdx.x * kFoldRows;
  const float* a = attn + ((size_t)b * heads + h) * d * d;
  if ((d & 3) == 0) {
    const float4* a4 = reinterpret_cast<const float4*>(a);
    float4* at4 = reinterpret_cast<float4*>(at);
    for (int e = tid; e < (d * d) >> 2; e += 256) at4[e] = a4[e];
  } else {
    for (int e = tid; e < d * d; e += 256) at[e] = a[e];
  }
  for (int r = warp; r < kFoldRows; r += 8) {
    const int o = o0 + r;
    const float* src = w_out + (size_t)o * C + h * d;
    for (int i = lane; i < d; i += 32) ws[r * (d + 1) + i] = o < C ? src[i] : 0.f;
  }
  __syncthreads();
  float acc[RT][KC];
#pragma unroll
  for (int r = 0; r < RT; ++r)
#pragma unroll
    for (int k = 0; k < KC; ++k) acc[r][k] = 0.f;
  const float* wrow = ws + (size_t)(ty * RT) * (d + 1);
#pragma unroll 4
  for (int i = 0; i < d; ++i) {
    float av[KC], wv[RT];
#pragma unroll
    for (int k = 0; k < KC; ++k) av[k] = (tx + 16 * k < d) ? at[i * d + tx + 16 * k] : 0.f;
#pragma unroll
    for (int r = 0; r < RT; ++r) wv[r] = wrow[r * (d + 1) + i];
#pragma unroll
    for (int r = 0; r < RT; ++r)
#pragma unroll
      for (int k = 0; k < KC; ++k) acc[r][k] = fmaf(wv[r], av[k], acc[r][k]);
  }
#pragma unroll
  for (int r = 0; r < RT; ++r) {
    const int o = o0 + ty * RT + r;
    if (o >= C) continue;
#pragma unroll
    for (int k = 0; k < KC; ++k) {
      const int j = tx + 16 * k;
      if (j >= d) continue;
      if (m) m[((size_t)b * C + o) * C + h * d + j] = acc[r][k];
      if (mp) mp[((size_t)b * nrows + o) * kpad + h * d + j] = __float2bfloat16_rn(acc[r][k]);
    }
  }
  if (mp) {  // zero padding: columns [C,kpad) of valid rows (head 0 blocks) and whole rows [C,nrows) (last row block)
    if (h == 0)
      for (int e = tid; e < kFoldRows * (kpad - C); e += 256) {
        const int o = o0 + e / (kpad - C), c = C + e % (kpad - C);
        if (o < C) mp[((size_t)b * nrows + o) * kpad + c] = __float2bfloat16_rn(0.f);
      }
    if (h == 0 && blockIdx.x == gridDim.x - 1)
      for (int e = tid; e < (nrows - C) * kpad; e += 256)
        mp[((size_t)b * nrows + C + e / kpad) * kpad + e % kpad] = __float2bfloat16_rn(0.f);
  }
}

int gram_slab_len(int B, int d, int heads, int64_t HW, int* nslab_out) {
  // (b, head, slab) work items = CTAs: as many as run CONCURRENTLY on 148 SMs and no more, so that there is no
  // nearly-empty last wave (19 slabs x 16 samples = 304 CTAs on 148 resident slots ran as 2.05 -> 3 waves).  Head dims
  // <= 64 fit two CTAs per SM (TMEM columns and shared memory), larger ones one.  Slabs are multiples of 256 pixels.
  const int64_t cap = d <= 64 ? 2 * num_sms() : num_sms();
  int64_t want = cap / ((int64_t)B * heads);
  int64_t maxslab = (HW + 255) / 256;
  if (want > maxslab) want = maxslab;
  if (want < 1) want = 1;
  int64_t len = (HW + want - 1) / want;
  len = (len + 255) / 256 * 256;
  int nslab = (int)((HW + len - 1) / len);
  if (nslab_out) *nslab_out = nslab;
  return (int)len;
}

}  // namespace b200ir

using namespace b200ir;

extern "C" int b200ir_mdta_gram_nslab(int B, int C, int heads, int64_t HW) {
  if (B <= 0 || C <= 0 || heads <= 0 || HW <= 0) return B200IR_E_BADARG;
  int n = 0;
  gram_slab_len(B, C / heads, heads, HW, &n);
  return n;
}

namespace b200ir {
int gram_tc_launch(const void* qkv, int64_t bstride, int B, int C, int heads, int64_t HW, float* partials, int nslab,
                   int64_t slab_len, cudaStream_t st);  // gram_tc.cu
bool gram_tc_supported(int dtype, int C, int heads, int64_t HW, int64_t bstride, const void* qkv);
}

extern "C" int b200ir_mdta_gram_fwd(const void* qkv, int64_t bstride, int dtype, int B, int C, int heads,
                                    int64_t HW, float* partials, int nslab, int impl, b200ir_stream_t stream) {
  B200IR_CHECK_ARG(qkv && partials && B > 0 && C > 0 && heads > 0 && HW > 0, "mdta_gram: bad args");
  B200IR_CHECK_ARG(C % heads == 0, "mdta_gram: C=%d not divisible by heads=%d", C, heads);
  B200IR_CHECK_ARG(dtype_ok(dtype), "mdta_gram: bad dtype");
  const int d = C / heads;
  B200IR_UNSUPPORTED(d <= 128, "mdta_gram: head dim %d > 128 unsupported", d);
  int want = 0;
  const int64_t slab_len = gram_slab_len(B, d, heads, HW, &want);
  B200IR_CHECK_ARG(nslab == want, "mdta_gram: nslab=%d, expected %d (b200ir_mdta_gram_nslab)", nslab, want);
  B200IR_CHECK_ARG(B <= 65535 && heads <= 65535, "mdta_gram: grid too large");
  cudaStream_t st = (cudaStream_t)stream;
  if (impl != B200IR_IMPL_SIMT && gram_tc_supported(dtype, C, heads, HW, bstride, qkv))
    return gram_tc_launch(qkv, bstride, B, C, heads, HW, partials, nslab, slab_len, st);
  B200IR_UNSUPPORTED(impl != B200IR_IMPL_TC, "mdta_gram: tcgen05 path unsupported for this shape/dtype");
  dim3 grid((unsigned)nslab, (unsigned)heads, (unsigned)B);
  const size_t smem = 2 * (size_t)d * (GPC + 1) * sizeof(float);
  const int ti = (d + 15) / 16;
#define LAUNCH(TI) gram_simt_kernel<TI><<<grid, 256, smem, st>>>(qkv, bstride, dtype, C, d, HW, slab_len, partials, nslab)
  if (ti <= 2) LAUNCH(2);
  else if (ti <= 3) LAUNCH(3);
  else if (ti <= 4) LAUNCH(4);
  else if (ti <= 6) LAUNCH(6);
  else LAUNCH(8);
#undef LAUNCH
  return check_launch("mdta_gram_simt");
}

extern "C" int b200ir_mdta_softmax_fold_fwd(const float* partials, int nslab, const float* temperature,
                                            const float* w_out, float* m, float* attn, void* mp, int kpad, int nrows,
                                            int B, int C, int heads, b200ir_stream_t stream) {
  return b200ir_mdta_softmax_fold_fwd_ex(partials, nslab, temperature, 0, w_out, m, attn, mp, kpad, nrows, B, C, heads,
                                         stream);
}

extern "C" int b200ir_mdta_softmax_fold_fwd_ex(const float* partials, int nslab, const float* temperature,
                                               int temp_per_sample, const float* w_out, float* m, float* attn, void* mp,
                                               int kpad, int nrows, int B, int C, int heads, b200ir_stream_t stream) {
  B200IR_CHECK_ARG(partials && temperature && w_out && attn && (m || mp) && nslab > 0 && B > 0 && C > 0 && heads > 0,
                   "mdta_softmax_fold: bad args");
  B200IR_CHECK_ARG(C % heads == 0 && B <= 65535 && heads <= 65535, "mdta_softmax_fold: bad C/heads/B");
  B200IR_CHECK_ARG(!mp || (kpad >= C && nrows >= C), "mdta_softmax_fold: bad packed geometry");
  const int d = C / heads;
  B200IR_UNSUPPORTED(d <= 128, "mdta_softmax_fold: head dim %d > 128 unsupported", d);
  cudaStream_t st = (cudaStream_t)stream;
  mdta_softmax_kernel<<<dim3((unsigned)((d + 7) / 8), (unsigned)heads, (unsigned)B), 256, (size_t)9 * d * sizeof(float),
                        st>>>(partials, nslab, temperature, temp_per_sample ? heads : 0, attn, d);
  int rc = check_launch("mdta_softmax");
  if (rc) return rc;
  // output rows per block (16 per register row RT): the largest of 64 / 32 / 16 that still gives every SM a block
  int RT = 1;
  for (int rt = 4; rt >= 1; rt >>= 1)
    if ((int64_t)((C + 16 * rt - 1) / (16 * rt)) * heads * B >= num_sms()) {
      RT = rt;
      break;
    }
  const int kFoldRows = 16 * RT;
  const int KC = (d + 15) / 16;   // 1..8
  const size_t smem = ((size_t)d * d + kFoldRows * (size_t)(d + 1)) * sizeof(float);
  const dim3 grid((unsigned)((C + kFoldRows - 1) / kFoldRows), (unsigned)heads, (unsigned)B);
#define FOLD_LAUNCH(RT_, KC_)                                                                                       \
  do {                                                                                                              \
    ensure_dyn_smem((const void*)mdta_fold_kernel<RT_, KC_>, 100 * 1024);                                                                                                             \
    mdta_fold_kernel<RT_, KC_><<<grid, 256, smem, st>>>(attn, w_out, m, (bf16*)mp, kpad, nrows, C, d);              \
  } while (0)
#define FOLD_KC(RT_)                                                                                                \
  do {                                                                                                              \
    if (KC <= 2) FOLD_LAUNCH(RT_, 2);                                                                               \
    else if (KC <= 3) FOLD_LAUNCH(RT_, 3);                                                                          \
    else if (KC <= 6) FOLD_LAUNCH(RT_, 6);                                                                          \
    else FOLD_LAUNCH(RT_, 8);                                                                                       \
  } while (0)
  if (RT == 4) FOLD_KC(4);
  else if (RT == 2) FOLD_KC(2);
  else FOLD_KC(1);
#undef FOLD_KC
#undef FOLD_LAUNCH
  return check_launch("mdta_fold");
}
