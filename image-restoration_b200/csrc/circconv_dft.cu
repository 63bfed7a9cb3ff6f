// circconv_dft.cu — the 32 x 32 and 16 x 16 FFT experts of MoCE-IR on tensor cores (bf16 path only).
//
// Reference: FFTAttention, /root/reference/moce_ir.py:406-412: per 32 x 32 patch and channel
//     out = irfft2(rfft2(q) * rfft2(k))            (mode 0; its gradients are mode 1: irfft2(rfft2(a) * conj(rfft2(b))))
// The direct circular convolution (moce_ops.cu: 1024 x 1024 MACs per patch on the FP32 pipe, ~75 % of its peak) is the
// largest kernel of the MoCE-IR forward (4.1 of 28 ms at cfg3).  Here the same product is taken in the Fourier domain
// with the 32-point DFT written as a MATRIX (F = Fr + i Fi, symmetric): a 2-D transform of a patch is F X F, i.e. small
// real 32 x 32 x 32 products — 18 of them per (patch, channel) instead of 1024^2 MACs:
//     forward  (x2)  A = F X        : Ar = Fr X, Ai = Fi X                                        2 products
//                    Y = A F        : Yr = Ar Fr - Ai Fi, Yi = Ar Fi + Ai Fr                      4
//     product        Z = Yq o Yk    (mode 1: Yq o conj(Yk)), element-wise in the accumulator registers
//     inverse        B = conj(F) Z  : Br = Fr Zr + Fi Zi, Bi = Fr Zi - Fi Zr                      4
//                    out = Re(B conj(F)) / 1024 = (Br Fr + Bi Fi) / 1024                          2
// Each product runs as mma.sync.m16n8k8 TF32 with fp32 accumulation, one WARP per (patch, channel); operands are rounded to
// TF32 (cvt.rna) when they are written to shared memory.  This is the numerics of the bf16 activation path only (TF32
// carries 11 significant bits against bf16's 8); the fp32 path keeps the exact direct kernel.
// The DFT is too small for tcgen05 (M = 64 / 128 tiles, one accumulator per CTA): 32 x 32 matrices per warp are what
// mma.sync is for.
#include "common.cuh"

namespace b200ir {

// P x P matrices (P = 32 or 16) as mma tiles; row strides (floats) chosen so that the fragment loads are conflict-free:
// a LEFT operand is read at (row g, col t): stride = 4 mod 32 -> bank 4g + t; a RIGHT operand at (k = t, n = g): stride = 8 mod 32
template <int P>
struct Dft {
  static constexpr int MT = P / 16, NT = P / 8, KT = P / 8;
  static constexpr int LDA = P == 32 ? 36 : 20;
  static constexpr int LDB = P == 32 ? 40 : 24;
  static constexpr int MAT = P * LDB;   // floats reserved per matrix (either stride fits)
};

__device__ __forceinline__ uint32_t to_tf32(float v) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
  return r;
}

__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// fragments of the mma (PTX ISA, m16n8k8 .tf32): g = lane / 4, t = lane % 4
//   left  16 x 8: a0 (g, t)  a1 (g + 8, t)  a2 (g, t + 4)  a3 (g + 8, t + 4)
//   right  8 x 8: b0 (k = t, n = g)  b1 (k = t + 4, n = g)
//   acc   16 x 8: c0 (g, 2t)  c1 (g, 2t + 1)  c2 (g + 8, 2t)  c3 (g + 8, 2t + 1)
template <int P>
__device__ __forceinline__ void ld_left(const uint32_t* m, int mt, int kt, int g, int t, uint32_t (&a)[4]) {
  constexpr int LD = Dft<P>::LDA;
  const uint32_t* p = m + (mt * 16 + g) * LD + kt * 8 + t;
  a[0] = p[0];
  a[1] = p[8 * LD];
  a[2] = p[4];
  a[3] = p[8 * LD + 4];
}
template <int P>
__device__ __forceinline__ void ld_right(const uint32_t* m, int kt, int nt, int g, int t, uint32_t& b0, uint32_t& b1) {
  constexpr int LD = Dft<P>::LDB;
  const uint32_t* p = m + (kt * 8 + t) * LD + nt * 8 + g;
  b0 = p[0];
  b1 = p[4 * LD];
}
// accumulator tile (mt, nt) -> matrix with row stride LD, rounded to TF32 (it is the operand of the next product)
template <int LD>
__device__ __forceinline__ void st_acc(uint32_t* m, int mt, int nt, int g, int t, const float (&c)[4]) {
  uint32_t* p = m + (mt * 16 + g) * LD + nt * 8 + 2 * t;
  *reinterpret_cast<uint2*>(p) = make_uint2(to_tf32(c[0]), to_tf32(c[1]));
  *reinterpret_cast<uint2*>(p + 8 * LD) = make_uint2(to_tf32(c[2]), to_tf32(c[3]));
}

// one real P x P matrix in the registers of a warp: [m tile][n tile][fragment]
template <int P>
struct Acc {
  float v[Dft<P>::MT][Dft<P>::NT][4];
  __device__ __forceinline__ void zero() {
#pragma unroll
    for (int i = 0; i < Dft<P>::MT; ++i)
#pragma unroll
      for (int j = 0; j < Dft<P>::NT; ++j)
#pragma unroll
        for (int k = 0; k < 4; ++k) v[i][j][k] = 0.f;
  }
};

// ---- the four kinds of product (all P x P x P, per warp) ----------------------------------------------------------------
// out1 (+)= L1 R1 [+ s12 L2 R2];  out2 = L1 R2' ... are all instances of: for every k tile load the RIGHT fragments of up to two
// matrices, for every m tile the LEFT fragments of up to two matrices, and issue the mma-s the caller's lambda names.
template <int P, typename F>
__device__ __forceinline__ void mm_pair(const uint32_t* L1, const uint32_t* L2, const uint32_t* R1, const uint32_t* R2, int g,
                                        int t, F&& issue) {
  constexpr int MT = Dft<P>::MT, NT = Dft<P>::NT, KT = Dft<P>::KT;
#pragma unroll
  for (int kt = 0; kt < KT; ++kt) {
    uint32_t r1a[NT], r1b[NT], r2a[NT] = {}, r2b[NT] = {};
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      ld_right<P>(R1, kt, nt, g, t, r1a[nt], r1b[nt]);
      if (R2) ld_right<P>(R2, kt, nt, g, t, r2a[nt], r2b[nt]);
    }
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) {
      uint32_t l1[4], l2[4];
      ld_left<P>(L1, mt, kt, g, t, l1);
      ld_left<P>(L2, mt, kt, g, t, l2);
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) issue(mt, nt, l1, l2, r1a[nt], r1b[nt], R2 ? r2a[nt] : 0u, R2 ? r2b[nt] : 0u);
    }
  }
}

constexpr uint32_t kNeg = 0x80000000u;   // sign flip of a TF32 operand

// rows:  Ur = Fr X, Ui = Fi X                      X = right operand in shared memory, F = left operand copies (FA)
template <int P>
__device__ __forceinline__ void dft_rows(const uint32_t* X, const uint32_t* FAr, const uint32_t* FAi, int g, int t, Acc<P>& ur,
                                         Acc<P>& ui) {
  ur.zero();
  ui.zero();
  mm_pair<P>(FAr, FAi, X, nullptr, g, t, [&](int mt, int nt, const uint32_t (&fr)[4], const uint32_t (&fi)[4], uint32_t b0,
                                             uint32_t b1, uint32_t, uint32_t) {
    mma_tf32(ur.v[mt][nt], fr, b0, b1);
    mma_tf32(ui.v[mt][nt], fi, b0, b1);
  });
}

// cols:  Yr = Ur Fr - Ui Fi, Yi = Ur Fi + Ui Fr    U = left operands in shared memory, F = right operand copies (FB)
template <int P>
__device__ __forceinline__ void dft_cols(const uint32_t* Ur, const uint32_t* Ui, const uint32_t* FBr, const uint32_t* FBi,
                                         int g, int t, Acc<P>& yr, Acc<P>& yi) {
  yr.zero();
  yi.zero();
  mm_pair<P>(Ur, Ui, FBr, FBi, g, t, [&](int mt, int nt, const uint32_t (&ar)[4], const uint32_t (&ai)[4], uint32_t r0,
                                         uint32_t r1, uint32_t i0, uint32_t i1) {
    mma_tf32(yr.v[mt][nt], ar, r0, r1);
    mma_tf32(yr.v[mt][nt], ai, i0 ^ kNeg, i1 ^ kNeg);
    mma_tf32(yi.v[mt][nt], ar, i0, i1);
    mma_tf32(yi.v[mt][nt], ai, r0, r1);
  });
}

// inverse rows:  Br = Fr Zr + Fi Zi, Bi = Fr Zi - Fi Zr     Z = right operands in shared memory
template <int P>
__device__ __forceinline__ void idft_rows(const uint32_t* Zr, const uint32_t* Zi, const uint32_t* FAr, const uint32_t* FAi,
                                          int g, int t, Acc<P>& br, Acc<P>& bi) {
  br.zero();
  bi.zero();
  mm_pair<P>(FAr, FAi, Zr, Zi, g, t, [&](int mt, int nt, const uint32_t (&fr)[4], const uint32_t (&fi)[4], uint32_t r0,
                                         uint32_t r1, uint32_t i0, uint32_t i1) {
    mma_tf32(br.v[mt][nt], fr, r0, r1);
    mma_tf32(br.v[mt][nt], fi, i0, i1);
    mma_tf32(bi.v[mt][nt], fr, i0, i1);
    mma_tf32(bi.v[mt][nt], fi, r0 ^ kNeg, r1 ^ kNeg);
  });
}

// inverse cols (real part only):  O = Br Fr + Bi Fi
template <int P>
__device__ __forceinline__ void idft_cols(const uint32_t* Br, const uint32_t* Bi, const uint32_t* FBr, const uint32_t* FBi,
                                          int g, int t, Acc<P>& o) {
  o.zero();
  mm_pair<P>(Br, Bi, FBr, FBi, g, t, [&](int mt, int nt, const uint32_t (&ar)[4], const uint32_t (&ai)[4], uint32_t r0,
                                         uint32_t r1, uint32_t i0, uint32_t i1) {
    mma_tf32(o.v[mt][nt], ar, r0, r1);
    mma_tf32(o.v[mt][nt], ai, i0, i1);
  });
}

template <int P, int LD>
__device__ __forceinline__ void st_mat(uint32_t* m, int g, int t, const Acc<P>& c) {
#pragma unroll
  for (int mt = 0; mt < Dft<P>::MT; ++mt)
#pragma unroll
    for (int nt = 0; nt < Dft<P>::NT; ++nt) st_acc<LD>(m, mt, nt, g, t, c.v[mt][nt]);
}

// one P x P bf16 patch -> X (right-operand layout, fp32 bits: bf16 values are exact in TF32); zero outside the plane
template <int P>
__device__ __forceinline__ void load_patch(const bf16* __restrict__ plane, int H, int W, int y0, int x0, int lane, uint32_t* X) {
  constexpr int CPR = P / 8, RPI = 32 / CPR, LD = Dft<P>::LDB;
  const int r0 = lane / CPR, ch = lane % CPR;
#pragma unroll
  for (int i = 0; i < P / RPI; ++i) {
    const int row = i * RPI + r0, gy = y0 + row, gx = x0 + ch * 8;
    uint4 q = make_uint4(0u, 0u, 0u, 0u);
    if (gy < H && gx < W) q = __ldg(reinterpret_cast<const uint4*>(plane + (int64_t)gy * W + gx));   // W % 8 == 0
    uint32_t* d = X + row * LD + ch * 8;
    *reinterpret_cast<uint4*>(d) = make_uint4(q.x << 16, q.x & 0xffff0000u, q.y << 16, q.y & 0xffff0000u);
    *reinterpret_cast<uint4*>(d + 4) = make_uint4(q.z << 16, q.z & 0xffff0000u, q.w << 16, q.w & 0xffff0000u);
  }
}

// MODE 0: y = a (*) b circular convolution;  MODE 1: y[s] = sum_i a[i] b[i - s] (spectrum: A o conj(B))
template <int P, int MODE>
__global__ void __launch_bounds__(128)
circconv_dft_kernel(const bf16* __restrict__ a, int64_t a_bs, const bf16* __restrict__ b, int64_t b_bs, bf16* __restrict__ y,
                    int64_t y_bs, int C, int H, int W, int npx, int npatch, int64_t total) {
  constexpr int MT = Dft<P>::MT, NT = Dft<P>::NT, LDA = Dft<P>::LDA, LDB = Dft<P>::LDB, MAT = Dft<P>::MAT;
  extern __shared__ __align__(16) uint32_t sm[];
  uint32_t* FAr = sm;                 // F as a left operand (stride LDA)
  uint32_t* FAi = sm + MAT;
  uint32_t* FBr = sm + 2 * MAT;       // F as a right operand (stride LDB)
  uint32_t* FBi = sm + 3 * MAT;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  uint32_t* X = sm + 4 * MAT + warp * 3 * MAT;   // patch / output staging
  uint32_t* U = X + MAT;
  uint32_t* V = U + MAT;
  for (int e = threadIdx.x; e < P * P; e += 128) {
    const int j = e / P, k = e % P;
    float s, c;
    sincospif((float)((j * k) & (P - 1)) * (2.0f / (float)P), &s, &c);   // exp(-2 pi i jk / P) = c - i s
    const uint32_t cr = to_tf32(c), ci = to_tf32(-s);
    FAr[j * LDA + k] = cr;
    FAi[j * LDA + k] = ci;
    FBr[j * LDB + k] = cr;
    FBi[j * LDB + k] = ci;
  }
  __syncthreads();
  constexpr float kInv = 1.0f / (float)(P * P);
  for (int64_t item = (int64_t)blockIdx.x * 4 + warp; item < total; item += (int64_t)gridDim.x * 4) {
    const int patch = (int)(item % npatch);
    const int64_t pc = item / npatch;
    const int c = (int)(pc % C), n = (int)(pc / C);
    const int y0 = (patch / npx) * P, x0 = (patch % npx) * P;
    const int64_t plane = (int64_t)c * H * W;
    Acc<P> qr, qi;
    {   // spectrum of a
      load_patch<P>(a + (int64_t)n * a_bs + plane, H, W, y0, x0, lane, X);
      __syncwarp();
      Acc<P> ur, ui;
      dft_rows<P>(X, FAr, FAi, g, t, ur, ui);
      st_mat<P, LDA>(U, g, t, ur);
      st_mat<P, LDA>(V, g, t, ui);
      __syncwarp();
      dft_cols<P>(U, V, FBr, FBi, g, t, qr, qi);
      __syncwarp();
    }
    {   // spectrum of b, product in place (the accumulator fragments of both spectra have the same layout)
      load_patch<P>(b + (int64_t)n * b_bs + plane, H, W, y0, x0, lane, X);
      __syncwarp();
      Acc<P> ur, ui;
      dft_rows<P>(X, FAr, FAi, g, t, ur, ui);
      st_mat<P, LDA>(U, g, t, ur);
      st_mat<P, LDA>(V, g, t, ui);
      __syncwarp();
      Acc<P> kr, ki;
      dft_cols<P>(U, V, FBr, FBi, g, t, kr, ki);
      __syncwarp();
#pragma unroll
      for (int mt = 0; mt < MT; ++mt)
#pragma unroll
        for (int nt = 0; nt < NT; ++nt)
#pragma unroll
          for (int f = 0; f < 4; ++f) {
            const float ar_ = qr.v[mt][nt][f], ai_ = qi.v[mt][nt][f], br_ = kr.v[mt][nt][f];
            const float bi_ = MODE == 0 ? ki.v[mt][nt][f] : -ki.v[mt][nt][f];
            qr.v[mt][nt][f] = ar_ * br_ - ai_ * bi_;
            qi.v[mt][nt][f] = ar_ * bi_ + ai_ * br_;
          }
    }
    st_mat<P, LDB>(U, g, t, qr);    // Z as the right operand of conj(F) Z
    st_mat<P, LDB>(V, g, t, qi);
    __syncwarp();
    {
      Acc<P> br, bi;
      idft_rows<P>(U, V, FAr, FAi, g, t, br, bi);
      __syncwarp();
      st_mat<P, LDA>(U, g, t, br);
      st_mat<P, LDA>(V, g, t, bi);
      __syncwarp();
    }
    Acc<P> o;
    idft_cols<P>(U, V, FBr, FBi, g, t, o);
    // result / P^2 -> X as plain fp32 [P][LDB], then 16-byte bf16 stores
    float* Xf = reinterpret_cast<float*>(X);
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        float* p = Xf + (mt * 16 + g) * LDB + nt * 8 + 2 * t;
        *reinterpret_cast<float2*>(p) = make_float2(o.v[mt][nt][0] * kInv, o.v[mt][nt][1] * kInv);
        *reinterpret_cast<float2*>(p + 8 * LDB) = make_float2(o.v[mt][nt][2] * kInv, o.v[mt][nt][3] * kInv);
      }
    __syncwarp();
    bf16* yp = y + (int64_t)n * y_bs + plane;
    constexpr int CPR = P / 8, RPI = 32 / CPR;
    const int r0 = lane / CPR, ch = lane % CPR;
#pragma unroll
    for (int i = 0; i < P / RPI; ++i) {
      const int row = i * RPI + r0, gy = y0 + row, gx = x0 + ch * 8;
      if (gy < H && gx < W) {
        const float4 v0 = *reinterpret_cast<const float4*>(Xf + row * LDB + ch * 8);
        const float4 v1 = *reinterpret_cast<const float4*>(Xf + row * LDB + ch * 8 + 4);
        __nv_bfloat162 h0 = __floats2bfloat162_rn(v0.x, v0.y), h1 = __floats2bfloat162_rn(v0.z, v0.w);
        __nv_bfloat162 h2 = __floats2bfloat162_rn(v1.x, v1.y), h3 = __floats2bfloat162_rn(v1.z, v1.w);
        uint4 q;
        q.x = *reinterpret_cast<uint32_t*>(&h0); q.y = *reinterpret_cast<uint32_t*>(&h1);
        q.z = *reinterpret_cast<uint32_t*>(&h2); q.w = *reinterpret_cast<uint32_t*>(&h3);
        *reinterpret_cast<uint4*>(yp + (int64_t)gy * W + gx) = q;
      }
    }
    __syncwarp();   // X is overwritten by the next item's patch
  }
}

template <int P, int MODE>
static int dft_launch(const void* a, int64_t a_bs, const void* b, int64_t b_bs, void* y, int64_t y_bs, int B, int C, int H, int W,
                      cudaStream_t st) {
  constexpr size_t smem = (size_t)(4 + 4 * 3) * Dft<P>::MAT * sizeof(uint32_t);   // F copies + 3 matrices per warp
  const int npx = (W + P - 1) / P, npy = (H + P - 1) / P;
  const int64_t total = (int64_t)B * C * npx * npy;
  int64_t grid = (total + 3) / 4;
  const int64_t cap = (int64_t)num_sms() * 8;   // persistent warps, several items each
  if (grid > cap) grid = cap;
  ensure_dyn_smem((const void*)circconv_dft_kernel<P, MODE>, (int)smem);
  circconv_dft_kernel<P, MODE><<<(unsigned)grid, 128, smem, st>>>((const bf16*)a, a_bs, (const bf16*)b, b_bs, (bf16*)y, y_bs, C, H,
                                                                 W, npx, npx * npy, total);
  return check_launch("circconv_dft");
}

// -> true if the tensor-core path took the call (bf16, 16 x 16 or 32 x 32 patches, 16-byte aligned rows)
bool circconv_dft_launch(const void* a, int64_t a_bs, const void* b, int64_t b_bs, void* y, int64_t y_bs, int B, int C, int H,
                         int W, int patch, int mode, cudaStream_t st, int* rc) {
  auto al = [](const void* p, int64_t bs) { return ((uintptr_t)p & 15) == 0 && bs % 8 == 0; };
  if ((patch != 32 && patch != 16) || W % 8 != 0 || !al(a, a_bs) || !al(b, b_bs) || !al(y, y_bs)) return false;
  if (patch == 32)
    *rc = mode == 0 ? dft_launch<32, 0>(a, a_bs, b, b_bs, y, y_bs, B, C, H, W, st) : dft_launch<32, 1>(a, a_bs, b, b_bs, y, y_bs, B, C, H, W, st);
  else
    *rc = mode == 0 ? dft_launch<16, 0>(a, a_bs, b, b_bs, y, y_bs, B, C, H, W, st) : dft_launch<16, 1>(a, a_bs, b, b_bs, y, y_bs, B, C, H, W, st);
  return true;
}

}  // namespace b200ir
