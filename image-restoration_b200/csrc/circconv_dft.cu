// circconv_dft.cu — the 32 x 32 FFT expert of MoCE-IR on tensor cores (bf16 path only).
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

constexpr int kLdA = 36;   // row stride (floats) of a matrix read as the LEFT mma operand  (bank = 4 * row + col: conflict-free)
constexpr int kLdB = 40;   // row stride of a matrix read as the RIGHT mma operand          (bank = 8 * k + n:    conflict-free)
constexpr int kMat = 32 * kLdB;   // floats reserved per 32 x 32 matrix (either stride fits)

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
__device__ __forceinline__ void ld_left(const uint32_t* m, int mt, int kt, int g, int t, uint32_t (&a)[4]) {
  const uint32_t* p = m + (mt * 16 + g) * kLdA + kt * 8 + t;
  a[0] = p[0];
  a[1] = p[8 * kLdA];
  a[2] = p[4];
  a[3] = p[8 * kLdA + 4];
}
__device__ __forceinline__ void ld_right(const uint32_t* m, int kt, int nt, int g, int t, uint32_t& b0, uint32_t& b1) {
  const uint32_t* p = m + (kt * 8 + t) * kLdB + nt * 8 + g;
  b0 = p[0];
  b1 = p[4 * kLdB];
}
// accumulator tile (mt, nt) -> matrix with row stride LD, rounded to TF32 (it is the operand of the next product)
template <int LD>
__device__ __forceinline__ void st_acc(uint32_t* m, int mt, int nt, int g, int t, const float (&c)[4]) {
  uint32_t* p = m + (mt * 16 + g) * LD + nt * 8 + 2 * t;
  *reinterpret_cast<uint2*>(p) = make_uint2(to_tf32(c[0]), to_tf32(c[1]));
  *reinterpret_cast<uint2*>(p + 8 * LD) = make_uint2(to_tf32(c[2]), to_tf32(c[3]));
}

typedef float Acc[2][4][4];   // [m tile][n tile][fragment]: one real 32 x 32 matrix in the registers of a warp

__device__ __forceinline__ void zero_acc(Acc& c) {
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int k = 0; k < 4; ++k) c[i][j][k] = 0.f;
}

// ---- the four kinds of product (all 32 x 32 x 32, per warp) -------------------------------------------------------
// rows:  Ur = Fr X, Ui = Fi X                      X = right operand in shared memory, F = left operand copies (FA)
__device__ __forceinline__ void dft_rows(const uint32_t* X, const uint32_t* FAr, const uint32_t* FAi, int g, int t, Acc& ur,
                                         Acc& ui) {
  zero_acc(ur);
  zero_acc(ui);
#pragma unroll
  for (int kt = 0; kt < 4; ++kt) {
    uint32_t b0[4], b1[4];
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) ld_right(X, kt, nt, g, t, b0[nt], b1[nt]);
#pragma unroll
    for (int mt = 0; mt < 2; ++mt) {
      uint32_t fr[4], fi[4];
      ld_left(FAr, mt, kt, g, t, fr);
      ld_left(FAi, mt, kt, g, t, fi);
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
        mma_tf32(ur[mt][nt], fr, b0[nt], b1[nt]);
        mma_tf32(ui[mt][nt], fi, b0[nt], b1[nt]);
      }
    }
  }
}

// cols:  Yr = Ur Fr - Ui Fi, Yi = Ur Fi + Ui Fr    U = left operands in shared memory, F = right operand copies (FB)
__device__ __forceinline__ void dft_cols(const uint32_t* Ur, const uint32_t* Ui, const uint32_t* FBr, const uint32_t* FBi,
                                         int g, int t, Acc& yr, Acc& yi) {
  zero_acc(yr);
  zero_acc(yi);
#pragma unroll
  for (int kt = 0; kt < 4; ++kt) {
    uint32_t r0[4], r1[4], i0[4], i1[4];
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) {
      ld_right(FBr, kt, nt, g, t, r0[nt], r1[nt]);
      ld_right(FBi, kt, nt, g, t, i0[nt], i1[nt]);
    }
#pragma unroll
    for (int mt = 0; mt < 2; ++mt) {
      uint32_t ar[4], ai[4];
      ld_left(Ur, mt, kt, g, t, ar);
      ld_left(Ui, mt, kt, g, t, ai);
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
        mma_tf32(yr[mt][nt], ar, r0[nt], r1[nt]);
        mma_tf32(yr[mt][nt], ai, i0[nt] ^ 0x80000000u, i1[nt] ^ 0x80000000u);
        mma_tf32(yi[mt][nt], ar, i0[nt], i1[nt]);
        mma_tf32(yi[mt][nt], ai, r0[nt], r1[nt]);
      }
    }
  }
}

// inverse rows:  Br = Fr Zr + Fi Zi, Bi = Fr Zi - Fi Zr     Z = right operands in shared memory
__device__ __forceinline__ void idft_rows(const uint32_t* Zr, const uint32_t* Zi, const uint32_t* FAr, const uint32_t* FAi,
                                          int g, int t, Acc& br, Acc& bi) {
  zero_acc(br);
  zero_acc(bi);
#pragma unroll
  for (int kt = 0; kt < 4; ++kt) {
    uint32_t r0[4], r1[4], i0[4], i1[4];
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) {
      ld_right(Zr, kt, nt, g, t, r0[nt], r1[nt]);
      ld_right(Zi, kt, nt, g, t, i0[nt], i1[nt]);
    }
#pragma unroll
    for (int mt = 0; mt < 2; ++mt) {
      uint32_t fr[4], fi[4];
      ld_left(FAr, mt, kt, g, t, fr);
      ld_left(FAi, mt, kt, g, t, fi);
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
        mma_tf32(br[mt][nt], fr, r0[nt], r1[nt]);
        mma_tf32(br[mt][nt], fi, i0[nt], i1[nt]);
        mma_tf32(bi[mt][nt], fr, i0[nt], i1[nt]);
        mma_tf32(bi[mt][nt], fi, r0[nt] ^ 0x80000000u, r1[nt] ^ 0x80000000u);
      }
    }
  }
}

// inverse cols (real part only):  O = Br Fr + Bi Fi
__device__ __forceinline__ void idft_cols(const uint32_t* Br, const uint32_t* Bi, const uint32_t* FBr, const uint32_t* FBi,
                                          int g, int t, Acc& o) {
  zero_acc(o);
#pragma unroll
  for (int kt = 0; kt < 4; ++kt) {
    uint32_t r0[4], r1[4], i0[4], i1[4];
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) {
      ld_right(FBr, kt, nt, g, t, r0[nt], r1[nt]);
      ld_right(FBi, kt, nt, g, t, i0[nt], i1[nt]);
    }
#pragma unroll
    for (int mt = 0; mt < 2; ++mt) {
      uint32_t ar[4], ai[4];
      ld_left(Br, mt, kt, g, t, ar);
      ld_left(Bi, mt, kt, g, t, ai);
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
        mma_tf32(o[mt][nt], ar, r0[nt], r1[nt]);
        mma_tf32(o[mt][nt], ai, i0[nt], i1[nt]);
      }
    }
  }
}

template <int LD>
__device__ __forceinline__ void st_mat(uint32_t* m, int g, int t, const Acc& c) {
#pragma unroll
  for (int mt = 0; mt < 2; ++mt)
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) st_acc<LD>(m, mt, nt, g, t, c[mt][nt]);
}

// one 32 x 32 bf16 patch -> X (right-operand layout, fp32 bits: bf16 values are exact in TF32); zero outside the plane
__device__ __forceinline__ void load_patch(const bf16* __restrict__ plane, int H, int W, int y0, int x0, int lane, uint32_t* X) {
  const int r8 = lane >> 2, ch = lane & 3;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int row = i * 8 + r8, gy = y0 + row, gx = x0 + ch * 8;
    uint4 q = make_uint4(0u, 0u, 0u, 0u);
    if (gy < H && gx < W) q = __ldg(reinterpret_cast<const uint4*>(plane + (int64_t)gy * W + gx));   // W % 8 == 0
    uint32_t* d = X + row * kLdB + ch * 8;
    *reinterpret_cast<uint4*>(d) = make_uint4(q.x << 16, q.x & 0xffff0000u, q.y << 16, q.y & 0xffff0000u);
    *reinterpret_cast<uint4*>(d + 4) = make_uint4(q.z << 16, q.z & 0xffff0000u, q.w << 16, q.w & 0xffff0000u);
  }
}

// MODE 0: y = a (*) b circular convolution;  MODE 1: y[s] = sum_i a[i] b[i - s] (spectrum: A o conj(B))
template <int MODE>
__global__ void __launch_bounds__(128)
circconv_dft32_kernel(const bf16* __restrict__ a, int64_t a_bs, const bf16* __restrict__ b, int64_t b_bs, bf16* __restrict__ y,
                      int64_t y_bs, int C, int H, int W, int npx, int npatch, int64_t total) {
  extern __shared__ __align__(16) uint32_t sm[];
  uint32_t* FAr = sm;                 // F as a left operand (stride kLdA)
  uint32_t* FAi = sm + kMat;
  uint32_t* FBr = sm + 2 * kMat;      // F as a right operand (stride kLdB)
  uint32_t* FBi = sm + 3 * kMat;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  uint32_t* X = sm + 4 * kMat + warp * 3 * kMat;   // patch / output staging
  uint32_t* U = X + kMat;
  uint32_t* V = U + kMat;
  for (int e = threadIdx.x; e < 1024; e += 128) {
    const int j = e >> 5, k = e & 31;
    float s, c;
    sincospif((float)((j * k) & 31) * (1.0f / 16.0f), &s, &c);   // exp(-2 pi i jk / 32) = c - i s
    const uint32_t cr = to_tf32(c), ci = to_tf32(-s);
    FAr[j * kLdA + k] = cr;
    FAi[j * kLdA + k] = ci;
    FBr[j * kLdB + k] = cr;
    FBi[j * kLdB + k] = ci;
  }
  __syncthreads();
  for (int64_t item = (int64_t)blockIdx.x * 4 + warp; item < total; item += (int64_t)gridDim.x * 4) {
    const int patch = (int)(item % npatch);
    const int64_t pc = item / npatch;
    const int c = (int)(pc % C), n = (int)(pc / C);
    const int y0 = (patch / npx) * 32, x0 = (patch % npx) * 32;
    const int64_t plane = (int64_t)c * H * W;
    Acc qr, qi;
    {   // spectrum of a
      load_patch(a + (int64_t)n * a_bs + plane, H, W, y0, x0, lane, X);
      __syncwarp();
      Acc ur, ui;
      dft_rows(X, FAr, FAi, g, t, ur, ui);
      st_mat<kLdA>(U, g, t, ur);
      st_mat<kLdA>(V, g, t, ui);
      __syncwarp();
      dft_cols(U, V, FBr, FBi, g, t, qr, qi);
      __syncwarp();
    }
    {   // spectrum of b, product in place (the accumulator fragments of both spectra have the same layout)
      load_patch(b + (int64_t)n * b_bs + plane, H, W, y0, x0, lane, X);
      __syncwarp();
      Acc ur, ui;
      dft_rows(X, FAr, FAi, g, t, ur, ui);
      st_mat<kLdA>(U, g, t, ur);
      st_mat<kLdA>(V, g, t, ui);
      __syncwarp();
      Acc kr, ki;
      dft_cols(U, V, FBr, FBi, g, t, kr, ki);
      __syncwarp();
#pragma unroll
      for (int mt = 0; mt < 2; ++mt)
#pragma unroll
        for (int nt = 0; nt < 4; ++nt)
#pragma unroll
          for (int f = 0; f < 4; ++f) {
            const float ar_ = qr[mt][nt][f], ai_ = qi[mt][nt][f], br_ = kr[mt][nt][f], bi_ = MODE == 0 ? ki[mt][nt][f] : -ki[mt][nt][f];
            qr[mt][nt][f] = ar_ * br_ - ai_ * bi_;
            qi[mt][nt][f] = ar_ * bi_ + ai_ * br_;
          }
    }
    st_mat<kLdB>(U, g, t, qr);    // Z as the right operand of conj(F) Z
    st_mat<kLdB>(V, g, t, qi);
    __syncwarp();
    {
      Acc br, bi;
      idft_rows(U, V, FAr, FAi, g, t, br, bi);
      __syncwarp();
      st_mat<kLdA>(U, g, t, br);
      st_mat<kLdA>(V, g, t, bi);
      __syncwarp();
    }
    Acc o;
    idft_cols(U, V, FBr, FBi, g, t, o);
    // result / 1024 -> X as plain fp32 [32][kLdB], then 16-byte bf16 stores (64 B per patch row)
    float* Xf = reinterpret_cast<float*>(X);
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
        float* p = Xf + (mt * 16 + g) * kLdB + nt * 8 + 2 * t;
        *reinterpret_cast<float2*>(p) = make_float2(o[mt][nt][0] * (1.0f / 1024.0f), o[mt][nt][1] * (1.0f / 1024.0f));
        *reinterpret_cast<float2*>(p + 8 * kLdB) = make_float2(o[mt][nt][2] * (1.0f / 1024.0f), o[mt][nt][3] * (1.0f / 1024.0f));
      }
    __syncwarp();
    bf16* yp = y + (int64_t)n * y_bs + plane;
    const int r8 = lane >> 2, ch = lane & 3;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int row = i * 8 + r8, gy = y0 + row, gx = x0 + ch * 8;
      if (gy < H && gx < W) {
        const float4 v0 = *reinterpret_cast<const float4*>(Xf + row * kLdB + ch * 8);
        const float4 v1 = *reinterpret_cast<const float4*>(Xf + row * kLdB + ch * 8 + 4);
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

constexpr size_t kDftSmem = (size_t)(4 + 4 * 3) * kMat * sizeof(uint32_t);   // F copies + 3 matrices per warp = 80 KB

// -> true if the tensor-core path took the call (bf16, 32 x 32 patches, 16-byte aligned rows)
bool circconv_dft32_launch(const void* a, int64_t a_bs, const void* b, int64_t b_bs, void* y, int64_t y_bs, int B, int C, int H,
                           int W, int mode, cudaStream_t st, int* rc) {
  auto al = [](const void* p, int64_t bs) { return ((uintptr_t)p & 15) == 0 && bs % 8 == 0; };
  if (W % 8 != 0 || !al(a, a_bs) || !al(b, b_bs) || !al(y, y_bs)) return false;
  const int npx = (W + 31) / 32, npy = (H + 31) / 32;
  const int64_t total = (int64_t)B * C * npx * npy;
  int64_t grid = (total + 3) / 4;
  const int64_t cap = (int64_t)num_sms() * 2 * 4;   // 2 CTAs per SM resident (shared memory); a few waves of items each
  if (grid > cap) grid = cap;
  if (mode == 0) {
    ensure_dyn_smem((const void*)circconv_dft32_kernel<0>, (int)kDftSmem);
    circconv_dft32_kernel<0><<<(unsigned)grid, 128, kDftSmem, st>>>((const bf16*)a, a_bs, (const bf16*)b, b_bs, (bf16*)y, y_bs, C,
                                                                   H, W, npx, npx * npy, total);
  } else {
    ensure_dyn_smem((const void*)circconv_dft32_kernel<1>, (int)kDftSmem);
    circconv_dft32_kernel<1><<<(unsigned)grid, 128, kDftSmem, st>>>((const bf16*)a, a_bs, (const bf16*)b, b_bs, (bf16*)y, y_bs, C,
                                                                   H, W, npx, npx * npy, total);
  }
  *rc = check_launch("circconv_dft32");
  return true;
}

}  // namespace b200ir
