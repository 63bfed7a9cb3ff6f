// Issue-rate probe for the CUDA-core instructions the fused depthwise conv leans on (B200, sm_100a):
// warp-instructions per cycle per SM sub-partition for FFMA (3 registers), FFMA2 (fma.rn.f32x2), HFMA2 (f16x2 / bf16x2),
// FMUL2, F2FP (cvt.rn.f16x2.f32 / bf16x2), MUFU.TANH, LOP3.  8 independent chains per thread, 16 warps per SM (4 / SMSP).
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipe_probe pipe_probe.cu && ./pipe_probe
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

constexpr int kIters = 4096, kChains = 8;

template <int OP>
__global__ void __launch_bounds__(512) probe(float* out, long long* cycles, float seed) {
  float a[kChains], b[kChains];
  uint64_t p[kChains];
  uint32_t h[kChains];
  for (int i = 0; i < kChains; ++i) {
    a[i] = seed + i + threadIdx.x;
    b[i] = seed * 0.5f + i;
    float2 t = make_float2(a[i], b[i]);
    p[i] = *reinterpret_cast<uint64_t*>(&t);
    h[i] = 0x3c003c00u + i;
  }
  const float w = seed * 0.999f, c = seed * 1e-3f;
  float2 w2f = make_float2(w, w), c2f = make_float2(c, c);
  const uint64_t w2 = *reinterpret_cast<uint64_t*>(&w2f), c2 = *reinterpret_cast<uint64_t*>(&c2f);
  const uint32_t wh = 0x3bff3bffu, ch = 0x10001000u;
  __syncthreads();
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < kIters; ++it) {
#pragma unroll
    for (int i = 0; i < kChains; ++i) {
      if (OP == 0) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[i]) : "f"(w), "f"(c));
      if (OP == 1) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(w2), "l"(c2));
      if (OP == 2) asm volatile("fma.rn.f16x2 %0, %0, %1, %2;" : "+r"(h[i]) : "r"(wh), "r"(ch));
      if (OP == 3) asm volatile("fma.rn.bf16x2 %0, %0, %1, %2;" : "+r"(h[i]) : "r"(wh), "r"(ch));
      if (OP == 4) asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(p[i]) : "l"(w2));
      if (OP == 5) asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(h[i]) : "f"(a[i]), "f"(b[i]));
      if (OP == 6) asm volatile("tanh.approx.f32 %0, %0;" : "+f"(a[i]));
      if (OP == 7) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(h[i]) : "r"(wh), "r"(ch));
      if (OP == 8) asm volatile("fma.rn.f32 %0, %0, %1, 0f3F000000;" : "+f"(a[i]) : "f"(w));   // immediate addend
      if (OP == 9) asm volatile("tanh.approx.f16x2 %0, %0;" : "+r"(h[i]));
      if (OP == 10) asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(h[i]) : "f"(a[i]), "f"(b[i]));
    }
    if (OP == 5 || OP == 10) for (int i = 0; i < kChains; ++i) a[i] += __uint_as_float(h[i] & 1u);   // keep a dependency, ~free
  }
  const long long t1 = clock64();
  float s = 0.f;
  for (int i = 0; i < kChains; ++i) {
    float2 t = *reinterpret_cast<float2*>(&p[i]);
    s += a[i] + t.x + t.y + __uint_as_float(h[i]);
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int OP>
void run(const char* name, float* out, long long* cyc, int sms) {
  probe<OP><<<sms, 512>>>(out, cyc, 1.0f);
  cudaDeviceSynchronize();
  probe<OP><<<sms, 512>>>(out, cyc, 1.0f);
  cudaDeviceSynchronize();
  long long h[256];
  cudaMemcpy(h, cyc, sizeof(long long) * sms, cudaMemcpyDeviceToHost);
  double avg = 0;
  for (int i = 0; i < sms; ++i) avg += (double)h[i];
  avg /= sms;
  const double inst_per_smsp = 4.0 * kIters * kChains;   // 4 warps per sub-partition
  printf("%-28s %7.3f warp-instr / cycle / SMSP   (%.2f cycles per instruction)\n", name, inst_per_smsp / avg, avg / inst_per_smsp);
}

int main() {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  float* out; long long* cyc;
  cudaMalloc(&out, sizeof(float) * sms * 512);
  cudaMalloc(&cyc, sizeof(long long) * sms);
  run<0>("FFMA  (3 registers)", out, cyc, sms);
  run<8>("FFMA  (immediate addend)", out, cyc, sms);
  run<1>("FFMA2 fma.rn.f32x2", out, cyc, sms);
  run<4>("FMUL2 mul.rn.f32x2", out, cyc, sms);
  run<2>("HFMA2 fma.rn.f16x2", out, cyc, sms);
  run<3>("HFMA2.BF16 fma.rn.bf16x2", out, cyc, sms);
  run<5>("F2FP  cvt.rn.f16x2.f32", out, cyc, sms);
  run<10>("F2FP  cvt.rn.bf16x2.f32", out, cyc, sms);
  run<6>("MUFU.TANH f32", out, cyc, sms);
  run<9>("MUFU.TANH f16x2", out, cyc, sms);
  run<7>("LOP3", out, cyc, sms);
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
