// pdl_probe.cu — what does programmatic dependent launch buy for a chain of small dependent kernels inside a CUDA graph?
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pdl_probe tools/pdl_probe.cu && ./pdl_probe
#include <cstdio>
#include <cuda_runtime.h>

template <bool PDL>
__global__ void __launch_bounds__(256) step_kernel(const float* __restrict__ in, float* __restrict__ out, int n) {
  if (PDL) {
    asm volatile("griddepcontrol.launch_dependents;");
    asm volatile("griddepcontrol.wait;" ::: "memory");
  }
  for (int i = blockIdx.x * 256 + threadIdx.x; i < n; i += gridDim.x * 256) out[i] = in[i] * 1.0001f + 1.0f;
}

template <bool PDL>
static void launch(const float* in, float* out, int n, int grid, cudaStream_t st) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid); cfg.blockDim = dim3(256); cfg.dynamicSmemBytes = 0; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = PDL ? 1 : 0;
  cudaLaunchKernelEx(&cfg, step_kernel<PDL>, in, out, n);
}

template <bool PDL>
static float run(int n, int grid, int chain) {
  float *a, *b;
  cudaMalloc(&a, n * 4); cudaMalloc(&b, n * 4);
  cudaMemset(a, 0, n * 4);
  cudaStream_t st; cudaStreamCreate(&st);
  cudaGraph_t g; cudaGraphExec_t ge;
  cudaStreamBeginCapture(st, cudaStreamCaptureModeGlobal);
  for (int i = 0; i < chain; ++i) { launch<PDL>(a, b, n, grid, st); float* t = a; a = b; b = t; }
  cudaStreamEndCapture(st, &g);
  cudaGraphInstantiate(&ge, g, 0);
  cudaGraphLaunch(ge, st); cudaStreamSynchronize(st);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0, st);
  for (int r = 0; r < 5; ++r) cudaGraphLaunch(ge, st);
  cudaEventRecord(e1, st); cudaStreamSynchronize(st);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  float h; cudaMemcpy(&h, a, 4, cudaMemcpyDeviceToHost);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) printf("CUDA error: %s\n", cudaGetErrorString(e));
  cudaFree(a); cudaFree(b);
  printf("  %s n=%8d grid=%4d: %.2f us per kernel (check %.3f)\n", PDL ? "PDL   " : "normal", n, grid, ms * 1e3f / (5 * chain), h);
  return ms;
}

int main() {
  const int chain = 1000;
  for (int n : {1 << 14, 1 << 18, 1 << 22}) {
    for (int grid : {16, 148, 592}) {
      run<false>(n, grid, chain);
      run<true>(n, grid, chain);
    }
  }
  return 0;
}
