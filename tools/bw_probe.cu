// Hand-written HBM bandwidth probes for B200 (the write roof quoted in DESIGN.md rested on
// torch's fill_).  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o bw_probe bw_probe.cu && ./bw_probe
// Four kernels over 2 GiB buffers (>> 126 MB L2), grid = 8 CTAs per SM, 128-bit accesses, CUDA-event timed:
//   write-only  st.global.v4           read-only  ld.global.v4 (+ xor sink)
//   copy        ld + st                write-only through TMA bulk stores (cp.async.bulk shared -> global, 16 KB per CTA step)
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__global__ void k_write(uint4* __restrict__ p, size_t n) {
  const uint4 v = make_uint4(1, 2, 3, 4);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = v;
}
__global__ void k_read(const uint4* __restrict__ p, size_t n, uint32_t* sink) {
  uint32_t acc = 0;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const uint4 v = p[i];
    acc ^= v.x ^ v.y ^ v.z ^ v.w;
  }
  if (acc == 0x12345678u) *sink = acc;
}
__global__ void k_copy(const uint4* __restrict__ a, uint4* __restrict__ b, size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) b[i] = a[i];
}
// TMA bulk store: each CTA fills a 16 KB shared tile once and streams it out with cp.async.bulk (no tensor map needed)
__global__ void k_write_tma(char* __restrict__ p, size_t bytes) {
  constexpr int kTile = 16384;
  __shared__ __align__(128) char tile[kTile];
  for (int i = threadIdx.x; i < kTile / 16; i += blockDim.x) reinterpret_cast<uint4*>(tile)[i] = make_uint4(1, 2, 3, 4);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(tile);
    int inflight = 0;
    for (size_t off = (size_t)blockIdx.x * kTile; off + kTile <= bytes; off += (size_t)gridDim.x * kTile) {
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(p + off), "r"(s), "r"(kTile) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      if (++inflight >= 8) asm volatile("cp.async.bulk.wait_group.read 4;" ::: "memory");
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
}

int main() {
  const size_t bytes = (size_t)2 << 30, n = bytes / 16;
  char *a, *b;
  uint32_t* sink;
  CK(cudaMalloc(&a, bytes)); CK(cudaMalloc(&b, bytes)); CK(cudaMalloc(&sink, 4));
  CK(cudaMemset(a, 1, bytes)); CK(cudaMemset(b, 2, bytes));
  int sms = 0;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  const int grid = sms * 8, it = 10;
  for (int which = 0; which < 4; ++which) {
    float best = 1e30f, sum = 0.f;
    for (int r = 0; r < it + 3; ++r) {
      CK(cudaEventRecord(e0));
      if (which == 0) k_write<<<grid, 256>>>((uint4*)a, n);
      else if (which == 1) k_read<<<grid, 256>>>((const uint4*)a, n, sink);
      else if (which == 2) k_copy<<<grid, 256>>>((const uint4*)a, (uint4*)b, n);
      else k_write_tma<<<sms * 4, 128>>>(a, bytes);
      CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); CK(cudaGetLastError());
      float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
      if (r >= 3) { sum += ms; best = ms < best ? ms : best; }
    }
    const double moved = (which == 2 ? 2.0 : 1.0) * (double)bytes;
    const char* name[] = {"write-only st.global.v4", "read-only  ld.global.v4", "copy       ld+st v4   ", "write-only TMA bulk st "};
    printf("%s  mean %.0f GB/s  best %.0f GB/s\n", name[which], moved / (sum / it) / 1e6, moved / best / 1e6);
  }
  return 0;
}
