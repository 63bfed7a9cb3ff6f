// Probe: 4D TMA tile loads (SWIZZLE_128B, bf16) with negative / out-of-range coordinates and a box wider than
// the tensor.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -I image-restoration_b200/csrc tools/tma4d_probe.cu \
//                          image-restoration_b200/csrc/build/gemm_tc.o ... (see tools/build_probe.sh)
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../image-restoration_b200/csrc/tc_common.cuh"
using namespace b200ir;
using namespace b200ir::tc;

__global__ void probe(const __grid_constant__ CUtensorMap tm, int c0, int c1, int c2, int c3, int bk, bf16* out) {
  extern __shared__ uint8_t raw[];
  uint8_t* sm = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    fence_barrier_init();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    mbar_arrive_expect_tx(&bar, bk * 128);
    tma_load_4d(sm, &tm, &bar, c0, c1, c2, c3);
  }
  mbar_wait(&bar, 0);
  for (int i = threadIdx.x; i < bk * 64; i += blockDim.x) {
    const int r = i / 64, p = i % 64;
    out[i] = *(bf16*)(sm + r * 128 + (((p >> 3) ^ (r & 7)) << 4) + (p & 7) * 2);
  }
}

int main(int argc, char** argv) {
  int a0 = argc > 1 ? atoi(argv[1]) : 0, a1 = argc > 2 ? atoi(argv[2]) : 3;
  const int B = 2, C = 48, H = 16;
  for (int W : {64, 40}) {
    std::vector<bf16> h((size_t)B * C * H * W);
    for (size_t i = 0; i < h.size(); ++i) h[i] = __float2bfloat16((float)(i % 251));
    bf16 *d, *o;
    cudaMalloc(&d, h.size() * 2);
    cudaMalloc(&o, 64 * 64 * 2);
    cudaMemcpy(d, h.data(), h.size() * 2, cudaMemcpyHostToDevice);
    CUtensorMap tm;
    const uint64_t dims[4] = {(uint64_t)W, (uint64_t)H, (uint64_t)C, (uint64_t)B};
    const uint64_t str[4] = {2, (uint64_t)W * 2, (uint64_t)H * W * 2, (uint64_t)C * H * W * 2};
    const uint32_t box[4] = {64, 1, 48, 1};
    int rc = make_tmap_bf16(&tm, d, 4, dims, str, box);
    printf("W=%d encode rc=%d (%s)\n", W, rc, rc ? b200ir_last_error_string() : "ok");
    if (rc) continue;
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768);
    const int coords[][4] = {{a0, a1, 0, 1}};
    for (auto& c : coords) {
      cudaMemset(o, 0xff, 64 * 64 * 2);
      probe<<<1, 128, 16384 + 1024>>>(tm, c[0], c[1], c[2], c[3], 48, o);
      cudaError_t e = cudaDeviceSynchronize();
      std::vector<bf16> r(64 * 48);
      cudaMemcpy(r.data(), o, r.size() * 2, cudaMemcpyDeviceToHost);
      int bad = 0;
      for (int ch = 0; ch < 48 && e == cudaSuccess; ++ch)
        for (int p = 0; p < 64; ++p) {
          const int x = c[0] + p, y = c[1], cc = c[2] + ch, b = c[3];
          float want = 0.f;
          if (x >= 0 && x < W && y >= 0 && y < H && cc < C && b < B)
            want = (float)(((((size_t)b * C + cc) * H + y) * W + x) % 251);
          if (__bfloat162float(r[ch * 64 + p]) != want) ++bad;
        }
      printf("  coords (%d,%d,%d,%d): %s, mismatches %d\n", c[0], c[1], c[2], c[3], cudaGetErrorString(e), bad);
      if (e != cudaSuccess) return 1;
    }
    cudaFree(d);
    cudaFree(o);
  }
  return 0;
}
