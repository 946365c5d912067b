// Random-gather throughput of HBM3e on B200: every thread reads CH-byte chunks at pseudo-random STRIDE-byte slots of a 6 GB table.
// Sizing evidence for the batched-affine level-1 kernel.  nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/gather_probe tools/gather_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
template <int CH16, int ILP>
__global__ void gather(const uint4* __restrict__ tab, uint64_t slots, uint32_t stride16, int iters, uint4* out) {
  uint64_t s = (blockIdx.x * (uint64_t)blockDim.x + threadIdx.x) * 0x9E3779B97F4A7C15ull + 12345;
  uint4 acc = make_uint4(0, 0, 0, 0);
  for (int it = 0; it < iters; it++) {
    uint4 v[ILP][CH16];
#pragma unroll
    for (int k = 0; k < ILP; k++) {
      s = s * 6364136223846793005ull + 1442695040888963407ull;
      const uint64_t slot = (s >> 20) % slots;
      const uint4* p = tab + slot * stride16;
#pragma unroll
      for (int j = 0; j < CH16; j++) v[k][j] = __ldg(p + j);
    }
#pragma unroll
    for (int k = 0; k < ILP; k++)
#pragma unroll
      for (int j = 0; j < CH16; j++) { acc.x ^= v[k][j].x; acc.y ^= v[k][j].y; acc.z ^= v[k][j].z; acc.w ^= v[k][j].w; }
  }
  out[blockIdx.x * (uint64_t)blockDim.x + threadIdx.x] = acc;
}
template <int CH16, int ILP> void run(const uint4* tab, size_t bytes, uint32_t stride, uint4* out, int blocks_per_sm) {
  const uint64_t slots = bytes / stride;
  const int blocks = 148 * blocks_per_sm, threads = 128, iters = 256;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  gather<CH16, ILP><<<blocks, threads>>>(tab, slots, stride / 16, 8, out);
  cudaEventRecord(e0);
  gather<CH16, ILP><<<blocks, threads>>>(tab, slots, stride / 16, iters, out);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double n = (double)blocks * threads * iters * ILP;
  printf("chunk %3d B  stride %3u B  ILP %d  CTAs/SM %d: %7.2f G gathers/s  %8.1f GB/s useful\n", CH16 * 16, stride, ILP, blocks_per_sm, n / ms / 1e6, n * CH16 * 16 / ms / 1e6);
}
int main() {
  const size_t bytes = 6ull << 30;
  uint4 *tab, *out;
  cudaMalloc(&tab, bytes); cudaMemset(tab, 1, bytes);
  cudaMalloc(&out, 148 * 16 * 128 * sizeof(uint4));
  for (int bps : {3, 8, 16}) {
    run<3, 1>(tab, bytes, 96, out, bps);    // x coordinate only of 96-byte points
    run<6, 1>(tab, bytes, 96, out, bps);    // 96-byte points, packed
    run<3, 1>(tab, bytes, 128, out, bps);   // x only, padded points
    run<6, 1>(tab, bytes, 128, out, bps);   // padded to 128
    run<6, 2>(tab, bytes, 96, out, bps);
    run<6, 4>(tab, bytes, 96, out, bps);
    run<6, 4>(tab, bytes, 128, out, bps);
    run<12, 2>(tab, bytes, 192, out, bps);  // G2 points
    run<4, 4>(tab, bytes, 64, out, bps);
  }
  return 0;
}
