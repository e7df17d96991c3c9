// Micro-benchmark: throughput of REDG.ADD.64 flushes from many blocks into one 114 KB histogram, for two
// lane->address assignments: (A) lane = (channel, feature) -> 16 sectors per instruction, (B) lanes = consecutive words.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k_red(unsigned long long* out, int pattern, int nonzero_every) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const int fl = lane & 15, ch = lane >> 4;
  for (int b = warp; b < 256; b += nw) {
    unsigned long long* p;
    if (pattern == 0) p = out + ((size_t)fl * 256 + b) * 2 + ch;
    else p = out + (size_t)b * 32 + lane;
    if ((b % nonzero_every) == 0) atomicAdd(p, (unsigned long long)(threadIdx.x + 1));
  }
}
int main() {
  unsigned long long* out; cudaMalloc(&out, 16 * 256 * 16); cudaMemset(out, 0, 16 * 256 * 16);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int blocks : {37, 74, 148, 296, 592}) for (int pattern = 0; pattern < 2; ++pattern) {
    k_red<<<blocks, 512>>>(out, pattern, 1);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    for (int r = 0; r < 20; ++r) k_red<<<blocks, 512>>>(out, pattern, 1);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    printf("blocks %4d pattern %c: %.2f us per launch, %.1f REDs/ns\n", blocks, pattern ? 'B' : 'A', ms * 1e3 / 20, blocks * 8192.0 / (ms * 1e6 / 20));
  }
  return 0;
}
