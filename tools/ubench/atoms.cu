// Micro-benchmark: shared-memory 32-bit atomic add pairs (lo + carry -> hi) with lane = column (conflict-free banks),
// against the non-atomic LDS.64 / IADD / STS.64 read-modify-write, for 1..32 warps per SM.
#include <cstdio>
#include <cuda_runtime.h>
#define ROWS 256
__global__ void k_atoms(const unsigned char* bins, int iters, unsigned long long* out, long long* cyc) {
  extern __shared__ unsigned smem[];
  unsigned* lo = smem; unsigned* hi = smem + ROWS * 32;
  const int lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < ROWS * 64; i += blockDim.x) smem[i] = 0;
  __syncthreads();
  long long t0 = clock64();
  unsigned long long v = 0x1234567890ull + threadIdx.x;
  const unsigned char* b = bins + (threadIdx.x >> 5) * 4096 + lane;
  for (int it = 0; it < iters; ++it) {
#pragma unroll 8
    for (int j = 0; j < 64; ++j) {
      const unsigned bin = b[j * 32 + (it & 1) * 2048];
      const unsigned vl = (unsigned)v, vh = (unsigned)(v >> 32);
      const unsigned old = atomicAdd(&lo[bin * 32 + lane], vl);
      const unsigned carry = (old + vl) < old;
      atomicAdd(&hi[bin * 32 + lane], vh + carry);
    }
  }
  long long t1 = clock64();
  __syncthreads();
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
  if (threadIdx.x < 32) out[blockIdx.x * 32 + lane] = lo[lane] + ((unsigned long long)hi[lane] << 32);
}
__global__ void k_rmw(const unsigned char* bins, int iters, unsigned long long* out, long long* cyc) {
  extern __shared__ unsigned smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned long long* H = reinterpret_cast<unsigned long long*>(smem) + (warp % 3) * ROWS * 32;
  for (int i = threadIdx.x; i < 3 * ROWS * 64; i += blockDim.x) smem[i] = 0;
  __syncthreads();
  long long t0 = clock64();
  unsigned long long v = 0x1234567890ull + threadIdx.x;
  const unsigned char* b = bins + warp * 4096 + lane;
  for (int it = 0; it < iters; ++it) {
#pragma unroll 8
    for (int j = 0; j < 64; ++j) {
      const unsigned bin = b[j * 32 + (it & 1) * 2048];
      H[bin * 32 + lane] += v;
    }
  }
  long long t1 = clock64();
  __syncthreads();
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
  if (threadIdx.x < 32) out[blockIdx.x * 32 + lane] = H[lane];
}
int main() {
  unsigned char* bins; unsigned long long* out; long long* cyc;
  cudaMalloc(&bins, 32 * 4096); cudaMalloc(&out, 148 * 32 * 8); cudaMalloc(&cyc, 148 * 8);
  unsigned char* h = new unsigned char[32 * 4096];
  srand(1); for (int i = 0; i < 32 * 4096; ++i) h[i] = rand() & 255;
  cudaMemcpy(bins, h, 32 * 4096, cudaMemcpyHostToDevice);
  cudaFuncSetAttribute(k_atoms, cudaFuncAttributeMaxDynamicSharedMemorySize, ROWS * 64 * 4);
  cudaFuncSetAttribute(k_rmw, cudaFuncAttributeMaxDynamicSharedMemorySize, 3 * ROWS * 64 * 4);
  const int iters = 200;
  long long hc[148];
  for (int warps = 1; warps <= 32; warps *= 2) {
    k_atoms<<<148, warps * 32, ROWS * 64 * 4>>>(bins, iters, out, cyc);
    cudaDeviceSynchronize();
    k_atoms<<<148, warps * 32, ROWS * 64 * 4>>>(bins, iters, out, cyc);
    cudaError_t e = cudaDeviceSynchronize();
    cudaMemcpy(hc, cyc, sizeof(hc), cudaMemcpyDeviceToHost);
    printf("atoms pair : %2d warps/SM  %.2f cycles per row-item per SM (%s)\n", warps, (double)hc[0] / (iters * 64.0 * warps), cudaGetErrorString(e));
  }
  for (int warps = 1; warps <= 3; ++warps) {
    k_rmw<<<148, warps * 32, 3 * ROWS * 64 * 4>>>(bins, iters, out, cyc);
    cudaDeviceSynchronize();
    k_rmw<<<148, warps * 32, 3 * ROWS * 64 * 4>>>(bins, iters, out, cyc);
    cudaError_t e = cudaDeviceSynchronize();
    cudaMemcpy(hc, cyc, sizeof(hc), cudaMemcpyDeviceToHost);
    printf("rmw private: %2d warps/SM  %.2f cycles per row-item per SM (%s)\n", warps, (double)hc[0] / (iters * 64.0 * warps), cudaGetErrorString(e));
  }
  return 0;
}
