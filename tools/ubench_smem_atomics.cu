// Micro-benchmark: shared-memory atomic (ATOMS.ADD) throughput on sm_100a under the address
// patterns a histogram kernel can produce.  Addresses are precomputed per thread (8 per thread,
// advanced by a cheap add + mask) so the loop body is atomics plus ~2 ALU ops each.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t mix(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x;
}
constexpr int kWords = 16384;  // words per array (G) ; H array follows

// MODE 0 consecutive (lane-distinct, one 128B line per instruction)
//      1 random word in the array (random banks)
//      2 one address per warp
//      3 two addresses per warp
//      4 own bank (word = row*32 + lane), random row per lane
//      5 own bank, each lane alternates between 2 rows
//      6 random, g/h interleaved (adjacent words)
template <int MODE>
__global__ void __launch_bounds__(1024, 1) k_atoms(int iters, unsigned long long* cycles, int* sink) {
  extern __shared__ int sm[];
  for (int i = threadIdx.x; i < kWords * 2; i += blockDim.x) sm[i] = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int a[8];
  for (int k = 0; k < 8; ++k) {
    uint32_t r = mix(threadIdx.x * 8 + k + blockIdx.x * 8192);
    uint32_t rw = mix(warp * 8 + k + 77);
    if (MODE == 0) a[k] = (warp * 256 + k * 32 + lane) & (kWords - 1);
    else if (MODE == 1 || MODE == 6) a[k] = r & (kWords - 1);
    else if (MODE == 2) a[k] = rw & (kWords - 1);
    else if (MODE == 3) a[k] = ((rw & (kWords - 1)) & ~1) | (r & 1);
    else if (MODE == 4) a[k] = ((r >> 5) & (kWords / 32 - 1)) * 32 + lane;
    else a[k] = (((rw >> 5) + (r & 1) * 3) & (kWords / 32 - 1)) * 32 + lane;
  }
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      if (MODE == 6) { int x = (a[k] & ~1); atomicAdd(&sm[x], lane); atomicAdd(&sm[x + 1], it); }
      else { atomicAdd(&sm[a[k]], lane); atomicAdd(&sm[a[k] + kWords], it); }
      if (MODE == 1 || MODE == 6) a[k] = (a[k] * 5 + 1) & (kWords - 1);          // stays pseudo-random
      else if (MODE == 4) a[k] = (a[k] + 32 * 37) & (kWords - 1);                // next row, same bank
    }
  }
  __syncthreads();
  long long t1 = clock64();
  if (threadIdx.x == 0) cycles[blockIdx.x] = (unsigned long long)(t1 - t0);
  if (sm[threadIdx.x] == 0x7fffffff) sink[0] = 1;
}

template <int MODE>
void run(const char* name, int sms) {
  unsigned long long* cyc; int* sink;
  cudaMalloc(&cyc, sizeof(unsigned long long) * sms); cudaMalloc(&sink, 4);
  int iters = 512, smem = kWords * 2 * 4;
  cudaFuncSetAttribute(k_atoms<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  k_atoms<MODE><<<sms, 1024, smem>>>(16, cyc, sink);
  k_atoms<MODE><<<sms, 1024, smem>>>(iters, cyc, sink);
  cudaDeviceSynchronize();
  unsigned long long h[256]; cudaMemcpy(h, cyc, sizeof(unsigned long long) * sms, cudaMemcpyDeviceToHost);
  double avg = 0; for (int i = 0; i < sms; ++i) avg += h[i]; avg /= sms;
  double atoms = 16.0 * iters * 1024;
  printf("%-42s %6.2f lane-atomics/clk/SM  (%5.2f clk per warp ATOMS)  err=%s\n", name, atoms / avg,
         avg / (16.0 * iters * 32), cudaGetErrorString(cudaGetLastError()));
  cudaFree(cyc); cudaFree(sink);
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  printf("%s  SMs=%d  smem optin=%zu  (1024 threads/SM, 2 x 64 KB arrays)\n", p.name, p.multiProcessorCount, p.sharedMemPerBlockOptin);
  int sms = p.multiProcessorCount;
  run<0>("consecutive words (1 line, 32 banks)", sms);
  run<1>("random words, separate G/H arrays", sms);
  run<6>("random words, g/h interleaved", sms);
  run<4>("own bank per lane, random row", sms);
  run<5>("own bank per lane, 2 rows per lane", sms);
  run<2>("one address per warp", sms);
  run<3>("two addresses per warp (binary feature)", sms);
  return 0;
}
