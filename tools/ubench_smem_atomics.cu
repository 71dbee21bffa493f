// Micro-benchmark: shared-memory atomic (ATOMS.ADD) throughput on sm_100a under the address
// patterns the histogram kernel produces.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t mix(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x;
}

// mode 0: lane-distinct consecutive words; 1: random in `range` words; 2: one address per warp;
// 3: two addresses per warp; 4: random but g/h interleaved (stride-2 slots, +1 for second add)
template <int MODE>
__global__ void __launch_bounds__(1024, 1) k_atoms(int iters, int range, unsigned long long* cycles, int* sink) {
  extern __shared__ int sm[];
  for (int i = threadIdx.x; i < range * 2; i += blockDim.x) sm[i] = 0;
  __syncthreads();
  uint32_t s = mix(threadIdx.x * 2654435761u + blockIdx.x);
  const int lane = threadIdx.x & 31;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    int a;
    if (MODE == 0) a = (threadIdx.x + it * 32) % range;
    else if (MODE == 1 || MODE == 4) { s = s * 1664525u + 1013904223u; a = (s >> 8) % range; }
    else if (MODE == 2) a = (it * 7) % range;
    else { s = s * 1664525u + 1013904223u; a = ((it * 7) % (range - 1)) + ((s >> 20) & 1); }
    if (MODE == 4) { atomicAdd(&sm[2 * a], lane); atomicAdd(&sm[2 * a + 1], it); }
    else { atomicAdd(&sm[a], lane); atomicAdd(&sm[a + range], it); }
  }
  __syncthreads();
  long long t1 = clock64();
  if (threadIdx.x == 0) cycles[blockIdx.x] = (unsigned long long)(t1 - t0);
  if (sm[threadIdx.x % range] == 0x7fffffff) sink[0] = 1;
}

template <int MODE>
void run(const char* name, int range, int sms) {
  unsigned long long* cyc; int* sink;
  cudaMalloc(&cyc, sizeof(unsigned long long) * sms); cudaMalloc(&sink, 4);
  int iters = 2048, smem = range * 2 * 4;
  cudaFuncSetAttribute(k_atoms<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  k_atoms<MODE><<<sms, 1024, smem>>>(64, range, cyc, sink);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  k_atoms<MODE><<<sms, 1024, smem>>>(iters, range, cyc, sink);
  cudaEventRecord(e1); cudaDeviceSynchronize();
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  unsigned long long h[256]; cudaMemcpy(h, cyc, sizeof(unsigned long long) * sms, cudaMemcpyDeviceToHost);
  double avg = 0; for (int i = 0; i < sms; ++i) avg += h[i]; avg /= sms;
  double atoms = 2.0 * iters * 1024;
  printf("%-34s range=%6d  %.2f lane-atomics/clk/SM  (%.1f clk per warp-instr)  chip %.1f G atomics/s  err=%s\n", name, range,
         atoms / avg, avg / (2.0 * iters * 32) * 1.0, atoms * sms / (ms * 1e6), cudaGetErrorString(cudaGetLastError()));
  cudaFree(cyc); cudaFree(sink);
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  printf("%s  SMs=%d  smem optin=%zu\n", p.name, p.multiProcessorCount, p.sharedMemPerBlockOptin);
  int sms = p.multiProcessorCount;
  run<0>("distinct consecutive", 18500, sms);
  run<1>("random, separate G/H arrays", 18500, sms);
  run<1>("random, separate G/H arrays", 1024, sms);
  run<4>("random, interleaved g/h", 18500, sms);
  run<2>("one address per warp", 18500, sms);
  run<3>("two addresses per warp (binary)", 18500, sms);
  return 0;
}
