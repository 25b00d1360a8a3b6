// Micro-benchmark (GPU box): latency and issue throughput of FFMA vs the packed FFMA2 (fma.rn.f32x2) on sm_100a.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ffma2 ffma2.cu ; run: ./ffma2
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ unsigned long long ffma2(unsigned long long a, unsigned long long b, unsigned long long c) {
  unsigned long long d;
  asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ float ffma1(float a, float b, float c) {
  float d;
  asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}

template <int CHAINS, bool PACKED>
__global__ void k(float* out, long long* cyc, int iters) {
  float a = 1.0001f + threadIdx.x * 1e-7f, b = 0.9999f;
  unsigned long long a2, b2;
  { float2 t = {a, a}; a2 = *reinterpret_cast<unsigned long long*>(&t); t = {b, b}; b2 = *reinterpret_cast<unsigned long long*>(&t); }
  float acc[CHAINS];
  unsigned long long acc2[CHAINS];
  for (int i = 0; i < CHAINS; i++) { acc[i] = i; float2 t = {float(i), float(i)}; acc2[i] = *reinterpret_cast<unsigned long long*>(&t); }
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < CHAINS; i++) {
      if (PACKED) acc2[i] = ffma2(acc2[i], a2, b2);
      else acc[i] = ffma1(acc[i], a, b);
    }
  }
  long long t1 = clock64();
  float s = 0;
  for (int i = 0; i < CHAINS; i++) { s += acc[i]; float2 t = *reinterpret_cast<float2*>(&acc2[i]); s += t.x + t.y; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int CHAINS, bool PACKED>
void run(int threads, const char* what) {
  float* out; long long* cyc; long long h;
  cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 8);
  const int iters = 4096;
  k<CHAINS, PACKED><<<148, threads>>>(out, cyc, iters);
  k<CHAINS, PACKED><<<148, threads>>>(out, cyc, iters);
  cudaDeviceSynchronize();
  cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  const double per = double(h) / iters / CHAINS;   // cycles per instruction of one warp
  const double warps = threads / 32.0;
  printf("%-6s chains=%d warps/SM=%2.0f: %.2f cycles per instr per warp -> %.2f warp-instr/clk/SM, %.1f FMA lanes/clk/SM\n", what,
         CHAINS, warps, per, warps / per, warps / per * 32 * (PACKED ? 2 : 1));
  cudaFree(out); cudaFree(cyc);
}

int main() {
  run<1, false>(32, "FFMA");  run<1, true>(32, "FFMA2");      // latency
  run<8, false>(32, "FFMA");  run<8, true>(32, "FFMA2");      // one warp, 8 chains
  run<8, false>(128, "FFMA"); run<8, true>(128, "FFMA2");     // one warp per scheduler
  run<8, false>(512, "FFMA"); run<8, true>(512, "FFMA2");     // 4 warps per scheduler
  run<8, false>(1024, "FFMA"); run<8, true>(1024, "FFMA2");
  return 0;
}
