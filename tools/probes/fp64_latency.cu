// Latency probes (one warp): dependent DFMA / DMUL / MUFU.RCP64H chains, LDS after a block barrier, named barrier.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void probe(double* out, long long* t, double x, int warps_sync) {
  __shared__ double sh[64];
  double a = x, b = x * 0.5 + 1.0;
  long long t0 = clock64();
#pragma unroll
  for (int i = 0; i < 256; ++i) a = fma(a, b, 1e-9);
  long long t1 = clock64();
#pragma unroll
  for (int i = 0; i < 256; ++i) a = a * b;
  long long t2 = clock64();
#pragma unroll
  for (int i = 0; i < 64; ++i) { double r; asm volatile("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(a)); a = r; }
  long long t3 = clock64();
  // independent DFMA throughput: 8 chains
  double c[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) c[j] = a + j;
#pragma unroll
  for (int i = 0; i < 64; ++i) {
#pragma unroll
    for (int j = 0; j < 8; ++j) c[j] = fma(c[j], b, 1e-9);
  }
  long long t4 = clock64();
  for (int j = 0; j < 8; ++j) a += c[j];
  // barrier + STS + LDS round trip
  long long t5 = clock64();
#pragma unroll 1
  for (int i = 0; i < 64; ++i) {
    sh[threadIdx.x & 63] = a;
    __syncthreads();
    a += sh[(threadIdx.x + 1) & 63];
  }
  long long t6 = clock64();
  // float chain for comparison
  float f = (float)x, g = f * 0.5f + 1.0f;
  long long t7 = clock64();
#pragma unroll
  for (int i = 0; i < 256; ++i) f = fmaf(f, g, 1e-9f);
  long long t8 = clock64();
  if (threadIdx.x == 0) {
    t[0] = t1 - t0; t[1] = t2 - t1; t[2] = t3 - t2; t[3] = t4 - t3; t[4] = t6 - t5; t[5] = t8 - t7;
  }
  out[threadIdx.x] = a + f;
}
int main() {
  double* out; long long* t;
  cudaMalloc(&out, 8 * 1024); cudaMalloc(&t, 64);
  for (int threads : {32, 128, 256, 512}) {
    probe<<<1, threads>>>(out, t, 1.0000001, 0);
    cudaDeviceSynchronize();
    long long h[8];
    cudaMemcpy(h, t, 64, cudaMemcpyDeviceToHost);
    printf("threads %3d: DFMA dep %.1f clk  DMUL dep %.1f  MUFU.RCP64H dep %.1f  DFMA 8 indep chains %.1f clk/instr  STS+BAR+LDS+DADD round %.1f  FFMA dep %.1f\n",
           threads, h[0] / 256.0, h[1] / 256.0, h[2] / 64.0, h[3] / 512.0, h[4] / 64.0, h[5] / 256.0);
  }
  return 0;
}
