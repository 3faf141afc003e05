// Cost of one "publish -> block barrier -> read" round as a function of warps and of stores in flight.
#include <cstdio>
#include <cuda_runtime.h>
template <int NSTS, bool NAMED>
__global__ void probe(double* out, long long* t) {
  __shared__ __align__(16) double sh[2][1024];
  double a = threadIdx.x;
  const int lane = threadIdx.x & 31;
  long long t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < 64; ++i) {
    double* w = sh[i & 1];
    if (lane < 2) {
#pragma unroll
      for (int s = 0; s < NSTS; ++s) *reinterpret_cast<double2*>(w + ((threadIdx.x * NSTS + s) * 2 & 1022)) = make_double2(a, a + s);
    }
    if (NAMED) asm volatile("bar.sync 1, %0;" ::"r"((int)blockDim.x) : "memory"); else __syncthreads();
    const double2 v = *reinterpret_cast<const double2*>(w + ((threadIdx.x >> 4) * 4));
    a = fma(a, 0.5, v.x + v.y);
  }
  long long t1 = clock64();
  if (threadIdx.x == 0) t[0] = t1 - t0;
  out[threadIdx.x] = a;
}
template <int NSTS, bool NAMED> void run(int threads, double* out, long long* t) {
  probe<NSTS, NAMED><<<1, threads>>>(out, t);
  cudaDeviceSynchronize();
  long long h; cudaMemcpy(&h, t, 8, cudaMemcpyDeviceToHost);
  printf("threads %3d  sts/lane %d  %s: %.1f clk per round\n", threads, NSTS, NAMED ? "bar.sync 1" : "__syncthreads", h / 64.0);
}
int main() {
  double* out; long long* t;
  cudaMalloc(&out, 8 * 1024); cudaMalloc(&t, 64);
  for (int threads : {32, 64, 128, 256}) {
    run<1, false>(threads, out, t); run<2, false>(threads, out, t); run<5, false>(threads, out, t);
    run<2, true>(threads, out, t); run<5, true>(threads, out, t);
  }
  return 0;
}
