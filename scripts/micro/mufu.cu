// MUFU throughput probe: tanh.approx.f32 / ex2.approx / rcp.approx, lane-ops per clock per SM.
#include <cstdio>
#include <cuda_runtime.h>
template <int OP> __device__ __forceinline__ float f(float x) {
  float r;
  if (OP == 0) asm volatile("tanh.approx.f32 %0, %1;" : "=f"(r) : "f"(x));
  else if (OP == 1) asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  else asm volatile("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
template <int OP> __global__ void k(float* out, int iters, long long* clk) {
  float a[8];
  for (int i = 0; i < 8; ++i) a[i] = threadIdx.x * 0.001f + i;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it)
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = f<OP>(a[i]);
  long long t1 = clock64();
  float s = 0; for (int i = 0; i < 8; ++i) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *clk = t1 - t0;
}
int main() {
  float* o; long long* c; cudaMalloc(&o, 148 * 1024 * 4); cudaMallocManaged(&c, 8);
  const int iters = 4096;
  for (int op = 0; op < 3; ++op) {
    if (op == 0) k<0><<<148, 1024>>>(o, iters, c); else if (op == 1) k<1><<<148, 1024>>>(o, iters, c); else k<2><<<148, 1024>>>(o, iters, c);
    cudaDeviceSynchronize();
    printf("op %d: %.2f lane-ops/clk/SM\n", op, 1024.0 * 8 * iters / (double)*c);
  }
  return 0;
}
