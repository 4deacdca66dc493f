// Microbenchmark: latency / throughput of redux.sync.min.f32 (CREDUX) vs a 5-step shuffle butterfly on sm_100a.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/redux_rate tools/redux_rate.cu && ./tools/redux_rate
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ float redux_min_f32(float v) {
  float r;
  asm volatile("redux.sync.min.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
  return r;
}
__device__ __forceinline__ float shfl_min(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
template <int MODE>
__global__ void k(float* out, long long* cyc, int iters) {
  float v = threadIdx.x * 1.5f + 1.f, w = v + 3.f, x = v * 0.5f, y = v - 7.f;
  long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
    if (MODE == 0) { v = redux_min_f32(v) + threadIdx.x; }                       // dependent chain
    if (MODE == 1) { v = shfl_min(v) + threadIdx.x; }
    if (MODE == 2) { v = redux_min_f32(v) + threadIdx.x; w = redux_min_f32(w) + threadIdx.x; x = redux_min_f32(x) + threadIdx.x; y = redux_min_f32(y) + threadIdx.x; }
    if (MODE == 3) { v = shfl_min(v) + threadIdx.x; w = shfl_min(w) + threadIdx.x; x = shfl_min(x) + threadIdx.x; y = shfl_min(y) + threadIdx.x; }
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = v + w + x + y;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main() {
  float* out; long long* cyc; cudaMalloc(&out, 1 << 22); cudaMalloc(&cyc, 8);
  const int iters = 2000;
  for (int warps : {1, 4, 8, 16, 32}) {
    long long c[4];
    k<0><<<1, warps * 32>>>(out, cyc, iters); cudaMemcpy(&c[0], cyc, 8, cudaMemcpyDeviceToHost);
    k<1><<<1, warps * 32>>>(out, cyc, iters); cudaMemcpy(&c[1], cyc, 8, cudaMemcpyDeviceToHost);
    k<2><<<1, warps * 32>>>(out, cyc, iters); cudaMemcpy(&c[2], cyc, 8, cudaMemcpyDeviceToHost);
    k<3><<<1, warps * 32>>>(out, cyc, iters); cudaMemcpy(&c[3], cyc, 8, cudaMemcpyDeviceToHost);
    printf("warps/SM %2d: cycles per reduction  redux chain %.1f  shfl chain %.1f | 4 independent: redux %.1f  shfl %.1f\n", warps,
           (double)c[0] / iters, (double)c[1] / iters, (double)c[2] / iters / 4, (double)c[3] / iters / 4);
  }
  return 0;
}
