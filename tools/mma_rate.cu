// Microbenchmark: cycles per tcgen05.mma kind::tf32 (SS mode, K-major SWIZZLE_128B operands) for several (M, N),
// issued back to back by one thread on resident (garbage-free, zeroed) shared memory.  Guides the tile-shape choices in
// csrc/ (DESIGN.md section 4).   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_rate mma_rate.cu
#include <cstdio>
#include <cuda_runtime.h>

#include "../uuo-mocap_b200/csrc/umma.cuh"

__global__ void __launch_bounds__(128, 1) rate_kernel(int M, int N, int n_mma, int n_acc, int same_ab, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = umma::smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* sm = smem_raw + (base - raw);
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  for (int i = threadIdx.x; i < 96 * 1024 / 4; i += blockDim.x) reinterpret_cast<float*>(sm)[i] = 0.f;
  umma::fence_proxy_async_smem();
  if (threadIdx.x == 0) { umma::mbar_init(umma::smem_u32(&bar), 1); umma::fence_barrier_init(); }
  if (threadIdx.x < 32) umma::tmem_alloc(umma::smem_u32(&slot), 512);
  umma::tc_fence_before();
  __syncthreads();
  umma::tc_fence_after();
  const uint32_t tm = slot;
  if (threadIdx.x < 32) {
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (((uint32_t)N >> 3) << 17) | (((uint32_t)M >> 4) << 24);
    const uint32_t a0 = base, b0 = base + 32 * 1024;
    uint64_t da[2][4], db[2][4];
    for (int t = 0; t < 2; ++t)
      for (int k = 0; k < 4; ++k) {
        da[t][k] = umma::smem_desc_sw128(a0 + (same_ab ? 0 : t * 16384) + k * 32);
        db[t][k] = umma::smem_desc_sw128(b0 + (same_ab ? 0 : t * 32768) + k * 32);
      }
    const uint32_t d0 = tm, d1 = tm + (n_acc > 1 ? 256 : 0);
    long long t0 = clock64();
    if (umma::elect_one()) {
      for (int i = 0; i < n_mma; i += 8) {
#pragma unroll
        for (int t = 0; t < 2; ++t)
#pragma unroll
          for (int k = 0; k < 4; ++k) umma::mma_tf32((k & 1) ? d1 : d0, da[t][k], db[t][k], idesc, 1u);
      }
      umma::mma_commit(umma::smem_u32(&bar));
    }
    __syncwarp();
    umma::mbar_wait(umma::smem_u32(&bar), 0);
    long long t1 = clock64();
    if (threadIdx.x == 0) out[0] = t1 - t0;
  }
  umma::tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) umma::tmem_dealloc(tm, 512);
}

int main() {
  long long* d;
  cudaMalloc(&d, 8);
  const int smem = 1024 + 96 * 1024;
  cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int Ms[] = {64, 128}, Ns[] = {16, 32, 64, 128, 192, 240, 256};
  printf("M   N   n_acc  cycles/MMA (2048 MMAs, 1 CTA)   [same A/B tiles reused vs alternating]\n");
  for (int M : Ms)
    for (int N : Ns)
      for (int n_acc : {1, 2})
        for (int same : {1, 0}) {
          if (N * n_acc > 512) continue;
          const int n = 2048;
          rate_kernel<<<1, 128, smem>>>(M, N, n, n_acc, same, d);
          long long c = 0;
          cudaError_t e = cudaMemcpy(&c, d, 8, cudaMemcpyDeviceToHost);
          if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
          printf("%3d %3d %5d  %8.1f  %s\n", M, N, n_acc, (double)c / n, same ? "same" : "alt");
        }
  // all SMs busy at once: does the per-MMA cost change (power / clocks)?
  rate_kernel<<<148, 128, smem>>>(128, 256, 4096, 2, 0, d);
  long long c = 0;
  cudaMemcpy(&c, d, 8, cudaMemcpyDeviceToHost);
  printf("148 CTAs M=128 N=256: %.1f cycles/MMA\n", (double)c / 4096);
  return 0;
}
