// Micro-benchmark: cycles per tcgen05.mma.kind::tf32 (M=128, K=8) as a function of N, issued back to back by one
// thread into one TMEM accumulator (dependent) or round-robin over four (independent), operands in shared memory.
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I../multidisttorch_b200/csrc -o mma_rate mma_rate.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "gemm_tcgen05.cuh"
using namespace mdt;

template <int N>
__global__ void __launch_bounds__(128, 1) rate_kernel(long long* out, int reps, int nacc) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint32_t slot;
  __shared__ __align__(8) unsigned long long bar;
  const uint32_t sbase = ptx::smem_u32(smem);
  for (int i = threadIdx.x; i < (128 * 128 + 256 * 128) / 4; i += blockDim.x) reinterpret_cast<float*>(smem)[i] = 1.0f;
  if (threadIdx.x == 0) { ptx::mbar_init(ptx::smem_u32(&bar), 1); ptx::fence_mbar_init(); }
  if (threadIdx.x < 32) { ptx::tmem_alloc(ptx::smem_u32(&slot), 512); ptx::tmem_relinquish(); }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  ptx::fence_proxy_async();
  const uint32_t tm = slot;
  if (threadIdx.x == 0) {
    constexpr uint32_t idesc = umma_idesc_tf32(128, N);
    const uint64_t a = umma_desc_sw128(sbase), b = umma_desc_sw128(sbase + 128 * 128);
    const int stride = (nacc > 1) ? (N < 128 ? N : 128) : 0;
    const long long t0 = clock64();
    for (int r = 0; r < reps; ++r) {
#pragma unroll
      for (int k = 0; k < 4; ++k)
        ptx::mma_tf32(tm + (uint32_t)(((r * 4 + k) % nacc) * stride), a + (uint64_t)(k * 2), b + (uint64_t)(k * 2), idesc, 1u);
    }
    const long long t1 = clock64();
    ptx::mma_commit(ptx::smem_u32(&bar));
    ptx::mbar_wait(ptx::smem_u32(&bar), 0);
    const long long t2 = clock64();
    out[0] = t1 - t0;
    out[1] = t2 - t0;
  }
  __syncthreads();
  if (threadIdx.x < 32) ptx::tmem_dealloc(tm, 512);
}

template <int N>
void run(long long* d, int nacc) {
  const int reps = 64;
  const size_t smem = 128 * 128 + 256 * 128 + 1024;
  cudaFuncSetAttribute(rate_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  long long h[2];
  for (int it = 0; it < 2; ++it) {
    rate_kernel<N><<<1, 128, smem>>>(d, reps, nacc);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("N=%d: %s\n", N, cudaGetErrorString(e)); return; }
  }
  cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  printf("N=%3d acc=%d: issue %.1f cycles/MMA, issue+complete %.1f cycles/MMA (%d MMAs)\n", N, nacc, h[0] / (reps * 4.0),
         h[1] / (reps * 4.0), reps * 4);
}

int main() {
  long long* d;
  cudaMalloc(&d, 64);
  for (int nacc : {1, 4}) {
    run<16>(d, nacc); run<32>(d, nacc); run<64>(d, nacc); run<128>(d, nacc); run<256>(d, nacc);
  }
  return 0;
}
