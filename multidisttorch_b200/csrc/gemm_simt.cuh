// fp32 FFMA GEMM tiles (MDT_GEMM_FP32): the tight-tolerance parity path and the
// on-device comparator for the tcgen05 kernels.  True fp32 multiply-add, like the
// reference's sgemm with TF32 disabled.
//
//   C[m,n] = epilogue( sum_k A(m,k) * B(k,n) )
//   A(m,k) = kAK ? A[m*lda + k] : A[k*lda + m]      (K-contiguous or M-contiguous)
//   B(k,n) = kBK ? B[n*ldb + k] : B[k*ldb + n]      (K-contiguous or N-contiguous)
//
// forward  y = x W^T        : kAK=1, kBK=1   (ref:vae-hpo.py:30-31,39-40)
// dX = dY W                 : kAK=1, kBK=0
// dW = dY^T X               : kAK=0, kBK=0   (contraction over the batch)
#pragma once

#include "common.cuh"

namespace mdt {

struct EpiBiasAct {  // y = act(acc + b[n])
  float* C;
  int ldc;
  const float* bias;  // may be NULL
  int act;
  __device__ __forceinline__ void operator()(int m, int n, float v) const {
    if (bias) v += bias[n];
    if (act == MDT_ACT_RELU) v = fmaxf(v, 0.f);
    else if (act == MDT_ACT_SIGMOID) v = sigmoidf_ref(v);
    C[(size_t)m * ldc + n] = v;
  }
};

struct EpiStore {  // plain store (weight gradients, dz)
  float* C;
  int ldc;
  __device__ __forceinline__ void operator()(int m, int n, float v) const { C[(size_t)m * ldc + n] = v; }
};

struct EpiReluMask {  // threshold_backward: pass gradient where the forward output was > 0
  float* C;
  int ldc;
  const float* H;
  int ldh;
  __device__ __forceinline__ void operator()(int m, int n, float v) const {
    C[(size_t)m * ldc + n] = H[(size_t)m * ldh + n] > 0.f ? v : 0.f;
  }
};

template <int BM, int BN, int BK, int TM, int TN, bool kAK, bool kBK, class Epi>
__global__ void __launch_bounds__((BM / TM) * (BN / TN))
gemm_simt_kernel(const float* __restrict__ A, const float* __restrict__ B, int M, int N, int K, int lda, int ldb,
                 Epi epi) {
  constexpr int NT = (BM / TM) * (BN / TN);
  __shared__ float As[BK][BM + 4];
  __shared__ float Bs[BK][BN + 4];
  const int tid = threadIdx.x;
  const int tx = tid % (BN / TN), ty = tid / (BN / TN);
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;

  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

  for (int k0 = 0; k0 < K; k0 += BK) {
#pragma unroll
    for (int idx = tid; idx < BM * BK; idx += NT) {
      int m, k;
      if (kAK) { k = idx % BK; m = idx / BK; } else { m = idx % BM; k = idx / BM; }
      const int gm = m0 + m, gk = k0 + k;
      float v = 0.f;
      if (gm < M && gk < K) v = kAK ? A[(size_t)gm * lda + gk] : A[(size_t)gk * lda + gm];
      As[k][m] = v;
    }
#pragma unroll
    for (int idx = tid; idx < BN * BK; idx += NT) {
      int n, k;
      if (kBK) { k = idx % BK; n = idx / BK; } else { n = idx % BN; k = idx / BN; }
      const int gn = n0 + n, gk = k0 + k;
      float v = 0.f;
      if (gn < N && gk < K) v = kBK ? B[(size_t)gn * ldb + gk] : B[(size_t)gk * ldb + gn];
      Bs[k][n] = v;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      float a[TM], b[TN];
#pragma unroll
      for (int i = 0; i < TM; ++i) a[i] = As[k][ty * TM + i];
#pragma unroll
      for (int j = 0; j < TN; ++j) b[j] = Bs[k][tx * TN + j];
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    const int m = m0 + ty * TM + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < TN; ++j) {
      const int n = n0 + tx * TN + j;
      if (n < N) epi(m, n, acc[i][j]);
    }
  }
}

template <bool kAK, bool kBK, class Epi>
inline cudaError_t launch_gemm_simt(const float* A, const float* B, int M, int N, int K, int lda, int ldb, Epi epi,
                                    cudaStream_t s) {
  if (M <= 0 || N <= 0) return cudaSuccess;
  constexpr int BM = 32, BN = 64, BK = 16, TM = 4, TN = 4;
  dim3 grid((N + BN - 1) / BN, (M + BM - 1) / BM);
  gemm_simt_kernel<BM, BN, BK, TM, TN, kAK, kBK, Epi><<<grid, (BM / TM) * (BN / TN), 0, s>>>(A, B, M, N, K, lda, ldb, epi);
  return cudaGetLastError();
}

}  // namespace mdt
