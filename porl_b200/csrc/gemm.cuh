// Grouped contraction interface shared by the FFMA (gemm_simt.cu) and tcgen05 (gemm_tc.cu) paths.
#pragma once
#include "common.cuh"

namespace porl {

// One problem of a grouped launch:  C[M,N] = opA(A)[M,K] * opB(B)[K,N]  (+ epilogue)
//   a_kmajor: A(m,k) = A[m*lda + k]   else A(m,k) = A[k*lda + m]
//   b_kmajor: B(k,n) = B[n*ldb + k]   else B(k,n) = B[k*ldb + n]
// Linear forward   y = x W^T      : A = x (kmajor),  B = W (kmajor)
// Linear input grad dx = dy W     : A = dy (kmajor), B = W (not kmajor)
// Linear weight grad dW = dy^T x  : A = dy (not kmajor), B = x (not kmajor), K = batch
struct GemmGroup {
  const float* A;
  const float* B;
  float* C;
  const float* bias;      // (N) added once, or null
  const float* relu_src;  // (M, ld_relu): C is zeroed where relu_src <= 0 (ReLU backward), or null
  float* colsum;          // (M): sum_k A(m,k) (bias gradient in the dW form), or null
  int M, N, K;
  int lda, ldb, ldc, ld_relu;
  int n_store;            // >= N; columns N..n_store-1 are written as zeros
  int epilogue;           // 0 none, 1 relu, 2 tanh
};

struct GemmBatch {
  GemmGroup g[kMaxGroups];
  int n_groups;
  int splits;             // split-K factor (serialised through turnstile words, deterministic)
  int* sem;
  int sem_stride;         // turnstile words reserved per group
};

// Validates and launches one grouped contraction; all groups share the trans flags.
int launch_gemm(porl_ctx* ctx, bool a_kmajor, bool b_kmajor, GemmBatch& batch, cudaStream_t stream);
int launch_gemm_simt(porl_ctx* ctx, bool a_kmajor, bool b_kmajor, GemmBatch& batch, cudaStream_t stream);
// tcgen05 3xTF32 path; returns PORL_ERR_UNSUPPORTED when the shapes do not fit it.
int launch_gemm_tc(porl_ctx* ctx, bool a_kmajor, bool b_kmajor, GemmBatch& batch, cudaStream_t stream);

}  // namespace porl
