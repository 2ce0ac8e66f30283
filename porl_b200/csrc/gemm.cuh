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
//
// A_lo / B_lo / C_lo: optional second planes the FFMA kernel adds on load / splits on store
// (legacy of the first 3xTF32 design, which kept (hi, lo) planes in HBM; the tcgen05 kernel now
// splits in shared memory and takes plain fp32 operands, so these stay null on the product path).
struct GemmGroup {
  const float* A;
  const float* B;
  float* C;
  const float* A_lo;
  const float* B_lo;
  float* C_lo;
  const float* bias;      // (N) added once, or null
  const float* relu_src;  // (M, ld_relu): C is zeroed where relu_src <= 0 (ReLU backward), or null
  float* colsum;          // (M): sum_k A(m,k) (bias gradient in the dW form), or null (FFMA path only)
  float* Ct;              // optional transposed copy of the result: Ct[n * ldct + m] (tcgen05 path only), or null
  uint32_t* bits_out;     // tcgen05 path, epilogue 1: (M, ld_bits) words, bit j of word w = C[m, 32 w + j] > 0, or null
  const uint32_t* bits_in;  // tcgen05 path: ReLU mask as such bits (takes precedence over relu_src), or null
  int ld_bits;            // words per row, a multiple of 4 (rows are read with one 16-byte load per 128-column tile)
  // tcgen05 path: BatchNorm + ReLU of the layer that produced A, applied on the way in (encoders)
  const float* a_affine;  // (2, K): A is read as relu(A * a_affine[k] + a_affine[K + k]) (K-major A, K % 32 == 0), or null
  int M, N, K;
  int lda, ldb, ldc, ld_relu, ldct;
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
// hi = x rounded to tf32 (10-bit mantissa, round half away), exactly representable in tf32
__host__ __device__ inline float tf32_hi(float x) {
#ifdef __CUDA_ARCH__
  return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u);
#else
  union { float f; unsigned u; } v;
  v.f = x;
  v.u = (v.u + 0x1000u) & 0xffffe000u;
  return v.f;
#endif
}

// ---- helpers of the tensor-core mode (transpose.cu) ----
constexpr int kMaxSegs = 16;
// dst (cols, ld_dst) = src (rows, ld_src)^T, plain fp32
struct TransposeSeg { const float* src; float* dst; int rows, cols, ld_src, ld_dst; };
struct TransposeBatch { TransposeSeg seg[kMaxSegs]; int n_seg; };
struct ColsumSeg { const float* src; float* out; int64_t rows; int n, ld; };   // out[j] = sum_r src[r, j]
struct ColsumBatch { ColsumSeg seg[kMaxGroups]; int n_seg; int max_n; };
int launch_transpose(porl_ctx* ctx, const TransposeBatch& tb, cudaStream_t st);
int launch_colsum(porl_ctx* ctx, ColsumBatch& cb, cudaStream_t st);

// Several grouped contractions in ONE persistent tcgen05 launch (gemm_tc.cu): phase p may read what phase p-1 wrote when
// barrier_before is set (grid-wide barrier); phases without it share the CTAs freely.
struct TcPhaseSpec {
  GemmBatch* batch;
  bool a_kmajor, b_kmajor;
  bool barrier_before;
};
int launch_gemm_tc_phases(porl_ctx* ctx, int n_phases, TcPhaseSpec* specs, cudaStream_t stream);
// the same with the argument checks and the profiling scope of launch_gemm (api.cu); PORL_ERR_UNSUPPORTED when some
// problem needs the FFMA kernel or the context is in FFMA mode -- the caller then launches phase by phase
int launch_gemm_phases(porl_ctx* ctx, int n_phases, TcPhaseSpec* specs, cudaStream_t stream);

// tcgen05 3xTF32 path; returns PORL_ERR_UNSUPPORTED when the shapes do not fit it.
bool gemm_tc_eligible(const GemmGroup& g, bool a_kmajor, bool b_kmajor);
int launch_gemm_tc(porl_ctx* ctx, bool a_kmajor, bool b_kmajor, GemmBatch& batch, cudaStream_t stream);

}  // namespace porl
