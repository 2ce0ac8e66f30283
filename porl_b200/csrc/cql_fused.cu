// Fused discrete-CQL gradient: replay-buffer gather + target forward + online forward + TD/CQL head + backward of the
// QNetwork S -> 64 -> 128 -> 64 -> A (src/porl/net/q_network.py:8-30) in ONE persistent kernel, weights staged in
// shared memory (the whole net is 160 KB: src/porl/train/cql_trainer.py:88-124 is launch-latency bound when every
// layer is its own launch).  A second small kernel sums the per-CTA gradient partials in a fixed order
// (deterministic) into the optimizer bucket and finalises the three loss scalars.
//
// Two tile variants of the contractions, chosen by the context's gemm mode:
//   mode 1  warp-level mma.sync m16n8k8 TF32 with the 3xTF32 hi/lo split done in registers (warp_gemm, *_mma)
//   mode 0  register-tiled fp32 FFMA (fwd_layer, dh_layer, dw_layer) -- the exact-arithmetic yardstick
//
//   tile = 32 transitions per CTA pass; 256 threads (512 with PORL_CQL_WARPS=16); shared memory (floats, KP = pad32(S)):
//     XS  [32][KP+4]     next_obs rows (target pass), then obs rows (kept for dW0)
//     WS  max(64*(KP+4), 17696)   W0, then {W1 [128][68], W2 [64][132], W3 [8][68]} of the net being evaluated
//     H1  [32][68]  H2 [32][132]  H3 [32][68]    activations, overwritten in place by dz on the way back
//     BS  [264] biases, QS/DQ [32][8], NMAX [32], ROW [32]
//   Row strides are = 4 (mod 32) floats: an MMA fragment's 8 x 4 (row, k) reads cover the 32 banks once, and the 8
//   sample rows / 4 neuron rows a warp touches per LDS.128 of the FFMA variant fall in distinct banks.
// Measurements, the per-phase profile and what bounds the kernel: DESIGN.md section 3, profiles/r01_cql_fused_ncu.txt.
#include <math.h>
#include <stdlib.h>

#include "common.cuh"
#include "optim.cuh"

extern "C" int porl_mlp_desc(const porl_mlp_t* mlp, porl_mlp_desc_t* out);

namespace porl {
namespace {

constexpr int kT = 32;            // transitions per tile
constexpr int kThreads = 256;
constexpr int kH1 = 64, kH2 = 128, kH3 = 64;
constexpr int kLd1 = kH1 + 4, kLd2 = kH2 + 4, kLd3 = kH3 + 4;
constexpr int kAP = 8;            // padded action count (A <= 8)
constexpr int kW1Floats = kH2 * kLd1, kW2Floats = kH3 * kLd2, kW3Floats = kAP * kLd3;
constexpr int kW123Floats = kW1Floats + kW2Floats + kW3Floats;
constexpr int kBiasFloats = kH1 + kH2 + kH3 + kAP;

struct CqlArgs {
  const float* params;
  const float* target;
  int64_t w_off[4], b_off[4];
  const float* states;
  const float* next_states;
  int64_t ld_states;
  const int64_t* actions;
  const float* rewards;
  const float* dones;
  const int64_t* idx;      // rows to take (null: row i = i)
  int64_t n_rows;
  int64_t B;
  int KP;                  // pad32(S)
  int A;
  float gamma, alpha, inv_bn;
  float* partials;         // [gridDim.x][part_stride]: gradient in the flat parameter layout, then td, penalty
  int64_t part_stride;
  int64_t n_floats;
  uint32_t* status;
};

__device__ __forceinline__ void cp_async16(float* dst_smem, const float* src) {
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(dst_smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_but_last() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }

__device__ __forceinline__ float dot4(const float4& a, const float4& w, float acc) {
  acc = fmaf(a.x, w.x, acc);
  acc = fmaf(a.y, w.y, acc);
  acc = fmaf(a.z, w.z, acc);
  acc = fmaf(a.w, w.w, acc);
  return acc;
}

// rows of the tile -> XS (asynchronous; rows past B are zero)
__device__ __forceinline__ void load_rows(float* XS, int ldx, const float* table, int64_t ld, const int* rowsel,
                                          int KP, int k0 = 0) {
  const int c4 = KP >> 2, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int s = warp; s < kT; s += (int)(blockDim.x >> 5)) {
    const int r = rowsel[s];
    const float* src = table + (int64_t)(r >= 0 ? r : 0) * ld + k0;
    float* d = XS + s * ldx + k0;
    for (int c = lane; c < c4; c += 32) {
      if (r >= 0) cp_async16(d + 4 * c, src + 4 * c);
      else *reinterpret_cast<float4*>(d + 4 * c) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
}

// dense (rows, cols) -> shared (rows, lds), cols % 4 == 0 (asynchronous)
__device__ __forceinline__ void load_mat(float* dst, int lds, const float* src, int rows, int cols) {
  const int c4 = cols >> 2;
  if (c4 >= 32) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int r = warp; r < rows; r += (int)(blockDim.x >> 5))
      for (int c = lane; c < c4; c += 32) cp_async16(dst + r * lds + 4 * c, src + (int64_t)r * cols + 4 * c);
  } else {          // 16 float4 per row (64-wide matrices): half a warp per row
    const int lane = threadIdx.x & 15, hw = threadIdx.x >> 4;
    for (int r = hw; r < rows; r += (int)(blockDim.x >> 4))
      for (int c = lane; c < c4; c += 16) cp_async16(dst + r * lds + 4 * c, src + (int64_t)r * cols + 4 * c);
  }
}
// columns [k0, k0 + kc) of a dense (rows, ld_src) matrix
__device__ __forceinline__ void load_mat_cols(float* dst, int lds, const float* src, int ld_src, int rows, int k0, int kc) {
  const int c4 = kc >> 2, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int r = warp; r < rows; r += (int)(blockDim.x >> 5))
    for (int c = lane; c < c4; c += 32) cp_async16(dst + r * lds + k0 + 4 * c, src + (int64_t)r * ld_src + k0 + 4 * c);
}

__device__ __forceinline__ void load_w123(float* WS, const float* p, const int64_t* w_off, int A) {
  load_mat(WS, kLd1, p + w_off[1], kH2, kH1);
  load_mat(WS + kW1Floats, kLd2, p + w_off[2], kH3, kH2);
  load_mat(WS + kW1Floats + kW2Floats, kLd3, p + w_off[3], A, kH3);
  for (int e = threadIdx.x; e < (kAP - A) * kLd3; e += (int)blockDim.x) WS[kW1Floats + kW2Floats + A * kLd3 + e] = 0.f;
}

__device__ __forceinline__ void load_bias(float* BS, const float* p, const int64_t* b_off, int A) {
  for (int e = threadIdx.x; e < kBiasFloats; e += (int)blockDim.x) {
    float v;
    if (e < kH1) v = p[b_off[0] + e];
    else if (e < kH1 + kH2) v = p[b_off[1] + e - kH1];
    else if (e < kH1 + kH2 + kH3) v = p[b_off[2] + e - kH1 - kH2];
    else v = (e - kH1 - kH2 - kH3) < A ? p[b_off[3] + e - kH1 - kH2 - kH3] : 0.f;
    BS[e] = v;
  }
}

// out[s][n] = relu(b[n] + sum_k a[s][k] w[n][k]); thread tile 4 samples (sg + 8i) x NJ neurons (ng + 32j)
template <int N>
__device__ __forceinline__ void fwd_layer(const float* a, int lda, int K, const float* w, int ldw, const float* b,
                                          float* out, int ldo) {
  constexpr int NJ = N / 32;
  const int sg = threadIdx.x & 7, ng = threadIdx.x >> 3;
  float acc[4][NJ];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < NJ; ++j) acc[i][j] = 0.f;
  const float* ap = a + sg * lda;
  const float* wp = w + ng * ldw;
#pragma unroll 2
  for (int k = 0; k < K; k += 4) {
    float4 av[4], wv[NJ];
#pragma unroll
    for (int i = 0; i < 4; ++i) av[i] = *reinterpret_cast<const float4*>(ap + i * 8 * lda + k);
#pragma unroll
    for (int j = 0; j < NJ; ++j) wv[j] = *reinterpret_cast<const float4*>(wp + j * 32 * ldw + k);
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < NJ; ++j) acc[i][j] = dot4(av[i], wv[j], acc[i][j]);
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < NJ; ++j) out[(sg + 8 * i) * ldo + ng + 32 * j] = fmaxf(acc[i][j] + b[ng + 32 * j], 0.f);
}

// q[s][a] = b[a] + sum_k h3[s][k] w3[a][k]   (no activation), a < kAP (rows >= A of w3 and b are zero)
__device__ __forceinline__ void head_fwd(const float* h3, const float* w3, const float* b, float* q) {
  if (threadIdx.x >= kT * kAP) return;
  const int s = threadIdx.x >> 3, a = threadIdx.x & 7;
  float acc = 0.f;
#pragma unroll 4
  for (int k = 0; k < kH3; k += 4)
    acc = dot4(*reinterpret_cast<const float4*>(h3 + s * kLd3 + k), *reinterpret_cast<const float4*>(w3 + a * kLd3 + k),
               acc);
  q[s * kAP + a] = acc + b[a];
}

// h[s][k] <- (sum_n dz[s][n] w[n][k]) * [h[s][k] > 0], in place; thread tile 2 samples (sg, sg+16) x CJ float4 columns
template <int K, int N>
__device__ __forceinline__ void dh_layer(const float* dz, int ldz, const float* w, int ldw, float* h, int ldh) {
  constexpr int CJ = (K / 4) / 16;
  const int cg = threadIdx.x & 15, sg = threadIdx.x >> 4;
  float4 acc[2][CJ];
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < CJ; ++j) acc[i][j] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 2
  for (int n = 0; n < N; n += 4) {
    float4 d[2];
#pragma unroll
    for (int i = 0; i < 2; ++i) d[i] = *reinterpret_cast<const float4*>(dz + (sg + 16 * i) * ldz + n);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
#pragma unroll
      for (int j = 0; j < CJ; ++j) {
        const float4 wv = *reinterpret_cast<const float4*>(w + (n + q) * ldw + 4 * (cg + 16 * j));
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          const float dv = q == 0 ? d[i].x : q == 1 ? d[i].y : q == 2 ? d[i].z : d[i].w;
          acc[i][j].x = fmaf(dv, wv.x, acc[i][j].x);
          acc[i][j].y = fmaf(dv, wv.y, acc[i][j].y);
          acc[i][j].z = fmaf(dv, wv.z, acc[i][j].z);
          acc[i][j].w = fmaf(dv, wv.w, acc[i][j].w);
        }
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < CJ; ++j) {
      float4* hp = reinterpret_cast<float4*>(h + (sg + 16 * i) * ldh + 4 * (cg + 16 * j));
      const float4 hv = *hp;
      float4 o;
      o.x = hv.x > 0.f ? acc[i][j].x : 0.f;
      o.y = hv.y > 0.f ? acc[i][j].y : 0.f;
      o.z = hv.z > 0.f ? acc[i][j].z : 0.f;
      o.w = hv.w > 0.f ? acc[i][j].w : 0.f;
      *hp = o;
    }
}

// gw[n][k] (+)= sum_s dz[s][n] a[s][k];   thread tile 4 n x CJ float4 columns of k
template <int N, int CJ>
__device__ __forceinline__ void dw_layer(const float* dz, int ldz, const float* a, int lda, int K, float* gw,
                                         bool accumulate) {
  constexpr int NG = N / 4, TK = kThreads / NG;
  const int kg = threadIdx.x % TK, ngp = threadIdx.x / TK;
  const int kc = K >> 2;
  float4 acc[4][CJ];
#pragma unroll
  for (int q = 0; q < 4; ++q)
#pragma unroll
    for (int j = 0; j < CJ; ++j) acc[q][j] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 2
  for (int s = 0; s < kT; ++s) {
    const float4 d = *reinterpret_cast<const float4*>(dz + s * ldz + 4 * ngp);
#pragma unroll
    for (int j = 0; j < CJ; ++j) {
      const int c = kg + TK * j;
      if (c < kc) {
        const float4 av = *reinterpret_cast<const float4*>(a + s * lda + 4 * c);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float dv = q == 0 ? d.x : q == 1 ? d.y : q == 2 ? d.z : d.w;
          acc[q][j].x = fmaf(dv, av.x, acc[q][j].x);
          acc[q][j].y = fmaf(dv, av.y, acc[q][j].y);
          acc[q][j].z = fmaf(dv, av.z, acc[q][j].z);
          acc[q][j].w = fmaf(dv, av.w, acc[q][j].w);
        }
      }
    }
  }
#pragma unroll
  for (int q = 0; q < 4; ++q)
#pragma unroll
    for (int j = 0; j < CJ; ++j) {
      const int c = kg + TK * j;
      if (c < kc) {
        float4* gp = reinterpret_cast<float4*>(gw + (int64_t)(4 * ngp + q) * K + 4 * c);
        float4 o = acc[q][j];
        if (accumulate) {
          const float4 p = *gp;
          o.x += p.x; o.y += p.y; o.z += p.z; o.w += p.w;
        }
        *gp = o;
      }
    }
}


// ---------------------------------------------------------------- warp-level 3xTF32 contractions (mma.sync m16n8k8)
// fp32 operands are split in registers into tf32 hi + tf32 lo; acc += lo*hi + hi*lo + hi*hi keeps fp32-grade accuracy
// (same scheme as the tcgen05 contraction of gemm_tc.cu).  Fragment reads are 32-bit LDS: with row strides = 4 (mod 32)
// the 8 x 4 (row, k) pattern of a fragment covers the 32 banks exactly once.
__device__ __forceinline__ void split_tf32(float x, uint32_t& hi, uint32_t& lo) {
  // hi = x rounded to nearest tf32 (10 mantissa bits) by integer arithmetic on the bit pattern -- the carry of the
  // rounding add propagates into the exponent as it should; lo = x - hi is exact in fp32 and is handed to the tensor
  // core as is (it reads the upper 19 bits).  cvt.rna.tf32.f32 compiles to 4 instructions with an Inf/NaN guard that
  // finite activations do not need: 3 full-rate ALU operations per element instead of 9.
  hi = (__float_as_uint(x) + 0x1000u) & 0xffffe000u;
  lo = __float_as_uint(x - __uint_as_float(hi));
}
__device__ __forceinline__ void mma_tf32(float (&c)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
  asm(
      "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

// c (16 x 8 NB) += sum_k A(m, k) B(k, n) with A(m, k) = a[m a_ms + k a_ks], B(k, n) = b[k b_ks + n b_ns];
// `a` / `b` already point at the tile's first row / column.  Thread (g = lane / 4, t = lane % 4) ends up with
// c[nb][0..3] = C(g, 8 nb + 2t), C(g, 8 nb + 2t + 1), C(g + 8, 8 nb + 2t), C(g + 8, 8 nb + 2t + 1).
// The three partial products go to NC separate accumulator chains (lo*hi, hi*lo, hi*hi -> chain 0, 1 % NC, 2 % NC):
// consecutive HMMAs then do not wait on each other, and the small terms are summed among themselves first.
template <int NB, int NC>
struct Acc {
  float c[NC][NB][4];
  __device__ __forceinline__ void zero() {
#pragma unroll
    for (int q = 0; q < NC; ++q)
#pragma unroll
      for (int nb = 0; nb < NB; ++nb)
#pragma unroll
        for (int i = 0; i < 4; ++i) c[q][nb][i] = 0.f;
  }
  // total of the chains, small-term chains first
  __device__ __forceinline__ float get(int nb, int i) const {
    float v = c[0][nb][i];
#pragma unroll
    for (int q = 1; q < NC; ++q) v += c[q][nb][i];
    return v;
  }
};

template <int NB, int NC, int UNROLL>
__device__ __forceinline__ void warp_gemm(Acc<NB, NC>& acc, const float* a, int a_ms, int a_ks, const float* b,
                                          int b_ks, int b_ns, int K) {
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const float* ap = a + g * a_ms + t * a_ks;
  const float* bp = b + t * b_ks + g * b_ns;
#pragma unroll UNROLL
  for (int k = 0; k < K; k += 8) {
    uint32_t ah[4], al[4];
    split_tf32(ap[k * a_ks], ah[0], al[0]);
    split_tf32(ap[8 * a_ms + k * a_ks], ah[1], al[1]);
    split_tf32(ap[(k + 4) * a_ks], ah[2], al[2]);
    split_tf32(ap[8 * a_ms + (k + 4) * a_ks], ah[3], al[3]);
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) {
      uint32_t bh[2], bl[2];
      split_tf32(bp[k * b_ks + nb * 8 * b_ns], bh[0], bl[0]);
      split_tf32(bp[(k + 4) * b_ks + nb * 8 * b_ns], bh[1], bl[1]);
      mma_tf32(acc.c[0][nb], al, bh);
      mma_tf32(acc.c[1 % NC][nb], ah, bl);
      mma_tf32(acc.c[2 % NC][nb], ah, bh);
    }
  }
}

// out (32 x N) = relu(a (32 x K) w^T + bias), w (N x K); 8 warps: rows 16 (w & 1), columns (N / 4) (w >> 1)
template <int N, int U, int NW>
__device__ __forceinline__ void fwd_layer_mma(const float* a, int lda, int K, const float* w, int ldw, const float* b,
                                              float* out, int ldo) {
  constexpr int NB = N / (4 * NW);         // NW / 2 column groups of 8 NB columns
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int m0 = 16 * (warp & 1), n0 = 8 * NB * (warp >> 1);
  Acc<NB, (NB <= 2 ? 3 : 2)> c;
  c.zero();
  warp_gemm<NB, (NB <= 2 ? 3 : 2), U>(c, a + m0 * lda, lda, 1, w + n0 * ldw, 1, ldw, K);
#pragma unroll
  for (int nb = 0; nb < NB; ++nb) {
    const int n = n0 + 8 * nb + 2 * t;
    const float b0 = b[n], b1 = b[n + 1];
    *reinterpret_cast<float2*>(out + (m0 + g) * ldo + n) =
        make_float2(fmaxf(c.get(nb, 0) + b0, 0.f), fmaxf(c.get(nb, 1) + b1, 0.f));
    *reinterpret_cast<float2*>(out + (m0 + g + 8) * ldo + n) =
        make_float2(fmaxf(c.get(nb, 2) + b0, 0.f), fmaxf(c.get(nb, 3) + b1, 0.f));
  }
}

// wait until at most n of the most recently committed cp.async groups are still in flight
__device__ __forceinline__ void cp_async_wait_pending(int n) {
  if (n <= 0) asm volatile("cp.async.wait_group 0;" ::: "memory");
  else if (n == 1) asm volatile("cp.async.wait_group 1;" ::: "memory");
  else if (n == 2) asm volatile("cp.async.wait_group 2;" ::: "memory");
  else asm volatile("cp.async.wait_group 3;" ::: "memory");
}

// Layer 0 with its operands still arriving: rows and W0 are loaded as column chunks of kChunk floats, one cp.async group
// each; chunk c is multiplied while chunks c+1.. are in flight.  `younger` = groups committed after the chunk groups.
constexpr int kChunk = 128;
template <int N, int U, int NW>
__device__ __forceinline__ void fwd_layer0_mma(const float* a, int lda, int K, const float* w, int ldw, const float* b,
                                               float* out, int ldo, int younger) {
  constexpr int NB = N / (4 * NW);         // NW / 2 column groups of 8 NB columns
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int m0 = 16 * (warp & 1), n0 = 8 * NB * (warp >> 1);
  const int n_chunks = (K + kChunk - 1) / kChunk;
  Acc<NB, 3> c;
  c.zero();
  for (int ch = 0; ch < n_chunks; ++ch) {
    cp_async_wait_pending(n_chunks - 1 - ch + younger);
    __syncthreads();
    const int k0 = ch * kChunk, kc = min(kChunk, K - k0);
    warp_gemm<NB, 3, U>(c, a + m0 * lda + k0, lda, 1, w + n0 * ldw + k0, 1, ldw, kc);
  }
#pragma unroll
  for (int nb = 0; nb < NB; ++nb) {
    const int n = n0 + 8 * nb + 2 * t;
    const float b0 = b[n], b1 = b[n + 1];
    *reinterpret_cast<float2*>(out + (m0 + g) * ldo + n) =
        make_float2(fmaxf(c.get(nb, 0) + b0, 0.f), fmaxf(c.get(nb, 1) + b1, 0.f));
    *reinterpret_cast<float2*>(out + (m0 + g + 8) * ldo + n) =
        make_float2(fmaxf(c.get(nb, 2) + b0, 0.f), fmaxf(c.get(nb, 3) + b1, 0.f));
  }
}

// h (32 x K) <- (dz (32 x N) w) * [h > 0] in place, w (N x K): contraction over the layer's OUTPUT index n
template <int K, int N, int U, int NW>
__device__ __forceinline__ void dh_layer_mma(const float* dz, int ldz, const float* w, int ldw, float* h, int ldh) {
  constexpr int NB = K / (4 * NW);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int m0 = 16 * (warp & 1), n0 = 8 * NB * (warp >> 1);
  Acc<NB, (NB <= 2 ? 3 : 2)> c;
  c.zero();
  warp_gemm<NB, (NB <= 2 ? 3 : 2), U>(c, dz + m0 * ldz, ldz, 1, w + n0, ldw, 1, N);
#pragma unroll
  for (int nb = 0; nb < NB; ++nb) {
    const int n = n0 + 8 * nb + 2 * t;
    float2* p0 = reinterpret_cast<float2*>(h + (m0 + g) * ldh + n);
    float2* p1 = reinterpret_cast<float2*>(h + (m0 + g + 8) * ldh + n);
    const float2 h0 = *p0, h1 = *p1;
    *p0 = make_float2(h0.x > 0.f ? c.get(nb, 0) : 0.f, h0.y > 0.f ? c.get(nb, 1) : 0.f);
    *p1 = make_float2(h1.x > 0.f ? c.get(nb, 2) : 0.f, h1.y > 0.f ? c.get(nb, 3) : 0.f);
  }
}

// gw (N x K) (+)= dz^T (N x 32) a (32 x K).  Work units of 16 rows x 8 NB columns.
template <int N, int NB, int U, int NW>
__device__ __forceinline__ void dw_layer_mma(const float* dz, int ldz, const float* a, int lda, int K, float* gw,
                                             bool accumulate) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int col_units = K / (8 * NB), units = (N / 16) * col_units;
  for (int u = warp; u < units; u += NW) {
    const int m0 = 16 * (u / col_units), n0 = 8 * NB * (u % col_units);
    Acc<NB, (NB <= 4 ? 2 : 1)> c;
    c.zero();
    warp_gemm<NB, (NB <= 4 ? 2 : 1), (U > 2 ? 4 : 1)>(c, dz + m0, 1, ldz, a + n0, lda, 1, kT);
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) {
      const int n = n0 + 8 * nb + 2 * t;
      float2* p0 = reinterpret_cast<float2*>(gw + (int64_t)(m0 + g) * K + n);
      float2* p1 = reinterpret_cast<float2*>(gw + (int64_t)(m0 + g + 8) * K + n);
      float2 o0 = make_float2(c.get(nb, 0), c.get(nb, 1)), o1 = make_float2(c.get(nb, 2), c.get(nb, 3));
      if (accumulate) {
        const float2 q0 = *p0, q1 = *p1;
        o0.x += q0.x; o0.y += q0.y; o1.x += q1.x; o1.y += q1.y;
      }
      *p0 = o0;
      *p1 = o1;
    }
  }
}

template <bool MMA, int U, int NW>
__global__ void __launch_bounds__(NW * 32, 1) cql_fused_kernel(const CqlArgs p) {
  pdl_entry();
  extern __shared__ __align__(16) float smem[];
  const int KP = p.KP, ldx = KP + 4;
  const int ws_floats = max(kH1 * ldx, kW123Floats);
  float* XS = smem;
  float* WS = XS + kT * ldx;
  float* H1 = WS + ws_floats;
  float* H2 = H1 + kT * kLd1;
  float* H3 = H2 + kT * kLd2;
  float* BS = H3 + kT * kLd3;
  float* QS = BS + kBiasFloats;
  float* DQ = QS + kT * kAP;
  float* NMAX = DQ + kT * kAP;
  int* ROW = reinterpret_cast<int*>(NMAX + kT);
  float* W1s = WS;
  float* W2s = WS + kW1Floats;
  float* W3s = WS + kW1Floats + kW2Floats;
  float* B1 = BS, *B2 = BS + kH1, *B3 = BS + kH1 + kH2, *B4 = BS + kH1 + kH2 + kH3;

  float* part = p.partials + (int64_t)blockIdx.x * p.part_stride;
  const int64_t n_tiles = (p.B + kT - 1) / kT;
  float td_sum = 0.f, pen_sum = 0.f;     // thread 0 only
  bool acc = false;

  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, acc = true) {
    // ---- which rows
    if (threadIdx.x < kT) {
      const int64_t i = tile * kT + threadIdx.x;
      int r = -1;
      if (i < p.B) {
        int64_t rr = p.idx ? p.idx[i] : i;
        if (rr < 0 || rr >= p.n_rows) {
          atomicOr(p.status, (uint32_t)PORL_FLAG_BAD_INDEX);
          rr = rr < 0 ? 0 : p.n_rows - 1;
        }
        r = (int)rr;
      }
      ROW[threadIdx.x] = r;
    }
    __syncthreads();
    // ---- target network on next_states: max_a Q_tgt(s', a)     (cql_trainer.py:94-98)
    if (MMA) {
      for (int k0 = 0; k0 < KP; k0 += kChunk) {
        const int kc = min(kChunk, KP - k0);
        load_rows(XS, ldx, p.next_states, p.ld_states, ROW, kc, k0);
        load_mat_cols(WS, ldx, p.target + p.w_off[0], KP, kH1, k0, kc);
        cp_async_commit();
      }
      load_bias(BS, p.target, p.b_off, p.A);
      fwd_layer0_mma<kH1, U, NW>(XS, ldx, KP, WS, ldx, B1, H1, kLd1, 0);
    } else {
      load_rows(XS, ldx, p.next_states, p.ld_states, ROW, KP);
      load_mat(WS, ldx, p.target + p.w_off[0], kH1, KP);
      load_bias(BS, p.target, p.b_off, p.A);
      cp_async_wait_all();
      __syncthreads();
      fwd_layer<kH1>(XS, ldx, KP, WS, ldx, B1, H1, kLd1);
    }
    __syncthreads();
    load_w123(WS, p.target, p.w_off, p.A);
    cp_async_commit();
    load_rows(XS, ldx, p.states, p.ld_states, ROW, KP);      // obs rows arrive behind the target's upper layers
    cp_async_commit();
    cp_async_wait_but_last();
    __syncthreads();
    if (MMA) fwd_layer_mma<kH2, U, NW>(H1, kLd1, kH1, W1s, kLd1, B2, H2, kLd2); else fwd_layer<kH2>(H1, kLd1, kH1, W1s, kLd1, B2, H2, kLd2);
    __syncthreads();
    if (MMA) fwd_layer_mma<kH3, U, NW>(H2, kLd2, kH2, W2s, kLd2, B3, H3, kLd3); else fwd_layer<kH3>(H2, kLd2, kH2, W2s, kLd2, B3, H3, kLd3);
    __syncthreads();
    head_fwd(H3, W3s, B4, QS);
    __syncthreads();
    if (threadIdx.x < kT) {
      float m = -INFINITY;
      for (int a = 0; a < p.A; ++a) m = fmaxf(m, QS[threadIdx.x * kAP + a]);
      NMAX[threadIdx.x] = m;
    }
    // ---- online network on states                                (cql_trainer.py:100)
    if (MMA) {
      for (int k0 = 0; k0 < KP; k0 += kChunk) {
        load_mat_cols(WS, ldx, p.params + p.w_off[0], KP, kH1, k0, min(kChunk, KP - k0));
        cp_async_commit();
      }
      load_bias(BS, p.params, p.b_off, p.A);
      fwd_layer0_mma<kH1, U, NW>(XS, ldx, KP, WS, ldx, B1, H1, kLd1, 0);   // the obs rows' group is older than the chunks
    } else {
      load_mat(WS, ldx, p.params + p.w_off[0], kH1, KP);
      load_bias(BS, p.params, p.b_off, p.A);
      cp_async_wait_all();
      __syncthreads();
      fwd_layer<kH1>(XS, ldx, KP, WS, ldx, B1, H1, kLd1);
    }
    __syncthreads();
    load_w123(WS, p.params, p.w_off, p.A);
    cp_async_wait_all();
    __syncthreads();
    if (MMA) fwd_layer_mma<kH2, U, NW>(H1, kLd1, kH1, W1s, kLd1, B2, H2, kLd2); else fwd_layer<kH2>(H1, kLd1, kH1, W1s, kLd1, B2, H2, kLd2);
    __syncthreads();
    if (MMA) fwd_layer_mma<kH3, U, NW>(H2, kLd2, kH2, W2s, kLd2, B3, H3, kLd3); else fwd_layer<kH3>(H2, kLd2, kH2, W2s, kLd2, B3, H3, kLd3);
    __syncthreads();
    head_fwd(H3, W3s, B4, QS);
    __syncthreads();
    // ---- TD + conservative penalty and dL/dq                     (cql_trainer.py:72-86,101-108)
    if (threadIdx.x < kT) {
      const int s = threadIdx.x;
      const int r = ROW[s];
      float td_i = 0.f, pen_i = 0.f;
      float dq[kAP];
#pragma unroll
      for (int a = 0; a < kAP; ++a) dq[a] = 0.f;
      if (r >= 0) {
        const float* qr = QS + s * kAP;
        float qmax = -INFINITY;
        for (int a = 0; a < p.A; ++a) qmax = fmaxf(qmax, qr[a]);
        const float t = p.rewards[r] + (p.gamma * NMAX[s]) * (1.f - p.dones[r]);
        int64_t ai = p.actions[r];
        if (ai < 0 || ai >= p.A) {
          atomicOr(p.status, (uint32_t)PORL_FLAG_BAD_ACTION);
          ai = ai < 0 ? 0 : p.A - 1;
        }
        float se = 0.f;
        for (int a = 0; a < p.A; ++a) se += expf(qr[a] - qmax);
        const float lse = qmax + logf(se);
        const float qa = qr[ai];
        const float diff = qa - t;
        td_i = diff * diff;
        pen_i = lse - logf((float)p.A) - qa;
#pragma unroll
        for (int a = 0; a < kAP; ++a)
          if (a < p.A) {
            const float onehot = (a == (int)ai) ? 1.f : 0.f;
            dq[a] = p.inv_bn * (2.f * diff * onehot + p.alpha * (expf(qr[a] - lse) - onehot));
          }
      }
#pragma unroll
      for (int a = 0; a < kAP; ++a) DQ[s * kAP + a] = dq[a];
      td_i = warp_sum(td_i);
      pen_i = warp_sum(pen_i);
      if (s == 0) {
        td_sum += td_i;
        pen_sum += pen_i;
      }
    }
    __syncthreads();
    // ---- backward
    {  // dW3 (A x 64), db3
      float* gw = part + p.w_off[3];
      for (int o = threadIdx.x; o < p.A * kH3; o += (int)blockDim.x) {
        const int a = o / kH3, k = o - a * kH3;
        float sacc = 0.f;
        for (int s = 0; s < kT; ++s) sacc = fmaf(DQ[s * kAP + a], H3[s * kLd3 + k], sacc);
        gw[o] = acc ? gw[o] + sacc : sacc;
      }
      if (threadIdx.x < kAP) {   // the bias block is pad4(A) <= 8 floats; pad entries stay zero
        const int a = threadIdx.x;
        if (a < ((p.A + 3) & ~3)) {
          float sb = 0.f;
          for (int s = 0; s < kT; ++s) sb += DQ[s * kAP + a];
          float* gb = part + p.b_off[3];
          gb[a] = acc ? gb[a] + sb : sb;
        }
      }
    }
    __syncthreads();
    if (MMA) dh_layer_mma<kH3, kAP, U, NW>(DQ, kAP, W3s, kLd3, H3, kLd3); else dh_layer<kH3, kAP>(DQ, kAP, W3s, kLd3, H3, kLd3);                       // H3 <- dz3
    __syncthreads();
    if (MMA) dw_layer_mma<kH3, (NW > 8 ? 4 : 8), U, NW>(H3, kLd3, H2, kLd2, kH2, part + p.w_off[2], acc);
    else dw_layer<kH3, 2>(H3, kLd3, H2, kLd2, kH2, part + p.w_off[2], acc);
    __syncthreads();
    if (MMA) dh_layer_mma<kH2, kH3, U, NW>(H3, kLd3, W2s, kLd2, H2, kLd2); else dh_layer<kH2, kH3>(H3, kLd3, W2s, kLd2, H2, kLd2);                      // H2 <- dz2
    __syncthreads();
    if (MMA) dw_layer_mma<kH2, (NW > 8 ? 4 : 8), U, NW>(H2, kLd2, H1, kLd1, kH1, part + p.w_off[1], acc);
    else dw_layer<kH2, 2>(H2, kLd2, H1, kLd1, kH1, part + p.w_off[1], acc);
    __syncthreads();
    if (MMA) dh_layer_mma<kH1, kH2, U, NW>(H2, kLd2, W1s, kLd1, H1, kLd1); else dh_layer<kH1, kH2>(H2, kLd2, W1s, kLd1, H1, kLd1);                      // H1 <- dz1
    __syncthreads();
    if (MMA) dw_layer_mma<kH1, 4, U, NW>(H1, kLd1, XS, ldx, KP, part + p.w_off[0], acc);
    else dw_layer<kH1, 6>(H1, kLd1, XS, ldx, KP, part + p.w_off[0], acc);
    {  // bias gradients of the hidden layers: column sums of dz1 | dz2 | dz3, which the backward pass left in
       // H1 | H2 | H3 -- 64 + 128 + 64 columns, one per thread
      static_assert(kH1 + kH2 + kH3 == kThreads, "one thread per hidden unit");
      const int tid = threadIdx.x < kThreads ? threadIdx.x : 0;      // warps past the first 8 repeat column 0 and skip the store
      const float* src;
      int ld;
      float* gb;
      if (tid < kH1) { src = H1 + tid; ld = kLd1; gb = part + p.b_off[0] + tid; }
      else if (tid < kH1 + kH2) { src = H2 + tid - kH1; ld = kLd2; gb = part + p.b_off[1] + tid - kH1; }
      else { src = H3 + tid - kH1 - kH2; ld = kLd3; gb = part + p.b_off[2] + tid - kH1 - kH2; }
      float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
      for (int r = 0; r < kT; r += 4) {
        s0 += src[r * ld];
        s1 += src[(r + 1) * ld];
        s2 += src[(r + 2) * ld];
        s3 += src[(r + 3) * ld];
      }
      const float sb = (s0 + s1) + (s2 + s3);
      if (threadIdx.x < kThreads) *gb = acc ? *gb + sb : sb;
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    part[p.n_floats] = td_sum;
    part[p.n_floats + 1] = pen_sum;
  }
}

// grads[j] = sum over the CTAs' partials, in a fixed order (32 interleaved groups of CTAs summed in CTA order, then the
// groups in group order): deterministic.  256 threads = 8 float4 columns x 32 groups, so every thread has at most
// ceil(n_part / 32) independent 16-byte loads in flight (the partials are L2-resident: the fused kernel has just
// written them); block 0 also finalises the losses.
constexpr int kRedCols = 8, kRedGroups = 32;

// optimizer folded into the reduction (porl_cql_fused_step): the thread that finishes a float4 of the gradient also runs
// Adam on those four parameters (+ the hard target copy), the last block of the launch advances the optimizer clock --
// the whole CQL update is then two launches (fused forward/backward, reduce + Adam)
constexpr int kTicketGroup = 32;
struct CqlOpt {
  uint32_t* tickets;      // group ticket words, 32 words apart (ctx->counters + kTicketBase)
  float* p;
  float* m;
  float* v;
  float* target;          // hard-copy destination or null
  porl_opt_state_t* st;
  float beta1, beta2, eps;
  int hard_every;
};

template <bool ADAM>
__global__ void __launch_bounds__(kRedCols * kRedGroups)
cql_reduce_kernel(const float* __restrict__ partials, int n_part, int64_t part_stride, int64_t n_floats,
                  float* __restrict__ grads, float* __restrict__ losses_out, float inv_bn, float alpha,
                  uint32_t* status, const CqlOpt o) {
  // ADAM: launched with one block more than the columns need; that block has no column (j >= n_floats) and publishes
  // the scalars of the next update on the side (optim.cuh: published scalars)
  pdl_entry();
  const bool scalar_thread = ADAM && threadIdx.x == kRedCols * kRedGroups - 1;
  __shared__ float4 red[kRedGroups][kRedCols];
  __shared__ AdamScalars sh;
  const int col = threadIdx.x % kRedCols, grp = threadIdx.x / kRedCols;
  const int64_t j = ((int64_t)blockIdx.x * kRedCols + col) * 4;
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
  float4 pp = s, mm = s, vv = s;
  __shared__ long long sh_t;
  __shared__ double sh_lr;
  if (ADAM && grp == 0 && j < n_floats) {      // in flight during the partial sums
    pp = *reinterpret_cast<const float4*>(o.p + j);
    mm = *reinterpret_cast<const float4*>(o.m + j);
    vv = *reinterpret_cast<const float4*>(o.v + j);
  }
  AdamClock clk = {};
  if (scalar_thread) clk = adam_clock_load(o.st);      // in flight with the partial loads: one round trip
  if (j < n_floats) {
#pragma unroll 4
    for (int c = grp; c < n_part; c += kRedGroups) {
      const float4 v = __ldcg(reinterpret_cast<const float4*>(partials + c * part_stride + j));
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
  }
  red[grp][col] = s;
  if (scalar_thread) {
    sh = adam_scalars_from(clk, o.st, o.beta1, o.beta2, o.hard_every);
    sh_t = clk.t;
    sh_lr = clk.lr;
  }
  __syncthreads();
  if (ADAM && blockIdx.x == gridDim.x - 1 && threadIdx.x == 32) adam_publish_next(o.st, sh_t, sh_lr, o.beta1, o.beta2);
  if (grp == 0 && j < n_floats) {
#pragma unroll
    for (int q = 1; q < kRedGroups; ++q) {
      const float4 v = red[q][col];
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
    *reinterpret_cast<float4*>(grads + j) = s;
    if (ADAM) {
      const AdamScalars a = sh;
      adam_elem(pp.x, s.x, mm.x, vv.x, a, o.beta1, o.beta2, o.eps);
      adam_elem(pp.y, s.y, mm.y, vv.y, a, o.beta1, o.beta2, o.eps);
      adam_elem(pp.z, s.z, mm.z, vv.z, a, o.beta1, o.beta2, o.eps);
      adam_elem(pp.w, s.w, mm.w, vv.w, a, o.beta1, o.beta2, o.eps);
      *reinterpret_cast<float4*>(o.p + j) = pp;
      *reinterpret_cast<float4*>(o.m + j) = mm;
      *reinterpret_cast<float4*>(o.v + j) = vv;
      if (o.target != nullptr && a.do_hard) *reinterpret_cast<float4*>(o.target + j) = pp;
    }
  }
  if (scalar_thread) {
    // The optimizer clock.  This block has READ the step count and the learning rate (the only things the clock update
    // changes): it takes its ticket here, off the path of the warp that does the Adam arithmetic; the block that draws
    // the last ticket knows every block has read them and advances the clock.  Two levels: 1,300 atomics on ONE
    // address serialise in the L2, so the blocks draw tickets in groups of kTicketGroup on words 128 bytes apart and
    // only the last block of a group goes on to the optimizer state's counter.  All counters reset themselves.
    __threadfence();               // this thread's reads of the clock are complete before its ticket is visible
    const uint32_t grp_id = blockIdx.x / kTicketGroup;
    const uint32_t grp_size = min((uint32_t)kTicketGroup, gridDim.x - grp_id * kTicketGroup);
    uint32_t* gt = o.tickets + grp_id * 32u;
    if (atomicAdd(gt, 1u) == grp_size - 1) {
      *gt = 0;
      const uint32_t n_grp = (gridDim.x + kTicketGroup - 1) / kTicketGroup;
      if (atomicAdd(&o.st->blocks_done, 1u) == n_grp - 1) {
        o.st->blocks_done = 0;
        adam_advance_clock(o.st);
      }
    }
  }
  if (blockIdx.x == 0 && threadIdx.x < 32) {
    float st = 0.f, sp = 0.f;
    for (int c = threadIdx.x; c < n_part; c += 32) {
      st += __ldcg(partials + c * part_stride + n_floats);
      sp += __ldcg(partials + c * part_stride + n_floats + 1);
    }
    st = warp_sum(st);
    sp = warp_sum(sp);
    if (threadIdx.x == 0) {
      const float td = st * inv_bn, pen = sp * inv_bn;
      const float loss = td + alpha * pen;
      losses_out[0] = loss;
      losses_out[1] = td;
      losses_out[2] = pen;
      if (!isfinite(loss)) atomicOr(status, (uint32_t)PORL_FLAG_NONFINITE_LOSS);
    }
  }
}

size_t cql_smem_bytes(int KP) {
  const int ldx = KP + 4;
  const int ws = kH1 * ldx > kW123Floats ? kH1 * ldx : kW123Floats;
  const size_t floats = (size_t)kT * ldx + ws + kT * (kLd1 + kLd2 + kLd3) + kBiasFloats + 2 * kT * kAP + kT + kT;
  return floats * sizeof(float);
}

}  // namespace
}  // namespace porl

using namespace porl;

extern "C" int porl_cql_fused_supported(const porl_mlp_t* mlp) {
  porl_mlp_desc_t d;
  if (!mlp || porl_mlp_desc(mlp, &d) != PORL_OK) return 0;
  if (d.n_nets != 1 || d.n_layers != 4 || d.layer_norm || d.out_act) return 0;
  if (d.dims[1] != kH1 || d.dims[2] != kH2 || d.dims[3] != kH3) return 0;
  if (d.dims[4] < 1 || d.dims[4] > kAP || d.dims[0] < 1) return 0;
  const int KP = (int)round_up(d.dims[0], PORL_PAD);
  return cql_smem_bytes(KP) <= 227 * 1024 ? 1 : 0;
}

static int cql_fused_launch(porl_ctx_t* ctx, const porl_mlp_t* mlp, const float* params,
                            const float* target_params, const float* states, const float* next_states,
                            int64_t ld_states, const int64_t* actions, const float* rewards,
                            const float* dones, const int64_t* idx, int64_t n_rows, float gamma, float alpha,
                            int64_t B, int64_t B_norm, float* grads, float* losses_out, const CqlOpt* opt,
                            porl_stream_t stream) {
  PORL_ON_DEVICE(ctx);
  PORL_CHECK_ARG(ctx && mlp && params && target_params && states && next_states && actions && rewards && dones &&
                 grads && losses_out, "porl_cql_fused_grad: null argument");
  if (!porl_cql_fused_supported(mlp)) {
    set_error("porl_cql_fused_grad: the fused kernel serves one ReLU net S -> 64 -> 128 -> 64 -> A (A <= 8, "
              "pad32(S) <= 384) without LayerNorm");
    return PORL_ERR_UNSUPPORTED;
  }
  porl_mlp_desc_t d;
  porl_mlp_desc(mlp, &d);
  const int KP = (int)round_up(d.dims[0], PORL_PAD);
  PORL_CHECK_ARG(B > 0 && B_norm >= B && n_rows > 0 && n_rows < ((int64_t)1 << 31) && (idx || n_rows >= B),
                 "porl_cql_fused_grad: B=%lld B_norm=%lld n_rows=%lld", (long long)B, (long long)B_norm,
                 (long long)n_rows);
  PORL_CHECK_ARG(ld_states >= KP && (ld_states & 3) == 0 && aligned16(states) && aligned16(next_states) &&
                 aligned16(params) && aligned16(target_params) && aligned16(grads),
                 "porl_cql_fused_grad: state rows must be padded to %d floats (pad columns zero) and 16-byte "
                 "aligned (ld=%lld)", KP, (long long)ld_states);
  CqlArgs a;
  a.params = params;
  a.target = target_params;
  for (int l = 0; l < 4; ++l) {
    int64_t ld = 0;
    int rc = porl_mlp_param_layout(mlp, 0, l, PORL_PARAM_WEIGHT, &a.w_off[l], &ld);
    if (rc == PORL_OK) rc = porl_mlp_param_layout(mlp, 0, l, PORL_PARAM_BIAS, &a.b_off[l], &ld);
    if (rc != PORL_OK) return rc;
  }
  a.states = states;
  a.next_states = next_states;
  a.ld_states = ld_states;
  a.actions = actions;
  a.rewards = rewards;
  a.dones = dones;
  a.idx = idx;
  a.n_rows = n_rows;
  a.B = B;
  a.KP = KP;
  a.A = d.dims[4];
  a.gamma = gamma;
  a.alpha = alpha;
  a.inv_bn = 1.0f / (float)B_norm;
  a.n_floats = porl_mlp_param_floats(mlp);
  a.part_stride = round_up(a.n_floats + 2, 4);
  a.partials = ctx->tc_ws;
  a.status = ctx->counters;
  const int64_t n_tiles = (B + kT - 1) / kT;
  const int grid = (int)(n_tiles < ctx->sm_count ? n_tiles : ctx->sm_count);
  if ((int64_t)grid * a.part_stride > (int64_t)kTcWsFloats) {
    set_error("porl_cql_fused_grad: %d gradient partials of %lld floats exceed the workspace", grid,
              (long long)a.part_stride);
    return PORL_ERR_WORKSPACE;
  }
  const size_t smem = cql_smem_bytes(KP);
  static bool configured = false;
  if (!configured) {
    PORL_CUDA(cudaFuncSetAttribute(cql_fused_kernel<false, 4, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    PORL_CUDA(cudaFuncSetAttribute(cql_fused_kernel<true, 4, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    PORL_CUDA(cudaFuncSetAttribute(cql_fused_kernel<true, 2, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    configured = true;
  }
  cudaStream_t st = as_stream(stream);
  {
    // algorithmic flop of SURVEY 8d: 2 forwards + dW of every layer + dX of layers 2..4
    double mac = 0.0, dx = 0.0;
    for (int l = 0; l < 4; ++l) {
      mac += (double)d.dims[l] * d.dims[l + 1];
      if (l > 0) dx += (double)d.dims[l] * d.dims[l + 1];
    }
    ProfScope prof(ctx, PORL_CAT_CONTRACTION, 2.0 * (double)B * (3.0 * mac + dx), st);
    // contraction mode of the context: 0 = fp32 FFMA, 1 = tensor cores (mma.sync 3xTF32 here; PORL_CQL_MMA overrides)
    const int env_mma = getenv("PORL_CQL_MMA") ? atoi(getenv("PORL_CQL_MMA")) : -1;
    const bool use_mma = env_mma >= 0 ? env_mma != 0 : ctx->gemm_mode == 1;
    // PORL_CQL_WARPS: 8 (256 threads, 16 x 16 warp tiles, default) or 16 (512 threads, 16 x 8 tiles, <= 128 registers).
    // Measured equal within 3 % (53 vs 55 us at batch 4096): the contraction phases are bound by the legacy HMMA rate --
    // 48 HMMA.1688.TF32 per SM take ~205 cycles with either warp count, i.e. ~17 cycles per instruction and
    // sub-partition -- not by latency, so more resident warps do not help.
    const int env_warps = getenv("PORL_CQL_WARPS") ? atoi(getenv("PORL_CQL_WARPS")) : 8;
    if (use_mma && env_warps >= 16) PORL_CUDA(launch_pdl(ctx, cql_fused_kernel<true, 2, 16>, grid, 512, smem, st, a));
    else if (use_mma) PORL_CUDA(launch_pdl(ctx, cql_fused_kernel<true, 4, 8>, grid, kThreads, smem, st, a));
    else PORL_CUDA(launch_pdl(ctx, cql_fused_kernel<false, 4, 8>, grid, kThreads, smem, st, a));
    PORL_LAUNCHED(ctx);
  }
  {
    ProfScope prof(ctx, PORL_CAT_LOSS, (double)(grid + 1) * a.n_floats * 4.0, st);
    const int blocks = (int)((a.n_floats / 4 + kRedCols - 1) / kRedCols);
    PORL_CHECK_ARG(!opt || (blocks + kTicketGroup) / kTicketGroup * 32 <= kTickets, "porl_cql_fused_step: network too large");
    if (opt)
      PORL_CUDA(launch_pdl(ctx, cql_reduce_kernel<true>, blocks + 1 /* publisher block */, kRedCols * kRedGroups, 0, st, (const float*)a.partials, grid,
                           a.part_stride, a.n_floats, grads, losses_out, a.inv_bn, alpha, ctx->counters, *opt));
    else
      PORL_CUDA(launch_pdl(ctx, cql_reduce_kernel<false>, blocks, kRedCols * kRedGroups, 0, st, (const float*)a.partials, grid,
                           a.part_stride, a.n_floats, grads, losses_out, a.inv_bn, alpha, ctx->counters, CqlOpt{}));
    PORL_LAUNCHED(ctx);
  }
  return PORL_OK;
}

extern "C" int porl_cql_fused_grad(porl_ctx_t* ctx, const porl_mlp_t* mlp, const float* params,
                                   const float* target_params, const float* states, const float* next_states,
                                   int64_t ld_states, const int64_t* actions, const float* rewards,
                                   const float* dones, const int64_t* idx, int64_t n_rows, float gamma, float alpha,
                                   int64_t B, int64_t B_norm, float* grads, float* losses_out,
                                   porl_stream_t stream) {
  return cql_fused_launch(ctx, mlp, params, target_params, states, next_states, ld_states, actions, rewards, dones, idx,
                          n_rows, gamma, alpha, B, B_norm, grads, losses_out, nullptr, stream);
}

extern "C" int porl_cql_fused_step(porl_ctx_t* ctx, const porl_mlp_t* mlp, float* params, float* target_params,
                                   const float* states, const float* next_states, int64_t ld_states,
                                   const int64_t* actions, const float* rewards, const float* dones, const int64_t* idx,
                                   int64_t n_rows, float gamma, float alpha, int64_t B, float* grads, float* losses_out,
                                   float* m, float* v, porl_opt_state_t* state_dev, float beta1, float beta2, float eps,
                                   int32_t hard_every, porl_stream_t stream) {
  PORL_CHECK_ARG(m && v && state_dev && params, "porl_cql_fused_step: null optimizer argument");
  PORL_CHECK_ARG(aligned16(m) && aligned16(v), "porl_cql_fused_step: moment buffers must be 16-byte aligned");
  PORL_CHECK_ARG(hard_every >= 0, "porl_cql_fused_step: hard_every=%d", (int)hard_every);
  CqlOpt o;
  o.tickets = ctx->counters + kTicketBase;
  o.p = params;
  o.m = m;
  o.v = v;
  o.target = hard_every > 0 ? target_params : nullptr;
  o.st = state_dev;
  o.beta1 = beta1;
  o.beta2 = beta2;
  o.eps = eps;
  o.hard_every = hard_every;
  return cql_fused_launch(ctx, mlp, params, target_params, states, next_states, ld_states, actions, rewards, dones, idx,
                          n_rows, gamma, alpha, B, B, grads, losses_out, &o, stream);
}
