// K8/K9: the frozen scan encoder of the POR/SORL agents -- util/costmap.py:7-64 (state2costmap) followed by
// agent/fasternet.py:428-438 (FasterNet.forward_cls) in train mode (batch-statistic BatchNorm, DropPath).
//
// Layout: every activation is an NHWC matrix  X[(b, h, w), C]  (row-major, C contiguous), so a 1x1 convolution
// is the K-major contraction  Z = X W^T  the tcgen05 kernel already provides (gemm_tc.cu), BatchNorm is a
// per-column affine map, and the 3x3 partial convolution / 2x2 patch merging read neighbouring rows.
// The 1.1 MB/sample costmap is never materialised on the training path: it holds ~365 ones, so the 4x4/s4
// patch embedding is evaluated sparsely from the scan itself (porl_costmap writes the dense raster for the
// state2costmap API and the tests).
//
// No backward: no optimizer owns the encoder's parameters in the reference (agent/sorl.py:60,63,
// agent/por.py:61,64), so gradients that reach it are never applied.
#include "encoder.cuh"
#include "bn.cuh"

#include <math.h>

namespace porl {

namespace {

constexpr int kScan = 360, kBins = 256, kState = 362;
constexpr int kPH = 90, kPW = 64;                 // patch-embedding output grid (360/4, 256/4)

// Constants of util/costmap.py:19-20,34,44 in the precision torch uses them: python doubles cast to fp32
// when they meet an fp32 tensor.
__device__ __forceinline__ float dist_inc() { return (float)((4.0 + 1e-4) / 256.0); }
__device__ __forceinline__ float angle_inc() { return (float)((2.0 * 3.141592653589793 + 1e-4) / 360.0); }

__device__ __forceinline__ float clamp8(float v) { return v > 8.0f ? 0.0f : v; }   // costmap.py:17

// Range bin of a scan beam (costmap.py:25): trunc(v / dist_inc); -1 when the reference's scatter_ would raise.
__device__ __forceinline__ int scan_bin(float v) {
  const float q = __fdiv_rn(clamp8(v), dist_inc());
  if (!(q > -1.0f) || !(q < 256.0f)) return -1;
  return (int)q;                                   // q in (-1, 256): trunc toward zero gives 0..255
}

// Goal cross centre (costmap.py:33-46): row deg in [1, 358], column dist in [0, 254].
__device__ __forceinline__ void goal_cell(float gx, float gy, int& deg, int& dist) {
  gx = clamp8(gx);
  gy = clamp8(gy);
  const float pi = 3.14159265358979323846f;
  const float lim = (float)((2.0 * 3.141592653589793 + 2e-4) / 360.0);
  const float lo = (float)(-3.141592653589793 + (2.0 * 3.141592653589793 + 2e-4) / 360.0);
  const float hi = (float)(3.141592653589793 - (2.0 * 3.141592653589793 + 2e-4) / 360.0);
  (void)lim;
  float a = atan2f(gy, gx);
  a = fminf(fmaxf(a, lo), hi);
  deg = (int)__fdiv_rn(__fadd_rn(a, pi), angle_inc());
  float d = sqrtf(__fadd_rn(__fmul_rn(gx, gx), __fmul_rn(gy, gy)));
  d = fminf(d, (float)(4.0 - 4.0 / 256.0));
  dist = (int)__fdiv_rn(d, dist_inc());
}

// ------------------------------------------------------------------------------------ dense costmap
// grid (360 / 8, B), 256 threads: block = 8 raster rows of one sample (8 x 256 x 3 floats).
__global__ void __launch_bounds__(256) costmap_kernel(float* __restrict__ state, int64_t ld_state, float* __restrict__ out,
                                                      uint32_t* __restrict__ status) {
  const int b = blockIdx.y, r0 = blockIdx.x * 8;
  float* s = state + (size_t)b * ld_state;
  float* o = out + ((size_t)b * kScan + r0) * (kBins * 3);
  float4* o4 = reinterpret_cast<float4*>(o);
  for (int i = threadIdx.x; i < 8 * kBins * 3 / 4; i += 256) o4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  __syncthreads();
  if (threadIdx.x < 8) {
    const int r = r0 + threadIdx.x;
    const int beam = (r + 180) % kScan;            // torch.roll(idx, 180, 1): row r shows beam r - 180
    const int bin = scan_bin(s[beam]);
    if (bin < 0) atomicOr(status, (uint32_t)PORL_FLAG_SCAN_RANGE);
    else if (bin > 0) o[(threadIdx.x * kBins + bin) * 3] = 1.0f;       // bin 0 is cleared (costmap.py:30)
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int deg, dist;
    goal_cell(s[kState - 2], s[kState - 1], deg, dist);
    const int rr[5] = {deg, deg, deg, deg - 1, deg + 1};
    const int cc[5] = {(dist + 255) % kBins, dist, (dist + 1) % kBins, dist, dist};
    for (int i = 0; i < 5; ++i) {
      const int r = (rr[i] + kScan) % kScan;
      if (r >= r0 && r < r0 + 8) {
        float* p = o + ((r - r0) * kBins + cc[i]) * 3;
        p[0] = p[1] = p[2] = 1.0f;
      }
    }
  }
  // the in-place clamp of the caller's state (costmap.py:17); every reader above clamps on read
  if (blockIdx.x == 0)
    for (int i = threadIdx.x; i < kState; i += 256)
      if (s[i] > 8.0f) s[i] = 0.0f;
}

// ------------------------------------------------------------------------------------ sparse patch embedding
// Y[(b, ph, pw), o] = sum_{c,kh,kw} W[o, c, kh, kw] * costmap[b, c, 4 ph + kh, 4 pw + kw]   (Conv2d k4 s4, no bias)
// grid (90, B), 192 threads: one block = one patch row (64 patches x 96 channels).  The 4 x 256 x 3 raster
// window of the patch row is rebuilt as a byte map in shared memory; only patches that hold a one are summed,
// in the fixed order (c, kh, kw).  wk: (48, 96) = W transposed to [(c, kh, kw), o].
__global__ void __launch_bounds__(192) patch_embed_kernel(float* __restrict__ state, int64_t ld_state,
                                                          const float* __restrict__ wk, float* __restrict__ out,
                                                          uint32_t* __restrict__ status) {
  __shared__ unsigned char pix[3][4][kBins];
  __shared__ unsigned char hit[kPW];
  const int ph = blockIdx.x, b = blockIdx.y, t = threadIdx.x;
  float* s = state + (size_t)b * ld_state;
  for (int i = t; i < 3 * 4 * kBins / 4; i += 192) reinterpret_cast<uint32_t*>(&pix[0][0][0])[i] = 0u;
  if (t < kPW) hit[t] = 0;
  __syncthreads();
  if (t < 4) {
    const int r = 4 * ph + t;
    const int bin = scan_bin(s[(r + 180) % kScan]);
    if (bin < 0) atomicOr(status, (uint32_t)PORL_FLAG_SCAN_RANGE);
    else if (bin > 0) { pix[0][t][bin] = 1; hit[bin >> 2] = 1; }
  }
  __syncthreads();
  if (t == 0) {
    int deg, dist;
    goal_cell(s[kState - 2], s[kState - 1], deg, dist);
    const int rr[5] = {deg, deg, deg, deg - 1, deg + 1};
    const int cc[5] = {(dist + 255) % kBins, dist, (dist + 1) % kBins, dist, dist};
    for (int i = 0; i < 5; ++i) {
      const int r = (rr[i] + kScan) % kScan;
      if ((r >> 2) == ph) {
        pix[0][r & 3][cc[i]] = pix[1][r & 3][cc[i]] = pix[2][r & 3][cc[i]] = 1;
        hit[cc[i] >> 2] = 1;
      }
    }
  }
  __syncthreads();
  const int o = t % 96;
  float* orow = out + ((size_t)b * kPH + ph) * (size_t)(kPW * 96);
  for (int pw = t / 96; pw < kPW; pw += 2) {
    float acc = 0.f;
    if (hit[pw]) {
      for (int c = 0; c < 3; ++c)
        for (int kh = 0; kh < 4; ++kh)
          for (int kw = 0; kw < 4; ++kw)
            if (pix[c][kh][4 * pw + kw]) acc += wk[(c * 16 + kh * 4 + kw) * 96 + o];
    }
    orow[pw * 96 + o] = acc;
  }
  if (ph == 0)
    for (int i = t; i < kState; i += 192)
      if (s[i] > 8.0f) s[i] = 0.0f;
}

// x[(b, p), :] += keep[b] * t[(b, p), :]    (residual + DropPath, fasternet.py:183-187); keep may be null (all 1)
__global__ void __launch_bounds__(256) residual_kernel(float* __restrict__ x, const float* __restrict__ t, int64_t n4,
                                                       int64_t per_sample4, const float* __restrict__ keep) {
  float4* x4 = reinterpret_cast<float4*>(x);
  const float4* t4 = reinterpret_cast<const float4*>(t);
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n4; i += (int64_t)gridDim.x * 256) {
    const float k = keep ? __ldg(keep + i / per_sample4) : 1.0f;
    float4 v = x4[i];
    const float4 u = t4[i];
    // the reference multiplies the branch by the mask first, then adds the shortcut (two roundings)
    v.x = __fadd_rn(v.x, __fmul_rn(u.x, k)); v.y = __fadd_rn(v.y, __fmul_rn(u.y, k));
    v.z = __fadd_rn(v.z, __fmul_rn(u.z, k)); v.w = __fadd_rn(v.w, __fmul_rn(u.w, k));
    x4[i] = v;
  }
}

// ------------------------------------------------------------------------------------ partial 3x3 convolution
// Partial_conv3.forward_split_cat (fasternet.py:131-138) on an NHWC matrix: channels [0, c) of the output are
// the zero-padded 3x3 convolution of channels [0, c) of the input (this kernel), channels [c, C) are copied
// (copy_channels_kernel).  FFMA-bound (2 * 9 * c^2 flop per pixel against 8 C bytes; a tensor-core form would waste 3/4 of an N = 128 MMA on
// N = c <= 48 and costs the same).  grid (tiles_w, tiles_h, B), 128 threads = 16 x 8; one block = a 16-wide,
// 8 * PIX-tall pixel tile, one thread = PIX pixels (rows ty + 8 j) with all c output channels in registers: per (tap,
// input channel) PIX activation loads and c / 4 broadcast weight loads feed PIX * c FMAs (10 loads per 96 FMAs at c = 24,
// PIX = 4; the round-1 kernel had 16).  Input channels go through shared memory CI at a time (halo tile, pixel stride
// CI + 1 floats: conflict-free) together with the matching weight slab.  wk: (9, c, c) = [tap = kh * 3 + kw][ci][co].
template <int CC, int PIX, int CI>
struct PconvCfg {
  static constexpr int TH = 8 * PIX, XS = CI + 1, HALO = (TH + 2) * 18;
  static constexpr size_t SMEM = ((size_t)HALO * XS + 9 * CI * CC) * sizeof(float);
};

template <int CC, int PIX, int CI>
__global__ void __launch_bounds__(128) pconv3_kernel(const float* __restrict__ x, float* __restrict__ y,
                                                     const float* __restrict__ wk, int H, int W, int C) {
  using Cfg = PconvCfg<CC, PIX, CI>;
  constexpr int TH = Cfg::TH, XS = Cfg::XS, HALO = Cfg::HALO;
  static_assert(CC % CI == 0 && CI % 4 == 0 && CC % 4 == 0, "channel blocking");
  extern __shared__ __align__(16) float psm[];
  float* const ws = psm;                          // [9][CI][CC], 16-byte aligned rows
  float* const xs = psm + 9 * CI * CC;            // [HALO][XS]
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;       // ty 0..7
  const int w0 = blockIdx.x * 16, h0 = blockIdx.y * TH, b = blockIdx.z;
  const float* xb = x + (size_t)b * H * W * C;
  float* yb = y + (size_t)b * H * W * C;
  float acc[PIX][CC];
#pragma unroll
  for (int j = 0; j < PIX; ++j)
#pragma unroll
    for (int c = 0; c < CC; ++c) acc[j][c] = 0.f;

  for (int ci0 = 0; ci0 < CC; ci0 += CI) {
    __syncthreads();
    // halo pixels x float4 groups of the channel slab, kLd global loads in flight per thread (one at a time left the
    // block waiting ~30 round trips to L2 / HBM per stage: `long scoreboard` was 3 of every 4 stall cycles, ncu)
    constexpr int kLd = 8, kHaloVec = HALO * (CI / 4);
    for (int i0 = threadIdx.x; i0 < kHaloVec; i0 += 128 * kLd) {
      float4 v[kLd];
#pragma unroll
      for (int u = 0; u < kLd; ++u) {
        const int i = i0 + 128 * u;
        const int pix = i / (CI / 4), q4 = i % (CI / 4);
        const int hh = h0 - 1 + pix / 18, ww = w0 - 1 + pix % 18;
        v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (i < kHaloVec && hh >= 0 && hh < H && ww >= 0 && ww < W)
          v[u] = *reinterpret_cast<const float4*>(xb + ((size_t)hh * W + ww) * C + ci0 + 4 * q4);
      }
#pragma unroll
      for (int u = 0; u < kLd; ++u) {
        const int i = i0 + 128 * u;
        if (i < kHaloVec) {
          float* d = xs + (i / (CI / 4)) * XS + 4 * (i % (CI / 4));
          d[0] = v[u].x; d[1] = v[u].y; d[2] = v[u].z; d[3] = v[u].w;
        }
      }
    }
#pragma unroll 4
    for (int i = threadIdx.x; i < 9 * CI * CC / 4; i += 128) {
      const int tap = i / (CI * CC / 4), rem = i % (CI * CC / 4);
      reinterpret_cast<float4*>(ws)[i] = __ldg(reinterpret_cast<const float4*>(wk + ((size_t)tap * CC + ci0) * CC + 4 * rem));
    }
    __syncthreads();
#pragma unroll 1
    for (int tap = 0; tap < 9; ++tap) {
      const int dy = tap / 3, dx = tap % 3;
      const float* p0 = xs + ((ty + dy) * 18 + tx + dx) * XS;
      const float* wt = ws + tap * CI * CC;
#pragma unroll 4
      for (int ci = 0; ci < CI; ++ci) {
        float a[PIX];
#pragma unroll
        for (int j = 0; j < PIX; ++j) a[j] = p0[j * 8 * 18 * XS + ci];
#pragma unroll
        for (int co = 0; co < CC; co += 4) {
          const float4 wv = *reinterpret_cast<const float4*>(wt + ci * CC + co);
#pragma unroll
          for (int j = 0; j < PIX; ++j) {
            acc[j][co] = fmaf(a[j], wv.x, acc[j][co]); acc[j][co + 1] = fmaf(a[j], wv.y, acc[j][co + 1]);
            acc[j][co + 2] = fmaf(a[j], wv.z, acc[j][co + 2]); acc[j][co + 3] = fmaf(a[j], wv.w, acc[j][co + 3]);
          }
        }
      }
    }
  }
  const int w = w0 + tx;
  if (w < W) {
#pragma unroll
    for (int j = 0; j < PIX; ++j) {
      const int hh = h0 + ty + 8 * j;
      if (hh < H) {
        float4* o = reinterpret_cast<float4*>(yb + ((size_t)hh * W + w) * C);
#pragma unroll
        for (int co = 0; co < CC; co += 4) o[co >> 2] = make_float4(acc[j][co], acc[j][co + 1], acc[j][co + 2], acc[j][co + 3]);
      }
    }
  }
}

// channels [c0, C) of every row of x copied to y (the untouched part of Partial_conv3's output): a grid-stride float4
// copy with eight loads in flight per thread.  Its own launch on purpose: inside the convolution kernel (2-3 blocks of
// 128 threads per SM, one tile at a time) the copy ran at a third of the HBM rate (`long scoreboard`, ncu).
__global__ void __launch_bounds__(256) copy_channels_kernel(const float* __restrict__ x, float* __restrict__ y, int64_t rows,
                                                            int C, int c0) {
  const int rest4 = (C - c0) >> 2;
  const int64_t n = rows * rest4, stride = (int64_t)gridDim.x * 256;
  constexpr int kCp = 8;
  for (int64_t i0 = (int64_t)blockIdx.x * 256 + threadIdx.x; i0 < n; i0 += stride * kCp) {
    float4 v[kCp];
    int64_t off[kCp];
#pragma unroll
    for (int u = 0; u < kCp; ++u) {
      const int64_t i = i0 + stride * u;
      off[u] = i < n ? (i / rest4) * C + c0 + 4 * (i % rest4) : -1;
      if (off[u] >= 0) v[u] = __ldcs(reinterpret_cast<const float4*>(x + off[u]));      // read once here: streaming
    }
#pragma unroll
    for (int u = 0; u < kCp; ++u)
      if (off[u] >= 0) *reinterpret_cast<float4*>(y + off[u]) = v[u];
  }
}

// ------------------------------------------------------------------------------------ 2x2 space-to-depth
// out[(b, h2, w2), (kh * 2 + kw) * C + c] = in[(b, 2 h2 + kh, 2 w2 + kw), c]: the im2col of the k2 s2 patch-merging
// convolution (fasternet.py:264-277), so that it becomes one K = 4C contraction.
__global__ void __launch_bounds__(256) space_to_depth_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                             int64_t n4, int H2, int W2, int c4n) {
  const float4* in4 = reinterpret_cast<const float4*>(in);
  float4* out4 = reinterpret_cast<float4*>(out);
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n4; i += (int64_t)gridDim.x * 256) {
    const int cg = (int)(i % c4n);
    int64_t r = i / c4n;
    const int tap = (int)(r & 3);
    r >>= 2;
    const int w2 = (int)(r % W2);
    r /= W2;
    const int h2 = (int)(r % H2);
    const int64_t b = r / H2;
    const int64_t src = ((b * (2 * H2) + 2 * h2 + (tap >> 1)) * (2 * W2) + 2 * w2 + (tap & 1)) * c4n + cg;
    out4[i] = in4[src];
  }
}

// ------------------------------------------------------------------------------------ global average pool
// out[b, c] = mean_p x[(b, p), c]   (AdaptiveAvgPool2d(1), fasternet.py:372); grid B, block (C, 4): four row
// lanes per column, combined in lane order.
__global__ void avgpool_kernel(const float* __restrict__ x, float* __restrict__ out, int P, int C, int64_t ld_out) {
  extern __shared__ float spool[];                 // [4][C]
  const int c = threadIdx.x, lane = threadIdx.y;
  const float* xb = x + (size_t)blockIdx.x * P * C;
  float a0 = 0.f, a1 = 0.f;
  int p = lane;
  for (; p + 4 < P; p += 8) { a0 += xb[(size_t)p * C + c]; a1 += xb[(size_t)(p + 4) * C + c]; }
  if (p < P) a0 += xb[(size_t)p * C + c];
  spool[lane * C + c] = a0 + a1;
  __syncthreads();
  if (lane == 0)
    out[(size_t)blockIdx.x * ld_out + c] = __fdiv_rn((spool[c] + spool[C + c]) + (spool[2 * C + c] + spool[3 * C + c]), (float)P);
}

}  // namespace

// ------------------------------------------------------------------------------------ host side
namespace {

constexpr int kBnMaxBlocks = 592;                 // 4 x 148 slabs at most
constexpr int kBnTicket = 8;                      // ctx->counters word of the BatchNorm reduction
constexpr int64_t kHdrFloats = (int64_t)kBnMaxBlocks * 384 * 2 * 2 + 2048 + 1024;   // fp64 partials + totals + scale/shift
constexpr int64_t kPerSample = (int64_t)kPH * kPW * 384;                       // R0 + R1 + R2 floats per sample

struct EncWs {
  double* partial;
  double* totals;
  float* scale_shift;
  float *r0, *r1, *r2;
  bool bn_on_operand;    // BatchNorm + ReLU between a block's two 1x1 convolutions applied on the second one's operand
  const porl_fasternet_t* net;
  porl_stream_t stream;
};

void conv1x1_group(GemmGroup& g, const float* A, int64_t rows, int K, const float* W, int N, float* C, int64_t ldc,
                   const float* bias, int epilogue) {
  g = GemmGroup{};
  g.A = A; g.B = W; g.C = C; g.bias = bias;
  g.M = (int)rows; g.N = N; g.K = K;
  g.lda = K; g.ldb = K; g.ldc = (int)ldc; g.n_store = N; g.epilogue = epilogue;
}

int conv1x1(porl_ctx* ctx, const float* A, int64_t rows, int K, const float* W, int N, float* C, int64_t ldc,
            const float* bias, int epilogue, cudaStream_t st) {
  GemmBatch gb;
  gb.n_groups = 1;
  gb.splits = 1;
  conv1x1_group(gb.g[0], A, rows, K, W, N, C, ldc, bias, epilogue);
  return launch_gemm(ctx, true, true, gb, st);
}

// 1x1 convolution whose A operand is read through the BatchNorm + ReLU of the layer before: affine = scale | shift per
// input channel (w.scale_shift as bn_finalize_kernel leaves it), applied by the contraction's operand splitters
int conv1x1_affine(porl_ctx* ctx, const float* A, int64_t rows, int K, const float* W, int N, float* C, int64_t ldc,
                   const float* affine, cudaStream_t st) {
  GemmBatch gb;
  gb.n_groups = 1;
  gb.splits = 1;
  conv1x1_group(gb.g[0], A, rows, K, W, N, C, ldc, nullptr, 0);
  gb.g[0].a_affine = affine;
  return launch_gemm(ctx, true, true, gb, st);
}

// apply == 0: only scale / shift are produced (w.scale_shift); the consumer applies them on the way in
int batch_norm(porl_ctx* ctx, const EncWs& w, float* x, int64_t rows, int C, const porl_bn_t& bn, int relu, cudaStream_t st,
               int apply = 1) {
  PORL_CHECK_ARG(bn.weight && bn.bias && bn.running_mean && bn.running_var, "fasternet: null BatchNorm tensor");
  int64_t G = (rows + 255) / 256;
  if (G > kBnMaxBlocks) G = kBnMaxBlocks;
  const int64_t rpb = (rows + G - 1) / G;
  G = (rows + rpb - 1) / rpb;
  const size_t smem = (size_t)(kBnThreads / (C / 4)) * C * 2 * sizeof(double);
  {
    ProfScope prof(ctx, PORL_CAT_BATCHNORM, (double)rows * C * 4.0, st);
    bn_stats_kernel<kBnThreads><<<(unsigned)G, kBnThreads, smem, st>>>(x, rows, C, rpb, w.partial, ctx->counters + kBnTicket, w.totals);
    PORL_LAUNCHED(ctx);
  }
  double rows_global = (double)rows;
  if (w.net->world > 1) {
    const int rc = w.net->stats_allreduce(w.net->stats_user, w.totals, 2 * (int64_t)C, w.stream);
    if (rc != 0) {
      set_error("fasternet: stats_allreduce callback returned %d", rc);
      return PORL_ERR_INVALID;
    }
    rows_global *= (double)w.net->world;
  }
  bn_finalize_kernel<<<1, 128, 0, st>>>(w.totals, rows_global, C, bn.weight, bn.bias, bn.running_mean, bn.running_var,
                                        1e-5f, 0.1f, w.scale_shift);
  PORL_LAUNCHED(ctx);
  if (!apply) return PORL_OK;
  const int64_t n4 = rows * C / 4;
  int64_t grid = (n4 + 255) / 256;
  if (grid > (int64_t)ctx->sm_count * 16) grid = (int64_t)ctx->sm_count * 16;
  ProfScope prof(ctx, PORL_CAT_BATCHNORM, (double)rows * C * 8.0, st);
  bn_apply_kernel<<<(unsigned)grid, 256, 0, st>>>(x, n4, C / 4, w.scale_shift, C, relu);
  PORL_LAUNCHED(ctx);
  return PORL_OK;
}

// MLPBlock.forward: x (rows, C) is updated in place; xm (rows, C) and z (rows, 2C) are scratch
int mlp_block(porl_ctx* ctx, const EncWs& w, const porl_fn_block_t& blk, float* x, float* xm, float* z, int64_t B, int H,
              int W, int C, cudaStream_t st) {
  PORL_CHECK_ARG(blk.pconv_w && blk.w1 && blk.w2, "fasternet: null block weight");
  const int64_t rows = B * H * W;
  {
    const double cc = C / 4;
    ProfScope prof(ctx, PORL_CAT_PCONV, (double)rows * 18.0 * cc * cc, st);      // 2 * 9 * (C/4)^2 flop per pixel
    // C = 96: 4 pixels x 24 channels per thread, all 24 input channels staged at once (16 x 32 tiles);
    // C = 192: 2 pixels x 48 channels, input channels 24 at a time (16 x 16 tiles)
    using K24 = PconvCfg<24, 4, 24>;
    using K48 = PconvCfg<48, 2, 24>;
    static bool configured = false;
    if (!configured) {
      PORL_CUDA(cudaFuncSetAttribute(pconv3_kernel<24, 4, 24>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)K24::SMEM));
      PORL_CUDA(cudaFuncSetAttribute(pconv3_kernel<48, 2, 24>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)K48::SMEM));
      configured = true;
    }
    {
      const int64_t n4 = rows * (C - C / 4) / 4;
      int64_t cg = (n4 + 256 * 8 - 1) / (256 * 8);
      if (cg > (int64_t)ctx->sm_count * 8) cg = (int64_t)ctx->sm_count * 8;
      copy_channels_kernel<<<(unsigned)cg, 256, 0, st>>>(x, xm, rows, C, C / 4);
      PORL_LAUNCHED(ctx);
    }
    if (C == 96) {
      const dim3 grid((W + 15) / 16, (H + K24::TH - 1) / K24::TH, (unsigned)B);
      pconv3_kernel<24, 4, 24><<<grid, 128, K24::SMEM, st>>>(x, xm, blk.pconv_w, H, W, C);
    } else {
      const dim3 grid((W + 15) / 16, (H + K48::TH - 1) / K48::TH, (unsigned)B);
      pconv3_kernel<48, 2, 24><<<grid, 128, K48::SMEM, st>>>(x, xm, blk.pconv_w, H, W, C);
    }
    PORL_LAUNCHED(ctx);
  }
  int rc = conv1x1(ctx, xm, rows, C, blk.w1, 2 * C, z, 2 * C, nullptr, 0, st);
  if (rc != PORL_OK) return rc;
  // BatchNorm + ReLU between the two 1x1 convolutions: statistics of z in one pass; the normalised activation is never
  // written -- the second contraction reads z through scale / shift / ReLU (same arithmetic as bn_apply_kernel, in its
  // operand splitters), which saves a read and a write of the widest activation of the block
  rc = batch_norm(ctx, w, z, rows, 2 * C, blk.bn, 1, st, w.bn_on_operand ? 0 : 1);
  if (rc != PORL_OK) return rc;
  rc = w.bn_on_operand ? conv1x1_affine(ctx, z, rows, 2 * C, blk.w2, C, xm, C, w.scale_shift, st)
                       : conv1x1(ctx, z, rows, 2 * C, blk.w2, C, xm, C, nullptr, 0, st);      // branch output reuses xm
  if (rc != PORL_OK) return rc;
  const int64_t n4 = rows * C / 4;
  int64_t grid = (n4 + 255) / 256;
  if (grid > (int64_t)ctx->sm_count * 16) grid = (int64_t)ctx->sm_count * 16;
  ProfScope prof(ctx, PORL_CAT_ENCODER, (double)rows * C * 12.0, st);
  residual_kernel<<<(unsigned)grid, 256, 0, st>>>(x, xm, n4, (int64_t)H * W * C / 4, blk.keep);
  PORL_LAUNCHED(ctx);
  return PORL_OK;
}

}  // namespace
}  // namespace porl

using namespace porl;

extern "C" int porl_costmap(porl_ctx_t* ctx, float* state, int64_t ld_state, int64_t B, float* out, porl_stream_t stream) {
  PORL_ON_DEVICE(ctx);
  PORL_CHECK_ARG(ctx && state && out, "porl_costmap: null argument");
  PORL_CHECK_ARG(B >= 0 && B <= 65535 && ld_state >= kState, "porl_costmap: B=%lld ld_state=%lld", (long long)B,
                 (long long)ld_state);
  PORL_CHECK_ARG(aligned16(out), "porl_costmap: out must be 16-byte aligned");
  if (B == 0) return PORL_OK;
  cudaStream_t st = as_stream(stream);
  ProfScope prof(ctx, PORL_CAT_ENCODER, (double)B * kScan * kBins * 3 * 4.0, st);
  costmap_kernel<<<dim3(kScan / 8, (unsigned)B), 256, 0, st>>>(state, ld_state, out, ctx->counters);
  PORL_LAUNCHED(ctx);
  return PORL_OK;
}

extern "C" int64_t porl_fasternet_workspace_floats(int64_t B) { return B < 0 ? -1 : kHdrFloats + B * kPerSample; }

extern "C" int porl_fasternet_forward(porl_ctx_t* ctx, const porl_fasternet_t* net, float* state, int64_t ld_state,
                                      int64_t B, float* features, int64_t ld_feat, float* workspace,
                                      int64_t workspace_floats, porl_stream_t stream) {
  PORL_ON_DEVICE(ctx);
  PORL_CHECK_ARG(ctx && net && state && features && workspace, "porl_fasternet_forward: null argument");
  // rows = B * 5760 is a 32-bit extent of the contractions and B a grid dimension; offsets are 64-bit everywhere
  PORL_CHECK_ARG(B >= 1 && B <= 16384, "porl_fasternet_forward: B=%lld outside [1, 16384]", (long long)B);
  PORL_CHECK_ARG(ld_state >= kState, "porl_fasternet_forward: ld_state=%lld", (long long)ld_state);
  const int F = net->feature_dim;
  PORL_CHECK_ARG(F >= 1 && ld_feat >= F, "porl_fasternet_forward: feature_dim=%d ld_feat=%lld", F, (long long)ld_feat);
  PORL_CHECK_ARG(workspace_floats >= porl_fasternet_workspace_floats(B) && aligned16(workspace),
                 "porl_fasternet_forward: workspace of %lld floats, %lld needed (16-byte aligned)",
                 (long long)workspace_floats, (long long)porl_fasternet_workspace_floats(B));
  PORL_CHECK_ARG(net->pe_w && net->merge_w && net->head_conv_w && net->head_w && net->head_b,
                 "porl_fasternet_forward: null weight");
  PORL_CHECK_ARG(net->world <= 1 || net->stats_allreduce, "porl_fasternet_forward: world=%d needs stats_allreduce",
                 net->world);
  cudaStream_t st = as_stream(stream);
  const int64_t rows0 = B * kPH * kPW, rows1 = rows0 / 4;
  EncWs w;
  w.partial = reinterpret_cast<double*>(workspace);
  w.totals = net->stats_buf ? net->stats_buf : w.partial + (int64_t)kBnMaxBlocks * 384 * 2;
  w.scale_shift = workspace + (int64_t)kBnMaxBlocks * 384 * 2 * 2 + 2048;
  w.net = net;
  w.stream = stream;
  w.r0 = workspace + kHdrFloats;
  w.r1 = w.r0 + rows0 * 96;
  w.r2 = w.r1 + rows0 * 96;
  // PORL_ENC_FUSE_BN=0: BatchNorm apply as a pass of its own (the round-1 form, kept for comparison and the FFMA mode)
  w.bn_on_operand = ctx->gemm_mode == 1 && !(getenv("PORL_ENC_FUSE_BN") && atoi(getenv("PORL_ENC_FUSE_BN")) == 0);

  // patch embedding straight from the scan (costmap never materialised) + BatchNorm
  {
    ProfScope prof(ctx, PORL_CAT_ENCODER, (double)rows0 * 96 * 4.0, st);
    patch_embed_kernel<<<dim3(kPH, (unsigned)B), 192, 0, st>>>(state, ld_state, net->pe_w, w.r0, ctx->counters);
    PORL_LAUNCHED(ctx);
  }
  int rc = batch_norm(ctx, w, w.r0, rows0, 96, net->pe_bn, 0, st);
  if (rc != PORL_OK) return rc;
  // stage 0: one MLPBlock of width 96 on the 90 x 64 grid
  rc = mlp_block(ctx, w, net->block[0], w.r0, w.r1, w.r2, B, kPH, kPW, 96, st);
  if (rc != PORL_OK) return rc;
  // patch merging: k2 s2 convolution as space-to-depth + one K = 384 contraction, then BatchNorm
  float* a1 = w.r2;
  float* x1 = w.r1;
  {
    const int64_t n4 = rows1 * 384 / 4;
    int64_t grid = (n4 + 255) / 256;
    if (grid > (int64_t)ctx->sm_count * 16) grid = (int64_t)ctx->sm_count * 16;
    ProfScope prof(ctx, PORL_CAT_ENCODER, (double)rows1 * 384 * 8.0, st);
    space_to_depth_kernel<<<(unsigned)grid, 256, 0, st>>>(w.r0, a1, n4, kPH / 2, kPW / 2, 96 / 4);
    PORL_LAUNCHED(ctx);
  }
  rc = conv1x1(ctx, a1, rows1, 384, net->merge_w, 192, x1, 192, nullptr, 0, st);
  if (rc != PORL_OK) return rc;
  rc = batch_norm(ctx, w, x1, rows1, 192, net->merge_bn, 0, st);
  if (rc != PORL_OK) return rc;
  // stage 1: two MLPBlocks of width 192 on the 45 x 32 grid
  float* xm1 = w.r1 + rows1 * 192;
  float* z1 = w.r2 + rows1 * 384;
  for (int i = 1; i <= 2; ++i) {
    rc = mlp_block(ctx, w, net->block[i], x1, xm1, z1, B, kPH / 2, kPW / 2, 192, st);
    if (rc != PORL_OK) return rc;
  }
  // head: global average pool, 1x1 conv 192 -> 1280 + ReLU, Linear 1280 -> feature_dim
  float* pooled = w.r0;
  float* hidden = w.r0 + B * 192;
  {
    ProfScope prof(ctx, PORL_CAT_ENCODER, (double)rows1 * 192 * 4.0, st);
    avgpool_kernel<<<(unsigned)B, dim3(192, 4), 4 * 192 * sizeof(float), st>>>(x1, pooled, (kPH / 2) * (kPW / 2), 192, 192);
    PORL_LAUNCHED(ctx);
  }
  rc = conv1x1(ctx, pooled, B, 192, net->head_conv_w, 1280, hidden, 1280, nullptr, 1, st);
  if (rc != PORL_OK) return rc;
  return conv1x1(ctx, hidden, B, 1280, net->head_w, F, features, ld_feat, net->head_b, 0, st);
}
