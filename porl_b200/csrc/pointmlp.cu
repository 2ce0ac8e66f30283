// PointMLP scan encoder (agent/feature_pcl.py): forward pass in train mode on (rows x channels) matrices.
//
//   scan_to_points_kernel   (B, 362) states -> (B, 3, 361) point sets                      [feature_pcl.py:113-138, intent]
//   pm_embed_kernel         x (B, 3, N) -> xyz (B, N, 3) and Conv1d(3, E) rows (B*N, E)      [:366, :408-410]
//   fps_kernel              farthest-point sampling, one block per sample, points in smem   [:69-89]
//   knn_kernel              square_distance + top-k smallest, one warp per query            [:28-48, :163-174]
//   gather_anchor_kernel    new_xyz / new_points = index_points(., fps_idx)                 [:212-213]
//   group_stats_kernel      per-sample std of (grouped - mean)                              [:222-227]
//   group_write_kernel      normalise, affine, concat with the anchor feature              [:228-231]
//   pm_bn_apply_kernel      BatchNorm apply (+ residual) (+ ReLU) (* dropout multiplier)    [:263-301, :392-402]
//   maxpool_rows_kernel     max over k consecutive rows (adaptive_max_pool1d)               [:330, :417]
// Every Conv1d(kernel 1) / Linear is a K-major contraction through launch_gemm (tcgen05 3xTF32 for the big ones),
// BatchNorm statistics come from bn.cuh (fp64 partials, fixed order).
//
// Integer results (FPS and kNN indices) follow the reference's fp32 arithmetic operation by operation:
//   FPS distance  ((dx*dx + dy*dy) + dz*dz), no contraction into FMAs;  argmax = first maximal index
//   kNN distance  (-2 * fma(qz, z, fma(qy, y, qx * x)) + |q|^2) + |p|^2   (the K = 3 matmul of torch's CPU path)
#include "bn.cuh"
#include "gemm.cuh"

#include <math.h>

namespace porl {
namespace {

constexpr int kScanBeams = 360;

__global__ void __launch_bounds__(128) scan_to_points_kernel(const float* __restrict__ state, int64_t ld_state,
                                                             float* __restrict__ out) {
  const int64_t b = blockIdx.x;
  const float* s = state + b * ld_state;
  float* o = out + b * 3 * (kScanBeams + 1);
  for (int i = threadIdx.x; i <= kScanBeams; i += 128) {
    float x, y, f;
    if (i < kScanBeams) {
      float r = s[i];
      if (r > 8.0f) r = 0.0f;
      const float ang = __fmul_rn((float)i, (float)(M_PI / 180.0));
      x = __fmul_rn(cosf(ang), r);
      y = __fmul_rn(sinf(ang), r);
      f = 0.0f;
    } else {
      x = s[kScanBeams];
      y = s[kScanBeams + 1];
      f = 1.0f;
    }
    o[i] = x;
    o[(kScanBeams + 1) + i] = y;
    o[2 * (kScanBeams + 1) + i] = f;
  }
}

// one thread per (row, 4 output channels); the thread of channel group 0 also writes the transposed coordinates
__global__ void __launch_bounds__(256) pm_embed_kernel(const float* __restrict__ x, int N, int E, int64_t rows,
                                                       const float* __restrict__ w, const float* __restrict__ bias,
                                                       float* __restrict__ xyz, float* __restrict__ feat) {
  const int e4n = E >> 2;
  const int64_t total = rows * e4n;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (int64_t)gridDim.x * 256) {
    const int64_t r = i / e4n;
    const int eg = (int)(i % e4n);
    const int64_t b = r / N;
    const int n = (int)(r % N);
    const float* xb = x + b * 3 * N;
    const float px = xb[n], py = xb[N + n], pz = xb[2 * N + n];
    if (eg == 0) {
      xyz[r * 3] = px; xyz[r * 3 + 1] = py; xyz[r * 3 + 2] = pz;
    }
    float o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float* wr = w + (size_t)(4 * eg + j) * 4;
      float acc = fmaf(pz, wr[2], fmaf(py, wr[1], px * wr[0]));
      if (bias) acc += bias[4 * eg + j];
      o[j] = acc;
    }
    *reinterpret_cast<float4*>(feat + r * E + 4 * eg) = make_float4(o[0], o[1], o[2], o[3]);
  }
}

// ---- farthest point sampling: block per sample, SoA coordinates and running distances in shared memory ----
__global__ void __launch_bounds__(256) fps_kernel(const float* __restrict__ xyz, int N, int npoint,
                                                  const int64_t* __restrict__ start, int32_t* __restrict__ idx,
                                                  uint32_t* __restrict__ status) {
  extern __shared__ float sm[];
  float *sx = sm, *sy = sm + N, *sz = sm + 2 * N, *sd = sm + 3 * N;
  __shared__ float wv[2][8];
  __shared__ int wi[2][8];
  const int64_t b = blockIdx.x;
  const float* p = xyz + b * N * 3;
  for (int j = threadIdx.x; j < N; j += 256) {
    sx[j] = p[3 * j]; sy[j] = p[3 * j + 1]; sz[j] = p[3 * j + 2];
    sd[j] = 1e10f;
  }
  long long s0 = start[b];
  if (s0 < 0 || s0 >= N) {
    if (threadIdx.x == 0) atomicOr(status, (uint32_t)PORL_FLAG_BAD_INDEX);
    s0 = s0 < 0 ? 0 : N - 1;
  }
  int cur = (int)s0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int i = 0; i < npoint; ++i) {
    if (threadIdx.x == 0) idx[b * npoint + i] = cur;
    const float cx = sx[cur], cy = sy[cur], cz = sz[cur];
    float best = -1.0f;
    int bi = 0x7fffffff;
    for (int j = threadIdx.x; j < N; j += 256) {
      const float dx = __fsub_rn(sx[j], cx), dy = __fsub_rn(sy[j], cy), dz = __fsub_rn(sz[j], cz);
      const float d = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
      const float dd = fminf(sd[j], d);
      sd[j] = dd;
      if (dd > best) { best = dd; bi = j; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
    }
    const int buf = i & 1;
    if (lane == 0) { wv[buf][warp] = best; wi[buf][warp] = bi; }
    __syncthreads();
    best = wv[buf][0]; bi = wi[buf][0];
#pragma unroll
    for (int w = 1; w < 8; ++w) {
      const float ov = wv[buf][w];
      const int oi = wi[buf][w];
      if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
    }
    cur = bi;
  }
}

// ---- k nearest neighbours: grid (ceil(S / 8), B), one warp per query, distances of the warp in shared memory ----
__global__ void __launch_bounds__(256) knn_kernel(const float* __restrict__ xyz, const float* __restrict__ query, int N,
                                                  int S, int k, int32_t* __restrict__ idx) {
  extern __shared__ float sm[];
  float *sx = sm, *sy = sm + N, *sz = sm + 2 * N, *sn = sm + 3 * N;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* wd = sm + 4 * N + (size_t)warp * N;
  const int64_t b = blockIdx.y;
  const float* p = xyz + b * N * 3;
  for (int j = threadIdx.x; j < N; j += 256) {
    const float x = p[3 * j], y = p[3 * j + 1], z = p[3 * j + 2];
    sx[j] = x; sy[j] = y; sz[j] = z;
    sn[j] = __fadd_rn(__fadd_rn(__fmul_rn(x, x), __fmul_rn(y, y)), __fmul_rn(z, z));
  }
  __syncthreads();
  const int q = blockIdx.x * 8 + warp;
  if (q >= S) return;
  const float* qp = query + (b * S + q) * 3;
  const float qx = qp[0], qy = qp[1], qz = qp[2];
  const float qn = __fadd_rn(__fadd_rn(__fmul_rn(qx, qx), __fmul_rn(qy, qy)), __fmul_rn(qz, qz));
  for (int j = lane; j < N; j += 32) {
    const float dot = __fmaf_rn(qz, sz[j], __fmaf_rn(qy, sy[j], __fmul_rn(qx, sx[j])));
    wd[j] = __fadd_rn(__fadd_rn(__fmul_rn(-2.0f, dot), qn), sn[j]);
  }
  __syncwarp();
  int32_t* out = idx + (b * S + q) * k;
  for (int r = 0; r < k; ++r) {
    float best = INFINITY;
    int bi = 0x7fffffff;
    for (int j = lane; j < N; j += 32) {
      const float v = wd[j];
      if (v < best) { best = v; bi = j; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (ov < best || (ov == best && oi < bi)) { best = ov; bi = oi; }
    }
    if (bi == 0x7fffffff) bi = 0;                     // fewer than k finite distances (NaN input): keep indices valid
    if (lane == (bi & 31)) wd[bi] = INFINITY;
    if (lane == 0) out[r] = bi;
    __syncwarp();
  }
}

// new_xyz (B, S, 3) and new_points (B, S, d) = rows fps_idx of xyz / points
__global__ void __launch_bounds__(256) gather_anchor_kernel(const float* __restrict__ xyz, const float* __restrict__ pts,
                                                            const int32_t* __restrict__ fidx, int N, int S, int d,
                                                            int64_t total_rows, float* __restrict__ new_xyz,
                                                            float* __restrict__ new_pts) {
  const int d4 = d >> 2;
  const int per = d4 + 1;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < total_rows * per; i += (int64_t)gridDim.x * 256) {
    const int64_t r = i / per;
    const int c = (int)(i % per);
    const int64_t b = r / S;
    const int64_t src = b * N + fidx[r];
    if (c == d4) {
      new_xyz[r * 3] = xyz[src * 3]; new_xyz[r * 3 + 1] = xyz[src * 3 + 1]; new_xyz[r * 3 + 2] = xyz[src * 3 + 2];
    } else {
      *reinterpret_cast<float4*>(new_pts + r * d + 4 * c) = *reinterpret_cast<const float4*>(pts + src * d + 4 * c);
    }
  }
}

// value c of neighbour j: feature channel for c < d, coordinate c - d otherwise
__device__ __forceinline__ float group_value(const float* __restrict__ pts, const float* __restrict__ xyz, int64_t row,
                                             int d, int c) {
  return c < d ? pts[row * d + c] : xyz[row * 3 + (c - d)];
}

// grid (S, B): partial[(b, s)] = (sum, sum of squares) of grouped - mean over the anchor's k x D values (fp64)
__global__ void __launch_bounds__(256) group_stats_kernel(const float* __restrict__ pts, const float* __restrict__ xyz,
                                                          const float* __restrict__ new_pts, const float* __restrict__ new_xyz,
                                                          const int32_t* __restrict__ kidx, int N, int S, int k, int d, int D,
                                                          int mode, double* __restrict__ partial) {
  __shared__ double red[2][8];
  const int64_t b = blockIdx.y, a = b * S + blockIdx.x;
  const int32_t* nb = kidx + a * k;
  double s = 0.0, q = 0.0;
  for (int c = threadIdx.x; c < D; c += 256) {
    float mean;
    if (mode == 1) {
      float acc = 0.f;
      for (int j = 0; j < k; ++j) acc += group_value(pts, xyz, b * N + nb[j], d, c);
      mean = acc / (float)k;
    } else {
      mean = c < d ? new_pts[a * d + c] : new_xyz[a * 3 + (c - d)];
    }
    for (int j = 0; j < k; ++j) {
      const float df = __fsub_rn(group_value(pts, xyz, b * N + nb[j], d, c), mean);
      s += (double)df;
      q += (double)df * (double)df;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    q += __shfl_xor_sync(0xffffffffu, q, o);
  }
  if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = s; red[1][threadIdx.x >> 5] = q; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double ts = 0, tq = 0;
    for (int w = 0; w < 8; ++w) { ts += red[0][w]; tq += red[1][w]; }
    partial[a * 2] = ts;
    partial[a * 2 + 1] = tq;
  }
}

// std[b] = unbiased standard deviation over the sample's S * k * D values (torch.std of the flattened tensor)
__global__ void group_std_kernel(const double* __restrict__ partial, int S, double n, float* __restrict__ stdv) {
  const int64_t b = blockIdx.x;
  if (threadIdx.x != 0) return;
  double s = 0, q = 0;
  for (int a = 0; a < S; ++a) { s += partial[(b * S + a) * 2]; q += partial[(b * S + a) * 2 + 1]; }
  double var = n > 1 ? (q - s * s / n) / (n - 1) : 0.0;
  if (var < 0) var = 0;
  stdv[b] = (float)sqrt(var);
}

// grid (S, B): rows (b, s, j) of the grouped matrix: [ alpha * (v - mean) / (std + 1e-5) + beta | anchor feature | 0 pad ]
__global__ void __launch_bounds__(256) group_write_kernel(const float* __restrict__ pts, const float* __restrict__ xyz,
                                                          const float* __restrict__ new_pts, const float* __restrict__ new_xyz,
                                                          const int32_t* __restrict__ kidx, int N, int S, int k, int d, int D,
                                                          int mode, const float* __restrict__ stdv,
                                                          const float* __restrict__ alpha, const float* __restrict__ beta,
                                                          float* __restrict__ out, int ld) {
  const int64_t b = blockIdx.y, a = b * S + blockIdx.x;
  const int32_t* nb = kidx + a * k;
  const float denom = mode ? __fadd_rn(stdv[b], 1e-5f) : 1.0f;
  float* o = out + a * k * (int64_t)ld;
  for (int c = threadIdx.x; c < ld; c += 256) {
    if (c < D) {
      float mean = 0.f;
      if (mode == 1) {
        float acc = 0.f;
        for (int j = 0; j < k; ++j) acc += group_value(pts, xyz, b * N + nb[j], d, c);
        mean = acc / (float)k;
      } else if (mode == 2) {
        mean = c < d ? new_pts[a * d + c] : new_xyz[a * 3 + (c - d)];
      }
      const float al = mode ? alpha[c] : 1.0f, be = mode ? beta[c] : 0.0f;
      for (int j = 0; j < k; ++j) {
        float v = group_value(pts, xyz, b * N + nb[j], d, c);
        if (mode) v = __fadd_rn(__fmul_rn(al, __fdiv_rn(__fsub_rn(v, mean), denom)), be);
        o[(int64_t)j * ld + c] = v;
      }
    } else {
      const float v = c < D + d ? new_pts[a * d + (c - D)] : 0.0f;
      for (int j = 0; j < k; ++j) o[(int64_t)j * ld + c] = v;
    }
  }
}

// out = [relu]( t * scale[c] + shift[c] [+ res] ) [* mul];  t may alias out, res may alias out
__global__ void __launch_bounds__(256) pm_bn_apply_kernel(const float* t, const float* res, const float* __restrict__ mul,
                                                          float* out, int64_t n4, int c4n,
                                                          const float* __restrict__ scale_shift, int C, int relu) {
  const float4* sc4 = reinterpret_cast<const float4*>(scale_shift);
  const float4* sh4 = reinterpret_cast<const float4*>(scale_shift + C);
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n4; i += (int64_t)gridDim.x * 256) {
    const int cg = (int)(i % c4n);
    const float4 sc = __ldg(sc4 + cg), sh = __ldg(sh4 + cg);
    float4 v = reinterpret_cast<const float4*>(t)[i];
    v.x = fmaf(v.x, sc.x, sh.x); v.y = fmaf(v.y, sc.y, sh.y); v.z = fmaf(v.z, sc.z, sh.z); v.w = fmaf(v.w, sc.w, sh.w);
    if (res) {
      const float4 r = reinterpret_cast<const float4*>(res)[i];
      v.x += r.x; v.y += r.y; v.z += r.z; v.w += r.w;
    }
    if (relu) { v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f); }
    if (mul) {
      const float4 m = reinterpret_cast<const float4*>(mul)[i];
      v.x *= m.x; v.y *= m.y; v.z *= m.z; v.w *= m.w;
    }
    reinterpret_cast<float4*>(out)[i] = v;
  }
}

// out[r, :] = max over rows r*k .. r*k + k - 1 of in
__global__ void __launch_bounds__(256) maxpool_rows_kernel(const float* __restrict__ in, int64_t rows_out, int k, int C,
                                                           float* __restrict__ out) {
  const int c4n = C >> 2;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < rows_out * c4n; i += (int64_t)gridDim.x * 256) {
    const int64_t r = i / c4n;
    const int cg = (int)(i % c4n);
    const float4* src = reinterpret_cast<const float4*>(in + r * k * (int64_t)C) + cg;
    float4 m = src[0];
    for (int j = 1; j < k; ++j) {
      const float4 v = src[(int64_t)j * c4n];
      m.x = fmaxf(m.x, v.x); m.y = fmaxf(m.y, v.y); m.z = fmaxf(m.z, v.z); m.w = fmaxf(m.w, v.w);
    }
    reinterpret_cast<float4*>(out + r * C)[cg] = m;
  }
}


// ------------------------------------------------------------------------------------ host side
constexpr int kPmBnMaxBlocks = 592;
constexpr int kPmBnTicket = 8;                    // same ctx->counters word as the FasterNet BatchNorm (one stream)
constexpr int kPmMaxC = 1024;

inline int64_t pad4(int64_t x) { return (x + 3) & ~(int64_t)3; }
inline bool bn_width_ok(int C) { return C >= 4 && C <= kPmMaxC && (C & (C - 1)) == 0; }

struct PmPlan {
  // offsets in floats from the workspace base
  int64_t bn_partial, bn_totals, scale_shift, xyz0, feat0, G, T0, T1, T2, pooled, h1, h2, total;
  struct St { int64_t fidx, kidx, new_xyz, new_pts, std_partial, stdv, P0, P1, P2; int N, S, k, d, D, ld_g, c_out; } st[PORL_PM_MAX_STAGES];
  int c_last;
};

// lays the workspace out; returns false (with a message) when the configuration is outside what the kernels cover
bool pm_plan(const porl_pointmlp_t& net, int64_t B, PmPlan& pl) {
  if (B < 1 || net.points < 1 || net.points > 4096 || net.n_stages < 1 || net.n_stages > PORL_PM_MAX_STAGES ||
      net.normalize < 0 || net.normalize > 2 || net.class_num < 1 || !bn_width_ok(net.embed_dim)) {
    set_error("pointmlp: unsupported configuration (B=%lld points=%d stages=%d embed_dim=%d)", (long long)B, net.points,
              net.n_stages, net.embed_dim);
    return false;
  }
  int64_t off = 0;
  auto take = [&](int64_t n) { const int64_t o = off; off += pad4(n); return o; };
  pl.bn_partial = take((int64_t)kPmBnMaxBlocks * kPmMaxC * 4);
  pl.bn_totals = take(kPmMaxC * 4);
  pl.scale_shift = take(kPmMaxC * 2);
  pl.xyz0 = take(B * net.points * 3);
  pl.feat0 = take(B * net.points * net.embed_dim);
  int64_t g = 0, t0 = 0, t1 = 0;
  int N = net.points, d = net.embed_dim;
  for (int i = 0; i < net.n_stages; ++i) {
    const porl_pm_stage_t& s = net.stage[i];
    PmPlan::St& ps = pl.st[i];
    if (s.anchors < 1 || s.anchors > N || s.k < 1 || s.k > N || s.c_in != d || !bn_width_ok(s.c_out) || s.n_pre < 0 ||
        s.n_pre > PORL_PM_MAX_BLOCKS || s.n_pos < 0 || s.n_pos > PORL_PM_MAX_BLOCKS || (d & 3)) {
      set_error("pointmlp: stage %d unsupported (anchors=%d of %d points, k=%d, c_in=%d expected %d, c_out=%d)", i, s.anchors,
                N, s.k, s.c_in, d, s.c_out);
      return false;
    }
    int c_mid_max = 4;
    for (int j = 0; j < s.n_pre; ++j) c_mid_max = s.pre[j].c_mid > c_mid_max ? s.pre[j].c_mid : c_mid_max;
    for (int j = 0; j < s.n_pos; ++j) c_mid_max = s.pos[j].c_mid > c_mid_max ? s.pos[j].c_mid : c_mid_max;
    for (int j = 0; j < s.n_pre; ++j)
      if (!bn_width_ok(s.pre[j].c_mid)) { set_error("pointmlp: stage %d pre block %d width %d", i, j, s.pre[j].c_mid); return false; }
    for (int j = 0; j < s.n_pos; ++j)
      if (!bn_width_ok(s.pos[j].c_mid)) { set_error("pointmlp: stage %d pos block %d width %d", i, j, s.pos[j].c_mid); return false; }
    ps.N = N; ps.S = s.anchors; ps.k = s.k; ps.d = d; ps.D = d + (net.use_xyz ? 3 : 0);
    ps.ld_g = (int)pad4(ps.D + d); ps.c_out = s.c_out;
    const int64_t anchors = B * s.anchors, rows = anchors * s.k;
    ps.fidx = take(anchors); ps.kidx = take(rows);
    ps.new_xyz = take(anchors * 3); ps.new_pts = take(anchors * d);
    ps.std_partial = take(anchors * 4); ps.stdv = take(B);
    ps.P0 = take(anchors * s.c_out); ps.P1 = take(anchors * c_mid_max); ps.P2 = take(anchors * s.c_out);
    g = rows * ps.ld_g > g ? rows * ps.ld_g : g;
    t0 = rows * s.c_out > t0 ? rows * s.c_out : t0;
    t1 = rows * c_mid_max > t1 ? rows * c_mid_max : t1;
    N = s.anchors; d = s.c_out;
  }
  pl.c_last = d;
  pl.G = take(g); pl.T0 = take(t0); pl.T1 = take(t1); pl.T2 = take(t0);
  pl.pooled = take(B * d); pl.h1 = take(B * 512); pl.h2 = take(B * 256);
  pl.total = off;
  return true;
}

struct PmWs {
  porl_ctx* ctx;
  double *partial, *totals;
  float* scale_shift;
  cudaStream_t st;
};

inline unsigned ew_grid(porl_ctx* ctx, int64_t n) {
  int64_t g = (n + 255) / 256;
  const int64_t cap = (int64_t)ctx->sm_count * 16;
  return (unsigned)(g > cap ? cap : (g < 1 ? 1 : g));
}

int pm_gemm(porl_ctx* ctx, const float* A, int64_t rows, int K, int lda, const float* W, int ldw, int N, float* C, int64_t ldc,
            const float* bias, cudaStream_t st) {
  GemmBatch gb;
  gb.n_groups = 1;
  gb.splits = 1;
  GemmGroup& g = gb.g[0];
  g = GemmGroup{};
  g.A = A; g.B = W; g.C = C; g.bias = bias;
  g.M = (int)rows; g.N = N; g.K = K;
  g.lda = lda; g.ldb = ldw; g.ldc = (int)ldc; g.n_store = N; g.epilogue = 0;
  return launch_gemm(ctx, true, true, gb, st);
}

// BatchNorm1d training forward over x (rows, C): out = [relu](bn(x) [+ res]) [* mul]; out may be x
int pm_batch_norm(const PmWs& w, const float* x, int64_t rows, int C, const porl_bn_t& bn, int relu, const float* res,
                  const float* mul, float* out) {
  porl_ctx* ctx = w.ctx;
  PORL_CHECK_ARG(bn.weight && bn.bias && bn.running_mean && bn.running_var, "pointmlp: null BatchNorm tensor");
  PORL_CHECK_ARG(bn_width_ok(C), "pointmlp: BatchNorm width %d is not a power of two in [4, %d]", C, kPmMaxC);
  int64_t G = (rows + 255) / 256;
  if (G > kPmBnMaxBlocks) G = kPmBnMaxBlocks;
  const int64_t rpb = (rows + G - 1) / G;
  G = (rows + rpb - 1) / rpb;
  const size_t smem = (size_t)(256 / (C / 4)) * C * 2 * sizeof(double);
  {
    ProfScope prof(ctx, PORL_CAT_BATCHNORM, (double)rows * C * 4.0, w.st);
    bn_stats_kernel<256><<<(unsigned)G, 256, smem, w.st>>>(x, rows, C, rpb, w.partial, ctx->counters + kPmBnTicket, w.totals);
    PORL_LAUNCHED(ctx);
  }
  bn_finalize_kernel<<<1, 128, 0, w.st>>>(w.totals, (double)rows, C, bn.weight, bn.bias, bn.running_mean, bn.running_var, 1e-5f,
                                          0.1f, w.scale_shift);
  PORL_LAUNCHED(ctx);
  const int64_t n4 = rows * C / 4;
  ProfScope prof(ctx, PORL_CAT_BATCHNORM, (double)rows * C * (res ? 12.0 : 8.0), w.st);
  pm_bn_apply_kernel<<<ew_grid(ctx, n4), 256, 0, w.st>>>(x, res, mul, out, n4, C / 4, w.scale_shift, C, relu);
  PORL_LAUNCHED(ctx);
  return PORL_OK;
}

// Conv1d(k=1) + BatchNorm1d (+ ReLU): in (rows, ld_in) -> out (rows, c_out)
int pm_conv_bn(const PmWs& w, const porl_pm_conv_t& cv, const float* in, int64_t rows, int c_in, int ld_in, int c_out, float* out,
               int relu, const float* res, float* bn_out) {
  PORL_CHECK_ARG(cv.weight, "pointmlp: null convolution weight");
  const int rc = pm_gemm(w.ctx, in, rows, c_in, ld_in, cv.weight, (int)pad4(c_in), c_out, out, c_out, cv.bias, w.st);
  if (rc != PORL_OK) return rc;
  return pm_batch_norm(w, out, rows, c_out, cv.bn, relu, res, nullptr, bn_out);
}

// ConvBNReLURes1D: x (rows, C) <- relu(bn2(conv2(relu(bn1(conv1(x))))) + x); t1 (rows, c_mid), t2 (rows, C) scratch
int pm_res_block(const PmWs& w, const porl_pm_res_t& blk, float* x, int64_t rows, int C, float* t1, float* t2) {
  int rc = pm_conv_bn(w, blk.net1, x, rows, C, C, blk.c_mid, t1, 1, nullptr, t1);
  if (rc != PORL_OK) return rc;
  return pm_conv_bn(w, blk.net2, t1, rows, blk.c_mid, blk.c_mid, C, t2, 1, x, x);
}

int launch_fps(porl_ctx* ctx, const float* xyz, int64_t B, int N, int npoint, const int64_t* start, int32_t* idx,
               cudaStream_t st) {
  PORL_CHECK_ARG(N >= 1 && N <= 4096 && npoint >= 1 && npoint <= N, "fps: N=%d npoint=%d", N, npoint);
  ProfScope prof(ctx, PORL_CAT_ENCODER, (double)B * N * 12.0, st);
  fps_kernel<<<(unsigned)B, 256, (size_t)N * 16, st>>>(xyz, N, npoint, start, idx, ctx->counters);
  PORL_LAUNCHED(ctx);
  return PORL_OK;
}

int launch_knn(porl_ctx* ctx, const float* xyz, const float* query, int64_t B, int N, int S, int k, int32_t* idx,
               cudaStream_t st) {
  PORL_CHECK_ARG(N >= 1 && N <= 4096 && S >= 1 && k >= 1 && k <= N && B <= 65535, "knn: N=%d S=%d k=%d B=%lld", N, S, k,
                 (long long)B);
  const size_t smem = (size_t)N * 12 * sizeof(float);
  static bool raised = false;
  if (!raised) {
    PORL_CUDA(cudaFuncSetAttribute(knn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 4096 * 12 * 4));
    raised = true;
  }
  ProfScope prof(ctx, PORL_CAT_ENCODER, (double)B * (N * 12.0 + S * k * 4.0), st);
  knn_kernel<<<dim3((unsigned)((S + 7) / 8), (unsigned)B), 256, smem, st>>>(xyz, query, N, S, k, idx);
  PORL_LAUNCHED(ctx);
  return PORL_OK;
}

}  // namespace
}  // namespace porl

using namespace porl;

extern "C" int64_t porl_pointmlp_workspace_floats(const porl_pointmlp_t* net, int64_t B) {
  PmPlan pl;
  if (!net || !pm_plan(*net, B, pl)) return -1;
  return pl.total;
}

extern "C" int porl_fps(porl_ctx_t* ctx, const float* xyz, int64_t B, int32_t N, int32_t npoint, const int64_t* start,
                        int32_t* idx, porl_stream_t stream) {
  PORL_ON_DEVICE(ctx);
  PORL_CHECK_ARG(ctx && xyz && start && idx && B >= 1, "porl_fps: null argument");
  return launch_fps(ctx, xyz, B, N, npoint, start, idx, as_stream(stream));
}

extern "C" int porl_knn(porl_ctx_t* ctx, const float* xyz, const float* query, int64_t B, int32_t N, int32_t S, int32_t k,
                        int32_t* idx, porl_stream_t stream) {
  PORL_ON_DEVICE(ctx);
  PORL_CHECK_ARG(ctx && xyz && query && idx && B >= 1, "porl_knn: null argument");
  return launch_knn(ctx, xyz, query, B, N, S, k, idx, as_stream(stream));
}

extern "C" int porl_scan_to_points(porl_ctx_t* ctx, const float* state, int64_t ld_state, int64_t B, float* out,
                                   porl_stream_t stream) {
  PORL_ON_DEVICE(ctx);
  PORL_CHECK_ARG(ctx && state && out && B >= 1 && ld_state >= 362, "porl_scan_to_points: bad argument");
  ProfScope prof(ctx, PORL_CAT_ENCODER, (double)B * (362 + 3 * 361) * 4.0, as_stream(stream));
  scan_to_points_kernel<<<(unsigned)B, 128, 0, as_stream(stream)>>>(state, ld_state, out);
  PORL_LAUNCHED(ctx);
  return PORL_OK;
}

extern "C" int porl_pointmlp_forward(porl_ctx_t* ctx, const porl_pointmlp_t* netp, const float* x, const int64_t* fps_start,
                                     int64_t B, float* out, int64_t ld_out, float* workspace, int64_t workspace_floats,
                                     porl_stream_t stream) {
  PORL_ON_DEVICE(ctx);
  PORL_CHECK_ARG(ctx && netp && x && fps_start && out && workspace, "porl_pointmlp_forward: null argument");
  const porl_pointmlp_t& net = *netp;
  PmPlan pl;
  if (!pm_plan(net, B, pl)) return PORL_ERR_INVALID;
  PORL_CHECK_ARG(B <= 65535, "porl_pointmlp_forward: B=%lld above 65535", (long long)B);
  PORL_CHECK_ARG(ld_out >= net.class_num, "porl_pointmlp_forward: ld_out=%lld < class_num=%d", (long long)ld_out, net.class_num);
  if (workspace_floats < pl.total) {
    set_error("porl_pointmlp_forward: workspace of %lld floats, %lld needed", (long long)workspace_floats, (long long)pl.total);
    return PORL_ERR_WORKSPACE;
  }
  PORL_CHECK_ARG(aligned16(workspace) && aligned16(x), "porl_pointmlp_forward: buffers must be 16-byte aligned");
  PORL_CHECK_ARG(net.fc1_w && net.fc1_b && net.fc2_w && net.fc2_b && net.fc3_w && net.fc3_b && net.embedding.weight,
                 "porl_pointmlp_forward: null classifier / embedding tensor");
  cudaStream_t st = as_stream(stream);
  float* ws = workspace;
  PmWs w{ctx, reinterpret_cast<double*>(ws + pl.bn_partial), reinterpret_cast<double*>(ws + pl.bn_totals), ws + pl.scale_shift, st};

  // embedding: Conv1d(3, E) + BN + ReLU on (B*N, E) rows; coordinates transposed to (B, N, 3)
  const int E = net.embed_dim;
  float* xyz = ws + pl.xyz0;
  float* feat = ws + pl.feat0;
  {
    const int64_t rows = B * net.points;
    ProfScope prof(ctx, PORL_CAT_ENCODER, (double)rows * (24.0 + 4.0 * E), st);
    pm_embed_kernel<<<ew_grid(ctx, rows * (E / 4)), 256, 0, st>>>(x, net.points, E, rows, net.embedding.weight,
                                                                  net.embedding.bias, xyz, feat);
    PORL_LAUNCHED(ctx);
    const int rc = pm_batch_norm(w, feat, rows, E, net.embedding.bn, 1, nullptr, nullptr, feat);
    if (rc != PORL_OK) return rc;
  }

  float *G = ws + pl.G, *T0 = ws + pl.T0, *T1 = ws + pl.T1, *T2 = ws + pl.T2;
  for (int i = 0; i < net.n_stages; ++i) {
    const porl_pm_stage_t& s = net.stage[i];
    const PmPlan::St& ps = pl.st[i];
    PORL_CHECK_ARG(net.normalize == 0 || (s.affine_alpha && s.affine_beta), "pointmlp: stage %d affine parameters missing", i);
    int32_t* fidx = reinterpret_cast<int32_t*>(ws + ps.fidx);
    int32_t* kidx = reinterpret_cast<int32_t*>(ws + ps.kidx);
    float *new_xyz = ws + ps.new_xyz, *new_pts = ws + ps.new_pts, *stdv = ws + ps.stdv;
    double* std_partial = reinterpret_cast<double*>(ws + ps.std_partial);
    const int64_t anchors = B * ps.S, rows = anchors * ps.k;
    int rc = launch_fps(ctx, xyz, B, ps.N, ps.S, fps_start + (int64_t)i * B, fidx, st);
    if (rc != PORL_OK) return rc;
    {
      ProfScope prof(ctx, PORL_CAT_ENCODER, (double)anchors * (ps.d + 3) * 8.0, st);
      gather_anchor_kernel<<<ew_grid(ctx, anchors * (ps.d / 4 + 1)), 256, 0, st>>>(xyz, feat, fidx, ps.N, ps.S, ps.d, anchors,
                                                                                  new_xyz, new_pts);
      PORL_LAUNCHED(ctx);
    }
    rc = launch_knn(ctx, xyz, new_xyz, B, ps.N, ps.S, ps.k, kidx, st);
    if (rc != PORL_OK) return rc;
    const dim3 ggrid((unsigned)ps.S, (unsigned)B);
    if (net.normalize) {
      ProfScope prof(ctx, PORL_CAT_ENCODER, (double)rows * ps.D * 4.0, st);
      group_stats_kernel<<<ggrid, 256, 0, st>>>(feat, xyz, new_pts, new_xyz, kidx, ps.N, ps.S, ps.k, ps.d, ps.D, net.normalize,
                                                std_partial);
      PORL_LAUNCHED(ctx);
      group_std_kernel<<<(unsigned)B, 32, 0, st>>>(std_partial, ps.S, (double)ps.S * ps.k * ps.D, stdv);
      PORL_LAUNCHED(ctx);
    }
    {
      ProfScope prof(ctx, PORL_CAT_ENCODER, (double)rows * (ps.D + ps.ld_g) * 4.0, st);
      group_write_kernel<<<ggrid, 256, 0, st>>>(feat, xyz, new_pts, new_xyz, kidx, ps.N, ps.S, ps.k, ps.d, ps.D, net.normalize,
                                                stdv, s.affine_alpha, s.affine_beta, G, ps.ld_g);
      PORL_LAUNCHED(ctx);
    }
    // PreExtraction: transfer + residual blocks over (B*S*k) rows, then max over the k neighbours
    rc = pm_conv_bn(w, s.transfer, G, rows, ps.D + ps.d, ps.ld_g, s.c_out, T0, 1, nullptr, T0);
    if (rc != PORL_OK) return rc;
    for (int j = 0; j < s.n_pre; ++j) {
      rc = pm_res_block(w, s.pre[j], T0, rows, s.c_out, T1, T2);
      if (rc != PORL_OK) return rc;
    }
    float *P0 = ws + ps.P0, *P1 = ws + ps.P1, *P2 = ws + ps.P2;
    {
      ProfScope prof(ctx, PORL_CAT_ENCODER, (double)rows * s.c_out * 4.0, st);
      maxpool_rows_kernel<<<ew_grid(ctx, anchors * (s.c_out / 4)), 256, 0, st>>>(T0, anchors, ps.k, s.c_out, P0);
      PORL_LAUNCHED(ctx);
    }
    // PosExtraction over (B*S) rows
    for (int j = 0; j < s.n_pos; ++j) {
      rc = pm_res_block(w, s.pos[j], P0, anchors, s.c_out, P1, P2);
      if (rc != PORL_OK) return rc;
    }
    xyz = new_xyz;
    feat = P0;
  }

  // global max over the anchors, classifier
  const int Cl = pl.c_last, S_last = pl.st[net.n_stages - 1].S;
  float *pooled = ws + pl.pooled, *h1 = ws + pl.h1, *h2 = ws + pl.h2;
  {
    ProfScope prof(ctx, PORL_CAT_ENCODER, (double)B * S_last * Cl * 4.0, st);
    maxpool_rows_kernel<<<ew_grid(ctx, B * (Cl / 4)), 256, 0, st>>>(feat, B, S_last, Cl, pooled);
    PORL_LAUNCHED(ctx);
  }
  int rc = pm_gemm(ctx, pooled, B, Cl, Cl, net.fc1_w, Cl, 512, h1, 512, net.fc1_b, st);
  if (rc != PORL_OK) return rc;
  rc = pm_batch_norm(w, h1, B, 512, net.bn1, 1, nullptr, net.drop1, h1);
  if (rc != PORL_OK) return rc;
  rc = pm_gemm(ctx, h1, B, 512, 512, net.fc2_w, 512, 256, h2, 256, net.fc2_b, st);
  if (rc != PORL_OK) return rc;
  rc = pm_batch_norm(w, h2, B, 256, net.bn2, 1, nullptr, net.drop2, h2);
  if (rc != PORL_OK) return rc;
  return pm_gemm(ctx, h2, B, 256, 256, net.fc3_w, 256, net.class_num, out, ld_out, net.fc3_b, st);
}
