// K2 / K3: MLP forward and backward as sequences of grouped contractions with fused epilogues
// (bias + ReLU / tanh forward; ReLU mask, bias gradient and zero padding backward), plus the
// optional LayerNorm rows of util/util.py:34-39.
//
// Layout (all fp32, row-major):
//   flat parameters, per net, in the reference's parameters() order:
//       W_l (dims[l+1], pad32(dims[l]))   b_l (pad4(dims[l+1]))   [ln_w_l, ln_b_l (pad4 each)]
//   saved activations, per net, per hidden layer l = 1..L-1:
//       h_l (B, pad32(dims[l]))           [xhat_l (B, pad32(dims[l])), rstd_l (pad4(B))]
//   head output: (n_nets, B, pad4(dims[L]))
// Pad columns of weights, activations and gradients are zero, so contractions may run over the
// padded extents (the tcgen05 path does) without changing results.
#include <math.h>

#include "gemm.cuh"
#include "head.cuh"

struct porl_mlp {
  porl_ctx* ctx;
  porl_mlp_desc_t d;
  int64_t w_off[PORL_MAX_LAYERS], b_off[PORL_MAX_LAYERS], lnw_off[PORL_MAX_LAYERS], lnb_off[PORL_MAX_LAYERS];
  int64_t net_stride;     // floats per net in the flat buffer
  int pad[PORL_MAX_LAYERS + 1];  // pad32(dims[l])
  int ld_out;
};

namespace porl {

static inline int pad32(int x) { return (int)round_up(x, PORL_PAD); }
static inline int pad4(int64_t x) { return (int)round_up(x, 4); }

// Layout of the caller's `saved` buffer (the same in both contraction modes, so the mode may be switched between
// updates), per net, per hidden layer l = 1..L-1:
//     h_l (B x pad32(dims[l]))   relu bits of h_l (B x bits_ld(l) 32-bit words: bit j of word w = h_l[., 32 w + j] > 0)
//     [+ xhat (B x pad32), rstd (pad4(B)) with LayerNorm]
// The tcgen05 kernel reads MN-major operands in place (dz and a_l for the weight gradient, W_l for the input gradient),
// so no transposed copies are kept.  The relu bits are written by the forward epilogue that produces h_l and read by
// the input-gradient epilogue of layer l as one 16-byte load per row (instead of the 512-byte row of h_l).
static inline int bits_ld(const porl_mlp* m, int l) { return (int)round_up(m->pad[l] / 32, 4); }
static int64_t saved_layer_floats(const porl_mlp* m, int l /*hidden layer 1..L-1*/, int64_t B) {
  int64_t f = B * m->pad[l] + pad4(B * bits_ld(m, l));
  if (m->d.layer_norm) f += B * m->pad[l] + pad4(B);
  return f;
}
static int64_t saved_net_floats(const porl_mlp* m, int64_t B) {
  int64_t f = 0;
  for (int l = 1; l < m->d.n_layers; ++l) f += saved_layer_floats(m, l, B);
  return f;
}
static float* saved_h(const porl_mlp* m, float* saved, int net, int l, int64_t B) {
  float* p = saved + net * saved_net_floats(m, B);
  for (int i = 1; i < l; ++i) p += saved_layer_floats(m, i, B);
  return p;
}
static uint32_t* saved_bits(const porl_mlp* m, float* saved, int net, int l, int64_t B) {
  return reinterpret_cast<uint32_t*>(saved_h(m, saved, net, l, B) + B * m->pad[l]);
}
static float* saved_xhat(const porl_mlp* m, float* saved, int net, int l, int64_t B) {
  return saved_h(m, saved, net, l, B) + B * m->pad[l] + pad4(B * bits_ld(m, l));
}
static float* saved_rstd(const porl_mlp* m, float* saved, int net, int l, int64_t B) {
  return saved_xhat(m, saved, net, l, B) + B * m->pad[l];
}
static int64_t saved_total_floats(const porl_mlp* m, int64_t B) { return saved_net_floats(m, B) * m->d.n_nets; }
// the output layer is served by the row-wise head kernels (head.cu) instead of a contraction
static bool narrow_head(const porl_mlp* m) {
  return m->d.dims[m->d.n_layers] <= kHeadMax && m->ld_out <= 32;
}
// layer l (dims[l] -> dims[l+1]) runs its three contractions on the tcgen05 path
static bool tc_layer(const porl_ctx* ctx, const porl_mlp* m, int l, int64_t B) {
  if (ctx->gemm_mode != 1 || m->d.layer_norm || l < 0 || l >= m->d.n_layers) return false;
  const int64_t di = m->d.dims[l], dn = m->d.dims[l + 1];
  return di >= 32 && dn >= 32 && B >= 32 && (double)B * di * dn >= (double)(1 << 20);
}

// ---------------------------------------------------------------- LayerNorm rows
// z (B, ld) in place -> h = relu(xhat * gamma + beta); xhat and rstd saved.  One warp per row.
__global__ void __launch_bounds__(256)
ln_relu_forward_kernel(float* __restrict__ z, float* __restrict__ xhat, float* __restrict__ rstd,
                       const float* __restrict__ gamma, const float* __restrict__ beta, int64_t B, int d,
                       int ld) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= B) return;
  float* zr = z + row * ld;
  float* xr = xhat + row * ld;
  float s = 0.f;
  for (int j = lane; j < d; j += 32) s += zr[j];
  const float mean = warp_sum(s) / (float)d;
  float q = 0.f;
  for (int j = lane; j < d; j += 32) {
    const float c = zr[j] - mean;
    q = fmaf(c, c, q);
  }
  const float var = warp_sum(q) / (float)d;     // biased, as torch.nn.LayerNorm
  const float rs = rsqrtf(var + 1e-5f);
  for (int j = lane; j < d; j += 32) {
    const float xh = (zr[j] - mean) * rs;
    xr[j] = xh;
    zr[j] = fmaxf(fmaf(xh, gamma[j], beta[j]), 0.f);
  }
  for (int j = d + lane; j < ld; j += 32) {
    xr[j] = 0.f;
    zr[j] = 0.f;
  }
  if (lane == 0) rstd[row] = rs;
}

// dy (B, ld) in place -> dz = rstd * (g - mean_j(g) - xhat * mean_j(g * xhat)),  g = dy * gamma
__global__ void __launch_bounds__(256)
ln_backward_kernel(float* __restrict__ dy, const float* __restrict__ xhat, const float* __restrict__ rstd,
                   const float* __restrict__ gamma, int64_t B, int d, int ld) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= B) return;
  float* dr = dy + row * ld;
  const float* xr = xhat + row * ld;
  float s1 = 0.f, s2 = 0.f;
  for (int j = lane; j < d; j += 32) {
    const float g = dr[j] * gamma[j];
    s1 += g;
    s2 = fmaf(g, xr[j], s2);
  }
  s1 = warp_sum(s1) / (float)d;
  s2 = warp_sum(s2) / (float)d;
  const float rs = rstd[row];
  for (int j = lane; j < d; j += 32) {
    const float g = dr[j] * gamma[j];
    dr[j] = rs * (g - s1 - xr[j] * s2);
  }
}

// dgamma[j] = sum_m dy[m,j] * xhat[m,j];  dbeta[j] = sum_m dy[m,j].  One block per 32 columns,
// 8 warps stride the rows, fixed-order shared-memory reduction -> deterministic.
__global__ void __launch_bounds__(256)
ln_param_grad_kernel(const float* __restrict__ dy, const float* __restrict__ xhat, int64_t B, int d,
                     int ld, float* __restrict__ dgamma, float* __restrict__ dbeta) {
  __shared__ float sg[8][33], sb[8][33];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int j = blockIdx.x * 32 + lane;
  float ag = 0.f, ab = 0.f;
  if (j < d) {
    for (int64_t m = warp; m < B; m += 8) {
      const float g = dy[m * ld + j];
      ab += g;
      ag = fmaf(g, xhat[m * ld + j], ag);
    }
  }
  sg[warp][lane] = ag;
  sb[warp][lane] = ab;
  __syncthreads();
  if (warp == 0 && j < d) {
    float tg = 0.f, tb = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) {
      tg += sg[w][lane];
      tb += sb[w][lane];
    }
    dgamma[j] = tg;
    dbeta[j] = tb;
  }
}

}  // namespace porl

using namespace porl;

constexpr int kMaxTcPhases = 3;   // = kMaxPhases of gemm_tc.cu

extern "C" int porl_mlp_create(porl_ctx_t* ctx, const porl_mlp_desc_t* desc, porl_mlp_t** out) {
  PORL_CHECK_ARG(ctx && desc && out, "porl_mlp_create: null argument");
  PORL_CHECK_ARG(desc->n_nets >= 1 && desc->n_nets <= kMaxGroups, "porl_mlp_create: n_nets=%d", desc->n_nets);
  PORL_CHECK_ARG(desc->n_layers >= 1 && desc->n_layers <= PORL_MAX_LAYERS, "porl_mlp_create: n_layers=%d",
                 desc->n_layers);
  PORL_CHECK_ARG(desc->out_act == 0 || desc->out_act == 1, "porl_mlp_create: out_act=%d", desc->out_act);
  for (int l = 0; l <= desc->n_layers; ++l)
    PORL_CHECK_ARG(desc->dims[l] >= 1 && desc->dims[l] <= 65536, "porl_mlp_create: dims[%d]=%d", l, desc->dims[l]);
  porl_mlp* m = new porl_mlp();
  m->ctx = ctx;
  m->d = *desc;
  int64_t off = 0;
  for (int l = 0; l <= desc->n_layers; ++l) m->pad[l] = pad32(desc->dims[l]);
  for (int l = 0; l < desc->n_layers; ++l) {
    m->w_off[l] = off;
    off += (int64_t)desc->dims[l + 1] * m->pad[l];
    m->b_off[l] = off;
    off += pad4(desc->dims[l + 1]);
    m->lnw_off[l] = m->lnb_off[l] = -1;
    if (desc->layer_norm && l < desc->n_layers - 1) {
      m->lnw_off[l] = off;
      off += pad4(desc->dims[l + 1]);
      m->lnb_off[l] = off;
      off += pad4(desc->dims[l + 1]);
    }
  }
  m->net_stride = off;
  m->ld_out = pad4(desc->dims[desc->n_layers]);
  *out = m;
  return PORL_OK;
}

extern "C" int porl_mlp_destroy(porl_mlp_t* mlp) {
  delete mlp;
  return PORL_OK;
}

extern "C" int64_t porl_mlp_param_floats(const porl_mlp_t* mlp) {
  return mlp ? mlp->net_stride * mlp->d.n_nets : -1;
}

extern "C" int porl_mlp_param_layout(const porl_mlp_t* mlp, int32_t net, int32_t layer, int32_t kind,
                                     int64_t* offset, int64_t* ld) {
  PORL_CHECK_ARG(mlp && offset && ld, "porl_mlp_param_layout: null argument");
  PORL_CHECK_ARG(net >= 0 && net < mlp->d.n_nets && layer >= 0 && layer < mlp->d.n_layers,
                 "porl_mlp_param_layout: net=%d layer=%d", net, layer);
  int64_t o = -1;
  switch (kind) {
    case PORL_PARAM_WEIGHT: o = mlp->w_off[layer]; *ld = mlp->pad[layer]; break;
    case PORL_PARAM_BIAS: o = mlp->b_off[layer]; *ld = 1; break;
    case PORL_PARAM_LN_WEIGHT: o = mlp->lnw_off[layer]; *ld = 1; break;
    case PORL_PARAM_LN_BIAS: o = mlp->lnb_off[layer]; *ld = 1; break;
    default: break;
  }
  PORL_CHECK_ARG(o >= 0, "porl_mlp_param_layout: tensor kind %d absent at layer %d", kind, layer);
  *offset = net * mlp->net_stride + o;
  return PORL_OK;
}

extern "C" int64_t porl_mlp_saved_floats(const porl_mlp_t* mlp, int64_t B) {
  if (!mlp || B < 0) return -1;
  return saved_total_floats(mlp, B) + 4;
}

extern "C" int64_t porl_mlp_out_ld(const porl_mlp_t* mlp) { return mlp ? mlp->ld_out : -1; }

extern "C" int porl_mlp_desc(const porl_mlp_t* mlp, porl_mlp_desc_t* out) {
  PORL_CHECK_ARG(mlp && out, "porl_mlp_desc: null argument");
  *out = mlp->d;
  return PORL_OK;
}


static void clear_group(GemmGroup& g) {
  g.A = g.B = nullptr;
  g.C = nullptr;
  g.A_lo = g.B_lo = nullptr;
  g.C_lo = nullptr;
  g.bias = g.relu_src = nullptr;
  g.colsum = nullptr;
  g.Ct = nullptr;
  g.bits_out = nullptr;
  g.bits_in = nullptr;
  g.ld_bits = 0;
  g.a_affine = nullptr;
  g.M = g.N = g.K = 0;
  g.lda = g.ldb = g.ldc = g.ld_relu = g.ldct = 0;
  g.n_store = 0;
  g.epilogue = 0;
}

extern "C" int porl_mlp_forward(porl_ctx_t* ctx, int32_t n_calls, const porl_mlp_call_t* calls, int64_t B,
                                porl_stream_t stream) {
  PORL_ON_DEVICE(ctx);
  PORL_CHECK_ARG(ctx && calls && n_calls >= 1 && n_calls <= PORL_MAX_CALLS, "porl_mlp_forward: n_calls=%d", n_calls);
  PORL_CHECK_ARG(B > 0 && B < (1 << 24), "porl_mlp_forward: B=%lld", (long long)B);
  int max_layers = 0, total_nets = 0;
  for (int c = 0; c < n_calls; ++c) {
    const porl_mlp_call_t& k = calls[c];
    PORL_CHECK_ARG(k.mlp && k.params && k.x && k.saved && k.out, "porl_mlp_forward: null pointer in call %d", c);
    PORL_CHECK_ARG(k.ldx >= k.mlp->d.dims[0] && (k.ldx & 3) == 0 && aligned16(k.x) && aligned16(k.params) &&
                       aligned16(k.saved) && aligned16(k.out),
                   "porl_mlp_forward: call %d needs 16-byte aligned buffers and ldx %% 4 == 0 (ldx=%lld)", c,
                   (long long)k.ldx);
    max_layers = k.mlp->d.n_layers > max_layers ? k.mlp->d.n_layers : max_layers;
    total_nets += k.mlp->d.n_nets;
  }
  PORL_CHECK_ARG(total_nets <= kMaxGroups, "porl_mlp_forward: %d nets exceed the %d-problem group limit", total_nets,
                 kMaxGroups);
  cudaStream_t st = as_stream(stream);
  int rc;

  // tensor-core problems, FFMA problems and narrow output layers of layer l over all calls
  auto build_layer = [&](int l, GemmBatch& tc, GemmBatch& ff, HeadFwdBatch& hf) {
    tc.n_groups = ff.n_groups = 0;
    hf.n_seg = 0;
    hf.B = B;
    for (int c = 0; c < n_calls; ++c) {
      const porl_mlp_call_t& k = calls[c];
      const porl_mlp* m = k.mlp;
      if (l >= m->d.n_layers) continue;
      const bool last = (l == m->d.n_layers - 1);
      const bool on_tc = tc_layer(ctx, m, l, B);
      if (last && narrow_head(m) && (k.pad_ & PORL_MLP_SKIP_HEAD)) continue;   // the fused head step evaluates it
      if (last && narrow_head(m)) {
        for (int net = 0; net < m->d.n_nets; ++net) {
          const float* P = k.params + net * m->net_stride;
          HeadFwdSeg& h = hf.seg[hf.n_seg++];
          h.h = (l == 0) ? k.x : saved_h(m, k.saved, net, l, B);
          h.ldh = (l == 0) ? (int)k.ldx : m->pad[l];
          h.W = P + m->w_off[l];
          h.ldw = m->pad[l];
          h.bias = P + m->b_off[l];
          h.out = k.out + (int64_t)net * B * m->ld_out;
          h.n_out = m->d.dims[l + 1];
          h.H = m->d.dims[l];
          h.ld_out = m->ld_out;
          h.act = m->d.out_act == 1 ? 2 : 0;
        }
        continue;
      }
      for (int net = 0; net < m->d.n_nets; ++net) {
        GemmBatch& batch = on_tc ? tc : ff;
        GemmGroup& g = batch.g[batch.n_groups++];
        clear_group(g);
        const float* P = k.params + net * m->net_stride;
        g.A = (l == 0) ? k.x : saved_h(m, k.saved, net, l, B);
        g.lda = (l == 0) ? (int)k.ldx : m->pad[l];
        g.B = P + m->w_off[l];
        g.ldb = m->pad[l];
        g.bias = P + m->b_off[l];
        g.M = (int)B;
        g.N = m->d.dims[l + 1];
        g.K = m->d.dims[l];
        if (last) {
          g.C = k.out + (int64_t)net * B * m->ld_out;
          g.ldc = m->ld_out;
          g.n_store = m->ld_out;
          g.epilogue = m->d.out_act == 1 ? 2 : 0;
        } else {
          g.C = saved_h(m, k.saved, net, l + 1, B);
          g.ldc = m->pad[l + 1];
          g.n_store = m->pad[l + 1];
          g.epilogue = m->d.layer_norm ? 0 : 1;
          // the input gradient of layer l+1 masks with relu'(h_{l+1}): keep it as one bit per element
          if (on_tc && !k.inference && g.epilogue == 1 && tc_layer(ctx, m, l + 1, B)) {
            g.bits_out = saved_bits(m, k.saved, net, l + 1, B);
            g.ld_bits = bits_ld(m, l + 1);
          }
        }
      }
    }
  };

  // All hidden layers on the tensor cores: ONE persistent launch runs them back to back, a grid barrier between
  // consecutive layers instead of a kernel boundary (gemm_tc.cu: TcPhase); the narrow heads follow.
  bool any_ln = false;
  for (int c = 0; c < n_calls; ++c) any_ln = any_ln || calls[c].mlp->d.layer_norm;
  if (ctx->gemm_mode == 1 && ctx->tc_phases && !any_ln && max_layers >= 2) {
    GemmBatch tcs[PORL_MAX_LAYERS], ff;
    HeadFwdBatch hfs[PORL_MAX_LAYERS];
    bool ok = true;
    int n_tc = 0;
    for (int l = 0; l < max_layers && ok; ++l) {
      build_layer(l, tcs[l], ff, hfs[l]);
      if (ff.n_groups) ok = false;
      if (tcs[l].n_groups) {
        if (n_tc != l) ok = false;          // tensor-core layers must be the leading ones
        ++n_tc;
      }
    }
    if (ok && n_tc >= 2) {
      rc = PORL_OK;
      for (int l0 = 0; l0 < n_tc && rc == PORL_OK; l0 += kMaxTcPhases) {
        TcPhaseSpec specs[kMaxTcPhases];
        const int n = n_tc - l0 < kMaxTcPhases ? n_tc - l0 : kMaxTcPhases;
        for (int p = 0; p < n; ++p) specs[p] = TcPhaseSpec{&tcs[l0 + p], true, true, p > 0};
        rc = launch_gemm_phases(ctx, n, specs, st);
        if (rc == PORL_ERR_UNSUPPORTED && l0 > 0) {
          for (int p = 0; p < n && (rc == PORL_OK || rc == PORL_ERR_UNSUPPORTED); ++p) rc = launch_gemm(ctx, true, true, tcs[l0 + p], st);
        }
      }
      if (rc == PORL_OK) {
        for (int l = 0; l < max_layers; ++l) {
          rc = launch_head_forward(ctx, hfs[l], st);
          if (rc != PORL_OK) return rc;
        }
        return PORL_OK;
      }
      if (rc != PORL_ERR_UNSUPPORTED) return rc;
    }
  }

  for (int l = 0; l < max_layers; ++l) {
    GemmBatch tc, ff;
    HeadFwdBatch hf;
    build_layer(l, tc, ff, hf);
    if (tc.n_groups) {
      rc = launch_gemm(ctx, true, true, tc, st);
      if (rc != PORL_OK) return rc;
    }
    if (ff.n_groups) {
      rc = launch_gemm(ctx, true, true, ff, st);
      if (rc != PORL_OK) return rc;
    }
    rc = launch_head_forward(ctx, hf, st);
    if (rc != PORL_OK) return rc;
    // LayerNorm + ReLU rows
    for (int c = 0; c < n_calls; ++c) {
      const porl_mlp_call_t& k = calls[c];
      const porl_mlp* m = k.mlp;
      if (!m->d.layer_norm || l >= m->d.n_layers - 1) continue;
      for (int net = 0; net < m->d.n_nets; ++net) {
        const float* P = k.params + net * m->net_stride;
        ProfScope prof(ctx, PORL_CAT_NORM, (double)B * m->d.dims[l + 1] * 12.0, st);
        ln_relu_forward_kernel<<<(unsigned)((B + 7) / 8), 256, 0, st>>>(
            saved_h(m, k.saved, net, l + 1, B), saved_xhat(m, k.saved, net, l + 1, B),
            saved_rstd(m, k.saved, net, l + 1, B), P + m->lnw_off[l], P + m->lnb_off[l], B,
            m->d.dims[l + 1], m->pad[l + 1]);
        PORL_LAUNCHED(ctx);
      }
    }
  }
  return PORL_OK;
}

extern "C" int porl_mlp_backward(porl_ctx_t* ctx, const porl_mlp_call_t* call, int64_t B, float* dout,
                                 float* grads, porl_stream_t stream) {
  PORL_ON_DEVICE(ctx);
  PORL_CHECK_ARG(ctx && call && call->mlp && grads && call->params && call->x && call->saved &&
                     (dout || (call->pad_ & PORL_MLP_HEAD_DONE)),
                 "porl_mlp_backward: null argument");
  PORL_CHECK_ARG(B > 0 && B < (1 << 24), "porl_mlp_backward: B=%lld", (long long)B);
  PORL_CHECK_ARG((!dout || aligned16(dout)) && aligned16(grads) && aligned16(call->x) && (call->ldx & 3) == 0,
                 "porl_mlp_backward: buffers must be 16-byte aligned, ldx %% 4 == 0");
  PORL_CHECK_ARG(!call->inference, "porl_mlp_backward: the forward of this call ran with inference=1");
  const porl_mlp* m = call->mlp;
  const int L = m->d.n_layers;
  const int nn = m->d.n_nets;
  cudaStream_t st = as_stream(stream);
  float* saved = call->saved;
  int rc;

  // No transposed copies: the tcgen05 kernel takes MN-major operands through MN-major shared-memory descriptors, so
  // the weight gradient dW_l = dz^T a_l reads dz (B x dims[l+1]) and a_l (B x dims[l]) as they lie (K = batch runs
  // over rows), and the input gradient dz W_l reads W_l (dims[l+1] x dims[l]) as its (K x N) operand.
  auto layer_norm_backward = [&](int l) -> int {   // h_l holds dL/d(LayerNorm output) -> dL/d(Linear output)
    for (int net = 0; net < nn; ++net) {
      const float* P = call->params + net * m->net_stride;
      float* G = grads + net * m->net_stride;
      float* dy = saved_h(m, saved, net, l, B);
      const float* xh = saved_xhat(m, saved, net, l, B);
      const int d = m->d.dims[l];
      ProfScope prof(ctx, PORL_CAT_NORM, (double)B * d * 20.0, st);
      ln_param_grad_kernel<<<(d + 31) / 32, 256, 0, st>>>(dy, xh, B, d, m->pad[l], G + m->lnw_off[l - 1],
                                                          G + m->lnb_off[l - 1]);
      PORL_LAUNCHED(ctx);
      ln_backward_kernel<<<(unsigned)((B + 7) / 8), 256, 0, st>>>(dy, xh, saved_rstd(m, saved, net, l, B),
                                                                  P + m->lnw_off[l - 1], B, d, m->pad[l]);
      PORL_LAUNCHED(ctx);
    }
    return PORL_OK;
  };
  int l_first = L - 1;
  bool bias_done[PORL_MAX_LAYERS] = {false, false, false, false, false, false};
  if (narrow_head(m) && (call->pad_ & PORL_MLP_HEAD_DONE)) {
    // porl_value_head_step has run: dW / db of the head and db of the layer below are in `grads`, the last hidden
    // activation holds dz
    PORL_CHECK_ARG(L >= 2 && !m->d.layer_norm, "porl_mlp_backward: PORL_MLP_HEAD_DONE needs a hidden layer and no LayerNorm");
    bias_done[L - 2] = true;
    l_first = L - 2;
  } else if (narrow_head(m)) {
    // output layer: weight / bias gradient, ReLU-masked input gradient and (without LayerNorm) the bias
    // gradient of the layer below in one pass over the last hidden activation
    const int l = L - 1;
    HeadBwdBatch hb;
    hb.n_seg = 0;
    hb.B = B;
    hb.n_out = m->d.dims[L];
    hb.H = m->d.dims[l];
    hb.ldh = (l == 0) ? (int)call->ldx : m->pad[l];
    hb.ldw = m->pad[l];
    hb.ld_out = m->ld_out;
    hb.need_dz = l > 0 ? 1 : 0;
    const bool give_bias = l > 0 && !m->d.layer_norm;
    for (int net = 0; net < nn; ++net) {
      HeadBwdSeg& s = hb.seg[hb.n_seg++];
      float* G = grads + net * m->net_stride;
      s.h = (l == 0) ? const_cast<float*>(call->x) : saved_h(m, saved, net, l, B);
      s.dout = dout + (int64_t)net * B * m->ld_out;
      s.W = call->params + net * m->net_stride + m->w_off[l];
      s.dW = G + m->w_off[l];
      s.db = G + m->b_off[l];
      s.db_prev = give_bias ? G + m->b_off[l - 1] : nullptr;
    }
    rc = launch_head_backward(ctx, hb, st);
    if (rc != PORL_OK) return rc;
    if (give_bias) bias_done[l - 1] = true;
    if (l > 0 && m->d.layer_norm) {
      rc = layer_norm_backward(l);
      if (rc != PORL_OK) return rc;
    }
    l_first = L - 2;
  }

  // gradient of layer l's pre-activation output: the head gradient, else the (in-place rewritten) hidden activation
  // buffer of layer l+1
  auto dz_of = [&](int l, int net) -> float* {
    return (l == L - 1) ? dout + (int64_t)net * B * m->ld_out : saved_h(m, saved, net, l + 1, B);
  };
  // dW_l = dz^T a_l (contraction over the batch); db_l = sum_b dz
  auto build_dw = [&](int l, GemmBatch& bw, ColsumBatch& cb) {
    const bool on_tc = tc_layer(ctx, m, l, B);
    const int ld_dz = (l == L - 1) ? m->ld_out : m->pad[l + 1];
    bw.n_groups = 0;
    cb.n_seg = 0;
    for (int net = 0; net < nn; ++net) {
      GemmGroup& g = bw.g[bw.n_groups++];
      clear_group(g);
      float* G = grads + net * m->net_stride;
      g.A = dz_of(l, net);                                         // (B x dims[l+1]): MN-major A, K = batch
      g.lda = ld_dz;
      g.B = (l == 0) ? call->x : saved_h(m, saved, net, l, B);     // (B x dims[l]): MN-major B
      g.ldb = (l == 0) ? (int)call->ldx : m->pad[l];
      if (!bias_done[l]) {
        if (on_tc) cb.seg[cb.n_seg++] = ColsumSeg{dz_of(l, net), G + m->b_off[l], B, m->d.dims[l + 1], ld_dz};
        else g.colsum = G + m->b_off[l];
      }
      g.C = G + m->w_off[l];
      g.ldc = m->pad[l];
      g.n_store = m->pad[l];
      g.M = m->d.dims[l + 1];
      g.N = m->d.dims[l];
      g.K = (int)B;
    }
  };
  // da_l = dz W_l, masked by ReLU (h_l > 0), written in place over h_l
  auto build_dx = [&](int l, GemmBatch& bx) {
    const bool on_tc = tc_layer(ctx, m, l, B), prev_tc = tc_layer(ctx, m, l - 1, B);
    const int ld_dz = (l == L - 1) ? m->ld_out : m->pad[l + 1];
    bx.n_groups = 0;
    for (int net = 0; net < nn; ++net) {
      GemmGroup& g = bx.g[bx.n_groups++];
      clear_group(g);
      const float* P = call->params + net * m->net_stride;
      float* h = saved_h(m, saved, net, l, B);
      g.A = dz_of(l, net);
      g.lda = ld_dz;
      g.B = P + m->w_off[l];                                       // W_l (dims[l+1] x dims[l]) = (K x N): MN-major B
      g.ldb = m->pad[l];
      g.C = h;
      g.ldc = m->pad[l];
      g.n_store = m->pad[l];
      g.relu_src = h;
      g.ld_relu = m->pad[l];
      if (on_tc && prev_tc) {                                      // h_l came from a tcgen05 forward: its relu bits exist
        g.bits_in = saved_bits(m, saved, net, l, B);
        g.ld_bits = bits_ld(m, l);
      }
      g.M = (int)B;
      g.N = m->d.dims[l];
      g.K = m->d.dims[l + 1];
    }
  };

  // Every remaining layer on the tensor cores (no LayerNorm): the contractions dW_l, dx_l, dW_{l-1}, ... run as phases
  // of persistent launches (gemm_tc.cu: TcPhase).  dx_l overwrites h_l, which dW_l reads, and dW_{l-1} / dx_{l-1} read
  // what dx_l wrote: a grid barrier in front of every phase but the first of a launch.  The bias-gradient column sums
  // read buffers no later phase touches and follow the contractions.
  bool all_tc = ctx->gemm_mode == 1 && ctx->tc_phases && !m->d.layer_norm && l_first >= 1;
  for (int l = l_first; l >= 0 && all_tc; --l) all_tc = tc_layer(ctx, m, l, B);
  if (all_tc) {
    GemmBatch gb[2 * PORL_MAX_LAYERS];
    ColsumBatch cbs[PORL_MAX_LAYERS];
    TcPhaseSpec all[2 * PORL_MAX_LAYERS];
    int n = 0;
    for (int l = l_first; l >= 0; --l) {
      build_dw(l, gb[n], cbs[l]);
      all[n] = TcPhaseSpec{&gb[n], false, false, true};
      ++n;
      if (l > 0) {
        build_dx(l, gb[n]);
        all[n] = TcPhaseSpec{&gb[n], true, false, true};
        ++n;
      }
    }
    rc = PORL_OK;
    for (int p0 = 0; p0 < n && rc == PORL_OK; p0 += kMaxTcPhases) {
      const int k = n - p0 < kMaxTcPhases ? n - p0 : kMaxTcPhases;
      rc = launch_gemm_phases(ctx, k, all + p0, st);
      if (rc == PORL_ERR_UNSUPPORTED) {        // e.g. a problem too small for the tensor-core kernel: one launch each
        rc = PORL_OK;
        for (int p = p0; p < p0 + k && rc == PORL_OK; ++p) rc = launch_gemm(ctx, all[p].a_kmajor, all[p].b_kmajor, *all[p].batch, st);
      }
    }
    if (rc != PORL_OK) return rc;
    for (int l = l_first; l >= 0; --l) {
      rc = launch_colsum(ctx, cbs[l], st);
      if (rc != PORL_OK) return rc;
    }
    return PORL_OK;
  }

  for (int l = l_first; l >= 0; --l) {
    GemmBatch bw;
    ColsumBatch cb;
    build_dw(l, bw, cb);
    rc = launch_gemm(ctx, false, false, bw, st);
    if (rc != PORL_OK) return rc;
    rc = launch_colsum(ctx, cb, st);
    if (rc != PORL_OK) return rc;
    if (l == 0) break;
    GemmBatch bx;
    build_dx(l, bx);
    rc = launch_gemm(ctx, true, false, bx, st);
    if (rc != PORL_OK) return rc;
    if (m->d.layer_norm) {
      rc = layer_norm_backward(l);
      if (rc != PORL_OK) return rc;
    }
  }
  return PORL_OK;
}

// Fused head step of the twin value nets (head.cu: vhead_fused_kernel): forward of the target heads on next_obs and of
// the online heads on obs, TD target, twin expectile loss and the heads' backward in one pass over the last hidden
// activations.  `target` / `online`: the calls whose porl_mlp_forward ran with PORL_MLP_SKIP_HEAD.
extern "C" int porl_value_head_step(porl_ctx_t* ctx, const porl_mlp_call_t* target, const porl_mlp_call_t* online, int64_t B,
                                    const float* rewards, const float* dones, float gamma, float tau, int64_t B_norm,
                                    float* target_v, float* grads, float* loss_out, porl_stream_t stream) {
  PORL_ON_DEVICE(ctx);
  PORL_CHECK_ARG(ctx && target && online && target->mlp && online->mlp == target->mlp && rewards && dones && target_v &&
                     grads && loss_out && B > 0 && B_norm > 0,
                 "porl_value_head_step: null argument");
  PORL_CHECK_ARG(porl_value_head_step_supported(online->mlp), "porl_value_head_step: needs twin nets with a scalar head, "
                 "at least one hidden layer of width <= 1024 and no LayerNorm");
  const porl_mlp* m = online->mlp;
  const int L = m->d.n_layers, l = L - 1;
  VHeadArgs a;
  for (int k = 0; k < 2; ++k) {
    const float* Pt = target->params + k * m->net_stride;
    const float* P = online->params + k * m->net_stride;
    float* G = grads + k * m->net_stride;
    a.ht[k] = saved_h(m, target->saved, k, l, B);
    a.wt[k] = Pt + m->w_off[l];
    a.bt[k] = Pt + m->b_off[l];
    a.h[k] = saved_h(m, online->saved, k, l, B);
    a.w[k] = P + m->w_off[l];
    a.b[k] = P + m->b_off[l];
    a.dW[k] = G + m->w_off[l];
    a.db[k] = G + m->b_off[l];
    a.db_prev[k] = G + m->b_off[l - 1];
  }
  a.rewards = rewards;
  a.dones = dones;
  a.target_v = target_v;
  a.loss_out = loss_out;
  a.gamma = gamma;
  a.tau = tau;
  a.inv_bn = 1.f / (float)B_norm;
  a.B = B;
  a.H = m->d.dims[l];
  a.ldh = m->pad[l];
  a.ldw = m->pad[l];
  return launch_vhead_fused(ctx, a, as_stream(stream));
}

// Fused policy head step (head.cu: phead_fused_kernel): the scalar heads of the twin value nets (new weights, forward only),
// the advantage-weighted Gaussian NLL of a narrow policy head and that head's backward in one pass over the last hidden
// activations.  `value` / `policy`: the calls whose porl_mlp_forward ran with PORL_MLP_SKIP_HEAD.
extern "C" int porl_policy_head_step_supported(const porl_mlp_t* value_mlp, const porl_mlp_t* policy_mlp) {
  if (!value_mlp || !policy_mlp) return 0;
  const int Lv = value_mlp->d.n_layers, Lp = policy_mlp->d.n_layers;
  if (!(value_mlp->d.n_nets == 2 && Lv >= 2 && value_mlp->d.dims[Lv] == 1 && !value_mlp->d.layer_norm &&
        value_mlp->d.out_act == 0))
    return 0;
  if (!(policy_mlp->d.n_nets == 1 && Lp >= 2 && !policy_mlp->d.layer_norm)) return 0;
  // one row loop over the three activations: same last hidden width and leading dimension
  if (value_mlp->d.dims[Lv - 1] != policy_mlp->d.dims[Lp - 1] || value_mlp->pad[Lv - 1] != policy_mlp->pad[Lp - 1]) return 0;
  return phead_fused_supported(policy_mlp->d.dims[Lp - 1], policy_mlp->d.dims[Lp]) ? 1 : 0;
}

extern "C" int porl_policy_head_step(porl_ctx_t* ctx, const porl_mlp_call_t* value, const porl_mlp_call_t* policy, int64_t B,
                                     const float* log_std, const float* reg_target, int64_t ld_target, const float* target_v,
                                     float alpha, int32_t mode, int64_t B_norm, float* grads, float* dlog_std, float* loss_out,
                                     porl_stream_t stream) {
  PORL_ON_DEVICE(ctx);
  PORL_CHECK_ARG(ctx && value && policy && value->mlp && policy->mlp && log_std && reg_target && target_v && grads &&
                     dlog_std && loss_out && B > 0 && B_norm >= B,
                 "porl_policy_head_step: null argument");
  PORL_CHECK_ARG(porl_policy_head_step_supported(value->mlp, policy->mlp),
                 "porl_policy_head_step: needs twin scalar value nets and a policy net with a head of at most 2 outputs, "
                 "equal last hidden widths <= 1024, no LayerNorm");
  PORL_CHECK_ARG((mode == PORL_AWR_POR || mode == PORL_AWR_SORL) && alpha != 0.f, "porl_policy_head_step: mode=%d alpha=%g",
                 mode, (double)alpha);
  const porl_mlp* mv = value->mlp;
  const porl_mlp* mp = policy->mlp;
  const int lv = mv->d.n_layers - 1, lp = mp->d.n_layers - 1;
  PHeadArgs a;
  for (int k = 0; k < 2; ++k) {
    const float* P = value->params + k * mv->net_stride;
    a.hv[k] = saved_h(mv, value->saved, k, lv, B);
    a.wv[k] = P + mv->w_off[lv];
    a.bv[k] = P + mv->b_off[lv];
  }
  a.h = saved_h(mp, policy->saved, 0, lp, B);
  a.w = policy->params + mp->w_off[lp];
  a.b = policy->params + mp->b_off[lp];
  a.log_std = log_std;
  a.x = reg_target;
  a.ld_x = ld_target;
  a.target_v = target_v;
  a.dW = grads + mp->w_off[lp];
  a.db = grads + mp->b_off[lp];
  a.db_prev = grads + mp->b_off[lp - 1];
  a.dlog_std = dlog_std;
  a.loss_out = loss_out;
  a.alpha = alpha;
  a.inv_bn = 1.f / (float)B_norm;
  a.mode = mode;
  a.tanh_out = mp->d.out_act == 1 ? 1 : 0;
  a.B = B;
  a.H = mp->d.dims[lp];
  a.D = mp->d.dims[lp + 1];
  a.ldh = mp->pad[lp];
  a.ldw = mp->pad[lp];
  return launch_phead_fused(ctx, a, as_stream(stream));
}

extern "C" int porl_value_head_step_supported(const porl_mlp_t* mlp) {
  if (!mlp) return 0;
  const int L = mlp->d.n_layers;
  return (mlp->d.n_nets == 2 && L >= 2 && mlp->d.dims[L] == 1 && !mlp->d.layer_norm && mlp->d.out_act == 0 &&
          vhead_fused_supported(mlp->d.dims[L - 1])) ? 1 : 0;
}
