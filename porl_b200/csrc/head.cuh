// Narrow-head (n_out <= 8) forward / backward launch interface (head.cu).
#pragma once
#include "common.cuh"

namespace porl {

constexpr int kHeadMax = 8;        // widest output layer served by the head kernels
constexpr int kHeadRowChunks = 32; // most row chunks (block partials) of the backward reduction

struct HeadFwdSeg {
  const float* h;      // (B, ldh) input activation, pad columns zero
  const float* W;      // (n_out, ldw), pad columns zero
  const float* bias;   // (n_out) or null
  float* out;          // (B, ld_out); columns n_out..ld_out-1 written as zeros
  int n_out, H, ldh, ldw, ld_out;
  int act;             // 0 none, 1 relu, 2 tanh
};
struct HeadFwdBatch {
  HeadFwdSeg seg[kMaxGroups];
  int n_seg;
  int64_t B;
};

struct HeadBwdSeg {
  float* h;            // in: (B, ldh) activation feeding the head; out: dz = relu-masked input gradient
  const float* dout;   // (B, ld_out) gradient w.r.t. the head's pre-activation output
  const float* W;      // (n_out, ldw)
  float* dW;           // (n_out, ldw) overwritten (pad columns zero)
  float* db;           // (n_out) overwritten
  float* db_prev;      // (H) = sum_b dz[b, :], or null
};
struct HeadBwdBatch {
  HeadBwdSeg seg[kMaxGroups];
  int n_seg;
  int64_t B;
  int n_out, H, ldh, ldw, ld_out;
  int need_dz;         // 0: only dW / db (single-layer net)
};

int launch_head_forward(porl_ctx* ctx, const HeadFwdBatch& hb, cudaStream_t st);
int launch_head_backward(porl_ctx* ctx, const HeadBwdBatch& hb, cudaStream_t st);

}  // namespace porl

// ---- fused scalar-V head step (head.cu): forward of the target and online heads, TD target + twin expectile loss and
// the head's backward in ONE pass over the last hidden activations (agent/por.py:80-87, agent/sorl.py:85-92).
namespace porl {
struct VHeadArgs {
  const float* ht[2];   // target nets: last hidden activation of next_obs (B, ldh)
  const float* wt[2];   // target head weight rows (H) and bias (1)
  const float* bt[2];
  float* h[2];          // online nets: in last hidden activation of obs, out dz (ReLU-masked input gradient), in place
  const float* w[2];
  const float* b[2];
  float* dW[2];         // (ldw) head weight gradient, pad columns zero
  float* db[2];         // (1)
  float* db_prev[2];    // (H) bias gradient of the layer below (column sums of dz)
  const float* rewards;
  const float* dones;
  float* target_v;      // (B) out
  float* loss_out;      // scalar out
  float gamma, tau, inv_bn;
  int64_t B;
  int H, ldh, ldw;
};
bool vhead_fused_supported(int H);
int launch_vhead_fused(porl_ctx* ctx, const VHeadArgs& a, cudaStream_t st);
}  // namespace porl

// ---- fused policy head step (head.cu): value heads (new weights, forward only), advantage-weighted Gaussian NLL of a
// narrow policy head (D <= 2 means) and that head's backward in ONE pass (agent/sorl.py:96-111).
namespace porl {
struct PHeadArgs {
  const float* hv[2];   // value nets: last hidden activation of obs with the NEW value weights (B, ldh), read only
  const float* wv[2];   // value head weight rows (H) and bias (1)
  const float* bv[2];
  float* h;             // policy net: in last hidden activation of obs, out dz (ReLU-masked input gradient), in place
  const float* w;       // (D, ldw) policy head weight
  const float* b;       // (D)
  const float* log_std; // (D) raw parameter
  const float* x;       // (B, ld_x) regression target (actions / next_obs)
  const float* target_v;  // (B) TD target of the value step
  float* dW;            // (D, ldw) overwritten, pad columns zero
  float* db;            // (D)
  float* db_prev;       // (H) bias gradient of the layer below
  float* dlog_std;      // (D)
  float* loss_out;      // scalar
  float alpha, inv_bn;
  int mode, tanh_out;   // PORL_AWR_*, tanh on the mean
  int64_t B, ld_x;
  int H, D, ldh, ldw;
};
bool phead_fused_supported(int H, int D);
int launch_phead_fused(porl_ctx* ctx, const PHeadArgs& a, cudaStream_t st);
}  // namespace porl

