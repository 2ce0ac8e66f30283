/* porl_b200.h -- C ABI of libporl_b200.so: the B200 (sm_100a) offline-RL update step of porl.
 *
 * The reference (hadleyhzy34/porl) has no FFI for this path: it is eager PyTorch called from
 * Python agents.  This header is therefore the boundary a maintainer binds with ctypes from the
 * reference's own classes.  Each entry point cites the reference code it
 * replaces, paths relative to the reference root.
 *
 * Conventions
 *   - every function returns an int status (PORL_OK or a negative PORL_ERR_*), never throws;
 *     porl_last_error() gives the thread-local message of the last failure;
 *   - all tensors are caller-owned DEVICE pointers (fp32 unless stated), row-major; `ld*` are
 *     leading dimensions in elements; pointers handed to the GEMM paths must be 16-byte aligned
 *     and their leading dimensions multiples of 4 (the host side pads widths to PORL_PAD);
 *   - work is enqueued on the caller's stream (a cudaStream_t passed as void*) and is
 *     asynchronous: no call synchronises the device;
 *   - there is NO CPU fallback: every call fails with PORL_ERR_CUDA when no sm_100 device exists.
 */
#ifndef PORL_B200_H
#define PORL_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PORL_VERSION 100          /* 0.1.0 */
#define PORL_PAD 32               /* feature widths are padded to this many floats in HBM */
#define PORL_MAX_LAYERS 6         /* linear layers per MLP */
#define PORL_MAX_CALLS 4          /* MLP evaluations merged into one grouped launch sequence */

enum {
  PORL_OK = 0,
  PORL_ERR_INVALID = -1,          /* bad argument (shape, alignment, null pointer) */
  PORL_ERR_CUDA = -2,             /* CUDA runtime / driver error, message in porl_last_error() */
  PORL_ERR_UNSUPPORTED = -3,      /* valid request this build cannot serve */
  PORL_ERR_WORKSPACE = -4         /* internal workspace too small for the request */
};

/* bits of the device status word (porl_ctx_status_ptr): set by kernels, read by the host when it
 * reads the losses back.  Replaces the reference's pdb traps (agent/por.py:104-105). */
enum {
  PORL_FLAG_NONFINITE_LOSS = 1,   /* a loss became NaN/Inf */
  PORL_FLAG_NONPOSITIVE_NLL = 2,  /* some per-sample NLL <= 0: where agent/por.py:104 drops into pdb */
  PORL_FLAG_BAD_INDEX = 4,        /* gather index outside [0, n_rows) (clamped) */
  PORL_FLAG_BAD_ACTION = 8,       /* discrete action outside [0, A) (clamped) */
  PORL_FLAG_SCAN_RANGE = 16,      /* scan value outside [0, 4.0001) u (8, inf): where state2costmap's scatter_ raises
                                     (util/costmap.py:29); the beam is skipped */
  PORL_FLAG_PEER_TIMEOUT = 32     /* a data-parallel peer did not signal within the wait budget (porl_adam_step_peer) */
};

typedef struct porl_ctx porl_ctx_t;
typedef struct porl_mlp porl_mlp_t;
typedef void* porl_stream_t;      /* cudaStream_t */

/* ---- context ------------------------------------------------------------------------------- */
int porl_version(void);
const char* porl_last_error(void);
int porl_ctx_create(porl_ctx_t** out, int device);
int porl_ctx_destroy(porl_ctx_t* ctx);
/* kernels launched through this context so far (bench.py's gpu_launches) */
int64_t porl_launch_count(const porl_ctx_t* ctx);
/* of those, launches of the tcgen05 3xTF32 contraction kernel (evidence that the tensor-core path ran) */
int64_t porl_tc_launch_count(const porl_ctx_t* ctx);
/* device address of the 32-bit status word (PORL_FLAG_*), for an async read-back with the losses */
int porl_ctx_status_ptr(porl_ctx_t* ctx, uint32_t** out);
/* 0 = fp32 FFMA contractions; 1 = tcgen05 3xTF32 contractions (operands split into tf32 hi/lo in shared
 * memory, TMEM accumulator restarted every chunk and summed in fp32 registers) where shapes allow */
int porl_ctx_set_gemm_mode(porl_ctx_t* ctx, int mode);
int porl_ctx_get_gemm_mode(const porl_ctx_t* ctx);
/* k-blocks (32 k each) accumulated inside the tensor core before the fp32 register add; default 1 */
int porl_ctx_set_tc_chunk(porl_ctx_t* ctx, int k_blocks);
/* 1 (env PORL_TC_PHASES=1): the tensor-core contractions of one MLP pass -- the hidden layers of a forward, or
 * {dW_l, dx_l, dW_l-1, ...} of a backward -- run as phases of ONE persistent launch with grid-wide barriers where a phase
 * reads what the previous one wrote, instead of one launch per contraction (13 instead of 19 launches per SORL update).
 * 0 (default): one launch each -- measured 3 % FASTER on B200 (1833 vs 1777 updates/s at batch 4096, 4229 vs 4095 at
 * batch 256): under programmatic dependent launch a kernel boundary already overlaps the next kernel's set-up with the
 * tail of the previous one, while the in-kernel barrier has to wait for the TMA stores to complete and refill the
 * pipeline all the same.  Results are bitwise identical either way (same tiles, same split-K cuts, same order). */
int porl_ctx_set_tc_phases(porl_ctx_t* ctx, int enable);
/* Measurement aid (bench.py's measured TF32 ceiling): 3 (default) = the 3xTF32 product mode; 1 = the same kernel
 * issuing only the hi*hi MMAs with the operand split skipped, i.e. a plain single-pass TF32 contraction whose
 * rate bounds what the 3-pass mode can reach (1/3 of it).  Results in mode 1 are TF32-accurate only (~1e-3). */
int porl_ctx_set_tc_terms(porl_ctx_t* ctx, int terms);
/* Programmatic dependent launch of the update step's kernels (default on; env PORL_PDL=0 turns it off): each kernel
 * lets its successor be scheduled while it runs and waits for its predecessor's completion before touching global
 * memory, so launch latency hides behind the previous kernel's tail.  Results are unchanged. */
int porl_ctx_set_pdl(porl_ctx_t* ctx, int enable);
/* tcgen05 kernel variant (env PORL_TC_A_IN_TMEM sets the initial value): 0 (default) = all four operand planes in
 * shared memory, four TMEM chunk accumulators; 1 = the A operand's tf32 hi/lo planes are staged in tensor memory
 * (tcgen05.st) and the MMAs read A from TMEM; 2 = CTA pairs (clusters of 2, tcgen05.mma.cta_group::2, M = 256: each
 * CTA stages half of the B tile), used for launches whose problems are all at least 256 rows high.  All three give
 * bitwise identical results; 0 is the fastest on B200. */
int porl_ctx_set_tc_variant(porl_ctx_t* ctx, int variant);
/* profiling aid: device buffer of 16 x 148 uint64 that receives per-CTA globaltimer stamps (kernel start,
 * setup done, first/last TMA, first split, first/last MMA, chunk drain, epilogue, end; see gemm_tc.cu) of
 * every tcgen05 contraction launch (NULL disables) */
int porl_ctx_set_tc_debug(porl_ctx_t* ctx, uint64_t* stamps_dev);

/* ---- per-kernel timing (bench.py's roofline leg) --------------------------------------------- */
/* While enabled, every launch is bracketed by CUDA events on its own stream and attributed to a
 * category together with its ALGORITHMIC work (flops for contractions, bytes for the HBM-bound
 * kernels).  porl_profile_read synchronises the recorded events and returns per-category sums. */
enum { PORL_CAT_CONTRACTION = 0, PORL_CAT_GATHER = 1, PORL_CAT_LOSS = 2, PORL_CAT_OPTIMIZER = 3,
       PORL_CAT_NORM = 4, PORL_CAT_TRANSPOSE = 5, PORL_CAT_HEAD = 6, PORL_CAT_ENCODER = 7 /* HBM-bound encoder passes */,
       PORL_CAT_BATCHNORM = 8 /* encoder BatchNorm statistics + apply (bytes) */, PORL_CAT_PCONV = 9 /* 3x3 partial conv (flops) */,
       PORL_CAT_COUNT = 10 };
typedef struct {
  int64_t launches[PORL_CAT_COUNT];
  double ms[PORL_CAT_COUNT];       /* sum of per-launch device durations */
  double work[PORL_CAT_COUNT];     /* flops (contraction) or bytes (others) */
} porl_profile_t;
int porl_profile_enable(porl_ctx_t* ctx, int enable);   /* enabling clears earlier records */
int porl_profile_read(porl_ctx_t* ctx, porl_profile_t* out);

/* ---- K1: minibatch gather from the HBM-resident dataset ----------------------------------- */
/* out[b, 0:row_len] = table[idx[b], 0:row_len]; bit-exact.
 * Replaces DataLoader collate over CustomDataset.__getitem__ (dataloader/dataloader.py:34-52,
 * por_train.py:59-63,71-72). */
int porl_gather_rows_f32(porl_ctx_t* ctx, const float* table, int64_t n_rows, int64_t row_stride,
                         int64_t row_len, const int64_t* idx, int64_t B, float* out,
                         int64_t out_stride, porl_stream_t stream);
/* Gather AoS rows [obs(S) | r | next_obs(S) | done | action(A)] straight into the padded SoA
 * minibatch the update consumes: obs/next_obs (B, ld_obs) with columns S..ld_obs-1 zeroed,
 * rewards (B), dones (B), actions (B, ld_act).  Fuses the gather with the column slicing of
 * por_train.py:74-78 / sorl_train.py:54-58; payload bit-exact. */
int porl_gather_transitions(porl_ctx_t* ctx, const float* table, int64_t n_rows, int64_t row_stride,
                            int32_t S, int32_t A, const int64_t* idx, int64_t B, float* obs,
                            float* next_obs, int64_t ld_obs, float* rewards, float* dones,
                            float* actions, int64_t ld_act, porl_stream_t stream);
/* SoA replay-buffer gather: the five fancy-index gathers of ReplayBuffer.sample
 * (src/porl/buffer/replaybuffer.py:64-70).  states/next_states (cap, ld_state) -> (B, ld_out)
 * with columns S..ld_out-1 zeroed; actions int64; bit-exact. */
int porl_gather_soa(porl_ctx_t* ctx, const float* states, const float* next_states,
                    const int64_t* actions, const float* rewards, const float* dones, int64_t cap,
                    int32_t S, int64_t ld_state, const int64_t* idx, int64_t B, float* out_states,
                    float* out_next_states, int64_t ld_out, int64_t* out_actions,
                    float* out_rewards, float* out_dones, porl_stream_t stream);
/* (B, n) strided rows -> (B, ld_out) padded rows, pad columns zeroed (packs tensors a caller
 * passes to agent.update() into the layout the contractions need). */
int porl_pack_rows(porl_ctx_t* ctx, const float* src, int64_t ld_src, int64_t B, int32_t n,
                   float* dst, int64_t ld_dst, porl_stream_t stream);

/* Device-side shuffled indices: out[i] = P(first + i mod n), i < count, where P is a pseudo-random
 * bijection of [0, n) keyed by (seed, stream_id) -- an alternating Feistel network with cycle walking, O(1) per
 * index, no table.  Consecutive windows of one key are the minibatches of one shuffled epoch (distinct rows);
 * replaces the host-side torch.randperm of DataLoader(shuffle=True) (por_train.py:61) and the O(size)
 * np.random.choice(size, B, replace=False) of ReplayBuffer.sample (src/porl/buffer/replaybuffer.py:61) on the
 * throughput path.  Same distribution for SGD purposes, not the same index stream. */
int porl_sample_permutation(porl_ctx_t* ctx, int64_t n, uint64_t seed, uint64_t stream_id, int64_t first,
                            int64_t count, int64_t* out, porl_stream_t stream);

/* ---- K2/K3: fused MLP forward / backward --------------------------------------------------- */
/* util/util.py:29-47 mlp(): Linear[+LayerNorm]+ReLU per hidden layer, Linear[+Tanh] head.
 * n_nets > 1 evaluates structurally identical nets side by side on one input (TwinV.both,
 * agent/value_functions.py:31-39). */
typedef struct {
  int32_t n_nets;
  int32_t n_layers;                       /* linear layers */
  int32_t dims[PORL_MAX_LAYERS + 1];      /* dims[0] = input width ... dims[n_layers] = output width */
  int32_t layer_norm;                     /* LayerNorm after every hidden Linear (eps 1e-5) */
  int32_t out_act;                        /* 0 none, 1 tanh (BoundedGaussianPolicy, agent/policy.py:38-39) */
} porl_mlp_desc_t;

enum { PORL_PARAM_WEIGHT = 0, PORL_PARAM_BIAS = 1, PORL_PARAM_LN_WEIGHT = 2, PORL_PARAM_LN_BIAS = 3 };

int porl_mlp_create(porl_ctx_t* ctx, const porl_mlp_desc_t* desc, porl_mlp_t** out);
int porl_mlp_destroy(porl_mlp_t* mlp);
/* floats of the flat parameter (and gradient, and Adam moment) buffer of all n_nets nets */
int64_t porl_mlp_param_floats(const porl_mlp_t* mlp);
/* where tensor (net, layer, kind) lives in the flat buffer: offset in floats and, for weights,
 * the leading dimension (input width padded to PORL_PAD; pad columns stay zero). */
int porl_mlp_param_layout(const porl_mlp_t* mlp, int32_t net, int32_t layer, int32_t kind,
                          int64_t* offset, int64_t* ld);
/* floats of the saved-activation buffer a training forward at batch B needs */
int64_t porl_mlp_saved_floats(const porl_mlp_t* mlp, int64_t B);
/* padded leading dimension of the output rows, per net: out is (n_nets, B, ld_out) */
int64_t porl_mlp_out_ld(const porl_mlp_t* mlp);
/* copy of the descriptor the plan was created from */
int porl_mlp_desc(const porl_mlp_t* mlp, porl_mlp_desc_t* out);

typedef struct {
  porl_mlp_t* mlp;
  const float* params;    /* flat parameters */
  const float* x;         /* (B, ldx) input, pad columns zero */
  int64_t ldx;
  float* saved;           /* porl_mlp_saved_floats(B) floats; a training forward keeps the activations
                             here for porl_mlp_backward, an inference forward uses it as scratch */
  float* out;             /* (n_nets, B, ld_out) */
  int32_t inference;      /* 1: no backward follows; 0: training forward (keeps the ReLU bits the backward reads) */
  int32_t pad_;           /* flags: PORL_MLP_SKIP_HEAD | PORL_MLP_HEAD_DONE, else 0 */
} porl_mlp_call_t;
/* porl_mlp_forward stops before a scalar / narrow output layer (porl_value_head_step evaluates it) */
#define PORL_MLP_SKIP_HEAD 1
/* porl_mlp_backward: the output layer is already back-propagated (porl_value_head_step wrote its dW / db, the bias
 * gradient of the layer below and, over the last hidden activation, dz); dout is ignored and may be NULL */
#define PORL_MLP_HEAD_DONE 2

/* Forward of up to PORL_MAX_CALLS MLPs on the same batch size; layer l of every call runs in one
 * grouped launch.  Replaces TwinV.both/forward, GaussianPolicy.net, QNetwork.forward. */
int porl_mlp_forward(porl_ctx_t* ctx, int32_t n_calls, const porl_mlp_call_t* calls, int64_t B,
                     porl_stream_t stream);
/* Backward of one MLP: dout (n_nets, B, ld_out) is the loss gradient w.r.t. the PRE-activation head
 * output (the loss kernels fold tanh'), grads gets dW/db in the flat layout (every element
 * written, no accumulation).  dout and `saved` are consumed (overwritten).  Replaces autograd of
 * the nets (agent/por.py:89,108; agent/sorl.py:94,111; src/porl/train/cql_trainer.py:112). */
int porl_mlp_backward(porl_ctx_t* ctx, const porl_mlp_call_t* call, int64_t B, float* dout,
                      float* grads, porl_stream_t stream);

/* ---- K4: TD target + twin expectile loss (agent/por.py:15-17,85-87; agent/sorl.py:90-92) --- */
/* target = r + (1-done)*gamma*min(nv1,nv2); loss = sum_k sum_i |tau-1[u<0]|u^2 / (2*B_norm),
 * u = target - v_k; dv_k = -|tau-1[u<0]| u / B_norm.  B_norm is the GLOBAL batch (B * world).
 * loss_out[0] is overwritten. */
int porl_expectile_twin_loss(porl_ctx_t* ctx, const float* v1, const float* v2, const float* nv1,
                             const float* nv2, int64_t ldv, const float* rewards, const float* dones,
                             float gamma, float tau, int64_t B, int64_t B_norm, float* target_v,
                             float* dv1, float* dv2, float* loss_out, porl_stream_t stream);

/* No-allocation variant of porl_expectile_twin_loss.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  Same arithmetic, same kernel and the same order-fixed reduction as porl_expectile_twin_loss (results
 * are bitwise equal), but it takes no context and calls no cudaMalloc / cudaFree / allocating routine: the reduction
 * scratch (one partial per 256 rows + ticket and status words, 16.25 KB) is a file-scope static __device__ array that
 * every GPU loading the library holds until exit.  Limits: B <= PORL_NOALLOC_MAX_ROWS (larger B is rejected with
 * PORL_ERR_INVALID and a message in porl_last_error).  All calls on a device share the array, which keeps the last
 * call's contents (the ticket / status words are cleared by a cudaMemsetAsync on `stream` before each launch, the
 * partials are rewritten): calls that OVERLAP on the GPU (different streams without ordering) race; calls on one
 * stream are safe.  The array's device address is resolved with cudaGetSymbolAddress on the first call per device
 * (make that call outside graph capture, with the device current).  A non-finite loss shows in loss_out[0] only (there
 * is no context whose status flags could record it). */
#define PORL_NOALLOC_MAX_ROWS (1 << 20)
int porl_expectile_twin_loss_noalloc(const float* v1, const float* v2, const float* nv1, const float* nv2, int64_t ldv,
                                     const float* rewards, const float* dones, float gamma, float tau, int64_t B,
                                     int64_t B_norm, float* target_v, float* dv1, float* dv2, float* loss_out,
                                     porl_stream_t stream);

/* Fused head step of the twin value nets (agent/por.py:80-87, agent/sorl.py:85-92): the scalar heads of the target nets
 * on next_obs and of the online nets on obs, the TD target, the twin expectile loss (as porl_expectile_twin_loss) and the
 * heads' backward (dW, db, ReLU-masked dz in place, db of the layer below) in ONE pass over the last hidden activations.
 * `target` / `online` are the calls a porl_mlp_forward with pad_ = PORL_MLP_SKIP_HEAD has just evaluated (same plan);
 * follow with porl_mlp_backward(online with pad_ = PORL_MLP_HEAD_DONE, dout = NULL).  target_v (B) and loss_out[0] are
 * overwritten.  porl_value_head_step_supported: twin nets, scalar head, last hidden width <= 1024, no LayerNorm. */
int porl_value_head_step_supported(const porl_mlp_t* mlp);
int porl_value_head_step(porl_ctx_t* ctx, const porl_mlp_call_t* target, const porl_mlp_call_t* online, int64_t B,
                         const float* rewards, const float* dones, float gamma, float tau, int64_t B_norm,
                         float* target_v, float* grads, float* loss_out, porl_stream_t stream);

/* Fused policy head step (agent/sorl.py:96-111; agent/por.py:97-108 when the policy head is narrow): the scalar heads of
 * the twin value nets on obs with the NEW value weights (forward only), the advantage weight, the weighted Gaussian
 * NLL of the policy head's means (as porl_awr_gauss_nll, tanh folded when the policy net has a tanh output) and the
 * head's backward (dW, db, ReLU-masked dz in place, db of the layer below, dlog_std) in ONE pass over the last hidden
 * activations.  `value` / `policy` are the calls a porl_mlp_forward with pad_ = PORL_MLP_SKIP_HEAD has just evaluated;
 * follow with porl_mlp_backward(policy with pad_ = PORL_MLP_HEAD_DONE, dout = NULL).  grads is the policy net's
 * gradient buffer (flat layout), dlog_std (D) and loss_out[0] are overwritten.  Supported: twin scalar value nets and a
 * one-net policy with at most 2 outputs, equal last hidden widths <= 1024, no LayerNorm. */
int porl_policy_head_step_supported(const porl_mlp_t* value_mlp, const porl_mlp_t* policy_mlp);
int porl_policy_head_step(porl_ctx_t* ctx, const porl_mlp_call_t* value, const porl_mlp_call_t* policy, int64_t B,
                          const float* log_std, const float* reg_target, int64_t ld_target, const float* target_v,
                          float alpha, int32_t mode, int64_t B_norm, float* grads, float* dlog_std, float* loss_out,
                          porl_stream_t stream);

/* ---- K5: advantage-weighted diagonal-Gaussian NLL (agent/por.py:97-106; agent/sorl.py:102-109;
 *          agent/policy.py:18-23,42-49) ------------------------------------------------------- */
enum { PORL_AWR_POR = 0 /* w = exp(adv/alpha) */, PORL_AWR_SORL = 1 /* w = exp(alpha*adv) */ };
/* v = min(v1n, v2n); w = min(exp(..), 100); nll_i = 1/2 sum_j z_ij^2 + sum_j ls_j + D/2 log 2pi,
 * z = (x-mean)/exp(ls), ls = clamp(log_std,-5,2); loss = sum_i w_i nll_i / B_norm.
 * dz_out (B, ld_mean): gradient w.r.t. the pre-activation mean (times 1-mean^2 when tanh_out);
 * dlog_std (D) overwritten; loss_out[0] overwritten. */
int porl_awr_gauss_nll(porl_ctx_t* ctx, const float* mean, int64_t ld_mean, const float* log_std,
                       const float* x, int64_t ld_x, const float* target_v, const float* v1n,
                       const float* v2n, int64_t ldv, float alpha, int32_t mode, int32_t tanh_out,
                       int64_t B, int32_t D, int64_t B_norm, float* dz_out, float* dlog_std,
                       float* loss_out, porl_stream_t stream);

/* ---- K6: discrete CQL head (src/porl/train/cql_trainer.py:72-86,94-108) -------------------- */
/* t = r + gamma*max_a qn*(1-done); td = mean (q[a]-t)^2; pen = mean(logsumexp(q) - log A - q[a]);
 * loss = td + alpha*pen; dq = (2/Bn)(q[a]-t)1[a] + (alpha/Bn)(softmax(q) - 1[a]).
 * losses_out[0..2] = loss, td, penalty (overwritten). */
int porl_cql_head(porl_ctx_t* ctx, const float* q, const float* q_next_tgt, int64_t ldq,
                  const int64_t* actions, const float* rewards, const float* dones, float gamma,
                  float alpha, int64_t B, int32_t A, int64_t B_norm, float* dq, float* losses_out,
                  porl_stream_t stream);

/* ---- Fused CQL gradient (SURVEY.md 8b porl_cql_step; src/porl/train/cql_trainer.py:88-124) -- */
/* One persistent kernel per update for the QNetwork S -> 64 -> 128 -> 64 -> A of src/porl/net/q_network.py:8-30:
 * gathers rows idx[0..B) of the SoA replay arrays (replaybuffer.py:61-70; idx == NULL takes rows 0..B-1 of an already
 * gathered batch), evaluates the target net on next_states and the online net on states with the weights staged in
 * shared memory, applies the TD + conservative head of porl_cql_head and back-propagates; a second kernel sums the
 * per-CTA gradient partials in CTA order (deterministic) into `grads` (flat parameter layout of porl_mlp_param_layout)
 * and writes losses_out[0..2] = loss, td, penalty.  The caller follows with porl_adam_step on the same bucket.
 * states / next_states: (n_rows, ld_states) rows padded to PORL_PAD floats with zero pad columns.
 * PORL_ERR_UNSUPPORTED for any other network shape (porl_cql_fused_supported() == 0): use porl_mlp_* + porl_cql_head. */
int porl_cql_fused_supported(const porl_mlp_t* mlp);
int porl_cql_fused_grad(porl_ctx_t* ctx, const porl_mlp_t* mlp, const float* params, const float* target_params,
                        const float* states, const float* next_states, int64_t ld_states, const int64_t* actions,
                        const float* rewards, const float* dones, const int64_t* idx, int64_t n_rows, float gamma,
                        float alpha, int64_t B, int64_t B_norm, float* grads, float* losses_out,
                        porl_stream_t stream);

/* ---- K7: fused Adam (+ Polyak / hard target update) ---------------------------------------- */
/* Device-resident optimizer state so that a captured CUDA graph can be replayed: the step
 * counter and the (optionally cosine-annealed) learning rate advance on the device. */
typedef struct {
  int64_t step;          /* updates applied so far (t of the next step is step+1) */
  double lr;             /* learning rate the NEXT step uses */
  double base_lr;
  int64_t t_max;         /* > 0: torch CosineAnnealingLR(T_max=t_max, eta_min=0), recursive form */
  uint32_t blocks_done;  /* internal */
  uint32_t pad_;
  uint32_t pad2_[2];
  uint32_t rec_[2][4];   /* internal: two published 16-byte records of Adam scalars (parity of t); zero-initialise.
                          * The buffer must be 16-byte aligned. */
} porl_opt_state_t;

/* torch.optim.Adam.step (beta1 .9, beta2 .999, eps 1e-8 defaults passed explicitly; single-tensor
 * formula of torch/optim/adam.py) over a flat buffer of n floats, then optionally
 *   ema_beta > 0 : target = (1-ema_beta)*target + ema_beta*p   (util/util.py:54-56), or
 *   hard_every > 0 : target = p when (step % hard_every == 0) after the step
 *                    (src/porl/train/dqn_trainer.py:195-196),
 * then advances *state (step += 1, cosine LR recursion: agent/por.py:65,110). */
int porl_adam_step(porl_ctx_t* ctx, float* p, const float* g, float* m, float* v, int64_t n,
                   porl_opt_state_t* state_dev, float beta1, float beta2, float eps, float* target,
                   float ema_beta, int32_t hard_every, porl_stream_t stream);

/* The whole CQL / DQN gradient update in two launches (src/porl/train/cql_trainer.py:88-124: loss, backward,
 * optimizer.step, and the hard target copy of dqn_trainer.py:196 every `hard_every` steps): porl_cql_fused_grad with
 * the Adam step of porl_adam_step folded into the gradient reduction -- the thread that finishes a float4 of the gradient
 * updates those four parameters, the last block advances the device-side optimizer clock.  Bitwise equal to
 * porl_cql_fused_grad followed by porl_adam_step (same sums, same per-element update).  Single process only (B_norm = B):
 * under data parallelism use porl_cql_fused_grad + porl_adam_step_peer.  grads still receives the gradient. */
int porl_cql_fused_step(porl_ctx_t* ctx, const porl_mlp_t* mlp, float* params, float* target_params,
                        const float* states, const float* next_states, int64_t ld_states, const int64_t* actions,
                        const float* rewards, const float* dones, const int64_t* idx, int64_t n_rows, float gamma,
                        float alpha, int64_t B, float* grads, float* losses_out, float* m, float* v,
                        porl_opt_state_t* state_dev, float beta1, float beta2, float eps, int32_t hard_every,
                        porl_stream_t stream);

/* ---- data-parallel step: gradient sum over NVLink peer memory fused into the Adam kernel ----
 * (the reference has no distributed code.)  Every rank keeps its gradient bucket
 * (n gradients | tail_floats loss scalars) in memory the other ranks of the node have mapped (CUDA IPC,
 * NVLink P2P).  porl_adam_step_peer replaces   all_reduce(bucket); porl_adam_step(...)   by ONE kernel:
 *   1. every rank tells every peer "my bucket of epoch e is complete" (st.release.sys into the peers' pads)
 *      and waits for the same word from all of them,
 *   2. each thread loads its gradient float4 from ALL ranks' buckets (own one included) and adds them in
 *      rank order 0..world-1 -- the same order on every rank, so the replicas stay bitwise identical --
 *      and runs the Adam / Polyak update of porl_adam_step on the sum; block 0 sums the tails into tail_out,
 *   3. the last block tells the peers "I have read your bucket" and waits for theirs, so the kernel only
 *      completes once nobody still reads this rank's bucket (the next backward may overwrite it).
 * Traffic per rank: (world-1) * (n + tail) * 4 bytes over NVLink, no NCCL launch, no extra pass over HBM.
 * A peer that never signals makes the kernel give up after wait_ms, set PORL_FLAG_PEER_TIMEOUT and finish. */
#define PORL_MAX_PEERS 8
#define PORL_PEER_PAD_WORDS (2 * PORL_MAX_PEERS)   /* ready[PORL_MAX_PEERS] | done[PORL_MAX_PEERS], zero-initialised */
typedef struct porl_peer_group {
  int32_t world, rank;
  const float* bucket[PORL_MAX_PEERS];   /* bucket[r]: rank r's bucket as mapped in THIS process (bucket[rank] local) */
  uint32_t* pad[PORL_MAX_PEERS];         /* pad[r]: rank r's signal pad, PORL_PEER_PAD_WORDS words */
  uint32_t* epoch;                       /* local device word, zero-initialised, owned by this group */
  float* tail_out;                       /* local: sum over ranks of the tails (tail_floats values) */
  int32_t tail_floats;
  int32_t wait_ms;                       /* give-up budget of each wait (0: 20000) */
  /* Two-shot (sharded) mode, selected when params[0] != NULL: rank r sums shard r of the bucket over all ranks, runs
   * Adam / Polyak on that shard only and STORES the new parameters (and target-net values) into every rank's buffers.
   * Link traffic per rank and step: (world-1)/world * n floats read + as much (twice with a target net) written,
   * instead of (world-1) * n read; Adam's HBM traffic drops to 1/world; each rank keeps the Adam moments of ITS shard
   * only (gather them for a checkpoint).  Replicas are bitwise identical by construction.  All NULL: one-shot mode. */
  float* params[PORL_MAX_PEERS];         /* params[r]: rank r's flat parameter buffer as mapped here (params[rank] = p) */
  float* target[PORL_MAX_PEERS];         /* same for the target net; NULL entries when the group has none */
} porl_peer_group_t;

/* device memory other processes of the node can map: cudaMalloc'ed (not from a pooled allocator), zero-filled */
int porl_peer_alloc(porl_ctx_t* ctx, size_t bytes, void** out);
int porl_peer_free(porl_ctx_t* ctx, void* ptr);
/* handle: 64 opaque bytes (cudaIpcMemHandle_t) to send to the peers through any host channel */
int porl_peer_export(porl_ctx_t* ctx, void* ptr, uint8_t handle[64]);
/* maps a peer's allocation into this process (enables P2P access to its device); PORL_ERR_UNSUPPORTED when the
 * two devices cannot reach each other */
int porl_peer_open(porl_ctx_t* ctx, const uint8_t handle[64], void** out);
int porl_peer_close(porl_ctx_t* ctx, void* ptr);
int porl_adam_step_peer(porl_ctx_t* ctx, float* p, const porl_peer_group_t* peers, float* m, float* v, int64_t n,
                        porl_opt_state_t* state_dev, float beta1, float beta2, float eps, float* target,
                        float ema_beta, int32_t hard_every, porl_stream_t stream);

/* Test aid: `world` ranks of the two-shot step on ONE GPU as a single cooperative launch (blockIdx.y = rank), so the
 * blocks that wait for one another are co-resident by construction.  ranks[r].peers->rank must be r. */
typedef struct {
  float* p;
  float* m;
  float* v;
  float* target;                         /* or NULL */
  porl_opt_state_t* state;
  const porl_peer_group_t* peers;
} porl_peer_rank_t;
int porl_adam_step_peer_emulated(porl_ctx_t* ctx, int32_t world, const porl_peer_rank_t* ranks, int64_t n, float beta1,
                                 float beta2, float eps, float ema_beta, int32_t hard_every, porl_stream_t stream);

/* ---- whole-update entry points (porl_sorl_step / porl_por_step) ----------------
 * One POR / SORL gradient update as ONE call: the launch sequence of agent/por.py:73-112 and
 * agent/sorl.py:78-128 (target + online TwinV forward, TD target and twin expectile loss, value backward,
 * Adam + Polyak, online TwinV with the new weights + policy forward, advantage-weighted Gaussian NLL, policy
 * backward, Adam with the cosine LR recursion) enqueued on `stream`; nothing synchronises, so the call can be
 * captured into a CUDA graph.  All buffers are caller-owned device memory. */
typedef struct {
  float* params;               /* flat parameters of the optimizer group (policy group: MLP parameters | log_std) */
  float* grads;                /* gradient bucket, same layout (every element rewritten each update) */
  float* m;                    /* Adam exp_avg */
  float* v;                    /* Adam exp_avg_sq */
  int64_t n;                   /* floats in the group, a multiple of 4 */
  porl_opt_state_t* state;     /* device-side optimizer clock */
  float* loss;                 /* device scalar the loss kernel overwrites (data parallel: the bucket's tail, so the
                                  loss rides the gradient sum) */
  float beta1, beta2, eps;     /* torch.optim.Adam defaults .9, .999, 1e-8 (agent/por.py:61,64) */
  int32_t pad_;
  const porl_peer_group_t* peers;  /* data parallel: gradient sum inside the Adam kernel (porl_adam_step_peer), or NULL */
} porl_opt_group_t;

typedef struct {
  porl_mlp_t* value_mlp;       /* TwinV (agent/value_functions.py:31-42): n_nets = 2, F -> H.. -> 1 */
  porl_mlp_t* policy_mlp;      /* GaussianPolicy / BoundedGaussianPolicy mean net (agent/policy.py:12-49): F -> H.. -> D */
  porl_opt_group_t value;      /* v_optimizer = Adam(vf.parameters(), value_lr) */
  porl_opt_group_t policy;     /* Adam(policy.parameters(), policy_lr) + CosineAnnealingLR(max_steps) */
  float* target_params;        /* v_target: deepcopy(vf), updated by the Polyak step only */
  int64_t log_std_offset;      /* floats from policy.params to log_std (D values) */
  float discount, tau, alpha, ema_beta;   /* gamma, expectile, AWR temperature, Polyak beta (util/util.py:54-56) */
  int32_t awr_mode;            /* PORL_AWR_POR / PORL_AWR_SORL */
  int32_t tanh_out;            /* 1: BoundedGaussianPolicy (SORL) */
  int32_t policy_out;          /* D: state_size for POR's guide policy, action_size for SORL */
  int32_t world;               /* data-parallel ranks sharing the global batch (losses are scaled by 1 / (B * world)) */
} porl_agent_step_t;

/* floats of scratch one update at batch B needs (saved activations of the four MLP evaluations, head outputs,
 * loss gradients, TD targets); the same workspace serves every update at that batch size */
int64_t porl_step_workspace_floats(const porl_agent_step_t* agent, int64_t B);
/* SORL.update (agent/sorl.py:78-128) / SORL.vf_update (:130-152, value_only = 1) on a packed minibatch:
 * x_obs / x_next (B, ld_x) are what the heads read (the padded states, or the scan encoder's features of obs and
 * next_obs), actions (B, ld_act) the behaviour actions the policy regresses on. */
int porl_sorl_step(porl_ctx_t* ctx, const porl_agent_step_t* agent, int64_t B, const float* x_obs,
                   const float* x_next, int64_t ld_x, const float* actions, int64_t ld_act, const float* rewards,
                   const float* dones, float* workspace, int64_t workspace_floats, int32_t value_only,
                   porl_stream_t stream);
/* POR.por_residual_update (agent/por.py:73-112): the guide policy regresses on the RAW next observations
 * next_obs (B, ld_next) (after state2costmap's in-place clamp when an encoder is used), whatever the heads read. */
int porl_por_step(porl_ctx_t* ctx, const porl_agent_step_t* agent, int64_t B, const float* x_obs,
                  const float* x_next, int64_t ld_x, const float* next_obs, int64_t ld_next, const float* rewards,
                  const float* dones, float* workspace, int64_t workspace_floats, porl_stream_t stream);

/* ---- low-level contraction (exposed for tests and profiling) ------------------------------- */
/* C[M,N] = opA(A)[M,K] * opB(B)[K,N] (+ bias[n]) with epilogue; trans flags:
 *   a_kmajor: A(m,k) = A[m*lda+k] else A[k*lda+m];  b_kmajor: B(k,n) = B[n*ldb+k] else B[k*ldb+n].
 * epilogue: 0 none, 1 relu, 2 tanh; relu_src (M, ld_relu): zero C where relu_src <= 0.
 * n_store >= N: columns N..n_store-1 of C are written as zeros.  colsum (M): sum_k A(m,k), or NULL. */
/* Ct (N, ldct): optional transposed copy of C (tcgen05 path only: mode 1, both operands K-major).
 * On the tcgen05 path `bias` is read in 16-byte groups: it must be 16-byte aligned and readable (zero)
 * through round_up(N, 4) floats. */
int porl_gemm(porl_ctx_t* ctx, int32_t a_kmajor, int32_t b_kmajor, int64_t M, int64_t N, int64_t K,
              const float* A, int64_t lda, const float* B, int64_t ldb, float* C, int64_t ldc,
              const float* bias, int32_t epilogue, const float* relu_src, int64_t ld_relu,
              int64_t n_store, float* colsum, float* Ct, int64_t ldct, porl_stream_t stream);
/* dst (cols, ld_dst) = src (rows, ld_src)^T */
int porl_transpose(porl_ctx_t* ctx, const float* src, int64_t rows, int64_t cols, int64_t ld_src, float* dst,
                   int64_t ld_dst, porl_stream_t stream);

/* ---- K8/K9: scan encoder (util/costmap.py + agent/fasternet.py), forward only ------------------- */
/* state2costmap (util/costmap.py:7-64): out (B, 360, 256, 3) fp32 NHWC -- the buffer the reference returns
 * as the permuted (B, 3, 360, 256) view -- from state rows (B, >= 362 floats, stride ld_state).  Like the
 * reference it clamps the caller's state IN PLACE (values > 8 become 0, costmap.py:17).  Beams whose bin
 * falls outside [0, 256) set PORL_FLAG_SCAN_RANGE (the reference raises) and are skipped. */
int porl_costmap(porl_ctx_t* ctx, float* state, int64_t ld_state, int64_t B, float* out, porl_stream_t stream);

/* BatchNorm2d parameters and running statistics (device pointers, C floats each) */
typedef struct {
  const float* weight;
  const float* bias;
  float* running_mean;   /* updated: (1 - 0.1) r + 0.1 batch mean */
  float* running_var;    /* updated with the UNBIASED batch variance */
} porl_bn_t;

/* MLPBlock (agent/fasternet.py:141-195) of width C: weights in the layouts the kernels read */
typedef struct {
  const float* pconv_w;  /* (9, C/4, C/4): partial_conv3.weight (co, ci, kh, kw) permuted to [kh*3+kw][ci][co] */
  const float* w1;       /* (2C, C): mlp.0.weight */
  porl_bn_t bn;          /* mlp.1 over 2C channels */
  const float* w2;       /* (C, 2C): mlp.3.weight */
  const float* keep;     /* (B) DropPath factors mask / keep_prob (fasternet.py:76-93), or NULL = keep all */
} porl_fn_block_t;

/* Data-parallel BatchNorm: called on the host between the statistics kernel and the normalisation of every
 * BatchNorm layer with a device buffer of n doubles (per-channel sum, sum of squares of THIS rank's rows); it must
 * enqueue, on `stream`, an in-place SUM over all ranks (e.g. ncclAllReduce / torch.distributed.all_reduce) and
 * return 0.  The statistics then cover the global batch, as in a single-process run of the reference. */
typedef int (*porl_allreduce_f64_fn)(void* user, double* buf_dev, int64_t n, porl_stream_t stream);

/* FasterNet(in_chans=3, num_classes=feature_dim, embed_dim=96, depths=(1, 2)) as the agents build it
 * (sorl_train.py:29-32) */
typedef struct {
  const float* pe_w;          /* (48, 96): patch_embed.proj.weight (o, c, kh, kw) permuted to [(c, kh, kw)][o] */
  porl_bn_t pe_bn;            /* 96 */
  porl_fn_block_t block[3];   /* stages.0.blocks.0 (C = 96), stages.2.blocks.0/1 (C = 192) */
  const float* merge_w;       /* (192, 384): stages.1.reduction.weight (o, c, kh, kw) permuted to [o][(kh, kw, c)] */
  porl_bn_t merge_bn;         /* 192 */
  const float* head_conv_w;   /* (1280, 192): avgpool_pre_head.1.weight */
  const float* head_w;        /* (feature_dim, 1280): head.weight */
  const float* head_b;        /* (round_up(feature_dim, 4)), 16-byte aligned: head.bias */
  int32_t feature_dim;
  int32_t world;              /* ranks sharing the batch (every rank passes the same B); <= 1: statistics stay local */
  porl_allreduce_f64_fn stats_allreduce;   /* required when world > 1, ignored otherwise */
  void* stats_user;
  double* stats_buf;          /* optional (768 doubles): where the per-layer sums are staged for stats_allreduce;
                                 NULL = inside the workspace */
} porl_fasternet_t;

/* floats of scratch porl_fasternet_forward needs for a batch of B states */
int64_t porl_fasternet_workspace_floats(int64_t B);
/* FasterNet.forward_cls (agent/fasternet.py:428-438) in TRAIN mode: features (B, ld_feat >= feature_dim)
 * from state rows (B, ld_state >= 362), which are clamped in place like porl_costmap does.  Every
 * BatchNorm uses the statistics of THIS batch and updates its running statistics once (the caller
 * counts num_batches_tracked).  No backward pass exists: no optimizer owns these parameters in the
 * reference (agent/sorl.py:60,63), so its gradients are never applied. */
int porl_fasternet_forward(porl_ctx_t* ctx, const porl_fasternet_t* net, float* state, int64_t ld_state,
                           int64_t B, float* features, int64_t ld_feat, float* workspace,
                           int64_t workspace_floats, porl_stream_t stream);

/* ---- PointMLP scan encoder (agent/feature_pcl.py), forward in train mode -----------------------
 * PointMLPModel.forward (feature_pcl.py:407-420): embedding Conv1d(3,E)-BN-ReLU, then per stage
 *   LocalGrouper (feature_pcl.py:204-229): farthest-point sampling of S anchors (:69-89), k nearest neighbours by
 *     square_distance + top-k (:28-48, :163-174), gather, "center"/"anchor" normalisation by the per-sample std,
 *     affine, concat with the repeated anchor feature -> (B*S*k, 2C [+3]) rows,
 *   PreExtraction (:304-333): Conv-BN-ReLU transfer + residual Conv-BN-ReLU blocks over those rows, max over k,
 *   PosExtraction (:336-355): residual blocks over the (B*S, C') rows,
 * global max over the anchors and the classifier Linear-BN-ReLU-Dropout x2 + Linear (:392-402).
 * Activations are row-major (rows, channels) matrices; every Conv1d(k=1) / Linear is a K-major contraction.
 * BatchNorm1d uses batch statistics (training mode) and updates running_mean / running_var in place.
 * groups = 1 and ReLU only (the reference's defaults).  Single process (no cross-rank statistics). */
typedef struct porl_pm_conv {
  const float* weight;   /* (c_out, ld_w): Conv1d weight (c_out, c_in, 1) with rows padded to ld_w = pad4(c_in) */
  const float* bias;     /* (pad4(c_out)) or null (bias=False) */
  porl_bn_t bn;
} porl_pm_conv_t;

typedef struct porl_pm_res {   /* ConvBNReLURes1D (feature_pcl.py:263-301), groups = 1 */
  porl_pm_conv_t net1, net2;
  int32_t c_mid, pad_;   /* int(channel * res_expansion) */
} porl_pm_res_t;

#define PORL_PM_MAX_STAGES 4
#define PORL_PM_MAX_BLOCKS 4
typedef struct porl_pm_stage {
  int32_t anchors, k;              /* S and kneighbors */
  int32_t c_in, c_out;
  const float* affine_alpha;       /* (c_in [+3]) or null when normalize is off */
  const float* affine_beta;
  porl_pm_conv_t transfer;         /* in: 2 c_in [+3] (ld_w = pad4 of it) -> c_out */
  int32_t n_pre, n_pos;
  porl_pm_res_t pre[PORL_PM_MAX_BLOCKS], pos[PORL_PM_MAX_BLOCKS];
} porl_pm_stage_t;

typedef struct porl_pointmlp {
  int32_t points, embed_dim, n_stages, use_xyz;
  int32_t normalize;               /* 0 none, 1 center, 2 anchor */
  int32_t class_num, pad0_, pad1_;
  porl_pm_conv_t embedding;        /* weight (embed_dim, 4): Conv1d(3, E) rows padded to 4 */
  porl_pm_stage_t stage[PORL_PM_MAX_STAGES];
  const float *fc1_w, *fc1_b;      /* (512, C_last), (512) */
  porl_bn_t bn1;
  const float *fc2_w, *fc2_b;      /* (256, 512), (256) */
  porl_bn_t bn2;
  const float *fc3_w, *fc3_b;      /* (class_num, 256), (pad4(class_num)) */
  const float *drop1, *drop2;      /* (B, 512), (B, 256) multipliers mask / (1 - p) of nn.Dropout, or null */
} porl_pointmlp_t;

/* floats of workspace porl_pointmlp_forward needs for batch B (< 0: invalid configuration) */
int64_t porl_pointmlp_workspace_floats(const porl_pointmlp_t* net, int64_t B);
/* x: (B, 3, points) as the reference passes it (feature_pcl.py:408); fps_start: (n_stages, B) int64, the index
 * each sample's farthest-point sampling starts from (the reference draws it with torch.randint, :81);
 * out: (B, ld_out) receives class_num features per row. */
int porl_pointmlp_forward(porl_ctx_t* ctx, const porl_pointmlp_t* net, const float* x, const int64_t* fps_start,
                          int64_t B, float* out, int64_t ld_out, float* workspace, int64_t workspace_floats,
                          porl_stream_t stream);
/* the building blocks, exposed for tests: xyz (B, N, 3) row-major.
 * porl_fps: idx (B, npoint) int32, bit-identical to farthest_point_sample on the same start indices.
 * porl_knn: idx (B, S, k) int32: the k points nearest to each query (ascending distance, ties by index). */
int porl_fps(porl_ctx_t* ctx, const float* xyz, int64_t B, int32_t N, int32_t npoint, const int64_t* start,
             int32_t* idx, porl_stream_t stream);
int porl_knn(porl_ctx_t* ctx, const float* xyz, const float* query, int64_t B, int32_t N, int32_t S, int32_t k,
             int32_t* idx, porl_stream_t stream);
/* (B, 362) laser-scan states -> (B, 3, 361) point sets: beam i -> (r cos i deg, r sin i deg, 0) with r > 8 (no
 * return) collapsed to 0 as util/costmap.py:17 does, goal -> (gx, gy, 1): the intent of feature_pcl.py:113-138 */
int porl_scan_to_points(porl_ctx_t* ctx, const float* state, int64_t ld_state, int64_t B, float* out, porl_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* PORL_B200_H */
