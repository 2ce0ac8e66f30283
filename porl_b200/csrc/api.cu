// Context, error reporting and the contraction dispatcher of libporl_b200.so.
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include "gemm.cuh"

namespace porl {

static thread_local char g_error[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
}

static int check_batch(bool a_kmajor, bool b_kmajor, const GemmBatch& batch) {
  PORL_CHECK_ARG(batch.n_groups >= 1 && batch.n_groups <= kMaxGroups, "contraction: %d groups", batch.n_groups);
  for (int i = 0; i < batch.n_groups; ++i) {
    const GemmGroup& g = batch.g[i];
    PORL_CHECK_ARG(g.A && g.B && g.C, "contraction group %d: null operand", i);
    PORL_CHECK_ARG(g.M > 0 && g.N > 0 && g.K > 0 && g.n_store >= g.N && g.ldc >= g.n_store,
                   "contraction group %d: bad shape M=%d N=%d K=%d n_store=%d ldc=%d", i, g.M, g.N, g.K, g.n_store,
                   g.ldc);
    PORL_CHECK_ARG(aligned16(g.A) && aligned16(g.B) && (g.lda & 3) == 0 && (g.ldb & 3) == 0,
                   "contraction group %d: operands must be 16-byte aligned with leading dimensions %% 4 == 0 "
                   "(lda=%d ldb=%d)", i, g.lda, g.ldb);
    PORL_CHECK_ARG(g.lda >= (a_kmajor ? g.K : g.M) && g.ldb >= (b_kmajor ? g.K : g.N),
                   "contraction group %d: leading dimension smaller than the contiguous extent (lda=%d ldb=%d)", i,
                   g.lda, g.ldb);
    PORL_CHECK_ARG(g.epilogue >= 0 && g.epilogue <= 2, "contraction group %d: epilogue=%d", i, g.epilogue);
    PORL_CHECK_ARG(!g.relu_src || g.ld_relu >= g.N, "contraction group %d: ld_relu=%d", i, g.ld_relu);
    PORL_CHECK_ARG((!g.A_lo || aligned16(g.A_lo)) && (!g.B_lo || aligned16(g.B_lo)) && (!g.C_lo || aligned16(g.C_lo)),
                   "contraction group %d: lo planes must be 16-byte aligned", i);
  }
  return PORL_OK;
}

static double batch_flops(const GemmBatch& batch) {
  double flops = 0.0;
  for (int i = 0; i < batch.n_groups; ++i)
    flops += 2.0 * (double)batch.g[i].M * (double)batch.g[i].N * (double)batch.g[i].K;
  return flops;
}

int launch_gemm(porl_ctx* ctx, bool a_kmajor, bool b_kmajor, GemmBatch& batch, cudaStream_t stream) {
  int rc = check_batch(a_kmajor, b_kmajor, batch);
  if (rc != PORL_OK) return rc;
  ProfScope prof(ctx, PORL_CAT_CONTRACTION, batch_flops(batch), stream);
  if (ctx->gemm_mode == 1) {
    rc = launch_gemm_tc(ctx, a_kmajor, b_kmajor, batch, stream);
    if (rc != PORL_ERR_UNSUPPORTED) return rc;
  }
  for (int i = 0; i < batch.n_groups; ++i) {
    PORL_CHECK_ARG(!batch.g[i].Ct, "contraction group %d: a transposed output needs the tcgen05 path "
                   "(gemm mode 1, M, N, K >= 32)", i);
    PORL_CHECK_ARG(!batch.g[i].a_affine, "contraction group %d: the operand affine needs the tcgen05 path (gemm mode 1, "
                   "K-major A with K %% 32 == 0)", i);
  }
  return launch_gemm_simt(ctx, a_kmajor, b_kmajor, batch, stream);
}

int launch_gemm_phases(porl_ctx* ctx, int n_phases, TcPhaseSpec* specs, cudaStream_t stream) {
  if (ctx->gemm_mode != 1 || ctx->tc_phases == 0) return PORL_ERR_UNSUPPORTED;
  double flops = 0.0;
  for (int p = 0; p < n_phases; ++p) {
    int rc = check_batch(specs[p].a_kmajor, specs[p].b_kmajor, *specs[p].batch);
    if (rc != PORL_OK) return rc;
    for (int i = 0; i < specs[p].batch->n_groups; ++i)
      if (!gemm_tc_eligible(specs[p].batch->g[i], specs[p].a_kmajor, specs[p].b_kmajor)) return PORL_ERR_UNSUPPORTED;
    flops += batch_flops(*specs[p].batch);
  }
  ProfScope prof(ctx, PORL_CAT_CONTRACTION, flops, stream);
  return launch_gemm_tc_phases(ctx, n_phases, specs, stream);
}

}  // namespace porl

using namespace porl;

static void prof_clear(porl_ctx* ctx) {
  for (auto& r : ctx->prof) {
    cudaEventDestroy(r.e0);
    cudaEventDestroy(r.e1);
  }
  ctx->prof.clear();
}

extern "C" int porl_version(void) { return PORL_VERSION; }
extern "C" const char* porl_last_error(void) { return g_error; }

extern "C" int porl_ctx_create(porl_ctx_t** out, int device) {
  PORL_CHECK_ARG(out, "porl_ctx_create: null out");
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n <= 0) {
    set_error("porl_ctx_create: no CUDA device (%s); libporl_b200 has no CPU fallback",
              e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
    return PORL_ERR_CUDA;
  }
  PORL_CHECK_ARG(device >= 0 && device < n, "porl_ctx_create: device %d of %d", device, n);
  cudaDeviceProp prop;
  PORL_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) {
    set_error("porl_ctx_create: device %d is sm_%d%d; this library is built for sm_100a (B200) only", device,
              prop.major, prop.minor);
    return PORL_ERR_UNSUPPORTED;
  }
  int prev = 0;
  PORL_CUDA(cudaGetDevice(&prev));
  PORL_CUDA(cudaSetDevice(device));
  porl_ctx* ctx = new porl_ctx();
  ctx->profiling = false;
  ctx->launches = 0;
  ctx->tc_launches = 0;
  ctx->sem = nullptr;
  ctx->partials = nullptr;
  ctx->counters = nullptr;
  {
    const char* e = getenv("PORL_TC_CHUNK");            // k-blocks per TMEM accumulation chunk (porl_ctx_set_tc_chunk)
    ctx->tc_chunk_kb = e && atoi(e) > 0 ? atoi(e) : 1;
  }
  ctx->tc_terms = 3;
  {
    const char* e = getenv("PORL_TC_PHASES");
    ctx->tc_phases = e ? atoi(e) : 0;
  }
  {
    const char* e = getenv("PORL_TC_A_IN_TMEM");
    ctx->tc_a_in_tmem = e ? atoi(e) : 1;      // A operand through tensor memory wherever A is K-major (gemm_tc.cu)
    const char* pe = getenv("PORL_PDL");
    ctx->pdl = pe ? (atoi(pe) != 0) : 1;
  }
  ctx->tc_ws = nullptr;
  ctx->tc_dbg = nullptr;
  ctx->device = device;
  ctx->sm_count = prop.multiProcessorCount;
  ctx->gemm_mode = 0;
  cudaError_t rc = cudaMalloc(&ctx->sem, kSemInts * sizeof(int));
  if (rc == cudaSuccess) rc = cudaMemset(ctx->sem, 0, kSemInts * sizeof(int));
  if (rc == cudaSuccess) rc = cudaMalloc(&ctx->partials, kPartialFloats * sizeof(float));
  if (rc == cudaSuccess) rc = cudaMalloc(&ctx->tc_ws, kTcWsFloats * sizeof(float));
  if (rc == cudaSuccess) rc = cudaMalloc(&ctx->counters, kCounterWords * sizeof(uint32_t));
  if (rc == cudaSuccess) rc = cudaMemset(ctx->counters, 0, kCounterWords * sizeof(uint32_t));
  cudaSetDevice(prev);
  if (rc != cudaSuccess) {
    set_error("porl_ctx_create: workspace allocation failed: %s", cudaGetErrorString(rc));
    porl_ctx_destroy(ctx);
    return PORL_ERR_CUDA;
  }
  *out = ctx;
  return PORL_OK;
}

extern "C" int porl_ctx_destroy(porl_ctx_t* ctx) {
  if (!ctx) return PORL_OK;
  if (ctx->sem) cudaFree(ctx->sem);
  if (ctx->partials) cudaFree(ctx->partials);
  if (ctx->tc_ws) cudaFree(ctx->tc_ws);
  if (ctx->counters) cudaFree(ctx->counters);
  prof_clear(ctx);
  delete ctx;
  return PORL_OK;
}

extern "C" int porl_profile_enable(porl_ctx_t* ctx, int enable) {
  PORL_CHECK_ARG(ctx, "porl_profile_enable: null ctx");
  if (enable) prof_clear(ctx);
  ctx->profiling = enable != 0;
  return PORL_OK;
}

extern "C" int porl_profile_read(porl_ctx_t* ctx, porl_profile_t* out) {
  PORL_CHECK_ARG(ctx && out, "porl_profile_read: null argument");
  memset(out, 0, sizeof(*out));
  for (auto& r : ctx->prof) {
    PORL_CUDA(cudaEventSynchronize(r.e1));
    float ms = 0.f;
    PORL_CUDA(cudaEventElapsedTime(&ms, r.e0, r.e1));
    if (r.cat >= 0 && r.cat < PORL_CAT_COUNT) {
      out->launches[r.cat] += 1;
      out->ms[r.cat] += ms;
      out->work[r.cat] += r.work;
    }
  }
  return PORL_OK;
}

extern "C" int64_t porl_launch_count(const porl_ctx_t* ctx) { return ctx ? ctx->launches : -1; }

extern "C" int64_t porl_tc_launch_count(const porl_ctx_t* ctx) { return ctx ? ctx->tc_launches : -1; }

extern "C" int porl_ctx_status_ptr(porl_ctx_t* ctx, uint32_t** out) {
  PORL_CHECK_ARG(ctx && out, "porl_ctx_status_ptr: null argument");
  *out = ctx->counters;
  return PORL_OK;
}

extern "C" int porl_ctx_set_gemm_mode(porl_ctx_t* ctx, int mode) {
  PORL_CHECK_ARG(ctx && (mode == 0 || mode == 1), "porl_ctx_set_gemm_mode: mode=%d", mode);
  ctx->gemm_mode = mode;
  return PORL_OK;
}

extern "C" int porl_ctx_get_gemm_mode(const porl_ctx_t* ctx) { return ctx ? ctx->gemm_mode : -1; }

extern "C" int porl_ctx_set_tc_phases(porl_ctx_t* ctx, int enable) {
  PORL_CHECK_ARG(ctx, "porl_ctx_set_tc_phases: null context");
  ctx->tc_phases = enable ? 1 : 0;
  return PORL_OK;
}

extern "C" int porl_ctx_set_tc_terms(porl_ctx_t* ctx, int terms) {
  PORL_CHECK_ARG(ctx && (terms == 1 || terms == 3), "porl_ctx_set_tc_terms: terms=%d (1 or 3)", terms);
  ctx->tc_terms = terms;
  return PORL_OK;
}

extern "C" int porl_ctx_set_tc_chunk(porl_ctx_t* ctx, int k_blocks) {
  PORL_CHECK_ARG(ctx && k_blocks >= 1 && k_blocks <= 64, "porl_ctx_set_tc_chunk: k_blocks=%d", k_blocks);
  ctx->tc_chunk_kb = k_blocks;
  return PORL_OK;
}

extern "C" int porl_ctx_set_pdl(porl_ctx_t* ctx, int enable) {
  PORL_CHECK_ARG(ctx, "porl_ctx_set_pdl: null context");
  ctx->pdl = enable ? 1 : 0;
  return PORL_OK;
}

extern "C" int porl_ctx_set_tc_variant(porl_ctx_t* ctx, int a_in_tmem) {
  PORL_CHECK_ARG(ctx && a_in_tmem >= 0 && a_in_tmem <= 2, "porl_ctx_set_tc_variant: variant=%d", a_in_tmem);
  ctx->tc_a_in_tmem = a_in_tmem;
  return PORL_OK;
}

extern "C" int porl_ctx_set_tc_debug(porl_ctx_t* ctx, uint64_t* stamps_dev) {
  PORL_CHECK_ARG(ctx, "porl_ctx_set_tc_debug: null ctx");
  ctx->tc_dbg = reinterpret_cast<unsigned long long*>(stamps_dev);
  return PORL_OK;
}

extern "C" int porl_gemm(porl_ctx_t* ctx, int32_t a_kmajor, int32_t b_kmajor, int64_t M, int64_t N, int64_t K,
                         const float* A, int64_t lda, const float* B, int64_t ldb, float* C, int64_t ldc,
                         const float* bias, int32_t epilogue, const float* relu_src, int64_t ld_relu,
                         int64_t n_store, float* colsum, float* Ct, int64_t ldct, porl_stream_t stream) {
  PORL_ON_DEVICE(ctx);
  PORL_CHECK_ARG(ctx, "porl_gemm: null ctx");
  PORL_CHECK_ARG(M > 0 && N > 0 && K > 0 && M < (1 << 24) && N < (1 << 24) && K < (1 << 24),
                 "porl_gemm: bad shape M=%lld N=%lld K=%lld", (long long)M, (long long)N, (long long)K);
  GemmBatch b;
  b.n_groups = 1;
  GemmGroup& g = b.g[0];
  g.A = A; g.B = B; g.C = C; g.A_lo = nullptr; g.B_lo = nullptr; g.C_lo = nullptr; g.bias = bias; g.relu_src = relu_src; g.colsum = colsum;
  g.Ct = Ct; g.ldct = (int)ldct;
  g.bits_out = nullptr; g.bits_in = nullptr; g.ld_bits = 0;
  g.a_affine = nullptr;
  g.M = (int)M; g.N = (int)N; g.K = (int)K;
  g.lda = (int)lda; g.ldb = (int)ldb; g.ldc = (int)ldc; g.ld_relu = (int)ld_relu;
  g.n_store = (int)(n_store > 0 ? n_store : N);
  g.epilogue = epilogue;
  return launch_gemm(ctx, a_kmajor != 0, b_kmajor != 0, b, as_stream(stream));
}
