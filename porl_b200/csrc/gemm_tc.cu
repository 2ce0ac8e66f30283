// tcgen05 3xTF32 grouped contraction with fp32-grade accuracy: C[M,N] = A[M,K] * B[N,K]^T, both
// operands plain fp32, K-major (K contiguous), optional transposed copy Ct[N,M] of the result.
//
// Persistent, warp-specialised kernel, one CTA per SM, 128x128 output tiles, 32-wide k-blocks:
//   warp 0      TMA producer: raw fp32 A / B tiles (128 rows x 128 B, SWIZZLE_128B) -> stage ring
//   warps 2..7  splitters: lo = RN_tf32(x - trunc_tf32(x)) written next to the raw tile (same
//               swizzled layout, so the split is a flat elementwise pass over the stage).  The raw
//               tile itself serves as the hi operand: kind::tf32 reads only the top 19 bits of
//               each fp32 container, i.e. it truncates, and the split is built on that truncation.
//   warp 1      TMEM allocator + MMA issuer: per k-block 8 "small" MMAs (A_lo*B_hi, A_hi*B_lo for
//               the four 8-wide k-steps) then 4 "big" ones (A_hi*B_hi), all into one fp32 TMEM
//               accumulator that is RESTARTED every chunk (chunk_kb k-blocks, default 1).
//   warps 8..15 accumulators + epilogue (two sets of four, 64 columns each): drain each chunk's TMEM accumulator and add it
//               to a register accumulator with round-to-nearest fp32 adds, in fixed order.  The
//               tensor core truncates (RZ) on every accumulate; restarting the TMEM accumulator
//               every 32..64 k keeps that error at the fp32 rounding level (measured: DESIGN.md).
//               After the last chunk: bias / ReLU / tanh / ReLU-mask epilogue straight from the
//               registers (thread = row): 16-byte row-major stores, coalesced transposed stores.
// TMEM holds two chunk accumulators (2 x 128 columns) so the MMAs of chunk c+1 overlap the drain
// of chunk c.  mbarriers: raw_full[s] (TMA -> splitters), split_done[s] (splitters -> MMA),
// empty[s] (tcgen05.commit -> TMA), tmem_full[b] (commit -> accumulators), tmem_empty[b].
// Split-K (weight gradients, K = batch): every slice parks its partial tile in a workspace and the last
// slice to arrive sums the partials in slice order, so results are bitwise reproducible and no slice
// waits for another one.
#include <cuda.h>

#include <stdlib.h>

#include <type_traits>
#include <unordered_map>

#include "gemm.cuh"

namespace porl {

constexpr int TC_BM = 128, TC_BN = 128, TC_BK = 32;      // tile; BK floats = one 128-byte swizzle span
constexpr int TC_TILE_BYTES = TC_BM * TC_BK * 4;          // 16 KB
constexpr int TC_THREADS = 512;                           // 16 warps: TMA, MMA, 6 splitters, 8 accumulators (<= 128 registers each)
constexpr int TC_SPLIT_THREADS = 192;
// Two variants.  ATM = false: A_hi, A_lo, B_hi, B_lo all in shared memory (stage = 4 tiles, 3 stages).
// ATM = true: the splitters put A_hi / A_lo into TENSOR MEMORY (tcgen05.st, 64 columns per stage) and the
// MMAs take A from TMEM, which removes the A reads from the shared-memory port (the measured limiter)
// and the A_lo tile from the stage (stage = A_raw, B_raw, B_lo: 3 tiles, 4 stages).
// PAIR (two CTAs per 256 x 128 tile): each CTA stages A (16 KB) and HALF of B (8 KB) raw plus their lo copies:
// 48 KB per stage, 4 stages -- the ring is latency bound (TMA + split + MMA + commit per stage), so depth matters.
template <bool ATM, bool PAIR = false>
struct TcCfg {
  static constexpr int STAGES = (ATM || PAIR) ? 4 : 3;
  static constexpr int B_BYTES = PAIR ? TC_TILE_BYTES / 2 : TC_TILE_BYTES;      // B rows staged per CTA: 64 or 128
  static constexpr int RAW_BYTES = TC_TILE_BYTES + B_BYTES;                     // raw A, raw B (adjacent)
  static constexpr int STAGE_BYTES = ATM ? 3 * TC_TILE_BYTES : 2 * RAW_BYTES;
  static constexpr int A_LO_OFF = RAW_BYTES;                                    // lo A, lo B follow in the same order
  static constexpr int B_LO_OFF = ATM ? 2 * TC_TILE_BYTES : RAW_BYTES + TC_TILE_BYTES;
  // chunk accumulators in TMEM: the MMAs may run NBUF chunks ahead of the accumulator warps, which matters while
  // those warps are busy with a tile's epilogue (short-K tiles: a whole next tile fits into four buffers)
  static constexpr int NBUF = ATM ? 2 : 4;
  static constexpr int ACC_COLS = NBUF * TC_BN;
  static constexpr int NUM_BARS = 3 * STAGES + 2 * NBUF + 1;   // + phase_go (grid barrier passed: producer may load)
  static constexpr int STAGING_BYTES = 2 * TC_TILE_BYTES;   // epilogue: two 128 x 32 fp32 slabs for the TMA stores of C
  static constexpr int SMEM_BYTES = 1024 /*align slack*/ + STAGES * STAGE_BYTES + STAGING_BYTES + 256 /*barriers*/ + 512 /*bias*/;
  static_assert(SMEM_BYTES <= 227 * 1024, "shared memory per CTA");
  static constexpr int TMEM_COLS = 512;
};
constexpr int TC_SPLIT_WARP0 = 2, TC_EPI_WARP0 = 8;

// One launch runs up to kMaxPhases grouped contractions ("phases") back to back in the same persistent grid: the layers
// of an MLP forward, or {weight gradient, input gradient, weight gradient of the layer below} of a backward.  A phase
// whose inputs were produced by an earlier phase of the launch starts behind a grid-wide barrier (barrier_before);
// phases that are independent of each other share the pool of CTAs without one.  This replaces a kernel boundary
// (launch, TMEM allocation, barrier set-up, tensor-map fetch, pipeline fill and drain: ~9 us each at these sizes) by a
// ~2 us barrier.
constexpr int kMaxPhases = 3;
struct alignas(64) TcPhase {
  CUtensorMap mapA[kMaxGroups];
  CUtensorMap mapB[kMaxGroups];
  CUtensorMap mapC[kMaxGroups];     // row-major C (n_store x M), box {32, 128}: the epilogue's TMA stores
  GemmBatch batch;                  // problems, split factor, ticket words of this phase
  int tile_begin[kMaxGroups + 1];   // prefix sum of tiles per group
  int tiles_n[kMaxGroups];          // n-tiles per group
  int total_tiles;
  int dist_fix;                     // split-K: one unit per CTA, the slices of a tile share its final sum (see the epilogue)
  int a_mn, b_mn;                   // operand is MN-major in memory (A[k*lda + m] / B[k*ldb + n]): no transposed copy needed
  int barrier_before;               // wait for every CTA to have finished the previous phases before loading operands
  float* ws;                        // split-K partial tiles of this phase: [slice][tile][col/4][row] float4
};
struct alignas(64) TcParams {
  TcPhase ph[kMaxPhases];
  CUtensorMap mapB2[kMaxGroups];    // pair mode (one phase): B with box {32, 64} (each CTA of a pair loads half of the B tile)
  int n_phases;
  int chunk_kb;                     // k-blocks per TMEM accumulation chunk
  int terms;                        // 3: 3xTF32; 1: hi*hi MMAs only, no operand split (ceiling probe)
  int tile_m;                       // rows per work unit: 128, or 256 in pair mode (two CTAs x 128)
  uint32_t* gbar;                   // grid barrier words: [2 * kMaxPhases] (arrived, departed per barrier), self-resetting
  uint32_t* status;                 // device status word (PORL_FLAG_*)
  unsigned long long* dbg;          // optional per-CTA timestamps (tools/gemm_probe.py), or null
};

// ------------------------------------------------------------------------------------------ PTX
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred P1;\n\t"
      "WAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
      "@P1 bra WAIT_DONE;\n\t"
      "bra WAIT_LOOP;\n\t"
      "WAIT_DONE:\n\t"
      "}" ::"r"(bar), "r"(parity)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, uint32_t src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(map)), "r"(src), "r"(c0), "r"(c1)
               : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
// ---- CTA-pair (cta_group::2) helpers
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t mapa_rank(uint32_t addr, uint32_t rank) {      // same smem offset in CTA `rank`
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tc_commit_pair(uint32_t bar) {       // arrives on the barrier at this offset in BOTH CTAs
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(bar), "h"((uint16_t)3)
               : "memory");
}
__device__ __forceinline__ void tc_mma_tf32_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                                 uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// one lane of a converged warp (elect.sync): the compiler knows the guarded block runs on exactly one thread and keeps
// warp-uniform operands of the tcgen05 instructions inside it in uniform registers
__device__ __forceinline__ bool elect_one() {
  uint32_t pred = 0;
  asm volatile("{\n\t.reg .pred px;\n\telect.sync _|px, 0xffffffff;\n\t@px mov.s32 %0, 1;\n\t}" : "+r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_mma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tc_ld32(uint32_t taddr, float (&v)[32]) {
  uint32_t r[32];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tc_st32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]),
      "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]),
      "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
__device__ __forceinline__ void tc_mma_tf32_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc,
                                               uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t"
      "}" ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ unsigned long long gtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// Shared-memory matrix descriptor (cute::UMMA::SmemDescriptor bit layout) of a K-major SWIZZLE_128B
// tile: start address, SBO = 1024 B (8-row groups of 128-byte rows), version 1 (sm_100), layout type 2.
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr) {
  return (uint64_t)((addr >> 4) & 0x3fffu) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) |
         (2ull << 61);
}

// MN-major operands (A[k][m] with m contiguous).  For 32-bit elements the tensor core accepts ONE shared-memory
// layout in this orientation: "128-byte swizzle with 32-byte atomicity" (descriptor layout type 1; TMA swizzle mode
// CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B): each 128-byte line holds 32 consecutive MN elements of one k, its four 32-byte
// chunks are XOR-ed with (line & 3), four k-lines form a swizzle atom.  Canonical form ((32 mn, atoms), (4 k, groups)) :
// ((1, LBO), (32, SBO)) in floats.  Here a 128 x 32 (MN x K) tile is four TMA boxes {32 mn, 32 k} of 4096 bytes each,
// lines contiguous: LBO = 4096 (next 32 MN elements), SBO = 512 (next four k-lines); one K = 8 MMA reads two k-groups
// starting at start address + 1024 * ks.
__device__ __forceinline__ uint64_t smem_desc_mn(uint32_t addr) {
  return (uint64_t)((addr >> 4) & 0x3fffu) | ((uint64_t)(4096 >> 4) << 16) | ((uint64_t)(512 >> 4) << 32) | (1ull << 46) |
         (1ull << 61);
}

// lo plane of the truncation split: x = trunc_tf32(x) + d exactly; lo = d rounded to tf32
// (adding half a tf32 ulp to the bit pattern; the tensor core's own truncation of the container completes
// the round-to-nearest, so no second mask is needed: 3 ALU instructions per element)
__device__ __forceinline__ float split_lo(float x) {
  const float d = x - __uint_as_float(__float_as_uint(x) & 0xffffe000u);
  return __uint_as_float(__float_as_uint(d) + 0x1000u);
}

struct TcUnit {
  int g, tile, m0, n0, kb0, kb1, split;
};

__device__ __forceinline__ bool tc_decode(const TcPhase& F, int tile_m, int u, TcUnit& w) {
  if (u >= F.total_tiles * F.batch.splits) return false;
  w.split = u / F.total_tiles;
  int t = u - w.split * F.total_tiles;
  int g = 0;
  while (g + 1 < F.batch.n_groups && t >= F.tile_begin[g + 1]) ++g;
  t -= F.tile_begin[g];
  w.g = g;
  w.tile = t;
  const int nt = F.tiles_n[g];
  w.m0 = (t / nt) * tile_m;
  w.n0 = (t % nt) * TC_BN;
  const int nkb = (F.batch.g[g].K + TC_BK - 1) / TC_BK;
  w.kb0 = (int)((long long)w.split * nkb / F.batch.splits);
  w.kb1 = (int)((long long)(w.split + 1) * nkb / F.batch.splits);
  return true;
}

// ------------------------------------------------------------------------------------------ kernel
// PAIR: two CTAs of a cluster work on one 256 x 128 tile with tcgen05.mma.cta_group::2 (M = 256): each CTA owns 128
// rows of A and of the accumulator and stages only HALF of the B tile (64 rows), which the hardware shares between
// the two tensor cores -- 25 % less shared-memory traffic per MMA, the measured limiter.  The leader CTA (rank 0)
// issues the MMAs for both; the peer's splitters and accumulator warps signal the leader's barriers through the
// cluster address space, tcgen05.commit multicasts completion to both CTAs.
template <bool ATM, bool PAIR>
__global__ void __launch_bounds__(TC_THREADS, 1) gemm_tc_kernel(const __grid_constant__ TcParams P) {
  static_assert(!(ATM && PAIR), "pair mode uses the shared-memory A operand");
  using C = TcCfg<ATM, PAIR>;
  constexpr int TC_STAGES = C::STAGES, TC_STAGE_BYTES = C::STAGE_BYTES, TC_NUM_BARS = C::NUM_BARS;
  constexpr int TC_TMEM_COLS = C::TMEM_COLS, TC_NBUF = C::NBUF, TC_ACC_COLS = C::ACC_COLS;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;                         // SWIZZLE_128B tiles: 1024-B aligned
  uint8_t* base_ptr = smem_raw + (base - raw);
  uint8_t* const staging = base_ptr + TC_STAGES * TC_STAGE_BYTES;       // 1024-byte aligned like the stages
  uint64_t* bars = reinterpret_cast<uint64_t*>(staging + C::STAGING_BYTES);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + TC_NUM_BARS);
  uint32_t* flag_slot = tmem_slot + 1;
  float* const sbias = reinterpret_cast<float*>(staging + C::STAGING_BYTES + 256);   // 128 floats, 16-byte aligned
  const uint32_t bar0 = smem_u32(bars);
  auto raw_full = [&](int s) { return bar0 + 8u * s; };
  auto split_done = [&](int s) { return bar0 + 8u * (TC_STAGES + s); };
  auto empty_bar = [&](int s) { return bar0 + 8u * (2 * TC_STAGES + s); };
  auto tmem_full = [&](int b) { return bar0 + 8u * (3 * TC_STAGES + b); };
  auto tmem_empty = [&](int b) { return bar0 + 8u * (3 * TC_STAGES + TC_NBUF + b); };
  const uint32_t phase_go = bar0 + 8u * (3 * TC_STAGES + 2 * TC_NBUF);

  // warp index through a shuffle: provably warp-uniform for the compiler, so that the role branches below are uniform
  // control flow and a role's warp-uniform values (descriptors, barrier addresses) can live in uniform registers
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const uint32_t rank = PAIR ? cluster_ctarank() : 0u;          // CTA of the pair; rank 0 leads
  const int unit0 = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int unit_stride = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  constexpr uint32_t B_BYTES = C::B_BYTES;
  unsigned long long* const dbg = P.dbg ? P.dbg + 16 * blockIdx.x : nullptr;   // [0] start [1] setup done
  // [2] first TMA issued [3] last TMA issued [6] last MMA commit; wait totals in SM clocks: [11] producer on empty
  // [12] splitters on raw_full [13] MMA on split_done [14] MMA on tmem_empty [4] accumulators on tmem_full [5] CTA clocks
  // [7] first chunk full [8] last chunk drained [9] first tile epilogue done [10] all tiles done [15] end
  const long long clk0 = clock64();
  // PDL: the next kernel of the stream may be scheduled from here on; this one sets up barriers and tensor memory
  // below and only then waits for the previous grid (nothing before the wait reads or writes its data)
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  if (dbg && threadIdx.x == 0) dbg[0] = gtime();

  if (warp == 0 && lane == 0) {
    for (int p = 0; p < P.n_phases; ++p) {
      for (int g = 0; g < P.ph[p].batch.n_groups; ++g) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&P.ph[p].mapA[g])) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&P.ph[p].mapB[g])) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&P.ph[p].mapC[g])) : "memory");
      }
    }
    mbar_init(phase_go, 1);
    for (int s = 0; s < TC_STAGES; ++s) {
      mbar_init(raw_full(s), 1);
      mbar_init(split_done(s), PAIR ? 2 * TC_SPLIT_THREADS : TC_SPLIT_THREADS);
      mbar_init(empty_bar(s), 1);
    }
    for (int b = 0; b < TC_NBUF; ++b) {
      mbar_init(tmem_full(b), 1);
      mbar_init(tmem_empty(b), PAIR ? 512 : 256);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 1) {
    if (PAIR) {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                   "r"(TC_TMEM_COLS)
                   : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    } else {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                   "r"(TC_TMEM_COLS)
                   : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
  }
  tc_fence_before();
  if (PAIR) cluster_sync_all();         // barriers of both CTAs initialised before any remote arrive / multicast
  else __syncthreads();
  tc_fence_after();
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_slot);
  if (dbg && threadIdx.x == 0) dbg[1] = gtime();
  const int chunk_kb = P.chunk_kb;
  const bool three = P.terms != 1;

  if (warp == 0) {
    // ===================== TMA producer =====================
    // like the MMA issuer: the whole warp runs the loops, one elected lane issues (uniform-register tensor-map
    // addresses and coordinates instead of an R2UR loop per copy -- up to eight copies per k-block with MN-major operands)
    {
      const bool leader = elect_one();
      int it = 0;
      long long w_prod = 0;
      TcUnit w;
      uint32_t n_bar = 0;
      for (int p = 0; p < P.n_phases; ++p) {
      const TcPhase& F = P.ph[p];
      if (F.barrier_before) {
        // the accumulator warps of this CTA pass the grid barrier (every CTA's stores of the earlier phases are
        // complete and visible) and release this thread; the operands are then read through the async proxy
        mbar_wait(phase_go, n_bar & 1u);
        ++n_bar;
        asm volatile("fence.proxy.async;" ::: "memory");
      }
      for (int u = unit0; tc_decode(F, P.tile_m, u, w); u += unit_stride) {
        for (int kb = w.kb0; kb < w.kb1; ++kb, ++it) {
          const int s = it % TC_STAGES;
          const uint32_t ph = (uint32_t)(it / TC_STAGES) & 1u;
          const long long tw = dbg ? clock64() : 0;
          mbar_wait(empty_bar(s), ph ^ 1u);
          if (dbg) w_prod += clock64() - tw;
          const uint32_t st = base + (uint32_t)s * TC_STAGE_BYTES;
          if (leader) {
            mbar_expect_tx(raw_full(s), TC_TILE_BYTES + B_BYTES);
            if (F.a_mn) {         // MN-major: four {32 m, 32 k} boxes, 4096 bytes apart
#pragma unroll
              for (int i = 0; i < 4; ++i) tma_load_2d(st + 4096u * i, &F.mapA[w.g], raw_full(s), w.m0 + 32 * i, kb * TC_BK);
            } else {
              tma_load_2d(st, &F.mapA[w.g], raw_full(s), kb * TC_BK, w.m0 + (int)rank * TC_BM);
            }
            if (F.b_mn) {
#pragma unroll
              for (int i = 0; i < 4; ++i)
                tma_load_2d(st + TC_TILE_BYTES + 4096u * i, &F.mapB[w.g], raw_full(s), w.n0 + 32 * i, kb * TC_BK);
            } else if (PAIR) {
              tma_load_2d(st + TC_TILE_BYTES, &P.mapB2[w.g], raw_full(s), kb * TC_BK, w.n0 + (int)rank * (TC_BN / 2));
            } else {
              tma_load_2d(st + TC_TILE_BYTES, &F.mapB[w.g], raw_full(s), kb * TC_BK, w.n0);
            }
            if (dbg && it == 0) dbg[2] = gtime();
          }
          __syncwarp();
        }
      }
      }
      if (dbg && leader) { dbg[3] = gtime(); dbg[11] = (unsigned long long)w_prod; }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (pair mode: the leader CTA only) =====================
    // The WHOLE warp runs the loops below in lockstep (uniform control flow, every lane waits on the barriers) and one
    // elected lane issues.  With a single-thread role (`if (lane == 0)`) the compiler treats every operand as
    // divergent: each tcgen05.mma became an ELECT / 7 x R2UR / branch loop of ~28 instructions and the issue path, not
    // the tensor pipe, set the pace -- 137 clk per MMA where the pipe needs 64 (tools/micro/umma_rate.cu measures
    // 67.6 clk for this exact stream issued this way, 146-226 for the divergent form).
    if (rank == 0) {
      const bool leader = elect_one();
      const uint32_t tmem_base_u = __shfl_sync(0xffffffffu, tmem_base, 0);
      // instruction descriptor: D fp32, A/B tf32, both K-major, N >> 3 at [17,23), M >> 4 at [24,29)
      int it = 0, chunk = 0;
      long long w_te = 0, w_sd = 0;
      TcUnit w;
      for (int p = 0; p < P.n_phases; ++p) {
      const TcPhase& F = P.ph[p];
      // bit 15 / 16: A / B is MN-major ("transposed") in shared memory
      const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(TC_BN >> 3) << 17) |
                             ((uint32_t)((PAIR ? 2 * TC_BM : TC_BM) >> 4) << 24) | ((F.a_mn && !ATM) ? (1u << 15) : 0u) |
                             (F.b_mn ? (1u << 16) : 0u);      // A from tensor memory has no shared-memory layout
      // 16-byte units the start address advances per 8-wide k-step: 32 bytes along a K-major row, one 1024-byte
      // k-group of an MN-major tile
      const uint64_t a_step = F.a_mn ? 64u : 2u, b_step = F.b_mn ? 64u : 2u;
      // descriptor templates without the start address (this thread alone paces the tensor pipe: nothing that can be
      // decided per phase is recomputed per k-block); stage addresses are 1024-byte aligned, so adding (bytes >> 4)
      // to a descriptor moves its 14-bit start-address field only
      const uint64_t tmpl_a = (F.a_mn ? smem_desc_mn(0u) : smem_desc(0u)) + (uint64_t)(base >> 4);
      const uint64_t tmpl_b = (F.b_mn ? smem_desc_mn(0u) : smem_desc(0u)) + (uint64_t)((base + TC_TILE_BYTES) >> 4);
      for (int u = unit0; tc_decode(F, P.tile_m, u, w); u += unit_stride) {
        int kb = w.kb0;
        while (kb < w.kb1) {
          const int buf = chunk % TC_NBUF;
          const long long tw0 = dbg ? clock64() : 0;
          mbar_wait(tmem_empty(buf), ((uint32_t)(chunk / TC_NBUF) & 1u) ^ 1u);
          if (dbg) w_te += clock64() - tw0;
          tc_fence_after();
          const uint32_t d_tmem = tmem_base_u + (uint32_t)(buf * TC_BN);
          const int nk = (w.kb1 - kb) < chunk_kb ? (w.kb1 - kb) : chunk_kb;
          for (int j = 0; j < nk; ++j, ++it) {
            const int s = it % TC_STAGES;
            const uint32_t ph = (uint32_t)(it / TC_STAGES) & 1u;
            const long long tw1 = dbg ? clock64() : 0;
            mbar_wait(split_done(s), ph);
            if (dbg) w_sd += clock64() - tw1;
            tc_fence_after();
            const uint32_t acc0 = j == 0 ? 0u : 1u;           // the chunk's first MMA starts a fresh accumulation
            const uint64_t st16 = (uint64_t)((uint32_t)s * (TC_STAGE_BYTES >> 4));
            const uint64_t b_hi = tmpl_b + st16;
            const uint64_t b_lo = b_hi + (uint64_t)((C::B_LO_OFF - TC_TILE_BYTES) >> 4);
            if (leader) {
              // small terms first (they meet the smallest accumulator), then the hi*hi terms
              if (ATM) {
                const uint32_t a_hi = tmem_base_u + (uint32_t)(TC_ACC_COLS + s * 64), a_lo = a_hi + 32u;
#pragma unroll
                for (int ks = 0; ks < TC_BK / 8; ++ks) {
                  const uint64_t o = b_step * ks;
                  tc_mma_tf32_ts(d_tmem, a_lo + 8u * ks, b_hi + o, idesc, ks == 0 ? acc0 : 1u);
                  tc_mma_tf32_ts(d_tmem, a_hi + 8u * ks, b_lo + o, idesc, 1u);
                }
#pragma unroll
                for (int ks = 0; ks < TC_BK / 8; ++ks)
                  tc_mma_tf32_ts(d_tmem, a_hi + 8u * ks, b_hi + b_step * ks, idesc, 1u);
              } else {
                const uint64_t a_hi = tmpl_a + st16;
                const uint64_t a_lo = a_hi + (uint64_t)(C::A_LO_OFF >> 4);
                if (three) {
#pragma unroll
                  for (int ks = 0; ks < TC_BK / 8; ++ks) {
                    const uint64_t oa = a_step * ks, ob = b_step * ks;
                    if (PAIR) {
                      tc_mma_tf32_pair(d_tmem, a_lo + oa, b_hi + ob, idesc, ks == 0 ? acc0 : 1u);
                      tc_mma_tf32_pair(d_tmem, a_hi + oa, b_lo + ob, idesc, 1u);
                    } else {
                      tc_mma_tf32(d_tmem, a_lo + oa, b_hi + ob, idesc, ks == 0 ? acc0 : 1u);
                      tc_mma_tf32(d_tmem, a_hi + oa, b_lo + ob, idesc, 1u);
                    }
                  }
                }
#pragma unroll
                for (int ks = 0; ks < TC_BK / 8; ++ks) {
                  const uint64_t oa = a_step * ks, ob = b_step * ks;
                  const uint32_t acc = (three || ks != 0) ? 1u : acc0;
                  if (PAIR) tc_mma_tf32_pair(d_tmem, a_hi + oa, b_hi + ob, idesc, acc);
                  else tc_mma_tf32(d_tmem, a_hi + oa, b_hi + ob, idesc, acc);
                }
              }
              if (PAIR) tc_commit_pair(empty_bar(s));   // frees the stage (in both CTAs) once these MMAs have read it
              else tc_commit(empty_bar(s));
              if (j == nk - 1) {                        // chunk accumulator complete (pair mode: both CTAs' halves)
                if (PAIR) tc_commit_pair(tmem_full(buf));
                else tc_commit(tmem_full(buf));
              }
            }
            __syncwarp();
          }
          ++chunk;
          kb += nk;
        }
      }
      }
      if (dbg && leader) { dbg[6] = gtime(); dbg[13] = (unsigned long long)w_sd; dbg[14] = (unsigned long long)w_te; }
    }
  } else if (warp < TC_EPI_WARP0) {
    // ===================== splitters: lo = RN_tf32(x - trunc_tf32(x)) =====================
    const int t = threadIdx.x - TC_SPLIT_WARP0 * 32;     // 0..191
    int it = 0;
    long long w_rf = 0;
    TcUnit w;
    for (int p = 0; p < P.n_phases; ++p)
    for (int u = unit0; tc_decode(P.ph[p], P.tile_m, u, w); u += unit_stride) {
      for (int kb = w.kb0; kb < w.kb1; ++kb, ++it) {
        const int s = it % TC_STAGES;
        const uint32_t ph = (uint32_t)(it / TC_STAGES) & 1u;
        const long long tw = dbg ? clock64() : 0;
        mbar_wait(raw_full(s), ph);
        if (dbg) w_rf += clock64() - tw;
        uint8_t* stage = base_ptr + (size_t)s * TC_STAGE_BYTES;
        if (ATM) {
          // B: flat elementwise pass raw -> lo (same swizzled layout), shared by all six splitter warps
          const float4* src = reinterpret_cast<const float4*>(stage + TC_TILE_BYTES);
          float4* dst = reinterpret_cast<float4*>(stage + C::B_LO_OFF);
#pragma unroll
          for (int i = 0; i < (TC_TILE_BYTES / 16 + TC_SPLIT_THREADS - 1) / TC_SPLIT_THREADS; ++i) {
            const int e = t + i * TC_SPLIT_THREADS;
            if (e < TC_TILE_BYTES / 16) {
              const float4 v = src[e];
              dst[e] = make_float4(split_lo(v.x), split_lo(v.y), split_lo(v.z), split_lo(v.w));
            }
          }
          if (warp < TC_SPLIT_WARP0 + 4) {
            // A (first four splitter warps): this thread owns tile row r = its TMEM lane; un-swizzle the row
            // (K-major: 16-byte chunk j of row r sits at chunk j ^ (r & 7)) and store hi / lo as 32 + 32 TMEM columns
            // (column = k index)
            const int r = 32 * (warp & 3) + lane;
            uint32_t hi[32], lo[32];
            if (P.ph[p].a_mn) {
              // MN-major A tile (four boxes of {32 m, 32 k}): box warp & 3 holds this warp's 32 rows; its line k is the
              // 128 bytes A[k][m0 .. m0 + 31] whose four 32-byte chunks are XOR-ed with (k & 3).  The warp reads one
              // whole line per step (lane = row): conflict-free.
              const uint8_t* abox = stage + (warp & 3) * 4096 + ((lane & 7) << 2);
              const int c = lane >> 3;
#pragma unroll
              for (int k = 0; k < 32; ++k) {
                const float v = *reinterpret_cast<const float*>(abox + k * 128 + ((c ^ (k & 3)) << 5));
                hi[k] = __float_as_uint(v);
                lo[k] = __float_as_uint(split_lo(v));
              }
            } else {
              const uint8_t* arow = stage + r * 128;
              // BatchNorm + ReLU of the layer that produced A, applied on the way in (a_affine: scale | shift per k, the
              // arithmetic of bn_apply_kernel): every thread needs the k-block's 32 + 32 values -- broadcast loads
              const float* const aff = P.ph[p].batch.g[w.g].a_affine;
              const int affK = P.ph[p].batch.g[w.g].K;
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                float4 v = *reinterpret_cast<const float4*>(arow + ((j ^ (r & 7)) << 4));
                if (aff != nullptr) {
                  const float4 sc = __ldg(reinterpret_cast<const float4*>(aff) + kb * (TC_BK / 4) + j);
                  const float4 sh = __ldg(reinterpret_cast<const float4*>(aff + affK) + kb * (TC_BK / 4) + j);
                  v.x = fmaxf(fmaf(v.x, sc.x, sh.x), 0.f); v.y = fmaxf(fmaf(v.y, sc.y, sh.y), 0.f);
                  v.z = fmaxf(fmaf(v.z, sc.z, sh.z), 0.f); v.w = fmaxf(fmaf(v.w, sc.w, sh.w), 0.f);
                }
                hi[4 * j + 0] = __float_as_uint(v.x); lo[4 * j + 0] = __float_as_uint(split_lo(v.x));
                hi[4 * j + 1] = __float_as_uint(v.y); lo[4 * j + 1] = __float_as_uint(split_lo(v.y));
                hi[4 * j + 2] = __float_as_uint(v.z); lo[4 * j + 2] = __float_as_uint(split_lo(v.z));
                hi[4 * j + 3] = __float_as_uint(v.w); lo[4 * j + 3] = __float_as_uint(split_lo(v.w));
              }
            }
            const uint32_t ta = tmem_base + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)(TC_ACC_COLS + s * 64);
            tc_st32(ta, hi);
            tc_st32(ta + 32u, lo);
            asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
            tc_fence_before();
          }
        } else {
          float4* src = reinterpret_cast<float4*>(stage);
          float4* dst = reinterpret_cast<float4*>(stage + C::A_LO_OFF);
          // raw A and raw B are adjacent (pair mode: only the first 64 rows of the B region are staged)
          constexpr int RAW16 = (int)(TC_TILE_BYTES + B_BYTES) / 16;
          // BatchNorm + ReLU of the layer that produced A, applied on the way in (a_affine: scale | shift per k): the A
          // tile is rewritten in place as relu(a * scale[k] + shift[k]) -- the arithmetic of bn_apply_kernel -- before
          // its lo plane is taken.  K-major A only: 16-byte slot c of row r holds k = 4 * (c ^ (r & 7)) .. + 3.
          // A thread's slots e = t + 192 i all hold the SAME four k (192 = 24 rows x 8 slots: slot index and row & 7 do
          // not change with i), so one scale and one shift vector per k-block serve all of its elements.
          const float* const aff = P.ph[p].batch.g[w.g].a_affine;
          float4 sc = make_float4(1.f, 1.f, 1.f, 1.f), sh = make_float4(0.f, 0.f, 0.f, 0.f);
          if (aff != nullptr) {
            static_assert(TC_SPLIT_THREADS % 64 == 0, "slot and row & 7 of a thread's elements must not depend on i");
            const int k4 = kb * (TC_BK / 4) + ((t & 7) ^ ((t >> 3) & 7));
            sc = __ldg(reinterpret_cast<const float4*>(aff) + k4);
            sh = __ldg(reinterpret_cast<const float4*>(aff + P.ph[p].batch.g[w.g].K) + k4);
          }
#pragma unroll
          for (int i = 0; i < (RAW16 + TC_SPLIT_THREADS - 1) / TC_SPLIT_THREADS; ++i) {
            const int e = t + i * TC_SPLIT_THREADS;
            if (e < RAW16 && three) {
              float4 v = src[e];
              if (aff != nullptr && e < TC_TILE_BYTES / 16) {
                v.x = fmaxf(fmaf(v.x, sc.x, sh.x), 0.f); v.y = fmaxf(fmaf(v.y, sc.y, sh.y), 0.f);
                v.z = fmaxf(fmaf(v.z, sc.z, sh.z), 0.f); v.w = fmaxf(fmaf(v.w, sc.w, sh.w), 0.f);
                src[e] = v;
              }
              dst[e] = make_float4(split_lo(v.x), split_lo(v.y), split_lo(v.z), split_lo(v.w));
            }
          }
        }
        if (PAIR) {
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); // generic writes -> tensor-core reads (either SM)
          if (rank == 0) mbar_arrive(split_done(s));
          else mbar_arrive_cluster(mapa_rank(split_done(s), 0));       // the leader's MMA thread waits for both CTAs
        } else {
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); // generic writes -> tensor-core reads
          mbar_arrive(split_done(s));
        }
      }
    }
    if (dbg && t == 0) dbg[12] = (unsigned long long)w_rf;
  } else {
    // ===================== accumulators + epilogue =====================
    // Eight warps in two sets: set `half` owns columns [64 half, 64 half + 64) of the tile, i.e. the slabs 2 half and
    // 2 half + 1 -- half the TMEM drain per chunk and half the epilogue per warp of a four-warp design, and 64 instead of
    // 128 accumulator registers per thread.  The sets run independently (named barrier 1 + half, own staging buffer,
    // own TMA-store issuer); the few points that concern the whole tile (split-K parking, grid barrier) use barrier 3.
    const int q = warp & 3;                 // TMEM lane quarter this warp may read
    const int half = (warp - TC_EPI_WARP0) >> 2;
    const int col0 = 64 * half;
    const bool issuer = threadIdx.x == (TC_EPI_WARP0 + 4 * half) * 32;      // this set's TMA-store thread
    const bool first = threadIdx.x == TC_EPI_WARP0 * 32;                    // one thread for tile-wide duties
    auto set_sync = [&]() { asm volatile("bar.sync %0, 128;" ::"r"(1 + half) : "memory"); };
    auto all_sync = [&]() { asm volatile("bar.sync 3, 256;" ::: "memory"); };
    uint8_t* const sbuf_ptr = staging + (size_t)half * TC_TILE_BYTES;
    int chunk = 0;
    long long w_tf = 0;
    TcUnit w;
    for (int p = 0; p < P.n_phases; ++p) {
    const TcPhase& F = P.ph[p];
    if (F.barrier_before) {
      // ---- grid barrier: every CTA's results of the earlier phases complete and visible before this phase loads them.
      // This CTA's TMA stores have finished WRITING (wait_group 0, not .read), a device-scope fence orders them before
      // the arrival; one thread arrives, waits for all CTAs of the (co-resident, persistent) grid, and releases the TMA
      // producer of this CTA through phase_go.  The two words of barrier p reset themselves: the last CTA to leave
      // zeroes them for the next launch.  A wait that exceeds 2 s raises PORL_FLAG_PEER_TIMEOUT and goes on (wrong
      // results, flagged) rather than hang the device.
      if (issuer) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
      __threadfence();
      all_sync();
      if (first) {
        int* const arrived = reinterpret_cast<int*>(P.gbar) + 2 * p;
        int* const departed = arrived + 1;
        const int n_cta = (int)gridDim.x;
        __threadfence();
        atomicAdd(arrived, 1);
        const unsigned long long t0 = gtime();
        while (ld_acquire(arrived) < n_cta) {
          __nanosleep(64);
          if (gtime() - t0 > 2000000000ull) {
            atomicOr(P.status, (uint32_t)PORL_FLAG_PEER_TIMEOUT);
            break;
          }
        }
        if (atomicAdd(departed, 1) == n_cta - 1) {
          *arrived = 0;
          *departed = 0;
        }
        __threadfence();
        mbar_arrive(phase_go);
      }
      all_sync();
    }
    for (int u = unit0; tc_decode(F, P.tile_m, u, w); u += unit_stride) {
      w.m0 += (int)rank * TC_BM;                            // pair mode: this CTA's 128 rows of the 256-row tile
      // this tile's problem, by value: loads through a reference into the parameter block would be
      // re-issued after every global store of the epilogue
      const GemmGroup& Gp = F.batch.g[w.g];
      float* const gCt = Gp.Ct;
      const float* const gBias = Gp.bias;
      const uint32_t* const gBitsIn = Gp.bits_in;
      uint32_t* const gBitsOut = Gp.bits_out;
      const float* const gRelu = gBitsIn ? nullptr : Gp.relu_src;      // the bit form of the mask takes precedence
      const size_t ldbits = (size_t)Gp.ld_bits;
      const int gM = Gp.M, gN = Gp.N, gNs = Gp.n_store, gEpi = Gp.epilogue;
      const size_t ldr = (size_t)Gp.ld_relu, ldct = (size_t)Gp.ldct;

      // ---- tile prologue, hidden behind the wait for the first chunk (these warps idle ~45 % of a tile):
      // (a) the tile's 128 bias values go to shared memory with ONE coalesced load (the epilogue then reads them as
      //     broadcast LDS; 32 predicated broadcast LDGs per thread cost 3 us per tile, measured);
      // (b) the ReLU mask of the tile as 4 x 32 bits per row (a thread reading its own row of the activation with 32
      //     row-strided 16-byte loads in the epilogue cost 6 us per tile).  In-place operation (relu_src == C) is safe:
      //     this tile's C is written by this CTA's (or, under split-K, the last-arriving slice's) epilogue only, after
      //     every slice has parked.
      const int r_in_tile = 32 * q + lane;
      if (gBias && r_in_tile < 64) {
        const int col = w.n0 + col0 + r_in_tile;
        sbias[col0 + r_in_tile] = (col < ((gN + 3) & ~3)) ? __ldg(gBias + col) : 0.f;
      }
      // ReLU mask of this thread's row.  Preferred form: one bit per element written by the forward epilogue that
      // produced the activation (bits_in): ONE 16-byte load per row, issued here and consumed only in the epilogue.
      // Fallback (relu_src given as values): the warp reads each of its 32 rows with one coalesced 512-byte request
      // and ballots v > 0 (all 32 loads in flight together).
      uint4 mw = make_uint4(0u, 0u, 0u, 0u);                 // word sl: bit j = column n0 + 32 sl + j may pass
      if (gBitsIn) {
        if (w.m0 + r_in_tile < gM) mw = *reinterpret_cast<const uint4*>(gBitsIn + (size_t)(w.m0 + r_in_tile) * ldbits + (w.n0 >> 5));
      } else if (gRelu) {
        const int col = w.n0 + 4 * lane;
        const float* mp = gRelu + (size_t)(w.m0 + 32 * q) * ldr + col;
        const bool col_ok = col < gNs;
        uint32_t rb0 = 0u, rb1 = 0u, rb2 = 0u, rb3 = 0u;      // this thread's row: bit l of rb<k> = column 4 l + k
#pragma unroll 1
        for (int r0 = 0; r0 < 32; r0 += 8) {                  // eight rows' loads in flight at a time
          float4 m4[8];
#pragma unroll
          for (int r = 0; r < 8; ++r) {
            m4[r] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (col_ok && w.m0 + 32 * q + r0 + r < gM) m4[r] = *reinterpret_cast<const float4*>(mp + (size_t)(r0 + r) * ldr);
          }
#pragma unroll
          for (int r = 0; r < 8; ++r) {
            const uint32_t b0 = __ballot_sync(0xffffffffu, m4[r].x > 0.f), b1 = __ballot_sync(0xffffffffu, m4[r].y > 0.f);
            const uint32_t b2 = __ballot_sync(0xffffffffu, m4[r].z > 0.f), b3 = __ballot_sync(0xffffffffu, m4[r].w > 0.f);
            if (lane == r0 + r) { rb0 = b0; rb1 = b1; rb2 = b2; rb3 = b3; }
          }
        }
        // into column order: word wd covers lanes 8 wd .. 8 wd + 7; spread4 moves bit i of a byte to bit 4 i
        auto spread4 = [](uint32_t x) {
          x = (x | (x << 12)) & 0x000f000fu;
          x = (x | (x << 6)) & 0x03030303u;
          return (x | (x << 3)) & 0x11111111u;
        };
        auto word_of = [&](int wd) {
          const int s = 8 * wd;
          return spread4((rb0 >> s) & 0xffu) | (spread4((rb1 >> s) & 0xffu) << 1) | (spread4((rb2 >> s) & 0xffu) << 2) |
                 (spread4((rb3 >> s) & 0xffu) << 3);
        };
        mw = make_uint4(word_of(0), word_of(1), word_of(2), word_of(3));
      }
      const bool masked = gBitsIn != nullptr || gRelu != nullptr;

      float acc[64];                                           // this set's 64 columns of this thread's row
#pragma unroll
      for (int j = 0; j < 64; ++j) acc[j] = 0.f;
      const int nchunks = (w.kb1 - w.kb0 + chunk_kb - 1) / chunk_kb;
      for (int c = 0; c < nchunks; ++c, ++chunk) {
        const int buf = chunk % TC_NBUF;
        const long long tw = dbg ? clock64() : 0;
        mbar_wait(tmem_full(buf), (uint32_t)(chunk / TC_NBUF) & 1u);
        if (dbg) w_tf += clock64() - tw;
        tc_fence_after();
        if (dbg && chunk == 0 && first) dbg[7] = gtime();
        const uint32_t taddr = tmem_base + ((uint32_t)(32 * q) << 16) + (uint32_t)(buf * TC_BN + col0);
#pragma unroll
        for (int c0 = 0; c0 < 64; c0 += 32) {
          float v[32];
          tc_ld32(taddr + (uint32_t)c0, v);
#pragma unroll
          for (int j = 0; j < 32; ++j) acc[c0 + j] += v[j];
        }
        tc_fence_before();
        if (PAIR && rank != 0) mbar_arrive_cluster(mapa_rank(tmem_empty(buf), 0));
        else mbar_arrive(tmem_empty(buf));
      }

      if (dbg && u == unit0 && first) dbg[8] = gtime();
      // ---- split-K: every slice parks its partial tile in the workspace; the LAST slice to arrive (ticket)
      // sums all partials in slice order -- a fixed order, so the result is bitwise reproducible whichever
      // CTA happens to be last -- and runs the epilogue.  No slice ever waits for another one.
      const bool split = F.batch.splits > 1;
      const int S = F.batch.splits;
      // pair mode: each CTA's 128 x 128 half of the tile is its own split-K unit (workspace slot, ticket)
      const int gt = (F.tile_begin[w.g] + w.tile) * (PAIR ? 2 : 1) + (int)rank;
      const int n_sub = F.total_tiles * (PAIR ? 2 : 1);
      float* const ws_p = F.ws;
      int* const sem_p = F.batch.sem;
      const CUtensorMap* const map_c = &F.mapC[w.g];
      // Who sums: with one unit per CTA (every split launch: units <= resident CTAs, so all slices of a tile run at the
      // same time) the FIX-UP IS DISTRIBUTED: slice s sums and finishes the slabs sl with sl % S == s (at most four
      // slices take part), after waiting for all S slices of the tile to have parked -- S x 16 KB per SM instead of
      // S x 64 KB through the one SM of the last arriver (measured: the L2 -> SM port made that 7-16 us per tile).
      // Otherwise (a CTA holds several units, a later slice may not be running yet) the last slice to arrive sums.
      const bool dist = split && F.dist_fix != 0;
      int fix_mod = 1, fix_me = 0;                               // this CTA finishes the slabs sl with sl % fix_mod == fix_me
      if (split) {
        float4* wp = reinterpret_cast<float4*>(ws_p) + ((size_t)w.split * n_sub + gt) * (32 * 128) + r_in_tile;
#pragma unroll
        for (int c4 = 0; c4 < 64; c4 += 4)
          wp[((col0 + c4) >> 2) * 128] = make_float4(acc[c4], acc[c4 + 1], acc[c4 + 2], acc[c4 + 3]);
        __threadfence();
        all_sync();                                              // both sets have parked their halves
        int* const ticket = sem_p + gt;
        if (dist) {
          if (first) {
            atomicAdd(ticket, 1);
            if (w.split < 4) {                                   // a fixer: wait until every slice has parked
              while (ld_acquire(ticket) < S) __nanosleep(32);
            }
          }
          if (w.split >= 4) continue;                            // nothing to sum for this slice
          all_sync();
          __threadfence();
          fix_mod = S;
          fix_me = w.split;
        } else {
          if (first) {
            const int old = atomicAdd(ticket, 1);
            const int last = (old == S - 1) ? 1 : 0;
            if (last) *ticket = 0;                               // self-reset for the next launch (stream-ordered)
            *flag_slot = (uint32_t)last;
          }
          all_sync();
          const bool is_last = *reinterpret_cast<volatile uint32_t*>(flag_slot) != 0;
          all_sync();                                            // flag consumed before the next unit rewrites it
          if (!is_last) continue;
          __threadfence();
        }
      }
      // ---- epilogue, one 128 x 32 slab at a time (thread = output row).  The slab comes from the register accumulator
      // or, for the last slice of a split tile, from the parked partials: all slices' loads of a slab are in flight
      // together (the accumulator registers are dead on that path; summing the whole tile slice by slice into 128 live
      // accumulators left the loads 4 at a time and cost ~18 us per tile, measured) and are added in slice order.
      // Then bias (shared memory), ReLU, ReLU mask (prefetched bits), the transposed copy as coalesced 128-byte
      // stores (lanes = consecutive rows of one Ct row), and row-major C through TMA: the slab goes to shared memory
      // (eight 16-byte slots per row XOR-swizzled like every SWIZZLE_128B tile: conflict-free) and one thread hands it
      // to the copy engine, which writes full 128-byte row segments and clips rows >= M and columns >= n_store.  The
      // stores are asynchronous: these warps return to the next tile's chunks at once.  One slab buffer per warp set:
      // before it is rewritten, wait_group.read 0 (on the set's issuing thread) guarantees the set's previous store has
      // finished reading it.
      // Pad columns (N <= col < n_store) need no special case: their B rows are out of bounds for TMA, so their
      // accumulators are exact zeros, C owns those columns (ldc >= n_store), the staged bias is zero there and relu
      // maps 0 to 0; only the transposed store is guarded.
      const int row = w.m0 + r_in_tile;
      const bool row_ok = row < gM;
      if (gBias) set_sync();                                             // this set's 64 staged bias values visible to its four warps
      auto epilogue = [&](auto split_tag) {
        constexpr bool SPLIT = decltype(split_tag)::value;
#pragma unroll
        for (int s2 = 0; s2 < 2; ++s2) {
          const int sl = 2 * half + s2;                            // slab of the tile: columns 32 sl .. 32 sl + 31
          if (w.n0 + 32 * sl >= gNs) break;
          float v[32];
          if constexpr (SPLIT) {
            if (sl % fix_mod != fix_me) continue;                  // another slice finishes this slab
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = 0.f;
            const float4* rp = reinterpret_cast<const float4*>(ws_p) + (size_t)gt * (32 * 128) + r_in_tile + (8 * sl) * 128;
            const size_t sstride = (size_t)n_sub * (32 * 128);
            for (int s0 = 0; s0 < S; s0 += 4) {
              float4 o[4][8];
#pragma unroll
              for (int d = 0; d < 4; ++d) {
#pragma unroll
                for (int j4 = 0; j4 < 8; ++j4)
                  o[d][j4] = (s0 + d < S) ? __ldcg(rp + (size_t)(s0 + d) * sstride + j4 * 128) : make_float4(0.f, 0.f, 0.f, 0.f);
              }
#pragma unroll
              for (int d = 0; d < 4; ++d) {
#pragma unroll
                for (int j4 = 0; j4 < 8; ++j4) {
                  v[4 * j4] += o[d][j4].x; v[4 * j4 + 1] += o[d][j4].y; v[4 * j4 + 2] += o[d][j4].z; v[4 * j4 + 3] += o[d][j4].w;
                }
              }
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = acc[32 * s2 + j];
          }
          if (gBias) {
#pragma unroll
            for (int j4 = 0; j4 < 8; ++j4) {
              const float4 b4 = *reinterpret_cast<const float4*>(sbias + 32 * sl + 4 * j4);
              v[4 * j4] += b4.x; v[4 * j4 + 1] += b4.y; v[4 * j4 + 2] += b4.z; v[4 * j4 + 3] += b4.w;
            }
          }
          if (gEpi == 1) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
          }
          if (masked) {
            const uint32_t word = sl == 0 ? mw.x : sl == 1 ? mw.y : sl == 2 ? mw.z : mw.w;
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (!((word >> j) & 1u)) v[j] = 0.f;
          }
          if (gBitsOut && row_ok) {                              // relu'(h) for the backward pass, one word per slab
            uint32_t word = 0u;
#pragma unroll
            for (int j = 0; j < 32; ++j) word |= (v[j] > 0.f ? 1u : 0u) << j;
            gBitsOut[(size_t)row * ldbits + (w.n0 >> 5) + sl] = word;
          }
          if (gCt && row_ok) {
            float* ct = gCt + (size_t)(w.n0 + 32 * sl) * ldct + row;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              if (w.n0 + 32 * sl + j < gN) *ct = v[j];
              ct += ldct;
            }
          }
          uint8_t* const buf = sbuf_ptr;                           // one staging slab per set: the previous store has read it
          if (issuer) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
          set_sync();
          uint8_t* const srow = buf + r_in_tile * 128;
#pragma unroll
          for (int j = 0; j < 8; ++j)
            *reinterpret_cast<float4*>(srow + ((j ^ (lane & 7)) << 4)) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic writes -> copy-engine reads
          set_sync();
          if (issuer) tma_store_2d(map_c, smem_u32(buf), w.n0 + 32 * sl, w.m0);
        }
      };
      if (split) epilogue(std::true_type{});
      else epilogue(std::false_type{});
      if (dist) all_sync();                                       // both sets have finished this fixer's slabs
      if (dist && first) {
        // second round on the ticket: the last fixer to finish resets it for the next launch
        const int nfix = S < 4 ? S : 4;
        if (atomicAdd(sem_p + gt, 1) == S + nfix - 1) sem_p[gt] = 0;
      }
      if (gBias) set_sync();                                             // bias slots free for the next tile
      if (dbg && u == unit0 && first) dbg[9] = gtime();
    }
    }
    if (issuer) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    if (dbg && first) { dbg[10] = gtime(); dbg[4] = (unsigned long long)w_tf; }
  }

  tc_fence_before();
  if (PAIR) cluster_sync_all();      // the peer may not exit (or free TMEM) while the leader's MMAs / multicasts target it
  else __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    if (PAIR)
      asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TC_TMEM_COLS)
                   : "memory");
    else
      asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TC_TMEM_COLS)
                   : "memory");
  }
  if (dbg && threadIdx.x == 0) { dbg[15] = gtime(); dbg[5] = (unsigned long long)(clock64() - clk0); }
}

// ------------------------------------------------------------------------------------------ host
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

struct MapKey {
  const void* ptr;
  int outer, inner, ld, box_rows, swz;
  bool operator==(const MapKey& o) const {
    return ptr == o.ptr && outer == o.outer && inner == o.inner && ld == o.ld && box_rows == o.box_rows && swz == o.swz;
  }
};
struct MapKeyHash {
  size_t operator()(const MapKey& k) const {
    size_t h = reinterpret_cast<size_t>(k.ptr);
    auto mix = [&](size_t v) { h ^= v + 0x9e3779b97f4a7c15ull + (h << 6) + (h >> 2); };
    mix((size_t)k.outer); mix((size_t)k.inner); mix((size_t)k.ld); mix((size_t)k.box_rows); mix((size_t)k.swz);
    return h;
  }
};
static std::unordered_map<MapKey, CUtensorMap, MapKeyHash>& map_cache() {
  static thread_local std::unordered_map<MapKey, CUtensorMap, MapKeyHash> c;
  return c;
}

// 2-D tensor map of a K-major fp32 operand (rows x K, K contiguous, row pitch ld floats):
// box {32 k, 128 rows}, SWIZZLE_128B; out-of-bounds elements (k >= K, row >= rows) read as zero.
// atom32: the MN-major operand form (128-byte swizzle with 32-byte atomicity), else the plain 128-byte swizzle.
static int make_map(CUtensorMap* out, const float* ptr, int k_extent, int rows, int ld, int box_rows = TC_BM,
                    bool atom32 = false) {
  MapKey key{ptr, rows, k_extent, ld, box_rows, atom32 ? 1 : 0};
  auto& cache = map_cache();
  auto it = cache.find(key);
  if (it != cache.end()) {
    *out = it->second;
    return PORL_OK;
  }
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled is not available from the driver");
    return PORL_ERR_CUDA;
  }
  cuuint64_t dims[2] = {(cuuint64_t)k_extent, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 4ull};
  cuuint32_t box[2] = {(cuuint32_t)TC_BK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(ptr), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, atom32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d) for extent %d x %d, ld %d", (int)r, rows, k_extent, ld);
    return PORL_ERR_CUDA;
  }
  if (cache.size() > 4096) cache.clear();
  cache.emplace(key, *out);
  return PORL_OK;
}

bool gemm_tc_eligible(const GemmGroup& g, bool a_kmajor, bool b_kmajor) {
  (void)b_kmajor;      // both operand forms are served: K-major and MN-major shared-memory descriptors
  if (g.A_lo || g.B_lo || g.C_lo || g.colsum) return false;
  if (g.N < 32 || g.K < 32 || g.M < 32) return false;
  if ((double)g.M * g.N * g.K < (double)(1 << 20)) return false;       // tiny problems: FFMA kernel
  if ((g.lda & 3) || (g.ldb & 3) || !aligned16(g.A) || !aligned16(g.B)) return false;
  if ((g.ldc & 3) || (g.n_store & 3) || !aligned16(g.C)) return false;   // 16-byte row-major stores
  if (g.epilogue == 2) return false;                                     // tanh heads are narrow: FFMA kernel
  if (g.bias && !aligned16(g.bias)) return false;                        // read as float4 through pad4(N)
  if (g.relu_src && ((g.ld_relu & 3) || g.ld_relu < g.n_store || !aligned16(g.relu_src))) return false;
  if (g.Ct && g.ldct < g.M) return false;
  if (g.a_affine && (!a_kmajor || (g.K % TC_BK) != 0 || !aligned16(g.a_affine))) return false;
  return true;
}

// clusters of two CTAs that can be resident at once (pair mode grid); 0 when cluster launch is unavailable
static int max_pairs(int smem_bytes) {
  static int cached = -1;
  if (cached >= 0) return cached;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2, 1, 1);
  cfg.blockDim = dim3(TC_THREADS, 1, 1);
  cfg.dynamicSmemBytes = (size_t)smem_bytes;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 2;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, gemm_tc_kernel<false, true>, &cfg) != cudaSuccess) {
    cudaGetLastError();
    n = 0;
  }
  cached = n;
  return cached;
}

// One persistent launch over up to kMaxPhases grouped contractions (see TcPhase).  Every phase decides its own split-K
// factor; phases get disjoint slices of the split-K workspace and of the ticket words.  PORL_ERR_UNSUPPORTED when a
// problem does not fit the kernel (the caller falls back to one launch per phase / the FFMA kernel).
int launch_gemm_tc_phases(porl_ctx* ctx, int n_phases, TcPhaseSpec* specs, cudaStream_t stream) {
  if (n_phases < 1 || n_phases > kMaxPhases) return PORL_ERR_UNSUPPORTED;
  // Split-K whenever the tiles alone would leave most SMs idle: weight gradients (K = batch) and, at small batches,
  // the forward / input-gradient contractions too -- the slices of a tile share its final sum and run the whole
  // epilogue (bias, ReLU, ReLU mask, transposed copy), so a batch-256 layer is 100+ short units instead of 32 tiles
  // with a 16-k-block serial loop each.  The split factor is decided from the 128-row tile count and the SM count
  // alone, so all kernel variants cut K at the same places (bitwise identical results).
  static TcParams P;                 // 13 KB: filled per launch (the host side of the library is single-threaded per context)
  bool any_mn = false;
  for (int p = 0; p < n_phases; ++p) any_mn = any_mn || !specs[p].a_kmajor || !specs[p].b_kmajor;
  // the A-in-TMEM variant un-swizzles K-major rows and the pair variant halves a K-major B tile: MN-major operands and
  // multi-phase launches take the all-in-shared-memory variant.  Pair mode (variant 2): every problem at least one
  // full 256-row tile high, otherwise the second CTA idles.
  bool pair = ctx->tc_a_in_tmem == 2 && !any_mn && n_phases == 1;
  for (int i = 0; i < specs[0].batch->n_groups && pair; ++i)
    if (specs[0].batch->g[i].M < 2 * TC_BM) pair = false;
  bool bn_fused = false;             // operand affine: in the shared-memory and the tensor-memory variant, not in pair mode
  for (int p = 0; p < n_phases; ++p)
    for (int i = 0; i < specs[p].batch->n_groups; ++i)
      bn_fused = bn_fused || specs[p].batch->g[i].a_affine;
  if (bn_fused) pair = false;
  // A in tensor memory: the splitters un-swizzle A rows of either form (K-major rows, or the lines of an MN-major tile);
  // B's split is a flat pass over the tile and the MMA's B descriptor carries its layout
  const bool atm = ctx->tc_a_in_tmem == 1 && n_phases == 1;
  P.tile_m = pair ? 2 * TC_BM : TC_BM;
  P.n_phases = n_phases;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(gemm_tc_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         TcCfg<false>::SMEM_BYTES);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(gemm_tc_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               TcCfg<true>::SMEM_BYTES);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(gemm_tc_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               TcCfg<false, true>::SMEM_BYTES);
    if (e != cudaSuccess) {
      set_error("cudaFuncSetAttribute(max dynamic smem) failed: %s", cudaGetErrorString(e));
      return PORL_ERR_CUDA;
    }
    configured = true;
  }
  int n_pairs = 0;
  if (pair) {
    n_pairs = max_pairs(TcCfg<false, true>::SMEM_BYTES);
    if (n_pairs <= 0) {
      set_error("pair mode: cluster launch of 2 CTAs is not available on this device");
      return PORL_ERR_UNSUPPORTED;
    }
  }
  // work slots: one per SM, or one per resident CTA pair; sub-tiles (128 x 128, one per CTA) for the split-K buffers
  const int sms = pair ? n_pairs : ctx->sm_count;
  const int sub = pair ? 2 : 1;
  long long ws_used = 0, sem_used = 0, max_units = 0;
  const long long ws_tiles = (long long)(kTcWsFloats / (TC_BM * TC_BN));
  for (int p = 0; p < n_phases; ++p) {
    TcPhase& F = P.ph[p];
    GemmBatch& batch = *specs[p].batch;
    const bool a_mn = !specs[p].a_kmajor, b_mn = !specs[p].b_kmajor;
    int min_nkb = 1 << 30, tiles = 0, tiles128 = 0;
    for (int i = 0; i < batch.n_groups; ++i) {
      const GemmGroup& g = batch.g[i];
      if (!gemm_tc_eligible(g, specs[p].a_kmajor, specs[p].b_kmajor)) return PORL_ERR_UNSUPPORTED;
      const int nkb = (g.K + TC_BK - 1) / TC_BK;
      min_nkb = nkb < min_nkb ? nkb : min_nkb;
      tiles128 += ((g.M + TC_BM - 1) / TC_BM) * ((g.n_store + TC_BN - 1) / TC_BN);
      F.tile_begin[i] = tiles;
      F.tiles_n[i] = (g.n_store + TC_BN - 1) / TC_BN;
      tiles += ((g.M + P.tile_m - 1) / P.tile_m) * F.tiles_n[i];
    }
    for (int i = batch.n_groups; i <= kMaxGroups; ++i) F.tile_begin[i] = tiles;
    for (int i = batch.n_groups; i < kMaxGroups; ++i) F.tiles_n[i] = 1;
    F.total_tiles = tiles;
    for (int i = 0; i < batch.n_groups; ++i) {
      const GemmGroup& g = batch.g[i];
      // K-major: (rows x K), box {32 k, 128 rows}; MN-major: (K x rows), box {32 rows, 32 k}
      int rc = a_mn ? make_map(&F.mapA[i], g.A, g.M, g.K, g.lda, TC_BK, true) : make_map(&F.mapA[i], g.A, g.K, g.M, g.lda);
      if (rc == PORL_OK) rc = b_mn ? make_map(&F.mapB[i], g.B, g.N, g.K, g.ldb, TC_BK, true) : make_map(&F.mapB[i], g.B, g.K, g.N, g.ldb);
      if (rc == PORL_OK) rc = make_map(&F.mapC[i], g.C, g.n_store, g.M, g.ldc);
      if (rc == PORL_OK && pair) rc = make_map(&P.mapB2[i], g.B, g.K, g.N, g.ldb, TC_BN / 2);
      if (rc != PORL_OK) return rc;
    }
    int splits = 1;
    if (2 * tiles128 <= ctx->sm_count) {
      int want = ctx->sm_count / tiles128;
      // at least 4 k-blocks (128 k) per slice.  PORL_TC_SPLIT_KB overrides: measured at batch 256, 4 -> 5106, 3 -> 5188,
      // 2 -> 5330, 1 -> 5225 updates/s (a slice's main loop shrinks, the fix-up's round trips do not grow); the default
      // stays 4 because the cut changes the summation grouping and with it the free-running trajectory the 1,000-update
      // drift test of tests/test_bench_configs.py was validated on (its per-step loss bound tripped at 2)
      static const int kb_per_slice = getenv("PORL_TC_SPLIT_KB") ? atoi(getenv("PORL_TC_SPLIT_KB")) : 4;
      int cap = min_nkb / (kb_per_slice > 0 ? kb_per_slice : 4);
      splits = want < cap ? want : cap;
      const long long room = ws_tiles - ws_used;
      if ((long long)splits * tiles * sub > room) splits = (int)(room / ((long long)tiles * sub));
      if (splits > 32) splits = 32;
      if (splits < 1) splits = 1;
    }
    if (sem_used + (long long)tiles * sub > kSemInts) splits = 1;
    batch.splits = splits;
    batch.sem = ctx->sem + sem_used;
    batch.sem_stride = 0;
    F.ws = ctx->tc_ws + ws_used * (TC_BM * TC_BN);
    if (splits > 1) {
      ws_used += (long long)splits * tiles * sub;
      sem_used += (long long)tiles * sub;
    }
    F.batch = batch;
    F.a_mn = a_mn ? 1 : 0;
    F.b_mn = b_mn ? 1 : 0;
    F.barrier_before = (p > 0 && specs[p].barrier_before) ? 1 : 0;
    const long long units = (long long)tiles * splits;
    max_units = units > max_units ? units : max_units;
    // every slice of the phase on its own CTA of the persistent (co-resident) grid: a fixer waits for parks of its own
    // phase only, and a CTA reaches a phase whatever happens in that phase elsewhere, so the waits cannot form a cycle
    F.dist_fix = (splits > 1 && units <= sms && !pair && getenv("PORL_TC_LAST_ARRIVER") == nullptr) ? 1 : 0;
  }
  P.chunk_kb = ctx->tc_chunk_kb > 0 ? ctx->tc_chunk_kb : 1;
  P.terms = ctx->tc_terms == 1 ? 1 : 3;
  P.dbg = ctx->tc_dbg;
  P.gbar = ctx->counters + kGridBarBase;
  P.status = ctx->counters;
  const int grid = (int)(max_units < sms ? max_units : sms);
  if (pair) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * grid, 1, 1);
    cfg.blockDim = dim3(TC_THREADS, 1, 1);
    cfg.dynamicSmemBytes = TcCfg<false, true>::SMEM_BYTES;
    cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, gemm_tc_kernel<false, true>, P);
    if (e != cudaSuccess) {
      set_error("cluster launch of the pair-mode contraction failed: %s", cudaGetErrorString(e));
      return PORL_ERR_CUDA;
    }
  } else if (atm) {
    launch_pdl(ctx, gemm_tc_kernel<true, false>, grid, TC_THREADS, TcCfg<true>::SMEM_BYTES, stream, P);
  } else {
    launch_pdl(ctx, gemm_tc_kernel<false, false>, grid, TC_THREADS, TcCfg<false>::SMEM_BYTES, stream, P);
  }
  ctx->tc_launches++;
  PORL_LAUNCHED(ctx);
  return PORL_OK;
}

int launch_gemm_tc(porl_ctx* ctx, bool a_kmajor, bool b_kmajor, GemmBatch& batch, cudaStream_t stream) {
  TcPhaseSpec spec{&batch, a_kmajor, b_kmajor, false};
  return launch_gemm_tc_phases(ctx, 1, &spec, stream);
}

}  // namespace porl
