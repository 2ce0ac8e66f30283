// Device-side index sampler: a keyed pseudo-random BIJECTION of [0, n), evaluated point-wise.
//
// The reference draws its minibatch indices on the host: DataLoader(shuffle=True) materialises a torch.randperm(N)
// per epoch (por_train.py:61) and ReplayBuffer.sample calls np.random.choice(size, B, replace=False), which is O(size)
// per call (35 ms at 1 M rows, replaybuffer.py:61).  Both want "B distinct uniformly shuffled rows".  Here position i
// of the shuffled order is computed directly, P(i), in O(1) and without any table: an alternating Feistel network on
// ceil(log2 n) bits (each round XORs one half with a hash of the other half and a round key: invertible, hence a
// bijection on the power-of-two domain) with cycle walking (re-apply while the value is >= n, which restricts a
// bijection of the larger domain to a bijection of [0, n)).  A minibatch is P(first .. first + B - 1): distinct by
// construction; an epoch is first = 0, B, 2B, ...; a new (seed, stream) gives a new permutation.  Same DISTRIBUTION
// as the reference's samplers for the purposes of SGD, not the same index stream (parity tests feed recorded indices).
#include "common.cuh"

namespace porl {

__host__ __device__ inline uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

__host__ __device__ inline uint32_t round_fn(uint32_t x, uint32_t k) {
  x ^= k;
  x *= 0x9E3779B1u;
  x ^= x >> 15;
  x *= 0x85EBCA77u;
  x ^= x >> 13;
  x *= 0xC2B2AE3Du;
  x ^= x >> 16;
  return x;
}

constexpr int kFeistelRounds = 8;

struct FeistelKey {
  uint32_t k[kFeistelRounds];
  int bits_lo, bits_hi;       // low half / high half widths (bits_lo + bits_hi = domain bits, each <= 31)
};

__host__ __device__ inline uint64_t feistel_once(uint64_t v, const FeistelKey& key) {
  const uint32_t mlo = (1u << key.bits_lo) - 1u, mhi = (1u << key.bits_hi) - 1u;
  uint32_t lo = (uint32_t)v & mlo, hi = (uint32_t)(v >> key.bits_lo) & mhi;
#pragma unroll
  for (int r = 0; r < kFeistelRounds; ++r) {
    if (r & 1) hi ^= round_fn(lo, key.k[r]) & mhi;
    else lo ^= round_fn(hi, key.k[r]) & mlo;
  }
  return ((uint64_t)hi << key.bits_lo) | lo;
}

__global__ void __launch_bounds__(256) permutation_kernel(int64_t n, FeistelKey key, int64_t first, int64_t count,
                                                          int64_t* __restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < count; i += (int64_t)gridDim.x * 256) {
    uint64_t v = (uint64_t)((first + i) % n);
    do v = feistel_once(v, key); while (v >= (uint64_t)n);
    out[i] = (int64_t)v;
  }
}

}  // namespace porl

using namespace porl;

extern "C" int porl_sample_permutation(porl_ctx_t* ctx, int64_t n, uint64_t seed, uint64_t stream_id, int64_t first,
                                       int64_t count, int64_t* out, porl_stream_t stream) {
  PORL_ON_DEVICE(ctx);
  PORL_CHECK_ARG(ctx && out, "porl_sample_permutation: null argument");
  PORL_CHECK_ARG(n >= 1 && n < ((int64_t)1 << 62) && first >= 0 && count >= 0, "porl_sample_permutation: n=%lld first=%lld count=%lld",
                 (long long)n, (long long)first, (long long)count);
  if (count == 0) return PORL_OK;
  int bits = 2;                                   // at least one bit per half
  while (((uint64_t)1 << bits) < (uint64_t)n) ++bits;
  FeistelKey key;
  key.bits_lo = bits / 2;
  key.bits_hi = bits - key.bits_lo;
  uint64_t s = splitmix64(seed ^ splitmix64(stream_id));
  for (int r = 0; r < kFeistelRounds; ++r) {
    s = splitmix64(s);
    key.k[r] = (uint32_t)(s >> 32);
  }
  cudaStream_t st = as_stream(stream);
  int64_t grid = (count + 255) / 256;
  if (grid > (int64_t)ctx->sm_count * 8) grid = (int64_t)ctx->sm_count * 8;
  ProfScope prof(ctx, PORL_CAT_GATHER, (double)count * 8.0, st);
  permutation_kernel<<<(unsigned)grid, 256, 0, st>>>(n, key, first, count, out);
  PORL_LAUNCHED(ctx);
  return PORL_OK;
}
